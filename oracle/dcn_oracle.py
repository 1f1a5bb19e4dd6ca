"""TEST INFRASTRUCTURE — NOT PRODUCT CODE.

ctypes/numpy front-end of the C DCNv2 oracle (oracle/dcn_oracle.c).  Only tests/,
``__graft_entry__.smoke()`` and bench.py's CPU-baseline legs may import this module.

Provenance and pinning are described in the header of dcn_oracle.c: the arithmetic is the DCNv2
submodule (pinned dee2e77e..., missing from /root/reference), call sites
/root/reference/src/lib/models/networks/pose_dla_dcn.py:16,354; pinned against
torchvision.ops.deform_conv2d fp64 through tests/golden/dcn_*.npz.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libdcn_oracle.so")
_lib = None


def build(force: bool = False) -> str:
    """Compile oracle/libdcn_oracle.so with gcc (a few seconds)."""
    src = [os.path.join(_HERE, f) for f in ("dcn_oracle.c", "dcn_oracle_impl.h")]
    stale = (not os.path.exists(_LIB_PATH)) or any(
        os.path.getmtime(s) > os.path.getmtime(_LIB_PATH) for s in src)
    if force or stale:
        subprocess.check_call(["make", "-s", "-C", _HERE, "-B", "libdcn_oracle.so"])
    return _LIB_PATH


def _load():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            build()
        _lib = ctypes.CDLL(_LIB_PATH)
    return _lib


def _ptr(a):
    return a.ctypes.data_as(ctypes.c_void_p) if a is not None else None


def _prep(arrs, dtype):
    return [None if a is None else np.ascontiguousarray(a, dtype=dtype) for a in arrs]


def _geom(x, weight, stride, padding, dilation):
    B, Cin, H, W = x.shape
    Cout, _, kh, kw = weight.shape
    sh, sw = (stride, stride) if isinstance(stride, int) else stride
    ph, pw = (padding, padding) if isinstance(padding, int) else padding
    dh, dw = (dilation, dilation) if isinstance(dilation, int) else dilation
    Ho = (H + 2 * ph - (dh * (kh - 1) + 1)) // sh + 1
    Wo = (W + 2 * pw - (dw * (kw - 1) + 1)) // sw + 1
    ints = [B, Cin, Cout, H, W, kh, kw, sh, sw, ph, pw, dh, dw]
    return ints, Ho, Wo


def dcn_forward(x, offset, mask, weight, bias=None, stride=1, padding=1, dilation=1,
                deformable_groups=1, dtype=np.float64):
    """y = DCNv2(x; offset, mask, weight, bias); dense NCHW numpy arrays in, NCHW array out."""
    lib = _load()
    x, offset, mask, weight, bias = _prep([x, offset, mask, weight, bias], dtype)
    ints, Ho, Wo = _geom(x, weight, stride, padding, dilation)
    assert offset.shape == (ints[0], deformable_groups * 2 * ints[5] * ints[6], Ho, Wo), offset.shape
    assert mask.shape == (ints[0], deformable_groups * ints[5] * ints[6], Ho, Wo), mask.shape
    y = np.empty((ints[0], ints[2], Ho, Wo), dtype=dtype)
    fn = lib.dcn_oracle_forward_f64 if dtype == np.float64 else lib.dcn_oracle_forward_f32
    fn.restype = None
    fn(_ptr(x), _ptr(offset), _ptr(mask), _ptr(weight), _ptr(bias), _ptr(y),
       *[ctypes.c_int(v) for v in ints + [deformable_groups]])
    return y


def dcn_backward(x, offset, mask, weight, grad_out, stride=1, padding=1, dilation=1,
                 deformable_groups=1, dtype=np.float64):
    """Returns (grad_x, grad_offset, grad_mask, grad_weight, grad_bias) as numpy arrays."""
    lib = _load()
    x, offset, mask, weight, grad_out = _prep([x, offset, mask, weight, grad_out], dtype)
    ints, Ho, Wo = _geom(x, weight, stride, padding, dilation)
    assert grad_out.shape == (ints[0], ints[2], Ho, Wo)
    gx = np.empty_like(x)
    go = np.empty_like(offset)
    gm = np.empty_like(mask)
    gw = np.empty_like(weight)
    gb = np.empty((ints[2],), dtype=dtype)
    fn = lib.dcn_oracle_backward_f64 if dtype == np.float64 else lib.dcn_oracle_backward_f32
    fn.restype = None
    fn(_ptr(x), _ptr(offset), _ptr(mask), _ptr(weight), _ptr(grad_out), _ptr(gx), _ptr(go), _ptr(gm),
       _ptr(gw), _ptr(gb), *[ctypes.c_int(v) for v in ints + [deformable_groups]])
    return gx, go, gm, gw, gb
