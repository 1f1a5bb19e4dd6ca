"""ctypes binding of libcpoly.so (the C ABI declared in include/cpoly.h).

There is deliberately NO fallback: if the shared library is missing or a call fails, we raise.
torch is used only for device memory, streams and dtypes.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, Structure, c_char_p, c_double, c_float, c_int32, c_int64, c_size_t, c_uint32, c_void_p

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libcpoly.so")

F32, BF16 = 0, 1
REP = {"cartesian": 0, "polar": 1, "polar_fixed": 2}
PATH_AUTO, PATH_SIMT, PATH_TCGEN05 = 0, 1, 2

EXPORTED_SYMBOLS = (
    "cpoly_abi_version", "cpoly_last_error", "cpoly_device_supports_tcgen05", "cpoly_dcn_tcgen05_eligible",
    "cpoly_dcn_packed_weight_bytes", "cpoly_dcn_pack_weights", "cpoly_dcn_forward_workspace_bytes", "cpoly_dcn_forward",
    "cpoly_idaup_upsample_forward", "cpoly_idaup_upsample_backward",
    "cpoly_dcn_backward_workspace_bytes", "cpoly_dcn_backward", "cpoly_decode_workspace_bytes", "cpoly_decode",
    "cpoly_rasterise_workspace_bytes", "cpoly_rasterise", "cpoly_poly_iou_forward", "cpoly_poly_iou_backward",
    "cpoly_stem_pack_weights", "cpoly_stem_conv7x7_forward", "cpoly_maxpool2x2_nhwc_forward",
    "cpoly_offset_conv_packed_bytes", "cpoly_offset_conv_pack_weights", "cpoly_offset_conv_forward",
    "cpoly_decode_static_max_bytes", "cpoly_decode_static",
    "cpoly_bn_eligible", "cpoly_bn_workspace_bytes", "cpoly_bn_train_forward", "cpoly_bn_train_backward",
    "cpoly_idaup_upsample_backward_workspace_bytes",
    "cpoly_rasterise_fill_workspace_bytes", "cpoly_rasterise_fill",
)


class Tensor4(Structure):
    _fields_ = [("ptr", c_void_p), ("sb", c_int64), ("sc", c_int64), ("sh", c_int64), ("sw", c_int64)]


class DcnFwdArgs(Structure):
    _fields_ = [(n, c_int32) for n in (
        "B", "Cin", "Cout", "H", "W", "kh", "kw", "stride_h", "stride_w", "pad_h", "pad_w", "dil_h", "dil_w",
        "deformable_groups", "dtype", "om_dtype", "mask_is_logit", "path")] + [
        ("x", Tensor4), ("offset", Tensor4), ("mask", Tensor4),
        ("weight", c_void_p), ("bias", c_void_p), ("weight_packed", c_void_p), ("y", Tensor4),
        ("out_scale", c_void_p), ("relu", c_int32), ("reserved0", c_int32),
        ("offset_bias", c_void_p), ("mask_bias", c_void_p),
        ("workspace", c_void_p), ("workspace_bytes", c_size_t)]


class DcnBwdArgs(Structure):
    _fields_ = [("f", DcnFwdArgs), ("gout", Tensor4), ("gx", Tensor4), ("goffset", Tensor4), ("gmask", Tensor4),
                ("gweight", c_void_p), ("gbias", c_void_p), ("workspace", c_void_p), ("workspace_bytes", c_size_t),
                ("gom_dtype", c_int32), ("reserved1", c_int32)]


class UpsampleArgs(Structure):
    _fields_ = [(n, c_int32) for n in ("B", "C", "H", "W", "f", "dtype", "weight_tap_major", "reserved0")] + [
        ("x", Tensor4), ("weight", c_void_p), ("skip", Tensor4), ("y", Tensor4)]


class DecodeArgs(Structure):
    _fields_ = [(n, c_int32) for n in ("B", "C", "H", "W", "n_poly", "K", "rep", "heat_is_logit", "dtype")] + [
        ("heat", Tensor4), ("polys", Tensor4), ("depth", Tensor4), ("reg", Tensor4),
        ("out", c_void_p), ("out_inds", c_void_p), ("workspace", c_void_p), ("workspace_bytes", c_size_t),
        ("post_affine", c_void_p)]


class RasterArgs(Structure):
    _fields_ = [(n, c_int32) for n in ("B", "K", "N", "W", "H")] + [
        ("thresh", c_double), ("occluder_score", c_double), ("min_area", c_int32), ("skip_classes", c_uint32),
        ("dets", c_void_p), ("order", c_void_p), ("count", c_void_p), ("verts", c_void_p), ("area", c_void_p),
        ("keep", c_void_p), ("conf", c_void_p), ("status", c_void_p), ("masks", c_void_p),
        ("workspace", c_void_p), ("workspace_bytes", c_size_t)]


class PolyIouArgs(Structure):
    _fields_ = [(n, c_int32) for n in ("B", "M", "N", "H", "W", "dtype")] + [
        ("poly", Tensor4), ("ind", c_void_p), ("mask", c_void_p), ("target", c_void_p), ("iou", c_void_p),
        ("dpred", c_void_p), ("status", c_void_p), ("stats", c_void_p), ("grad_loss", c_void_p), ("grad_poly", Tensor4)]


class StemArgs(Structure):
    _fields_ = [(n, c_int32) for n in ("B", "H", "W", "in_dtype", "out_dtype", "relu")] + [
        ("x", Tensor4), ("weight_packed", c_void_p), ("bias", c_void_p), ("y", c_void_p)]


class MaxPoolArgs(Structure):
    _fields_ = [(n, c_int32) for n in ("B", "C", "H", "W", "dtype", "reserved0")] + [("x", c_void_p), ("y", c_void_p)]


class OffsetConvArgs(Structure):
    _fields_ = [(n, c_int32) for n in ("B", "Cin", "H", "W", "om_dtype", "reserved0")] + [
        ("x", c_void_p), ("weight_packed", c_void_p), ("bias", c_void_p), ("om", c_void_p)]


class RasterFillArgs(Structure):
    _fields_ = [(n, c_int32) for n in ("B", "K", "N", "W", "H", "inclusive")] + [
        ("thresh", c_double), ("occluder_score", c_double), ("dets", c_void_p), ("count", c_void_p), ("slot_det", c_void_p),
        ("area", c_void_p), ("status", c_void_p), ("masks", c_void_p), ("workspace", c_void_p), ("workspace_bytes", c_size_t)]


class BnFwdArgs(Structure):
    _fields_ = [("N", c_int64), ("C", c_int32), ("dtype", c_int32), ("relu", c_int32), ("reserved0", c_int32),
                ("x", c_void_p), ("y", c_void_p), ("weight", c_void_p), ("bias", c_void_p), ("running_mean", c_void_p),
                ("running_var", c_void_p), ("momentum", c_float), ("eps", c_float), ("save", c_void_p),
                ("workspace", c_void_p), ("workspace_bytes", c_size_t)]


class BnBwdArgs(Structure):
    _fields_ = [("N", c_int64), ("C", c_int32), ("dtype", c_int32), ("relu", c_int32), ("reserved0", c_int32),
                ("x", c_void_p), ("gy", c_void_p), ("gx", c_void_p), ("save", c_void_p), ("gweight", c_void_p),
                ("gbias", c_void_p), ("workspace", c_void_p), ("workspace_bytes", c_size_t)]


class UpsampleBwdArgs(Structure):
    _fields_ = [(n, c_int32) for n in ("B", "C", "H", "W", "f", "dtype")] + [
        ("x", Tensor4), ("gy", Tensor4), ("weight", c_void_p), ("gx", Tensor4), ("gweight", c_void_p),
        ("workspace", c_void_p), ("workspace_bytes", c_size_t)]


_lib = None
# kernels launched by the shims since import (our kernels only; bench.py reports it as gpu_launches)
LAUNCHES = {"dcn_forward": 0, "dcn_backward": 0, "pack_weights": 0, "upsample": 0, "upsample_bwd": 0, "decode": 0, "rasterise": 0, "poly_iou": 0, "stem": 0, "maxpool": 0, "offset_conv": 0}


def lib():
    """The loaded library; raises (loudly) when it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -m centerpoly_b200.build` "
                "(centerpoly_b200 has no CPU or PyTorch fallback by design)")
        L = ctypes.CDLL(LIB_PATH)
        L.cpoly_abi_version.restype = c_int32
        L.cpoly_last_error.restype = c_char_p
        L.cpoly_device_supports_tcgen05.restype = c_int32
        L.cpoly_dcn_tcgen05_eligible.restype = c_int32
        L.cpoly_dcn_tcgen05_eligible.argtypes = [POINTER(DcnFwdArgs)]
        L.cpoly_dcn_packed_weight_bytes.restype = c_size_t
        L.cpoly_dcn_packed_weight_bytes.argtypes = [c_int32] * 5
        L.cpoly_dcn_pack_weights.restype = c_int32
        L.cpoly_dcn_pack_weights.argtypes = [c_void_p, c_void_p, c_int32, c_int32, c_int32, c_int32, c_int32, c_void_p]
        L.cpoly_dcn_forward_workspace_bytes.restype = c_size_t
        L.cpoly_dcn_forward_workspace_bytes.argtypes = [POINTER(DcnFwdArgs)]
        L.cpoly_idaup_upsample_forward.restype = c_int32
        L.cpoly_idaup_upsample_forward.argtypes = [POINTER(UpsampleArgs), c_void_p]
        L.cpoly_dcn_forward.restype = c_int32
        L.cpoly_dcn_forward.argtypes = [POINTER(DcnFwdArgs), c_void_p]
        L.cpoly_dcn_backward_workspace_bytes.restype = c_size_t
        L.cpoly_dcn_backward_workspace_bytes.argtypes = [POINTER(DcnBwdArgs)]
        L.cpoly_dcn_backward.restype = c_int32
        L.cpoly_dcn_backward.argtypes = [POINTER(DcnBwdArgs), c_void_p]
        L.cpoly_decode_workspace_bytes.restype = c_size_t
        L.cpoly_decode_workspace_bytes.argtypes = [c_int32] * 5
        L.cpoly_decode.restype = c_int32
        L.cpoly_decode.argtypes = [POINTER(DecodeArgs), c_void_p]
        L.cpoly_decode_static_max_bytes.restype = c_size_t
        L.cpoly_decode_static_max_bytes.argtypes = []
        L.cpoly_decode_static.restype = c_int32
        L.cpoly_decode_static.argtypes = [POINTER(DecodeArgs), c_void_p]
        L.cpoly_rasterise_fill_workspace_bytes.restype = c_size_t
        L.cpoly_rasterise_fill_workspace_bytes.argtypes = [c_int32] * 3
        L.cpoly_rasterise_fill.restype = c_int32
        L.cpoly_rasterise_fill.argtypes = [POINTER(RasterFillArgs), c_void_p]
        L.cpoly_bn_eligible.restype = c_int32
        L.cpoly_bn_eligible.argtypes = [c_int32, c_int32]
        L.cpoly_bn_workspace_bytes.restype = c_size_t
        L.cpoly_bn_workspace_bytes.argtypes = [c_int32]
        L.cpoly_bn_train_forward.restype = c_int32
        L.cpoly_bn_train_forward.argtypes = [POINTER(BnFwdArgs), c_void_p]
        L.cpoly_bn_train_backward.restype = c_int32
        L.cpoly_bn_train_backward.argtypes = [POINTER(BnBwdArgs), c_void_p]
        L.cpoly_rasterise_workspace_bytes.restype = c_size_t
        L.cpoly_rasterise_workspace_bytes.argtypes = [c_int32] * 2
        L.cpoly_rasterise.restype = c_int32
        L.cpoly_rasterise.argtypes = [POINTER(RasterArgs), c_void_p]
        L.cpoly_poly_iou_forward.restype = c_int32
        L.cpoly_poly_iou_forward.argtypes = [POINTER(PolyIouArgs), c_void_p]
        L.cpoly_poly_iou_backward.restype = c_int32
        L.cpoly_poly_iou_backward.argtypes = [POINTER(PolyIouArgs), c_void_p]
        L.cpoly_stem_pack_weights.restype = c_int32
        L.cpoly_stem_pack_weights.argtypes = [c_void_p, c_int32, c_void_p, c_void_p]
        L.cpoly_stem_conv7x7_forward.restype = c_int32
        L.cpoly_stem_conv7x7_forward.argtypes = [POINTER(StemArgs), c_void_p]
        L.cpoly_maxpool2x2_nhwc_forward.restype = c_int32
        L.cpoly_maxpool2x2_nhwc_forward.argtypes = [POINTER(MaxPoolArgs), c_void_p]
        L.cpoly_idaup_upsample_backward.restype = c_int32
        L.cpoly_idaup_upsample_backward.argtypes = [POINTER(UpsampleBwdArgs), c_void_p]
        L.cpoly_idaup_upsample_backward_workspace_bytes.restype = c_size_t
        L.cpoly_idaup_upsample_backward_workspace_bytes.argtypes = [POINTER(UpsampleBwdArgs)]
        L.cpoly_offset_conv_packed_bytes.restype = c_size_t
        L.cpoly_offset_conv_packed_bytes.argtypes = [c_int32]
        L.cpoly_offset_conv_pack_weights.restype = c_int32
        L.cpoly_offset_conv_pack_weights.argtypes = [c_void_p, c_void_p, c_int32, c_int32, c_void_p]
        L.cpoly_offset_conv_forward.restype = c_int32
        L.cpoly_offset_conv_forward.argtypes = [POINTER(OffsetConvArgs), c_void_p]
        if L.cpoly_abi_version() != 10:
            raise RuntimeError("libcpoly.so ABI version mismatch; rebuild with `python -m centerpoly_b200.build --force`")
        _lib = L
    return _lib


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = lib().cpoly_last_error()
        raise RuntimeError(f"{what} failed (code {rc}): {msg.decode() if msg else '?'}")


def dtype_code(t: torch.Tensor) -> int:
    if t.dtype == torch.float32:
        return F32
    if t.dtype == torch.bfloat16:
        return BF16
    raise RuntimeError(f"centerpoly_b200 supports float32 and bfloat16 tensors, got {t.dtype}")


def t4(t: torch.Tensor | None) -> Tensor4:
    """Base pointer + element strides of a 4-d tensor (any memory format)."""
    if t is None:
        return Tensor4(None, 0, 0, 0, 0)
    assert t.dim() == 4
    sb, sc, sh, sw = t.stride()
    return Tensor4(t.data_ptr(), sb, sc, sh, sw)


def require_cuda(*tensors) -> torch.device:
    dev = None
    for t in tensors:
        if t is None:
            continue
        if not t.is_cuda:
            raise RuntimeError("centerpoly_b200 runs on CUDA (B200) tensors only; there is no CPU path")
        if dev is None:
            dev = t.device
        elif t.device != dev:
            raise RuntimeError(f"tensors on different devices: {dev} vs {t.device}")
    return dev


def stream_ptr(dev: torch.device) -> c_void_p:
    return c_void_p(torch.cuda.current_stream(dev).cuda_stream)


def workspace(nbytes: int, dev: torch.device) -> torch.Tensor:
    return torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=dev)
