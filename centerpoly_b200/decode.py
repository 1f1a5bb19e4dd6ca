"""Signature-identical replacement of the reference's ``polydet_decode``
(src/lib/models/decode.py:512-670; callers src/lib/detectors/polydet.py:39 and
src/lib/trains/polydet.py:222-225), backed by ONE fused CUDA kernel in libcpoly.so
(centerpoly_b200/csrc/decode.cu).  ``heat`` is POST-sigmoid exactly as in the reference
(detectors/polydet.py:28 applies ``sigmoid_`` before the call); ``polydet_decode_from_logits``
additionally fuses that sigmoid.  No CPU path, no PyTorch fallback.
"""
from __future__ import annotations

from ctypes import byref

import torch

from . import _lib

__all__ = ["polydet_decode", "polydet_decode_from_logits"]


def _decode(heat, polys, depth, reg, cat_spec_poly, K, rep, heat_is_logit, return_inds=False, post_affine=None):
    if cat_spec_poly:
        # dead/buggy branch of the reference (decode.py:514,534-537 uses W as the vertex count)
        raise NotImplementedError("cat_spec_poly=True is not supported (it is broken in the reference as well)")
    dev = _lib.require_cuda(heat, polys, depth, reg)
    if heat.dim() != 4 or polys.dim() != 4 or depth.dim() != 4:
        raise RuntimeError("polydet_decode: heat, polys and depth must be 4-d (B,C,H,W)")
    B, C, H, W = heat.shape
    n_poly = polys.shape[1]
    if polys.shape[0] != B or tuple(polys.shape[2:]) != (H, W) or tuple(depth.shape) != (B, 1, H, W):
        raise RuntimeError("polydet_decode: head tensors disagree on (B,H,W)")
    if reg is not None and tuple(reg.shape) != (B, 2, H, W):
        raise RuntimeError("polydet_decode: reg must be (B,2,H,W)")
    K = int(K)
    if K > H * W:
        # the reference's per-class torch.topk raises for k > H*W (decode.py:120)
        raise RuntimeError(f"polydet_decode: K={K} is out of range for a {H}x{W} map")
    tensors = [heat, polys, depth] + ([reg] if reg is not None else [])
    if any(t.dtype != heat.dtype for t in tensors) or heat.dtype not in (torch.float32, torch.bfloat16):
        heat, polys, depth = heat.float(), polys.float(), depth.float()
        reg = reg.float() if reg is not None else None
    L = _lib.lib()
    with torch.cuda.device(dev):
        out = torch.empty((B, K, n_poly + 7), dtype=torch.float32, device=dev)
        inds = torch.empty((B, K), dtype=torch.int64, device=dev) if return_inds else None
        ws = _lib.workspace(L.cpoly_decode_workspace_bytes(B, C, H, W, K), dev)
        a = _lib.DecodeArgs()
        a.B, a.C, a.H, a.W, a.n_poly, a.K = B, C, H, W, n_poly, K
        a.rep = _lib.REP.get(rep, 0)  # any other string (e.g. 'polar_confined') decodes as cartesian, like the reference
        a.heat_is_logit = int(heat_is_logit)
        a.dtype = _lib.dtype_code(heat)
        a.heat, a.polys, a.depth, a.reg = _lib.t4(heat), _lib.t4(polys), _lib.t4(depth), _lib.t4(reg)
        a.out = out.data_ptr()
        if post_affine is not None:
            # (B,2,3) inverse affine of polydet_post_process; fp64 on the device, applied by the gather kernel
            if not (torch.is_tensor(post_affine) and post_affine.is_cuda):
                post_affine = torch.as_tensor(post_affine, dtype=torch.float64).to(dev, non_blocking=True)
            post_affine = post_affine.to(torch.float64).contiguous()
            if post_affine.numel() != B * 6:
                raise RuntimeError(f"polydet_decode: post_affine must hold {B} 2x3 matrices")
            a.post_affine = post_affine.data_ptr()
        a.out_inds = inds.data_ptr() if inds is not None else None
        a.workspace, a.workspace_bytes = ws.data_ptr(), ws.numel()
        _lib.check(L.cpoly_decode(byref(a), _lib.stream_ptr(dev)), "cpoly_decode")
        _lib.LAUNCHES["decode"] += 5   # histogram, threshold, collect, select, gather (csrc/decode.cu)
    return (out, inds) if return_inds else out


def polydet_decode(heat, polys, depth, reg=None, cat_spec_poly=False, K=100, rep='cartesian', post_affine=None):
    """(B, K, 2N+7) float32: [xmin, ymin, xmax, ymax, score, class, x0, y0, ..., pseudo_depth].

    ``post_affine`` (optional, (B,2,3), see ``centerpoly_b200.post_process.batch_inverse_affine``) additionally fuses the
    coordinate back-mapping of ``polydet_post_process`` (src/lib/utils/post_process.py:105-112) into the gather kernel."""
    return _decode(heat, polys, depth, reg, cat_spec_poly, K, rep, heat_is_logit=False, post_affine=post_affine)


def polydet_decode_from_logits(hm_logits, polys, depth, reg=None, cat_spec_poly=False, K=100, rep='cartesian', post_affine=None):
    """Same, but takes the raw ``hm`` head and fuses the ``sigmoid_`` of detectors/polydet.py:28."""
    return _decode(hm_logits, polys, depth, reg, cat_spec_poly, K, rep, heat_is_logit=True, post_affine=post_affine)
