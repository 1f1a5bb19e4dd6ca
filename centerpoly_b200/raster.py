"""Depth-ordered instance rasterisation on the GPU: the mask construction inside the reference's
``Cityscapes.format_and_write_to_cityscapes`` (src/lib/datasets/dataset/cityscapes.py:226-279), i.e. everything that
function does between reading ``all_bboxes`` and writing the PNG / txt files, backed by ``cpoly_rasterise``
(centerpoly_b200/csrc/raster.cu).  Input is the (B,K,2N+7) tensor ``polydet_decode(..., post_affine=...)`` leaves on
the device, so decode -> back-mapping -> masks never visits the host.  No CPU path, no PyTorch fallback.
"""
from __future__ import annotations

from ctypes import byref
from dataclasses import dataclass

import torch

from . import _lib

__all__ = ["rasterise_instances", "RasterResult", "CITYSCAPES_CLASSES", "NOT_DRAWN"]

# src/lib/datasets/dataset/cityscapes.py:88-89 without the leading '__background__' (classes here are 0-based)
CITYSCAPES_CLASSES = ("person", "rider", "car", "truck", "bus", "train", "motorcycle", "bicycle", "pole", "traffic sign", "traffic light")
NOT_DRAWN = ("pole", "traffic sign", "traffic light")   # cityscapes.py:242,273


@dataclass
class RasterResult:
    """Slot s of image b is the s-th nearest instance with score > thresh.  All tensors live on the device."""
    masks: torch.Tensor | None   # (B,K,H,W) uint8 0/255; slots >= count[b] are not written
    order: torch.Tensor          # (B,K) int32 detection index per slot, -1 past count
    count: torch.Tensor          # (B) int32
    verts: torch.Tensor          # (B,K,N,2) int16 truncated vertices
    area: torch.Tensor           # (B,K) int32 visible pixels
    keep: torch.Tensor           # (B,K) bool: the reference would write this instance (drawn class, area > 100)
    conf: torch.Tensor           # (B,K) float64 min(1, 1.2*score)
    status: torch.Tensor         # (B) int32 clamped vertices (0 => bit-exact to the reference)

    def instances(self, dets: torch.Tensor, b: int = 0, class_names=CITYSCAPES_CLASSES):
        """Host view of image ``b`` in the order the reference writes it: [(mask HxW uint8, label, confidence)]."""
        n = int(self.count[b])
        keep = self.keep[b, :n].cpu().numpy()
        order = self.order[b, :n].cpu().numpy()
        conf = self.conf[b, :n].cpu().numpy()
        cls = dets[b, :, 5].cpu().numpy().astype(int)
        out = []
        for s in range(n):
            if keep[s]:
                out.append((self.masks[b, s].cpu().numpy(), class_names[cls[order[s]]], float(conf[s])))
        return out


def rasterise_instances(dets: torch.Tensor, width: int = 2048, height: int = 1024, thresh: float = 0.05, *, class_names=CITYSCAPES_CLASSES,
                        occluder_score: float = 0.5, min_area: int = 100, want_masks: bool = True, out: torch.Tensor | None = None) -> RasterResult:
    """``dets``: (B,K,2N+7) float32 on the GPU, rows [x1,y1,x2,y2,score,class,x0,y0,...,depth] in ORIGINAL-image
    coordinates (``polydet_decode(..., post_affine=batch_inverse_affine(...))``).  ``thresh`` is ``opt.thresh``."""
    dev = _lib.require_cuda(dets)
    if dets.dim() != 3 or dets.dtype != torch.float32 or (dets.shape[2] - 7) % 2 or dets.shape[2] < 13:
        raise RuntimeError("rasterise_instances: dets must be float32 (B,K,2N+7)")
    dets = dets.contiguous()
    B, K, row = dets.shape
    N = (row - 7) // 2
    skip = 0
    for c, name in enumerate(class_names):
        if name in NOT_DRAWN:
            skip |= 1 << c
    L = _lib.lib()
    with torch.cuda.device(dev):
        masks = None
        if want_masks:
            masks = out if out is not None else torch.empty((B, K, height, width), dtype=torch.uint8, device=dev)
            if tuple(masks.shape) != (B, K, height, width) or masks.dtype != torch.uint8 or not masks.is_contiguous():
                raise RuntimeError("rasterise_instances: out must be a contiguous uint8 (B,K,H,W) tensor")
        r = RasterResult(masks=masks,
                         order=torch.empty((B, K), dtype=torch.int32, device=dev), count=torch.empty((B,), dtype=torch.int32, device=dev),
                         verts=torch.empty((B, K, N, 2), dtype=torch.int16, device=dev), area=torch.empty((B, K), dtype=torch.int32, device=dev),
                         keep=torch.empty((B, K), dtype=torch.bool, device=dev), conf=torch.empty((B, K), dtype=torch.float64, device=dev),
                         status=torch.empty((B,), dtype=torch.int32, device=dev))
        ws = _lib.workspace(L.cpoly_rasterise_workspace_bytes(B, K), dev)
        a = _lib.RasterArgs()
        a.B, a.K, a.N, a.W, a.H = B, K, N, int(width), int(height)
        a.thresh, a.occluder_score, a.min_area, a.skip_classes = float(thresh), float(occluder_score), int(min_area), skip
        a.dets = dets.data_ptr()
        a.order, a.count, a.verts, a.area = r.order.data_ptr(), r.count.data_ptr(), r.verts.data_ptr(), r.area.data_ptr()
        a.keep, a.conf, a.status = r.keep.data_ptr(), r.conf.data_ptr(), r.status.data_ptr()
        a.masks = masks.data_ptr() if masks is not None else None
        a.workspace, a.workspace_bytes = ws.data_ptr(), ws.numel()
        _lib.check(L.cpoly_rasterise(byref(a), _lib.stream_ptr(dev)), "cpoly_rasterise")
        _lib.LAUNCHES["rasterise"] += 3   # prep, cover, keep (csrc/raster.cu)
    return r
