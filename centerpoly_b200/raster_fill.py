"""IDD / KITTI instance masks on the GPU: the mask construction inside the reference's
``IDD.format_and_write_to_IDD`` (src/lib/datasets/dataset/IDD.py:123-170) and ``KITTIPOLY.format_and_write_to_kitti``
(src/lib/datasets/dataset/kitti_poly.py:95-136), backed by ``cpoly_rasterise_fill`` (csrc/raster_fill.cu), plus host
writers with the reference's file layout (``<frame>.txt`` lines ``<mask png> <label id> <score>`` and one 8-bit PNG per
instance).  Input is the (B,K,2N+7) tensor ``polydet_decode(..., post_affine=...)`` leaves on the device.  No CPU path.
"""
from __future__ import annotations

import os
import struct
import zlib
from ctypes import byref
from dataclasses import dataclass

import torch

from . import _lib

__all__ = ["rasterise_fill_instances", "FillRasterResult", "write_idd", "write_kitti", "IDD_CLASSES", "IDD_LABEL_IDS",
           "KITTI_CLASSES", "KITTI_LABEL_IDS"]

# IDD.py:41-43 / kitti_poly.py:40-43 without the leading '__background__' (classes here are 0-based)
IDD_CLASSES = ("person", "rider", "motorcycle", "bicycle", "autorickshaw", "car", "truck", "bus", "vehicle fallback")
IDD_LABEL_IDS = {"person": 6, "rider": 8, "motorcycle": 9, "bicycle": 10, "autorickshaw": 11, "car": 12, "truck": 13, "bus": 14,
                 "vehicle fallback": 18}
KITTI_CLASSES = ("person", "rider", "car", "truck", "bus", "train", "motorcycle", "bicycle")
KITTI_LABEL_IDS = {"person": 24, "rider": 25, "car": 26, "truck": 27, "bus": 28, "train": 31, "motorcycle": 32, "bicycle": 33}


@dataclass
class FillRasterResult:
    """Slot s of image b is the s-th mask file the reference writes (classes ascending, detection order inside a class)."""
    masks: torch.Tensor | None   # (B,K,H,W) uint8 0/255; slots >= count[b] are not written
    count: torch.Tensor          # (B) int32
    slot_det: torch.Tensor       # (B,K) int32 detection index per slot, -1 past count
    area: torch.Tensor           # (B,K) int32 visible pixels
    status: torch.Tensor         # (B) int32 clamped vertices (0 => exact)


def rasterise_fill_instances(dets: torch.Tensor, width: int, height: int, thresh: float, *, inclusive: bool,
                             occluder_score: float = 0.5, want_masks: bool = True, out: torch.Tensor | None = None) -> FillRasterResult:
    """``dets``: (B,K,2N+7) float32 on the GPU, rows [x1,y1,x2,y2,score,class,x0,y0,...,depth] in ORIGINAL-image coordinates.
    ``inclusive``: True = IDD's ``score >= thresh``, False = KITTI's ``score > thresh``."""
    dev = _lib.require_cuda(dets)
    if dets.dim() != 3 or dets.dtype != torch.float32 or (dets.shape[2] - 7) % 2 or dets.shape[2] < 13:
        raise RuntimeError("rasterise_fill_instances: dets must be float32 (B,K,2N+7)")
    dets = dets.contiguous()
    B, K, row = dets.shape
    N = (row - 7) // 2
    L = _lib.lib()
    with torch.cuda.device(dev):
        masks = None
        if want_masks:
            masks = out if out is not None else torch.empty((B, K, height, width), dtype=torch.uint8, device=dev)
            if tuple(masks.shape) != (B, K, height, width) or masks.dtype != torch.uint8 or not masks.is_contiguous():
                raise RuntimeError("rasterise_fill_instances: out must be a contiguous uint8 (B,K,H,W) tensor")
        r = FillRasterResult(masks=masks, count=torch.empty((B,), dtype=torch.int32, device=dev),
                             slot_det=torch.empty((B, K), dtype=torch.int32, device=dev),
                             area=torch.empty((B, K), dtype=torch.int32, device=dev), status=torch.empty((B,), dtype=torch.int32, device=dev))
        ws = _lib.workspace(L.cpoly_rasterise_fill_workspace_bytes(B, K, N), dev)
        a = _lib.RasterFillArgs()
        a.B, a.K, a.N, a.W, a.H, a.inclusive = B, K, N, int(width), int(height), int(bool(inclusive))
        a.thresh, a.occluder_score = float(thresh), float(occluder_score)
        a.dets = dets.data_ptr()
        a.count, a.slot_det, a.area, a.status = r.count.data_ptr(), r.slot_det.data_ptr(), r.area.data_ptr(), r.status.data_ptr()
        a.masks = masks.data_ptr() if masks is not None else None
        a.workspace, a.workspace_bytes = ws.data_ptr(), ws.numel()
        _lib.check(L.cpoly_rasterise_fill(byref(a), _lib.stream_ptr(dev)), "cpoly_rasterise_fill")
        _lib.LAUNCHES["rasterise_fill"] = _lib.LAUNCHES.get("rasterise_fill", 0) + (5 if masks is not None else 1)
    return r


def _png_gray8(path: str, img) -> None:
    """Minimal 8-bit greyscale PNG (zlib from the standard library): what `Image.fromarray(uint8).save(path)` stores."""
    h, w = img.shape
    raw = b"".join(b"\x00" + img[y].tobytes() for y in range(h))

    def chunk(tag, data):
        return struct.pack(">I", len(data)) + tag + data + struct.pack(">I", zlib.crc32(tag + data) & 0xffffffff)
    with open(path, "wb") as f:
        f.write(b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 0, 0, 0, 0)) +
                chunk(b"IDAT", zlib.compress(raw, 6)) + chunk(b"IEND", b""))


def _write(result: FillRasterResult, dets: torch.Tensor, b: int, frame_name: str, out_dir: str, class_names, label_ids):
    n = int(result.count[b])
    slot_det = result.slot_det[b, :n].cpu().numpy()
    rows = dets[b].cpu().numpy()
    os.makedirs(out_dir, exist_ok=True)
    stem = os.path.basename(frame_name)
    lines = []
    for s in range(n):
        d = int(slot_det[s])
        mask_name = stem.replace(".png", "_" + str(s) + ".png")
        # `str(bbox[4])` of a Python float that came out of a float32 array through .tolist(): the shortest repr of the double
        lines.append(mask_name + " " + str(label_ids[class_names[int(rows[d, 5])]]) + " " + str(float(rows[d, 4])) + "\n")
        _png_gray8(os.path.join(out_dir, mask_name), result.masks[b, s].cpu().numpy())
    with open(os.path.join(out_dir, stem.replace(".png", ".txt")), "w") as f:
        f.writelines(lines)
    return n


def write_idd(result: FillRasterResult, dets: torch.Tensor, b: int, image_name: str, save_dir: str, class_names=IDD_CLASSES,
              label_ids=IDD_LABEL_IDS) -> int:
    """Files of `format_and_write_to_IDD` for image ``b``: `<save_dir>/<parent dir of the image>/<frame>.txt` + one PNG per
    instance (IDD.py:131-168).  Returns the number of instances written."""
    sub_dir = image_name.split("/")[-2]
    return _write(result, dets, b, image_name, os.path.join(save_dir, sub_dir), class_names, label_ids)


def write_kitti(result: FillRasterResult, dets: torch.Tensor, b: int, image_name: str, save_dir: str, class_names=KITTI_CLASSES,
                label_ids=KITTI_LABEL_IDS) -> int:
    """Files of `format_and_write_to_kitti` for image ``b``: `<save_dir>/<frame>.txt` + one PNG per instance (kitti_poly.py:101-136)."""
    return _write(result, dets, b, image_name, save_dir, class_names, label_ids)
