"""Back-mapping of the decoded polygons to the original image: the reference's ``polydet_post_process``
(src/lib/utils/post_process.py:105-122, called from src/lib/detectors/polydet.py:45-60).

The reference walks every detection in Python: (N + 2) ``affine_transform`` calls (a 2x3 np.dot each) per detection,
3400 per 100-detection image.  Here the 2x3 inverse affine of every image is built once on the host
(``inverse_affine`` = ``get_affine_transform(c, s, 0, (out_w, out_h), inv=1)``, src/lib/utils/image.py:27-60) and handed
to the decode's gather kernel (``polydet_decode(..., post_affine=...)``, csrc/decode.cu), which writes the box corners
and vertices in image coordinates directly; ``group_by_class`` then only formats the result the way the detector
expects (``{class_id: rows}``, 1-based, src/lib/utils/post_process.py:113-120).
"""
from __future__ import annotations

import numpy as np

__all__ = ["inverse_affine", "batch_inverse_affine", "group_by_class", "polydet_post_process"]


def inverse_affine(c, s, out_w, out_h) -> np.ndarray:
    """(2,3) float64: feature-map (out_w x out_h) coordinates -> original image, for centre ``c`` and scale ``s``
    (scalar = fix_res convention, or [w, h]; src/lib/detectors/base_detector.py:55-63).  Same construction as
    ``get_affine_transform(..., rot=0, inv=1)``: three point pairs in float32, exact 3-point solve in float64
    (what ``cv2.getAffineTransform`` does)."""
    c = np.asarray(c, dtype=np.float32)
    if np.ndim(s) == 0:
        s = np.array([s, s], dtype=np.float32)
    s = np.asarray(s, dtype=np.float32)
    src_dir = np.array([0.0, float(s[0]) * -0.5])            # get_dir([0, src_w * -0.5], 0)
    dst_dir = np.array([0, out_w * -0.5], np.float32)
    src = np.zeros((3, 2), dtype=np.float32)
    dst = np.zeros((3, 2), dtype=np.float32)
    src[0] = c
    src[1] = c + src_dir
    dst[0] = [out_w * 0.5, out_h * 0.5]
    dst[1] = np.array([out_w * 0.5, out_h * 0.5], np.float32) + dst_dir
    for pts in (src, dst):                                   # get_3rd_point
        d = pts[0] - pts[1]
        pts[2] = pts[1] + np.array([-d[1], d[0]], dtype=np.float32)
    a = np.zeros((6, 6), dtype=np.float64)
    b = np.zeros(6, dtype=np.float64)
    for i in range(3):                                       # cv2.getAffineTransform(dst, src)
        a[2 * i, 0:3] = [dst[i, 0], dst[i, 1], 1.0]
        a[2 * i + 1, 3:6] = [dst[i, 0], dst[i, 1], 1.0]
        b[2 * i], b[2 * i + 1] = src[i, 0], src[i, 1]
    return np.linalg.solve(a, b).reshape(2, 3)


def batch_inverse_affine(c, s, out_w, out_h) -> np.ndarray:
    """(B,2,3) float64 for per-image lists ``c`` / ``s`` (the detector passes ``[meta['c']]``, ``[meta['s']]``)."""
    return np.stack([inverse_affine(ci, si, out_w, out_h) for ci, si in zip(c, s)])


def group_by_class(dets: np.ndarray, num_classes: int):
    """``dets`` (B,K,2N+7) already in image coordinates -> ``[{1: rows, ..., num_classes: rows}]`` with
    rows = [x1, y1, x2, y2, score, x0, y0, ..., depth] (the class column dropped), post_process.py:113-120."""
    ret = []
    for i in range(dets.shape[0]):
        classes = dets[i, :, 5]
        body = np.concatenate([dets[i, :, :5], dets[i, :, 6:]], axis=1).astype(np.float32)
        ret.append({j + 1: body[classes == j].tolist() for j in range(num_classes)})
    return ret


def polydet_post_process(dets, c, s, h, w, num_classes):
    """Signature of the reference function (src/lib/utils/post_process.py:105) for callers that already hold the
    feature-map-coordinate detections on the host (src/lib/detectors/polydet.py:46-50): one vectorised affine per image
    instead of (N + 2) * K Python-level np.dot calls.  The B200 path proper is ``polydet_decode(post_affine=...)``."""
    dets = np.array(dets, copy=True)
    trans = batch_inverse_affine(c, s, w, h)
    for i in range(dets.shape[0]):
        t = trans[i]
        for lo, hi in ((0, 4), (6, dets.shape[-1] - 1)):
            pts = dets[i, :, lo:hi].reshape(dets.shape[1], -1, 2).astype(np.float32).astype(np.float64)
            x, y = pts[..., 0], pts[..., 1]
            out = np.stack([t[0, 0] * x + t[0, 1] * y + t[0, 2], t[1, 0] * x + t[1, 1] * y + t[1, 2]], axis=-1)
            dets[i, :, lo:hi] = out.reshape(dets.shape[1], -1)
    return group_by_class(dets, num_classes)
