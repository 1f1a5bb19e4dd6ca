"""TEST INFRASTRUCTURE — NOT PRODUCT CODE.

numpy restatement of the reference's polydet decode.  Only tests/, ``__graft_entry__.smoke()``
and bench.py's CPU-baseline legs may import this module; the product path never does.

Follows (file:line under /root/reference):
  * ``_nms``                       src/lib/models/decode.py:13-19
  * ``_topk``                      src/lib/models/decode.py:117-133
  * ``_gather_feat`` / ``_transpose_and_gather_feat``   src/lib/models/utils.py:12-26
  * ``polydet_decode``             src/lib/models/decode.py:512-670
  * the sigmoid applied by the caller   src/lib/detectors/polydet.py:28

Pinned: tests/golden/decode_*.npz hold outputs of the UNMODIFIED reference ``polydet_decode``
(imported from /root/reference in the build container by tests/golden/make_golden.py);
tests/test_oracle_decode.py checks this restatement against them bit-for-bit.

Tie rule.  ``torch.topk`` leaves the order of equal scores unspecified; this oracle (and the CUDA
kernel) order equal scores by ascending flat index ``c*H*W + y*W + x``.  On tie-free inputs that
is indistinguishable from the reference.
"""
from __future__ import annotations

import math

import numpy as np


def sigmoid_f32(x: np.ndarray) -> np.ndarray:
    """float32 logistic, the oracle of ``hm.sigmoid_()`` (detectors/polydet.py:28)."""
    x = np.asarray(x, dtype=np.float32)
    return (np.float32(1.0) / (np.float32(1.0) + np.exp(-x, dtype=np.float32))).astype(np.float32)


def nms3x3(heat: np.ndarray) -> np.ndarray:
    """decode.py:13-19 — keep a pixel iff it equals the max of its 3x3 window (-inf padding)."""
    heat = np.asarray(heat, dtype=np.float32)
    B, C, H, W = heat.shape
    padded = np.full((B, C, H + 2, W + 2), -np.inf, dtype=np.float32)
    padded[:, :, 1:-1, 1:-1] = heat
    hmax = np.full_like(heat, -np.inf)
    for dy in range(3):
        for dx in range(3):
            hmax = np.maximum(hmax, padded[:, :, dy:dy + H, dx:dx + W])
    keep = (hmax == heat).astype(np.float32)
    return heat * keep


def _topk_desc(values: np.ndarray, K: int):
    """Per-row top-K, score descending, ties by ascending position (stable)."""
    order = np.argsort(-values, axis=-1, kind="stable")[..., :K]
    return np.take_along_axis(values, order, axis=-1), order


def topk_two_stage(scores: np.ndarray, K: int):
    """decode.py:117-133 — per-class top-K, then top-K over the C*K survivors."""
    B, C, H, W = scores.shape
    s1, i1 = _topk_desc(scores.reshape(B, C, H * W), K)          # (B, C, K)
    s2, i2 = _topk_desc(s1.reshape(B, C * K), K)                 # (B, K) positions in the C*K list
    clses = (i2 // K).astype(np.int32)
    inds = np.take_along_axis(i1.reshape(B, C * K), i2, axis=1).astype(np.int64)
    ys = (inds // W).astype(np.float32)
    xs = (inds % W).astype(np.float32)
    return s2, inds, clses, ys, xs


def _gather(feat: np.ndarray, inds: np.ndarray) -> np.ndarray:
    """utils.py:22-26 — feat (B,D,H,W), inds (B,K) flat y*W+x  ->  (B,K,D)."""
    B, D, H, W = feat.shape
    flat = feat.reshape(B, D, H * W)
    return np.stack([flat[b][:, inds[b]].T for b in range(B)], axis=0)


def polydet_decode_oracle(heat, polys, depth, reg=None, cat_spec_poly=False, K=100, rep="cartesian"):
    """decode.py:512-670.  ``heat`` is POST-sigmoid.  Returns (B, K, 2N+7) float32:
    [xmin, ymin, xmax, ymax, score, class, x0, y0, ..., x_{N-1}, y_{N-1}, pseudo_depth]."""
    if cat_spec_poly:
        raise NotImplementedError("cat_spec_poly=True is a dead branch in the reference (decode.py:514,534-537)")
    heat = np.asarray(heat, dtype=np.float32)
    polys = np.asarray(polys, dtype=np.float32)
    depth = np.asarray(depth, dtype=np.float32)
    B, C, H, W = heat.shape
    scores, inds, clses, ys, xs = topk_two_stage(nms3x3(heat), K)
    if reg is not None:
        r = _gather(np.asarray(reg, dtype=np.float32), inds)      # decode.py:523-527
        xs = xs[..., None] + r[..., 0:1]
        ys = ys[..., None] + r[..., 1:2]
    else:
        xs = xs[..., None] + np.float32(0.5)                      # decode.py:529-530
        ys = ys[..., None] + np.float32(0.5)
    p = _gather(polys, inds).astype(np.float32)                   # (B, K, 2N)
    d = _gather(depth, inds).astype(np.float32)                   # (B, K, 1)
    n2 = p.shape[-1]
    if rep in ("polar", "polar_fixed"):                           # decode.py:582-614
        # the reference evaluates cos/sin in Python doubles, then multiplies the fp32 radius by the
        # fp32-rounded result (tensor * python-float stays float32).
        rad = p[..., 0:n2 - 1:2].copy()
        if rep == "polar_fixed":
            j = np.arange(0, n2 - 1, 2, dtype=np.float64)
            ang = np.broadcast_to(2 * 3.14 - 2 * 3.14 / n2 * j, rad.shape)   # literal 3.14 as in the reference
        else:
            ang = p[..., 1:n2:2].astype(np.float64)
        cosv = np.vectorize(math.cos, otypes=[np.float64])(ang).astype(np.float32)
        sinv = np.vectorize(math.sin, otypes=[np.float64])(ang).astype(np.float32)
        p[..., 0:n2 - 1:2] = rad * cosv
        p[..., 1:n2:2] = rad * sinv
    p[..., 0::2] += xs                                            # decode.py:624-625
    p[..., 1::2] += ys
    px, py = p[..., 0::2], p[..., 1::2]
    bbox = np.stack([px.min(-1), py.min(-1), px.max(-1), py.max(-1)], axis=-1)   # decode.py:647-659
    return np.concatenate([bbox, scores[..., None].astype(np.float32),
                           clses[..., None].astype(np.float32), p, d], axis=-1).astype(np.float32)
