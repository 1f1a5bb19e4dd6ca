"""Seeded star-polygon cases for the polygon-IoU loss tests (no reference import; usable on the GPU box)."""
import numpy as np
import torch


def star_case(seed, B, M, N, n_obj, r_noise, t_noise, shift=0.0, shuffle=True, neg_r=False):
    """Targets: star polygons with increasing angles (what sample/polydet.py:250-263 encodes); predictions: the
    target plus noise, rows optionally shuffled (the loss sorts by angle) and some radii negated (it takes abs)."""
    g = torch.Generator().manual_seed(seed)
    H, W = 16, 24
    pred = torch.zeros(B, M, 2 * N)
    tgt = torch.zeros(B, M, 2 * N)
    mask = torch.zeros(B, M, dtype=torch.uint8)
    ind = torch.zeros(B, M, dtype=torch.int64)
    for b in range(B):
        cells = torch.randperm(H * W, generator=g)[:M]
        ind[b] = cells
        for k in range(n_obj):
            ang = torch.linspace(0, 2 * np.pi, N + 1)[:-1] + 0.05 + 0.5 * float(torch.rand(1, generator=g)) / N
            r = 4 + 10 * torch.rand(N, generator=g)
            tgt[b, k, 0::2], tgt[b, k, 1::2] = r, ang
            rp = r * (1 + shift) + torch.randn(N, generator=g) * r_noise
            tp = ang + torch.randn(N, generator=g) * t_noise
            if neg_r:
                rp = rp * torch.where(torch.rand(N, generator=g) < 0.3, -1.0, 1.0)
            perm = torch.randperm(N, generator=g) if shuffle else torch.arange(N)
            pred[b, k, 0::2], pred[b, k, 1::2] = rp[perm], tp[perm]
            mask[b, k] = 1
    return pred, mask, ind, tgt, (H, W)
