"""Generates tests/golden/polyiou_*.npz by running the UNMODIFIED reference ``PolyLoss`` (poly_loss='iou',
rep='polar'; /root/reference/src/lib/models/losses.py:833-959) in this container.

    python tests/golden/make_polyiou_golden.py

matplotlib / seaborn / progress are absent here and only imported at module top of losses.py (:18-19), so they are
stubbed with empty modules; nothing of the reference is modified.  Stored per case: the gathered prediction rows
``pred`` (B,M,2N) (the head map holds them at ``ind``), ``mask``, ``target``, the reference's scalar ``loss`` and
``grad`` = d loss / d pred read back from the head map's gradient.
"""
import os
import sys
import types

import numpy as np
import torch

for _n in ["matplotlib", "matplotlib.pyplot", "seaborn", "progress", "progress.bar"]:
    sys.modules.setdefault(_n, types.ModuleType(_n))
sys.modules["progress.bar"].Bar = object
sys.path.insert(0, "/root/reference/src/lib")
from models.losses import PolyLoss  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)


class _Opt:
    poly_loss = "iou"
    poly_order = False
    rep = "polar"


from make_polyiou_cases import star_case  # noqa: E402


def run_reference(pred, mask, ind, tgt, hw):
    B, M, C = pred.shape
    H, W = hw
    out = torch.zeros(B, C, H * W)
    out.scatter_(2, ind.unsqueeze(1).expand(B, C, M), pred.transpose(1, 2))
    out = out.view(B, C, H, W).requires_grad_(True)
    loss = PolyLoss(_Opt())(out, mask, ind, tgt)
    loss.backward()
    grad = out.grad.view(B, C, H * W).gather(2, ind.unsqueeze(1).expand(B, C, M)).transpose(1, 2)
    return float(loss), grad.contiguous().numpy()


CASES = {
    # name: (seed, B, M, N, n_obj, r_noise, t_noise, shift, shuffle, neg_r)
    "polyiou_n8": (1, 2, 6, 8, 4, 1.5, 0.08, 0.0, True, False),
    "polyiou_n16_negr": (2, 1, 8, 16, 6, 1.0, 0.04, 0.0, True, True),
    "polyiou_n32": (3, 1, 6, 32, 5, 0.8, 0.02, 0.0, True, False),
    "polyiou_n8_contained": (4, 1, 4, 8, 3, 0.0, 0.0, -0.5, False, False),   # no crossings -> min(area) branch
    "polyiou_n16_noisy": (5, 1, 8, 16, 8, 3.0, 0.15, 0.0, True, False),
}

if __name__ == "__main__":
    for name, (seed, B, M, N, n_obj, rn, tn, shift, shuffle, neg_r) in CASES.items():
        pred, mask, ind, tgt, hw = star_case(seed, B, M, N, n_obj, rn, tn, shift, shuffle, neg_r)
        try:
            loss, grad = run_reference(pred, mask, ind, tgt, hw)
        except Exception as e:  # the reference itself fails on some degenerate walks
            print(name, "REFERENCE RAISED", type(e).__name__, e)
            continue
        np.savez_compressed(os.path.join(HERE, name + ".npz"), pred=pred.numpy(), mask=mask.numpy(),
                            target=tgt.numpy(), loss=np.float32(loss), grad=grad)
        print(name, "loss", loss, "|grad|max", np.abs(grad).max())
