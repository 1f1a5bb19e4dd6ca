"""Generates tests/golden/polyorder_*.npz by running the UNMODIFIED reference ``PolyLoss`` with ``opt.poly_order = True``
(/root/reference/src/lib/models/losses.py:833-959; the order arm is :891-906,908 and its in-place ``angles[j] += 2*3.14``
on a view of the gathered prediction is seen by the regression arm that follows) in this container.

    python tests/golden/make_polyorder_golden.py

poly_loss='l1' so that the (separately pinned, slow) IoU arm stays out; rep 'polar' (angles enter the 1 - cos term) and
'polar_fixed'.  Stored: pred rows, mask, target, the reference's ``loss``, ``loss_order`` and d(loss + loss_order)/d pred.
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
from make_polyiou_cases import star_case  # noqa: E402


def order_case(seed, B, M, N, n_obj, t_noise):
    """Star case whose predicted angles are wrapped to (-pi, pi] (what a regression head produces around the branch
    cut) and partly shuffled, so that the seen-a-positive-angle rule and the pair sum are both exercised."""
    pred, mask, ind, tgt, hw = star_case(seed, B, M, N, n_obj, 1.0, t_noise, 0.0, False, False)
    g = torch.Generator().manual_seed(1000 + seed)
    a = pred[..., 1::2]
    a = torch.where(a > np.pi, a - 2 * np.pi, a)
    swap = torch.rand(a.shape, generator=g) < 0.2
    a = torch.where(swap, a.roll(1, dims=-1), a)
    pred[..., 1::2] = a
    return pred, mask, ind, tgt, hw


def run_reference(pred, mask, ind, tgt, hw, rep):
    class _Opt:
        poly_loss, poly_order = "l1", True
    _Opt.rep = rep
    B, M, C = pred.shape
    H, W = hw
    out = torch.zeros(B, C, H * W)
    out.scatter_(2, ind.unsqueeze(1).expand(B, C, M), pred.transpose(1, 2))
    out = out.view(B, C, H, W).requires_grad_(True)
    loss, order = PolyLoss(_Opt())(out, mask, ind, tgt)
    (loss + order).backward()
    grad = out.grad.view(B, C, H * W).gather(2, ind.unsqueeze(1).expand(B, C, M)).transpose(1, 2)
    return float(loss), float(order), grad.contiguous().numpy()


CASES = {"polyorder_n8_polar": (11, 2, 6, 8, 4, 0.3, "polar"), "polyorder_n16_polar": (12, 1, 8, 16, 6, 0.2, "polar"),
         "polyorder_n32_fixed": (13, 1, 6, 32, 5, 0.1, "polar_fixed")}

if __name__ == "__main__":
    for name, (seed, B, M, N, n_obj, tn, rep) in CASES.items():
        pred, mask, ind, tgt, hw = order_case(seed, B, M, N, n_obj, tn)
        loss, order, grad = run_reference(pred.clone(), mask, ind, tgt, hw, rep)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), pred=pred.numpy(), mask=mask.numpy(), ind=ind.numpy(),
                            target=tgt.numpy(), hw=np.array(hw), rep=np.array(rep), loss=np.float32(loss),
                            loss_order=np.float32(order), grad=grad)
        print(name, "loss", loss, "order", order, "|grad|max", np.abs(grad).max())
