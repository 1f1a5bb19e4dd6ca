"""The order arm of PolyLoss (centerpoly_b200/poly_loss.py::poly_order_loss, PolyLoss with opt.poly_order) against
outputs of the UNMODIFIED reference PolyLoss (tests/golden/polyorder_*.npz, made by make_polyorder_golden.py):
loss, loss_order and the gradient of their sum, including the reference's in-place +2*3.14 shift that the
regression arm sees.  Pure tensor code, so the same test runs on the CPU here and on cuda:0 on the GPU box."""
import glob
import os

import numpy as np
import pytest
import torch

GOLD = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "polyorder_*.npz")))


def _run(z, device):
    from centerpoly_b200.poly_loss import PolyLoss

    class _Opt:
        poly_loss, poly_order = "l1", True
    _Opt.rep = str(z["rep"])
    pred, mask, ind, tgt = (torch.from_numpy(z[k]).to(device) for k in ("pred", "mask", "ind", "target"))
    H, W = (int(v) for v in z["hw"])
    B, M, C = pred.shape
    out = torch.zeros(B, C, H * W, device=device)
    out.scatter_(2, ind.unsqueeze(1).expand(B, C, M), pred.transpose(1, 2))
    out = out.view(B, C, H, W).requires_grad_(True)
    loss, order = PolyLoss(_Opt())(out, mask, ind, tgt)
    (loss + order).backward()
    grad = out.grad.view(B, C, H * W).gather(2, ind.unsqueeze(1).expand(B, C, M)).transpose(1, 2)
    return float(loss), float(order), grad.cpu().numpy()


def _check(z, got):
    loss, order, grad = got
    assert abs(loss - float(z["loss"])) <= 2e-6 * max(1.0, abs(float(z["loss"])))
    assert abs(order - float(z["loss_order"])) <= 1e-5 * max(1e-3, abs(float(z["loss_order"])))
    assert np.abs(grad - z["grad"]).max() <= 1e-5 * np.abs(z["grad"]).max()


@pytest.mark.parametrize("path", GOLD, ids=[os.path.basename(p)[:-4] for p in GOLD])
def test_order_loss_matches_reference_cpu(path):
    z = np.load(path)
    _check(z, _run(z, "cpu"))


@pytest.mark.gpu
@pytest.mark.parametrize("path", GOLD, ids=[os.path.basename(p)[:-4] for p in GOLD])
def test_order_loss_matches_reference_gpu(path):
    z = np.load(path)
    _check(z, _run(z, "cuda"))


def test_goldens_present_and_shift_is_exercised():
    assert len(GOLD) >= 3
    from centerpoly_b200.poly_loss import poly_order_loss
    z = np.load(GOLD[0])
    pred, mask = torch.from_numpy(z["pred"]), torch.from_numpy(z["mask"])
    shifted, order = poly_order_loss(pred, mask)
    moved = (shifted != pred).any(-1)
    assert moved.any() and not moved[mask == 0].any() and float(order) > 0
    assert torch.equal(shifted[..., 0::2], pred[..., 0::2])     # radii untouched
