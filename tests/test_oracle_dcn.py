"""CPU: pins the C DCNv2 oracle against the committed torchvision-fp64 golden vectors and, live,
against torchvision.ops.deform_conv2d (the stand-in BASELINE.json names for the missing submodule)."""
import glob
import os

import numpy as np
import pytest
import torch

from oracle import dcn_oracle

from conftest import GOLDEN

DCN_GOLDEN = sorted(glob.glob(os.path.join(GOLDEN, "dcn_*.npz")))


def taps_on_minus_one(offset, H, W, s, p, d, dg, k=3):
    """Boolean array shaped like ``offset``: taps whose sample point has p_y == -1 or p_x == -1 exactly."""
    B, _, Ho, Wo = offset.shape
    ys, xs = np.meshgrid(np.arange(Ho), np.arange(Wo), indexing="ij")
    edge = np.zeros(offset.shape, dtype=bool)
    T = k * k
    for g in range(dg):
        for t in range(T):
            i, j = divmod(t, k)
            cy, cx = g * 2 * T + 2 * t, g * 2 * T + 2 * t + 1
            py = ys * s - p + i * d + offset[:, cy]
            px = xs * s - p + j * d + offset[:, cx]
            e = (py == -1) | (px == -1)
            edge[:, cy] = e
            edge[:, cx] = e
    return edge


def test_golden_files_present():
    assert len(DCN_GOLDEN) >= 6


@pytest.mark.parametrize("path", DCN_GOLDEN, ids=lambda p: os.path.basename(p)[:-4])
def test_oracle_matches_golden_f64(path):
    g = np.load(path)
    s, p, d, dg = (int(v) for v in g["geom"])
    y = dcn_oracle.dcn_forward(g["x"], g["offset"], g["mask"], g["weight"], g["bias"], s, p, d, dg)
    np.testing.assert_allclose(y, g["y"], rtol=0, atol=1e-12)
    gx, go, gm, gw, gb = dcn_oracle.dcn_backward(g["x"], g["offset"], g["mask"], g["weight"], g["gout"], s, p, d, dg)
    edge = taps_on_minus_one(g["offset"], g["x"].shape[2], g["x"].shape[3], s, p, d, dg)
    for name, val in (("gx", gx), ("goffset", go), ("gmask", gm), ("gw", gw), ("gb", gb)):
        want = g[name]
        if name == "goffset":
            # Known, documented divergence of the stand-in: at p_y == -1 or p_x == -1 EXACTLY upstream
            # DCNv2 treats the tap as dead (dmcn_get_coordinate_weight returns 0 for argmax <= -1) while
            # torchvision returns the one-sided difference.  The oracle follows DCNv2.
            assert np.all(val[edge] == 0.0)
            val, want = np.where(edge, 0.0, val), np.where(edge, 0.0, want)
        np.testing.assert_allclose(val, want, rtol=0, atol=1e-11, err_msg=name)


@pytest.mark.parametrize("path", DCN_GOLDEN[:2], ids=lambda p: os.path.basename(p)[:-4])
def test_oracle_f32_close_to_f64(path):
    g = np.load(path)
    s, p, d, dg = (int(v) for v in g["geom"])
    y = dcn_oracle.dcn_forward(g["x"], g["offset"], g["mask"], g["weight"], g["bias"], s, p, d, dg, dtype=np.float32)
    assert y.dtype == np.float32
    np.testing.assert_allclose(y, g["y"], rtol=1e-4, atol=1e-5)


def test_zero_offset_is_half_conv():
    """The reference's true init state: zero conv_offset_mask => offsets 0, mask 0.5."""
    rng = np.random.default_rng(5)
    x = rng.standard_normal((1, 3, 8, 10))
    w = rng.standard_normal((4, 3, 3, 3)) * 0.2
    b = rng.standard_normal(4)
    off = np.zeros((1, 18, 8, 10))
    msk = np.full((1, 9, 8, 10), 0.5)
    y = dcn_oracle.dcn_forward(x, off, msk, w, b)
    ref = torch.nn.functional.conv2d(torch.from_numpy(x), torch.from_numpy(w), None, padding=1).numpy() * 0.5
    ref = ref + b[None, :, None, None]
    np.testing.assert_allclose(y, ref, atol=1e-12)


def test_oracle_vs_torchvision_live_idaup_shape():
    """A (reduced-resolution) IDAUp shape class, random offsets with 5% far out-of-bounds taps."""
    from torchvision.ops import deform_conv2d
    gen = torch.Generator().manual_seed(317)
    B, Cin, Cout, H, W = 1, 64, 64, 12, 16
    x = torch.randn(B, Cin, H, W, generator=gen, dtype=torch.float64)
    off = torch.randn(B, 18, H, W, generator=gen, dtype=torch.float64) * 2
    far = torch.rand(B, 18, H, W, generator=gen) < 0.05
    off[far] = (torch.rand(int(far.sum()), generator=gen, dtype=torch.float64) * 2 - 1) * W
    msk = torch.sigmoid(torch.randn(B, 9, H, W, generator=gen, dtype=torch.float64))
    bound = 1.0 / np.sqrt(9 * Cin)
    w = (torch.rand(Cout, Cin, 3, 3, generator=gen, dtype=torch.float64) * 2 - 1) * bound
    b = torch.randn(Cout, generator=gen, dtype=torch.float64) * 0.1
    gout = torch.randn(B, Cout, H, W, generator=gen, dtype=torch.float64)
    leaves = [t.clone().requires_grad_(True) for t in (x, off, msk, w, b)]
    y = deform_conv2d(leaves[0], leaves[1], leaves[3], leaves[4], padding=1, mask=leaves[2])
    y.backward(gout)
    yo = dcn_oracle.dcn_forward(x.numpy(), off.numpy(), msk.numpy(), w.numpy(), b.numpy())
    np.testing.assert_allclose(yo, y.detach().numpy(), atol=1e-11)
    grads = dcn_oracle.dcn_backward(x.numpy(), off.numpy(), msk.numpy(), w.numpy(), gout.numpy())
    for got, leaf in zip(grads, leaves):
        np.testing.assert_allclose(got, leaf.grad.numpy(), atol=1e-10)
