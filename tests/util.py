"""Shared helpers for the parity tests."""
import numpy as np
import torch


def gate(got, want, tol, name=""):
    """SURVEY §8 G2 gate: max|a-b| <= tol*max|b| and ||a-b||_2 <= tol*||b||_2."""
    got = np.asarray(got, dtype=np.float64)
    want = np.asarray(want, dtype=np.float64)
    assert got.shape == want.shape, (name, got.shape, want.shape)
    scale = max(np.abs(want).max(), 1e-30)
    err_max = np.abs(got - want).max() / scale
    err_l2 = np.linalg.norm(got - want) / max(np.linalg.norm(want), 1e-30)
    assert err_max <= tol and err_l2 <= tol, f"{name}: max-rel {err_max:.3e}, l2-rel {err_l2:.3e} > {tol}"
    return err_max, err_l2


def make_dcn_case(seed, B, Cin, Cout, H, W, k=3, stride=1, pad=1, dil=1, dg=1, offset_mode="random", dtype=torch.float32):
    """SURVEY §8(d) config 3 distributions: x~N(0,1); offset~N(0,2^2) with 5% far taps; mask=sigmoid(N(0,1));
    W~U(+-1/sqrt(9Cin)); b~N(0,0.1^2); grad_out~N(0,1)."""
    g = torch.Generator().manual_seed(seed)
    Ho = (H + 2 * pad - (dil * (k - 1) + 1)) // stride + 1
    Wo = (W + 2 * pad - (dil * (k - 1) + 1)) // stride + 1
    x = torch.randn(B, Cin, H, W, generator=g)
    if offset_mode == "zero":
        off = torch.zeros(B, dg * 2 * k * k, Ho, Wo)
    else:
        off = torch.randn(B, dg * 2 * k * k, Ho, Wo, generator=g) * 2.0
        far = torch.rand(off.shape, generator=g) < 0.05
        off[far] = (torch.rand(int(far.sum()), generator=g) * 2 - 1) * W
    mask = torch.sigmoid(torch.randn(B, dg * k * k, Ho, Wo, generator=g))
    bound = 1.0 / (Cin * k * k) ** 0.5
    w = (torch.rand(Cout, Cin, k, k, generator=g) * 2 - 1) * bound
    b = torch.randn(Cout, generator=g) * 0.1
    gout = torch.randn(B, Cout, Ho, Wo, generator=g)
    if dtype == torch.bfloat16:
        # bf16 mode: the oracle sees exactly the bf16-rounded activations; weights stay fp32 masters
        x, off, mask, gout = (t.bfloat16().float() for t in (x, off, mask, gout))
    return dict(x=x, offset=off, mask=mask, weight=w, bias=b, gout=gout)


def tie_free_heat(gen, B, C, H, W):
    n = B * C * H * W
    logits = torch.linspace(-6.0, 1.5, n)[torch.randperm(n, generator=gen)]
    heat = torch.sigmoid(logits)
    assert torch.unique(heat).numel() == n
    return heat.view(B, C, H, W).contiguous()


def make_decode_case(seed, B, C, H, W, N, rep="cartesian", with_reg=True):
    gen = torch.Generator().manual_seed(seed)
    heat = tie_free_heat(gen, B, C, H, W)
    if rep == "polar":
        polys = torch.empty(B, 2 * N, H, W)
        polys[:, 0::2] = (torch.randn(B, N, H, W, generator=gen) * 10).abs()
        polys[:, 1::2] = torch.rand(B, N, H, W, generator=gen) * 2 * np.pi
    else:
        polys = torch.randn(B, 2 * N, H, W, generator=gen) * 6
    depth = torch.rand(B, 1, H, W, generator=gen) * 50
    reg = torch.rand(B, 2, H, W, generator=gen) if with_reg else None
    return heat, polys, depth, reg
