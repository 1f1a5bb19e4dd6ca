"""GPU parity: the fused decode kernel (through the C ABI) against the reference's golden outputs and the
numpy oracle.  Gate (north_star): top-K indices and classes bit-exact; scores / vertices <= 1e-3 relative
(they are in fact bit-identical in cartesian mode)."""
import glob
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN
from util import make_decode_case

pytestmark = pytest.mark.gpu

DEC_GOLDEN = sorted(glob.glob(os.path.join(GOLDEN, "decode_*.npz")))


def ours(heat, polys, depth, reg, K, rep, fn="polydet_decode", dtype=None, channels_last=False):
    import centerpoly_b200 as cp
    def prep(t):
        if t is None:
            return None
        t = torch.as_tensor(t).cuda()
        if dtype is not None:
            t = t.to(dtype)
        if channels_last:
            t = t.contiguous(memory_format=torch.channels_last)
        return t
    out = getattr(cp, fn)(prep(heat), prep(polys), prep(depth), reg=prep(reg), K=K, rep=rep)
    torch.cuda.synchronize()
    return out.cpu().numpy()


def check(got, want, rep):
    assert got.shape == want.shape and got.dtype == np.float32
    np.testing.assert_array_equal(got[..., 5], want[..., 5], err_msg="class")
    np.testing.assert_array_equal(got[..., 4], want[..., 4], err_msg="score / order")
    if rep in ("polar", "polar_fixed"):
        np.testing.assert_allclose(got, want, rtol=1e-6, atol=2e-6)   # device vs libm cos/sin: <= 1 ulp
    else:
        np.testing.assert_array_equal(got, want)


@pytest.mark.parametrize("path", DEC_GOLDEN, ids=lambda p: os.path.basename(p)[:-4])
def test_decode_vs_reference_golden(path):
    g = np.load(path)
    reg = g["reg"] if g["reg"].size else None
    rep = str(g["rep"])
    got = ours(g["heat"], g["polys"], g["depth"], reg, int(g["K"]), rep)
    check(got, g["out"], rep)


CASES = [  # (B, C, H, W, N, K, rep, reg)
    (1, 8, 128, 256, 32, 100, "cartesian", True),     # config 1 head shape
    (1, 8, 256, 512, 32, 100, "cartesian", True),     # config 2 head shape (2048x1024 frame)
    (4, 9, 272, 488, 32, 128, "polar", True),         # config 5 (IDD) head shape, ragged width
    (2, 3, 37, 131, 8, 50, "polar_fixed", False),     # odd sizes: partial band, unaligned rows
    (2, 1, 64, 64, 16, 1024, "cartesian", True),      # K at the supported maximum
    (1, 2, 5, 3, 4, 15, "cartesian", False),          # K == H*W
]


@pytest.mark.parametrize("case", CASES, ids=lambda c: "-".join(map(str, c)))
def test_decode_vs_oracle(case):
    from oracle.decode_oracle import polydet_decode_oracle
    B, C, H, W, N, K, rep, with_reg = case
    heat, polys, depth, reg = make_decode_case(4242 + H, B, C, H, W, N, rep, with_reg)
    want = polydet_decode_oracle(heat.numpy(), polys.numpy(), depth.numpy(), reg=None if reg is None else reg.numpy(), K=K, rep=rep)
    got = ours(heat, polys, depth, reg, K, rep)
    check(got, want, rep)


def test_decode_channels_last_heads():
    from oracle.decode_oracle import polydet_decode_oracle
    heat, polys, depth, reg = make_decode_case(9, 2, 8, 64, 96, 32)
    want = polydet_decode_oracle(heat.numpy(), polys.numpy(), depth.numpy(), reg=reg.numpy(), K=100)
    got = ours(heat, polys, depth, reg, 100, "cartesian", channels_last=True)
    check(got, want, "cartesian")


def test_decode_plateau_and_sparse_edge_cases():
    """Tie-aware edge cases under the documented tie rule (score desc, flat index asc): a constant heat map
    (every pixel is a peak) and a map with fewer than K positive peaks (zero-score fillers)."""
    from oracle.decode_oracle import polydet_decode_oracle
    gen = torch.Generator().manual_seed(1)
    B, C, H, W, N, K = 2, 3, 20, 140, 4, 40
    polys = torch.randn(B, 2 * N, H, W, generator=gen)
    depth = torch.rand(B, 1, H, W, generator=gen)
    reg = torch.rand(B, 2, H, W, generator=gen)
    flat = torch.full((B, C, H, W), 0.25)
    want = polydet_decode_oracle(flat.numpy(), polys.numpy(), depth.numpy(), reg=reg.numpy(), K=K)
    check(ours(flat, polys, depth, reg, K, "cartesian"), want, "cartesian")
    sparse = torch.zeros(B, C, H, W)
    sparse[0, 1, 3, 5] = 0.9
    sparse[0, 2, 10, 100] = 0.8
    sparse[1, 0, 19, 139] = 0.7
    want = polydet_decode_oracle(sparse.numpy(), polys.numpy(), depth.numpy(), reg=reg.numpy(), K=K)
    got = ours(sparse, polys, depth, reg, K, "cartesian")
    np.testing.assert_array_equal(got[..., 4], want[..., 4])
    pos = want[..., 4] > 0
    np.testing.assert_array_equal(got[pos], want[pos])         # real detections: exact
    np.testing.assert_array_equal(got[~pos], want[~pos])       # fillers: same rule as the oracle


def test_decode_from_logits_matches_sigmoid_then_decode():
    """polydet_decode_from_logits (the entry point the engine and bench.py time) == sigmoid followed by polydet_decode.
    (a) logits on a tie-free grid (neighbouring sigmoid values >= 180 ulp apart, so a last-ulp difference between the
        device expf and ATen's sigmoid cannot reorder anything): every row, in order, is identical - class, box,
        vertices, depth bit-exact, scores within 2 ulp.
    (b) N(-2.19, 1) logits as the network produces them, tie-aware: every row whose score is separated from its
        neighbours (and from the K+1-th candidate) by more than 4 ulp must be identical; the rest must be the same
        multiset of classes."""
    import centerpoly_b200 as cp
    _, polys, depth, reg = make_decode_case(3, 2, 8, 64, 96, 16)
    polys, depth, reg = polys.cuda(), depth.cuda(), reg.cuda()
    n = 2 * 8 * 64 * 96
    gen = torch.Generator().manual_seed(2)
    grid = torch.linspace(-6.0, 1.5, n)[torch.randperm(n, generator=gen)].view(2, 8, 64, 96).cuda()
    a = cp.polydet_decode_from_logits(grid, polys, depth, reg=reg, K=100).cpu().numpy()
    b = cp.polydet_decode(torch.sigmoid(grid), polys, depth, reg=reg, K=100).cpu().numpy()
    np.testing.assert_allclose(a[..., 4], b[..., 4], rtol=2.4e-7)
    np.testing.assert_array_equal(np.delete(a, 4, axis=2), np.delete(b, 4, axis=2))
    logits = (torch.randn(2, 8, 64, 96, generator=gen) - 2.19).cuda()
    a = cp.polydet_decode_from_logits(logits, polys, depth, reg=reg, K=100).cpu().numpy()
    b = cp.polydet_decode(torch.sigmoid(logits), polys, depth, reg=reg, K=101).cpu().numpy()   # one more: the K boundary
    compared = 0
    for img in range(2):
        sb = b[img, :, 4].astype(np.float64)
        np.testing.assert_allclose(a[img, :, 4], sb[:100], rtol=2.4e-7)
        for i in range(100):
            gap = min(sb[i - 1] - sb[i] if i else np.inf, sb[i] - sb[i + 1])
            if gap > 4 * 6e-8 * sb[i]:
                np.testing.assert_array_equal(np.delete(a[img, i], 4), np.delete(b[img, i], 4), err_msg=f"image {img} row {i}")
                compared += 1
        if np.all(sb[99] - sb[100] > 4 * 6e-8 * sb[99]):
            assert sorted(a[img, :, 5].tolist()) == sorted(b[img, :100, 5].tolist())
    assert compared >= 180, compared


def test_decode_bf16_heads_computed_in_fp32():
    from oracle.decode_oracle import polydet_decode_oracle
    heat, polys, depth, reg = make_decode_case(21, 1, 8, 64, 96, 32)
    polys16, depth16, reg16 = polys.bfloat16(), depth.bfloat16(), reg.bfloat16()
    # heat stays tie-free only in fp32; feed bf16 heads + fp32 heat through the fp32 upcast path
    want = polydet_decode_oracle(heat.numpy(), polys16.float().numpy(), depth16.float().numpy(), reg=reg16.float().numpy(), K=100)
    import centerpoly_b200 as cp
    got = cp.polydet_decode(heat.cuda(), polys16.cuda(), depth16.cuda(), reg=reg16.cuda(), K=100).cpu().numpy()
    check(got, want, "cartesian")


def test_decode_errors():
    import centerpoly_b200 as cp
    z = torch.zeros(1, 2, 4, 4, device="cuda")
    with pytest.raises(NotImplementedError):
        cp.polydet_decode(z, torch.zeros(1, 8, 4, 4, device="cuda"), z[:, :1], cat_spec_poly=True, K=4)
    with pytest.raises(RuntimeError):
        cp.polydet_decode(z, torch.zeros(1, 8, 4, 4, device="cuda"), z[:, :1], K=100)  # K > H*W
    with pytest.raises(RuntimeError):
        cp.polydet_decode(z.cpu(), torch.zeros(1, 8, 4, 4), z[:, :1].cpu(), K=4)


POST_GOLDEN = sorted(__import__("glob").glob(os.path.join(os.path.dirname(__file__), "golden", "post_*.npz")))


@pytest.mark.gpu
@pytest.mark.parametrize("path", POST_GOLDEN, ids=lambda p: os.path.basename(p)[:-4])
def test_fused_post_process_vs_reference_golden(path):
    """decode + fused back-mapping (cpoly_decode with post_affine) == reference polydet_decode followed by the
    reference polydet_post_process (golden vectors of both, tests/golden/make_golden.py)."""
    import centerpoly_b200 as cp
    d = np.load(path)
    src = {"post_fixres_small": "decode_small_cart_reg", "post_keepres_small": "decode_small_polar_reg",
           "post_fixres_mid": "decode_mid_cart_reg", "post_keepres_mid": "decode_mid_polar_reg"}[os.path.basename(path)[:-4]]
    g = np.load(os.path.join(os.path.dirname(path), src + ".npz"))
    B = d["dets"].shape[0]
    c = [d["c"][i] for i in range(B)]
    s = [float(d["s"][i][0]) if bool(d["s_is_scalar"]) else d["s"][i] for i in range(B)]
    h, w, nc = int(d["out_h"]), int(d["out_w"]), int(d["num_classes"])
    trans = cp.batch_inverse_affine(c, s, w, h)
    t = lambda k: torch.from_numpy(g[k]).cuda()
    out = cp.polydet_decode(t("heat"), t("polys"), t("depth"), reg=t("reg"), K=int(g["K"]), rep=str(g["rep"]), post_affine=trans)
    ret = cp.group_by_class(out.cpu().numpy(), nc)
    D = d["dets"].shape[2] - 1
    for i in range(B):
        for j in range(1, nc + 1):
            got = np.array(ret[i][j], dtype=np.float32).reshape(-1, D)
            want = d[f"ret_{i}_{j}"]
            assert got.shape == want.shape
            assert np.array_equal(got[:, 4], want[:, 4]) and np.array_equal(got[:, -1], want[:, -1])   # score, depth
            np.testing.assert_allclose(got, want, rtol=2e-6, atol=0)   # polar vertices: <= 1 ulp of cos/sin, then the affine
    # and without the affine the call is still the plain decode
    plain = cp.polydet_decode(t("heat"), t("polys"), t("depth"), reg=t("reg"), K=int(g["K"]), rep=str(g["rep"]))
    np.testing.assert_allclose(plain.cpu().numpy(), g["out"], rtol=2e-6, atol=0)
