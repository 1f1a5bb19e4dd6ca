"""GPU parity: the CUDA DCN (through the C ABI) against the CPU oracle and the committed golden vectors.
Tolerances (BASELINE.json north_star): 1e-3 relative for fp32 I/O (tf32 tensor-core products), 1e-2 for bf16."""
import glob
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN
from util import gate, make_dcn_case

pytestmark = pytest.mark.gpu

DCN_GOLDEN = sorted(glob.glob(os.path.join(GOLDEN, "dcn_*.npz")))


def _cuda(t, dtype=None, channels_last=False):
    t = t.cuda()
    if dtype is not None:
        t = t.to(dtype)
    if channels_last:
        t = t.contiguous(memory_format=torch.channels_last)
    return t


def run_ours(case, path, dtype=torch.float32, channels_last=False, stride=1, pad=1, dil=1, dg=1, backward=True):
    import centerpoly_b200 as cp
    cp.set_dcn_path(path)
    try:
        x = _cuda(case["x"], dtype, channels_last).requires_grad_(True)
        off = _cuda(case["offset"], dtype).requires_grad_(True)
        msk = _cuda(case["mask"], dtype).requires_grad_(True)
        w = _cuda(case["weight"]).requires_grad_(True)
        b = _cuda(case["bias"]).requires_grad_(True)
        y = cp.dcn_v2_conv(x, off, msk, w, b, stride, pad, dil, dg)
        out = {"y": y.detach().float().cpu().numpy()}
        if backward:
            y.backward(_cuda(case["gout"], dtype, channels_last))
            for name, leaf in (("gx", x), ("goffset", off), ("gmask", msk), ("gw", w), ("gb", b)):
                out[name] = leaf.grad.detach().float().cpu().numpy()
        torch.cuda.synchronize()
        return out
    finally:
        cp.set_dcn_path("auto")


def oracle(case, stride=1, pad=1, dil=1, dg=1, backward=True):
    from oracle import dcn_oracle
    a = {k: v.double().numpy() for k, v in case.items()}
    out = {"y": dcn_oracle.dcn_forward(a["x"], a["offset"], a["mask"], a["weight"], a["bias"], stride, pad, dil, dg)}
    if backward:
        gx, go, gm, gw, gb = dcn_oracle.dcn_backward(a["x"], a["offset"], a["mask"], a["weight"], a["gout"], stride, pad, dil, dg)
        out.update(gx=gx, goffset=go, gmask=gm, gw=gw, gb=gb)
    return out


@pytest.mark.parametrize("path", DCN_GOLDEN, ids=lambda p: os.path.basename(p)[:-4])
def test_simt_vs_golden_all_geometries(path):
    """Generic path against the committed torchvision-fp64 vectors (stride 2, dilation 2, 2 deformable groups,
    far out-of-bounds offsets, zero offsets, integer offsets)."""
    g = np.load(path)
    s, p, d, dg = (int(v) for v in g["geom"])
    case = {k: torch.from_numpy(g[k]).float() for k in ("x", "offset", "mask", "weight", "bias", "gout")}
    got = run_ours(case, "simt", stride=s, pad=p, dil=d, dg=dg)
    want = oracle(case, s, p, d, dg)  # C oracle == golden except the documented p == -1 grad_offset convention
    np.testing.assert_allclose(want["y"], g["y"], atol=1e-5)
    for name in ("y", "gx", "goffset", "gmask", "gw", "gb"):
        gate(got[name], want[name], 1e-4, name)


TC_SHAPES = [  # (B, Cin, Cout, H, W) — reduced-resolution versions of the 7 IDAUp shape classes + ragged sizes
    (1, 64, 64, 16, 32), (2, 128, 64, 24, 40), (1, 128, 128, 16, 16), (1, 256, 64, 8, 16),
    (1, 256, 128, 9, 21), (1, 256, 256, 8, 16), (1, 512, 256, 8, 16), (3, 64, 64, 5, 7),
    (1, 128, 64, 72, 100),    # 63 tiles  -> tap split 3
    (1, 64, 64, 100, 200),    # 169 tiles -> unsplit persistent schedule, ragged right/bottom tiles
    # three tap-split items per CTA whose items have fewer steps than there are sampler groups: a group's next step lies
    # two items ahead, which the look-ahead must not wait for (it deadlocked on its own item before the fix)
    (1, 64, 64, 88, 144),     # 99 tiles, split 3: 3 steps per item, 297 items on 148 CTAs
    (1, 128, 128, 40, 128),   # 40 tiles, split 9: 2 steps per item, 360 items
]


@pytest.mark.parametrize("sampler", ["gather", "window"])   # csrc/dcn_tc.cu (default) / csrc/dcn_tc_win.cu (TMA-staged window)
@pytest.mark.parametrize("shape", TC_SHAPES, ids=lambda s: "x".join(map(str, s)))
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32], ids=["bf16", "fp32"])
def test_tcgen05_forward_vs_oracle(shape, dtype, sampler, monkeypatch):
    monkeypatch.setenv("CPOLY_TC_SAMPLER", sampler)   # read by libcpoly.so at every call
    B, Cin, Cout, H, W = shape
    case = make_dcn_case(317 + Cin + Cout, B, Cin, Cout, H, W, dtype=dtype)
    got = run_ours(case, "tcgen05", dtype=dtype, channels_last=True, backward=False)
    want = oracle(case, backward=False)
    gate(got["y"], want["y"], 1e-2 if dtype == torch.bfloat16 else 1e-3, "y")


@pytest.mark.parametrize("sampler", ["gather", "window"])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32], ids=["bf16", "fp32"])
def test_tcgen05_zero_offsets(dtype, sampler, monkeypatch):
    monkeypatch.setenv("CPOLY_TC_SAMPLER", sampler)
    case = make_dcn_case(5, 1, 64, 64, 16, 32, offset_mode="zero", dtype=dtype)
    got = run_ours(case, "tcgen05", dtype=dtype, channels_last=True, backward=False)
    want = oracle(case, backward=False)
    gate(got["y"], want["y"], 1e-2 if dtype == torch.bfloat16 else 1e-3, "y")


BWD_SHAPES = [(1, 64, 64, 12, 20), (2, 128, 64, 9, 11), (1, 128, 128, 20, 36), (1, 256, 256, 9, 17), (1, 64, 64, 70, 90),
              (1, 512, 256, 8, 16)]


@pytest.mark.parametrize("shape", BWD_SHAPES, ids=lambda s: "x".join(map(str, s)))
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32], ids=["bf16", "fp32"])
def test_backward_vs_oracle(shape, dtype):
    B, Cin, Cout, H, W = shape
    case = make_dcn_case(99 + Cin, B, Cin, Cout, H, W, dtype=dtype)
    got = run_ours(case, "auto", dtype=dtype, channels_last=True)
    want = oracle(case)
    tol = 1e-2 if dtype == torch.bfloat16 else 1e-3
    for name in ("y", "gx", "goffset", "gmask", "gw", "gb"):
        gate(got[name], want[name], tol, name)


@pytest.mark.parametrize("shape", [(2, 64, 64, 12, 20), (1, 128, 64, 40, 56), (1, 64, 64, 70, 90)], ids=lambda s: "x".join(map(str, s)))
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32], ids=["bf16", "fp32"])
def test_backward_is_bitwise_deterministic(shape, dtype):
    """Every output of the backward, grad_input included, is bit-identical from run to run: the col2im is a
    gather over a sorted contribution list (csrc/dcn_col2im.cu), there is no floating-point atomic on the path."""
    case = make_dcn_case(7, *shape, dtype=dtype)
    a = run_ours(case, "auto", dtype=dtype, channels_last=True)
    for _ in range(2):
        b = run_ours(case, "auto", dtype=dtype, channels_last=True)
        for name in ("y", "gx", "goffset", "gmask", "gw", "gb"):
            np.testing.assert_array_equal(a[name], b[name], err_msg=name)


FULL_GOLDEN = sorted(glob.glob(os.path.join(GOLDEN, "dcnfull_*.npz")))


def _take(t, flat):
    B, C, H, W = t.shape
    return t.permute(0, 2, 3, 1).reshape(B * H * W, C)[flat]


@pytest.mark.parametrize("path", FULL_GOLDEN, ids=lambda p: os.path.basename(p)[:-4])
def test_full_size_vs_torchvision_golden(path):
    """BASELINE.json configs[2] shapes at FULL size (64->64 @256x512, 128->128 @128x256; B = 1 and 4), random offsets
    with 5 % far taps: forward + all five gradients against torchvision.deform_conv2d float64 (fixtures made by
    tests/golden/make_dcn_fullsize_golden.py: expected values at 2048 seeded pixels, gw / gb in full, norms of the full
    tensors).  Inputs are regenerated from the stored seed.  bf16-rounded inputs are run in bf16 (1e-2) and fp32
    (1e-3) mode, fp32 inputs in fp32 mode.  The full tensors are additionally cross-checked against the generic
    CUDA-core path, which sees every pixel."""
    g = np.load(path)
    B, Cin, Cout, H, W = (int(v) for v in g["shape"])
    rounding = torch.bfloat16 if str(g["rounding"]) == "bfloat16" else torch.float32
    case = make_dcn_case(int(g["seed"]), B, Cin, Cout, H, W, dtype=rounding)
    flat = torch.from_numpy(g["positions"])
    # A sample coordinate that fp32 arithmetic rounds onto an integer (y - 1 + i + dy == 266.0f exactly, 265.99999976 in
    # float64) sits on a kink of the piecewise-bilinear function: the kernels (fp32 positions, like upstream) and the
    # float64 stand-in then take the derivative from different sides.  Those grad_offset entries (about one in 10^6)
    # are exempt; everything else is gated.
    off_s = _take(case["offset"], flat).double()
    yy, xx = ((flat % (H * W)) // W).double(), (flat % W).double()
    kink = torch.zeros_like(off_s, dtype=torch.bool)
    for k in range(9):
        py, px = yy - 1 + k // 3 + off_s[:, 2 * k], xx - 1 + k % 3 + off_s[:, 2 * k + 1]
        near = ((py - py.round()).abs() < 1e-4) | ((px - px.round()).abs() < 1e-4)
        kink[:, 2 * k] = near
        kink[:, 2 * k + 1] = near
    kink = kink.numpy()
    modes = [torch.float32] if rounding == torch.float32 else [torch.bfloat16, torch.float32]
    ref_full = None
    for dtype in modes:
        tol = 1e-2 if dtype == torch.bfloat16 else 1e-3
        got = run_ours(case, "tcgen05", dtype=dtype, channels_last=True)
        for name in ("y", "gx", "goffset", "gmask"):
            want = g[name].astype(np.float64)
            have = _take(torch.from_numpy(got[name]), flat).numpy().astype(np.float64)
            if name == "goffset":
                have = np.where(kink, want, have)
            err_max = np.abs(have - want).max() / float(g[name + "_max"])
            err_l2 = np.linalg.norm(have - want) / np.linalg.norm(want)
            assert err_max <= tol and err_l2 <= tol, f"{name} [{dtype}]: max-rel {err_max:.3e}, l2-rel {err_l2:.3e} > {tol}"
            l2 = np.linalg.norm(got[name].astype(np.float64))
            assert abs(l2 - float(g[name + "_l2"])) <= tol * float(g[name + "_l2"]), f"{name} [{dtype}]: full-tensor norm"
        gate(got["gw"], g["gw"], tol, "gw")
        gate(got["gb"], g["gb"], tol, "gb")
        if dtype == torch.float32 and B == 1:   # every pixel: tensor-core path vs the CUDA-core path (IEEE fp32)
            ref_full = run_ours(case, "simt", dtype=torch.float32, channels_last=True)
            for name in ("y", "gx", "goffset", "gmask", "gw", "gb"):
                gate(got[name], ref_full[name], 1e-3, "full-tensor " + name)


def test_fused_adam_step_is_seen_by_the_next_forward():
    """torch._fused_adam_ updates parameters without bumping Tensor._version: the packed weight image must not be
    served from the cache in a differentiated forward (dcn_v2._PackedWeights)."""
    import centerpoly_b200 as cp
    torch.manual_seed(5)
    m = cp.DCN(64, 64, (3, 3), 1, 1).cuda().to(memory_format=torch.channels_last)
    torch.nn.init.normal_(m.conv_offset_mask.weight, std=0.05)
    opt = torch.optim.Adam(m.parameters(), lr=0.05, fused=True)
    x = torch.randn(2, 64, 16, 32, device="cuda").contiguous(memory_format=torch.channels_last)
    y0 = m(x)
    y0.square().mean().backward()
    w_before = m.weight.detach().clone()
    opt.step()
    assert not torch.equal(w_before, m.weight.detach())
    y1 = m(x)                                   # training-mode forward after the fused step
    with torch.no_grad():
        om = m.conv_offset_mask(x)
        want = cp.dcn_v2_conv(x, om[:, :18], torch.sigmoid(om[:, 18:]), m.weight.detach().clone(), m.bias.detach().clone(), 1, 1, 1, 1)
    # (a stale weight image is an O(1) error; 1e-4 leaves room for the in-kernel sigmoid flipping a TF32 rounding)
    gate(y1.detach().float().cpu().numpy(), want.float().cpu().numpy(), 1e-4, "forward after fused Adam")
    assert not torch.allclose(y1.detach(), y0.detach())
    # no-grad inference after an out-of-band update: the cache is dropped by hand
    with torch.no_grad():
        m.weight.data.mul_(2.0)                 # .data: no version bump
        cp.invalidate_packed_weights(m)
        y2 = m(x)
        want2 = cp.dcn_v2_conv(x, om[:, :18], torch.sigmoid(om[:, 18:]), m.weight.detach().clone(), m.bias.detach().clone(), 1, 1, 1, 1)
    gate(y2.float().cpu().numpy(), want2.float().cpu().numpy(), 1e-4, "forward after invalidate")


def test_nchw_input_is_accepted():
    case = make_dcn_case(11, 1, 64, 64, 10, 12)
    got = run_ours(case, "auto", channels_last=False)
    want = oracle(case)
    for name in ("y", "gx", "gw"):
        gate(got[name], want[name], 1e-3, name)


def test_full_size_zero_offset_identity_and_linearity():
    """BASELINE full size (64->64 @ 256x512, the stride-4 node layers): size-independent properties.
    (1) zero conv_offset_mask => DCN(x) == 0.5*conv2d(x, W) + b  (the reference's true init state);
    (2) linearity in x for fixed offsets/mask."""
    import centerpoly_b200 as cp
    torch.manual_seed(0)
    m = cp.DCN(64, 64, (3, 3), 1, 1).cuda().to(memory_format=torch.channels_last)
    x = torch.randn(1, 64, 256, 512, device="cuda").contiguous(memory_format=torch.channels_last)
    with torch.no_grad():
        y = m(x)
        ref = 0.5 * torch.nn.functional.conv2d(x.double(), m.weight.double(), None, padding=1) + m.bias.double().view(1, -1, 1, 1)
        gate(y.double().cpu().numpy(), ref.cpu().numpy(), 1e-3, "identity")
        m.conv_offset_mask.weight.normal_(std=0.05)
        m.conv_offset_mask.bias.normal_(std=1.0)
        om = m.conv_offset_mask(x)
        x2 = torch.randn_like(x)
        f = lambda t: cp.dcn_v2_conv(t, om[:, :18].float(), torch.sigmoid(om[:, 18:]).float(), m.weight, None, 1, 1, 1, 1)
        lhs = f(x + 2 * x2)
        rhs = f(x) + 2 * f(x2)
        gate(lhs.double().cpu().numpy(), rhs.double().cpu().numpy(), 2e-3, "linearity")


def test_module_matches_torchvision_module(monkeypatch):
    """DCN module (fused sigmoid/chunk path) vs the torchvision-backed oracle module, same state_dict."""
    import centerpoly_b200 as cp
    from oracle.tv_dcn import TVDCN
    # cuDNN's default TF32 convolution would perturb the OFFSETS (conv_offset_mask output) by ~1e-3 relative and
    # move the sampling points; with the flag off both the offset conv and (by the same flag, dcn_v2._tc_allowed)
    # our DCN run in IEEE fp32.
    monkeypatch.setattr(torch.backends.cudnn, "allow_tf32", False)
    torch.manual_seed(3)
    ref = TVDCN(64, 64, (3, 3), 1, 1)
    torch.nn.init.normal_(ref.conv_offset_mask.weight, std=0.05)
    torch.nn.init.normal_(ref.conv_offset_mask.bias, std=1.0)
    ours = cp.DCN(64, 64, (3, 3), 1, 1)
    assert list(ours.state_dict().keys()) == list(ref.state_dict().keys())
    ours.load_state_dict(ref.state_dict())
    ours = ours.cuda()
    x = torch.randn(2, 64, 14, 18)
    xr = x.clone().requires_grad_(True)
    yr = ref(xr)
    gout = torch.randn_like(yr)
    yr.backward(gout)
    xo = x.cuda().requires_grad_(True)
    yo = ours(xo)
    yo.backward(gout.cuda())
    gate(yo.detach().cpu().numpy(), yr.detach().numpy(), 1e-3, "y")
    gate(xo.grad.cpu().numpy(), xr.grad.numpy(), 1e-3, "gx")
    for (n, po), (_, pr) in zip(ours.named_parameters(), ref.named_parameters()):
        gate(po.grad.cpu().numpy(), pr.grad.numpy(), 2e-3, n)


def test_frozen_weights_get_no_grad():
    import centerpoly_b200 as cp
    m = cp.DCN(64, 64, (3, 3), 1, 1).cuda()
    m.weight.requires_grad_(False)
    x = torch.randn(1, 64, 8, 8, device="cuda", requires_grad=True)
    m(x).sum().backward()
    assert m.weight.grad is None and x.grad is not None and m.bias.grad is not None


def test_errors_are_runtime_errors():
    import centerpoly_b200 as cp
    x = torch.randn(1, 4, 8, 8, device="cuda")
    w = torch.randn(4, 4, 3, 3, device="cuda")
    with pytest.raises(RuntimeError):
        cp.dcn_v2_conv(x, torch.zeros(1, 17, 8, 8, device="cuda"), torch.zeros(1, 9, 8, 8, device="cuda"), w, None, 1, 1, 1, 1)
    with pytest.raises(RuntimeError):
        cp.dcn_v2_conv(x.cpu(), torch.zeros(1, 18, 8, 8), torch.zeros(1, 9, 8, 8), w.cpu(), None, 1, 1, 1, 1)


def test_pack_cache_is_not_confused_by_recycled_addresses():
    """A freed weight tensor's address is routinely recycled by the caching allocator for the next weight of the
    same shape: the packed-weight cache must follow the tensor object, not the address."""
    import centerpoly_b200 as cp
    case = make_dcn_case(3, 1, 64, 64, 16, 32)
    x = case["x"].cuda().contiguous(memory_format=torch.channels_last)
    off, msk = case["offset"].cuda(), case["mask"].cuda()
    outs, ptrs = [], []
    for scale in (1.0, -2.0, 0.5):
        w = (case["weight"] * scale).cuda()
        ptrs.append(w.data_ptr())
        outs.append(cp.dcn_v2_conv(x, off, msk, w, None, 1, 1, 1, 1).float().cpu())
        del w
    gate(outs[1].numpy(), (-2.0 * outs[0]).numpy(), 2e-3, "recycled weight 1")
    gate(outs[2].numpy(), (0.5 * outs[0]).numpy(), 2e-3, "recycled weight 2")
    # in-place update of the same tensor (optimizer step) must also be seen
    w = case["weight"].cuda()
    y0 = cp.dcn_v2_conv(x, off, msk, w, None, 1, 1, 1, 1).float().cpu()
    w.mul_(3.0)
    y1 = cp.dcn_v2_conv(x, off, msk, w, None, 1, 1, 1, 1).float().cpu()
    gate(y1.numpy(), (3.0 * y0).numpy(), 2e-3, "in-place update")


def test_fused_module_backward_returns_offset_gradients_in_the_producers_dtype():
    """bf16 `om` (autocast training): d/d(om) comes back as bf16 straight from the kernels (cpoly_dcn_bwd_args.gom_dtype)
    and equals the fp32 result up to one bf16 rounding."""
    from centerpoly_b200.dcn_v2 import _DCNFused
    torch.manual_seed(23)
    B, C, H, W = 2, 64, 24, 40
    x = torch.randn(B, C, H, W, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    om32 = torch.randn(B, 27, H, W, device="cuda")
    om32[:, :18] *= 1.5
    om32 = om32.to(torch.bfloat16).float().contiguous(memory_format=torch.channels_last)
    w = (torch.rand(C, C, 3, 3, device="cuda") - 0.5) * 0.1
    bias = torch.zeros(C, device="cuda")
    gy = torch.randn(B, C, H, W, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    res = []
    for om in (om32, om32.to(torch.bfloat16)):
        omr = om.clone().requires_grad_(True)
        y = _DCNFused.apply(x, omr, w, bias, (1, 1), (1, 1), (1, 1), 1)
        y.backward(gy)
        assert omr.grad.dtype == om.dtype
        res.append(omr.grad.float().cpu().numpy())
    gate(res[1], res[0], 1e-2, "d/d(om) bf16 vs fp32")


def test_tcgen05_schedules_over_random_shapes_match_the_simt_path():
    """Every (tile count, tap split, channel chunks, items per CTA) combination takes a different walk through the forward
    and backward kernels' barrier rings; 16 seeded random shapes against the oracle-verified CUDA-core path (both ours, same
    inputs), so that a schedule that hangs or mis-orders shows up here and not at a user's image size."""
    import random
    rng = random.Random(20261105)
    for i in range(16):
        Cin = rng.choice([64, 64, 128, 256])
        Cout = rng.choice([16, 64, 64, 128, 256])
        B = rng.choice([1, 1, 2, 3])
        H, W = rng.randint(5, 150), rng.randint(5, 170)
        if B * H * W > 40000:
            H = max(5, 40000 // (B * W))
        case = make_dcn_case(1000 + i, B, Cin, Cout, H, W, dtype=torch.bfloat16)
        bwd = i % 4 == 0 and B * H * W < 12000
        tc = run_ours(case, "tcgen05", dtype=torch.bfloat16, channels_last=True, backward=bwd)
        ref = run_ours(case, "simt", dtype=torch.bfloat16, channels_last=True, backward=bwd)
        gate(tc["y"], ref["y"], 1e-2, f"y {B}x{Cin}->{Cout}@{H}x{W}")
        if bwd:
            gate(tc["gw"], ref["gw"], 1e-2, f"gw {B}x{Cin}->{Cout}@{H}x{W}")
            gate(tc["gx"], ref["gx"], 1e-2, f"gx {B}x{Cin}->{Cout}@{H}x{W}")
