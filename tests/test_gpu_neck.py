"""GPU parity of the neck rewrites: IDAUp up-sampling kernel, fused DeformConv epilogue, and the
whole optimised network against the unoptimised one.  Oracle for the up-sampler = torch's own CPU
``conv_transpose2d`` with the reference's arguments (pose_dla_dcn.py:372-375); for the fused DeformConv = the
unfused DCN -> BatchNorm2d(eval) -> ReLU sequence of pose_dla_dcn.py:347-359 evaluated on the oracle's DCN output."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from util import gate, make_dcn_case

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("f", [2, 4, 8])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16], ids=["fp32", "bf16"])
@pytest.mark.parametrize("layout", ["nhwc", "nchw"])
def test_idaup_upsample_vs_conv_transpose(f, dtype, layout):
    from centerpoly_b200.neck import idaup_upsample
    g = torch.Generator().manual_seed(10 + f)
    B, C, H, W = 2, 64, 9, 13
    x = torch.randn(B, C, H, W, generator=g)
    w = torch.randn(C, 1, 2 * f, 2 * f, generator=g) * 0.3
    skip = torch.randn(B, C, H * f, W * f, generator=g)
    if dtype == torch.bfloat16:
        x, skip = x.bfloat16().float(), skip.bfloat16().float()
    want = F.conv_transpose2d(x.double(), w.double(), stride=f, padding=f // 2, groups=C)
    fmt = torch.channels_last if layout == "nhwc" else torch.contiguous_format
    xd = x.cuda().to(dtype).contiguous(memory_format=fmt)
    sd = skip.cuda().to(dtype).contiguous(memory_format=fmt)
    tol = 1e-2 if dtype == torch.bfloat16 else 1e-5
    got = idaup_upsample(xd, w.cuda(), f)
    assert got.shape == want.shape
    gate(got.float().cpu().numpy(), want.numpy(), tol, "up")
    got2 = idaup_upsample(xd, w.cuda(), f, sd)
    gate(got2.float().cpu().numpy(), (want + skip.double()).numpy(), tol, "up+skip")
    # tap-major weights (what the IDAUpsample module passes): on channels_last this is the 2x2-quad kernel, which must
    # give the same fp32 sums in the same order as the one-pixel kernel
    wt = w.reshape(C, -1).t().contiguous().cuda()
    got3 = idaup_upsample(xd, wt, f, sd, tap_major=True)
    assert torch.equal(got3, got2)
    got4 = idaup_upsample(xd, wt, f, tap_major=True)
    assert torch.equal(got4, got)


def test_idaup_upsample_odd_channels_generic_path():
    from centerpoly_b200.neck import idaup_upsample
    g = torch.Generator().manual_seed(3)
    x = torch.randn(1, 6, 5, 7, generator=g)
    w = torch.randn(6, 1, 4, 4, generator=g)
    want = F.conv_transpose2d(x, w, stride=2, padding=1, groups=6)
    got = idaup_upsample(x.cuda(), w.cuda(), 2)
    gate(got.cpu().numpy(), want.numpy(), 1e-5, "generic")


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16], ids=["fp32", "bf16"])
@pytest.mark.parametrize("shape", [(1, 64, 64, 16, 32), (1, 128, 64, 80, 96), (1, 64, 64, 128, 160)],
                         ids=["split9", "split3", "split1"])
def test_fused_deformconv_epilogue_vs_oracle(dtype, shape):
    """DCN -> BN(eval) -> ReLU fused in the epilogue, on the three tap-split schedules."""
    import centerpoly_b200 as cp
    from centerpoly_b200 import dcn_v2 as D
    from oracle import dcn_oracle
    B, Cin, Cout, H, W = shape
    case = make_dcn_case(77 + H, B, Cin, Cout, H, W, dtype=dtype)
    g = torch.Generator().manual_seed(5)
    scale = torch.rand(Cout, generator=g) + 0.5
    shift = torch.randn(Cout, generator=g) * 0.3
    a = {k: v.double().numpy() for k, v in case.items()}
    y = dcn_oracle.dcn_forward(a["x"], a["offset"], a["mask"], a["weight"], None)
    want = np.maximum(y * scale.double().numpy()[None, :, None, None] + shift.double().numpy()[None, :, None, None], 0.0)
    x = case["x"].cuda().to(dtype).contiguous(memory_format=torch.channels_last)
    got, _ = D._forward_impl(x, case["offset"].cuda().to(dtype), case["mask"].cuda().to(dtype), case["weight"].cuda(),
                             shift.cuda(), (1, 1), (1, 1), (1, 1), 1, False, out_scale=scale.cuda(), relu=True)
    gate(got.float().cpu().numpy(), want, 1e-2 if dtype == torch.bfloat16 else 1e-3, "fused")


def _neck_inputs(seed=0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    mk = lambda c, h, w: torch.randn(1, c, h, w, device="cuda", generator=g).contiguous(memory_format=torch.channels_last)
    return [mk(64, 32, 64), mk(128, 16, 32), mk(256, 8, 16)]


def test_rewrites_are_exact_per_module(monkeypatch):
    """Every rewrite of optimize_for_inference against the module it replaces, on identical inputs.  (The
    whole random-init network amplifies a 1e-6 perturbation to ~1e-2 through its 16 data-dependent samplers -
    tools/fusion_debug.py - so the module is the level at which exactness can be gated.)"""
    import copy
    from centerpoly_b200 import fuse
    from centerpoly_b200.harness.dla34 import PolydetDLA34, calibrate_random_init
    monkeypatch.setattr(torch.backends.cudnn, "allow_tf32", False)   # IEEE fp32 on both sides (convs and DCN)
    torch.manual_seed(317)
    net = PolydetDLA34().cuda().to(memory_format=torch.channels_last)
    calibrate_random_init(net, 128, 256)
    opt = copy.deepcopy(net)
    with torch.no_grad():
        stats = fuse.optimize_for_inference(opt)
        assert stats["deform_convs"] == 16 and stats["upsamplers"] == 8 and stats["conv_bn"] > 30
        # (a) DeformConv: DCN -> BN -> ReLU vs fused epilogue
        x = _neck_inputs(1)[1]
        gate(opt.ida_up.proj_1(x).cpu().numpy(), net.ida_up.proj_1(x).cpu().numpy(), 1e-5, "DeformConv")
        # (b) up_i + skip add
        a, skip = _neck_inputs(2)[0][:, :, :16, :32].contiguous(memory_format=torch.channels_last), _neck_inputs(3)[0]
        gate(opt.ida_up.up_1(a, skip).cpu().numpy(), (net.ida_up.up_1(a) + skip).cpu().numpy(), 1e-5, "IDAUp.up")
        # (c) whole IDAUp.forward (fused schedule) vs the reference schedule
        la, lb = _neck_inputs(4), _neck_inputs(4)
        net.ida_up(la, 0, 3)
        opt.ida_up(lb, 0, 3)
        gate(lb[2].cpu().numpy(), la[2].cpu().numpy(), 2e-3, "IDAUp.forward")
        # (d) conv+BN folding of the backbone: first levels (no data-dependent sampling -> no amplification)
        img = torch.randn(1, 3, 64, 128, device="cuda").contiguous(memory_format=torch.channels_last)
        fa, fb = net.base(img), opt.base(img)
        for i, (u, v) in enumerate(zip(fa, fb)):
            gate(v.cpu().numpy(), u.cpu().numpy(), 1e-4, f"base level {i}")
        # (e) the whole optimised network runs and stays close (loose: chaotic amplification, see docstring)
        want, got = net(img)[-1], opt(img)[-1]
        for k in want:
            gate(got[k].cpu().numpy(), want[k].cpu().numpy(), 0.25, k)
    with pytest.raises(RuntimeError):
        opt(img.requires_grad_(True))


@pytest.mark.parametrize("f", [2, 4, 8])
@pytest.mark.parametrize("dtype,C", [(torch.float32, 64), (torch.bfloat16, 64), (torch.bfloat16, 128)], ids=["fp32c64", "bf16c64", "bf16c128"])
def test_idaup_upsample_backward_vs_conv_transpose(f, dtype, C):
    """Data and weight gradients of the depthwise transposed convolution against torch's CPU autograd in float64."""
    from centerpoly_b200.neck import IDAUpsampleTrain
    g = torch.Generator().manual_seed(20 + f)
    B, H, W = 2, 9, 13
    x = torch.randn(B, C, H, W, generator=g)
    w = torch.randn(C, 1, 2 * f, 2 * f, generator=g) * 0.3
    gy = torch.randn(B, C, H * f, W * f, generator=g)
    if dtype == torch.bfloat16:
        x, gy = x.bfloat16().float(), gy.bfloat16().float()
    xr, wr = x.double().requires_grad_(True), w.double().requires_grad_(True)
    F.conv_transpose2d(xr, wr, stride=f, padding=f // 2, groups=C).backward(gy.double())
    deconv = torch.nn.ConvTranspose2d(C, C, 2 * f, stride=f, padding=f // 2, groups=C, bias=False)
    with torch.no_grad():
        deconv.weight.copy_(w)
    m = IDAUpsampleTrain(deconv.cuda())
    xd = x.cuda().to(dtype).contiguous(memory_format=torch.channels_last).requires_grad_(True)
    y = m(xd)
    y.backward(gy.cuda().to(dtype))
    tol = 1e-2 if dtype == torch.bfloat16 else 1e-5
    gate(xd.grad.float().cpu().numpy(), xr.grad.numpy(), tol, "grad_x")
    gate(m.weight.grad.float().cpu().numpy(), wr.grad.numpy(), 1e-4 if dtype == torch.float32 else 1e-3, "grad_w")
    assert list(m.state_dict().keys()) == ["weight"]
    # frozen weight: only the data gradient is produced
    m.weight.requires_grad_(False)
    m.weight.grad = None
    xd.grad = None
    m(xd).backward(gy.cuda().to(dtype))
    assert m.weight.grad is None
    gate(xd.grad.float().cpu().numpy(), xr.grad.numpy(), tol, "grad_x (frozen weight)")


def test_idaup_upsample_weight_gradient_is_bit_reproducible_at_training_size():
    """The weight gradient with a workspace is summed in a fixed order: two runs give the same bits (the atomics form, kept
    for callers without a workspace, varies at the 1e-7 level), and it matches fp64 at a training-size map."""
    from centerpoly_b200.neck import IDAUpsampleTrain
    torch.manual_seed(9)
    B, C, H, W, f = 4, 64, 32, 64, 4
    deconv = torch.nn.ConvTranspose2d(C, C, 2 * f, stride=f, padding=f // 2, groups=C, bias=False).cuda()
    m = IDAUpsampleTrain(deconv)
    x = torch.randn(B, C, H, W, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    gy = torch.randn(B, C, H * f, W * f, device="cuda").to(torch.bfloat16)
    grads = []
    for _ in range(2):
        m.weight.grad = None
        m(x).backward(gy)
        grads.append(m.weight.grad.clone())
    assert torch.equal(grads[0], grads[1])
    wr = m.weight.detach().double().requires_grad_(True)
    F.conv_transpose2d(x.double(), wr, stride=f, padding=f // 2, groups=C).backward(gy.double())
    gate(grads[0].double().cpu().numpy(), wr.grad.cpu().numpy(), 1e-4, "grad_w")
