"""Stem convolution and 2x2 max-pool kernels (csrc/backbone.cu) against plain PyTorch fp32 references of the same op."""
import pytest
import torch
import torch.nn.functional as F

from util import gate

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("B,H,W", [(1, 64, 96), (2, 37, 53), (1, 128, 256)])
@pytest.mark.parametrize("in_dtype", [torch.float32, torch.bfloat16], ids=["x_fp32", "x_bf16"])
@pytest.mark.parametrize("fmt", [torch.contiguous_format, torch.channels_last], ids=["nchw", "nhwc"])
def test_stem_matches_torch(B, H, W, in_dtype, fmt):
    from centerpoly_b200.backbone import StemConv7x7
    g = torch.Generator().manual_seed(B * 1000 + H)
    conv = torch.nn.Conv2d(3, 16, 7, 1, 3, bias=True)
    with torch.no_grad():
        conv.weight.copy_(torch.randn(16, 3, 7, 7, generator=g) * 0.1)
        conv.bias.copy_(torch.randn(16, generator=g))
    x = torch.randn(B, 3, H, W, generator=g)
    # the reference sees exactly the bf16-rounded operands (what a bf16 model feeds its tensor cores), fp32 accumulation
    want = F.relu(F.conv2d(x.to(in_dtype).bfloat16().float(), conv.weight.bfloat16().float(), conv.bias.bfloat16().float(), 1, 3))
    m = StemConv7x7(conv.cuda().bfloat16()).cuda()
    with torch.no_grad():
        y = m(x.to(in_dtype).cuda().contiguous(memory_format=fmt))
    assert y.dtype == torch.bfloat16 and y.is_contiguous(memory_format=torch.channels_last) and tuple(y.shape) == (B, 16, H, W)
    gate(y.float().cpu().numpy(), want.detach().numpy(), 1e-2, "stem bf16")      # bf16 output rounding: 2^-9 relative
    assert list(m.state_dict().keys()) == ["conv.weight", "conv.bias"]


def test_stem_fp32_output_and_cache_invalidation():
    from centerpoly_b200.backbone import StemConv7x7
    conv = torch.nn.Conv2d(3, 16, 7, 1, 3, bias=False).cuda()
    x = torch.randn(1, 3, 40, 72, device="cuda")
    m = StemConv7x7(conv, relu=False)
    with torch.no_grad():
        y = m(x)
        want = F.conv2d(x.bfloat16().float(), conv.weight.bfloat16().float(), None, 1, 3)
        gate(y.cpu().numpy(), want.cpu().numpy(), 1e-3, "stem fp32 out (bf16 operands)")
        conv.weight.mul_(2.0)                       # in-place update must invalidate the packed-weight cache
        gate(m(x).cpu().numpy(), (2 * want).cpu().numpy(), 1e-3, "stem after weight update")
    with pytest.raises(RuntimeError):
        m(x.requires_grad_(True))


@pytest.mark.parametrize("dtype,C", [(torch.bfloat16, 32), (torch.bfloat16, 8), (torch.float32, 16), (torch.bfloat16, 128)])
@pytest.mark.parametrize("H,W", [(64, 128), (33, 51)])
def test_maxpool_bit_exact(dtype, C, H, W):
    from centerpoly_b200.backbone import MaxPool2x2
    x = torch.randn(2, C, H, W, device="cuda").to(dtype).contiguous(memory_format=torch.channels_last)
    x[0, 0, 4, 6] = float("nan")
    with torch.no_grad():
        y = MaxPool2x2()(x)
    want = F.max_pool2d(x.float(), 2, 2).to(dtype)
    assert y.is_contiguous(memory_format=torch.channels_last)
    assert torch.equal(torch.nan_to_num(y.float(), nan=-7.0), torch.nan_to_num(want.float(), nan=-7.0))
    assert torch.isnan(y[0, 0, 2, 3])


def test_engine_uses_the_kernels():
    from centerpoly_b200 import _lib
    from centerpoly_b200.harness.engine import PolydetEngine
    eng = PolydetEngine(height=128, width=256, dtype="bf16")
    assert eng.fusion["stem"] == 1 and eng.fusion["maxpool"] >= 4, eng.fusion


@pytest.mark.parametrize("cout", [1, 2, 8, 64])
def test_head_1x1_gemm_matches_conv(cout):
    from centerpoly_b200.fuse import _Conv1x1Gemm
    torch.manual_seed(cout)
    conv = torch.nn.Conv2d(256, cout, 1, bias=True).cuda().bfloat16()
    x = torch.randn(2, 256, 24, 40, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last)
    m = _Conv1x1Gemm(conv)
    with torch.no_grad():
        y = m(x)
        want = F.conv2d(x.float(), conv.weight.float(), conv.bias.float())
    assert tuple(y.shape) == (2, cout, 24, 40) and y.dtype == torch.bfloat16
    gate(y.float().cpu().numpy(), want.cpu().numpy(), 1e-2, "1x1 head as GEMM")
    assert list(m.state_dict().keys()) == ["conv.weight", "conv.bias"]
