"""GPU parity of the DCN module's offset / mask producer on our kernel (csrc/offset_conv.cu):
``DCN.offset_conv(x)`` == ``conv_offset_mask(x)`` (regular 3x3 conv Cin -> 27, bias), computed from bf16 activations and
bf16-rounded weights with fp32 accumulation; and the module's inference forward through it == the C oracle fed the
same offsets."""
import numpy as np
import pytest
import torch

from util import gate

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("shape", [(1, 64, 16, 32), (2, 128, 9, 21), (1, 256, 24, 40), (1, 512, 8, 16), (3, 64, 5, 7), (1, 64, 70, 90)],
                         ids=lambda s: "x".join(map(str, s)))
def test_offset_conv_matches_conv2d(shape):
    import centerpoly_b200 as cp
    B, C, H, W = shape
    torch.manual_seed(C + H)
    m = cp.DCN(C, 64, (3, 3), 1, 1).cuda()
    torch.nn.init.normal_(m.conv_offset_mask.weight, std=0.05)
    torch.nn.init.normal_(m.conv_offset_mask.bias, std=1.0)
    x = torch.randn(B, C, H, W, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last)
    om = m.offset_conv(x)
    assert om is not None and om.shape == (B, 32, H, W) and om.dtype == torch.float32
    want = torch.nn.functional.conv2d(x.double(), m.conv_offset_mask.weight.detach().bfloat16().double(),
                                      m.conv_offset_mask.bias.detach().double(), padding=1)
    gate(om[:, :27].double().cpu().numpy(), want.cpu().numpy(), 1e-4, "om")
    assert float(om[:, 27:].abs().max()) == 0.0
    # the weights the reference would use are fp32: the bf16 rounding of the weight image stays inside the bf16 gate
    full = torch.nn.functional.conv2d(x.double(), m.conv_offset_mask.weight.detach().double(), m.conv_offset_mask.bias.detach().double(), padding=1)
    gate(om[:, :27].double().cpu().numpy(), full.cpu().numpy(), 1e-2, "om vs fp32 weights")


def test_module_inference_forward_uses_it_and_matches_oracle():
    import centerpoly_b200 as cp
    from centerpoly_b200 import _lib
    from oracle import dcn_oracle
    torch.manual_seed(9)
    m = cp.DCN(64, 64, (3, 3), 1, 1).cuda().to(memory_format=torch.channels_last)
    torch.nn.init.normal_(m.conv_offset_mask.weight, std=0.05)
    torch.nn.init.normal_(m.conv_offset_mask.bias, std=1.0)
    x = torch.randn(2, 64, 20, 36, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last)
    mb = m.bfloat16()
    before = _lib.LAUNCHES["offset_conv"]
    with torch.no_grad():
        y = mb(x)
        om = mb.offset_conv(x)
    assert _lib.LAUNCHES["offset_conv"] == before + 2
    omc = om.float().cpu()
    want = dcn_oracle.dcn_forward(x.float().cpu().double().numpy(), omc[:, :18].double().numpy(), torch.sigmoid(omc[:, 18:27].double()).numpy(),
                                  mb.weight.detach().float().cpu().double().numpy(), mb.bias.detach().float().cpu().double().numpy())
    gate(y.float().cpu().numpy(), want, 1e-2, "DCN(x) through the own offset conv")
    # switching it off gives the cuDNN producer: same result within the bf16 gate (offsets rounded to bf16 there)
    cp.dcn_v2.set_offset_conv(False)
    try:
        with torch.no_grad():
            y2 = mb(x)
    finally:
        cp.dcn_v2.set_offset_conv(True)
    assert _lib.LAUNCHES["offset_conv"] == before + 2
    gate(y2.float().cpu().numpy(), y.float().cpu().numpy(), 5e-2, "cuDNN producer vs own producer")


@pytest.mark.parametrize("shape", [(1, 64, 64, 16, 32), (1, 128, 64, 72, 100), (1, 512, 256, 8, 16), (3, 64, 64, 5, 7), (2, 128, 128, 24, 40)],
                         ids=lambda s: "x".join(map(str, s)))
@pytest.mark.parametrize("om_dtype", [torch.float32, torch.bfloat16], ids=["om-fp32", "om-bf16"])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32], ids=["bf16", "fp32"])
def test_fused_om_tensor_staging_vs_oracle(shape, om_dtype, dtype):
    """Offsets + mask logits as ONE dense 32-channel NHWC tensor (what the producer kernels write): the forward kernel's
    coalesced staging path (dcn_tc.cu, kOmFast) against the C oracle, all tap-split schedules, ragged tiles."""
    from centerpoly_b200 import dcn_v2 as D
    from oracle import dcn_oracle
    from util import make_dcn_case
    B, Cin, Cout, H, W = shape
    case = make_dcn_case(41 + Cin + H, B, Cin, Cout, H, W, dtype=dtype)
    g = torch.Generator().manual_seed(5)
    om = torch.zeros(B, 32, H, W)
    om[:, :18] = case["offset"]
    om[:, 18:27] = torch.randn(B, 9, H, W, generator=g)
    om[:, 27:] = 7.0                       # padding channels must be ignored
    om = om.to(om_dtype).cuda().contiguous(memory_format=torch.channels_last)
    x = case["x"].to(dtype).cuda().contiguous(memory_format=torch.channels_last)
    y, _ = D._forward_impl(x, om[:, :18], om[:, 18:27], case["weight"].cuda(), case["bias"].cuda(), (1, 1), (1, 1), (1, 1), 1, True)
    omf = om.float().cpu().double()
    want = dcn_oracle.dcn_forward(x.float().cpu().double().numpy(), omf[:, :18].numpy(), torch.sigmoid(omf[:, 18:27]).numpy(),
                                  case["weight"].double().numpy(), case["bias"].double().numpy())
    gate(y.float().cpu().numpy(), want, 1e-2 if dtype == torch.bfloat16 else 1e-3, "y")
