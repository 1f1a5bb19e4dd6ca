"""CPU: the C-ABI library loads and exports every symbol include/cpoly.h declares; host-side logic."""
import ctypes
import os
import re
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    from centerpoly_b200 import _lib
    header = open(os.path.join(ROOT, "include", "cpoly.h")).read()
    declared = set(re.findall(r"\b(cpoly_[a-z0-9_]+)\s*\(", header))
    assert declared == set(_lib.EXPORTED_SYMBOLS), declared ^ set(_lib.EXPORTED_SYMBOLS)
    L = ctypes.CDLL(_lib.LIB_PATH)
    for sym in declared:
        assert hasattr(L, sym), sym
    assert _lib.lib().cpoly_abi_version() == 10


def test_struct_layouts_match_header(tmp_path):
    """sizeof of every ctypes mirror == sizeof of the C struct, as compiled by gcc from include/cpoly.h."""
    import subprocess
    from centerpoly_b200 import _lib
    names = {"cpoly_tensor4": _lib.Tensor4, "cpoly_dcn_fwd_args": _lib.DcnFwdArgs, "cpoly_dcn_bwd_args": _lib.DcnBwdArgs,
             "cpoly_upsample_args": _lib.UpsampleArgs, "cpoly_decode_args": _lib.DecodeArgs,
             "cpoly_raster_args": _lib.RasterArgs, "cpoly_poly_iou_args": _lib.PolyIouArgs, "cpoly_stem_args": _lib.StemArgs,
             "cpoly_maxpool_args": _lib.MaxPoolArgs, "cpoly_upsample_bwd_args": _lib.UpsampleBwdArgs,
             "cpoly_offset_conv_args": _lib.OffsetConvArgs}
    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include "cpoly.h"\nint main(void){' +
                   "".join(f'printf("{n} %zu\\n", sizeof({n}));' for n in names) + "return 0;}\n")
    exe = tmp_path / "sz"
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    out = dict(l.split() for l in subprocess.check_output([str(exe)], text=True).splitlines())
    for n, cls in names.items():
        assert ctypes.sizeof(cls) == int(out[n]), (n, ctypes.sizeof(cls), out[n])
    assert ctypes.sizeof(_lib.Tensor4) == 40


def test_size_queries_need_no_gpu():
    from centerpoly_b200 import _lib
    L = _lib.lib()
    assert L.cpoly_dcn_packed_weight_bytes(64, 64, 3, 3, _lib.BF16) == 9 * 64 * 64 * 2
    assert L.cpoly_dcn_packed_weight_bytes(64, 64, 3, 3, _lib.F32) == 9 * 64 * 64 * 4
    assert L.cpoly_dcn_packed_weight_bytes(48, 64, 3, 3, _lib.BF16) == 0      # Cin not a multiple of 64
    assert L.cpoly_dcn_packed_weight_bytes(64, 64, 5, 5, _lib.BF16) == 0
    assert L.cpoly_decode_workspace_bytes(1, 8, 256, 512, 100) >= 256
    assert L.cpoly_offset_conv_packed_bytes(128) == 2 * 9 * 32 * 64 * 2 and L.cpoly_offset_conv_packed_bytes(48) == 0


def test_module_api_and_state_dict():
    import centerpoly_b200 as cp
    m = cp.DCN(128, 64, kernel_size=(3, 3), stride=1, padding=1, dilation=1, deformable_groups=1)
    sd = m.state_dict()
    assert {k: tuple(v.shape) for k, v in sd.items()} == {
        "weight": (64, 128, 3, 3), "bias": (64,),
        "conv_offset_mask.weight": (27, 128, 3, 3), "conv_offset_mask.bias": (27,)}
    assert float(m.conv_offset_mask.weight.abs().max()) == 0.0 and float(m.bias.abs().max()) == 0.0
    bound = 1.0 / (128 * 9) ** 0.5
    assert float(m.weight.abs().max()) <= bound
    assert all(p.requires_grad for p in m.parameters())
    m2 = cp.DCNv2(8, 8, 3, 1, 1)
    assert set(m2.state_dict().keys()) == {"weight", "bias"}


def test_no_cpu_fallback():
    import centerpoly_b200 as cp
    m = cp.DCN(64, 64, (3, 3), 1, 1)
    with pytest.raises(RuntimeError, match="no CPU path"):
        m(torch.randn(1, 64, 8, 8))
    z = torch.zeros(1, 2, 8, 8)
    with pytest.raises(RuntimeError, match="no CPU path"):
        cp.polydet_decode(z, torch.zeros(1, 8, 8, 8), z[:, :1], K=4)


def test_product_never_imports_oracle():
    """The oracle is test infrastructure: nothing under centerpoly_b200/ may import or load it."""
    pkg = os.path.join(ROOT, "centerpoly_b200")
    pat = re.compile(r"^\s*(from|import)\s+oracle\b|libdcn_oracle|oracle/", re.M)
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(".py"):
                assert not pat.search(open(os.path.join(dirpath, f)).read()), f


@pytest.mark.skipif(not os.path.isdir("/root/reference/src/lib"), reason="reference tree only exists in the build container")
def test_overlay_resolves_reference_import():
    """The reference's own `from .DCNv2.dcn_v2 import DCN` (pose_dla_dcn.py:16) must pick up our module."""
    import subprocess
    code = (
        "import sys; sys.path.insert(0, %r)\n"
        "import centerpoly_b200.install as inst; inst.install_dcn_overlay('/root/reference/src/lib')\n"
        "from models.networks import pose_dla_dcn as m\n"
        "import centerpoly_b200\n"
        "assert m.DCN is centerpoly_b200.DCN, m.DCN\n"
        "net = m.DLASeg('dla34', {'hm': 8, 'poly': 64, 'pseudo_depth': 1, 'reg': 2}, pretrained=False, down_ratio=4, final_kernel=1, last_level=5, head_conv=256)\n"
        "n = sum(isinstance(x, centerpoly_b200.DCN) for x in net.modules())\n"
        "assert n == 16, n\n"
        "print('ok', n)\n" % ROOT)
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    assert "ok 16" in out.stdout


def test_poly_loss_mirror_rejects_what_it_does_not_build():
    """Host-side contract of the PolyLoss mirror (no GPU needed): unsupported options raise at construction, CPU
    tensors raise instead of falling back."""
    import pytest
    import torch
    from centerpoly_b200.poly_loss import PolyLoss, poly_iou_loss

    class Opt:
        poly_loss, rep, poly_order = "iou", "polar", False
    PolyLoss(Opt())
    PolyLoss(type("O", (), dict(poly_loss="iou", rep="polar", poly_order=True)))      # order arm is built (tests/test_polyorder.py)
    for attr, val in (("poly_loss", "relu"), ("poly_loss", "bce"), ("rep", "cartesian")):
        o = type("O", (), dict(poly_loss="iou", rep="polar", poly_order=False))
        setattr(o, attr, val)
        with pytest.raises(NotImplementedError):
            PolyLoss(o)
    with pytest.raises(RuntimeError):
        poly_iou_loss(torch.zeros(1, 16, 4, 4), torch.ones(1, 2), torch.zeros(1, 2, dtype=torch.int64), torch.zeros(1, 2, 16))


def test_training_upsampler_swap_keeps_the_state_dict():
    """use_training_upsamplers (host logic, no GPU): same keys and tensors before and after; a non-CUDA input raises
    like every other entry point (no PyTorch fallback in the product)."""
    import torch
    from centerpoly_b200.harness.dla34 import PolydetDLA34
    from centerpoly_b200.neck import IDAUpsampleTrain, use_training_upsamplers
    net = PolydetDLA34()
    before = {k: v.clone() for k, v in net.state_dict().items()}
    assert use_training_upsamplers(net) == 8
    after = net.state_dict()
    assert list(after.keys()) == list(before.keys()) and all(torch.equal(after[k], before[k]) for k in before)
    up = next(m for m in net.modules() if isinstance(m, IDAUpsampleTrain))
    with pytest.raises(RuntimeError):
        up(torch.randn(1, up.weight.shape[0], 5, 7))
