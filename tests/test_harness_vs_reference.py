"""CPU, build container only: the measurement harness (centerpoly_b200/harness/dla34.py) is the same network
as the reference's DLASeg('dla34', ...) — identical state_dict keys/shapes and identical outputs when both
run the torchvision-backed oracle DCN on the same weights."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CODE = r'''
import sys, types, torch
sys.path.insert(0, %(root)r)
# route the reference's `from .DCNv2.dcn_v2 import DCN` to the torchvision oracle module
shim = types.ModuleType("models.networks.DCNv2.dcn_v2")
from oracle.tv_dcn import TVDCN
shim.DCN = TVDCN
import importlib
sys.path.insert(1, "/root/reference/src/lib")
import models.networks
pkg = types.ModuleType("models.networks.DCNv2"); pkg.__path__ = []
sys.modules["models.networks.DCNv2"] = pkg
sys.modules["models.networks.DCNv2.dcn_v2"] = shim
from models.networks.pose_dla_dcn import DLASeg
from centerpoly_b200.harness.dla34 import PolydetDLA34, calibrate_random_init
heads = {"hm": 8, "poly": 64, "pseudo_depth": 1, "reg": 2}
torch.manual_seed(317)
ref = DLASeg("dla34", heads, pretrained=False, down_ratio=4, final_kernel=1, last_level=5, head_conv=256)
ours = PolydetDLA34(heads, dcn_cls=TVDCN)
a, b = ref.state_dict(), ours.state_dict()
assert list(a.keys()) == list(b.keys()), set(a) ^ set(b)
assert all(a[k].shape == b[k].shape for k in a)
# bilinear deconv init must agree before loading
for k in a:
    if ".up_" in k:
        assert torch.equal(a[k], b[k]), k
calibrate_random_init(ref, 64, 96)
ours.load_state_dict(ref.state_dict())
ours.eval()
x = torch.randn(1, 3, 64, 96)
with torch.no_grad():
    ya, yb = ref(x)[-1], ours(x)[-1]
for h in heads:
    assert torch.equal(ya[h], yb[h]), h
print("harness == reference on", sum(p.numel() for p in ours.parameters()), "params")
'''


@pytest.mark.skipif(not os.path.isdir("/root/reference/src/lib"), reason="reference tree only exists in the build container")
def test_harness_matches_reference_dlaseg():
    out = subprocess.run([sys.executable, "-c", CODE % {"root": ROOT}], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-3000:]
    assert "harness == reference on 19821291 params" in out.stdout, out.stdout
