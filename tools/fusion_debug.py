"""Which inference rewrite moves the heads, and by how much (fp32, cuDNN TF32 off)."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from centerpoly_b200 import fuse
from centerpoly_b200.harness.dla34 import PolydetDLA34, calibrate_random_init
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
def build():
    torch.manual_seed(317)
    net = PolydetDLA34().cuda().to(memory_format=torch.channels_last)
    calibrate_random_init(net, 128, 256)
    return net
x = torch.randn(1, 3, 128, 256, device="cuda", generator=torch.Generator(device="cuda").manual_seed(1)).contiguous(memory_format=torch.channels_last)
def heads(net):
    with torch.no_grad():
        return {k: v.float().cpu().numpy() for k, v in net(x)[-1].items()}
base = heads(build())
again = heads(build())
def report(name, got):
    print(name, {k: "%.2e" % (np.abs(got[k] - base[k]).max() / np.abs(base[k]).max()) for k in base})
report("rebuild", again)
import centerpoly_b200 as cp
for name, fn in (("deform", fuse.fuse_deform_convs), ("idaup", fuse.fuse_idaup), ("conv_bn", fuse.fold_conv_bn)):
    net = build()
    with torch.no_grad():
        fn(net)
    report(name, heads(net))
cp.set_dcn_path("simt")
net = build(); ref_simt = heads(net); report("simt-unfused", ref_simt)
with torch.no_grad():
    fuse.optimize_for_inference(net)
got = heads(net)
print("simt fused vs simt unfused", {k: "%.2e" % (np.abs(got[k] - ref_simt[k]).max() / np.abs(ref_simt[k]).max()) for k in base})
