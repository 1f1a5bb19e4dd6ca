"""Debug aid: per-tile error map of the tcgen05 DCN forward against the SIMT path (same device, no oracle needed).
Usage: python tools/tc_err_map.py B Cin Cout H W [bf16|fp32]"""
import os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import centerpoly_b200 as cp
from util import make_dcn_case

B, Cin, Cout, H, W = map(int, sys.argv[1:6])
dtype = torch.bfloat16 if (len(sys.argv) > 6 and sys.argv[6] == "bf16") else torch.float32
case = make_dcn_case(317 + Cin + Cout, B, Cin, Cout, H, W, dtype=dtype)
x = case["x"].cuda().to(dtype).contiguous(memory_format=torch.channels_last)
args = (case["offset"].cuda().to(dtype), case["mask"].cuda().to(dtype), case["weight"].cuda(), case["bias"].cuda(), 1, 1, 1, 1)
outs = {}
for path in ("simt", "tcgen05", "tcgen05"):
    cp.set_dcn_path(path)
    y = cp.dcn_v2_conv(x, *args).float()
    torch.cuda.synchronize()
    outs.setdefault(path, []).append(y)
ref, a, b = outs["simt"][0], outs["tcgen05"][0], outs["tcgen05"][1]
print("deterministic:", bool((a == b).all()), " max|ref|", float(ref.abs().max()))
err = (a - ref).abs().amax(dim=1)[0]          # (H, W)
th, tw = (H + 7) // 8, (W + 15) // 16
pad = torch.zeros(th * 8, tw * 16, device=err.device); pad[:H, :W] = err
tiles = pad.view(th, 8, tw, 16).amax(dim=(1, 3)).cpu().numpy()
bad = tiles > 1e-2 * float(ref.abs().max())
print("bad tiles:", int(bad.sum()), "of", bad.size)
print("bad tile ids (row-major):", np.flatnonzero(bad.reshape(-1))[:60])
rows = pad.view(th, 8, tw, 16).permute(0, 2, 1, 3).reshape(th * tw, 128)
if bad.any():
    t0 = int(np.flatnonzero(bad.reshape(-1))[0])
    print("first bad tile", t0, "per-pixel err:", np.round(rows[t0].cpu().numpy()[:32], 3))
    chan_err = (a - ref).abs()[0].amax(dim=(1, 2)).cpu().numpy()
    print("per-channel max err:", np.round(chan_err[:16], 3))
