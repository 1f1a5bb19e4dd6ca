"""Debug (GPU): (1) fp32 tensor-core wgrad vs the CUDA-core path on a small case; (2) full-size goffset outliers vs the golden."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import centerpoly_b200 as cp  # noqa: E402
from util import make_dcn_case  # noqa: E402


def run(case, path, dtype):
    cp.set_dcn_path(path)
    try:
        cl = torch.channels_last
        x = case["x"].cuda().to(dtype).contiguous(memory_format=cl).requires_grad_(True)
        off = case["offset"].cuda().to(dtype).requires_grad_(True)
        msk = case["mask"].cuda().to(dtype).requires_grad_(True)
        w = case["weight"].cuda().requires_grad_(True)
        b = case["bias"].cuda().requires_grad_(True)
        y = cp.dcn_v2_conv(x, off, msk, w, b, 1, 1, 1, 1)
        y.backward(case["gout"].cuda().to(dtype).contiguous(memory_format=cl))
        torch.cuda.synchronize()
        return dict(y=y.detach().float().cpu(), gx=x.grad.float().cpu(), goffset=off.grad.float().cpu(), gmask=msk.grad.float().cpu(),
                    gw=w.grad.float().cpu(), gb=b.grad.float().cpu())
    finally:
        cp.set_dcn_path("auto")


print("== fp32 wgrad, tensor-core vs CUDA-core")
for shape in [(1, 64, 64, 12, 20), (1, 32, 64, 12, 20), (1, 64, 128, 12, 20), (1, 64, 32, 16, 16)]:
    case = make_dcn_case(99, *shape)
    try:
        a, b = run(case, "tcgen05", torch.float32), run(case, "simt", torch.float32)
    except Exception as e:
        print(shape, "ERR", str(e)[:120])
        continue
    gw_a, gw_b = a["gw"], b["gw"]
    print(shape, "gw tc norm %.4g simt norm %.4g  nonzero frac %.3f  rel err %.3g" % (gw_a.norm(), gw_b.norm(), (gw_a != 0).float().mean(),
          (gw_a - gw_b).norm() / gw_b.norm()))
    # per tap / per channel-chunk error pattern
    d = (gw_a - gw_b)
    print("   per tap rel err", [round(float(d[:, :, i // 3, i % 3].norm() / gw_b[:, :, i // 3, i % 3].norm()), 3) for i in range(9)])
    print("   per cin-chunk(32) rel err", [round(float(d[:, c:c + 32].norm() / gw_b[:, c:c + 32].norm()), 3) for c in range(0, shape[1], 32)])
    print("   per cout-chunk(32) rel err", [round(float(d[c:c + 32].norm() / gw_b[c:c + 32].norm()), 3) for c in range(0, shape[2], 32)])
    for name in ("gx", "goffset", "gmask", "gb"):
        print("   %s rel err tc vs simt %.3g" % (name, (a[name] - b[name]).norm() / b[name].norm()))

print("== full-size goffset outliers (64->64 @256x512, fp32 inputs)")
g = np.load(os.path.join(ROOT, "tests", "golden", "dcnfull_64x64_256x512_b1_f32.npz"))
B, Cin, Cout, H, W = (int(v) for v in g["shape"])
case = make_dcn_case(int(g["seed"]), B, Cin, Cout, H, W)
flat = torch.from_numpy(g["positions"])
tk = lambda t: t.permute(0, 2, 3, 1).reshape(B * H * W, -1)[flat]
a, b = run(case, "tcgen05", torch.float32), run(case, "simt", torch.float32)
for name in ("y", "gx", "goffset", "gmask"):
    want = torch.from_numpy(g[name]).double()
    for tag, got in (("tc", a), ("simt", b)):
        e = (tk(got[name]).double() - want).abs()
        print(name, tag, "max abs err %.4g at sampled pos (max|want| %.4g)" % (e.max(), float(g[name + "_max"])))
    e = (a[name] - b[name]).abs()
    print(name, "tc vs simt full tensor max abs %.4g" % e.max())
want = torch.from_numpy(g["goffset"]).double()
e = (tk(a["goffset"]).double() - want).abs()
off_s, msk_s = tk(case["offset"]), tk(case["mask"])
idx = torch.nonzero(e > 1e-2 * float(g["goffset_max"]))
print("outliers:", idx.shape[0])
for i, c in idx[:12].tolist():
    p = int(flat[i]); yy, xx = (p % (H * W)) // W, p % W
    k = c // 2
    py = yy - 1 + k // 3 + float(off_s[i, 2 * k]); px = xx - 1 + k % 3 + float(off_s[i, 2 * k + 1])
    print("  pix (%d,%d) ch %d tap %d: sample pos (%.4f, %.4f)  got tc %.5g simt %.5g want %.5g" % (yy, xx, c, k, py, px, float(tk(a["goffset"])[i, c]),
          float(tk(b["goffset"])[i, c]), float(want[i, c])))
