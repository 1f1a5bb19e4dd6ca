"""Backward of one DCN layer through the C ABI, N times (for `ncu --metrics gpu__time_duration.sum`: per-kernel split of
pack / dgrad / bbox / col2im / wgrad / reduce / bias).  python tools/bwd_breakdown.py [Cin Cout H W B dtype iters]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from centerpoly_b200 import dcn_v2 as D  # noqa: E402

a = sys.argv[1:]
Cin, Cout, H, W, B = (int(v) for v in (a[:5] if len(a) >= 5 else (64, 64, 256, 512, 1)))
dtype = torch.float32 if len(a) > 5 and a[5] == "fp32" else torch.bfloat16
iters = int(a[6]) if len(a) > 6 else 3
g = torch.Generator(device="cuda").manual_seed(317)
x = torch.randn(B, Cin, H, W, device="cuda", generator=g).to(dtype).contiguous(memory_format=torch.channels_last)
om = torch.randn(B, 27, H, W, device="cuda", generator=g)
om[:, :18] *= 2.0
om = om.to(dtype).contiguous(memory_format=torch.channels_last)
gout = torch.randn(B, Cout, H, W, device="cuda", generator=g).to(dtype).contiguous(memory_format=torch.channels_last)
w = (torch.rand(Cout, Cin, 3, 3, device="cuda", generator=g) - 0.5) * 0.1
for i in range(iters):
    D._forward_impl(x, om[:, :18], om[:, 18:], w, None, (1, 1), (1, 1), (1, 1), 1, True)
    D._backward_impl(x, om[:, :18], om[:, 18:], w, gout, (1, 1), (1, 1), (1, 1), 1, True, (True,) * 5)
torch.cuda.synchronize()
ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
ev[0].record()
for i in range(10):
    D._backward_impl(x, om[:, :18], om[:, 18:], w, gout, (1, 1), (1, 1), (1, 1), 1, True, (True,) * 5)
ev[1].record()
torch.cuda.synchronize()
print("bwd ms", ev[0].elapsed_time(ev[1]) / 10)
