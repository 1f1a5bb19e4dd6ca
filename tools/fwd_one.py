"""A few forward launches of one DCN layer through the C ABI (for ncu).  python tools/fwd_one.py [Cin Cout H W B dtype fused_om]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from centerpoly_b200 import dcn_v2 as D  # noqa: E402

a = sys.argv[1:]
Cin, Cout, H, W, B = (int(v) for v in (a[:5] if len(a) >= 5 else (64, 64, 256, 512, 1)))
dtype = torch.float32 if len(a) > 5 and a[5] == "fp32" else torch.bfloat16
fused = len(a) > 6 and a[6] == "1"
g = torch.Generator(device="cuda").manual_seed(317)
x = torch.randn(B, Cin, H, W, device="cuda", generator=g).to(dtype).contiguous(memory_format=torch.channels_last)
om = torch.randn(B, 32 if fused else 27, H, W, device="cuda", generator=g)
om[:, :18] *= float(a[7]) if len(a) > 7 else 2.0
om = (om if fused else om.to(dtype)).contiguous(memory_format=torch.channels_last)
w = (torch.rand(Cout, Cin, 3, 3, device="cuda", generator=g) - 0.5) * 0.1
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
ev = []
for i in range(6):
    flush.zero_()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    D._forward_impl(x, om[:, :18], om[:, 18:27], w, None, (1, 1), (1, 1), (1, 1), 1, True)
    e1.record()
    ev.append((e0, e1))
torch.cuda.synchronize()
print("fwd us", sorted(a.elapsed_time(b) * 1000 for a, b in ev)[len(ev) // 2])

if os.environ.get("CPOLY_TC_TRACE"):
    sys.exit(0)
# back-to-back launches replayed from a CUDA graph: no host launch gap inside the timed region (x stays in L2)
N = 20
call = lambda: D._forward_impl(x, om[:, :18], om[:, 18:27], w, None, (1, 1), (1, 1), (1, 1), 1, True)
side = torch.cuda.Stream()
with torch.cuda.stream(side):
    call(); call()
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr, stream=side):
        for _ in range(N):
            y = call()
torch.cuda.synchronize()
ts = []
for i in range(5):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); gr.replay(); e1.record(); torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1) * 1000 / N)
print("fwd us", sorted(a.elapsed_time(b) * 1000 for a, b in ev)[len(ev) // 2], "graph us/launch", round(sorted(ts)[2], 2))
