"""IDAUp up-sampling kernel timing at the frame's shapes (CUDA events, L2 flushed between calls)."""
import os, sys, json
import torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from centerpoly_b200.neck import idaup_upsample
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for (C, H, W, f) in [(64, 128, 256, 2), (64, 64, 128, 4), (64, 32, 64, 8), (128, 64, 128, 2), (256, 32, 64, 2)]:
    x = torch.randn(1, C, H, W, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last)
    s = torch.randn(1, C, H * f, W * f, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last)
    wt = torch.randn(4 * f * f, C, device="cuda")
    ts = []
    for _ in range(12):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); idaup_upsample(x, wt, f, s, tap_major=True); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1000)
    ts = sorted(ts[2:])
    mb = (C * H * W + 2 * C * H * f * W * f) * 2 / 1e6
    print(json.dumps(dict(C=C, H=H, W=W, f=f, us=round(ts[len(ts) // 2], 1), mb=round(mb, 1), gbs=round(mb / ts[len(ts) // 2] * 1e3)), ))
