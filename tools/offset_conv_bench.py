"""Offset/mask producer timing per neck layer shape (CUDA events, L2 flushed): python tools/offset_conv_bench.py"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import centerpoly_b200 as cp  # noqa: E402

flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
tot = 0.0
for (cin, h, w, n) in [(64, 256, 512, 5), (128, 128, 256, 4), (256, 64, 128, 4), (512, 32, 64, 1), (128, 256, 512, 1), (256, 128, 256, 1)]:
    m = cp.DCN(cin, 64, (3, 3), 1, 1).cuda().to(torch.bfloat16).to(memory_format=torch.channels_last).eval()
    torch.nn.init.normal_(m.conv_offset_mask.weight, std=0.05); torch.nn.init.normal_(m.conv_offset_mask.bias, std=0.1)
    x = torch.randn(1, cin, h, w, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    with torch.no_grad():
        for _ in range(3):
            om = m.offset_conv(x)
        ref = torch.nn.functional.conv2d(x.float(), m.conv_offset_mask.weight.float(), m.conv_offset_mask.bias.float(), padding=1)
        err = (om[:, :27].float() - ref).abs().max().item() / ref.abs().max().item()
        ts = []
        for _ in range(7):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); m.offset_conv(x); e1.record(); torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1) * 1000)
        # cuDNN for comparison
        tc = []
        for _ in range(7):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); m.conv_offset_mask(x); e1.record(); torch.cuda.synchronize()
            tc.append(e0.elapsed_time(e1) * 1000)
    t = sorted(ts)[3]
    tot += t * n
    print(f"Cin {cin:4d} {h}x{w}: ours {t:7.1f} us  cudnn {sorted(tc)[3]:7.1f} us  rel err {err:.2e}  (x{n} per frame)")
print(f"per frame (sum over the 16 sites, flushed): {tot:.0f} us")
