"""DCN layer sweep at the IDAUp shapes of a 2048x1024 frame (BASELINE.json configs[2]).
CUDA-event timing, L2 flushed between iterations.  Usage: python tools/layer_sweep.py [--bwd] [--batch B] [--simt]"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import centerpoly_b200 as cp  # noqa: E402

SHAPES = [(64, 64, 256, 512), (128, 64, 128, 256), (128, 128, 128, 256), (256, 64, 64, 128),
          (256, 128, 64, 128), (256, 256, 64, 128), (512, 256, 32, 64)]
COUNT_PER_FRAME = {(64, 64): 5, (128, 64): 4, (128, 128): 2, (256, 64): 1, (256, 128): 2, (256, 256): 1, (512, 256): 1}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d["hbm_gbs"], d["bf16_tflops"], "measured"
    return 6650.0, 1590.0, "fallback"


def time_fn(fn, iters, flush):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    evs = []
    for _ in range(iters):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        evs.append((a, b))
    torch.cuda.synchronize()
    ts = sorted(a.elapsed_time(b) for a, b in evs)
    return ts[len(ts) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--bwd", action="store_true")
    ap.add_argument("--simt", action="store_true")
    ap.add_argument("--batch", type=int, default=1)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--dtypes", default="bf16,fp32")
    ap.add_argument("--offset-std", type=float, default=2.0, help="std of the synthetic offsets in pixels (SURVEY 8d: 2)")
    ap.add_argument("--shapes", type=int, default=len(SHAPES), help="only the first N shape classes")
    args = ap.parse_args()
    hbm, tf, src = peaks()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    B = args.batch
    rows = []
    for dname in args.dtypes.split(","):
        dtype = torch.bfloat16 if dname == "bf16" else torch.float32
        es = 2 if dname == "bf16" else 4
        total_ms = 0.0
        for (Cin, Cout, H, W) in SHAPES[:args.shapes]:
            g = torch.Generator(device="cuda").manual_seed(317)
            x = torch.randn(B, Cin, H, W, device="cuda", generator=g).to(dtype).contiguous(memory_format=torch.channels_last)
            om = torch.randn(B, 27, H, W, device="cuda", generator=g)
            om[:, :18] *= args.offset_std
            om = om.to(dtype).contiguous(memory_format=torch.channels_last)
            m = cp.DCN(Cin, Cout, (3, 3), 1, 1).cuda()
            w, bias = m.weight.detach(), m.bias.detach()
            off, msk = om[:, :18], om[:, 18:]
            from centerpoly_b200 import dcn_v2 as D
            cp.set_dcn_path("simt" if args.simt else "auto")
            fwd = lambda: D._forward_impl(x, off, msk, w, bias, (1, 1), (1, 1), (1, 1), 1, True)
            ms = time_fn(fwd, args.iters, flush)
            flops = 2.0 * Cout * Cin * 9 * H * W * B
            byts = es * B * H * W * (Cin + 27 + Cout)
            row = dict(dtype=dname, Cin=Cin, Cout=Cout, H=H, W=W, B=B, fwd_ms=round(ms, 4), tflops=round(flops / ms / 1e9, 1),
                       gbs=round(byts / ms / 1e6, 1), frac_tensor=round(flops / ms / 1e9 / tf, 3), frac_hbm=round(byts / ms / 1e6 / hbm, 3))
            if args.bwd:
                gout = torch.randn(B, Cout, H, W, device="cuda", generator=g).to(dtype).contiguous(memory_format=torch.channels_last)
                bwd = lambda: D._backward_impl(x, off, msk, w, gout, (1, 1), (1, 1), (1, 1), 1, True, (True,) * 5)
                row["bwd_ms"] = round(time_fn(bwd, max(3, args.iters // 4), flush), 4)
            total_ms += ms * COUNT_PER_FRAME[(Cin, Cout)]
            rows.append(row)
            print(json.dumps(row), flush=True)
        print(json.dumps(dict(dtype=dname, B=B, dcn_fwd_ms_per_frame=round(total_ms, 4), peaks=src, hbm_gbs=hbm, bf16_tflops=tf)), flush=True)
    cp.set_dcn_path("auto")


if __name__ == "__main__":
    main()
