"""polydet_decode timing against the HBM roofline (SURVEY 8(d) "Decode - HBM bound").

    python tools/decode_bench.py [--iters 50]

Two workloads: the Cityscapes frame of BASELINE.json configs[1] (B=1, 8 classes, 256x512 maps, cartesian, K=100;
4.44 MB algorithmic -> latency-bound, reported in us) and the polar IDD batch of configs[4] (B=64, 9 classes,
272x488 maps, rep='polar'; 321 MB algorithmic -> bandwidth-bound, reported as a fraction of the measured HBM peak).
Algorithmic bytes = s*B*C*H*W (heat once) + B*K*(2N+3)*32 (sector-granular gathers from NCHW heads) + B*K*(2N+7)*4.
CUDA events around a CUDA-graph replay of one call, L2 flushed before every replay.  Prints one JSON line per workload."""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import centerpoly_b200 as cp  # noqa: E402


def hbm_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p))["hbm_gbs"], "measured"
    return 6650.0, "fallback"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=50)
    args = ap.parse_args()
    peak, src = hbm_peak()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    g = torch.Generator(device="cuda").manual_seed(317)
    for name, (B, C, H, W, N, K, rep) in {"cityscapes_b1_cartesian": (1, 8, 256, 512, 32, 100, "cartesian"),
                                          "idd_b64_polar": (64, 9, 272, 488, 32, 100, "polar")}.items():
        heat = torch.sigmoid(torch.randn(B, C, H, W, device="cuda", generator=g) - 2.19)
        polys = torch.randn(B, 2 * N, H, W, device="cuda", generator=g).abs() * 10
        depth = torch.rand(B, 1, H, W, device="cuda", generator=g)
        reg = torch.rand(B, 2, H, W, device="cuda", generator=g)
        fn = lambda: cp.polydet_decode(heat, polys, depth, reg=reg, K=K, rep=rep)
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        # replayed from a CUDA graph like the inference engine does (5 small launches: without it the B=1 number is
        # the host's launch rate, not the kernels)
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            fn()
        ts = []
        for _ in range(args.iters):
            flush.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            graph.replay()
            b.record()
            torch.cuda.synchronize()
            ts.append(a.elapsed_time(b))
        ts.sort()
        ms = ts[len(ts) // 2]
        alg = 4 * B * C * H * W + B * K * (2 * N + 3) * 32 + B * K * (2 * N + 7) * 4
        print(json.dumps(dict(workload=name, B=B, C=C, H=H, W=W, K=K, rep=rep, us=round(ms * 1000, 1),
                              algorithmic_mb=round(alg / 1e6, 2), gbs=round(alg / ms / 1e6, 1), hbm_peak_gbs=peak,
                              frac_hbm=round(alg / ms / 1e6 / peak, 3), peak_source=src, launches_per_call=5, timed="CUDA graph replay, L2 flushed")), flush=True)


if __name__ == "__main__":
    main()
