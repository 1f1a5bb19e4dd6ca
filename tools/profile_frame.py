"""One eager (no CUDA graph) polydet frame between cudaProfilerStart/Stop, for the ncu launch list:

    ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv \
        --log-file gpurun_out/launches.csv python tools/profile_frame.py [--dtype bf16|fp32] [--frames N]

Prints the CUDA-event time of the profiled frames so that the kernel SHARES of the launch list can be
compared with a number taken without the profiler."""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from centerpoly_b200.harness.engine import PolydetEngine  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--dtype", default="bf16", choices=["bf16", "fp32"])
    ap.add_argument("--frames", type=int, default=1)
    ap.add_argument("--height", type=int, default=1024)
    ap.add_argument("--width", type=int, default=2048)
    args = ap.parse_args()
    eng = PolydetEngine(args.height, args.width, batch=1, dtype=args.dtype, use_graph=False, n_slots=1)
    eng.slots[0]["x"].normal_()
    for _ in range(3):
        eng.infer_resident(0)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.profiler.start()
    with torch.cuda.stream(eng.stream):
        a.record(eng.stream)
    for _ in range(args.frames):
        eng.infer_resident(0)
    with torch.cuda.stream(eng.stream):
        b.record(eng.stream)
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()
    print(f"frames {args.frames} dtype {args.dtype} eager ms/frame {a.elapsed_time(b) / args.frames:.3f}")


if __name__ == "__main__":
    main()
