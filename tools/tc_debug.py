"""Bring-up diagnostics for the GPU box: each stage runs in its own subprocess with a timeout so that a
device trap in one stage cannot take the others down.  Usage: python tools/tc_debug.py [stage ...]"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def stage_simt():
    import numpy as np, torch
    from util import make_dcn_case, gate
    from test_gpu_dcn import run_ours, oracle
    case = make_dcn_case(1, 1, 64, 64, 12, 20)
    got, want = run_ours(case, "simt"), oracle(case)
    for k in ("y", "gx", "goffset", "gmask", "gw", "gb"):
        e = np.abs(got[k] - want[k]).max() / np.abs(want[k]).max()
        print(f"simt {k}: max-rel {e:.3e}")


def _tc(dtype_name, shape, zero=False):
    import numpy as np, torch
    from util import make_dcn_case
    from test_gpu_dcn import run_ours, oracle
    dtype = torch.bfloat16 if dtype_name == "bf16" else torch.float32
    B, Cin, Cout, H, W = shape
    case = make_dcn_case(2, B, Cin, Cout, H, W, offset_mode="zero" if zero else "random", dtype=dtype)
    want = oracle(case, backward=False)["y"]
    got = run_ours(case, "tcgen05", dtype=dtype, channels_last=True, backward=False)["y"]
    err = np.abs(got - want)
    scale = np.abs(want).max()
    print(f"tc {dtype_name} {shape} zero={zero}: max-rel {err.max() / scale:.3e}  l2-rel {np.linalg.norm(got - want) / np.linalg.norm(want):.3e}")
    if err.max() / scale > 2e-2:
        e = err[0] / scale                       # (Cout, H, W)
        print("  per-channel-block(16) max err:", np.round(e.reshape(Cout // 16, 16, H, W).max(axis=(1, 2, 3)), 3))
        print("  per-row max err:", np.round(e.max(axis=(0, 2)), 3))
        print("  per-col max err:", np.round(e.max(axis=(0, 1)), 3))
        print("  got[0,:4,0,:4]\n", got[0, :4, 0, :4], "\n  want\n", want[0, :4, 0, :4])
        print("  nan:", np.isnan(got).sum(), " zeros:", (got == 0).mean())


def stage_tc_bf16_zero(): _tc("bf16", (1, 64, 64, 8, 16), zero=True)
def stage_tc_bf16(): _tc("bf16", (1, 64, 64, 16, 32))
def stage_tc_fp32(): _tc("fp32", (1, 64, 64, 16, 32))
def stage_tc_big(): _tc("bf16", (1, 256, 256, 16, 32)); _tc("fp32", (2, 128, 64, 24, 40)); _tc("bf16", (1, 512, 256, 8, 16))


def stage_decode():
    import numpy as np, torch
    import centerpoly_b200 as cp
    from util import make_decode_case
    from oracle.decode_oracle import polydet_decode_oracle
    heat, polys, depth, reg = make_decode_case(1, 2, 8, 64, 96, 32)
    want = polydet_decode_oracle(heat.numpy(), polys.numpy(), depth.numpy(), reg=reg.numpy(), K=100)
    got = cp.polydet_decode(heat.cuda(), polys.cuda(), depth.cuda(), reg=reg.cuda(), K=100).cpu().numpy()
    print("decode: exact =", bool((got == want).all()), " score-order exact =", bool((got[..., 4] == want[..., 4]).all()),
          " max abs diff", float(np.abs(got - want).max()))


STAGES = ["simt", "tc_bf16_zero", "tc_bf16", "tc_fp32", "tc_big", "decode"]

if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "--run":
        globals()["stage_" + sys.argv[2]]()
        sys.exit(0)
    for st in (sys.argv[1:] or STAGES):
        print(f"===== stage {st}", flush=True)
        try:
            r = subprocess.run([sys.executable, os.path.abspath(__file__), "--run", st], timeout=240, capture_output=True, text=True)
            print(r.stdout[-4000:])
            if r.returncode != 0:
                print(f"  [exit {r.returncode}] stderr tail:\n{r.stderr[-3000:]}")
        except subprocess.TimeoutExpired:
            print("  [TIMEOUT]")
