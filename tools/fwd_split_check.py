"""Forward kernel: tap-split schedules (CPOLY_TC_SPLIT=3|9, many items per CTA) against the unsplit one.  Debug aid."""
import os, subprocess, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
if len(sys.argv) > 1 and sys.argv[1] == "child":
    from centerpoly_b200 import dcn_v2 as D
    Cin, Cout, H, W, B = (int(v) for v in sys.argv[2:7])
    g = torch.Generator(device="cuda").manual_seed(5)
    x = torch.randn(B, Cin, H, W, device="cuda", generator=g).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    om = torch.randn(B, 32, H, W, device="cuda", generator=g).contiguous(memory_format=torch.channels_last)
    w = (torch.rand(Cout, Cin, 3, 3, device="cuda", generator=g) - 0.5) * 0.1
    y, _ = D._forward_impl(x, om[:, :18], om[:, 18:27], w, None, (1, 1), (1, 1), (1, 1), 1, True)
    torch.cuda.synchronize()
    torch.save(y.float().cpu(), sys.argv[7])
    sys.exit(0)
for cfg in ["64 64 64 128 1", "128 128 128 256 1", "128 64 64 64 2", "64 64 40 72 3", "256 128 32 64 1", "64 64 88 144 1", "128 128 40 128 1", "64 128 96 160 1"]:
    outs = {}
    for sp in ("1", "3", "9"):
        path = f"/tmp/split_{sp}.pt"
        env = dict(os.environ, CPOLY_TC_SPLIT=sp)
        r = subprocess.run([sys.executable, __file__, "child", *cfg.split(), path], env=env, capture_output=True, text=True, timeout=120)
        outs[sp] = torch.load(path) if r.returncode == 0 else None
        if r.returncode != 0:
            print("cfg", cfg, "split", sp, "FAILED:", r.stderr.strip().splitlines()[-1][:200])
    for sp in ("3", "9"):
        if outs[sp] is not None and outs["1"] is not None:
            print("cfg", cfg, "split", sp, "max |diff| vs unsplit", float((outs[sp] - outs["1"]).abs().max()), "ref max", float(outs["1"].abs().max()))
