"""Generates tests/golden/dcnfull_*.npz: BASELINE.json configs[2] shapes at FULL size (64->64 @256x512 and
128->128 @128x256, B = 1 and 4), offsets ~ N(0, 2^2) with 5 % far taps (tests/util.py::make_dcn_case), forward and
all five gradients of ``torchvision.ops.deform_conv2d`` in float64 on the CPU (the stand-in BASELINE.json names for
the missing DCNv2 submodule, /root/reference/.SUBMODULES.json:8-14).

    python tests/golden/make_dcn_fullsize_golden.py          # ~10 min of CPU, build container only

The full tensors are too large to commit (y alone is 33 MB per frame), so a fixture stores, per case:
  * the SEED: inputs are regenerated bit-for-bit by make_dcn_case on the GPU box (CPU generator);
  * expected values of y / gx / goffset / gmask at 2048 seeded pixel positions (all channels), gw and gb in full;
  * max|.| and the 2-norm of every full expected tensor (the gate's denominators) and its plain sum (a checksum that
    sees every element).
"""
import os
import sys
import time

import numpy as np
import torch
from torchvision.ops import deform_conv2d

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from util import make_dcn_case  # noqa: E402

N_SAMPLES = 2048
CASES = [  # name, seed, B, Cin, Cout, H, W, input rounding
    ("dcnfull_64x64_256x512_b1_f32", 401, 1, 64, 64, 256, 512, torch.float32),
    ("dcnfull_64x64_256x512_b1_bf16", 402, 1, 64, 64, 256, 512, torch.bfloat16),
    ("dcnfull_128x128_128x256_b1_f32", 403, 1, 128, 128, 128, 256, torch.float32),
    ("dcnfull_128x128_128x256_b4_bf16", 404, 4, 128, 128, 128, 256, torch.bfloat16),
    ("dcnfull_64x64_256x512_b4_bf16", 405, 4, 64, 64, 256, 512, torch.bfloat16),
]


def sample_positions(seed, B, H, W):
    g = torch.Generator().manual_seed(9000 + seed)
    flat = torch.randperm(B * H * W, generator=g)[:N_SAMPLES].sort().values
    return flat


def take(t, flat):
    """(B,C,H,W) -> (n, C) at flat (b*H*W + y*W + x) positions."""
    B, C, H, W = t.shape
    return t.permute(0, 2, 3, 1).reshape(B * H * W, C)[flat]


if __name__ == "__main__":
    only = sys.argv[1:]
    torch.set_num_threads(os.cpu_count() or 1)
    for name, seed, B, Cin, Cout, H, W, rnd in CASES:
        if only and name not in only:
            continue
        t0 = time.time()
        case = make_dcn_case(seed, B, Cin, Cout, H, W, dtype=rnd)
        leaves = {k: case[k].double().requires_grad_(True) for k in ("x", "offset", "mask", "weight", "bias")}
        y = deform_conv2d(leaves["x"], leaves["offset"], leaves["weight"], leaves["bias"], stride=1, padding=1, dilation=1,
                          mask=leaves["mask"])
        y.backward(case["gout"].double())
        full = dict(y=y.detach(), gx=leaves["x"].grad, goffset=leaves["offset"].grad, gmask=leaves["mask"].grad)
        flat = sample_positions(seed, B, H, W)
        out = dict(seed=np.int64(seed), shape=np.array([B, Cin, Cout, H, W]), rounding=np.array(str(rnd).split(".")[-1]),
                   positions=flat.numpy(), gw=leaves["weight"].grad.numpy().astype(np.float32),
                   gb=leaves["bias"].grad.numpy().astype(np.float32))
        for k, t in full.items():
            out[k] = take(t, flat).numpy().astype(np.float32)
            out[k + "_max"] = np.float64(t.abs().max())
            out[k + "_l2"] = np.float64(t.norm())
            out[k + "_sum"] = np.float64(t.sum())
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
        print(name, "%.0f s" % (time.time() - t0), {k: float(out[k + "_max"]) for k in full}, flush=True)
