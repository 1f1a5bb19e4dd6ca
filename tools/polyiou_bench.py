"""Polygon-IoU loss: GPU kernel time (forward incl. gradient, and the backward scatter) on the training shape of
BASELINE.json configs[3] (16 crops x 128 slots, ~20 objects each, N = 32 vertices) next to the CPU oracle port on
a bounded sample of the same objects (the reference's own PolyLoss costs about the same per object).

    python tools/polyiou_bench.py [--objects 20] [--cpu-objects 8]
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=16)
    ap.add_argument("--slots", type=int, default=128)
    ap.add_argument("--objects", type=int, default=20)
    ap.add_argument("--vertices", type=int, default=32)
    ap.add_argument("--cpu-objects", type=int, default=8)
    ap.add_argument("--reps", type=int, default=50)
    args = ap.parse_args()
    import centerpoly_b200 as cp
    from make_polyiou_cases import star_case
    from oracle import polyloss_oracle as po
    B, M, N = args.batch, args.slots, args.vertices
    H, W = 128, 256
    pred, mask, _, tgt, _ = star_case(7, B, M, N, args.objects, 1.0, 0.03)
    g = torch.Generator().manual_seed(0)
    ind = torch.stack([torch.randperm(H * W, generator=g)[:M] for _ in range(B)])
    out = torch.randn(B, 2 * N, H * W, generator=g)
    out.scatter_(2, ind.unsqueeze(1).expand(B, 2 * N, M), pred.transpose(1, 2))
    out = out.view(B, 2 * N, H, W).cuda().contiguous(memory_format=torch.channels_last).requires_grad_(True)
    mask_d, ind_d, tgt_d = mask.cuda(), ind.cuda(), tgt.cuda()

    def step():
        out.grad = None
        loss = cp.poly_iou_loss(out, mask_d, ind_d, tgt_d)
        loss.backward()
        return loss

    for _ in range(5):
        loss = step()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(args.reps):
        step()
    b.record()
    torch.cuda.synchronize()
    us_step = a.elapsed_time(b) * 1000 / args.reps
    # kernel alone (forward launch: clip + gradient + reduce), no autograd / allocation around it
    with torch.no_grad():
        a.record()
        for _ in range(args.reps):
            cp.poly_iou_loss(out.detach(), mask_d, ind_d, tgt_d)
        b.record()
        torch.cuda.synchronize()
    us_fwd = a.elapsed_time(b) * 1000 / args.reps
    n_obj = int(mask.sum())
    # CPU: oracle port on a bounded sample
    k = args.cpu_objects
    pm = np.zeros((1, M), np.uint8)
    pm[0, :k] = 1
    t0 = time.perf_counter()
    want = po.poly_iou_loss_oracle(pred[:1].numpy(), pm, tgt[:1].numpy())
    cpu_s = time.perf_counter() - t0
    _, iou, status = cp.poly_iou_loss(out.detach(), mask_d, ind_d, tgt_d, return_details=True)
    ok = bool(np.allclose(iou[0, :k].cpu().numpy(), want["iou"][0, :k], rtol=1e-3, atol=1e-6))
    print(json.dumps(dict(workload=f"poly_iou B={B} M={M} N={N} objects={n_obj}", us_fwd_bwd_autograd=round(us_step, 1),
                          us_forward_call=round(us_fwd, 1), objects_per_s_gpu=round(n_obj / (us_step * 1e-6)),
                          cpu_oracle_ms_per_object=round(1000 * cpu_s / k, 1), cpu_objects=k, cpu_threads=torch.get_num_threads(),
                          objects_per_s_cpu=round(k / cpu_s, 1), speedup=round((n_obj / (us_step * 1e-6)) / (k / cpu_s)),
                          bad_status=int((status != 0).sum()), sample_matches_oracle=ok, loss=float(loss))), flush=True)


if __name__ == "__main__":
    main()
