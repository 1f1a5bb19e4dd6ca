"""Data-parallel polydet training step (BASELINE.json configs[3]): 1024x512 crops, 16 per GPU, focal + L1 losses,
Adam, bucketed gradient all-reduce over NCCL.  One process per GPU:

    python tools/train_bench.py [--batch 16] [--steps 10] [--dtype bf16|fp32]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/train_bench.py

Prints one JSON line on rank 0: images/s over all ranks (CUDA events, barrier + max over ranks)."""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def _launches(key):
    from centerpoly_b200 import _lib
    return _lib.LAUNCHES.get(key, 0)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=16)
    ap.add_argument("--height", type=int, default=512)
    ap.add_argument("--width", type=int, default=1024)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--dtype", default="bf16", choices=["bf16", "fp32"])
    ap.add_argument("--dcn-path", default="auto", choices=["auto", "simt"])
    ap.add_argument("--poly-loss", default="l1", choices=["l1", "l1+iou"],
                    help="l1: cartesian targets, masked L1; l1+iou: polar targets, the reference's PolyLoss l1+iou (IoU arm on the GPU kernel)")
    ap.add_argument("--cudnn-upsample", action="store_true", help="keep nn.ConvTranspose2d for the IDAUp up-samplers (cuDNN grouped kernels)")
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    import centerpoly_b200 as cp
    from centerpoly_b200.harness.dla34 import PolydetDLA34, calibrate_random_init
    from centerpoly_b200.harness.train import DataParallelTrainer, polydet_l1_iou_loss, polydet_l1_loss, synthetic_batch
    cp.set_dcn_path(args.dcn_path)
    torch.backends.cudnn.benchmark = True
    torch.manual_seed(317)                                   # same initial weights on every rank
    net = PolydetDLA34().cuda().to(memory_format=torch.channels_last)
    calibrate_random_init(net, 128, 256)
    n_up = 0
    if not args.cudnn_upsample:
        from centerpoly_b200.neck import use_training_upsamplers
        n_up = use_training_upsamplers(net)     # IDAUp up_i: forward + both gradients on our kernels
    iou = args.poly_loss == "l1+iou"
    trainer = DataParallelTrainer(net, lr=2e-4, dtype=torch.bfloat16 if args.dtype == "bf16" else None,
                                  loss_fn=polydet_l1_iou_loss if iou else polydet_l1_loss)
    batch = synthetic_batch(args.batch, args.height, args.width, device="cuda", seed=1000 + rank, rep="polar" if iou else "cartesian")
    batch["input"] = batch["input"].contiguous(memory_format=torch.channels_last)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    losses = []
    for _ in range(args.warmup):
        losses.append(float(trainer.step(batch)))
    barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(args.steps):
        loss = trainer.step(batch)
    b.record()
    barrier()
    ms = a.elapsed_time(b)
    losses.append(float(loss))
    if world > 1:
        t = torch.tensor([ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t)
    if rank == 0:
        print(json.dumps(dict(metric=f"polydet dla_34 training images/s (1024x512 crops, focal + {'L1 + polygon IoU (polar)' if iou else 'L1'}, Adam, data-parallel)",
                              poly_loss=args.poly_loss, poly_iou_launches=_launches("poly_iou"), upsamplers_on_our_kernels=n_up,
                              value=world * args.batch * args.steps / (ms / 1000.0), unit="images/s", n_gpus=world,
                              ms_per_step=ms / args.steps, batch_per_gpu=args.batch, dtype=args.dtype, dcn_path=args.dcn_path,
                              loss_first=losses[0], loss_last=losses[-1], scaling="weak", data="synthetic")), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
