"""Which kernels a polydet training step (configs[3], one GPU) spends its time in (torch.profiler, debug aid)."""
import os, sys
import torch
from torch.profiler import profile, ProfilerActivity
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import centerpoly_b200 as cp
from centerpoly_b200.harness.dla34 import PolydetDLA34, calibrate_random_init
from centerpoly_b200.harness.train import DataParallelTrainer, polydet_l1_iou_loss, synthetic_batch
torch.backends.cudnn.benchmark = True
torch.manual_seed(317)
net = PolydetDLA34().cuda().to(memory_format=torch.channels_last)
calibrate_random_init(net, 128, 256)
from centerpoly_b200.neck import use_training_upsamplers
from centerpoly_b200.batchnorm import use_training_batchnorm
use_training_upsamplers(net)
if os.environ.get("CPOLY_NATIVE_BN") != "1":
    use_training_batchnorm(net)
tr = DataParallelTrainer(net, lr=2e-4, dtype=torch.bfloat16, loss_fn=polydet_l1_iou_loss)
batch = synthetic_batch(16, 512, 1024, device="cuda", seed=1, rep="polar")
batch["input"] = batch["input"].contiguous(memory_format=torch.channels_last)
for _ in range(3):
    tr.step(batch)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    tr.step(batch)
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=40, max_name_column_width=90))
if os.environ.get("TRACE_FULL_NAMES"):
    for e in sorted(prof.key_averages(), key=lambda e: -e.device_time_total)[:45]:
        print(f"{e.device_time_total / 1000:8.3f} ms  n={e.count:4d}  {e.key[:420]}")
