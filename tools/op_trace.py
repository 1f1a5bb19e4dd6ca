"""Which ATen ops are behind the framework kernels of one eager frame (torch.profiler, debug aid)."""
import os, sys
import torch
from torch.profiler import profile, ProfilerActivity
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from centerpoly_b200.harness.engine import PolydetEngine
eng = PolydetEngine(1024, 2048, batch=1, dtype="bf16", use_graph=False, n_slots=1)
eng.slots[0]["x"].normal_()
for _ in range(3):
    eng.infer_resident(0)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA], record_shapes=True, with_stack=False) as prof:
    eng.infer_resident(0)
    torch.cuda.synchronize()
print(prof.key_averages(group_by_input_shape=True).table(sort_by="cuda_time_total", row_limit=45, max_name_column_width=50, max_shapes_column_width=70))
