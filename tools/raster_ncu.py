"""One cpoly_rasterise call on the Cityscapes workload of tools/raster_bench.py, for ncu captures."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import centerpoly_b200 as cp  # noqa: E402
from raster_bench import street  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1
dets = torch.from_numpy(np.stack([street(40 + b) for b in range(B)])).cuda()
buf = torch.empty((B, 100, 1024, 2048), dtype=torch.uint8, device="cuda")
for _ in range(2):
    r = cp.rasterise_instances(dets, 2048, 1024, 0.05, out=buf)
torch.cuda.synchronize()
print(int(r.count.sum()), int(r.keep.sum()))
