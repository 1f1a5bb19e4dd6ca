import sys, time, torch
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from centerpoly_b200.harness.engine import PolydetEngine
for rast in (False, True):
    eng = PolydetEngine(rasterise=rast, n_slots=2)
    g = torch.Generator(device="cuda").manual_seed(1)
    eng.set_resident_frames([torch.randn(1,3,1024,2048, device="cuda", generator=g) for _ in range(2)])
    for i in range(5): eng.infer_resident(i)
    eng.stream.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(eng.stream):
        a.record(eng.stream)
        for i in range(40): eng.infer_resident(i)
        b.record(eng.stream)
    eng.stream.synchronize()
    ms = a.elapsed_time(b)/40
    extra = ''
    if rast:
        r = eng.slots[0]["raster"]; extra = f' instances={int(r.count[0])} kept={int(r.keep.sum())}'
    print(f'rasterise={rast}: {ms:.3f} ms/frame = {1000/ms:.1f} fps{extra}', flush=True)
    del eng; torch.cuda.empty_cache()
