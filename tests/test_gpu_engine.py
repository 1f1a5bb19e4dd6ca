"""GPU: the captured inference engine (network -> decode + back-mapping -> instance masks) replays to the same bytes
as the eager calls, and its masks are what the CPU oracle draws from its detections."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

NAMES = ['__background__', 'person', 'rider', 'car', 'truck', 'bus', 'train', 'motorcycle', 'bicycle', 'pole', 'traffic sign', 'traffic light']


def test_engine_graph_with_masks_matches_eager_and_oracle():
    import centerpoly_b200 as cp
    from centerpoly_b200.harness.engine import PolydetEngine
    from oracle import raster_oracle as ro
    H, W, K = 256, 512, 40
    eng = PolydetEngine(height=H, width=W, batch=1, K=K, n_slots=1, rasterise=True, thresh=0.02)
    g = torch.Generator(device="cuda").manual_seed(5)
    frame = torch.randn(1, 3, H, W, device="cuda", generator=g)
    eng.set_resident_frames([frame])
    eng.infer_resident(0)
    eng.stream.synchronize()
    slot = eng.slots[0]
    dets, r = slot["dets"].clone(), slot["raster"]
    assert dets.shape == (1, K, 71) and bool(torch.isfinite(dets).all())
    # replay == eager on the same frame
    with torch.cuda.stream(eng.stream):
        eager = eng._step(slot["x"]).clone()
    eng.stream.synchronize()
    assert torch.equal(eager, dets)
    # masks of the replay == a standalone call on the replay's detections == the CPU oracle
    alone = cp.rasterise_instances(dets, W, H, 0.02)
    n = int(r.count[0])
    assert n == int(alone.count[0]) and n > 0
    assert torch.equal(alone.masks[0, :n], r.masks[0, :n]) and torch.equal(alone.keep, r.keep) and torch.equal(alone.area, r.area)
    d = dets[0].cpu().numpy()
    cls = d[:, 5].astype(int)
    ab = {j + 1: np.concatenate([d[cls == j][:, :5], d[cls == j][:, 6:]], axis=1).astype(np.float32).tolist() for j in range(11)}
    _, records = ro.rasterise_oracle(ab, NAMES, 0.02, width=W, height=H)
    assert len(records) == n
    masks = r.masks[0, :n].cpu().numpy()
    for s, rec in enumerate(records):
        assert np.array_equal(masks[s], rec[4]), s
    # a second replay on another frame rewrites the same buffers
    eng.set_resident_frames([torch.randn(1, 3, H, W, device="cuda", generator=g)])
    eng.infer_resident(0)
    eng.stream.synchronize()
    assert not torch.equal(slot["dets"], dets)
    again = cp.rasterise_instances(slot["dets"], W, H, 0.02)
    m = int(again.count[0])
    assert torch.equal(again.masks[0, :m], slot["raster"].masks[0, :m]) and torch.equal(again.area, slot["raster"].area)
