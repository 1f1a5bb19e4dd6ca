"""GPU parity: cpoly_rasterise (through the C ABI) against the reference's own rasterisation output
(tests/golden/raster_*.npz, written by the unmodified `format_and_write_to_cityscapes`, cityscapes.py:196-283) and
against oracle/raster_oracle.py on seeded scenes.  Byte masks, areas, order and keep flags are bit-exact; the
confidence is the same fp64 product."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN

pytestmark = pytest.mark.gpu

NAMES = ['__background__', 'person', 'rider', 'car', 'truck', 'bus', 'train', 'motorcycle', 'bicycle', 'pole', 'traffic sign', 'traffic light']
LABEL_TO_ID = {'person': 24, 'rider': 25, 'car': 26, 'truck': 27, 'bus': 28, 'train': 31, 'motorcycle': 32, 'bicycle': 33}


def all_bboxes_of(rows, num_classes=11):
    cls = rows[:, 5].astype(int)
    return {j + 1: np.concatenate([rows[cls == j][:, :5], rows[cls == j][:, 6:]], axis=1).astype(np.float32).tolist() for j in range(num_classes)}


def scene(seed, K, N, W, H, wild=False):
    rng = np.random.default_rng(seed)
    rows = np.zeros((K, 2 * N + 7), np.float32)
    for i in range(K):
        cx, cy = rng.uniform(-0.1 * W, 1.1 * W), rng.uniform(-0.1 * H, 1.1 * H)
        ang = np.linspace(0, 2 * np.pi, N, endpoint=False) + rng.uniform(0, 1)
        rad = rng.uniform(1, 0.3 * min(W, H)) * rng.uniform(0.5, 1.0, N)
        if wild and i % 3 == 0:
            ang = rng.permutation(ang)
        x, y = cx + 1.4 * rad * np.cos(ang), cy + rad * np.sin(ang)
        if wild and i % 4 == 1:
            x, y = np.round(x), np.floor(y) + 0.9951
        if wild and i % 7 == 2:
            x[:], y[:] = np.round(cx), np.round(cy)     # a single point: 21-pixel disc, never kept
        rows[i, 6:-1:2], rows[i, 7:-1:2] = x, y
        rows[i, 4] = rng.choice([0.03, 0.05, 0.3, 0.5, 0.9]) if i % 3 == 0 else rng.uniform(0, 1)
        rows[i, 5] = rng.integers(0, 11)
        rows[i, -1] = rng.choice([0.25, 0.75]) if i % 5 == 0 else rng.uniform(0, 1)
    return rows


def check_against_oracle(r, dets, b, W, H, thresh):
    from oracle import raster_oracle as ro
    _, records = ro.rasterise_oracle(all_bboxes_of(dets[b]), NAMES, thresh, width=W, height=H)
    n = int(r.count[b])
    assert n == len(records)
    order = r.order[b].cpu().numpy()
    assert np.all(order[n:] == -1)
    masks = r.masks[b, :n].cpu().numpy() if r.masks is not None else None
    area, keep, conf, verts = (t[b].cpu().numpy() for t in (r.area, r.keep, r.conf, r.verts))
    for s, (points, score, label, depth, vis, kept) in enumerate(records):
        i = order[s]
        assert dets[b, i, 4] == np.float32(score) and dets[b, i, -1] == np.float32(depth) and NAMES[int(dets[b, i, 5]) + 1] == label, s
        assert [tuple(v) for v in verts[s]] == points, s
        assert area[s] == np.count_nonzero(vis) and bool(keep[s]) == kept, (s, area[s], np.count_nonzero(vis))
        assert conf[s] == min(1, score * 1.2)
        if masks is not None:
            assert np.array_equal(masks[s], vis), s


@pytest.fixture(params=["px16", "px8"])
def store_width(request, monkeypatch):
    """both store variants of raster_cover_kernel (16-byte stores by default, 8-byte when forced or W % 16 != 0)"""
    if request.param == "px8":
        monkeypatch.setenv("CPOLY_RASTER_PX", "8")
    return request.param


@pytest.mark.parametrize("name", ["raster_street", "raster_tangled", "raster_tiny"])
def test_vs_reference_golden(name, store_width):
    import centerpoly_b200 as cp
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    dets = torch.from_numpy(g["dets"])[None].cuda()
    r = cp.rasterise_instances(dets, 2048, 1024, float(g["thresh"]))
    torch.cuda.synchronize()
    assert int(r.status[0]) == 0
    inst = r.instances(dets)
    want = np.unpackbits(g["packed"], axis=-1).astype(bool)
    assert len(inst) == len(want)
    assert [LABEL_TO_ID[label] for _, label, _ in inst] == list(g["label_ids"])
    assert np.array_equal(np.array([c for _, _, c in inst]), g["confs"])
    for i, (m, _, _) in enumerate(inst):
        assert set(np.unique(m)) <= {0, 255}
        assert np.array_equal(m > 0, want[i]), i
    n = int(r.count[0])
    assert np.array_equal(r.area[0, :n][r.keep[0, :n]].cpu().numpy(), g["areas"])


@pytest.mark.parametrize("B,K,N,W,H,wild", [(1, 100, 32, 256, 128, False), (3, 40, 32, 203, 97, True), (2, 64, 16, 517, 64, True),
                                            (1, 256, 8, 96, 33, True), (2, 17, 64, 64, 200, True), (1, 5, 3, 9, 9, True), (2, 30, 32, 1040, 48, True), (1, 250, 64, 160, 64, True)])
def test_vs_oracle(B, K, N, W, H, wild, store_width):
    import centerpoly_b200 as cp
    dets = np.stack([scene(100 * B + b, K, N, W, H, wild) for b in range(B)])
    d = torch.from_numpy(dets).cuda()
    r = cp.rasterise_instances(d, W, H, 0.05)
    torch.cuda.synchronize()
    for b in range(B):
        check_against_oracle(r, dets, b, W, H, 0.05)
    # areas / keep without materialising the masks, and into a caller-provided buffer with stale contents
    r2 = cp.rasterise_instances(d, W, H, 0.05, want_masks=False)
    buf = torch.full((B, K, H, W), 77, dtype=torch.uint8, device="cuda")
    r3 = cp.rasterise_instances(d, W, H, 0.05, out=buf)
    for k in ("area", "keep", "order", "count", "conf"):
        assert torch.equal(getattr(r, k), getattr(r2, k)) and torch.equal(getattr(r, k), getattr(r3, k)), k
    for b in range(B):
        n = int(r.count[b])
        assert torch.equal(r3.masks[b, :n], r.masks[b, :n]) and bool((r3.masks[b, n:] == 77).all())   # slots past count untouched


def test_nothing_above_threshold_and_clamping():
    import centerpoly_b200 as cp
    dets = scene(3, 20, 32, 128, 64)
    dets[:, 4] = 0.01
    r = cp.rasterise_instances(torch.from_numpy(dets)[None].cuda(), 128, 64, 0.05)
    assert int(r.count[0]) == 0 and not bool(r.keep.any()) and bool((r.order == -1).all())
    dets[:, 4] = 0.9
    dets[0, 6] = 1e6        # one x far outside the exact range: clamped and reported
    dets[1, 7] = float("nan")
    r = cp.rasterise_instances(torch.from_numpy(dets)[None].cuda(), 128, 64, 0.05)
    assert int(r.count[0]) == 20 and int(r.status[0]) == 1


def test_argument_errors():
    import centerpoly_b200 as cp
    with pytest.raises(RuntimeError):
        cp.rasterise_instances(torch.zeros(1, 10, 70, device="cuda"), 64, 64)            # 2N+7 must be odd
    with pytest.raises(RuntimeError, match="K must be"):
        cp.rasterise_instances(torch.zeros(1, 300, 71, device="cuda"), 64, 64)
    with pytest.raises(RuntimeError, match="mask size"):
        cp.rasterise_instances(torch.zeros(1, 10, 71, device="cuda"), 9000, 64, want_masks=False)
    with pytest.raises(RuntimeError):
        cp.rasterise_instances(torch.zeros(1, 10, 71), 64, 64)                           # host tensor: no CPU path


def test_chained_from_decode():
    """heads -> cpoly_decode(post_affine) -> cpoly_rasterise, all on the device, == host post-process + oracle."""
    import centerpoly_b200 as cp
    g = torch.Generator(device="cuda").manual_seed(9)
    B, C, H, W, N, K = 2, 8, 64, 128, 16, 50
    heat = torch.sigmoid(torch.randn(B, C, H, W, device="cuda", generator=g) * 2 - 3)
    polys = torch.randn(B, 2 * N, H, W, device="cuda", generator=g) * 6
    depth = torch.rand(B, 1, H, W, device="cuda", generator=g)
    reg = torch.rand(B, 2, H, W, device="cuda", generator=g)
    trans = cp.batch_inverse_affine([np.array([W * 2.0, H * 2.0], np.float32)] * B, [float(max(H, W) * 4)] * B, W, H)
    dets = cp.polydet_decode(heat, polys, depth, reg=reg, K=K, post_affine=trans)
    r = cp.rasterise_instances(dets, W * 4, H * 4, 0.3)
    torch.cuda.synchronize()
    for b in range(B):
        check_against_oracle(r, dets.cpu().numpy(), b, W * 4, H * 4, 0.3)
