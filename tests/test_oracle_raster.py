"""CPU: pins oracle/raster_oracle.py to the reference's own rasterisation output (tests/golden/raster_*.npz, written by
the unmodified `format_and_write_to_cityscapes`, cityscapes.py:196-283) and proves the closed form the CUDA kernel
evaluates against the real PIL on random polygons."""
import os

import numpy as np
import pytest

from oracle import raster_oracle as ro

GOLD = os.path.join(os.path.dirname(__file__), "golden")
NAMES = ['__background__', 'person', 'rider', 'car', 'truck', 'bus', 'train', 'motorcycle', 'bicycle', 'pole', 'traffic sign', 'traffic light']
LABEL_TO_ID = {'person': 24, 'rider': 25, 'car': 26, 'truck': 27, 'bus': 28, 'train': 31, 'motorcycle': 32, 'bicycle': 33}


def all_bboxes_of(rows, num_classes=11):
    cls = rows[:, 5].astype(int)
    return {j + 1: np.concatenate([rows[cls == j][:, :5], rows[cls == j][:, 6:]], axis=1).astype(np.float32).tolist() for j in range(num_classes)}


def test_bresenham_known_answers():
    # the example of the PyPI package's README, and its symmetric / degenerate cases
    assert list(ro.bresenham(-1, -4, 3, 2)) == [(-1, -4), (0, -3), (0, -2), (1, -1), (2, 0), (2, 1), (3, 2)]
    assert list(ro.bresenham(0, 0, 0, 0)) == [(0, 0)]
    assert list(ro.bresenham(2, 5, 2, 2)) == [(2, 5), (2, 4), (2, 3), (2, 2)]
    assert list(ro.bresenham(0, 0, 3, 3)) == [(0, 0), (1, 1), (2, 2), (3, 3)]
    assert list(ro.bresenham(0, 0, 5, 2)) == [(0, 0), (1, 0), (2, 1), (3, 1), (4, 2), (5, 2)]


def test_bresenham_closed_form():
    rng = np.random.default_rng(5)
    for _ in range(3000):
        x0, y0, x1, y1 = (int(v) for v in rng.integers(-40, 40, 4))
        dx, dy = abs(x1 - x0), abs(y1 - y0)
        sx, sy = (1 if x1 > x0 else -1), (1 if y1 > y0 else -1)
        if dx > dy:
            want = [(x0 + sx * k, y0 + sy * ((2 * dy * k + dx) // (2 * dx))) for k in range(dx + 1)]
        else:
            want = [(x0 + sx * ((2 * dx * k + dy) // (2 * dy) if dy else 0), y0 + sy * k) for k in range(dy + 1)]
        assert list(ro.bresenham(x0, y0, x1, y1)) == want


def test_round2_trunc_matches_string_round_trip():
    rng = np.random.default_rng(6)
    base = rng.integers(-3000, 3000, 20000).astype(np.float64)
    frac = np.concatenate([rng.uniform(0, 1, 10000), rng.uniform(0.99, 1.0, 5000), 0.995 + rng.uniform(-1e-4, 1e-4, 5000)])
    x = (base + np.sign(base + 0.5) * frac).astype(np.float32)
    x = np.concatenate([x, np.float32([0.0, -0.0, 0.995, -0.995, 0.9950001, 12.996, -12.996, 0.004, -0.004, 2047.9951, 1e-30])])
    want = np.array([int(ro._to_float(float(v))) for v in x])
    assert np.array_equal(ro.round2_trunc(x), want)


@pytest.mark.parametrize("seed", [0, 1])
def test_closed_form_equals_pil(seed):
    rng = np.random.default_rng(seed)
    for t in range(240):
        W, H = int(rng.integers(20, 130)), int(rng.integers(20, 100))
        n = int(rng.choice([3, 4, 5, 8, 16, 32]))
        kind = t % 6
        if kind == 0:     # star-shaped, the detector's typical output
            c = rng.uniform(0, [W, H]); r = rng.uniform(0, 40, n); a = np.linspace(0, 2 * np.pi, n, endpoint=False) + rng.uniform(0, 6)
            p = np.stack([c[0] + 1.5 * r * np.cos(a), c[1] + r * np.sin(a)], 1)
        elif kind == 1:   # arbitrary, self-intersecting, partly outside
            p = rng.uniform(-30, [W + 30, H + 30], (n, 2))
        elif kind == 2:   # degenerate: repeated points, tiny
            p = rng.integers(5, 12, (n, 2)).astype(float)
        elif kind == 3:   # vertical runs
            p = rng.integers(0, [W, H], (n, 2)).astype(float); p[1::2, 0] = p[0::2, 0][:len(p[1::2])]
        elif kind == 4:   # horizontal runs
            p = rng.integers(0, [W, H], (n, 2)).astype(float); p[1::2, 1] = p[0::2, 1][:len(p[1::2])]
        else:             # exact diagonals (Bresenham ties)
            p = rng.integers(0, [W, H], (n, 2)).astype(float)
            for i in range(1, n, 2):
                d = int(rng.integers(-20, 20)); p[i] = p[i - 1] + [d, d * rng.choice([-1, 1])]
        pts = [(int(x), int(y)) for x, y in p]
        assert np.array_equal(ro.instance_mask(pts, W, H), ro.closed_form_mask(pts, W, H)), (t, pts)


@pytest.mark.parametrize("name", ["raster_street", "raster_tangled", "raster_tiny"])
def test_oracle_reproduces_reference_output(name):
    g = np.load(os.path.join(GOLD, name + ".npz"))
    written, records = ro.rasterise_oracle(all_bboxes_of(g["dets"]), NAMES, float(g["thresh"]))
    assert len(written) == len(g["label_ids"])
    assert [LABEL_TO_ID[label] for _, label, _ in written] == list(g["label_ids"])
    assert np.array_equal(np.array([c for _, _, c in written], np.float64), g["confs"])   # str(float) round-trips exactly
    masks = np.unpackbits(g["packed"], axis=-1).astype(bool)
    for i, (m, _, _) in enumerate(written):
        assert np.array_equal(m > 0, masks[i]), i
    depths = [r[3] for r in records]
    assert depths == sorted(depths)
