"""GPU: cpoly_rasterise_fill (csrc/raster_fill.cu) against the goldens of the unmodified reference writers (byte-exact
masks, file order, label ids, score strings) and against the oracle on random polygons."""
import os
import struct
import zlib

import numpy as np
import pytest
import torch

import centerpoly_b200  # noqa: F401
from centerpoly_b200 import raster_fill as RF
from oracle import raster_fill_oracle as R
from test_oracle_raster_fill import CASES, load, to_all_bboxes

pytestmark = pytest.mark.gpu


def read_png_gray8(path):
    data = open(path, "rb").read()
    assert data[:8] == b"\x89PNG\r\n\x1a\n"
    pos, idat, w, h = 8, b"", 0, 0
    while pos < len(data):
        n, tag = struct.unpack(">I4s", data[pos:pos + 8])
        body = data[pos + 8:pos + 8 + n]
        if tag == b"IHDR":
            w, h, depth, ctype = struct.unpack(">IIBB", body[:10])
            assert (depth, ctype) == (8, 0)
        elif tag == b"IDAT":
            idat += body
        pos += 12 + n
    raw = zlib.decompress(idat)
    rows = np.frombuffer(raw, np.uint8).reshape(h, w + 1)
    assert (rows[:, 0] == 0).all()
    return rows[:, 1:]


@pytest.mark.parametrize("name", CASES)
def test_masks_match_reference_writers(name, tmp_path):
    g, w, h, which, names, ids, want = load(name)
    dets = torch.from_numpy(g["dets"]).cuda()[None]
    r = RF.rasterise_fill_instances(dets, w, h, float(g["thresh"]), inclusive=(which == "idd"))
    torch.cuda.synchronize()
    n = int(r.count[0])
    assert n == len(want) and int(r.status[0]) == 0
    got = r.masks[0, :n].cpu().numpy()
    assert set(np.unique(got)) <= {0, 255}
    for i in range(n):
        assert np.array_equal(got[i] > 0, want[i]), (name, i, int(((got[i] > 0) != want[i]).sum()))
    assert r.area[0, :n].cpu().tolist() == [int(a) for a in g["areas"]]
    # the host writer: same file names, label ids and score strings as the reference's text file; PNGs hold the masks
    frame = "/data/seq7/frame0007.png"
    if which == "idd":
        RF.write_idd(r, dets, 0, frame, str(tmp_path))
        out_dir = tmp_path / "seq7"
    else:
        RF.write_kitti(r, dets, 0, frame, str(tmp_path))
        out_dir = tmp_path
    lines = open(out_dir / "frame0007.txt").read().splitlines()
    assert len(lines) == n
    for i, ln in enumerate(lines):
        path, label_id, score = ln.split(" ")
        assert path == f"frame0007_{i}.png" and int(label_id) == int(g["label_ids"][i]) and score == str(g["scores"][i])
    assert np.array_equal(read_png_gray8(out_dir / "frame0007_3.png") > 0, want[3])


def test_random_polygons_match_the_oracle_including_degenerate_ones():
    rng = np.random.default_rng(77)
    W, H, N, K = 208, 144, 16, 24
    rows = np.zeros((2, K, 2 * N + 7), np.float32)
    for b in range(2):
        for i in range(K):
            kind = i % 4
            if kind == 0:      # star
                ang = np.linspace(0, 2 * np.pi, N, endpoint=False)
                rad = rng.uniform(2, 50) * rng.uniform(0.3, 1.0, N)
                x, y = rng.uniform(20, 190) + rad * np.cos(ang), rng.uniform(20, 120) + rad * np.sin(ang)
            elif kind == 1:    # arbitrary, self-intersecting, partly outside
                x, y = rng.uniform(-40, W + 40, N), rng.uniform(-40, H + 40, N)
            elif kind == 2:    # repeated vertices and horizontal runs
                bx, by = rng.integers(5, W - 5, N // 2), rng.integers(5, H - 5, N // 2)
                x, y = np.repeat(bx, 2).astype(float), np.repeat(by, 2).astype(float)
                y[::4] = y[1::4]
            else:              # thin spikes
                x, y = rng.integers(60, 70, N).astype(float), rng.integers(10, 130, N).astype(float)
            rows[b, i, 6::2][:N], rows[b, i, 7::2][:N] = x, y
            rows[b, i, 4] = rng.choice([0.2, 0.5, 0.9])
            rows[b, i, 5] = rng.integers(0, 3)
            rows[b, i, -1] = rng.choice([0.1, 0.4, 0.4, 0.8])
    dets = torch.from_numpy(rows).cuda()
    r = RF.rasterise_fill_instances(dets, W, H, 0.2, inclusive=True)
    torch.cuda.synchronize()
    for b in range(2):
        inst = R.writer_masks(to_all_bboxes(rows[b], 3), W, H, 0.2, inclusive=True)
        assert int(r.count[b]) == len(inst) == K
        got = r.masks[b, :K].cpu().numpy() > 0
        for i, (_, _, m) in enumerate(inst):
            assert np.array_equal(got[i], m), (b, i, int((got[i] != m).sum()))


def test_kitti_threshold_is_exclusive_and_empty_images_are_fine():
    rows = np.zeros((1, 4, 2 * 8 + 7), np.float32)
    rows[0, :, 6:22] = np.tile(np.array([10, 10, 30, 12, 40, 30, 35, 50, 20, 55, 8, 40, 5, 25, 7, 15], np.float32), (4, 1))
    rows[0, :, 4] = [0.3, 0.3, 0.29, 0.31]
    dets = torch.from_numpy(rows).cuda()
    t = float(np.float32(0.3))      # the score as the reference sees it: the double of the float32 (0.30000001192...)
    assert int(RF.rasterise_fill_instances(dets, 64, 64, t, inclusive=True).count[0]) == 3
    assert int(RF.rasterise_fill_instances(dets, 64, 64, t, inclusive=False).count[0]) == 1
    assert int(RF.rasterise_fill_instances(dets, 64, 64, 0.3, inclusive=False).count[0]) == 3     # 0.3 < float32(0.3)
    r = RF.rasterise_fill_instances(dets, 64, 64, 0.9, inclusive=False)
    assert int(r.count[0]) == 0 and r.slot_det[0].cpu().tolist() == [-1, -1, -1, -1]
