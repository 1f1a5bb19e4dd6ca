"""CPU: the restatement of the IDD / KITTI mask writers (oracle/raster_fill_oracle.py: Pillow's polygon_generic fill, line8
outline, per-class depth-ordered removal) against goldens produced by the UNMODIFIED reference writers with the real Pillow
(tests/golden/make_idd_raster_golden.py)."""
import os

import numpy as np
import pytest

from oracle import raster_fill_oracle as R

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = ["fillraster_idd_street", "fillraster_idd_tiny", "fillraster_kitti_street"]
IDD_NAMES = ['__background__', 'person', 'rider', 'motorcycle', 'bicycle', 'autorickshaw', 'car', 'truck', 'bus', 'vehicle fallback']
IDD_IDS = {'person': 6, 'rider': 8, 'motorcycle': 9, 'bicycle': 10, 'autorickshaw': 11, 'car': 12, 'truck': 13, 'bus': 14, 'vehicle fallback': 18}
KITTI_NAMES = ['__background__', 'person', 'rider', 'car', 'truck', 'bus', 'train', 'motorcycle', 'bicycle']
KITTI_IDS = {'person': 24, 'rider': 25, 'car': 26, 'truck': 27, 'bus': 28, 'train': 31, 'motorcycle': 32, 'bicycle': 33}


def to_all_bboxes(rows, num_classes):
    out = {}
    cls = rows[:, 5].astype(int)
    for j in range(num_classes):
        sel = rows[cls == j]
        out[j + 1] = np.concatenate([sel[:, :5], sel[:, 6:]], axis=1).astype(np.float32).tolist()
    return out


def load(name):
    g = np.load(os.path.join(GOLD, name + ".npz"))
    w, h = (int(v) for v in g["size"])
    which = str(g["which"])
    names, ids = (IDD_NAMES, IDD_IDS) if which == "idd" else (KITTI_NAMES, KITTI_IDS)
    want = np.unpackbits(g["packed"], axis=-1)[:, :, :w].astype(bool)
    return g, w, h, which, names, ids, want


@pytest.mark.parametrize("name", CASES)
def test_writer_oracle_matches_reference_writers(name):
    g, w, h, which, names, ids, want = load(name)
    inst = R.writer_masks(to_all_bboxes(g["dets"], len(names) - 1), w, h, float(g["thresh"]), inclusive=(which == "idd"))
    assert len(inst) == len(want) > 10
    for i, (c, score, m) in enumerate(inst):
        assert ids[names[c]] == int(g["label_ids"][i])
        assert str(score) == str(g["scores"][i])
        assert np.array_equal(m, want[i]), (name, i, int((m != want[i]).sum()))
    assert [int(m.sum()) for _, _, m in inst] == [int(a) for a in g["areas"]]


def test_line8_excludes_the_end_point_and_matches_known_walks():
    assert R.line8(0, 0, 4, 0) == [(0, 0), (1, 0), (2, 0), (3, 0)]
    assert R.line8(0, 0, 0, -3) == [(0, 0), (0, -1), (0, -2)]
    assert R.line8(0, 0, 5, 2) == [(0, 0), (1, 0), (2, 1), (3, 1), (4, 2)]
    assert R.line8(3, 3, 1, -2) == [(3, 3), (3, 2), (2, 1), (2, 0), (1, -1)]
    assert R.line8(2, 2, 2, 2) == []


def test_fill_rounding_and_corner_rules():
    # spans include pixels within half a pixel of the edge, ties excluded on both sides (Pillow's ROUND_UP / ROUND_DOWN)
    # (edges at x = 34.5 and 35.5 on row 4: the pixels exactly half a pixel away, 34 and 36, stay empty)
    tri = R.pil_polygon_fill([(35, 3), (55, 43), (15, 43)], 64, 48)
    assert np.nonzero(tri[4])[0].tolist() == [35] and np.nonzero(tri[5])[0].tolist() == [34, 35, 36]
    # a bottom vertex whose row above ends short of it is connected (Pillow's "discontiguous corners" patch)
    m = R.pil_polygon_fill([(203, 37), (16, 36), (17, 22)], 220, 60)
    assert np.nonzero(m[36])[0].tolist() == list(range(16, 192)) and np.nonzero(m[37])[0].tolist() == list(range(192, 204))
    # ... and is left alone when the rows already touch
    m = R.pil_polygon_fill([(185, 46), (105, 132), (126, 22)], 220, 140)
    assert np.nonzero(m[132])[0].tolist() == [105] and np.nonzero(m[131])[0].tolist() == [105, 106]
