"""The polygon-IoU oracle (oracle/polyloss_oracle.py) against outputs of the reference's own PolyLoss
(tests/golden/polyiou_*.npz, made by tests/golden/make_polyiou_golden.py)."""
import glob
import os

import numpy as np
import pytest

from oracle import polyloss_oracle as po

GOLD = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "polyiou_*.npz")))


@pytest.mark.parametrize("path", GOLD, ids=[os.path.basename(p)[:-4] for p in GOLD])
def test_oracle_matches_reference(path):
    z = np.load(path)
    got = po.poly_iou_loss_oracle(z["pred"], z["mask"], z["target"])
    assert (got["status"] == 0).all()
    assert abs(float(got["loss"]) - float(z["loss"])) <= 2e-6 * max(1.0, abs(float(z["loss"])))
    scale = np.abs(z["grad"]).max()
    assert np.abs(got["grad"] - z["grad"]).max() <= 1e-4 * scale
    # unmasked rows never receive gradient
    assert np.all(got["grad"][z["mask"] == 0] == 0)


def test_goldens_present():
    assert len(GOLD) >= 5


def test_area_extra_wrap_term():
    """losses.py:25-41 sums rows 0..n of the doubled polygon: closed shoelace + the first edge once more."""
    import torch
    sq = torch.tensor([[1.0, 0.0], [1.0, np.pi / 2], [1.0, np.pi], [1.0, 3 * np.pi / 2]])
    # closed shoelace of this diamond = 2.0; extra term cross(p0,p1)/2 = 0.5
    assert abs(float(po.area(sq)) - 2.5) < 1e-6


def test_walk_statuses():
    assert po.walk([], [], 8, 8) == (po.STATUS_OK, [])
    st, _ = po.walk([(1, 2)], [], 8, 8)
    assert st == po.STATUS_NO_OUTBOUND
    # an in-bound whose clip index can never be reached again after deletion -> the reference would spin
    st, _ = po.walk([(1, 2), (5, 6)], [(3, 4)], 8, 8)
    assert st in (po.STATUS_OK, po.STATUS_RUNAWAY)
