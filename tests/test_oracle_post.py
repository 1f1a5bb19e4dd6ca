"""CPU: the post-process oracle and the host mirror against golden vectors produced by the UNMODIFIED reference
``polydet_post_process`` (src/lib/utils/post_process.py:105-122; tests/golden/make_golden.py::make_post)."""
import glob
import os

import numpy as np
import pytest

from oracle.postprocess_oracle import get_affine_transform, polydet_post_process_oracle

GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "post_*.npz")))


def load(path):
    d = np.load(path)
    B = d["dets"].shape[0]
    c = [d["c"][i] for i in range(B)]
    s = [float(d["s"][i][0]) if bool(d["s_is_scalar"]) else d["s"][i] for i in range(B)]
    return d, c, s, int(d["out_h"]), int(d["out_w"]), int(d["num_classes"])


def check(ret, d, nc, tol):
    D = d["dets"].shape[2] - 1
    for i in range(d["dets"].shape[0]):
        for j in range(1, nc + 1):
            got = np.array(ret[i][j], dtype=np.float32).reshape(-1, D)
            want = d[f"ret_{i}_{j}"]
            assert got.shape == want.shape          # same detections in the same order per class
            # columns that are never transformed are bit-exact: score and pseudo-depth
            assert np.array_equal(got[:, 4], want[:, 4]) and np.array_equal(got[:, -1], want[:, -1])
            np.testing.assert_allclose(got, want, rtol=tol, atol=0)


@pytest.mark.parametrize("path", GOLDEN, ids=lambda p: os.path.basename(p)[:-4])
def test_oracle_matches_reference(path):
    d, c, s, h, w, nc = load(path)
    for i in range(len(c)):   # the restated cv2.getAffineTransform: same matrix as the reference's to fp64 round-off
        np.testing.assert_allclose(get_affine_transform(c[i], s[i], 0, (w, h), inv=1), d["trans"][i], rtol=0, atol=1e-12)
    check(polydet_post_process_oracle(d["dets"], c, s, h, w, nc), d, nc, 1e-7)   # 1 fp32 ulp


@pytest.mark.parametrize("path", GOLDEN, ids=lambda p: os.path.basename(p)[:-4])
def test_host_mirror_matches_reference(path):
    import centerpoly_b200.post_process as pp   # host logic only: no CUDA needed
    d, c, s, h, w, nc = load(path)
    np.testing.assert_allclose(pp.batch_inverse_affine(c, s, w, h), d["trans"], rtol=0, atol=1e-12)
    check(pp.polydet_post_process(d["dets"], c, s, h, w, nc), d, nc, 1e-7)


def test_class_outside_range_is_dropped():
    import centerpoly_b200.post_process as pp
    dets = np.zeros((1, 3, 11), dtype=np.float32)
    dets[0, :, 5] = [0, 1, 7]
    ret = pp.polydet_post_process(dets, [np.array([8, 8], np.float32)], [16.0], 4, 4, 2)
    assert [len(ret[0][j]) for j in (1, 2)] == [1, 1]
    want = polydet_post_process_oracle(dets, [np.array([8, 8], np.float32)], [16.0], 4, 4, 2)
    assert ret == want
