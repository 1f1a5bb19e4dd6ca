"""CPU restatement of the reference's ``polydet_post_process`` — TEST INFRASTRUCTURE ONLY (imported by tests/,
__graft_entry__.smoke() and bench.py's CPU baseline; never by the product path).

Follows /root/reference/src/lib/utils/post_process.py:105-122 and the helpers it calls in
/root/reference/src/lib/utils/image.py: ``transform_preds`` :19-24, ``get_affine_transform`` :27-60 (inv=1),
``affine_transform`` :63-66, ``get_3rd_point`` :69-71, ``get_dir`` :74-81.  The one third-party call,
``cv2.getAffineTransform(dst, src)`` (opencv-python, version unpinned by the reference's requirements.txt), is restated
as the exact solve of the 3-point system in float64 (OpenCV solves the same 6x6 system in double).

Pinned: tests/test_oracle_post.py checks it against tests/golden/post_*.npz, which tests/golden/make_golden.py produced
by running the UNMODIFIED reference function (with the real cv2) in the build container."""
import numpy as np


def get_dir(src_point, rot_rad):          # image.py:74-81
    sn, cs = np.sin(rot_rad), np.cos(rot_rad)
    return [src_point[0] * cs - src_point[1] * sn, src_point[0] * sn + src_point[1] * cs]


def get_3rd_point(a, b):                  # image.py:69-71
    direct = a - b
    return b + np.array([-direct[1], direct[0]], dtype=np.float32)


def solve_affine(src, dst):
    """cv2.getAffineTransform(src, dst): the 2x3 matrix M with M @ [x, y, 1] = dst for the three point pairs."""
    a = np.zeros((6, 6), dtype=np.float64)
    b = np.zeros(6, dtype=np.float64)
    for i in range(3):
        a[2 * i, 0:3] = [src[i, 0], src[i, 1], 1.0]
        a[2 * i + 1, 3:6] = [src[i, 0], src[i, 1], 1.0]
        b[2 * i], b[2 * i + 1] = dst[i, 0], dst[i, 1]
    return np.linalg.solve(a, b).reshape(2, 3)


def get_affine_transform(center, scale, rot, output_size, shift=np.array([0, 0], dtype=np.float32), inv=0):   # image.py:27-60
    if not isinstance(scale, np.ndarray) and not isinstance(scale, list):
        scale = np.array([scale, scale], dtype=np.float32)
    scale_tmp = scale
    src_w = scale_tmp[0]
    dst_w, dst_h = output_size[0], output_size[1]
    rot_rad = np.pi * rot / 180
    src_dir = get_dir([0, src_w * -0.5], rot_rad)
    dst_dir = np.array([0, dst_w * -0.5], np.float32)
    src = np.zeros((3, 2), dtype=np.float32)
    dst = np.zeros((3, 2), dtype=np.float32)
    src[0, :] = center + scale_tmp * shift
    src[1, :] = center + src_dir + scale_tmp * shift
    dst[0, :] = [dst_w * 0.5, dst_h * 0.5]
    dst[1, :] = np.array([dst_w * 0.5, dst_h * 0.5], np.float32) + dst_dir
    src[2:, :] = get_3rd_point(src[0, :], src[1, :])
    dst[2:, :] = get_3rd_point(dst[0, :], dst[1, :])
    if inv:
        return solve_affine(np.float32(dst), np.float32(src))
    return solve_affine(np.float32(src), np.float32(dst))


def affine_transform(pt, t):              # image.py:63-66
    new_pt = np.array([pt[0], pt[1], 1.], dtype=np.float32).T
    return np.dot(t, new_pt)[:2]


def transform_preds(coords, center, scale, output_size):   # image.py:19-24
    target_coords = np.zeros(coords.shape)
    trans = get_affine_transform(center, scale, 0, output_size, inv=1)
    for p in range(coords.shape[0]):
        target_coords[p, 0:2] = affine_transform(coords[p, 0:2], trans)
    return target_coords


def polydet_post_process_oracle(dets, c, s, h, w, num_classes):   # post_process.py:105-122
    dets = dets.copy()
    ret = []
    for i in range(dets.shape[0]):
        top_preds = {}
        dets[i, :, :2] = transform_preds(dets[i, :, 0:2], c[i], s[i], (w, h))
        dets[i, :, 2:4] = transform_preds(dets[i, :, 2:4], c[i], s[i], (w, h))
        for j in range(6, dets.shape[-1] - 1, 2):
            dets[i, :, j:j + 2] = transform_preds(dets[i, :, j:j + 2], c[i], s[i], (w, h))
        classes = dets[i, :, 5]
        for j in range(num_classes):
            inds = (classes == j)
            top_preds[j + 1] = np.concatenate([
                dets[i, inds, :4].astype(np.float32),
                dets[i, inds, 4:5].astype(np.float32),
                dets[i, inds, 6:].astype(np.float32)], axis=1).tolist()
        ret.append(top_preds)
    return ret
