"""Golden vectors for the IDD / KITTI mask writers, produced by the UNMODIFIED reference methods
``IDD.format_and_write_to_IDD`` (/root/reference/src/lib/datasets/dataset/IDD.py:123-170) and
``KITTIPOLY.format_and_write_to_kitti`` (kitti_poly.py:95-136) run in this container with the real Pillow:

    python tests/golden/make_idd_raster_golden.py

Packages the modules import but the methods never touch (pycocotools, cv2, ...) are stubbed with empty modules.  The
methods read the image size with ``Image.open(file_name).size``: a blank PNG of the wanted size is written to a temporary
directory and named in a temporary annotation file.  The PNG masks the methods write are read back and stored bit-packed
together with the text-file lines.  Inputs: ``all_bboxes`` in the layout polydet_post_process returns."""
import json
import os
import sys
import tempfile
import types

import numpy as np
from PIL import Image

HERE = os.path.dirname(os.path.abspath(__file__))
N = 32


def load_reference():
    for name in ("pycocotools", "pycocotools.coco", "pycocotools.cocoeval", "cv2"):
        if name not in sys.modules:
            try:
                __import__(name)
            except Exception:
                sys.modules[name] = types.ModuleType(name)
    sys.modules["pycocotools.cocoeval"].__dict__.setdefault("COCOeval", object)
    sys.path.insert(0, "/root/reference/src/lib")
    from datasets.dataset.IDD import IDD
    from datasets.dataset.kitti_poly import KITTIPOLY
    return IDD, KITTIPOLY


def scene(seed, n_det, n_classes, w, h, kind):
    """(K, 2N+7) decode-layout rows in image coordinates: box4, score, class (0-based), 2N vertices, depth."""
    rng = np.random.default_rng(seed)
    rows = np.zeros((n_det, 2 * N + 7), np.float32)
    for i in range(n_det):
        cx, cy = rng.uniform(-30, w + 30), rng.uniform(0.3 * h, 0.95 * h)
        size = rng.uniform(4, 0.14 * w) if kind != "tiny" else rng.uniform(1, 9)
        ang = np.linspace(0, 2 * np.pi, N, endpoint=False) + rng.uniform(0, 0.2)
        rad = size * rng.uniform(0.75, 1.0, N)
        x, y = cx + 1.5 * rad * np.cos(ang), cy + rad * np.sin(ang)
        if i % 5 == 0:
            x = np.round(x)                     # integers and x.995+ fractions exercise the 2-decimal rounding
            y = np.floor(y) + 0.996
        rows[i, 6::2][:N], rows[i, 7::2][:N] = x, y
        rows[i, 0:4] = x.min(), y.min(), x.max(), y.max()
        rows[i, 4] = rng.choice([0.03, 0.3, 0.49, 0.5, 0.7, 0.95]) if i % 4 == 0 else rng.uniform(0, 1)
        rows[i, 5] = rng.integers(0, n_classes)
        rows[i, -1] = rng.choice([0.25, 0.5]) if i % 6 == 0 else rng.uniform(0, 1)   # ties in depth
    return rows


def to_all_bboxes(rows, num_classes):
    """polydet_post_process layout (utils/post_process.py:113-121): {1-based class: [[box4, score, 2N, depth], ...]}"""
    out = {}
    cls = rows[:, 5].astype(int)
    for j in range(num_classes):
        sel = rows[cls == j]
        out[j + 1] = np.concatenate([sel[:, :5], sel[:, 6:]], axis=1).astype(np.float32).tolist()
    return out


IDD_NAMES = ['__background__', 'person', 'rider', 'motorcycle', 'bicycle', 'autorickshaw', 'car', 'truck', 'bus', 'vehicle fallback']
IDD_IDS = {'person': 6, 'rider': 8, 'motorcycle': 9, 'bicycle': 10, 'autorickshaw': 11, 'car': 12, 'truck': 13, 'bus': 14, 'vehicle fallback': 18}
KITTI_NAMES = ['__background__', 'person', 'rider', 'car', 'truck', 'bus', 'train', 'motorcycle', 'bicycle']
KITTI_IDS = {'person': 24, 'rider': 25, 'car': 26, 'truck': 27, 'bus': 28, 'train': 31, 'motorcycle': 32, 'bicycle': 33}


def run_reference(cls, which, rows, thresh, w, h):
    self = cls.__new__(cls)
    self.class_name, self.label_to_id = (IDD_NAMES, IDD_IDS) if which == "idd" else (KITTI_NAMES, KITTI_IDS)
    self.opt = types.SimpleNamespace(thresh=thresh)
    with tempfile.TemporaryDirectory() as tmp:
        os.makedirs(os.path.join(tmp, "img", "seq7"))
        png = os.path.join(tmp, "img", "seq7", "frame0007.png")
        Image.new("L", (w, h), 0).save(png)
        self.annot_path = os.path.join(tmp, "anno.json")
        json.dump({"images": [{"id": 7, "file_name": png}]}, open(self.annot_path, "w"))
        save = os.path.join(tmp, "out")
        os.mkdir(save)
        boxes = {7: to_all_bboxes(rows, len(self.class_name) - 1)}
        if which == "idd":
            self.format_and_write_to_IDD(boxes, save)
            out_dir = os.path.join(save, "seq7")
        else:
            self.format_and_write_to_kitti(boxes, save)
            out_dir = save
        lines = open(os.path.join(out_dir, "frame0007.txt")).read().splitlines()
        masks, ids, scores = [], [], []
        for ln in lines:
            path, label_id, score = ln.split(" ")
            masks.append(np.array(Image.open(os.path.join(out_dir, path))))
            ids.append(int(label_id))
            scores.append(score)
    masks = np.stack(masks) if masks else np.zeros((0, h, w), np.uint8)
    assert set(np.unique(masks)) <= {0, 255}
    return masks, np.array(ids), np.array(scores)


def main():
    IDD, KITTIPOLY = load_reference()
    cases = {"fillraster_idd_street": (IDD, "idd", 21, 48, 9, 960, 540, "street", 0.3),
             "fillraster_idd_tiny": (IDD, "idd", 22, 30, 9, 480, 270, "tiny", 0.49),
             "fillraster_kitti_street": (KITTIPOLY, "kitti", 23, 40, 8, 1242, 375, "street", 0.3)}
    for name, (cls, which, seed, n_det, n_cls, w, h, kind, thresh) in cases.items():
        rows = scene(seed, n_det, n_cls, w, h, kind)
        masks, ids, scores = run_reference(cls, which, rows, thresh, w, h)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), dets=rows, thresh=np.float64(thresh), size=np.array([w, h]), which=which,
                            label_ids=ids, scores=scores, packed=np.packbits(masks > 0, axis=-1), areas=(masks > 0).sum((1, 2)))
        print(name, "written instances:", len(ids), "areas", (masks > 0).sum((1, 2))[:10])


if __name__ == "__main__":
    main()
