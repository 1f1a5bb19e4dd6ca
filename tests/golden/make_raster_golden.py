"""Golden vectors for the instance rasterisation, produced by the UNMODIFIED reference method
``Cityscapes.format_and_write_to_cityscapes`` (/root/reference/src/lib/datasets/dataset/cityscapes.py:196-283) run in
this container:  python tests/golden/make_raster_golden.py

The module imports packages that are absent here and that the method never touches (pycocotools, shapely, wandb,
matplotlib, seaborn): those are stubbed with empty modules.  ``bresenham`` (PyPI, absent) is the one stub that IS on
the path: it is replaced by oracle/raster_oracle.bresenham, the restatement of that package's published generator -
so these goldens pin everything but that generator to the reference's own code and to the real PIL.
The method hard-codes 2048x1024 masks; the PNGs it writes are read back and stored bit-packed (np.packbits) together
with the text file lines.  Inputs: ``all_bboxes`` in the layout polydet_post_process returns."""
import json
import os
import sys
import tempfile
import types

import numpy as np
from PIL import Image

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import raster_oracle  # noqa: E402

N = 32


def load_reference():
    for name in ("pycocotools", "pycocotools.coco", "pycocotools.cocoeval", "shapely", "shapely.geometry", "wandb", "matplotlib",
                 "matplotlib.pyplot", "seaborn", "cv2"):
        if name not in sys.modules:
            try:
                __import__(name)
            except Exception:
                sys.modules[name] = types.ModuleType(name)
    sys.modules["pycocotools.cocoeval"].__dict__.setdefault("COCOeval", object)
    sys.modules["shapely.geometry"].__dict__.setdefault("Polygon", object)
    b = types.ModuleType("bresenham")
    b.bresenham = raster_oracle.bresenham
    sys.modules["bresenham"] = b
    sys.path.insert(0, "/root/reference/src/lib")
    from datasets.dataset.cityscapes import CITYSCAPES
    return CITYSCAPES


def scene(seed, n_det, kind):
    """(K, 2N+7) decode-layout rows in image coordinates: box4, score, class(0-based), 2N vertices, depth."""
    rng = np.random.default_rng(seed)
    rows = np.zeros((n_det, 2 * N + 7), np.float32)
    for i in range(n_det):
        cx, cy = rng.uniform(-50, 2100), rng.uniform(300, 1000)
        size = rng.uniform(5, 260) if kind != "tiny" else rng.uniform(1, 12)
        ang = np.linspace(0, 2 * np.pi, N, endpoint=False) + rng.uniform(0, 0.2)
        rad = size * rng.uniform(0.6, 1.0, N)
        if kind == "tangled" and i % 3 == 0:
            ang = rng.permutation(ang)          # self-intersecting
        x, y = cx + 1.6 * rad * np.cos(ang), cy + rad * np.sin(ang)
        if kind == "tangled" and i % 5 == 0:
            x = np.round(x)                     # integers and x.995+ fractions exercise the 2-decimal rounding
            y = np.floor(y) + 0.996
        rows[i, 6::2][:N], rows[i, 7::2][:N] = x, y
        rows[i, 0:4] = x.min(), y.min(), x.max(), y.max()
        rows[i, 4] = rng.choice([0.03, 0.05, 0.2, 0.49, 0.5, 0.7, 0.95]) if i % 4 == 0 else rng.uniform(0, 1)
        rows[i, 5] = rng.integers(0, 11)
        rows[i, -1] = rng.choice([0.25, 0.5]) if i % 6 == 0 else rng.uniform(0, 1)   # ties in depth
    return rows


def to_all_bboxes(rows, num_classes=11):
    """polydet_post_process layout (utils/post_process.py:113-121): {1-based class: [[box4, score, 2N, depth], ...]}"""
    out = {}
    cls = rows[:, 5].astype(int)
    for j in range(num_classes):
        sel = rows[cls == j]
        out[j + 1] = np.concatenate([sel[:, :5], sel[:, 6:]], axis=1).astype(np.float32).tolist()
    return out


def run_reference(CITYSCAPES, rows, thresh):
    self = CITYSCAPES.__new__(CITYSCAPES)
    self.class_name = ['__background__', 'person', 'rider', 'car', 'truck', 'bus', 'train', 'motorcycle', 'bicycle', 'pole',
                       'traffic sign', 'traffic light']
    self.label_to_id = {'person': 24, 'rider': 25, 'car': 26, 'truck': 27, 'bus': 28, 'train': 31, 'motorcycle': 32, 'bicycle': 33,
                        'pole': -1, 'traffic sign': -1, 'traffic light': -1}
    self.opt = types.SimpleNamespace(thresh=thresh)
    with tempfile.TemporaryDirectory() as tmp:
        self.annot_path = os.path.join(tmp, "anno.json")
        json.dump({"images": [{"id": 7, "file_name": "frankfurt/f_000007_leftImg8bit.png"}]}, open(self.annot_path, "w"))
        save = os.path.join(tmp, "out")
        os.mkdir(save)
        self.format_and_write_to_cityscapes({7: to_all_bboxes(rows)}, save)
        lines = open(os.path.join(save, "f_000007_leftImg8bit.txt")).read().splitlines()
        masks, ids, confs = [], [], []
        for ln in lines:
            path, label_id, conf = ln.split(" ")
            masks.append(np.array(Image.open(os.path.join(save, path))))
            ids.append(int(label_id))
            confs.append(float(conf))
    masks = np.stack(masks) if masks else np.zeros((0, 1024, 2048), np.uint8)
    assert set(np.unique(masks)) <= {0, 255}
    return masks, np.array(ids), np.array(confs, np.float64)


def main():
    CITYSCAPES = load_reference()
    for name, (seed, n_det, kind, thresh) in {"raster_street": (11, 40, "street", 0.05), "raster_tangled": (12, 30, "tangled", 0.05),
                                              "raster_tiny": (13, 24, "tiny", 0.3)}.items():
        rows = scene(seed, n_det, kind)
        masks, ids, confs = run_reference(CITYSCAPES, rows, thresh)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), dets=rows, thresh=np.float64(thresh), label_ids=ids, confs=confs,
                            packed=np.packbits(masks > 0, axis=-1), areas=(masks > 0).sum((1, 2)))
        print(name, "written instances:", len(ids), "areas", (masks > 0).sum((1, 2))[:8])


if __name__ == "__main__":
    main()
