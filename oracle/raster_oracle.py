"""CPU restatement of the reference's depth-ordered polygon rasterisation — TEST INFRASTRUCTURE ONLY (imported by
tests/ and by the CPU leg of tools/raster_bench.py; never by the product path).

Follows /root/reference/src/lib/datasets/dataset/cityscapes.py:196-283 (``format_and_write_to_cityscapes``) statement
by statement, with the file / annotation I/O removed: score filter :228, ``int()`` truncation of the vertices :232,
stable sort by pseudo-depth :239 (vertices first pass through ``_to_float``, :120-121,
231 - rounding to two decimals BEFORE the truncation, so 12.996 becomes 13), PIL polygon fill :253, Bresenham contour + radius-2 discs :256-264, occlusion by the
masks of earlier instances with score >= 0.5 :267-271, area > 100 filter and the min(1, 1.2*score) confidence :273-279.

Third-party pieces: ``PIL.ImageDraw`` is used as is (the reference does the same; Pillow is present in this image).
``bresenham`` (PyPI package "bresenham", unpinned by the reference and ABSENT here) is restated below from its published
algorithm (an integer error-accumulating line walk that yields both end points) - so this oracle is pinned to the
reference for everything except that one generator, which tests/test_oracle_raster.py checks against its closed form
and known-answer lines."""
import numpy as np
from PIL import Image, ImageDraw

NOT_DRAWN = ("pole", "traffic sign", "traffic light")   # cityscapes.py:242,273


def _to_float(x):
    """cityscapes.py:120-121"""
    return float("{:.2f}".format(x))


def bresenham(x0, y0, x1, y1):
    """PyPI ``bresenham`` 0.2.1: yields every integer point of the line from (x0, y0) to (x1, y1), both included."""
    dx = x1 - x0
    dy = y1 - y0
    xsign = 1 if dx > 0 else -1
    ysign = 1 if dy > 0 else -1
    dx = abs(dx)
    dy = abs(dy)
    if dx > dy:
        xx, xy, yx, yy = xsign, 0, 0, ysign
    else:
        dx, dy = dy, dx
        xx, xy, yx, yy = 0, ysign, xsign, 0
    D = 2 * dy - dx
    y = 0
    for x in range(dx + 1):
        yield x0 + x * xx + y * yx, y0 + x * xy + y * yy
        if D >= 0:
            y += 1
            D -= 2 * dx
        D += 2 * dy


def instance_mask(points, width, height):
    """cityscapes.py:241-264: filled polygon plus a radius-2 disc on every contour point (uint8, 0 / 255)."""
    polygon_mask = Image.new('L', (width, height), 0)
    ImageDraw.Draw(polygon_mask).polygon(points, outline=255, fill=255)
    contour = list(bresenham(points[-1][0], points[-1][1], points[0][0], points[0][1]))
    for i in range(len(points) - 1):
        contour += bresenham(points[i][0], points[i][1], points[i + 1][0], points[i + 1][1])
    radius = 2
    draw = ImageDraw.Draw(polygon_mask)
    for point in set(contour):
        draw.ellipse([(point[0] - radius, point[1] - radius), (point[0] + radius, point[1] + radius)], outline=255, fill=255)
    return np.array(polygon_mask)


def rasterise_oracle(all_bboxes, class_name, thresh, width=2048, height=1024):
    """``all_bboxes`` = {class index (1-based): rows [x1, y1, x2, y2, score, x0, y0, ..., depth]} of ONE image, as produced
    by polydet_post_process.  Returns the list of (mask uint8 HxW, label, confidence) the reference would write, in
    order, plus the per-instance records (for tests): (points, score, label, depth, visible mask, kept)."""
    param_list = []
    for cls_ind in all_bboxes:
        if cls_ind == 'fg':
            continue
        for bbox in all_bboxes[cls_ind]:
            if bbox[4] > thresh:
                depth = bbox[-1]
                label = class_name[cls_ind]
                polygon = list(map(_to_float, bbox[5:-1]))
                polygon = [(int(x), int(y)) for x, y in zip(polygon[0::2], polygon[1::2])]
                param_list.append((polygon, bbox[4], label, depth))
    ones = np.ones((height, width))
    to_remove_mask = np.zeros((height, width))
    written, records = [], []
    for points, score, label, depth in sorted(param_list, key=lambda x: x[-1]):
        polygon_mask = np.zeros((height, width), dtype=np.uint8)
        if label not in NOT_DRAWN:
            polygon_mask = instance_mask(points, width, height)
            polygon_mask = polygon_mask * (ones - to_remove_mask).astype(np.uint8)
        if score >= 0.5:
            to_remove_mask += polygon_mask
            to_remove_mask[to_remove_mask > 0] = 1
        kept = label not in NOT_DRAWN and np.count_nonzero(polygon_mask) > 100
        if kept:
            written.append((polygon_mask, label, min(1, score * 1.2)))
        records.append((points, score, label, depth, polygon_mask, kept))
    return written, records


# ---- the closed form the CUDA kernel evaluates (centerpoly_b200/csrc/raster.cu), in numpy ---------------------------
# Not part of the reference: kept here so the CPU suite can prove, against PIL, the identity the kernel relies on:
#   PIL fill+outline  U  discs on the Bresenham contour  ==  even-odd(pixel centres)  U  21-px discs on the closed-form steps
def round2_trunc(x):
    """int(_to_float(x)) without the string round trip: the fraction rounds up to 1.00 iff it exceeds 0.995."""
    a = np.abs(np.asarray(x, dtype=np.float64))
    fl = np.floor(a)
    n = fl + ((a - fl) > 0.995)
    return (np.sign(x) * n).astype(np.int64)


def closed_form_mask(points, width, height):
    pts = np.asarray(points, dtype=np.int64)
    ys, xs = np.mgrid[0:height, 0:width]
    inside = np.zeros((height, width), bool)
    near = np.zeros((height, width), bool)
    for i in range(len(pts)):
        x0, y0 = pts[i - 1]
        x1, y1 = pts[i]
        dx, dy = x1 - x0, y1 - y0
        if dy != 0:
            lhs = (ys - y0) * dx - (xs - x0) * dy
            inside ^= ((y0 <= ys) != (y1 <= ys)) & (lhs > 0 if dy > 0 else lhs < 0)
        adx, ady = abs(dx), abs(dy)
        sx, sy = (1 if dx > 0 else -1), (1 if dy > 0 else -1)
        for o in range(-2, 3):          # offset of the pixel from the contour point along the major axis
            lim = 1 if abs(o) == 2 else 2
            if adx > ady:
                k = (xs - o - x0) * sx
                m = (2 * ady * k + adx) // (2 * adx)
                near |= (k >= 0) & (k <= adx) & (np.abs(ys - (y0 + sy * m)) <= lim)
            else:
                k = (ys - o - y0) * sy
                m = (2 * adx * k + ady) // (2 * ady) if ady else np.zeros_like(k)
                near |= (k >= 0) & (k <= ady) & (np.abs(xs - (x0 + sx * m)) <= lim)
    return ((inside | near) * 255).astype(np.uint8)
