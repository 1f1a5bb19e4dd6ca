"""CPU restatement of the mask construction inside the reference's IDD / KITTI writers - TEST INFRASTRUCTURE ONLY (the
product never imports this file).

    IDD.format_and_write_to_IDD            /root/reference/src/lib/datasets/dataset/IDD.py:123-170
    KITTIPOLY.format_and_write_to_kitti    /root/reference/src/lib/datasets/dataset/kitti_poly.py:95-136

Both walk the classes of one image in dict order; inside a class they (1) keep detections with `score >= thresh` (IDD,
:141) or `score > thresh` (KITTI, :113), name their mask files with a counter that runs over the classes in detection
order, (2) draw them in stable depth order: `polygon(points, outline=0, fill=255)` on a fresh 'L' image, multiplied by the
class's running `to_remove_mask`, and (3) let instances with `score >= 0.5` clear `to_remove_mask` with
`polygon(points, outline=0, fill=0)`.  The points are `int(float("{:.2f}".format(v)))` of the decoded vertices.

What Pillow draws (third-party dependency, Pillow 12.2.0 in the build container; its C source is not in /root/reference):
  * `fill=255, outline=0`: `ImagingDrawPolygon(fill=1)` = `polygon_generic` below, then - because the outline ink differs
    from the fill ink - `ImagingDrawPolygon(fill=0)` = one `line8` Bresenham per edge (end point excluded) with ink 0;
  * `fill=0, outline=0`: the inks are equal, ImageDraw skips the outline: the bare `polygon_generic` fill.
`polygon_generic` is restated from its published algorithm and pinned against the real Pillow by
tests/test_oracle_raster_fill.py (goldens made by tests/golden/make_idd_raster_golden.py from the unmodified reference
writers): edge table with float32 `dx`, per scan line the float32 intersections of the edges with `ymin <= y <= ymax`,
an end-of-edge duplicate on rows below the polygon's last, sorted pairs filled from `floor(x0 + 0.5)` to
`ceil(x1 - 0.5)` with Pillow's `x_pos` overlap rule, horizontal edges drawn as spans, and the "connect discontiguous
corners" patch: where two edges of the same slope sign end in a common vertex on a row and both continue to the adjacent
row (the previous one on the polygon's last row, else the next), the vertex's span is extended up to the pixel before
the adjacent row's first pixel (`floor(min + 0.5) - 1`) or from `floor(max + 1.5)` on the other side.  Known residue:
0.2 % of a corpus of 3000 random, spiky star polygons differ from Pillow by one pixel at doubly degenerate vertices (the
ring visits the same point twice, or runs back along its own edge); none of the committed goldens does.
"""
import math

import numpy as np

f32 = np.float32


def round_up(f):
    f = f32(f)
    return int(math.floor(f + f32(0.5))) if f >= 0 else int(-math.floor(abs(f) + f32(0.5)))


def round_down(f):
    f = f32(f)
    return int(math.ceil(f - f32(0.5))) if f >= 0 else int(-math.ceil(abs(f) - f32(0.5)))


def _edge_x(e, y):
    return f32(f32(y - e["y0"]) * e["dx"] + f32(e["x0"]))


def _ends_at(e, x, y):
    return (e["x0"] == x and e["y0"] == y) or (e["x1"] == x and e["y1"] == y)


def pil_polygon_fill(pts, w, h):
    """Boolean (h, w) mask of `ImageDraw.polygon(pts, fill=ink)` without outline (Pillow's polygon_generic)."""
    m = np.zeros((h, w), bool)
    n = len(pts)
    edges = []
    for i in range(n):
        x0, y0 = pts[i]
        x1, y1 = pts[(i + 1) % n]
        if (x0, y0) == (x1, y1) and i == n - 1:
            continue                               # Pillow closes the ring only when last != first
        ymin, ymax = (y0, y1) if y0 <= y1 else (y1, y0)
        dx = f32(0) if y0 == y1 else f32(f32(x1 - x0) / f32(y1 - y0))
        edges.append(dict(x0=x0, y0=y0, x1=x1, y1=y1, xmin=min(x0, x1), xmax=max(x0, x1), ymin=ymin, ymax=ymax, dx=dx))

    def hline(xa, y, xb):
        if 0 <= y < h:
            xa = max(xa, 0)
            xb = min(xb, w - 1)
            if xa <= xb:
                m[y, xa:xb + 1] = True

    table = []
    gymin, gymax = h - 1, 0
    for e in edges:
        gymin = min(gymin, e["ymin"])
        gymax = max(gymax, e["ymax"])
        if e["ymin"] == e["ymax"]:
            hline(e["xmin"], e["ymin"], e["xmax"])
            continue
        table.append(e)
    for y in range(max(gymin, 0), min(gymax, h) + 1):
        xx, owner = [], []
        for i, cur in enumerate(table):
            if not (cur["ymin"] <= y <= cur["ymax"]):
                continue
            xv = _edge_x(cur, y)
            xx.append(xv)
            owner.append(i)
            if y == cur["ymax"] and y < gymax:
                xx.append(xv)                     # "needed to draw consistent polygons"
                owner.append(-1)
                continue
            if cur["dx"] != 0 and f32(np.round(xv)) == xv:
                for jx in range(len(xx) - 1):     # connect discontiguous corners
                    k = owner[jx]
                    if k < 0 or k == i:
                        continue
                    oth = table[k]
                    if (cur["dx"] > 0 and oth["dx"] <= 0) or (cur["dx"] < 0 and oth["dx"] >= 0):
                        continue
                    if xv != _edge_x(oth, y):
                        continue
                    if not (_ends_at(cur, xv, y) and _ends_at(oth, xv, y)):
                        continue                # two edges that merely cross at an integer x are not a corner
                    ya = y - 1 if y == gymax else y + 1
                    if not (cur["ymin"] <= ya <= cur["ymax"] and oth["ymin"] <= ya <= oth["ymax"]):
                        continue
                    a, b = _edge_x(cur, ya), _edge_x(oth, ya)
                    lo, hi = min(a, b), max(a, b)
                    if round_up(lo) - 1 > xv:
                        xx[jx] = f32(round_up(lo) - 1)
                    elif round_up(f32(hi + 1)) < xv:
                        xx[jx] = f32(round_up(f32(hi + 1)))
                    break
        s = sorted(xx)
        x_pos = 0 if s else -1
        for i in range(1, len(s), 2):
            x_end = round_down(s[i])
            if x_end < x_pos:
                continue
            x_start = round_up(s[i - 1])
            if x_pos > x_start:
                x_start = x_pos
                if x_end < x_start:
                    continue
            hline(x_start, y, x_end)
            x_pos = x_end + 1
    return m


def line8(x0, y0, x1, y1):
    """Pillow's 8-bit Bresenham: the points from (x0, y0) towards (x1, y1), the end point NOT included."""
    pts = []
    dx, xs = x1 - x0, 1
    if dx < 0:
        dx, xs = -dx, -1
    dy, ys = y1 - y0, 1
    if dy < 0:
        dy, ys = -dy, -1
    if dx == 0:
        for _ in range(dy):
            pts.append((x0, y0)); y0 += ys
    elif dy == 0:
        for _ in range(dx):
            pts.append((x0, y0)); x0 += xs
    elif dx > dy:
        n = dx; dy += dy; e = dy - dx; dx += dx
        for _ in range(n):
            pts.append((x0, y0))
            if e >= 0:
                y0 += ys; e -= dx
            e += dy; x0 += xs
    else:
        n = dy; dx += dx; e = dx - dy; dy += dy
        for _ in range(n):
            pts.append((x0, y0))
            if e >= 0:
                x0 += xs; e -= dy
            e += dx; y0 += ys
    return pts


def pil_polygon_outline(pts, w, h):
    m = np.zeros((h, w), bool)
    n = len(pts)
    for i in range(n):
        a, b = pts[i], pts[(i + 1) % n]
        for x, y in line8(a[0], a[1], b[0], b[1]):
            if 0 <= x < w and 0 <= y < h:
                m[y, x] = True
    return m


def to_points(vals):
    """`int(float("{:.2f}".format(v)))` per coordinate (IDD.py:62-63,143,159-160)."""
    v = [int(float("{:.2f}".format(float(x)))) for x in vals]
    return list(zip(v[0::2], v[1::2]))


def writer_masks(all_bboxes_image, w, h, thresh, inclusive, occluder_score=0.5):
    """One image.  `all_bboxes_image`: {1-based class: [[box4, score, 2N vertices, depth], ...]} (polydet_post_process layout).
    Returns the instances in FILE order (the counter of the reference): [(class index, score, mask bool (h, w))]."""
    out = {}
    count = 0
    for cls_ind in all_bboxes_image:
        to_remove = np.ones((h, w), bool)
        params = []
        for bbox in all_bboxes_image[cls_ind]:
            score = bbox[4]
            if (score >= thresh) if inclusive else (score > thresh):
                params.append((to_points(bbox[5:-1]), count, score, bbox[-1]))
                count += 1
        for pts, idx, score, depth in sorted(params, key=lambda t: t[-1]):
            fill = pil_polygon_fill(pts, w, h)
            mask = fill & ~pil_polygon_outline(pts, w, h) & to_remove
            if float(score) >= occluder_score:
                to_remove &= ~fill
            out[idx] = (cls_ind, score, mask)
    return [out[i] for i in range(count)]
