import numpy as np, math, random
from PIL import Image, ImageDraw

def pil_mask(pts, w, h):
    im = Image.new('L', (w, h), 0)
    ImageDraw.Draw(im).polygon(pts, outline=0, fill=255)
    return np.array(im) > 0

def pil_fill_only(pts, w, h):
    im = Image.new('L', (w, h), 0)
    ImageDraw.Draw(im).polygon(pts, fill=255)
    return np.array(im) > 0

def pil_outline_only(pts, w, h):
    im = Image.new('L', (w, h), 0)
    ImageDraw.Draw(im).polygon(pts, outline=255)
    return np.array(im) > 0

def line8(x0, y0, x1, y1):
    pts = []
    dx = x1 - x0; xs = 1
    if dx < 0: dx = -dx; xs = -1
    dy = y1 - y0; ys = 1
    if dy < 0: dy = -dy; ys = -1
    if dx == 0:
        for i in range(dy): pts.append((x0, y0)); y0 += ys
    elif dy == 0:
        for i in range(dx): pts.append((x0, y0)); x0 += xs
    elif dx > dy:
        n = dx; dy += dy; e = dy - dx; dx += dx
        for i in range(n):
            pts.append((x0, y0))
            if e >= 0: y0 += ys; e -= dx
            e += dy; x0 += xs
    else:
        n = dy; dx += dx; e = dx - dy; dy += dy
        for i in range(n):
            pts.append((x0, y0))
            if e >= 0: x0 += xs; e -= dy
            e += dx; y0 += ys
    return pts

def my_outline(pts, w, h):
    m = np.zeros((h, w), bool)
    n = len(pts)
    for i in range(n):
        a = pts[i]; b = pts[(i + 1) % n]
        for (x, y) in line8(a[0], a[1], b[0], b[1]):
            if 0 <= x < w and 0 <= y < h: m[y, x] = True
    return m

def my_evenodd(pts, w, h):
    # pixel centre (x, y) integer coordinates; crossing rule half-open in y
    ys, xs = np.mgrid[0:h, 0:w]
    inside = np.zeros((h, w), bool)
    n = len(pts)
    for i in range(n):
        x0, y0 = pts[i]; x1, y1 = pts[(i + 1) % n]
        if y0 == y1: continue
        cond = ((y0 <= ys) & (ys < y1)) | ((y1 <= ys) & (ys < y0))
        xi = x0 + (ys - y0) * (x1 - x0) / (y1 - y0)
        inside ^= cond & (xs < xi)
    return inside

random.seed(1)
W, H = 200, 160
tot = 0; bad_out = 0; bad_final = 0; npx = 0
for t in range(300):
    n = 16
    cx, cy = random.uniform(40, 160), random.uniform(40, 120)
    pts = []
    for k in range(n):
        ang = 2 * math.pi * k / n
        r = random.uniform(3, 70)
        pts.append((int(cx + r * math.cos(ang)), int(cy + r * math.sin(ang))))
    po = pil_outline_only(pts, W, H); mo = my_outline(pts, W, H)
    bad_out += int((po != mo).sum())
    pm = pil_mask(pts, W, H)
    mine = my_evenodd(pts, W, H) & ~mo
    d = int((pm != mine).sum()); bad_final += d; npx += int(pm.sum())
    if d and tot < 3:
        tot += 1
        ys, xs = np.nonzero(pm != mine)
        print('case', t, 'diff px', list(zip(xs[:8], ys[:8])), 'pil', pm[ys[:8], xs[:8]], 'pts', pts)
print('outline mismatches', bad_out, 'final mismatches', bad_final, 'of', npx)
