import numpy as np, math, random
from pil_rules import *
f32 = np.float32
def RU(f):
    f = f32(f)
    return int(math.floor(f + f32(0.5))) if f >= 0 else int(-math.floor(abs(f) + f32(0.5)))
def RD(f):
    f = f32(f)
    return int(math.ceil(f - f32(0.5))) if f >= 0 else int(-math.ceil(abs(f) - f32(0.5)))
def edge_x(e, y):
    return f32(f32(y - e['y0']) * e['dx'] + f32(e['x0']))
def pil_fill_v2(pts, w, h, jodd=False, guard=0):
    m = np.zeros((h, w), bool)
    n = len(pts)
    edges = []
    for i in range(n):
        x0, y0 = pts[i]; x1, y1 = pts[(i + 1) % n]
        ymin, ymax = (y0, y1) if y0 <= y1 else (y1, y0)
        xmin, xmax = min(x0, x1), max(x0, x1)
        dx = f32(0) if y0 == y1 else f32(f32(x1 - x0) / f32(y1 - y0))
        edges.append(dict(x0=x0, y0=y0, xmin=xmin, xmax=xmax, ymin=ymin, ymax=ymax, dx=dx))
    def hline(xa, y, xb):
        if 0 <= y < h:
            if xa > xb: xa, xb = xb, xa
            xa = max(xa, 0); xb = min(xb, w - 1)
            if xa <= xb: m[y, xa:xb + 1] = True
    table = []
    ymin, ymax = h - 1, 0
    for e in edges:
        ymin = min(ymin, e['ymin']); ymax = max(ymax, e['ymax'])
        if e['ymin'] == e['ymax']:
            hline(e['xmin'], e['ymin'], e['xmax']); continue
        table.append(e)
    ymin = max(ymin, 0); ymax = min(ymax, h)
    xx = [f32(0)] * (2 * len(table) + 2)
    for y in range(ymin, ymax + 1):
        j = 0
        for i, cur in enumerate(table):
            if cur['ymin'] <= y <= cur['ymax']:
                xx[j] = edge_x(cur, y); j += 1
                if y == cur['ymax'] and y < ymax:
                    xx[j] = xx[j - 1]; j += 1
                elif cur['dx'] != 0 and (not jodd or j % 2 == 1) and f32(np.round(xx[j - 1])) == xx[j - 1]:
                    for k in range(i):
                        oth = table[k]
                        if (cur['dx'] > 0 and oth['dx'] <= 0) or (cur['dx'] < 0 and oth['dx'] >= 0): continue
                        if xx[j - 1] == edge_x(oth, y):
                            off = -1 if y == ymax else 1
                            a = edge_x(cur, y + off); b = edge_x(oth, y + off)
                            if y == cur['ymax']:
                                v = f32(max(a, b) + 1) if cur['dx'] > 0 else f32(min(a, b) - 1)
                            else:
                                v = f32(min(a, b)) if cur['dx'] > 0 else f32(max(a, b) + 1)
                            ok = True
                            if guard == 1: ok = abs(cur['dx']) > 1 or abs(oth['dx']) > 1
                            if guard == 2:
                                if y == cur['ymax']: ok = (v < xx[j - 1]) if cur['dx'] > 0 else (v > xx[j - 1])
                                else: ok = (v < xx[j - 1]) if cur['dx'] > 0 else (v > xx[j - 1])
                            if guard == 3: ok = abs(cur['dx']) > 1 and abs(oth['dx']) > 1
                            if ok: xx[k] = v
                            break
        s = sorted(xx[:j])
        x_pos = 0 if j else -1
        for i in range(1, j, 2):
            x_end = RD(s[i])
            if x_end < x_pos: continue
            x_start = RU(s[i - 1])
            if x_pos > x_start:
                x_start = x_pos
                if x_end < x_start: continue
            hline(x_start, y, x_end)
            x_pos = x_end + 1
    return m
def gen(kind, n, W, H):
    if kind == 'random': return [(random.randint(5, W - 5), random.randint(5, H - 5)) for _ in range(n)]
    if kind == 'star':
        cx, cy = random.uniform(40, 160), random.uniform(40, 120)
        return [(int(cx + random.uniform(1, 60) * math.cos(2 * math.pi * k / n)), int(cy + random.uniform(1, 60) * math.sin(2 * math.pi * k / n))) for k in range(n)]
    return [(random.randint(-60, W + 60), random.randint(-60, H + 60)) for _ in range(n)]
if __name__ == '__main__':
    W, H = 220, 180
    for jodd, guard in ((False, 1), (False, 2), (False, 3), (True, 2)):
        random.seed(3)
        tot = 0; cases = 0; shown = 0
        for t in range(300):
            n = random.choice([3, 4, 8, 16, 32]); kind = random.choice(['random', 'star', 'offimage'])
            pts = gen(kind, n, W, H)
            pf = pil_fill_only(pts, W, H); mf = pil_fill_v2(pts, W, H, jodd, guard)
            d = int((pf != mf).sum()); tot += d; cases += d > 0
            if d and shown < 5 and n <= 4:
                shown += 1
                ys, xs = np.nonzero(pf != mf)
                print(kind, n, 'diff', [(int(x), int(y), 'pil' if pf[y, x] else 'mine') for x, y in zip(xs[:6], ys[:6])], pts)
        print('jodd', jodd, 'guard', guard, 'mismatched px', tot, 'cases', cases)
