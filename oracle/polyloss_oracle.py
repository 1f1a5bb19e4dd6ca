"""CPU oracle for the polygon-IoU term of the reference's ``PolyLoss``.  TEST INFRASTRUCTURE ONLY:
nothing under ``centerpoly_b200/`` imports this file; ``tests/`` and ``tools/`` use it as the checker.

Restates /root/reference/src/lib/models/losses.py
  * ``area``                     :25-41   (shoelace over the polar polygon, with the reference's extra wrap term)
  * ``WeilPolygonClipper``       :373-628 (``is_inside`` :378-400, ``compute_intersection`` :404-478, ``clip`` :480-617)
  * ``PolyLoss.forward`` iou arm :870-888, :909-910
as: classification of every (target edge, prediction edge) pair in vectorised fp32 (same operation order as the
reference's 0-dim tensor arithmetic), the reference's in/out-bound traversal on plain integer lists, and the area /
IoU arithmetic on torch CPU tensors so that autograd supplies the gradient the reference's own graph would.
Pinned by ``tests/golden/polyiou_*.npz``: loss and d loss / d output produced by the UNMODIFIED reference class
(``tests/golden/make_polyiou_golden.py``).

Where the reference does not return (it raises ``IndexError`` when there are in-bound but no out-bound crossings,
or loops forever when its in-bound list runs empty mid-walk) the object is reported with ``status != 0`` and
contributes IoU 0 — the CUDA kernel reports the same status codes.
"""
from __future__ import annotations

import math

import numpy as np
import torch

STATUS_OK = 0
STATUS_NO_OUTBOUND = 1      # reference: IndexError on outbounds[:,0] (losses.py:590)
STATUS_RUNAWAY = 2          # reference: walk never returns to its start / in-bound list exhausted (:596)
STATUS_TOO_MANY = 3         # more than 4n in-bound or out-bound crossings (a limit of the CUDA kernel, mirrored here)
STATUS_NONFINITE = 4        # NaN / Inf in a row (the reference would produce NaN or spin)


def _cart(p):
    """(n,2) polar (r, theta) fp32 -> x, y fp32, each a separate rounded product like the reference (:383-385)."""
    return p[:, 0] * torch.cos(p[:, 1]), p[:, 0] * torch.sin(p[:, 1])


def _side(p1x, p1y, p2x, p2y, qx, qy):
    """is_inside (:378-400): 1 if R < 0, 2 if R == 0, else 0; R evaluated op by op in fp32."""
    r = (p2x - p1x) * (qy - p1y) - (p2y - p1y) * (qx - p1x)
    out = np.zeros(r.shape, np.int8)
    out[r < 0] = 1
    out[r == 0] = 2
    return out


def crossings(subject, clipping):
    """The double loop of ``clip`` (:497-566).  subject / clipping: (n,2) / (m,2) fp32 polar tensors.
    Returns (inbounds, outbounds): lists of (j, i) in the reference's append order; the k-th crossing overall is
    the k-th element of the merged order, which only matters through these two lists."""
    sx, sy = (t.numpy() for t in _cart(subject))
    cx, cy = (t.numpy() for t in _cart(clipping))
    n, m = len(sx), len(cx)
    ins, outs = [], []
    for i in range(m):
        c0x, c0y, c1x, c1y = cx[i - 1], cy[i - 1], cx[i], cy[i]
        side = _side(c0x, c0y, c1x, c1y, sx, sy)          # side of every subject vertex w.r.t. clip edge i
        for j in range(n):
            to_end, to_start = side[j], side[j - 1]
            if to_start == 0 and to_end in (1, 2):
                kind = "out"
            elif to_end == 0 and to_start in (1, 2):
                kind = "in"
            else:
                continue
            a = _side(sx[j - 1], sy[j - 1], sx[j], sy[j], c1x, c1y)
            b = _side(sx[j - 1], sy[j - 1], sx[j], sy[j], c0x, c0y)
            if a != b:
                (outs if kind == "out" else ins).append((j, i))
    return ins, outs


def walk(ins, outs, n, m, max_len=None):
    """The traversal of ``clip`` (:577-612) on integer lists.  Returns (status, vertices) with vertices a list of
    ('s', j) | ('c', i) | ('x', j, i)."""
    max_len = max_len or 4 * (n + m)
    final = []
    if len(ins) > 4 * n or len(outs) > 4 * n:
        return STATUS_TOO_MANY, final
    if not ins:
        return STATUS_OK, final
    if not outs:
        return STATUS_NO_OUTBOUND, final
    ins = list(ins)
    first_out = {}
    for k, (j, i) in enumerate(outs):
        first_out.setdefault(j, k)
    used = 0
    while used < len(ins):
        stop = ins[0]
        j, i = stop
        start = True
        while j != stop[0] or i != stop[1] or start:
            start = False
            while j not in first_out:
                final.append(("s", j))
                j = (j + 1) % n
                if len(final) > max_len:
                    return STATUS_RUNAWAY, final
            oj, oi = outs[first_out[j]]
            final.append(("x", oj, oi))
            i = oi
            while True:
                k = next((k for k, e in enumerate(ins) if e[1] == i), None)
                if k is not None:
                    break
                final.append(("c", i))
                i = (i + 1) % m
                if len(final) > max_len or not ins:
                    return STATUS_RUNAWAY, final
            j = ins[k][0]
            final.append(("x", ins[k][0], ins[k][1]))
            del ins[k]
            used += 1
            if len(final) > max_len:
                return STATUS_RUNAWAY, final
    return STATUS_OK, final


def _intersection(subject, clipping, j, i):
    """compute_intersection (:404-478) on torch scalars (differentiable w.r.t. ``subject``)."""
    c1, c2, c3, c4 = subject[j - 1], subject[j], clipping[i - 1], clipping[i]
    p1 = c1[0] * torch.cos(c1[1]), c1[0] * torch.sin(c1[1])
    p2 = c2[0] * torch.cos(c2[1]), c2[0] * torch.sin(c2[1])
    p3 = c3[0] * torch.cos(c3[1]), c3[0] * torch.sin(c3[1])
    p4 = c4[0] * torch.cos(c4[1]), c4[0] * torch.sin(c4[1])
    if p2[0] - p1[0] == 0:
        x = p1[0]
        m2 = (p4[1] - p3[1]) / (p4[0] - p3[0])
        b2 = p3[1] - m2 * p3[0]
        y = m2 * x + b2
    elif p4[0] - p3[0] == 0:
        x = p3[0]
        m1 = (p2[1] - p1[1]) / (p2[0] - p1[0])
        b1 = p1[1] - m1 * p1[0]
        y = m1 * x + b1
    else:
        m1 = (p2[1] - p1[1]) / (p2[0] - p1[0])
        b1 = p1[1] - m1 * p1[0]
        m2 = (p4[1] - p3[1]) / (p4[0] - p3[0])
        b2 = p3[1] - m2 * p3[0]
        x = (b2 - b1) / (m1 - m2)
        y = m1 * x + b1
    r = torch.sqrt(x * x + y * y)
    theta = torch.atan((y + 1e-8) / (x + 1e-8))
    if x < 0:
        theta = theta + math.pi
    elif y < 0:
        theta = theta + 2 * math.pi
    return torch.stack((r, theta))


def area(poly):
    """``area`` (:25-41): |0.5 (sum right - sum left)| over rows 0..n of the twice-repeated cartesian polygon —
    i.e. the closed shoelace sum PLUS the first edge's term once more."""
    n = poly.shape[0]
    if n == 0:
        return poly.sum() * 0.0
    x, y = poly[:, 0] * torch.cos(poly[:, 1]), poly[:, 0] * torch.sin(poly[:, 1])
    dx, dy = x.repeat(2), y.repeat(2)
    left = (dx[0:n + 1] * dy[1:n + 2]).sum()
    right = (dy[0:n + 1] * dx[1:n + 2]).sum()
    return torch.abs(0.5 * (right - left))


def object_iou(pred_row, target_row, f64=False):
    """One object of PolyLoss.forward (:874-888).  pred_row: (2N,) torch fp32 (may require grad), target_row (2N,).
    Returns (iou tensor, status, n_vertices_of_clipped_polygon).  ``f64``: the DECISIONS (sort, side tests, walk) stay
    fp32 like the reference's, the differentiable arithmetic (crossing points, areas, IoU) runs in float64."""
    p = pred_row.view(-1, 2)
    if not (bool(torch.isfinite(pred_row).all()) and bool(torch.isfinite(target_row).all())):
        return torch.zeros(()), STATUS_NONFINITE, 0
    order = torch.sort(p[:, 1].detach(), stable=True, dim=0)[1]
    sp = p[order]
    sp = torch.cat((torch.abs(sp[:, 0]).unsqueeze(1), sp[:, 1].unsqueeze(1)), 1)
    tg = target_row.view(-1, 2)
    ins, outs = crossings(sp.detach(), tg)
    status, verts = walk(ins, outs, sp.shape[0], tg.shape[0])
    if status != STATUS_OK:
        return pred_row.sum() * 0.0, status, len(verts)
    if f64:
        sp, tg = sp.double(), tg.double()
    rows = []
    for v in verts:
        if v[0] == "s":
            rows.append(sp[v[1]])
        elif v[0] == "c":
            rows.append(tg[v[1]])
        else:
            rows.append(_intersection(sp, tg, v[1], v[2]))
    clipped = torch.stack(rows) if rows else torch.empty((0, 2), dtype=sp.dtype)
    a_int = area(clipped)
    a_s, a_t = area(sp), area(tg)
    inter = float(a_int.item() == 0.0) * torch.min(a_s, a_t) + a_int
    union = a_t + a_s - inter
    return inter / (union + 1e-6), status, len(verts)


def poly_iou_loss_oracle(pred, mask, target, f64=False):
    """pred (B,M,2N) fp32 numpy (already gathered at ``ind``), mask (B,M), target (B,M,2N).
    Returns dict(loss, iou (B,M), status (B,M), grad (B,M,2N) = d loss / d pred), loss = 1 - sum iou / (sum mask + 1e-6)
    (:909-910)."""
    p = torch.tensor(np.asarray(pred, np.float32), requires_grad=True)
    t = torch.tensor(np.asarray(target, np.float32))
    mk = np.asarray(mask)
    B, M, _ = p.shape
    iou = np.zeros((B, M), np.float64 if f64 else np.float32)
    status = np.zeros((B, M), np.int32)
    total = torch.zeros(())
    for b in range(B):
        for k in range(M):
            if mk[b, k]:
                v, st, _ = object_iou(p[b, k], t[b, k], f64)
                iou[b, k], status[b, k] = float(v), st
                total = total + v
    loss = 1 - total / (float(mk.astype(np.float32).sum()) + 1e-6)
    loss.backward()
    return dict(loss=np.float64(loss.item()) if f64 else np.float32(loss.item()), iou=iou, status=status, grad=p.grad.numpy().copy())
