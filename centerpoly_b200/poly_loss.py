"""Polygon loss of CenterPoly's polydet trainer on the GPU — host mirror of the reference's ``PolyLoss``
(src/lib/models/losses.py:833-959) for the polar representations, with the polygon-IoU arm (``poly_loss`` =
'iou' / 'l1+iou': ``WeilPolygonClipper`` :373-628 + ``area`` :25-41, ~25-400 ms of host Python per object in the
reference) computed by ``cpoly_poly_iou_forward`` / ``_backward`` (centerpoly_b200/csrc/poly_iou.cu): value and
analytic gradient, one warp per object, the ``_transpose_and_gather_feat`` (:855) fused.  No CPU path.

    crit = PolyLoss(opt)                                   # opt.poly_loss, opt.rep, opt.poly_order as in opts.py:72,166-169
    loss = crit(output['poly'], batch['reg_mask'], batch['ind'], batch['poly'])

``poly_order`` (:891-906, ``opt.poly_order``): ``poly_order_loss`` below; like the reference, ``forward`` then returns
``(loss, loss_order)``.  Not built (raise): ``poly_loss='relu'`` / 'bce', and the IoU arm for ``rep='cartesian'`` (the
reference's ``area`` hard-codes POLAR = True).

Objects whose clip the reference cannot finish (``IOU_STATUS`` != 0: it raises IndexError or never returns) contribute
IoU 0 and no gradient but still count in the denominator; ``PolyLoss.last_status`` keeps the per-object status tensor
on the device (no sync), and ``opt.poly_strict = True`` (or ``poly_iou_loss(..., strict=True)``) raises like the reference.
"""
from __future__ import annotations

from ctypes import byref

import torch
from torch import nn
from torch.autograd.function import once_differentiable

from . import _lib

__all__ = ["poly_iou_loss", "poly_order_loss", "PolyLoss", "IOU_STATUS"]

IOU_STATUS = {0: "ok", 1: "no out-bound crossing (reference raises IndexError)", 2: "walk does not terminate in the reference",
              3: "more than 4N crossings", 4: "non-finite row"}


class _PolyIoU(torch.autograd.Function):
    """loss = 1 - sum_obj IoU(pred_obj, target_obj) / (sum(mask) + 1e-6)    (losses.py:870-888, :909-910)"""

    @staticmethod
    def forward(ctx, poly, mask, ind, target):
        dev = _lib.require_cuda(poly, mask, ind, target)
        if poly.dim() != 4 or poly.shape[1] % 2 or not 2 <= poly.shape[1] // 2 <= 64:
            raise RuntimeError("poly_iou_loss: output must be (B,2N,H,W) with 2 <= N <= 64")
        B, C, H, W = poly.shape
        if ind.dim() != 2 or ind.shape[0] != B or tuple(target.shape) != (B, ind.shape[1], C) or tuple(mask.shape) != tuple(ind.shape):
            raise RuntimeError("poly_iou_loss: mask/ind must be (B,M) and target (B,M,2N)")
        M = ind.shape[1]
        mask_u8 = (mask != 0).to(torch.uint8).contiguous()
        ind = ind.to(torch.int64).contiguous()
        target = target.to(torch.float32).contiguous()
        L = _lib.lib()
        with torch.cuda.device(dev):
            iou = torch.empty((B, M), dtype=torch.float32, device=dev)
            dpred = torch.empty((B, M, C), dtype=torch.float32, device=dev)
            status = torch.empty((B, M), dtype=torch.int32, device=dev)
            stats = torch.empty((4,), dtype=torch.float32, device=dev)
            a = _lib.PolyIouArgs()
            a.B, a.M, a.N, a.H, a.W, a.dtype = B, M, C // 2, H, W, _lib.dtype_code(poly)
            a.poly = _lib.t4(poly)
            a.ind, a.mask, a.target = ind.data_ptr(), mask_u8.data_ptr(), target.data_ptr()
            a.iou, a.dpred, a.status, a.stats = iou.data_ptr(), dpred.data_ptr(), status.data_ptr(), stats.data_ptr()
            _lib.check(L.cpoly_poly_iou_forward(byref(a), _lib.stream_ptr(dev)), "cpoly_poly_iou_forward")
            _lib.LAUNCHES["poly_iou"] += 2      # poly_iou_kernel, poly_iou_reduce_kernel
        ctx.save_for_backward(mask_u8, ind, dpred, stats)
        ctx.shape, ctx.in_dtype = tuple(poly.shape), poly.dtype
        ctx.channels_last = poly.is_contiguous(memory_format=torch.channels_last) and not poly.is_contiguous()
        ctx.mark_non_differentiable(iou, status)
        return stats[0].clone(), iou, status

    @staticmethod
    @once_differentiable
    def backward(ctx, grad_loss, _gi, _gs):
        mask_u8, ind, dpred, stats = ctx.saved_tensors
        dev = dpred.device
        B, C, H, W = ctx.shape
        L = _lib.lib()
        with torch.cuda.device(dev):
            fmt = torch.channels_last if ctx.channels_last else torch.contiguous_format
            gmap = torch.zeros(ctx.shape, dtype=torch.float32, device=dev).contiguous(memory_format=fmt)
            g = grad_loss.to(torch.float32).reshape(1).contiguous()
            a = _lib.PolyIouArgs()
            a.B, a.M, a.N, a.H, a.W, a.dtype = B, ind.shape[1], C // 2, H, W, _lib.F32
            a.ind, a.mask, a.dpred, a.stats = ind.data_ptr(), mask_u8.data_ptr(), dpred.data_ptr(), stats.data_ptr()
            a.grad_loss, a.grad_poly = g.data_ptr(), _lib.t4(gmap)
            _lib.check(L.cpoly_poly_iou_backward(byref(a), _lib.stream_ptr(dev)), "cpoly_poly_iou_backward")
            _lib.LAUNCHES["poly_iou"] += 1      # poly_iou_scatter_kernel
        return gmap.to(ctx.in_dtype), None, None, None


def poly_iou_loss(output, mask, ind, target, return_details: bool = False, strict: bool = False):
    """The IoU arm of ``PolyLoss.forward``: ``output`` (B,2N,H,W) polar poly head (fp32 / bf16, NCHW or channels_last),
    ``mask`` (B,M), ``ind`` (B,M) int64, ``target`` (B,M,2N).  Returns the scalar loss (differentiable w.r.t.
    ``output``), and with ``return_details`` also the per-object IoU and status (see ``IOU_STATUS``).  ``strict``:
    raise when an object's clip fails where the reference raises / hangs (costs one device->host sync)."""
    loss, iou, status = _PolyIoU.apply(output, mask, ind, target)
    if strict:
        bad = status[status != 0]
        if bad.numel():
            kinds = {IOU_STATUS.get(int(k), str(int(k))): int((bad == k).sum()) for k in bad.unique()}
            raise RuntimeError(f"poly_iou_loss: {bad.numel()} object(s) could not be clipped (the reference fails on them too): {kinds}")
    return (loss, iou, status) if return_details else loss


def poly_order_loss(pred, mask):
    """The order arm of the reference's ``PolyLoss.forward`` (losses.py:891-906, 908), vectorised on the tensors' device.

    Per selected object, over its N predicted angles ``a`` (``pred[..., 1::2]``), in vertex order:
      * once an angle > 0 has been seen, every later negative angle gets ``+ 2*3.14`` - IN PLACE on a view of the
        gathered prediction in the reference, so the regression arm that follows sees the shifted angles too;
      * ``loss_order += sum_{j=0}^{N-2} sum_{k=j}^{N-1} max(a_j - a_k, 0)``; finally ``/ (10 * mask.sum() + 1e-4)``.
    Returns ``(pred_shifted, loss_order)``; ``pred_shifted`` is a new tensor (same gradient as the in-place add)."""
    a = pred[..., 1::2]
    sel = (mask != 0).unsqueeze(-1)
    seen_pos = torch.cumsum((a > 0).to(torch.int32), dim=-1) > 0          # inclusive == exclusive where a_j < 0
    shift = ((a < 0) & seen_pos & sel).to(pred.dtype) * pred.new_tensor(2 * 3.14)
    a = a + shift
    N = a.shape[-1]
    d = a.unsqueeze(-1) - a.unsqueeze(-2)                                  # [.., j, k] = a_j - a_k
    jj = torch.arange(N, device=pred.device)
    pair = (jj.unsqueeze(1) <= jj.unsqueeze(0)) & (jj.unsqueeze(1) < N - 1)   # k >= j, j <= N-2 (range((2N-1)//2))
    order = (torch.relu(d) * pair * sel.unsqueeze(-1)).sum()
    order = order / (10 * mask.sum() + 1e-4)
    shifted = torch.stack((pred[..., 0::2], a), dim=-1).flatten(-2)
    return shifted, order


def _gather(feat, ind):
    """_transpose_and_gather_feat (src/lib/models/utils.py:22-26) without the transposed copy of the map."""
    B, C, H, W = feat.shape
    return feat.reshape(B, C, H * W).gather(2, ind.unsqueeze(1).expand(B, C, ind.shape[1])).transpose(1, 2)


class PolyLoss(nn.Module):
    """Same constructor and call signature as the reference's ``PolyLoss`` (losses.py:833-959)."""

    def __init__(self, opt):
        super().__init__()
        self.opt = opt
        self.last_status = None
        if opt.poly_loss not in ("l1", "iou", "l1+iou"):
            raise NotImplementedError(f"poly_loss={opt.poly_loss!r} (built: l1, iou, l1+iou)")
        if opt.poly_loss != "l1" and opt.rep not in ("polar", "polar_fixed"):
            raise NotImplementedError("the reference's IoU arm assumes polar vertices (losses.py:28, :380)")

    def forward(self, output, mask, ind, target, freq_mask=None, peak=None, hm=None):
        opt = self.opt
        loss = 0.0
        if opt.poly_loss in ("iou", "l1+iou"):                      # per object BEFORE its angles are shifted (:870-888)
            loss, _, self.last_status = poly_iou_loss(output, mask, ind, target, return_details=True,
                                                      strict=bool(getattr(opt, "poly_strict", False)))
        order = None
        pred = None
        if getattr(opt, "poly_order", False):                       # :891-906
            pred, order = poly_order_loss(_gather(output.float(), ind), mask)
        if opt.poly_loss in ("l1", "l1+iou"):                       # :911-945
            if pred is None:
                pred = _gather(output.float(), ind)
            m = mask.unsqueeze(2).expand_as(pred).float()
            if opt.rep == "cartesian":
                reg = torch.nn.functional.l1_loss(pred * m, target * m, reduction="sum")
            else:
                radii = torch.zeros(pred.shape[2], device=pred.device)
                radii[0::2] = 1.0
                reg = torch.nn.functional.l1_loss(pred * m * radii, target * m * radii, reduction="sum")
                if opt.rep == "polar":
                    reg = reg + torch.sum(1 - torch.cos(pred * m * (1 - radii) - target * m * (1 - radii)))
                elif opt.rep != "polar_fixed":
                    reg = 0.0 * reg                                  # any other rep string: no regression arm in the reference
            loss = loss + reg / (m.sum() + 1e-6)
        if order is not None:
            return loss, order                                      # :951-954
        return loss
