"""Data-parallel polydet training step (BASELINE.json configs[3]) — the CALLER of the DCN backward.

The reference trains with single-process ``nn.DataParallel`` (src/lib/trains/base_trainer.py:51-57,
src/lib/models/data_parallel.py:119-128): every step it re-broadcasts all parameters, scatters the batch, gathers
outputs and reduce-adds the replica gradients onto GPU 0.  Here: one process per GPU (torchrun), parameters
live on every rank, and the ONLY exchange step is a bucketed gradient all-reduce (sum, then 1/world) launched
from post-accumulate-grad hooks as soon as a bucket's last gradient is written, on a side stream, so that it
overlaps the rest of backward (NCCL over NVLink/NVSwitch on the GPU box; gloo in the CPU tests).  BatchNorm
statistics stay per rank, exactly like the reference's per-replica BN (no SyncBN), and the loss is averaged over
ranks like ``loss.mean()`` over replicas (base_trainer.py:95).

Loss: focal loss on the heat map (src/lib/models/losses.py:146-171 ``_neg_loss``) + masked L1 on the
poly / pseudo_depth / reg heads gathered at the ground-truth centres (``RegL1Loss`` :817-830 via
``_transpose_and_gather_feat``) = ``polydet_l1_loss``.  With ``rep='polar'`` targets, ``polydet_l1_iou_loss`` adds the
reference's polygon-IoU term (``PolyLoss`` with ``poly_loss='l1+iou'``, losses.py:870-945; ``WeilPolygonClipper``,
host Python at ~80 ms per object in the reference) through ``centerpoly_b200.poly_loss`` — the configuration
BASELINE.json configs[3] names ("L1 + polygon IoU loss").
"""
from __future__ import annotations

import torch
import torch.distributed as dist
from torch import nn


# ------------------------------------------------------------------------------------------------ loss
def focal_loss(pred, gt):
    """CornerNet focal loss on a post-sigmoid, clamped heat map (losses.py:146-171)."""
    pos = gt.eq(1).float()
    neg = gt.lt(1).float()
    neg_w = torch.pow(1 - gt, 4)
    pos_loss = (torch.log(pred) * torch.pow(1 - pred, 2) * pos).sum()
    neg_loss = (torch.log(1 - pred) * torch.pow(pred, 2) * neg_w * neg).sum()
    num_pos = pos.sum()
    return torch.where(num_pos > 0, -(pos_loss + neg_loss) / num_pos.clamp(min=1), -neg_loss)


def gather_at(feat, ind):
    """_transpose_and_gather_feat (models/utils.py:22-26) without materialising the transposed map."""
    B, C, H, W = feat.shape
    flat = feat.reshape(B, C, H * W)
    return flat.gather(2, ind.unsqueeze(1).expand(B, C, ind.shape[1])).transpose(1, 2)   # (B, K, C)


def reg_l1(feat, mask, ind, target):
    """RegL1Loss (losses.py:817-830): sum |pred*m - target*m| / (sum m + 1e-4)."""
    pred = gather_at(feat.float(), ind)
    m = mask.unsqueeze(2).expand_as(pred).float()
    return torch.nn.functional.l1_loss(pred * m, target * m, reduction="sum") / (m.sum() + 1e-4)


def polydet_l1_loss(out, batch, w_hm=1.0, w_poly=1.0, w_depth=0.1, w_off=1.0):
    hm = torch.clamp(out["hm"].float().sigmoid(), min=1e-4, max=1 - 1e-4)   # _sigmoid, models/utils.py:8-10
    loss = w_hm * focal_loss(hm, batch["hm"])
    loss = loss + w_poly * reg_l1(out["poly"], batch["reg_mask"], batch["ind"], batch["poly"])
    loss = loss + w_depth * reg_l1(out["pseudo_depth"], batch["reg_mask"], batch["ind"], batch["pseudo_depth"])
    loss = loss + w_off * reg_l1(out["reg"], batch["reg_mask"], batch["ind"], batch["reg"])
    return loss


def polydet_l1_iou_loss(out, batch, w_hm=1.0, w_poly=1.0, w_depth=0.1, w_off=1.0):
    """PolydetLoss.forward (src/lib/trains/polydet.py:38-162) with poly_loss='l1+iou', rep='polar': the poly term is
    the GPU ``PolyLoss`` mirror (IoU arm on our kernel, polar L1 + (1 - cos) angle arm as losses.py:927-936)."""
    from ..poly_loss import PolyLoss

    class _Opt:
        poly_loss, rep, poly_order = "l1+iou", "polar", False
    hm = torch.clamp(out["hm"].float().sigmoid(), min=1e-4, max=1 - 1e-4)
    loss = w_hm * focal_loss(hm, batch["hm"])
    loss = loss + w_poly * PolyLoss(_Opt())(out["poly"], batch["reg_mask"], batch["ind"], batch["poly"])
    loss = loss + w_depth * reg_l1(out["pseudo_depth"], batch["reg_mask"], batch["ind"], batch["pseudo_depth"])
    loss = loss + w_off * reg_l1(out["reg"], batch["reg_mask"], batch["ind"], batch["reg"])
    return loss


def synthetic_batch(batch, height, width, n_cls=8, n_vtx=32, max_objs=128, n_obj=20, device="cpu", seed=0, rep="cartesian"):
    """Targets shaped like PolydetDataset.__getitem__ (src/lib/datasets/sample/polydet.py:138-150,425-449)."""
    g = torch.Generator().manual_seed(seed)
    h, w = height // 4, width // 4
    x = torch.randn(batch, 3, height, width, generator=g)
    hm = torch.zeros(batch, n_cls, h, w)
    ind = torch.zeros(batch, max_objs, dtype=torch.int64)
    reg_mask = torch.zeros(batch, max_objs)
    poly = torch.zeros(batch, max_objs, 2 * n_vtx)
    depth = torch.zeros(batch, max_objs, 1)
    reg = torch.zeros(batch, max_objs, 2)
    yy, xx = torch.meshgrid(torch.arange(h, dtype=torch.float32), torch.arange(w, dtype=torch.float32), indexing="ij")
    for b in range(batch):
        for k in range(n_obj):
            cy, cx = int(torch.randint(2, h - 2, (1,), generator=g)), int(torch.randint(2, w - 2, (1,), generator=g))
            c = int(torch.randint(0, n_cls, (1,), generator=g))
            sigma = 1.0 + 3.0 * float(torch.rand(1, generator=g))
            hm[b, c] = torch.maximum(hm[b, c], torch.exp(-((yy - cy) ** 2 + (xx - cx) ** 2) / (2 * sigma ** 2)))
            hm[b, c, cy, cx] = 1.0
            ind[b, k] = cy * w + cx
            reg_mask[b, k] = 1
            ang = torch.linspace(0, 2 * 3.14159, n_vtx + 1)[:-1]
            r = 4.0 + 10.0 * torch.rand(n_vtx, generator=g)
            if rep == "polar":           # (r, theta) pairs, theta in [0, 2 pi) increasing (sample/polydet.py:250-263)
                poly[b, k, 0::2] = r
                poly[b, k, 1::2] = ang + 0.5 * 6.28318 / n_vtx * float(torch.rand(1, generator=g))
            else:
                poly[b, k, 0::2] = r * torch.cos(ang)
                poly[b, k, 1::2] = r * torch.sin(ang)
            depth[b, k, 0] = float(torch.rand(1, generator=g))
            reg[b, k] = torch.rand(2, generator=g)
    out = dict(input=x, hm=hm, ind=ind, reg_mask=reg_mask, poly=poly, pseudo_depth=depth, reg=reg)
    return {k: v.to(device) for k, v in out.items()}


# ------------------------------------------------------------------------------------------------ gradient exchange
class BucketedAllReduce:
    """Flat gradient buckets filled in reverse parameter order (= roughly the order backward produces them) and
    all-reduced asynchronously as soon as they are complete.

    Data movement per step and parameter: ONE copy, gradient -> its slice of the flat bucket, issued on the side stream
    from the post-accumulate hook; ``p.grad`` is then re-pointed at that slice, the all-reduce averages in place
    (``ReduceOp.AVG`` on NCCL; sum + one scale of the flat bucket on gloo), and the optimizer reads the bucket directly -
    there is no copy back and no per-parameter multiply.  With one rank nothing is registered at all."""

    def __init__(self, params, bucket_bytes=25 << 20, group=None):
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.params = [p for p in params if p.requires_grad]
        self.buckets = []     # dict(flat, views[(param, view)], pending, work)
        self.stream = None
        self._hooks = []
        self.launch_events = []   # (bucket index, CUDA event recorded when its all-reduce was enqueued): overlap evidence
        self.last_launch_events = []
        if self.world == 1:
            return
        cur, cur_bytes = [], 0
        for p in reversed(self.params):
            cur.append(p)
            cur_bytes += p.numel() * 4
            if cur_bytes >= bucket_bytes:
                self._close(cur)
                cur, cur_bytes = [], 0
        if cur:
            self._close(cur)
        self._where = {}
        for bi, b in enumerate(self.buckets):
            for p, _ in b["views"]:
                self._where[p] = bi
        self.stream = torch.cuda.Stream() if self.params and self.params[0].is_cuda else None
        self._avg = dist.ReduceOp.AVG if dist.get_backend(group) == "nccl" else None
        self._hooks = [p.register_post_accumulate_grad_hook(self._on_grad) for p in self.params]

    def _close(self, plist):
        n = sum(p.numel() for p in plist)
        flat = torch.zeros(n, dtype=torch.float32, device=plist[0].device)
        views, off = [], 0
        for p in plist:
            seg = flat[off:off + p.numel()]
            # the slice takes the parameter's own (dense) layout - channels_last weights keep channels_last strides - so
            # that p.grad can BE the slice (fused optimizers require params and grads of the same layout)
            dense = p.is_contiguous() or (p.dim() == 4 and p.is_contiguous(memory_format=torch.channels_last))
            views.append((p, seg.as_strided(p.shape, p.stride()) if dense else seg.view_as(p)))
            off += p.numel()
        self.buckets.append(dict(flat=flat, views=views, pending=len(plist), work=None, n=len(plist), launched=False, idx=len(self.buckets)))

    def _on_grad(self, p):
        b = self.buckets[self._where[p]]
        b["pending"] -= 1
        if b["pending"] == 0:
            self._launch(b)

    def _launch(self, b):
        """Move the bucket's gradients into its flat buffer and start the all-reduce on the side stream."""
        import contextlib
        if self.stream is not None:
            self.stream.wait_stream(torch.cuda.current_stream())
        with (torch.cuda.stream(self.stream) if self.stream is not None else contextlib.nullcontext()):
            for p, v in b["views"]:
                if p.grad is None:
                    v.zero_()      # frozen / unused parameter: contributes zeros, keeps the buckets aligned across ranks
                elif p.grad.data_ptr() != v.data_ptr():
                    g = p.grad
                    v.copy_(g)
                    if g.dtype == v.dtype:
                        if g.is_cuda and self.stream is not None:
                            g.record_stream(self.stream)   # its memory must not be recycled before the side-stream copy ran
                        p.grad = v  # the optimizer reads the (averaged) bucket slice; the producer's tensor is released
            if self._avg is not None:
                b["work"] = dist.all_reduce(b["flat"], op=self._avg, group=self.group, async_op=True)
            else:
                b["work"] = dist.all_reduce(b["flat"], op=dist.ReduceOp.SUM, group=self.group, async_op=True)
            if self.stream is not None:
                ev = torch.cuda.Event(enable_timing=True)
                ev.record(self.stream)
                self.launch_events.append((b["idx"], ev))
        b["launched"] = True

    def finish(self):
        """Call after backward: flush incomplete buckets, wait; p.grad then holds the rank-averaged gradient."""
        if self.world == 1:
            return
        for b in self.buckets:
            if not b.get("launched"):
                self._launch(b)
        for b in self.buckets:
            if b["work"] is not None:
                b["work"].wait()
        if self.stream is not None:
            torch.cuda.current_stream().wait_stream(self.stream)
        for b in self.buckets:
            if self._avg is None:
                b["flat"].mul_(1.0 / self.world)
            for p, v in b["views"]:
                if p.grad is not None and p.grad.data_ptr() != v.data_ptr():
                    p.grad.copy_(v)        # gradient kept in another dtype: one cast back
            b["pending"], b["work"], b["launched"] = b["n"], None, False
        self.last_launch_events, self.launch_events = self.launch_events, []

    def remove(self):
        for h in self._hooks:
            h.remove()


def broadcast_module_state(model: nn.Module, src: int = 0, group=None) -> int:
    """What nn.DataParallel.replicate does every step (src/lib/models/data_parallel.py:70-75): every rank starts from
    rank ``src``'s parameters AND buffers (BN running statistics), so a rank-dependent init, a different seed or a
    checkpoint loaded on one rank cannot make the replicas diverge silently.  Returns the number of tensors sent."""
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return 0
    n = 0
    with torch.no_grad():
        for t in list(model.parameters()) + list(model.buffers()):
            dist.broadcast(t.data, src=src, group=group)
            n += 1
    return n


class DataParallelTrainer:
    """model + Adam + bucketed gradient all-reduce; ``step(batch)`` returns the rank-local loss (a tensor)."""

    def __init__(self, model: nn.Module, lr=2e-4, dtype=None, bucket_bytes=25 << 20, loss_fn=polydet_l1_loss):
        self.model = model
        self.dtype = dtype
        self.loss_fn = loss_fn
        broadcast_module_state(model)      # replicas start identical (reference: nn.DataParallel re-broadcasts GPU 0's)
        params = list(model.parameters())
        # src/main.py:44 builds torch.optim.Adam(model.parameters(), opt.lr); same update rule, single multi-tensor launch on CUDA
        self.opt = torch.optim.Adam(params, lr, fused=bool(params) and all(p.is_cuda for p in params))
        self.reducer = BucketedAllReduce(list(model.parameters()), bucket_bytes)
        self.timing = False          # set True to record the events `exchange_ms()` reads
        self._ev = None

    def exchange_ms(self):
        """(exposed, span) of the last step's gradient exchange in ms, CUDA events, after a synchronize():
        exposed = time the main stream waited for the all-reduces after backward had finished (what the step pays);
        span    = first all-reduce enqueued -> exchange complete (what NCCL was busy for, mostly under backward)."""
        if self._ev is None:
            return None
        bwd_end, done = self._ev
        evs = self.reducer.last_launch_events
        span = evs[0][1].elapsed_time(done) if evs else 0.0
        return bwd_end.elapsed_time(done), span

    def step(self, batch):
        self.model.train()
        self.opt.zero_grad(set_to_none=True)
        x = batch["input"]
        if self.dtype is not None and x.is_cuda:
            with torch.autocast("cuda", dtype=self.dtype):
                out = self.model(x)[-1]
        else:
            out = self.model(x)[-1]
        loss = self.loss_fn(out, batch)
        loss.backward()
        if self.timing and loss.is_cuda:
            bwd_end = torch.cuda.Event(enable_timing=True)
            bwd_end.record()
        self.reducer.finish()
        if self.timing and loss.is_cuda:
            done = torch.cuda.Event(enable_timing=True)
            done.record()
            self._ev = (bwd_end, done)
        self.opt.step()
        return loss.detach()
