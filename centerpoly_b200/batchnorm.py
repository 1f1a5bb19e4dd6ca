"""Training-mode ``nn.BatchNorm2d`` (+ fused ReLU) on libcpoly.so's streaming kernels (csrc/batchnorm.cu, ABI v10).

The reference network normalises with ``nn.BatchNorm2d(planes, momentum=BN_MOMENTUM)`` at 55 sites
(src/lib/models/networks/pose_dla_dcn.py:31-47 BasicBlock, :143-160 Root, :231-236 base_layer, :247-252 conv levels,
:345-359 DeformConv's ``actf = Sequential(BatchNorm2d, ReLU)``).  ``BatchNorm2dTrain`` is a drop-in subclass: same
parameters, buffers, state_dict keys and eval behaviour; in training mode on a CUDA channels_last bf16 / fp32 tensor
the statistics, the normalisation (+ ReLU when ``fuse_relu``) and the whole backward run on our kernels.
``use_training_batchnorm(model)`` swaps the modules of a live model (and fuses the ReLU of ``Sequential(BN, ReLU)``).
"""
from __future__ import annotations

from ctypes import byref

import torch
from torch import nn

from . import _lib

__all__ = ["BatchNorm2dTrain", "use_training_batchnorm", "batch_norm_train"]


def _workspace(C, dev):
    return _lib.workspace(_lib.lib().cpoly_bn_workspace_bytes(C), dev)


class _BnTrainFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, running_mean, running_var, momentum, eps, relu):
        B, C, H, W = x.shape
        dev = x.device
        L = _lib.lib()
        with torch.cuda.device(dev):
            y = torch.empty_like(x, memory_format=torch.channels_last)
            save = torch.empty(4, C, dtype=torch.float32, device=dev)
            ws = _workspace(C, dev)
            a = _lib.BnFwdArgs()
            a.N, a.C, a.dtype, a.relu = B * H * W, C, _lib.dtype_code(x), int(relu)
            a.x, a.y = x.data_ptr(), y.data_ptr()
            a.weight = weight.data_ptr() if weight is not None else None
            a.bias = bias.data_ptr() if bias is not None else None
            a.running_mean = running_mean.data_ptr() if running_mean is not None else None
            a.running_var = running_var.data_ptr() if running_var is not None else None
            a.momentum, a.eps = float(momentum), float(eps)
            a.save = save.data_ptr()
            a.workspace, a.workspace_bytes = ws.data_ptr(), ws.numel()
            _lib.check(L.cpoly_bn_train_forward(byref(a), _lib.stream_ptr(dev)), "cpoly_bn_train_forward")
            _lib.LAUNCHES["batchnorm"] = _lib.LAUNCHES.get("batchnorm", 0) + 3
        ctx.save_for_backward(x, save)
        ctx.relu = bool(relu)
        ctx.has_affine = (weight is not None, bias is not None)
        return y

    @staticmethod
    def backward(ctx, gy):
        x, save = ctx.saved_tensors
        B, C, H, W = x.shape
        dev = x.device
        L = _lib.lib()
        if gy.dtype != x.dtype:
            gy = gy.to(x.dtype)
        if not gy.is_contiguous(memory_format=torch.channels_last):
            gy = gy.contiguous(memory_format=torch.channels_last)
        need_x, need_w, need_b = ctx.needs_input_grad[0], ctx.needs_input_grad[1], ctx.needs_input_grad[2]
        with torch.cuda.device(dev):
            gx = torch.empty_like(x, memory_format=torch.channels_last) if need_x else None
            gw = torch.empty(C, dtype=torch.float32, device=dev) if need_w else None
            gb = torch.empty(C, dtype=torch.float32, device=dev) if need_b else None
            ws = _workspace(C, dev)
            a = _lib.BnBwdArgs()
            a.N, a.C, a.dtype, a.relu = B * H * W, C, _lib.dtype_code(x), int(ctx.relu)
            a.x, a.gy = x.data_ptr(), gy.data_ptr()
            a.gx = gx.data_ptr() if gx is not None else None
            a.save = save.data_ptr()
            a.gweight = gw.data_ptr() if gw is not None else None
            a.gbias = gb.data_ptr() if gb is not None else None
            a.workspace, a.workspace_bytes = ws.data_ptr(), ws.numel()
            _lib.check(L.cpoly_bn_train_backward(byref(a), _lib.stream_ptr(dev)), "cpoly_bn_train_backward")
            _lib.LAUNCHES["batchnorm_bwd"] = _lib.LAUNCHES.get("batchnorm_bwd", 0) + (3 if need_x else 2)
        return gx, gw, gb, None, None, None, None, None


def _fast_path_ok(x, C):
    return (x.is_cuda and x.dim() == 4 and x.dtype in (torch.bfloat16, torch.float32) and x.numel() > C and
            bool(_lib.lib().cpoly_bn_eligible(C, _lib.dtype_code(x))))


def batch_norm_train(x, weight, bias, running_mean, running_var, momentum=0.1, eps=1e-5, relu=False):
    """Functional form: training-mode batch norm (+ ReLU) of an NCHW-shaped tensor; fp32 parameters and buffers."""
    if not _fast_path_ok(x, x.shape[1]):
        raise RuntimeError(f"batch_norm_train: needs a CUDA bf16/fp32 (B,C,H,W) tensor with an eligible channel count, got "
                           f"{tuple(x.shape)} {x.dtype} {x.device}")
    if not x.is_contiguous(memory_format=torch.channels_last):
        x = x.contiguous(memory_format=torch.channels_last)
    return _BnTrainFn.apply(x, weight, bias, running_mean, running_var, momentum, eps, relu)


class BatchNorm2dTrain(nn.BatchNorm2d):
    """``nn.BatchNorm2d`` whose training-mode forward / backward run on csrc/batchnorm.cu; ``fuse_relu`` adds the ReLU of
    ``Sequential(BatchNorm2d, ReLU)``.  Eval mode, CPU tensors, momentum=None (cumulative average) and channel counts
    without a fast path behave exactly like the parent class (that is PyTorch's own module, not a fallback of ours)."""

    def __init__(self, *args, fuse_relu: bool = False, **kwargs):
        super().__init__(*args, **kwargs)
        self.fuse_relu = fuse_relu

    @classmethod
    def from_module(cls, bn: nn.BatchNorm2d, fuse_relu: bool = False):
        m = cls(bn.num_features, eps=bn.eps, momentum=bn.momentum, affine=bn.affine, track_running_stats=bn.track_running_stats,
                fuse_relu=fuse_relu)
        m._parameters = bn._parameters      # the same Parameter objects: optimizers and checkpoints see no difference
        m._buffers = bn._buffers
        m.training = bn.training
        return m

    def forward(self, input):
        use = (self.training and self.momentum is not None and input.is_cuda and
               (self.weight is None or self.weight.dtype == torch.float32) and _fast_path_ok(input, self.num_features))
        if not use:
            out = super().forward(input)
            return torch.relu(out) if self.fuse_relu else out
        if self.track_running_stats and self.num_batches_tracked is not None:
            self.num_batches_tracked.add_(1)
        x = input if input.is_contiguous(memory_format=torch.channels_last) else input.contiguous(memory_format=torch.channels_last)
        rm, rv = (self.running_mean, self.running_var) if self.track_running_stats else (None, None)
        return _BnTrainFn.apply(x, self.weight, self.bias, rm, rv, self.momentum, self.eps, self.fuse_relu)


def use_training_batchnorm(model: nn.Module) -> int:
    """Swap every plain ``nn.BatchNorm2d`` of a live model for ``BatchNorm2dTrain`` (parameters and buffers shared);
    a ``Sequential(BatchNorm2d, ReLU)`` (DeformConv's ``actf``, pose_dla_dcn.py:348-351) becomes one fused module followed
    by an ``Identity`` so the container keeps its length and state_dict keys.  Returns the number of modules swapped."""
    n = 0
    for parent in model.modules():
        items = list(parent._modules.items())
        for i, (name, c) in enumerate(items):
            if type(c) is not nn.BatchNorm2d:
                continue
            fuse = False
            if isinstance(parent, nn.Sequential) and i + 1 < len(items) and type(items[i + 1][1]) is nn.ReLU:
                fuse = True
                parent._modules[items[i + 1][0]] = nn.Identity()
            parent._modules[name] = BatchNorm2dTrain.from_module(c, fuse_relu=fuse)
            n += 1
    return n
