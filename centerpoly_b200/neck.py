"""The ``up_i`` of the reference's IDAUp — depthwise ``nn.ConvTranspose2d(o, o, 2f, stride=f, padding=f//2,
groups=o, bias=False)`` (src/lib/models/networks/pose_dla_dcn.py:372-375, applied at :385) — as one streaming
CUDA kernel, optionally fused with the skip add ``layers[i] + layers[i-1]`` of :387.

``IDAUpsample`` keeps the parameter name and shape of the module it replaces (``weight`` (C,1,2f,2f)), so
``state_dict`` keys do not change.  ``IDAUpsample`` is the inference module (optional fused skip add, no autograd);
``IDAUpsampleTrain`` is the training drop-in: same forward kernel plus ``cpoly_idaup_upsample_backward`` for the data
and weight gradients (cuDNN's grouped-deconvolution kernels cost 15 ms of the 86 ms configs[3] step).
"""
from __future__ import annotations

from ctypes import byref

import torch
from torch import nn

from . import _lib

from torch.autograd.function import once_differentiable

__all__ = ["idaup_upsample", "IDAUpsample", "IDAUpsampleTrain", "use_training_upsamplers"]


def idaup_upsample(x: torch.Tensor, weight: torch.Tensor, f: int, skip: torch.Tensor | None = None,
                   tap_major: bool = False) -> torch.Tensor:
    """conv_transpose2d(x, weight, stride=f, padding=f//2, groups=C) [+ skip]; x (B,C,H,W) -> (B,C,H*f,W*f).
    ``weight`` is the parameter (C,1,2f,2f), or with ``tap_major`` its derived (2f*2f, C) fp32 transpose."""
    dev = _lib.require_cuda(x, weight, skip)
    want = (4 * f * f, x.shape[1]) if tap_major else (x.shape[1], 1, 2 * f, 2 * f)
    if x.dim() != 4 or tuple(weight.shape) != want:
        raise RuntimeError(f"idaup_upsample: weight {tuple(weight.shape)} does not match C={x.shape[1]}, f={f}")
    B, C, H, W = x.shape
    if skip is not None and (tuple(skip.shape) != (B, C, H * f, W * f) or skip.dtype != x.dtype):
        raise RuntimeError("idaup_upsample: skip must be (B,C,H*f,W*f) with the dtype of x")
    w32 = weight.detach()
    if w32.dtype != torch.float32 or not w32.is_contiguous():
        w32 = w32.float().contiguous()
    with torch.cuda.device(dev):
        cl = x.is_contiguous(memory_format=torch.channels_last)
        y = torch.empty((B, C, H * f, W * f), dtype=x.dtype, device=dev,
                        memory_format=torch.channels_last if cl else torch.contiguous_format)
        a = _lib.UpsampleArgs()
        a.B, a.C, a.H, a.W, a.f = B, C, H, W, f
        a.dtype = _lib.dtype_code(x)
        a.weight_tap_major = int(tap_major)
        a.x, a.skip, a.y = _lib.t4(x), _lib.t4(skip), _lib.t4(y)
        a.weight = w32.data_ptr()
        _lib.check(_lib.lib().cpoly_idaup_upsample_forward(byref(a), _lib.stream_ptr(dev)), "cpoly_idaup_upsample_forward")
        _lib.LAUNCHES["upsample"] += 1
    return y


class IDAUpsample(nn.Module):
    """Drop-in for IDAUp's ``up_i`` (same ``weight`` parameter), forward(x, skip=None)."""

    def __init__(self, deconv: nn.ConvTranspose2d):
        super().__init__()
        k, s = deconv.kernel_size[0], deconv.stride[0]
        ok = (deconv.groups == deconv.in_channels == deconv.out_channels and deconv.bias is None and
              deconv.kernel_size == (2 * s, 2 * s) and deconv.stride == (s, s) and deconv.padding == (s // 2, s // 2) and
              deconv.output_padding == (0, 0) and deconv.dilation == (1, 1) and s >= 2 and s % 2 == 0)
        if not ok:
            raise RuntimeError(f"IDAUpsample: {deconv} is not the depthwise 2f/f/f//2 deconvolution of IDAUp")
        self.f = s
        self.weight = deconv.weight
        self._wt = None

    def forward(self, x, skip=None):
        w = self.weight
        tag = (w.data_ptr(), w._version, w.device)
        if self._wt is None or self._wt[0] != tag:   # derived (2f*2f, C) fp32 copy of the (tiny) weights
            self._wt = (tag, w.detach().float().reshape(w.shape[0], -1).t().contiguous())
        return idaup_upsample(x, self._wt[1], self.f, skip, tap_major=True)


class _UpsampleFn(torch.autograd.Function):
    """y = conv_transpose2d(x, weight, stride=f, padding=f//2, groups=C) with both gradients on our kernels."""

    @staticmethod
    def forward(ctx, x, weight, f):
        wt = weight.detach().float().reshape(weight.shape[0], -1).t().contiguous()      # (4f^2, C) fp32
        y = idaup_upsample(x, wt, f, None, tap_major=True)
        ctx.save_for_backward(x, wt)
        ctx.f, ctx.wshape, ctx.wdtype = f, tuple(weight.shape), weight.dtype
        return y

    @staticmethod
    @once_differentiable
    def backward(ctx, gy):
        x, wt = ctx.saved_tensors
        f = ctx.f
        dev = x.device
        B, C, H, W = x.shape
        gy = gy.to(x.dtype).contiguous(memory_format=torch.channels_last)
        need_x, need_w = ctx.needs_input_grad[0], ctx.needs_input_grad[1]
        with torch.cuda.device(dev):
            gx = torch.empty_like(x, memory_format=torch.channels_last) if need_x else None
            gwt = torch.empty_like(wt) if need_w else None
            a = _lib.UpsampleBwdArgs()
            a.B, a.C, a.H, a.W, a.f, a.dtype = B, C, H, W, f, _lib.dtype_code(x)
            a.x, a.gy, a.gx = _lib.t4(x), _lib.t4(gy), _lib.t4(gx)
            a.weight = wt.data_ptr()
            a.gweight = gwt.data_ptr() if need_w else None
            if need_w:      # fixed-order (bit-reproducible) weight gradient: per-block partial sums in a workspace, no atomics
                ws = _lib.workspace(_lib.lib().cpoly_idaup_upsample_backward_workspace_bytes(byref(a)), dev)
                a.workspace, a.workspace_bytes = ws.data_ptr(), ws.numel()
            _lib.check(_lib.lib().cpoly_idaup_upsample_backward(byref(a), _lib.stream_ptr(dev)), "cpoly_idaup_upsample_backward")
            _lib.LAUNCHES["upsample_bwd"] += int(need_x) + int(need_w)
        gw = gwt.t().reshape(ctx.wshape).to(ctx.wdtype) if need_w else None
        return gx, gw, None


class IDAUpsampleTrain(nn.Module):
    """Training drop-in for IDAUp's ``up_i``: keeps the ``nn.ConvTranspose2d`` as the sub-module ``deconv`` is NOT done -
    the parameter is re-registered under the same name ``weight`` so checkpoints and optimizers see no difference."""

    def __init__(self, deconv: nn.ConvTranspose2d):
        super().__init__()
        s = deconv.stride[0]
        ok = (deconv.groups == deconv.in_channels == deconv.out_channels and deconv.bias is None and
              deconv.kernel_size == (2 * s, 2 * s) and deconv.stride == (s, s) and deconv.padding == (s // 2, s // 2) and
              deconv.output_padding == (0, 0) and deconv.dilation == (1, 1) and s >= 2 and s % 2 == 0)
        if not ok:
            raise RuntimeError(f"IDAUpsampleTrain: {deconv} is not the depthwise 2f/f/f//2 deconvolution of IDAUp")
        self.f = s
        self.weight = deconv.weight

    def forward(self, x):
        if torch.is_autocast_enabled():
            x = x.to(torch.get_autocast_dtype("cuda"))
        C = x.shape[1]
        vec = 8 if x.dtype == torch.bfloat16 else 4
        pieces = C // vec
        if not (x.is_cuda and x.dtype in (torch.bfloat16, torch.float32) and C % vec == 0 and pieces <= 32 and pieces & (pieces - 1) == 0):
            # no CPU / PyTorch fallback in the product (north_star): same behaviour as every other entry point
            raise RuntimeError(f"IDAUpsampleTrain: needs a CUDA bf16/fp32 tensor whose channel count is {vec} x a power of two "
                               f"<= {32 * vec} (got {tuple(x.shape)}, {x.dtype}, {x.device}); keep nn.ConvTranspose2d for other inputs")
        if not x.is_contiguous(memory_format=torch.channels_last):
            x = x.contiguous(memory_format=torch.channels_last)
        return _UpsampleFn.apply(x, self.weight, self.f)


def use_training_upsamplers(model: nn.Module) -> int:
    """Swap every IDAUp ``up_i`` (depthwise ConvTranspose2d 2f/f/f//2, pose_dla_dcn.py:372-375) for ``IDAUpsampleTrain``."""
    n = 0
    for parent in model.modules():
        for name, c in list(parent._modules.items()):
            if (type(c) is nn.ConvTranspose2d and c.groups == c.in_channels == c.out_channels and c.bias is None and
                    c.kernel_size == (2 * c.stride[0],) * 2 and c.padding == (c.stride[0] // 2,) * 2 and c.stride[0] % 2 == 0):
                parent._modules[name] = IDAUpsampleTrain(c)
                n += 1
    return n
