"""Stem convolution and 2x2 max-pool of the DLA-34 backbone on our kernels (centerpoly_b200/csrc/backbone.cu).

The backbone is not on the reference's hot path and otherwise stays in cuDNN; these two pieces are the exceptions
because the framework kernels handle them badly at 2048x1024 (a legacy tf32 kernel plus layout conversions for the
3-channel 7x7 stem, `max_pool_forward_nhwc` at ~10 % of HBM speed): together 19 % of the measured bf16 frame.

    StemConv7x7(conv)     <- DLA.base_layer  Conv2d(3,16,7,1,3)+BN(folded)+ReLU   pose_dla_dcn.py:231-236
    MaxPool2x2()          <- nn.MaxPool2d(2, stride=2)                           pose_dla_dcn.py:201,262

Inference only (no autograd); bf16 model weights only for the stem (it computes in bf16 with fp32 accumulation,
which is what autocast does to that convolution).  No CPU path.
"""
from __future__ import annotations

from ctypes import byref

import torch
from torch import nn

from . import _lib

__all__ = ["StemConv7x7", "MaxPool2x2", "stem_eligible"]


def stem_eligible(c) -> bool:
    return (type(c) is nn.Conv2d and c.in_channels == 3 and c.out_channels == 16 and c.kernel_size == (7, 7) and
            c.stride == (1, 1) and c.padding == (3, 3) and c.dilation == (1, 1) and c.groups == 1 and c.padding_mode == "zeros")


class StemConv7x7(nn.Module):
    """conv (BN already folded into weight/bias) + ReLU in one kernel; keeps ``conv`` as a sub-module so the
    state_dict keys do not change.  Output: channels_last, dtype of the conv weight."""

    def __init__(self, conv: nn.Conv2d, relu: bool = True):
        super().__init__()
        if not stem_eligible(conv):
            raise RuntimeError("StemConv7x7 needs Conv2d(3, 16, 7, stride 1, padding 3)")
        self.conv = conv
        self.relu = relu
        self._cache = None

    def _packed(self):
        w, b = self.conv.weight, self.conv.bias
        key = (w._version, w.data_ptr(), w.dtype, w.device, None if b is None else (b._version, b.data_ptr()))
        if self._cache is None or self._cache[0] != key:
            dev = w.device
            packed = torch.empty(16 * 224, dtype=torch.bfloat16, device=dev)
            with torch.cuda.device(dev):
                _lib.check(_lib.lib().cpoly_stem_pack_weights(w.detach().contiguous().data_ptr(), _lib.dtype_code(w), packed.data_ptr(),
                                                              _lib.stream_ptr(dev)), "cpoly_stem_pack_weights")
            bias = None if b is None else b.detach().float().contiguous()
            self._cache = (key, packed, bias)
        return self._cache[1], self._cache[2]

    def forward(self, x):
        dev = _lib.require_cuda(x, self.conv.weight)
        if x.dim() != 4 or x.shape[1] != 3:
            raise RuntimeError("StemConv7x7: input must be (B,3,H,W)")
        if torch.is_grad_enabled() and (x.requires_grad or self.conv.weight.requires_grad):
            raise RuntimeError("StemConv7x7 is an inference kernel (no autograd); run under torch.no_grad()")
        packed, bias = self._packed()
        B, _, H, W = x.shape
        out_dtype = self.conv.weight.dtype
        y = torch.empty((B, 16, H, W), dtype=out_dtype, device=dev, memory_format=torch.channels_last)
        a = _lib.StemArgs()
        a.B, a.H, a.W, a.in_dtype, a.out_dtype, a.relu = B, H, W, _lib.dtype_code(x), _lib.dtype_code(y), int(self.relu)
        a.x = _lib.t4(x)
        a.weight_packed, a.bias, a.y = packed.data_ptr(), None if bias is None else bias.data_ptr(), y.data_ptr()
        with torch.cuda.device(dev):
            _lib.check(_lib.lib().cpoly_stem_conv7x7_forward(byref(a), _lib.stream_ptr(dev)), "cpoly_stem_conv7x7_forward")
        _lib.LAUNCHES["stem"] += 1
        return y


class MaxPool2x2(nn.Module):
    """nn.MaxPool2d(2, stride=2) for channels_last bf16 / fp32 CUDA tensors."""

    def forward(self, x):
        dev = _lib.require_cuda(x)
        B, C, H, W = x.shape
        if torch.is_grad_enabled() and x.requires_grad:
            raise RuntimeError("MaxPool2x2 is an inference kernel (no autograd); run under torch.no_grad()")
        if not x.is_contiguous(memory_format=torch.channels_last):
            x = x.contiguous(memory_format=torch.channels_last)
        y = torch.empty((B, C, H // 2, W // 2), dtype=x.dtype, device=dev, memory_format=torch.channels_last)
        a = _lib.MaxPoolArgs()
        a.B, a.C, a.H, a.W, a.dtype = B, C, H, W, _lib.dtype_code(x)
        a.x, a.y = x.data_ptr(), y.data_ptr()
        with torch.cuda.device(dev):
            _lib.check(_lib.lib().cpoly_maxpool2x2_nhwc_forward(byref(a), _lib.stream_ptr(dev)), "cpoly_maxpool2x2_nhwc_forward")
        _lib.LAUNCHES["maxpool"] += 1
        return y
