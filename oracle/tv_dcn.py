"""TEST INFRASTRUCTURE — NOT PRODUCT CODE.

``DCN`` with the upstream DCNv2 module API, computed by ``torchvision.ops.deform_conv2d`` on the
CPU.  This is (a) the independent implementation the C oracle is pinned against and (b) the
"reference arm": BASELINE.json's north_star names torchvision's CPU ``deform_conv2d`` as the CPU
stand-in because the reference's DCNv2 submodule is missing
(/root/reference/.SUBMODULES.json:8-14; call sites pose_dla_dcn.py:16,354, resnet_dcn.py:18,221-223).

Only tests/, ``__graft_entry__.smoke()`` and bench.py's CPU-baseline legs may import this module.
"""
from __future__ import annotations

import math

import torch
from torch import nn
from torchvision.ops import deform_conv2d


def _two(v):
    return (v, v) if isinstance(v, int) else tuple(v)


def tv_dcn_v2_conv(input, offset, mask, weight, bias, stride=1, padding=1, dilation=1,
                   deformable_groups=1):
    """DCNv2 convention == torchvision convention: offset channel 2k = dy, 2k+1 = dx of tap k."""
    del deformable_groups  # torchvision infers the group count from offset.shape[1]
    return deform_conv2d(input, offset, weight, bias, stride=_two(stride), padding=_two(padding),
                         dilation=_two(dilation), mask=mask)


class TVDCN(nn.Module):
    """state_dict: weight, bias, conv_offset_mask.weight, conv_offset_mask.bias (upstream layout)."""

    def __init__(self, in_channels, out_channels, kernel_size, stride, padding, dilation=1,
                 deformable_groups=1):
        super().__init__()
        self.kernel_size, self.stride = _two(kernel_size), _two(stride)
        self.padding, self.dilation = _two(padding), _two(dilation)
        self.deformable_groups = deformable_groups
        kh, kw = self.kernel_size
        self.weight = nn.Parameter(torch.empty(out_channels, in_channels, kh, kw))
        self.bias = nn.Parameter(torch.zeros(out_channels))
        bound = 1.0 / math.sqrt(in_channels * kh * kw)
        nn.init.uniform_(self.weight, -bound, bound)
        self.conv_offset_mask = nn.Conv2d(in_channels, deformable_groups * 3 * kh * kw, self.kernel_size,
                                          stride=self.stride, padding=self.padding, bias=True)
        nn.init.zeros_(self.conv_offset_mask.weight)
        nn.init.zeros_(self.conv_offset_mask.bias)

    def forward(self, x):
        om = self.conv_offset_mask(x)
        n_off = om.shape[1] // 3 * 2
        return tv_dcn_v2_conv(x, om[:, :n_off], torch.sigmoid(om[:, n_off:]), self.weight, self.bias,
                              self.stride, self.padding, self.dilation, self.deformable_groups)
