"""Measurement harness: the CALLER of the hot path — a DLA-34 + DLAUp/IDAUp polydet network whose
``state_dict`` is key-for-key compatible with the reference's ``DLASeg('dla34', heads, ...)``
(src/lib/models/networks/pose_dla_dcn.py:225-307 DLA, :347-413 DeformConv/IDAUp/DLAUp, :427-482 DLASeg).

The reference tree cannot travel to the GPU box, so bench.py and the GPU tests need a network that
schedules the 16 DCN calls at their true shapes without importing it.  This is NOT part of the product
(the product is the DCN module + decode that drop into the reference's own network); it exists so that
frames/s is measured on the full polydet inference and so the data-parallel step has a model to train.
tests/test_harness_vs_reference.py loads the reference's state_dict into this network (build container
only) and checks that both produce the same heads.

``dcn_cls`` selects the DCN implementation: ``centerpoly_b200.DCN`` (CUDA, default) or the oracle's
torchvision-backed module for the CPU reference arm of bench.py.
"""
from __future__ import annotations

import math

import torch
from torch import nn

BN_MOMENTUM = 0.1
DLA34_LEVELS = (1, 1, 1, 2, 2, 1)
DLA34_CHANNELS = (16, 32, 64, 128, 256, 512)


def _bn(c):
    return nn.BatchNorm2d(c, momentum=BN_MOMENTUM)


class _ResBlock(nn.Module):
    """Two 3x3 conv+BN with an externally supplied residual (DLA passes the projected tree input)."""

    def __init__(self, cin, cout, stride=1):
        super().__init__()
        self.conv1 = nn.Conv2d(cin, cout, 3, stride, 1, bias=False)
        self.bn1 = _bn(cout)
        self.relu = nn.ReLU(inplace=True)
        self.conv2 = nn.Conv2d(cout, cout, 3, 1, 1, bias=False)
        self.bn2 = _bn(cout)

    def forward(self, x, residual=None):
        r = x if residual is None else residual
        y = self.relu(self.bn1(self.conv1(x)))
        y = self.bn2(self.conv2(y))
        return self.relu(y + r)


class _Root(nn.Module):
    """1x1 aggregation node over the concatenated children."""

    def __init__(self, cin, cout):
        super().__init__()
        self.conv = nn.Conv2d(cin, cout, 1, 1, 0, bias=False)
        self.bn = _bn(cout)
        self.relu = nn.ReLU(inplace=True)

    def forward(self, *xs):
        return self.relu(self.bn(self.conv(torch.cat(xs, 1))))


class _Tree(nn.Module):
    """Hierarchical deep aggregation tree (levels 1 or 2 for DLA-34)."""

    def __init__(self, levels, cin, cout, stride=1, level_root=False, root_dim=0):
        super().__init__()
        root_dim = root_dim or 2 * cout
        if level_root:
            root_dim += cin
        self.levels, self.level_root = levels, level_root
        if levels == 1:
            self.tree1 = _ResBlock(cin, cout, stride)
            self.tree2 = _ResBlock(cout, cout, 1)
            self.root = _Root(root_dim, cout)
        else:
            self.tree1 = _Tree(levels - 1, cin, cout, stride)
            self.tree2 = _Tree(levels - 1, cout, cout, root_dim=root_dim + cout)
        self.downsample = nn.MaxPool2d(stride, stride=stride) if stride > 1 else None
        self.project = nn.Sequential(nn.Conv2d(cin, cout, 1, 1, bias=False), _bn(cout)) if cin != cout else None

    def forward(self, x, residual=None, children=None):
        children = [] if children is None else children
        bottom = self.downsample(x) if self.downsample is not None else x
        residual = self.project(bottom) if self.project is not None else bottom
        if self.level_root:
            children.append(bottom)
        x1 = self.tree1(x, residual)
        if self.levels == 1:
            return self.root(self.tree2(x1), x1, *children)
        children.append(x1)
        return self.tree2(x1, children=children)


def _conv_level(cin, cout, n, stride=1):
    mods = []
    for i in range(n):
        mods += [nn.Conv2d(cin, cout, 3, stride if i == 0 else 1, 1, bias=False), _bn(cout), nn.ReLU(inplace=True)]
        cin = cout
    return nn.Sequential(*mods)


class DLA34Base(nn.Module):
    def __init__(self):
        super().__init__()
        L, C = DLA34_LEVELS, DLA34_CHANNELS
        self.channels = list(C)
        self.base_layer = nn.Sequential(nn.Conv2d(3, C[0], 7, 1, 3, bias=False), _bn(C[0]), nn.ReLU(inplace=True))
        self.level0 = _conv_level(C[0], C[0], L[0])
        self.level1 = _conv_level(C[0], C[1], L[1], stride=2)
        self.level2 = _Tree(L[2], C[1], C[2], 2, level_root=False)
        self.level3 = _Tree(L[3], C[2], C[3], 2, level_root=True)
        self.level4 = _Tree(L[4], C[3], C[4], 2, level_root=True)
        self.level5 = _Tree(L[5], C[4], C[5], 2, level_root=True)

    def forward(self, x):
        x = self.base_layer(x)
        feats = []
        for i in range(6):
            x = getattr(self, f"level{i}")(x)
            feats.append(x)
        return feats


def _bilinear_upsample_weights(up: nn.ConvTranspose2d) -> None:
    """Depthwise deconv initialised as bilinear interpolation (what the reference's fill_up_weights does)."""
    k = up.weight.shape[2]
    f = math.ceil(k / 2)
    c = (2 * f - 1 - f % 2) / (2.0 * f)
    idx = torch.arange(k, dtype=torch.float32)
    tri = 1 - (idx / f - c).abs()
    with torch.no_grad():
        up.weight.copy_((tri[:, None] * tri[None, :]).expand_as(up.weight))


class DeformConv(nn.Module):
    """DCN -> BN -> ReLU (pose_dla_dcn.py:347-359); ``conv`` is the hot-path module."""

    def __init__(self, cin, cout, dcn_cls):
        super().__init__()
        self.actf = nn.Sequential(_bn(cout), nn.ReLU(inplace=True))
        self.conv = dcn_cls(cin, cout, kernel_size=(3, 3), stride=1, padding=1, dilation=1, deformable_groups=1)

    def forward(self, x):
        return self.actf(self.conv(x))


class IDAUp(nn.Module):
    def __init__(self, o, channels, up_f, dcn_cls):
        super().__init__()
        for i in range(1, len(channels)):
            f = int(up_f[i])
            setattr(self, f"proj_{i}", DeformConv(channels[i], o, dcn_cls))
            up = nn.ConvTranspose2d(o, o, f * 2, stride=f, padding=f // 2, output_padding=0, groups=o, bias=False)
            _bilinear_upsample_weights(up)
            setattr(self, f"up_{i}", up)
            setattr(self, f"node_{i}", DeformConv(o, o, dcn_cls))

    def forward(self, layers, startp, endp):
        for i in range(startp + 1, endp):
            j = i - startp
            layers[i] = getattr(self, f"up_{j}")(getattr(self, f"proj_{j}")(layers[i]))
            layers[i] = getattr(self, f"node_{j}")(layers[i] + layers[i - 1])


class DLAUp(nn.Module):
    def __init__(self, startp, channels, scales, dcn_cls):
        super().__init__()
        self.startp = startp
        channels = list(channels)
        in_channels = list(channels)
        scales = list(scales)
        for i in range(len(channels) - 1):
            j = -i - 2
            setattr(self, f"ida_{i}", IDAUp(channels[j], in_channels[j:], [s // scales[j] for s in scales[j:]], dcn_cls))
            scales[j + 1:] = [scales[j]] * len(scales[j + 1:])
            in_channels[j + 1:] = [channels[j]] * len(in_channels[j + 1:])

    def forward(self, layers):
        out = [layers[-1]]
        for i in range(len(layers) - self.startp - 1):
            getattr(self, f"ida_{i}")(layers, len(layers) - i - 2, len(layers))
            out.insert(0, layers[-1])
        return out


class PolydetDLA34(nn.Module):
    """Same graph and parameter names as the reference's DLASeg('dla34', heads, pretrained=False, down_ratio=4,
    final_kernel=1, last_level=5, head_conv=256)."""

    def __init__(self, heads=None, head_conv=256, dcn_cls=None, down_ratio=4, last_level=5):
        super().__init__()
        if dcn_cls is None:
            from ..dcn_v2 import DCN as dcn_cls
        heads = heads or {"hm": 8, "poly": 64, "pseudo_depth": 1, "reg": 2}
        self.heads = heads
        self.first_level = int(math.log2(down_ratio))
        self.last_level = last_level
        self.base = DLA34Base()
        ch = self.base.channels
        scales = [2 ** i for i in range(len(ch[self.first_level:]))]
        self.dla_up = DLAUp(self.first_level, ch[self.first_level:], scales, dcn_cls)
        oc = ch[self.first_level]
        self.ida_up = IDAUp(oc, ch[self.first_level:self.last_level],
                            [2 ** i for i in range(self.last_level - self.first_level)], dcn_cls)
        for head, classes in heads.items():
            fc = nn.Sequential(nn.Conv2d(oc, head_conv, 3, padding=1, bias=True), nn.ReLU(inplace=True),
                               nn.Conv2d(head_conv, classes, 1, bias=True))
            if "hm" in head:
                fc[-1].bias.data.fill_(-2.19)
            else:
                for m in fc.modules():
                    if isinstance(m, nn.Conv2d):
                        nn.init.constant_(m.bias, 0)
            setattr(self, head, fc)

    def forward(self, x):
        feats = self.dla_up(self.base(x))
        y = [feats[i].clone() for i in range(self.last_level - self.first_level)]
        self.ida_up(y, 0, len(y))
        return [{head: getattr(self, head)(y[-1]) for head in self.heads}]


def dcn_modules(net: nn.Module):
    return [m for m in net.modules() if hasattr(m, "conv_offset_mask")]


def calibrate_random_init(net: nn.Module, height: int, width: int, seed: int = 317, device="cpu", offset_std: float = 0.05) -> nn.Module:
    """SURVEY §A.5 recipe: random weights stay random, but (a) the DCN offset convs get small random weights so
    the offsets are non-trivial and (b) one train-mode forward with BN momentum 1.0 makes the running statistics
    consistent with the activations, so the heat map is tie-free instead of degenerate."""
    g = torch.Generator().manual_seed(seed)
    for m in dcn_modules(net):
        with torch.no_grad():
            m.conv_offset_mask.weight.copy_(torch.randn(m.conv_offset_mask.weight.shape, generator=g) * offset_std)
            m.conv_offset_mask.bias.copy_(torch.randn(m.conv_offset_mask.bias.shape, generator=g) * 1.0)
    bns = [m for m in net.modules() if isinstance(m, nn.BatchNorm2d)]
    for m in bns:
        m.momentum = 1.0
    net.train()
    with torch.no_grad():
        p = next(net.parameters())
        x = torch.randn(2, 3, height, width, generator=g).to(device=p.device, dtype=p.dtype)
        if p.is_cuda:
            x = x.contiguous(memory_format=torch.channels_last)
        net(x)
    for m in bns:
        m.momentum = BN_MOMENTUM
    net.eval()
    return net
