"""Post-construction rewrite of an (unmodified) reference ``DLASeg`` instance for inference — SURVEY §8(f)-2.

    model = create_model(...).eval().cuda()
    centerpoly_b200.fuse.optimize_for_inference(model)

No file of the reference is edited; every rewrite is on the live module tree and keeps ``state_dict`` keys:

* ``DeformConv`` (pose_dla_dcn.py:347-359: DCN -> BatchNorm2d -> ReLU): the eval-mode BN affine and the ReLU move
  into the DCN kernel's epilogue (``y = relu(scale * conv + shift)``), removing two elementwise passes per layer.
* ``IDAUp`` (pose_dla_dcn.py:362-387): ``up_i`` (depthwise ConvTranspose2d) becomes one streaming kernel that also
  performs the ``layers[i] + layers[i-1]`` of :387.
* plain ``Conv2d -> BatchNorm2d`` pairs of the DLA backbone (BasicBlock :42-69, Root :167-185, Tree.project,
  base_layer / level0 / level1) are folded into the convolution weights (exact in eval mode).

Inference only: the rewritten model raises if asked for gradients.
"""
from __future__ import annotations

import types

import torch
from torch import nn
from torch.nn.utils.fusion import fuse_conv_bn_eval

from .dcn_v2 import DCN
from .neck import IDAUpsample

__all__ = ["optimize_for_inference", "fuse_deform_convs", "fuse_idaup", "fold_conv_bn"]


def _bn_affine(bn: nn.BatchNorm2d):
    var = bn.running_var.detach().float()
    mean = bn.running_mean.detach().float()
    gamma = bn.weight.detach().float() if bn.weight is not None else torch.ones_like(var)
    beta = bn.bias.detach().float() if bn.bias is not None else torch.zeros_like(var)
    scale = gamma / torch.sqrt(var + bn.eps)
    return scale, beta - mean * scale


def fuse_deform_convs(model: nn.Module) -> int:
    """DeformConv modules: fold eval-mode BN + ReLU into the DCN epilogue.  Returns the number fused."""
    n = 0
    for m in model.modules():
        conv, actf = getattr(m, "conv", None), getattr(m, "actf", None)
        if not (isinstance(conv, DCN) and isinstance(actf, nn.Sequential) and len(actf) == 2 and
                isinstance(actf[0], nn.BatchNorm2d) and isinstance(actf[1], nn.ReLU)):
            continue
        if actf[0].training:
            raise RuntimeError("fuse_deform_convs needs model.eval(): BatchNorm must use its running statistics")
        scale, shift = _bn_affine(actf[0])
        if conv.bias is not None:
            shift = shift + scale * conv.bias.detach().float()
        conv._fused_epilogue = (scale.contiguous(), shift.contiguous(), True)
        m._cpoly_saved_actf = [actf]          # in a list: not registered as a sub-module
        m.actf = nn.Identity()
        n += 1
    return n


def _idaup_forward_fused(self, layers, startp, endp):
    """IDAUp.forward (pose_dla_dcn.py:381-387) with up + skip-add as one kernel."""
    for i in range(startp + 1, endp):
        j = str(i - startp)
        up, proj, node = getattr(self, "up_" + j), getattr(self, "proj_" + j), getattr(self, "node_" + j)
        layers[i] = node(up(proj(layers[i]), layers[i - 1]))


def fuse_idaup(model: nn.Module) -> int:
    n = 0
    for m in model.modules():
        if not (hasattr(m, "proj_1") and hasattr(m, "up_1") and hasattr(m, "node_1")):
            continue
        i = 1
        while hasattr(m, f"up_{i}"):
            up = getattr(m, f"up_{i}")
            if isinstance(up, nn.ConvTranspose2d):
                setattr(m, f"up_{i}", IDAUpsample(up))
                n += 1
            i += 1
        m.forward = types.MethodType(_idaup_forward_fused, m)
    return n


_PAIRS = (("conv", "bn"), ("conv1", "bn1"), ("conv2", "bn2"), ("conv3", "bn3"))


def fold_conv_bn(model: nn.Module) -> int:
    """Conv2d -> BatchNorm2d pairs (attribute pairs convN/bnN, and neighbours inside nn.Sequential)."""
    n = 0
    for m in model.modules():
        if isinstance(m, nn.Sequential):
            keys = list(m._modules.keys())
            for a, b in zip(keys, keys[1:]):
                ca, cb = m._modules[a], m._modules[b]
                if isinstance(ca, nn.Conv2d) and isinstance(cb, nn.BatchNorm2d) and not cb.training:
                    m._modules[a] = fuse_conv_bn_eval(ca, cb)
                    m._modules[b] = nn.Identity()
                    n += 1
            continue
        for cn, bnn in _PAIRS:
            c, b = getattr(m, cn, None), getattr(m, bnn, None)
            if type(c) is nn.Conv2d and isinstance(b, nn.BatchNorm2d) and not b.training:
                setattr(m, cn, fuse_conv_bn_eval(c, b))
                setattr(m, bnn, nn.Identity())
                n += 1
    return n


def optimize_for_inference(model: nn.Module) -> dict:
    """Apply every rewrite; the model must be in eval mode and on its CUDA device.  Returns counts."""
    if model.training:
        raise RuntimeError("optimize_for_inference needs model.eval()")
    with torch.no_grad():
        stats = dict(deform_convs=fuse_deform_convs(model), upsamplers=fuse_idaup(model), conv_bn=fold_conv_bn(model))
    return stats
