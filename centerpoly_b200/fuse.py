"""Post-construction rewrite of an (unmodified) reference ``DLASeg`` instance for inference — SURVEY §8(f)-2.

    model = create_model(...).eval().cuda()
    centerpoly_b200.fuse.optimize_for_inference(model)

No file of the reference is edited; every rewrite is on the live module tree.  Load the checkpoint FIRST: the DCN,
deconvolution and convolution parameters keep their names, but the folded BatchNorm tensors leave ``state_dict``.

* ``DeformConv`` (pose_dla_dcn.py:347-359: DCN -> BatchNorm2d -> ReLU): the eval-mode BN affine and the ReLU move
  into the DCN kernel's epilogue (``y = relu(scale * conv + shift)``), removing two elementwise passes per layer.
* ``IDAUp`` (pose_dla_dcn.py:362-387): ``up_i`` (depthwise ConvTranspose2d) becomes one streaming kernel that also
  performs the ``layers[i] + layers[i-1]`` of :387.
* plain ``Conv2d -> BatchNorm2d`` pairs of the DLA backbone (BasicBlock :42-69, Root :167-185, Tree.project,
  base_layer / level0 / level1) are folded into the convolution weights (exact in eval mode), and the resulting
  conv+bias+ReLU / conv+bias+residual+ReLU sites run as single cuDNN fused calls (backbone and heads are not the
  hot path and stay in cuDNN).

Inference only: the rewritten model raises if asked for gradients.
"""
from __future__ import annotations

import types

import torch
from torch import nn
from torch.nn.utils.fusion import fuse_conv_bn_eval

from .dcn_v2 import DCN
from .neck import IDAUpsample

__all__ = ["optimize_for_inference", "fuse_deform_convs", "fuse_idaup", "fold_conv_bn", "fuse_backbone_cudnn"]


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


CONCURRENT_PROJ = True      # run the mutually independent proj_i DCNs of one IDAUp call on side streams (see below)
_SIDE_STREAMS = {}


def _side_streams(device, n):
    pool = _SIDE_STREAMS.setdefault(device, [])
    while len(pool) < n:
        pool.append(torch.cuda.Stream(device))
    return pool[:n]


def _idaup_forward_fused(self, layers, startp, endp):
    """IDAUp.forward (pose_dla_dcn.py:381-387) with up + skip-add as one kernel.

    Inside one call every ``proj_i`` reads only its own ``layers[i]`` (the ``node_i`` form the serial chain through
    ``layers[i-1]``), so the proj DCNs - and their offset/mask producers - are forked onto side streams and joined
    before their ``up_i``.  At batch 1 a proj layer is 64..256 pixel tiles on 148 SMs: a persistent kernel's last wave
    leaves up to 100 SMs idle, which the next proj's CTAs now fill.  The fork/join is plain events, so it is captured
    into the engine's CUDA graph as parallel branches."""
    idx = list(range(startp + 1, endp))
    x0 = layers[idx[0]] if idx else None
    if not (CONCURRENT_PROJ and len(idx) >= 2 and x0.is_cuda and not torch.is_grad_enabled()):
        for i in idx:
            j = str(i - startp)
            up, proj, node = getattr(self, "up_" + j), getattr(self, "proj_" + j), getattr(self, "node_" + j)
            layers[i] = node(up(proj(layers[i]), layers[i - 1]))
        return
    cur = torch.cuda.current_stream(x0.device)
    side = _side_streams(x0.device, len(idx) - 1)
    fork = torch.cuda.Event()
    fork.record(cur)
    projected = {}
    for n, i in enumerate(idx):
        proj = getattr(self, "proj_" + str(i - startp))
        if n == 0:
            projected[i] = (proj(layers[i]), None)
        else:
            s = side[n - 1]
            s.wait_event(fork)
            with torch.cuda.stream(s):
                t = proj(layers[i])
            ev = torch.cuda.Event()
            ev.record(s)
            projected[i] = (t, ev)
    capturing = torch.cuda.is_current_stream_capturing()
    for i in idx:
        j = str(i - startp)
        t, ev = projected.pop(i)
        if ev is not None:
            cur.wait_event(ev)
            if not capturing:
                t.record_stream(cur)     # produced on a side stream, consumed (and later freed) under the main stream
        layers[i] = getattr(self, "node_" + j)(getattr(self, "up_" + j)(t, layers[i - 1]))


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


# ---- backbone (NOT the hot path; stays in cuDNN): conv+bias+ReLU and conv+bias+residual+ReLU as single cuDNN calls
def _conv_args(c: nn.Conv2d):
    return c.stride, c.padding, c.dilation, c.groups


def _conv_relu(c: nn.Conv2d, x):
    return torch.cudnn_convolution_relu(x, c.weight, c.bias, *_conv_args(c))


def _conv_add_relu(c: nn.Conv2d, x, z):
    return torch.cudnn_convolution_add_relu(x, c.weight, z, 1.0, c.bias, *_conv_args(c))


class _ConvReLU(nn.Module):
    """Conv2d (BN already folded) + ReLU as one cuDNN call; keeps the conv as a sub-module (same state_dict keys)."""

    def __init__(self, conv):
        super().__init__()
        self.conv = conv

    def forward(self, x):
        return _conv_relu(self.conv, x)


def _basic_block_forward(self, x, residual=None):
    """BasicBlock.forward (pose_dla_dcn.py:55-69) after BN folding: two fused cuDNN calls."""
    if residual is None:
        residual = x
    return _conv_add_relu(self.conv2, _conv_relu(self.conv1, x), residual)


def _root_forward(self, *x):
    """Root.forward (pose_dla_dcn.py:176-185) after BN folding."""
    children = x
    cat = torch.cat(x, 1)
    if getattr(self, "residual", False):
        return _conv_add_relu(self.conv, cat, children[0])
    return _conv_relu(self.conv, cat)


def _cudnn_fused_ops_work(device, dtype) -> bool:
    try:
        x = torch.randn(1, 16, 8, 8, device=device, dtype=dtype).contiguous(memory_format=torch.channels_last)
        w = torch.randn(16, 16, 3, 3, device=device, dtype=dtype).contiguous(memory_format=torch.channels_last)
        b = torch.zeros(16, device=device, dtype=dtype)
        y = torch.cudnn_convolution_relu(x, w, b, (1, 1), (1, 1), (1, 1), 1)
        torch.cudnn_convolution_add_relu(x, w, y, 1.0, b, (1, 1), (1, 1), (1, 1), 1)
        return True
    except Exception:
        return False


def _plain_conv(c):
    return type(c) is nn.Conv2d and c.bias is not None and c.padding_mode == "zeros" and isinstance(c.padding, tuple)


def fuse_backbone_cudnn(model: nn.Module) -> int:
    """After fold_conv_bn: route Conv+ReLU / Conv+residual+ReLU of the DLA backbone and heads through cuDNN's fused
    conv-bias-activation engines (removes the separate bias-add, residual-add and ReLU kernels).  Call it AFTER the
    model has its final dtype.  Returns the number of rewritten sites (0 if cuDNN lacks the fused engines here)."""
    p = next(model.parameters())
    if not p.is_cuda or not _cudnn_fused_ops_work(p.device, p.dtype):
        return 0
    n = 0
    for m in model.modules():
        if isinstance(m, nn.Sequential):
            keys = list(m._modules.keys())
            for i, a in enumerate(keys):
                ca = m._modules[a]
                if not _plain_conv(ca):
                    continue
                j = i + 1
                while j < len(keys) and isinstance(m._modules[keys[j]], nn.Identity):
                    j += 1
                if j < len(keys) and isinstance(m._modules[keys[j]], nn.ReLU):
                    m._modules[a] = _ConvReLU(ca)
                    m._modules[keys[j]] = nn.Identity()
                    n += 1
            continue
        ident = lambda name: isinstance(getattr(m, name, None), nn.Identity)
        if (_plain_conv(getattr(m, "conv1", None)) and _plain_conv(getattr(m, "conv2", None)) and ident("bn1") and
                ident("bn2") and not hasattr(m, "conv3") and isinstance(getattr(m, "relu", None), nn.ReLU)):
            m.forward = types.MethodType(_basic_block_forward, m)
            n += 2
        elif (_plain_conv(getattr(m, "conv", None)) and ident("bn") and isinstance(getattr(m, "relu", None), nn.ReLU) and
              not hasattr(m, "actf")):
            m.forward = types.MethodType(_root_forward, m)
            n += 1
    return n


def pad_stem_input_channels(model: nn.Module, multiple: int = 8) -> int:
    """Zero-pad the input channels of the stem convolution (3 -> 8) so cuDNN can use its NHWC tensor-core kernels
    for it (with 3 input channels it converts the frame to fp32 and runs a generic kernel: 12 % of a bf16 frame).
    Exact: the extra weights are zero.  The caller must feed inputs with the returned channel count (extra channels
    zero).  Returns the new channel count (unchanged if no stem was found)."""
    for parent in model.modules():
        for name, c in list(parent._modules.items()):
            if type(c) is nn.Conv2d and c.in_channels == 3 and c.groups == 1:
                cin = (3 + multiple - 1) // multiple * multiple
                new = nn.Conv2d(cin, c.out_channels, c.kernel_size, c.stride, c.padding, c.dilation, 1, c.bias is not None)
                new = new.to(device=c.weight.device, dtype=c.weight.dtype)
                with torch.no_grad():
                    new.weight.zero_()
                    new.weight[:, :3] = c.weight
                    if c.bias is not None:
                        new.bias.copy_(c.bias)
                new = new.to(memory_format=torch.channels_last)
                parent._modules[name] = new
                return cin
    return 3


class _Conv1x1Gemm(nn.Module):
    """A biased 1x1 Conv2d on a channels_last tensor as ONE cuBLASLt GEMM with the bias in its epilogue
    (``F.conv2d`` runs cuDNN + a separate bias-add kernel here).  Output channels are padded to a multiple of 8 so the
    bf16 rows stay 16-byte aligned; the result is returned as a (B,Cout,H,W) view of that buffer (the decode reads any
    strides).  Keeps ``conv`` as a sub-module: same state_dict keys; the padded copies are version-keyed caches."""

    def __init__(self, conv: nn.Conv2d):
        super().__init__()
        self.conv = conv
        self._cache = None

    def _wb(self):
        w, b = self.conv.weight, self.conv.bias
        key = (w._version, w.data_ptr(), w.dtype, w.device, b._version, b.data_ptr())
        if self._cache is None or self._cache[0] != key:
            co, ci = w.shape[0], w.shape[1]
            cp = (co + 7) // 8 * 8
            wt = torch.zeros((ci, cp), dtype=w.dtype, device=w.device)
            wt[:, :co] = w.detach().reshape(co, ci).t()
            bp = torch.zeros((cp,), dtype=w.dtype, device=w.device)
            bp[:co] = b.detach()
            self._cache = (key, wt, bp)
        return self._cache[1], self._cache[2]

    def forward(self, x):
        if not (x.is_cuda and x.is_contiguous(memory_format=torch.channels_last)) or torch.is_grad_enabled():
            return self.conv(x)
        wt, bp = self._wb()
        B, C, H, W = x.shape
        y = torch.addmm(bp, x.permute(0, 2, 3, 1).reshape(B * H * W, C), wt)
        return y.view(B, H, W, -1).permute(0, 3, 1, 2)[:, :self.conv.out_channels]


def fuse_head_1x1(model: nn.Module) -> int:
    """Biased 1x1 convolutions inside ``nn.Sequential`` containers: the closing convolutions of the heads
    (pose_dla_dcn.py:445-468: Conv3x3 -> ReLU -> Conv1x1 with bias) and, after BN folding, the ``project`` branches of
    the DLA trees (:203-207: Conv1x1 -> BatchNorm2d).  F.conv2d runs these as cuDNN + a separate bias-add kernel."""
    n = 0
    for m in model.modules():
        if not isinstance(m, nn.Sequential):
            continue
        for key, c in list(m._modules.items()):
            if (type(c) is nn.Conv2d and c.kernel_size == (1, 1) and c.stride == (1, 1) and c.padding == (0, 0) and c.groups == 1 and
                    c.bias is not None and c.weight.is_cuda and c.in_channels % 8 == 0):
                m._modules[key] = _Conv1x1Gemm(c)
                n += 1
    return n


def fuse_stem(model: nn.Module) -> int:
    """Conv2d(3,16,7,1,3) (BN folded) -> ReLU of ``DLA.base_layer`` (pose_dla_dcn.py:231-236) as ONE kernel of ours
    (centerpoly_b200/backbone.py).  bf16 models only: the kernel multiplies in bf16 (fp32 accumulation), which is what
    a bf16 model does to that convolution anyway.  Returns the number of rewritten sites."""
    from .backbone import StemConv7x7, stem_eligible
    n = 0
    for m in model.modules():
        if not isinstance(m, nn.Sequential):
            continue
        keys = list(m._modules.keys())
        for i, a in enumerate(keys):
            c = m._modules[a]
            if not (stem_eligible(c) and c.weight.is_cuda and c.weight.dtype == torch.bfloat16):
                continue
            j = i + 1
            while j < len(keys) and isinstance(m._modules[keys[j]], nn.Identity):
                j += 1
            if j < len(keys) and isinstance(m._modules[keys[j]], nn.ReLU):
                m._modules[a] = StemConv7x7(c, relu=True)
                m._modules[keys[j]] = nn.Identity()
                n += 1
    return n


def fuse_maxpool(model: nn.Module) -> int:
    """nn.MaxPool2d(2, stride=2) (Tree.downsample / project, pose_dla_dcn.py:201,262) -> our NHWC kernel."""
    from .backbone import MaxPool2x2
    n = 0
    for parent in model.modules():
        for name, c in list(parent._modules.items()):
            if (type(c) is nn.MaxPool2d and c.kernel_size in (2, (2, 2)) and c.stride in (2, (2, 2)) and c.padding in (0, (0, 0)) and
                    c.dilation in (1, (1, 1)) and not c.ceil_mode and not c.return_indices):
                parent._modules[name] = MaxPool2x2()
                n += 1
    return n


def optimize_for_inference(model: nn.Module, dtype: torch.dtype | None = None) -> dict:
    """Apply every rewrite; the model must be in eval mode and on its CUDA device (load the checkpoint first: the
    folded BatchNorm parameters leave the state_dict).  ``dtype`` (e.g. torch.bfloat16) casts the model after the
    fp32 folding and before the cuDNN fused-op rewrite.  Returns counts."""
    if model.training:
        raise RuntimeError("optimize_for_inference needs model.eval()")
    with torch.no_grad():
        stats = dict(deform_convs=fuse_deform_convs(model), upsamplers=fuse_idaup(model), conv_bn=fold_conv_bn(model))
        if dtype is not None:
            model.to(dtype)
        p = next(model.parameters())
        stats["stem"] = fuse_stem(model) if p.is_cuda else 0
        stats["maxpool"] = fuse_maxpool(model) if p.is_cuda and p.dtype in (torch.bfloat16, torch.float32) else 0
        stats["cudnn_fused"] = fuse_backbone_cudnn(model)
        stats["head_gemm"] = fuse_head_1x1(model) if p.is_cuda and p.dtype == torch.bfloat16 else 0
    return stats
