"""Drop-in replacement for the DCNv2 extension's Python module ``dcn_v2.py``.

The reference imports ``from .DCNv2.dcn_v2 import DCN`` (src/lib/models/networks/pose_dla_dcn.py:16,
resnet_dcn.py:18) and constructs ``DCN(chi, cho, kernel_size=(3,3), stride=1, padding=1, dilation=1,
deformable_groups=1)`` (pose_dla_dcn.py:354, resnet_dcn.py:221-223).  The DCNv2 sources themselves are
a git submodule that is not vendored in the reference; this module mirrors the upstream public API —
``DCN``, ``DCNv2``, ``dcn_v2_conv``, ``_DCNv2`` — with the same constructor signature, parameter names
(``weight``, ``bias``, ``conv_offset_mask.{weight,bias}``: checkpoints load unchanged,
src/lib/models/model.py:52-99), init (weight ~ U(+-1/sqrt(Cin*kh*kw)), bias 0, zero offset conv) and
error behaviour (RuntimeError on failure), and routes the arithmetic to libcpoly.so through the C ABI
of include/cpoly.h.  No CPU path, no PyTorch fallback.
"""
from __future__ import annotations

import math
from ctypes import byref, c_void_p

import torch
from torch import nn
from torch.autograd import Function
from torch.autograd.function import once_differentiable
from torch.nn.modules.utils import _pair

from . import _lib

__all__ = ["DCN", "DCNv2", "dcn_v2_conv", "_DCNv2", "set_dcn_path", "set_offset_conv", "invalidate_packed_weights"]

_FORCED_PATH = _lib.PATH_AUTO
_USE_OWN_OFFSET_CONV = True     # tests / profiling: False routes the fused inference path back to F.conv2d (cuDNN)


def set_offset_conv(own: bool) -> None:
    global _USE_OWN_OFFSET_CONV
    _USE_OWN_OFFSET_CONV = bool(own)



def set_dcn_path(path: str) -> None:
    """'auto' (default), 'simt' or 'tcgen05' — used by tests and profiling to pin a kernel family."""
    global _FORCED_PATH
    _FORCED_PATH = {"auto": _lib.PATH_AUTO, "simt": _lib.PATH_SIMT, "tcgen05": _lib.PATH_TCGEN05}[path]


def _tc_allowed(x) -> bool:
    """The tcgen05 path multiplies fp32 tensors as TF32 (10-bit mantissa, gated at 1e-3 by the tests).  Like every
    PyTorch convolution it therefore follows ``torch.backends.cudnn.allow_tf32`` (default True): with the flag off,
    fp32 tensors take the IEEE-fp32 CUDA-core kernels.  bf16 tensors are unaffected."""
    if _FORCED_PATH == _lib.PATH_SIMT:
        return False
    if _FORCED_PATH == _lib.PATH_TCGEN05:
        return True
    return x.dtype != torch.float32 or torch.backends.cudnn.allow_tf32


def _out_hw(H, W, k, s, p, d):
    return ((H + 2 * p[0] - (d[0] * (k[0] - 1) + 1)) // s[0] + 1,
            (W + 2 * p[1] - (d[1] * (k[1] - 1) + 1)) // s[1] + 1)


class _PackedWeights:
    """Derived, non-persistent cache of the tap-major / swizzled weight images (never in state_dict).
    The image hangs off the weight tensor OBJECT (a plain Python attribute, so it dies with the parameter and can
    never be confused with another tensor that later reuses the same address); it is rebuilt when the parameter's
    version counter, storage address, device or shape moves (load_state_dict, .to(), .bfloat16(), in-place ops).
    The version counter is NOT trusted while training: ``torch._fused_adam_`` (and any ``.data``-style update: EMA,
    apex optimizers) changes the values without bumping it, so every forward that will be differentiated w.r.t. the
    weight (``fresh=True``) re-packs (9*Cin*Cout elements, negligible next to the layer) and the cache serves
    no-grad inference only.  ``invalidate_packed_weights`` drops it by hand after an out-of-band update."""

    def get(self, weight: torch.Tensor, dtype_code: int, fresh: bool = False):
        Cout, Cin, kh, kw = weight.shape
        L = _lib.lib()
        nbytes = L.cpoly_dcn_packed_weight_bytes(Cin, Cout, kh, kw, dtype_code)
        if nbytes == 0:
            return None
        tag = (weight._version, weight.data_ptr(), weight.device, tuple(weight.shape), weight.dtype)
        cache = getattr(weight, "_cpoly_packed", None)
        if cache is None:
            cache = {}
            try:
                weight._cpoly_packed = cache
            except Exception:      # tensor subclass without a __dict__: pack every call
                pass
        hit = cache.get(dtype_code)
        if hit is not None and hit[0] == tag and not fresh:
            return hit[1]
        w32 = weight.detach()
        if w32.dtype != torch.float32 or not w32.is_contiguous():
            w32 = w32.float().contiguous()
        packed = torch.empty(nbytes, dtype=torch.uint8, device=weight.device)
        _lib.check(L.cpoly_dcn_pack_weights(w32.data_ptr(), packed.data_ptr(), Cin, Cout, kh, kw, dtype_code,
                                            _lib.stream_ptr(weight.device)), "cpoly_dcn_pack_weights")
        _lib.LAUNCHES["pack_weights"] += 1
        cache[dtype_code] = (tag, packed)
        return packed


_GLOBAL_PACK_CACHE = _PackedWeights()


def invalidate_packed_weights(obj) -> int:
    """Drop the derived weight images of a tensor, a DCN module or a whole model (after a ``.data`` / fused-optimizer
    update made outside autograd's version counting while the model is used for no-grad inference)."""
    tensors = [obj] if isinstance(obj, torch.Tensor) else [p for p in obj.parameters()]
    n = 0
    for t in tensors:
        if getattr(t, "_cpoly_packed", None):
            t._cpoly_packed.clear()
            n += 1
    if isinstance(obj, nn.Module):
        for m in obj.modules():
            if hasattr(m, "_om_cache"):
                m._om_cache = None
    return n


def _fwd_args(x, offset, mask, weight32, bias32, stride, padding, dilation, dg, mask_is_logit, packed, y,
              wshape=None, out_scale=None, relu=False, om_bias=None):
    a = _lib.DcnFwdArgs()
    a.B, a.Cin, a.H, a.W = x.shape
    a.Cout, _, a.kh, a.kw = wshape if wshape is not None else weight32.shape
    a.stride_h, a.stride_w = stride
    a.pad_h, a.pad_w = padding
    a.dil_h, a.dil_w = dilation
    a.deformable_groups = dg
    a.dtype = _lib.dtype_code(x)
    a.om_dtype = _lib.dtype_code(offset)
    a.mask_is_logit = int(mask_is_logit)
    a.path = _FORCED_PATH if _tc_allowed(x) else _lib.PATH_SIMT
    a.x, a.offset, a.mask = _lib.t4(x), _lib.t4(offset), _lib.t4(mask)
    a.weight = weight32.data_ptr() if weight32 is not None else None
    a.bias = bias32.data_ptr() if bias32 is not None else None
    a.weight_packed = packed.data_ptr() if packed is not None else None
    a.y = _lib.t4(y)
    a.out_scale = out_scale.data_ptr() if out_scale is not None else None
    a.relu = int(bool(relu))
    if om_bias is not None:   # fp32 (3*dg*kh*kw): offsets first, then mask logits
        n_off = a.deformable_groups * 2 * a.kh * a.kw
        a.offset_bias = om_bias.data_ptr()
        a.mask_bias = om_bias.data_ptr() + 4 * n_off
    return a


def _as_f32_contig(t):
    if t is None:
        return None
    t = t.detach()
    return t if (t.dtype == torch.float32 and t.is_contiguous()) else t.float().contiguous()


def _check_shapes(x, offset, mask, weight, stride, padding, dilation, dg):
    if x.dim() != 4 or weight.dim() != 4:
        raise RuntimeError("dcn_v2: input and weight must be 4-d")
    B, Cin, H, W = x.shape
    Cout, Cin_w, kh, kw = weight.shape
    if Cin != Cin_w:
        raise RuntimeError(f"dcn_v2: input has {Cin} channels but weight expects {Cin_w}")
    Ho, Wo = _out_hw(H, W, (kh, kw), stride, padding, dilation)
    if tuple(offset.shape) != (B, dg * 2 * kh * kw, Ho, Wo):
        raise RuntimeError(f"dcn_v2: offset shape {tuple(offset.shape)} != {(B, dg * 2 * kh * kw, Ho, Wo)}")
    if tuple(mask.shape) != (B, dg * kh * kw, Ho, Wo):
        raise RuntimeError(f"dcn_v2: mask shape {tuple(mask.shape)} != {(B, dg * kh * kw, Ho, Wo)}")
    if offset.dtype != mask.dtype:
        raise RuntimeError("dcn_v2: offset and mask must share a dtype")
    return Ho, Wo


def _forward_impl(x, offset, mask, weight, bias, stride, padding, dilation, dg, mask_is_logit, out_scale=None, relu=False,
                  om_bias=None, fresh_pack=False):
    """y = act(out_scale * dcn(x) + bias); out_scale/relu are the optional fused DeformConv epilogue."""
    dev = _lib.require_cuda(x, offset, mask, weight, bias)
    Ho, Wo = _check_shapes(x, offset, mask, weight, stride, padding, dilation, dg)
    L = _lib.lib()
    with torch.cuda.device(dev):
        packed = _GLOBAL_PACK_CACHE.get(weight, _lib.dtype_code(x), fresh_pack) if _tc_allowed(x) else None
        if packed is not None and not x.is_contiguous(memory_format=torch.channels_last):
            # the tensor-core path reads NHWC; run the model channels_last to make this a no-op
            x = x.contiguous(memory_format=torch.channels_last)
        y = torch.empty((x.shape[0], weight.shape[0], Ho, Wo), dtype=x.dtype, device=dev,
                        memory_format=torch.channels_last)
        b32 = _as_f32_contig(bias)
        s32 = _as_f32_contig(out_scale)
        a = _fwd_args(x, offset, mask, None, b32, stride, padding, dilation, dg, mask_is_logit, packed, y,
                      wshape=weight.shape, out_scale=s32, relu=relu, om_bias=om_bias)
        use_tc = packed is not None and L.cpoly_dcn_tcgen05_eligible(byref(a))
        if use_tc:
            nbytes = L.cpoly_dcn_forward_workspace_bytes(byref(a))
            if nbytes > 256:
                ws = _lib.workspace(nbytes, dev)
                a.workspace, a.workspace_bytes = ws.data_ptr(), ws.numel()
        else:
            w32 = _as_f32_contig(weight)      # the generic SIMT path reads the dense fp32 master weights
            a.weight = w32.data_ptr()
        _lib.check(L.cpoly_dcn_forward(byref(a), _lib.stream_ptr(dev)), "cpoly_dcn_forward")
        _lib.LAUNCHES["dcn_forward"] += 2 if a.workspace else 1   # tap-split schedules add the combine kernel
    return y, x


def _backward_impl(x, offset, mask, weight, gout, stride, padding, dilation, dg, mask_is_logit, needs):
    """needs = (x, offset, mask, weight, bias) booleans.  Returns fp32 gradients (None where not needed)."""
    dev = _lib.require_cuda(x, offset, mask, weight, gout)
    L = _lib.lib()
    w32 = _as_f32_contig(weight)
    gout = _grad_like(gout, x)
    with torch.cuda.device(dev):
        gx = torch.empty_like(x, dtype=torch.float32) if needs[0] else None
        goff = torch.empty_like(offset, dtype=torch.float32) if needs[1] else None
        gmask = torch.empty_like(mask, dtype=torch.float32) if needs[2] else None
        gw = torch.empty(weight.shape, dtype=torch.float32, device=dev) if needs[3] else None
        gb = torch.empty(weight.shape[0], dtype=torch.float32, device=dev) if needs[4] else None
        b = _lib.DcnBwdArgs()
        b.f = _fwd_args(x, offset, mask, w32, None, stride, padding, dilation, dg, mask_is_logit, None, None)
        b.gout = _lib.t4(gout)
        b.gx, b.goffset, b.gmask = _lib.t4(gx), _lib.t4(goff), _lib.t4(gmask)
        b.gweight = gw.data_ptr() if gw is not None else None
        b.gbias = gb.data_ptr() if gb is not None else None
        nbytes = L.cpoly_dcn_backward_workspace_bytes(byref(b))
        ws = _lib.workspace(nbytes, dev)
        b.workspace, b.workspace_bytes = ws.data_ptr(), ws.numel()
        _lib.check(L.cpoly_dcn_backward(byref(b), _lib.stream_ptr(dev)), "cpoly_dcn_backward")
    return gx, goff, gmask, gw, gb


def _grad_like(gout, x):
    """grad_output in x's dtype and memory format: the tensor-core backward reads dense NHWC rows, and autograd hands
    over whatever layout the consumer produced (a caller's ``y.backward(g)`` with an NCHW ``g`` would otherwise drop
    the whole backward to the generic CUDA-core kernels)."""
    if gout.dtype != x.dtype:
        gout = gout.to(x.dtype)
    if x.dim() == 4 and x.is_contiguous(memory_format=torch.channels_last) and not gout.is_contiguous(memory_format=torch.channels_last):
        gout = gout.contiguous(memory_format=torch.channels_last)
    return gout


def _cast(g, like):
    return None if g is None else (g if g.dtype == like.dtype else g.to(like.dtype))


class _DCNv2(Function):
    """Upstream-compatible autograd seam: forward(ctx, input, offset, mask, weight, bias, stride, padding,
    dilation, deformable_groups); backward returns 5 gradients + 4 Nones; once_differentiable."""

    @staticmethod
    def forward(ctx, input, offset, mask, weight, bias, stride, padding, dilation, deformable_groups):
        ctx.stride, ctx.padding, ctx.dilation = _pair(stride), _pair(padding), _pair(dilation)
        ctx.deformable_groups = deformable_groups
        y, x_used = _forward_impl(input, offset, mask, weight, bias, ctx.stride, ctx.padding, ctx.dilation,
                                  deformable_groups, mask_is_logit=False, fresh_pack=ctx.needs_input_grad[3])
        ctx.save_for_backward(x_used, offset, mask, weight, bias)
        return y

    @staticmethod
    @once_differentiable
    def backward(ctx, grad_output):
        x, offset, mask, weight, bias = ctx.saved_tensors
        needs = ctx.needs_input_grad[:5]
        gx, goff, gmask, gw, gb = _backward_impl(x, offset, mask, weight, grad_output, ctx.stride, ctx.padding,
                                                 ctx.dilation, ctx.deformable_groups, False, needs)
        return (_cast(gx, x), _cast(goff, offset), _cast(gmask, mask), _cast(gw, weight),
                _cast(gb, bias) if bias is not None else None, None, None, None, None)


class _DCNFused(Function):
    """DCN.forward's body with the chunk/cat/sigmoid of the offset-conv output fused into the kernels:
    takes the raw 3*dg*kh*kw-channel tensor ``om`` (offsets | mask logits) and returns d/d(om)."""

    @staticmethod
    def forward(ctx, input, om, weight, bias, stride, padding, dilation, deformable_groups):
        ctx.stride, ctx.padding, ctx.dilation = stride, padding, dilation
        ctx.deformable_groups = deformable_groups
        n_off = om.shape[1] // 3 * 2
        y, x_used = _forward_impl(input, om[:, :n_off], om[:, n_off:], weight, bias, stride, padding, dilation,
                                  deformable_groups, mask_is_logit=True, fresh_pack=ctx.needs_input_grad[2])
        ctx.save_for_backward(x_used, om, weight, bias)
        return y

    @staticmethod
    @once_differentiable
    def backward(ctx, grad_output):
        x, om, weight, bias = ctx.saved_tensors
        n_off = om.shape[1] // 3 * 2
        needs = ctx.needs_input_grad
        # d/d(om) in om's own dtype when that is bf16 (autocast): the kernels round once on store, no 27-channel cast pass
        gom_dt = torch.bfloat16 if om.dtype == torch.bfloat16 else torch.float32
        gom = torch.empty_like(om, dtype=gom_dt) if needs[1] else None
        # the kernels write straight into the two channel ranges of d/d(om)
        dev = x.device
        L = _lib.lib()
        w32 = _as_f32_contig(weight)
        gout = _grad_like(grad_output, x)
        with torch.cuda.device(dev):
            gx = torch.empty_like(x, dtype=torch.float32) if needs[0] else None
            gw = torch.empty(weight.shape, dtype=torch.float32, device=dev) if needs[2] else None
            gb = torch.empty(weight.shape[0], dtype=torch.float32, device=dev) if (needs[3] and bias is not None) else None
            b = _lib.DcnBwdArgs()
            b.f = _fwd_args(x, om[:, :n_off], om[:, n_off:], w32, None, ctx.stride, ctx.padding, ctx.dilation,
                            ctx.deformable_groups, True, None, None)
            b.gout = _lib.t4(gout)
            b.gx = _lib.t4(gx)
            b.goffset = _lib.t4(gom[:, :n_off] if gom is not None else None)
            b.gmask = _lib.t4(gom[:, n_off:] if gom is not None else None)
            b.gom_dtype = _lib.BF16 if gom_dt == torch.bfloat16 else _lib.F32
            b.gweight = gw.data_ptr() if gw is not None else None
            b.gbias = gb.data_ptr() if gb is not None else None
            ws = _lib.workspace(L.cpoly_dcn_backward_workspace_bytes(byref(b)), dev)
            b.workspace, b.workspace_bytes = ws.data_ptr(), ws.numel()
            _lib.check(L.cpoly_dcn_backward(byref(b), _lib.stream_ptr(dev)), "cpoly_dcn_backward")
        return (_cast(gx, x), _cast(gom, om), _cast(gw, weight), _cast(gb, bias) if bias is not None else None,
                None, None, None, None)


dcn_v2_conv = _DCNv2.apply


class DCNv2(nn.Module):
    """Modulated deformable convolution whose offsets and mask are supplied by the caller."""

    def __init__(self, in_channels, out_channels, kernel_size, stride, padding, dilation=1, deformable_groups=1):
        super().__init__()
        self.in_channels = in_channels
        self.out_channels = out_channels
        self.kernel_size = _pair(kernel_size)
        self.stride = _pair(stride)
        self.padding = _pair(padding)
        self.dilation = _pair(dilation)
        self.deformable_groups = deformable_groups
        self.weight = nn.Parameter(torch.empty(out_channels, in_channels, *self.kernel_size))
        self.bias = nn.Parameter(torch.empty(out_channels))
        self.reset_parameters()

    def reset_parameters(self):
        fan = self.in_channels * self.kernel_size[0] * self.kernel_size[1]
        bound = 1.0 / math.sqrt(fan)
        with torch.no_grad():
            self.weight.uniform_(-bound, bound)
            self.bias.zero_()

    def forward(self, input, offset, mask):
        return dcn_v2_conv(input, offset, mask, self.weight, self.bias, self.stride, self.padding, self.dilation,
                           self.deformable_groups)


class DCN(DCNv2):
    """DCNv2 that predicts its own offsets/mask with a zero-initialised regular conv (a2 in SURVEY §8)."""

    def __init__(self, in_channels, out_channels, kernel_size, stride, padding, dilation=1, deformable_groups=1):
        super().__init__(in_channels, out_channels, kernel_size, stride, padding, dilation, deformable_groups)
        n = self.deformable_groups * 3 * self.kernel_size[0] * self.kernel_size[1]
        self.conv_offset_mask = nn.Conv2d(self.in_channels, n, kernel_size=self.kernel_size, stride=self.stride,
                                          padding=self.padding, bias=True)
        self.init_offset()

    def init_offset(self):
        with torch.no_grad():
            self.conv_offset_mask.weight.zero_()
            self.conv_offset_mask.bias.zero_()

    def _offset_conv_image(self):
        """(packed bf16 weight image of conv_offset_mask for cpoly_offset_conv_forward, fp32 bias padded to 32) or
        None when the shape has no fast path: a derived cache, never in state_dict."""
        w, b = self.conv_offset_mask.weight, self.conv_offset_mask.bias
        L = _lib.lib()
        nbytes = L.cpoly_offset_conv_packed_bytes(self.in_channels)
        if (nbytes == 0 or w.shape[0] > 32 or tuple(w.shape[2:]) != (3, 3) or self.stride != (1, 1) or self.padding != (1, 1) or
                self.dilation != (1, 1)):
            return None
        tag = (w._version, w.data_ptr(), w.dtype, w.device, b._version, b.data_ptr())
        hit = getattr(self, "_oc_cache", None)
        if hit is None or hit[0] != tag:
            w32 = w.detach().float().contiguous()
            packed = torch.empty(nbytes, dtype=torch.uint8, device=w.device)
            _lib.check(L.cpoly_offset_conv_pack_weights(w32.data_ptr(), packed.data_ptr(), self.in_channels, w.shape[0],
                                                        _lib.stream_ptr(w.device)), "cpoly_offset_conv_pack_weights")
            bias = torch.zeros(32, dtype=torch.float32, device=w.device)
            bias[:w.shape[0]] = b.detach().float()
            hit = (tag, packed, bias)
            self._oc_cache = hit
        return hit[1], hit[2]

    def offset_conv(self, input):
        """``conv_offset_mask(input)`` on our kernel (csrc/offset_conv.cu): (B,32,H,W) fp32 channels_last, channels
        27..31 zero, bias included; None when the input does not qualify (then the caller uses the nn.Conv2d)."""
        if not (input.is_cuda and input.dtype == torch.bfloat16 and input.dim() == 4):
            return None
        img = self._offset_conv_image()
        if img is None:
            return None
        if not input.is_contiguous(memory_format=torch.channels_last):
            input = input.contiguous(memory_format=torch.channels_last)
        B, C, H, W = input.shape
        with torch.cuda.device(input.device):
            om = torch.empty((B, 32, H, W), dtype=torch.float32, device=input.device, memory_format=torch.channels_last)
            a = _lib.OffsetConvArgs()
            a.B, a.Cin, a.H, a.W, a.om_dtype = B, C, H, W, _lib.F32
            a.x, a.weight_packed, a.bias, a.om = input.data_ptr(), img[0].data_ptr(), img[1].data_ptr(), om.data_ptr()
            _lib.check(_lib.lib().cpoly_offset_conv_forward(byref(a), _lib.stream_ptr(input.device)), "cpoly_offset_conv_forward")
            _lib.LAUNCHES["offset_conv"] += 1
        return om

    def _padded_offset_conv(self):
        """(weight padded to a multiple of 8 output channels, fp32 bias): derived cache, never in state_dict."""
        w, b = self.conv_offset_mask.weight, self.conv_offset_mask.bias
        tag = (w._version, w.data_ptr(), w.dtype, w.device, b._version, b.data_ptr())
        hit = getattr(self, "_om_cache", None)
        if hit is None or hit[0] != tag:
            n = w.shape[0]
            npad = (n + 7) // 8 * 8
            wp = torch.zeros((npad,) + tuple(w.shape[1:]), dtype=w.dtype, device=w.device)
            wp[:n] = w.detach()
            wp = wp.contiguous(memory_format=torch.channels_last)
            hit = (tag, wp, b.detach().float().contiguous())
            self._om_cache = hit
        return hit[1], hit[2]

    def forward(self, input):
        fused = getattr(self, "_fused_epilogue", None)
        if fused is not None:
            # inference-only DeformConv fusion installed by centerpoly_b200.fuse: eval-mode BN + ReLU in the epilogue,
            # the offset conv runs bias-free with its 27 output channels padded to 32 (no cuDNN padding kernels, no
            # separate bias add) and the kernel adds conv_offset_mask.bias itself.
            if torch.is_grad_enabled() and (input.requires_grad or self.weight.requires_grad):
                raise RuntimeError("the fused DeformConv epilogue is inference-only; rebuild the model for training")
            scale, shift, relu = fused
            n_off = self.deformable_groups * 2 * self.kernel_size[0] * self.kernel_size[1]
            n_all = n_off // 2 * 3
            om = self.offset_conv(input) if _USE_OWN_OFFSET_CONV else None
            if om is not None:      # our kernel: fp32 offsets / mask logits, bias already added
                y, _ = _forward_impl(input, om[:, :n_off], om[:, n_off:n_all], self.weight, shift, self.stride, self.padding,
                                     self.dilation, self.deformable_groups, True, out_scale=scale, relu=relu)
                return y
            w_om, b_om = self._padded_offset_conv()
            om = torch.nn.functional.conv2d(input, w_om, None, self.stride, self.padding)
            y, _ = _forward_impl(input, om[:, :n_off], om[:, n_off:n_all], self.weight, shift, self.stride, self.padding,
                                 self.dilation, self.deformable_groups, True, out_scale=scale, relu=relu, om_bias=b_om)
            return y
        if not torch.is_grad_enabled() and _USE_OWN_OFFSET_CONV and not torch.is_autocast_enabled():
            om = self.offset_conv(input)       # inference through the plain module: same kernel as the fused path
            if om is not None:
                y, _ = _forward_impl(input, om[:, :18], om[:, 18:27], self.weight, self.bias, self.stride, self.padding,
                                     self.dilation, self.deformable_groups, True)
                return y
        om = self.conv_offset_mask(input)
        return _DCNFused.apply(input, om, self.weight, self.bias, self.stride, self.padding, self.dilation,
                               self.deformable_groups)
