"""GPU: `cpoly_decode_static` (the allocation-free decode entry point, ABI v9; no counterpart in the reference) against
`cpoly_decode`, both called through the C ABI.  Collected last on purpose (file name).
"""
from ctypes import byref

import pytest
import torch

from centerpoly_b200 import _lib

pytestmark = pytest.mark.gpu


def _heads(seed, B, C, H, W, n_poly, dtype):
    g = torch.Generator(device="cuda").manual_seed(seed)
    heat = torch.sigmoid(torch.randn(B, C, H, W, device="cuda", generator=g) * 2.0 - 3.0).to(dtype)
    polys = (torch.randn(B, n_poly, H, W, device="cuda", generator=g) * 10.0).to(dtype)
    depth = torch.rand(B, 1, H, W, device="cuda", generator=g).to(dtype)
    reg = torch.rand(B, 2, H, W, device="cuda", generator=g).to(dtype)
    return heat, polys, depth, reg


def _args(heads, K, rep, out):
    heat, polys, depth, reg = heads
    B, C, H, W = heat.shape
    a = _lib.DecodeArgs()
    a.B, a.C, a.H, a.W, a.n_poly, a.K = B, C, H, W, polys.shape[1], K
    a.rep = _lib.REP[rep]
    a.heat_is_logit = 0
    a.dtype = _lib.dtype_code(heat)
    a.heat, a.polys, a.depth, a.reg = _lib.t4(heat), _lib.t4(polys), _lib.t4(depth), _lib.t4(reg)
    a.out = out.data_ptr()
    a.out_inds = None
    a.post_affine = None
    return a


def _run_reference_entry(L, heads, K, rep):
    B, C, H, W = heads[0].shape
    out = torch.full((B, K, heads[1].shape[1] + 7), float("nan"), device="cuda")
    ws = torch.empty(L.cpoly_decode_workspace_bytes(B, C, H, W, K) + 256, dtype=torch.uint8, device="cuda")
    a = _args(heads, K, rep, out)
    base = (ws.data_ptr() + 255) & ~255
    a.workspace, a.workspace_bytes = base, ws.numel() - (base - ws.data_ptr())
    _lib.check(L.cpoly_decode(byref(a), _lib.stream_ptr(heads[0].device)), "cpoly_decode")
    torch.cuda.synchronize()
    return out


def _run_static_entry(L, heads, K, rep):
    B = heads[0].shape[0]
    out = torch.full((B, K, heads[1].shape[1] + 7), float("nan"), device="cuda")
    a = _args(heads, K, rep, out)
    a.workspace, a.workspace_bytes = None, 0          # ignored by the entry point
    _lib.check(L.cpoly_decode_static(byref(a), _lib.stream_ptr(heads[0].device)), "cpoly_decode_static")
    torch.cuda.synchronize()
    return out


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_static_decode_matches_decode_on_two_calls_in_a_row(dtype):
    L = _lib.lib()
    assert L.cpoly_decode_static_max_bytes() == 4 << 20
    # two different frames back to back: the second call sees the first call's leftovers in the shared scratch
    cases = [(3, 2, 8, 128, 256, 32, 100, "cartesian"), (4, 3, 8, 96, 160, 32, 40, "polar")]
    for seed, B, C, H, W, n_poly, K, rep in cases:
        heads = _heads(seed, B, C, H, W, n_poly, dtype)
        want = _run_reference_entry(L, heads, K, rep)
        got = _run_static_entry(L, heads, K, rep)
        assert not torch.isnan(got).any()
        # integer / index work and the same arithmetic in the same order: bit-exact (the bar of tests/test_gpu_decode.py)
        assert torch.equal(got, want), (seed, float((got - want).abs().max()))
    # and the first frame again, after a smaller one has overwritten part of the scratch
    heads = _heads(3, 2, 8, 128, 256, 32, dtype)
    assert torch.equal(_run_static_entry(L, heads, 100, "cartesian"), _run_reference_entry(L, heads, 100, "cartesian"))


def test_static_decode_rejects_shapes_beyond_its_scratch():
    L = _lib.lib()
    B, C, H, W, K = 64, 8, 128, 256, 100
    assert L.cpoly_decode_workspace_bytes(B, C, H, W, K) > L.cpoly_decode_static_max_bytes()
    heads = _heads(5, 1, C, H, W, 32, torch.float32)
    out = torch.empty(B, K, 39, device="cuda")
    a = _args(heads, K, "cartesian", out)
    a.B = B                                               # never launched: the size check comes first
    rc = L.cpoly_decode_static(byref(a), _lib.stream_ptr(heads[0].device))
    assert rc == 2                                        # CPOLY_EWORKSPACE
    msg = L.cpoly_last_error().decode()
    assert "cpoly_decode_static" in msg and "built-in array" in msg
