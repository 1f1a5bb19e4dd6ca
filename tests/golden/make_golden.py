"""Regenerates tests/golden/*.npz.  Run in the BUILD container only (needs /root/reference):

    python tests/golden/make_golden.py

* decode_*.npz : inputs + outputs of the UNMODIFIED reference ``polydet_decode``
  (/root/reference/src/lib/models/decode.py:512-670) imported from the read-only reference tree.
* dcn_*.npz    : inputs + outputs + all five gradients of ``torchvision.ops.deform_conv2d`` in fp64
  on the CPU — the stand-in BASELINE.json names for the missing DCNv2 submodule
  (/root/reference/.SUBMODULES.json:8-14).

* post_*.npz   : inputs + outputs of the UNMODIFIED reference ``polydet_post_process``
  (/root/reference/src/lib/utils/post_process.py:105-122) on the stored decode outputs.

The fixtures are what travels to the GPU box; /root/reference does not.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/src/lib"


def tie_free_heat(gen, B, C, H, W):
    """Post-sigmoid heat whose values are all distinct (so torch.topk's tie order never matters):
    a random permutation of evenly spaced logits."""
    n = B * C * H * W
    logits = torch.linspace(-6.0, 1.5, n)[torch.randperm(n, generator=gen)]
    heat = torch.sigmoid(logits)
    assert torch.unique(heat).numel() == n, "collision"
    return heat.view(B, C, H, W).contiguous()


def make_decode():
    sys.path.insert(0, REF)
    from models.decode import polydet_decode  # the reference, unmodified

    cases = [
        # name, B, C, H, W, N, K, rep, with_reg
        ("small_cart_reg", 2, 3, 12, 20, 5, 8, "cartesian", True),
        ("small_cart_noreg", 2, 3, 12, 20, 5, 8, "cartesian", False),
        ("small_polar_reg", 2, 3, 12, 20, 5, 8, "polar", True),
        ("small_polarfixed_noreg", 1, 4, 9, 11, 8, 16, "polar_fixed", False),
        ("small_polarconfined_reg", 1, 2, 8, 8, 4, 6, "polar_confined", True),   # decodes as cartesian
        ("mid_cart_reg", 1, 8, 32, 48, 32, 100, "cartesian", True),
        ("mid_polar_reg", 2, 9, 24, 40, 16, 128, "polar", True),
    ]
    for idx, (name, B, C, H, W, N, K, rep, with_reg) in enumerate(cases):
        gen = torch.Generator().manual_seed(1000 + idx)
        heat = tie_free_heat(gen, B, C, H, W)
        if rep == "polar":
            polys = torch.empty(B, 2 * N, H, W)
            polys[:, 0::2] = (torch.randn(B, N, H, W, generator=gen) * 10).abs()
            polys[:, 1::2] = torch.rand(B, N, H, W, generator=gen) * 2 * np.pi
        else:
            polys = torch.randn(B, 2 * N, H, W, generator=gen) * 6
        depth = torch.rand(B, 1, H, W, generator=gen) * 50
        reg = torch.rand(B, 2, H, W, generator=gen) if with_reg else None
        out = polydet_decode(heat.clone(), polys.clone(), depth.clone(),
                             reg=None if reg is None else reg.clone(), cat_spec_poly=False, K=K, rep=rep)
        assert out.shape == (B, K, 2 * N + 7)
        # the K-boundary must be tie-free too: all positive scores distinct
        sc = out[..., 4]
        for b in range(B):
            pos = sc[b][sc[b] > 0]
            assert torch.unique(pos).numel() == pos.numel()
        np.savez_compressed(
            os.path.join(HERE, f"decode_{name}.npz"), heat=heat.numpy(), polys=polys.numpy(),
            depth=depth.numpy(), reg=(reg.numpy() if reg is not None else np.zeros(0, np.float32)),
            K=np.int64(K), rep=np.array(rep), out=out.numpy())
        print("decode", name, tuple(out.shape))


def make_dcn():
    from torchvision.ops import deform_conv2d

    cases = [
        # name, B, Cin, Cout, H, W, k, stride, pad, dil, dg, offset_mode
        ("3x3_random", 2, 4, 6, 7, 9, 3, 1, 1, 1, 1, "random"),
        ("3x3_zero_offset", 1, 3, 5, 6, 6, 3, 1, 1, 1, 1, "zero"),
        ("3x3_far_oob", 1, 4, 4, 5, 8, 3, 1, 1, 1, 1, "far"),
        ("3x3_dg2_s2", 2, 4, 3, 9, 8, 3, 2, 1, 1, 2, "random"),
        ("3x3_dil2", 1, 2, 3, 8, 7, 3, 1, 2, 2, 1, "random"),
        ("3x3_integer_offsets", 1, 3, 4, 6, 7, 3, 1, 1, 1, 1, "integer"),
    ]
    for idx, (name, B, Cin, Cout, H, W, k, s, p, d, dg, mode) in enumerate(cases):
        gen = torch.Generator().manual_seed(2000 + idx)
        Ho = (H + 2 * p - (d * (k - 1) + 1)) // s + 1
        Wo = (W + 2 * p - (d * (k - 1) + 1)) // s + 1
        x = torch.randn(B, Cin, H, W, generator=gen, dtype=torch.float64)
        if mode == "zero":
            off = torch.zeros(B, dg * 2 * k * k, Ho, Wo, dtype=torch.float64)
        elif mode == "far":
            off = (torch.rand(B, dg * 2 * k * k, Ho, Wo, generator=gen, dtype=torch.float64) - 0.5) * 4 * W
        elif mode == "integer":
            off = torch.randint(-2, 3, (B, dg * 2 * k * k, Ho, Wo), generator=gen).double()
        else:
            off = torch.randn(B, dg * 2 * k * k, Ho, Wo, generator=gen, dtype=torch.float64) * 2.0
        msk = torch.sigmoid(torch.randn(B, dg * k * k, Ho, Wo, generator=gen, dtype=torch.float64))
        w = (torch.rand(Cout, Cin, k, k, generator=gen, dtype=torch.float64) - 0.5) * 0.6
        b = torch.randn(Cout, generator=gen, dtype=torch.float64) * 0.1
        gout = torch.randn(B, Cout, Ho, Wo, generator=gen, dtype=torch.float64)
        leaves = [t.clone().requires_grad_(True) for t in (x, off, msk, w, b)]
        y = deform_conv2d(leaves[0], leaves[1], leaves[3], leaves[4], stride=s, padding=p, dilation=d, mask=leaves[2])
        y.backward(gout)
        np.savez_compressed(
            os.path.join(HERE, f"dcn_{name}.npz"), x=x.numpy(), offset=off.numpy(), mask=msk.numpy(),
            weight=w.numpy(), bias=b.numpy(), gout=gout.numpy(), y=y.detach().numpy(),
            gx=leaves[0].grad.numpy(), goffset=leaves[1].grad.numpy(), gmask=leaves[2].grad.numpy(),
            gw=leaves[3].grad.numpy(), gb=leaves[4].grad.numpy(),
            geom=np.array([s, p, d, dg], dtype=np.int64))
        print("dcn", name, tuple(y.shape))


def make_post():
    """post_*.npz: the UNMODIFIED reference ``polydet_post_process`` (src/lib/utils/post_process.py:105-122, which
    calls ``transform_preds`` / ``get_affine_transform`` / ``affine_transform`` of src/lib/utils/image.py) applied to
    the outputs of the reference decode stored in decode_*.npz, for the two (c, s) conventions of
    src/lib/detectors/base_detector.py:55-63 (fix_res: scalar s; keep_res: s = [inp_w, inp_h])."""
    if REF not in sys.path:
        sys.path.insert(0, REF)
    from utils.post_process import polydet_post_process  # the reference, unmodified
    from utils.image import get_affine_transform

    cases = [
        # name, decode fixture, num_classes, original (h, w), fix_res, down_ratio
        ("fixres_small", "decode_small_cart_reg", 3, (37, 91), True, 4),
        ("keepres_small", "decode_small_polar_reg", 3, (45, 77), False, 4),
        ("fixres_mid", "decode_mid_cart_reg", 8, (1024, 2048), True, 4),
        ("keepres_mid", "decode_mid_polar_reg", 8, (125, 190), False, 4),
    ]
    for name, fix, ncls, (h, w), fix_res, down in cases:
        d = np.load(os.path.join(HERE, fix + ".npz"))
        dets = d["out"].astype(np.float32)
        B = dets.shape[0]
        fh, fw = d["heat"].shape[2:]
        inp_h, inp_w = fh * down, fw * down
        cs, ss = [], []
        for i in range(B):   # detectors/base_detector.py:51-63 (scale 1); a different image size per batch entry
            hh, ww = h + 3 * i, w + 5 * i
            if fix_res:
                cs.append(np.array([ww / 2., hh / 2.], dtype=np.float32))
                ss.append(max(hh, ww) * 1.0)
            else:
                cs.append(np.array([ww // 2, hh // 2], dtype=np.float32))
                ss.append(np.array([inp_w, inp_h], dtype=np.float32))
        out_h, out_w = inp_h // down, inp_w // down
        ret = polydet_post_process(dets.copy(), cs, ss, out_h, out_w, ncls)
        trans = np.stack([get_affine_transform(cs[i], ss[i], 0, (out_w, out_h), inv=1) for i in range(B)])
        blob = dict(dets=dets, c=np.stack(cs), s=np.stack([np.broadcast_to(np.asarray(x, dtype=np.float32), (2,)) for x in ss]),
                    s_is_scalar=np.array(fix_res), out_h=np.array(out_h), out_w=np.array(out_w), num_classes=np.array(ncls), trans=trans)
        for i in range(B):
            for j in range(1, ncls + 1):
                blob[f"ret_{i}_{j}"] = np.array(ret[i][j], dtype=np.float32).reshape(-1, dets.shape[2] - 1)
        np.savez_compressed(os.path.join(HERE, f"post_{name}.npz"), **blob)
        print("post", name, dets.shape, [len(ret[0][j]) for j in range(1, ncls + 1)])


if __name__ == "__main__":
    torch.set_num_threads(4)
    if "--post-only" not in sys.argv:
        make_decode()
        make_dcn()
    make_post()
