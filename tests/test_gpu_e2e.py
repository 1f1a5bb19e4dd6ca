"""G3 (parity protocol): the BENCHMARKED pipeline end to end - the inference engine (module rewrites
of centerpoly_b200.fuse -> our DCN / up-sampler / stem kernels -> polydet_decode_from_logits with the fused
polydet_post_process back-mapping, replayed from its CUDA graph) - against the UNFUSED network with the torchvision CPU
DCN, ``hm.sigmoid_()``, the decode oracle and the affine back-mapping (src/lib/detectors/polydet.py:25-60), on one
256x512 BN-calibrated frame with the same state_dict.  Three gates:

 (a) network: every head map of the engine against the CPU fp32 reference network, dense (all pixels), relative L2 - and
     beside it the SAME measure for the framework's own GPU pipeline at the same precision (the unfused network with
     torchvision's CUDA deform_conv2d; bf16 = torch.autocast, fp32 = cuDNN's default TF32 convolutions).  The random-init
     network is ill-conditioned (~55 layers, 16 data-dependent samplers: bf16 rounding alone moves the poly head by
     40 %), so the engine is gated RELATIVE to the framework: its error may not exceed 2x the framework's;
 (b) composition: the engine's detections (graph replay) == sigmoid -> decode oracle -> affine applied to the ENGINE's
     own head maps: class / cell / order exact, scores to 2 ulp, coordinates to 2e-6 - tie-aware (rows whose score is
     closer than 4 ulp to a neighbour are exempt).  (a) and (b) together pin what bench.py times;
 (c) detections MATCHED by class and position against the reference's detections: reported, loosely gated.
     A random-init heat map has std 0.19 in its logits and adjacent top-K scores ~5e-6 apart, so any network-level
     difference above that reshuffles the top-K SET (75 of 100 survive in fp32, 31 in bf16): set overlap measures the
     network's conditioning, not the kernels.

The gated network has offset producers of std 0.01 (offsets dominated by their bias, ~N(0,1) px): with the std-0.05
recipe of SURVEY A.5 the random-init neck amplifies a 1e-6 input perturbation to ~1e-2 at the heads
(tools/fusion_debug.py); that net is run too and its numbers are printed (informational)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

H, W, K = 256, 512, 100
HEADS = {"hm": 8, "poly": 64, "pseudo_depth": 1, "reg": 2}
NET_FACTOR = 2.0      # gate (a): engine error <= NET_FACTOR x the framework's error at the same precision (+ 1e-3)


def _decode_cpu(heads):
    """sigmoid -> polydet_decode oracle -> inverse affine of polydet_post_process, on CPU float arrays."""
    from oracle.decode_oracle import polydet_decode_oracle, sigmoid_f32
    import centerpoly_b200 as cp
    dets = polydet_decode_oracle(sigmoid_f32(heads["hm"]), heads["poly"], heads["pseudo_depth"], reg=heads["reg"], K=K + 1)
    t = cp.batch_inverse_affine([[W / 2.0, H / 2.0]], [float(max(H, W))], W // 4, H // 4)[0]     # detectors/polydet.py:45-50
    d = dets[0].astype(np.float64).copy()
    for cx, cy in [(0, 1), (2, 3)] + [(6 + 2 * v, 7 + 2 * v) for v in range(32)]:
        x, y = d[:, cx].copy(), d[:, cy].copy()
        d[:, cx] = t[0, 0] * x + t[0, 1] * y + t[0, 2]
        d[:, cy] = t[1, 0] * x + t[1, 1] * y + t[1, 2]
    return d


def _match(ours, ref):
    """Greedy one-to-one matching by class and vertex-0 position (within 2 image pixels = half a feature cell)."""
    pairs, used = [], set()
    for i in range(ref.shape[0]):
        cand = [(abs(ours[j, 6] - ref[i, 6]) + abs(ours[j, 7] - ref[i, 7]), j) for j in range(ours.shape[0])
                if j not in used and ours[j, 5] == ref[i, 5]]
        if cand:
            dist, j = min(cand)
            if dist < 4.0:
                pairs.append((j, i))
                used.add(j)
    return pairs


def _matched_stats(ours, ref):
    pairs = _match(ours, ref)
    if not pairs:
        return dict(matched=0)
    jo, ir = (np.array(v) for v in zip(*pairs))
    score = np.abs(ours[jo, 4] - ref[ir, 4]) / ref[ir, 4]
    ext = np.maximum(ref[ir, 2] - ref[ir, 0], ref[ir, 3] - ref[ir, 1])[:, None]      # polygon extent in pixels
    geom = np.abs(ours[jo, 6:70] - ref[ir, 6:70]) / np.maximum(ext, 1.0)
    return dict(matched=len(pairs), score_max=float(score.max()), score_med=float(np.median(score)), geom_max=float(geom.max()),
                geom_med=float(np.median(geom)))


@pytest.mark.parametrize("dtype", ["fp32", "bf16"])
@pytest.mark.parametrize("offset_std", [0.01, 0.05], ids=["gated-std0.01", "informational-std0.05"])
def test_g3_engine_vs_unfused_cpu_reference(dtype, offset_std):
    from centerpoly_b200.harness.dla34 import PolydetDLA34, calibrate_random_init
    from centerpoly_b200.harness.engine import PolydetEngine
    from oracle.tv_dcn import TVDCN
    torch.manual_seed(317)
    ref_net = PolydetDLA34(HEADS, dcn_cls=TVDCN)
    calibrate_random_init(ref_net, H, W, offset_std=offset_std)
    ours_net = PolydetDLA34(HEADS)
    ours_net.load_state_dict(ref_net.state_dict())
    frame = torch.randn(1, 3, H, W, generator=torch.Generator().manual_seed(11))
    with torch.no_grad():
        ref_heads = {k: v.numpy() for k, v in ref_net(frame)[-1].items()}
    eng = PolydetEngine(height=H, width=W, batch=1, dtype=dtype, K=K, n_slots=1, net=ours_net)
    eng.set_resident_frames([frame.cuda()])
    eng.infer_resident(0)
    eng.stream.synchronize()
    ours = eng.slots[0]["dets"][0].double().cpu().numpy()
    assert ours.shape == (K, 71) and np.isfinite(ours).all()

    # ---- (a) network: dense head maps
    heads = {k: v.float().cpu().numpy() for k, v in eng.heads_eager(0).items()}
    net_err = {}
    for k, want in ref_heads.items():
        got = heads[k]
        assert got.shape == want.shape
        net_err[k] = (float(np.linalg.norm(got - want) / np.linalg.norm(want)), float(np.abs(got - want).max() / np.abs(want).max()))
    print(f"\nG3 {dtype} offset_std={offset_std}: head maps (rel L2, rel max) {net_err}")
    # the framework's own GPU pipeline at the same precision, same measure
    fw_net = PolydetDLA34(HEADS, dcn_cls=TVDCN)
    fw_net.load_state_dict(ref_net.state_dict())
    fw_net = fw_net.cuda().eval()
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16, enabled=dtype == "bf16"):
        fw_heads = {k: v.float().cpu().numpy() for k, v in fw_net(frame.cuda())[-1].items()}
    fw_err = {k: float(np.linalg.norm(fw_heads[k] - ref_heads[k]) / np.linalg.norm(ref_heads[k])) for k in ref_heads}
    print(f"G3 {dtype} offset_std={offset_std}: framework GPU pipeline at the same precision, rel L2 {fw_err}")

    # ---- (b) composition: replayed detections == oracle decode + affine of the engine's OWN head maps
    want = _decode_cpu(heads)                       # K + 1 rows: the last one is the K boundary
    sw = want[:, 4]
    np.testing.assert_allclose(ours[:, 4], sw[:K], rtol=2.4e-7)
    compared = 0
    for i in range(K):
        gap = min(sw[i - 1] - sw[i] if i else np.inf, sw[i] - sw[i + 1])
        if gap > 4 * 6e-8 * sw[i]:
            assert ours[i, 5] == want[i, 5], f"class of row {i}"
            np.testing.assert_allclose(np.delete(ours[i], [4, 5]), np.delete(want[i], [4, 5]), rtol=2e-6, atol=2e-4, err_msg=f"row {i}")
            compared += 1
    print(f"G3 {dtype} offset_std={offset_std}: composition exact on {compared}/{K} tie-free rows")
    assert compared >= (K // 2 if dtype == "fp32" else 1)      # bf16 logits: 8 mantissa bits, most top scores tie

    # ---- (c) matched detections against the reference network's detections
    st = _matched_stats(ours, _decode_cpu(ref_heads)[:K])
    print(f"G3 {dtype} offset_std={offset_std}: matched {st}")
    if offset_std == 0.01:
        for k, (l2, mx) in net_err.items():
            assert l2 <= NET_FACTOR * fw_err[k] + 1e-3, f"{k}: engine rel L2 {l2:.3e} vs framework {fw_err[k]:.3e}"
        assert st["matched"] >= (50 if dtype == "fp32" else 15), st
