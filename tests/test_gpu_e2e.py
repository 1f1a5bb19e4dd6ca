"""G3 (parity protocol): the BENCHMARKED pipeline end to end - the inference engine (module rewrites
of centerpoly_b200.fuse -> our DCN / up-sampler / stem kernels -> polydet_decode_from_logits with the fused
polydet_post_process back-mapping, replayed from its CUDA graph) - against the UNFUSED network with the torchvision CPU
DCN, ``hm.sigmoid_()``, the decode oracle and the affine back-mapping (src/lib/detectors/polydet.py:25-60), on one
256x512 BN-calibrated frame with the same state_dict.

Because adjacent top-K scores can be 5e-6 apart while every DCN layer is allowed 1e-3 / 1e-2, rank swaps and
K-boundary substitutions are expected: the gate is on detections MATCHED by class and position.  The
network used for the gate has offset producers of std 0.01 (offsets dominated by their bias, ~N(0,1) px): with the
std-0.05 recipe of SURVEY A.5 the random-init neck amplifies a 1e-6 input perturbation to ~1e-2 at the heads
(tools/fusion_debug.py), which measures the network's conditioning, not the kernels; that net is run as well and its
numbers are printed (informational)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

H, W, K = 256, 512, 100
HEADS = {"hm": 8, "poly": 64, "pseudo_depth": 1, "reg": 2}


def _reference(net_cpu, frame):
    from oracle.decode_oracle import polydet_decode_oracle, sigmoid_f32
    import centerpoly_b200 as cp
    with torch.no_grad():
        out = net_cpu(frame)[-1]
    heat = sigmoid_f32(out["hm"].numpy())
    dets = polydet_decode_oracle(heat, out["poly"].numpy(), out["pseudo_depth"].numpy(), reg=out["reg"].numpy(), K=K)
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


def _stats(ours, ref):
    pairs = _match(ours, ref)
    if not pairs:
        return dict(matched=0, score=np.inf, geom=np.inf)
    jo, ir = (np.array(v) for v in zip(*pairs))
    score = np.abs(ours[jo, 4] - ref[ir, 4]) / ref[ir, 4]
    ext = np.maximum(ref[ir, 2] - ref[ir, 0], ref[ir, 3] - ref[ir, 1])[:, None]      # polygon extent in pixels
    geom = np.abs(ours[jo, 6:70] - ref[ir, 6:70]) / np.maximum(ext, 1.0)
    return dict(matched=len(pairs), score=float(score.max()), score_med=float(np.median(score)), geom=float(geom.max()),
                geom_med=float(np.median(geom)), depth=float(np.abs(ours[jo, 70] - ref[ir, 70]).max()))


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
    ref = _reference(ref_net, frame)
    eng = PolydetEngine(height=H, width=W, batch=1, dtype=dtype, K=K, n_slots=1, net=ours_net)
    eng.set_resident_frames([frame.cuda()])
    eng.infer_resident(0)
    eng.stream.synchronize()
    ours = eng.slots[0]["dets"][0].double().cpu().numpy()
    assert ours.shape == ref.shape == (K, 71) and np.isfinite(ours).all()
    st = _stats(ours, ref)
    print(f"\nG3 {dtype} offset_std={offset_std}: {st}")
    if offset_std == 0.01:
        tol = 1e-3 if dtype == "fp32" else 1e-2
        # top-K overlap: substitutions only at the K boundary; matched detections: score / vertices / depth
        assert st["matched"] >= (90 if dtype == "fp32" else 70), st
        assert st["score_med"] <= tol and st["geom_med"] <= tol, st
        assert st["score"] <= 20 * tol and st["geom"] <= 20 * tol, st
    else:
        assert st["matched"] >= 10, st
