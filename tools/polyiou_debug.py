"""Per-object comparison of the GPU polygon-IoU loss with the CPU oracle (debug aid)."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden"))
from oracle import polyloss_oracle as po
import centerpoly_b200 as cp
from make_polyiou_cases import star_case

def one(pred, mask, tgt):
    B, M, C = pred.shape
    H, W = 16, 24
    g = torch.Generator().manual_seed(0)
    ind = torch.stack([torch.randperm(H * W, generator=g)[:M] for _ in range(B)])
    out = torch.zeros(B, C, H * W)
    out.scatter_(2, ind.unsqueeze(1).expand(B, C, M), torch.as_tensor(pred).transpose(1, 2))
    out = out.view(B, C, H, W).cuda().requires_grad_(True)
    loss, iou, status = cp.poly_iou_loss(out, torch.as_tensor(mask).cuda(), ind.cuda(), torch.as_tensor(tgt).cuda(), return_details=True)
    loss.backward()
    grad = out.grad.reshape(B, C, H * W).gather(2, ind.cuda().unsqueeze(1).expand(B, C, M)).transpose(1, 2).cpu().numpy()
    want = po.poly_iou_loss_oracle(pred, mask, tgt)
    for b in range(B):
        for k in range(M):
            if not mask[b, k]:
                continue
            e = np.abs(grad[b, k] - want["grad"][b, k]).max() / max(np.abs(want["grad"][b, k]).max(), 1e-12)
            flag = "  <<<<" if e > 1e-3 else ""
            print(f"obj {b},{k}: iou {iou[b,k].item():.6f} vs {want['iou'][b,k]:.6f}  grad err {e:.2e}{flag}")
            if flag:
                p = torch.tensor(pred[b, k]).view(-1, 2)
                order = torch.sort(p[:, 1], stable=True, dim=0)[1]
                sp = torch.stack((p[order, 0].abs(), p[order, 1]), 1)
                tg = torch.tensor(tgt[b, k]).view(-1, 2)
                ins, outs = po.crossings(sp, tg)
                st, verts = po.walk(ins, outs, len(sp), len(tg))
                print("   ins", ins, "outs", outs)
                print("   verts", verts)
                bad = np.argsort(-np.abs(grad[b, k] - want["grad"][b, k]))[:6]
                print("   worst channels (orig order)", bad, grad[b, k][bad], want["grad"][b, k][bad], "order", order.tolist())

for path in sys.argv[1:]:
    if path.endswith(".npz"):
        z = np.load(path)
        print("==", path)
        one(z["pred"], z["mask"], z["target"])
    else:
        seed, B, M, N, n_obj, rn, tn = [float(v) for v in path.split(",")]
        pred, mask, _, tgt, _ = star_case(int(seed), int(B), int(M), int(N), int(n_obj), rn, tn)
        print("== case", path)
        one(pred.numpy(), mask.numpy(), tgt.numpy())
