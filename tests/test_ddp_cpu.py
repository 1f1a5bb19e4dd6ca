"""CPU, world_size 2, gloo: the data-parallel step's only exchange (bucketed gradient all-reduce) gives every rank
the gradients of the concatenated batch — the reference's nn.DataParallel semantics (src/lib/models/data_parallel.py,
loss averaged over replicas base_trainer.py:95).  The DCN kernels need a GPU, so the host logic is exercised on a
small conv net; the GPU path of the same trainer is covered by tests/test_gpu_train.py."""
import os
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp
from torch import nn

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _net():
    torch.manual_seed(7)
    return nn.Sequential(nn.Conv2d(3, 8, 3, padding=1), nn.ReLU(), nn.Conv2d(8, 8, 3, padding=1), nn.ReLU(),
                         nn.Conv2d(8, 4, 1), nn.Flatten(), nn.LazyLinear(5))


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    from centerpoly_b200.harness.train import BucketedAllReduce, broadcast_module_state
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.manual_seed(0)
    x = torch.randn(4, 3, 8, 8)
    y = torch.randn(4, 5)
    net = _net()
    net(x)  # materialise the lazy layer on every rank
    if rank == 1:   # a rank-dependent init / a checkpoint loaded on one rank only: the broadcast must repair it
        with torch.no_grad():
            for p in net.parameters():
                p.add_(1.0)
    assert broadcast_module_state(net) == len(list(net.parameters()))
    red = BucketedAllReduce(list(net.parameters()), bucket_bytes=2048)   # several small buckets
    assert len(red.buckets) > 1
    for step in range(2):                                               # buckets must re-arm between steps
        net.zero_grad(set_to_none=True)
        xs, ys = x[rank * 2:(rank + 1) * 2], y[rank * 2:(rank + 1) * 2]
        loss = ((net(xs) - ys) ** 2).mean()
        loss.backward()
        red.finish()
    assert all(p.grad.data_ptr() == v.data_ptr() for b in red.buckets for p, v in b["views"])   # no copy back: grad IS the bucket slice
    q.put((rank, [p.grad.clone().numpy() for p in net.parameters()]))   # by value: the worker may exit before the parent reads
    dist.barrier()
    dist.destroy_process_group()


def test_bucketed_allreduce_matches_full_batch_gradients():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    torch.manual_seed(0)
    x = torch.randn(4, 3, 8, 8)
    y = torch.randn(4, 5)
    net = _net()
    net(x)
    # mean over ranks of per-rank mean losses == mean over the full batch (equal shard sizes)
    ((net(x) - y) ** 2).mean().backward()
    for g0, g1, p in zip(got[0], got[1], net.parameters()):
        torch.testing.assert_close(torch.from_numpy(g0), torch.from_numpy(g1), rtol=0, atol=0)
        torch.testing.assert_close(torch.from_numpy(g0), p.grad, rtol=1e-5, atol=1e-6)


def test_single_process_is_a_no_op():
    sys.path.insert(0, ROOT)
    from centerpoly_b200.harness.train import BucketedAllReduce
    net = _net()
    x = torch.randn(2, 3, 8, 8)
    net(x)
    red = BucketedAllReduce(list(net.parameters()))
    net(x).sum().backward()
    want = [p.grad.clone() for p in net.parameters()]
    red.finish()
    for w, p in zip(want, net.parameters()):
        assert torch.equal(w, p.grad)


def test_loss_pieces_match_reference_definitions():
    """focal / RegL1 restatements vs direct formulas on a tiny case."""
    sys.path.insert(0, ROOT)
    from centerpoly_b200.harness.train import focal_loss, gather_at, reg_l1, synthetic_batch, polydet_l1_loss
    b = synthetic_batch(2, 32, 64, n_obj=3, seed=1)
    feat = torch.randn(2, 6, 8, 16)
    ind = b["ind"][:, :5]
    want = feat.permute(0, 2, 3, 1).contiguous().view(2, -1, 6).gather(1, ind.unsqueeze(2).expand(2, 5, 6))  # utils.py:12-26
    assert torch.equal(gather_at(feat, ind), want)
    out = dict(hm=torch.randn(2, 8, 8, 16), poly=torch.randn(2, 64, 8, 16), pseudo_depth=torch.randn(2, 1, 8, 16), reg=torch.randn(2, 2, 8, 16))
    loss = polydet_l1_loss(out, b)
    assert loss.ndim == 0 and torch.isfinite(loss)
