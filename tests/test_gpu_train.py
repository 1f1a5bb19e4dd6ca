"""GPU: the data-parallel trainer drives the full network (16 DCN forwards + backwards through the C ABI) and learns."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dtype", [None, torch.bfloat16], ids=["fp32", "bf16"])
def test_training_steps_reduce_the_loss(dtype):
    from centerpoly_b200.harness.dla34 import PolydetDLA34, calibrate_random_init
    from centerpoly_b200.harness.train import DataParallelTrainer, synthetic_batch
    torch.manual_seed(317)
    net = PolydetDLA34().cuda().to(memory_format=torch.channels_last)
    calibrate_random_init(net, 64, 128)
    trainer = DataParallelTrainer(net, lr=5e-4, dtype=dtype)
    batch = synthetic_batch(2, 128, 256, n_obj=6, device="cuda", seed=3)
    batch["input"] = batch["input"].contiguous(memory_format=torch.channels_last)
    losses = [float(trainer.step(batch)) for _ in range(12)]
    assert all(l == l and abs(l) < 1e6 for l in losses), losses
    assert min(losses[-3:]) < losses[0], losses
    # every DCN parameter received a gradient through the tensor-core backward
    for n, p in net.named_parameters():
        if ".conv.weight" in n or "conv_offset_mask" in n:
            assert p.grad is not None and torch.isfinite(p.grad).all(), n


def test_training_with_polygon_iou_loss():
    """configs[3] names "L1 + polygon IoU": polar targets, PolyLoss l1+iou with the IoU arm on cpoly_poly_iou_*."""
    from centerpoly_b200 import _lib
    from centerpoly_b200.harness.dla34 import PolydetDLA34, calibrate_random_init
    from centerpoly_b200.harness.train import DataParallelTrainer, polydet_l1_iou_loss, synthetic_batch
    torch.manual_seed(317)
    net = PolydetDLA34().cuda().to(memory_format=torch.channels_last)
    calibrate_random_init(net, 64, 128)
    from centerpoly_b200.neck import use_training_upsamplers
    keys = list(net.state_dict().keys())
    assert use_training_upsamplers(net) == 8 and list(net.state_dict().keys()) == keys    # IDAUp up_i on our fwd + bwd kernels
    trainer = DataParallelTrainer(net, lr=5e-4, dtype=torch.bfloat16, loss_fn=polydet_l1_iou_loss)
    batch = synthetic_batch(2, 128, 256, n_obj=6, device="cuda", seed=3, rep="polar")
    batch["input"] = batch["input"].contiguous(memory_format=torch.channels_last)
    before = _lib.LAUNCHES["poly_iou"]
    losses = [float(trainer.step(batch)) for _ in range(12)]
    assert _lib.LAUNCHES["poly_iou"] == before + 12 * 3          # forward (2 kernels) + backward (1) every step
    assert _lib.LAUNCHES["upsample_bwd"] >= 12 * 16              # data + weight gradient of 8 up-samplers every step
    for n, p in net.named_parameters():
        if ".up_" in n:
            assert p.grad is not None and torch.isfinite(p.grad).all(), n
    assert all(l == l and abs(l) < 1e6 for l in losses), losses
    assert min(losses[-3:]) < losses[0], losses
