"""GPU tests of the per-rank training step (deep_vision_b200.train.Trainer) — the restatement of train()
(ResNet/pytorch/train.py:431-485) with flat buffers, direct gradient writes, fused optimiser and CUDA-graph capture."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

import deep_vision_b200 as dv
from deep_vision_b200.train import TRAINING_CONFIG, Trainer
from helpers import rel_l2

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def _restore():
    yield
    dv.set_compute_dtype(torch.bfloat16)


def _batch(n=4, size=64, seed=5):
    g = torch.Generator().manual_seed(seed)
    return torch.randn(n, 3, size, size, generator=g).cuda(), torch.randint(0, 1000, (n,), generator=g).cuda()


@pytest.mark.parametrize("name,ctor", [("resnet34", dv.models.ResNet34), ("mobilenet1", dv.models.MobileNetV1)])
def test_trainer_gradients_equal_plain_autograd_and_update_matches_optimizer(name, ctor):
    """The trainer's direct gradient writes give the gradients plain autograd accumulates, and its fused optimiser
    kernel applies the reference's update rule (SGD momentum+wd / RMSprop) to them."""
    dv.set_compute_dtype(torch.float32)
    x, y = _batch()
    torch.manual_seed(0)
    plain = ctor().cuda().train()
    F.cross_entropy(plain(x), y).backward()
    ref_grads = {k: p.grad.detach().clone() for k, p in plain.named_parameters()}
    ref_params = {k: p.detach().clone() for k, p in plain.named_parameters()}

    torch.manual_seed(0)
    net = ctor().cuda().train()
    cfg = dict(TRAINING_CONFIG[name])
    cfg.pop("batch_size")
    tr = Trainer(net, **cfg)
    loss = tr.step(x, y)
    torch.cuda.synchronize()
    assert torch.isfinite(loss)
    for k, p in net.named_parameters():
        assert rel_l2(p.grad.cpu().numpy(), ref_grads[k].cpu().numpy()) < 1e-4, k   # same kernels; fp32 atomics aside
    for k, p in net.named_parameters():
        g, w = ref_grads[k].double(), ref_params[k].double()
        if cfg["optimizer"] == "sgd":       # first step: buf = g + wd*w ; w -= lr*buf        (train.py:166-184)
            expect = w - cfg["lr"] * (g + cfg["weight_decay"] * w)
        else:                               # RMSprop: sq = (1-a) g^2 ; w -= lr*g/(sqrt(sq)+eps)   (train.py:199-205)
            sq = (1 - cfg["alpha"]) * g * g
            expect = w - cfg["lr"] * g / (sq.sqrt() + cfg["eps"])
        assert rel_l2(p.detach().double().cpu().numpy(), expect.cpu().numpy()) < 1e-5, k
    # BN counters advanced once, through the shared vector
    bn = [m for m in net.modules() if isinstance(m, torch.nn.BatchNorm2d)]
    assert all(int(m.num_batches_tracked) == 1 for m in bn)


def test_trainer_cuda_graph_replays_the_eager_step():
    """bf16, tensor-core path: the captured step follows the eager step's loss trajectory.  Capture runs two eager
    warm-up steps (real training steps) first, so replay i is training step i+2."""
    x, y = _batch(n=8, size=64)
    losses = {}
    for graph, nsteps in ((False, 7), (True, 5)):
        torch.manual_seed(0)
        net = dv.models.ResNet34().cuda()
        tr = Trainer(net, use_cuda_graph=graph, optimizer="sgd", lr=0.01, momentum=0.9, weight_decay=1e-4)
        ls = []
        for _ in range(nsteps):
            ls.append(float(tr.step(x, y)))
        torch.cuda.synchronize()
        losses[graph] = ls
        assert tr.steps_done == 7
    assert all(np.isfinite(losses[True])) and all(np.isfinite(losses[False]))
    # bf16 training of a random-init network is chaotic; the trajectories agree to ~10 % per step
    assert rel_l2(np.array(losses[True]), np.array(losses[False][2:])) < 0.2
    assert losses[False][-1] < 0.2 * losses[False][0]      # it learns the fixed batch
