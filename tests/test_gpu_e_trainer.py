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
    # Same kernels on both sides; split reductions are combined with fp32 atomics, whose order varies run to run, and a
    # random-init BN network amplifies that last-bit noise towards the first layers: gate the whole gradient tightly
    # and every tensor loosely.
    flat_t = torch.cat([p.grad.flatten() for _, p in net.named_parameters()]).cpu().numpy()
    flat_r = torch.cat([ref_grads[k].flatten() for k, _ in net.named_parameters()]).cpu().numpy()
    assert rel_l2(flat_t, flat_r) < 2e-3
    for k, p in net.named_parameters():
        assert rel_l2(p.grad.cpu().numpy(), ref_grads[k].cpu().numpy()) < 5e-2, k
    for k, p in net.named_parameters():
        g, w = ref_grads[k].double(), ref_params[k].double()
        if cfg["optimizer"] == "sgd":       # first step: buf = g + wd*w ; w -= lr*buf        (train.py:166-184)
            expect = w - cfg["lr"] * (g + cfg["weight_decay"] * w)
        else:                               # RMSprop: sq = (1-a) g^2 ; w -= lr*g/(sqrt(sq)+eps)   (train.py:199-205)
            sq = (1 - cfg["alpha"]) * g * g
            expect = w - cfg["lr"] * g / (sq.sqrt() + cfg["eps"])
        g = p.grad.detach().double()   # the update rule is checked against the gradient the trainer itself produced
        if cfg["optimizer"] == "sgd":
            expect = w - cfg["lr"] * (g + cfg["weight_decay"] * w)
        else:
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


def test_validate_matches_oracle_eval_pass():
    """validate() (train.py:488-520): eval-mode forward, mean CE, top-1/top-5 — against the functional oracle on CPU."""
    from deep_vision_b200.evaluate import accuracy, validate
    from oracle import ref_models
    dv.set_compute_dtype(torch.float32)
    torch.manual_seed(0)
    net = dv.models.MobileNetV1()
    sd = ref_models.clone_state(net.state_dict(), dtype=torch.float64)
    g = torch.Generator().manual_seed(9)
    batches = [(torch.randn(4, 3, 64, 64, generator=g), torch.randint(0, 1000, (4,), generator=g)) for _ in range(2)]
    loss, top1, top5 = validate(net.cuda(), batches)
    ref_loss = ref1 = ref5 = 0.0
    for x, y in batches:
        out = ref_models.model_forward("mobilenet1", x.double(), sd, bn_train=False)
        ref_loss += float(F.cross_entropy(out, y))
        a1, a5 = accuracy(out, y, topk=(1, 5))
        ref1 += float(a1)
        ref5 += float(a5)
    assert abs(loss - ref_loss / 2) < 1e-4 * abs(ref_loss / 2)
    assert abs(top1 - ref1 / 2) < 1e-6 and abs(top5 - ref5 / 2) < 1e-6


def test_run_epochs_schedules_checkpoints_and_resumes(tmp_path):
    """epochs.run_epochs (train.py:310-428) on a toy loader: validate at epoch 0, train/validate per epoch, StepLR moves the
    rate of the captured step, one checkpoint per epoch in the reference's format, and a resumed run continues from it."""
    import glob
    import os
    from torch import optim
    torch.manual_seed(0)
    dv.set_compute_dtype(torch.bfloat16)
    g = torch.Generator().manual_seed(5)
    batches = [{"image": torch.randn(8, 3, 64, 64, generator=g), "annotation": torch.randint(0, 1000, (8,), generator=g)} for _ in range(4)]
    cfg = dict(dv.epochs.training_config["mobilenet1"])
    cfg.update(scheduler=optim.lr_scheduler.StepLR, scheduler_params={"step_size": 1, "gamma": 0.5}, total_epochs=2)
    net, trainer, sched, loggers = dv.epochs.run_epochs(cfg, None, batches, batches[:2], model_dir=str(tmp_path) + "/", verbose=False)
    assert abs(trainer.lr - 0.045 * 0.25) < 1e-12
    assert loggers["val_loss"]["epochs"] == [0, 1, 2] and len(loggers["val_top5_acc"]["value"]) == 3
    files = sorted(glob.glob(os.path.join(tmp_path, "mobilenet1-*-epoch-*.pt")))
    assert len(files) == 2
    ck = torch.load(files[-1], weights_only=False)
    assert set(ck) == {"epoch", "model", "optimizer", "scheduler", "loggers"} and ck["epoch"] == 2
    assert set(ck["model"]) == set(net.state_dict())
    assert all(np.isfinite(v) for v in loggers["val_loss"]["value"])
    # resume: epoch 3 only, starting from the saved rate and optimiser state
    cfg["total_epochs"] = 3
    net2, trainer2, sched2, loggers2 = dv.epochs.run_epochs(cfg, files[-1], batches, batches[:2], model_dir=str(tmp_path) + "/r/",
                                                             verbose=False)
    assert loggers2["val_loss"]["epochs"] == [0, 1, 2, 0, 3]
    assert abs(trainer2.lr - 0.045 * 0.125) < 1e-12
    assert trainer2.steps_done > 0 and float(trainer2.flat_s.abs().sum()) > 0


def test_inference_after_graph_training_sees_the_new_statistics():
    """Training steps replayed from the CUDA graph rewrite weights and BN running statistics through raw pointers; the folded
    inference path caches scale/shift per BatchNorm2d and must not serve the values from before the steps."""
    torch.manual_seed(0)
    dv.set_compute_dtype(torch.bfloat16)
    net = dv.models.ResNet34().cuda()
    cfg = dict(TRAINING_CONFIG["resnet34"])
    cfg.pop("batch_size")
    tr = Trainer(net, use_cuda_graph=True, **cfg)
    x, y = _batch(8, 64)

    def eval_logits(fold):
        dv.set_fold_bn(fold)
        try:
            net.eval()
            with torch.no_grad():
                return net(x).float().cpu().numpy()
        finally:
            dv.set_fold_bn(True)

    before = eval_logits(True)                 # fills the per-BN caches
    for _ in range(4):
        tr.step(x, y)
    torch.cuda.synchronize()
    folded, plain = eval_logits(True), eval_logits(False)
    assert rel_l2(folded, plain) < 3e-2        # same statistics on both paths
    assert rel_l2(folded, before) > 1e-2       # and they are the new ones
