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
    """bf16, tensor-core path: the captured step follows the eager step's loss trajectory.  The warm-up that precedes the
    capture is not training (state is put back), so replay i is training step i, exactly as in the eager run."""
    x, y = _batch(n=8, size=64)
    losses = {}
    for graph in (False, True):
        torch.manual_seed(0)
        net = dv.models.ResNet34().cuda()
        tr = Trainer(net, use_cuda_graph=graph, optimizer="sgd", lr=0.01, momentum=0.9, weight_decay=1e-4)
        ls = []
        for _ in range(6):
            ls.append(float(tr.step(x, y)))
        torch.cuda.synchronize()
        losses[graph] = ls
        assert tr.steps_done == 6
        bn = [m for m in net.modules() if isinstance(m, torch.nn.BatchNorm2d)]
        assert all(int(m.num_batches_tracked) == 6 for m in bn)
    assert all(np.isfinite(losses[True])) and all(np.isfinite(losses[False]))
    assert abs(losses[True][0] - losses[False][0]) < 2e-2 * losses[False][0]   # same first batch from the same initial weights
    # bf16 training of a random-init network is chaotic; the trajectories agree to ~10 % per step
    assert rel_l2(np.array(losses[True]), np.array(losses[False])) < 0.2
    assert losses[False][-1] < 0.2 * losses[False][0]      # it learns the fixed batch


def test_first_captured_step_is_one_update_not_three():
    """fp32, fixed-order reductions: the first step through the CUDA graph leaves exactly the parameters, momentum buffers,
    running statistics and counters that one eager step leaves (the two warm-up passes before the capture are undone)."""
    dv.set_compute_dtype(torch.float32)
    dv.set_deterministic(True)
    try:
        x, y = _batch(n=4, size=64)
        out = {}
        for graph in (False, True):
            torch.manual_seed(0)
            net = dv.models.ResNet34().cuda()
            tr = Trainer(net, use_cuda_graph=graph, optimizer="sgd", lr=0.1, momentum=0.9, weight_decay=1e-4)
            tr.step(x, y)
            torch.cuda.synchronize()
            out[graph] = (tr.flat_p.clone(), tr.flat_s.clone(), net.bn1.running_var.clone(), int(net.bn1.num_batches_tracked), tr.steps_done)
        assert out[True][3] == out[False][3] == 1 and out[True][4] == out[False][4] == 1
        for a, b in zip(out[True][:3], out[False][:3]):
            assert rel_l2(a.cpu().numpy(), b.cpu().numpy()) < 1e-6
    finally:
        dv.set_deterministic(False)


def test_ragged_last_batch_runs_eagerly_beside_the_captured_step():
    """The reference's loader has no drop_last (train.py:229-234): the last batch of an epoch is smaller.  Under
    use_cuda_graph=True it runs eagerly through the same kernels and state; the next full batch replays the graph."""
    x, y = _batch(n=8, size=64)
    torch.manual_seed(0)
    net = dv.models.MobileNetV1().cuda()
    tr = Trainer(net, use_cuda_graph=True, optimizer="rmsprop", lr=0.01, alpha=0.9, eps=1.0, weight_decay=0.0)
    l0 = float(tr.step(x, y))
    before = tr.flat_p.clone()
    l1 = float(tr.step(x[:5], y[:5]))          # ragged
    assert np.isfinite(l1) and tr.steps_done == 2
    assert float((tr.flat_p - before).abs().max()) > 0          # it trained
    l2 = float(tr.step(x, y))
    assert np.isfinite(l2) and tr.steps_done == 3 and np.isfinite(l0)
    assert all(int(m.num_batches_tracked) == 3 for m in net.modules() if isinstance(m, torch.nn.BatchNorm2d))


ORACLE_STEP_CASES = [
    ("resnet34", dv.models.ResNet34, dict(optimizer="sgd", lr=0.1, momentum=0.9, weight_decay=0.05)),
    ("mobilenet1", dv.models.MobileNetV1, dict(optimizer="rmsprop", lr=0.045, alpha=0.9, eps=1.0, weight_decay=0.05)),
]


@pytest.mark.parametrize("name,ctor,kw", ORACLE_STEP_CASES)
def test_trainer_steps_match_oracle_train_step(name, ctor, kw):
    """Trainer.step against oracle.ref_models.train_step (the restatement of train.py:452-467 with SGD :166-184,362-365 /
    RMSprop :199-205), four consecutive steps in fp32 with fixed-order reductions.  The oracle advances in fp64; before every
    step the trainer is put into the oracle's state (parameters, momentum / squared-average buffers, BN running statistics and
    counters), because a free-running fp32 copy of a random-init BN network leaves an fp64 one chaotically (torch's own fp32
    kernels: 1.4e-2 after four steps).  That pins, step by step: the first-step momentum rule (buf = g), later steps with a
    non-zero buffer, weight decay on EVERY parameter incl. BN gamma/beta and the classifier bias (weight_decay is 0.05 here so
    that the term is a visible share of the update), RMSprop's eps = 1.0 outside the square root, and the running-statistics
    update.  Bars: error relative to the size of the update <= 2e-2 over all parameters, <= 5e-2 per tensor (torch fp32 vs fp64
    under the same protocol: <= 4.5e-3 / 8.4e-3); running statistics <= 1e-4; counters equal."""
    from deep_vision_b200.train import flat_view
    from oracle import ref_models
    dv.set_compute_dtype(torch.float32)
    dv.set_deterministic(True)
    try:
        torch.manual_seed(0)
        net = ctor().cuda().train()
        tr = Trainer(net, **kw)
        sd = ref_models.clone_state({k: v.cpu() for k, v in net.state_dict().items()}, dtype=torch.float64, requires_grad=True)
        opt_state = {}
        g = torch.Generator().manual_seed(5)
        x = torch.randn(8, 3, 64, 64, generator=g)
        y = torch.randint(0, 1000, (8,), generator=g)
        names = [k for k, _ in net.named_parameters()]
        for step in range(4):
            # trainer <- oracle state
            net.load_state_dict({k: v.detach().float() for k, v in sd.items()})
            for (k, p), off in zip(net.named_parameters(), tr.flat.offsets):
                buf = opt_state.get(k)
                view = flat_view(tr.flat_s, off, p.shape)
                view.zero_() if buf is None else view.copy_(buf.float())
            tr.steps_done = step
            tr.refresh_from_parameters()
            before = {k: sd[k].detach().clone() for k in names}
            loss_ref = ref_models.train_step(name, sd, x.double(), y, opt_state, **kw)
            loss = float(tr.step(x.cuda(), y.cuda()))
            torch.cuda.synchronize()
            assert abs(loss - loss_ref) < 1e-4 * abs(loss_ref), (step, loss, loss_ref)
            got = {k: p.detach().double().cpu() for k, p in net.named_parameters()}
            num = torch.cat([(got[k] - sd[k].detach()).flatten() for k in names]).norm()
            den = torch.cat([(sd[k].detach() - before[k]).flatten() for k in names]).norm()
            assert float(num / den) < 2e-2, (step, float(num / den))
            for k in names:
                upd = float((sd[k].detach() - before[k]).norm())
                assert float((got[k] - sd[k].detach()).norm()) <= 5e-2 * upd + 1e-12, (step, k)
            # the decay term really is part of what was compared: for BN gammas it is a visible share of the update
            k_bn = next(k for k in names if k.endswith(("bn1.weight", "features.1.weight")))
            assert kw["weight_decay"] * float(before[k_bn].norm()) * kw["lr"] > 1e-2 * float((sd[k_bn].detach() - before[k_bn]).norm())
            state = net.state_dict()
            for k, v in sd.items():
                if k.endswith(("running_mean", "running_var")):
                    assert rel_l2(state[k].cpu().numpy(), v.detach().numpy()) < 1e-4, (step, k)
                elif k.endswith("num_batches_tracked"):
                    assert int(state[k]) == int(v), (step, k)
            # momentum / squared-average buffers as well
            for (k, p), off in zip(net.named_parameters(), tr.flat.offsets):
                ref_buf = opt_state[k]
                mine = flat_view(tr.flat_s, off, p.shape).double().cpu()
                assert float((mine - ref_buf).norm()) <= 5e-2 * float(ref_buf.norm()) + 1e-12, (step, k)
    finally:
        dv.set_deterministic(False)


def test_reference_optimizer_state_loads_into_krsc_flat_state_and_back(tmp_path):
    """A checkpoint written the reference's way (train.py:417-428: torch.optim.SGD.state_dict(), OIHW momentum buffers keyed by
    parameter index) loads into the trainer's flat KRSC state without scrambling the 3x3 / 7x7 filters, and the trainer's own
    checkpoint is accepted by torch.optim.SGD.load_state_dict (train.py:295) with the same buffers."""
    from deep_vision_b200.checkpoint import load_checkpoint, save_checkpoint
    from deep_vision_b200.train import flat_view
    dv.set_compute_dtype(torch.float32)
    torch.manual_seed(0)
    net = dv.models.ResNet34().cuda()
    tr = Trainer(net, optimizer="sgd", lr=0.1, momentum=0.9, weight_decay=1e-4)
    ref_params = [torch.nn.Parameter(p.detach().cpu().contiguous().clone()) for p in net.parameters()]
    opt = torch.optim.SGD(ref_params, lr=0.05, momentum=0.9, weight_decay=1e-4)
    gen = torch.Generator().manual_seed(3)
    for p in ref_params:
        p.grad = torch.randn(p.shape, generator=gen)
    opt.step()                                    # every parameter now has a momentum_buffer
    path = str(tmp_path / "ref.pt")
    torch.save({"epoch": 3, "model": {"module." + k: v.cpu() for k, v in net.state_dict().items()}, "optimizer": opt.state_dict(),
                "scheduler": None, "loggers": {}}, path)
    next_epoch, _ = load_checkpoint(path, net, tr)
    assert next_epoch == 4 and abs(tr.lr - 0.05) < 1e-12 and tr.steps_done >= 1
    for idx, ((k, p), off) in enumerate(zip(net.named_parameters(), tr.flat.offsets)):
        want = opt.state[ref_params[idx]]["momentum_buffer"]
        got = flat_view(tr.flat_s, off, p.shape).cpu()
        assert torch.equal(got, want), k              # incl. conv1.weight (7x7) and every 3x3
    # and back: our file -> torch.optim
    path2 = str(tmp_path / "ours.pt")
    save_checkpoint(path2, net, 4, tr)
    ck = torch.load(path2, weights_only=True)
    opt2 = torch.optim.SGD([torch.nn.Parameter(p.detach().cpu().contiguous().clone()) for p in net.parameters()], lr=0.1, momentum=0.9)
    opt2.load_state_dict(ck["optimizer"])
    assert abs(opt2.param_groups[0]["lr"] - 0.05) < 1e-12
    for idx, p2 in enumerate(opt2.param_groups[0]["params"]):
        assert torch.equal(opt2.state[p2]["momentum_buffer"], opt.state[ref_params[idx]]["momentum_buffer"])


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
