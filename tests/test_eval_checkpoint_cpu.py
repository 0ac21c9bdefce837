"""CPU tests of the 'next' rows of the scope table that are pure host logic: accuracy() (train.py:524-538) and the
checkpoint format (train.py:293-307,417-428)."""
import os

import torch

import deep_vision_b200 as dv
from deep_vision_b200.evaluate import accuracy


def test_accuracy_topk_matches_manual_count():
    g = torch.Generator().manual_seed(0)
    out = torch.randn(64, 1000, generator=g)
    tgt = torch.randint(0, 1000, (64,), generator=g)
    tgt[:10] = out[:10].argmax(1)                       # 10 certain top-1 hits
    p1, p5 = accuracy(out, tgt, topk=(1, 5))
    top5 = out.topk(5, 1).indices
    assert abs(float(p1) - 100.0 * float((out.argmax(1) == tgt).float().mean())) < 1e-5
    assert abs(float(p5) - 100.0 * float((top5 == tgt[:, None]).any(1).float().mean())) < 1e-5
    assert float(p1) >= 100.0 * 10 / 64


def test_checkpoint_roundtrip_and_dataparallel_prefix(tmp_path):
    torch.manual_seed(0)
    a = dv.models.MobileNetV1()
    path = os.path.join(tmp_path, "ck.pt")
    dv.checkpoint.save_checkpoint(path, a, epoch=7, loggers={"train_loss": [1.0, 0.5]})
    ck = torch.load(path, weights_only=False)
    assert set(ck) == {"epoch", "model", "optimizer", "scheduler", "loggers"}          # the reference's keys (train.py:417-428)
    b = dv.models.MobileNetV1()
    nxt, loggers = dv.checkpoint.load_checkpoint(path, b)
    assert nxt == 8 and loggers["train_loss"] == [1.0, 0.5]
    for (k, v), (_, w) in zip(a.state_dict().items(), b.state_dict().items()):
        assert torch.equal(v, w), k
    # a file written by the reference under nn.DataParallel: every key prefixed with 'module.'
    ck["model"] = {"module." + k: v for k, v in ck["model"].items()}
    torch.save(ck, path)
    c = dv.models.MobileNetV1()
    dv.checkpoint.load_checkpoint(path, c)
    assert torch.equal(c.linear.weight, a.linear.weight)


def test_training_config_matches_reference_recipes():
    """epochs.training_config restates the hot-path entries of ResNet/pytorch/train.py:141-214 (values frozen here)."""
    from torch import optim
    tc = dv.epochs.training_config
    for name in ("resnet34", "resnet50"):
        c = tc[name]
        assert c["batch_size"] == 256 and c["optimizer"] is optim.SGD and c["total_epochs"] == 200
        assert c["optimizer_params"] == {"lr": 0.1, "momentum": 0.9, "weight_decay": 0.0001}
        assert c["scheduler"] is optim.lr_scheduler.ReduceLROnPlateau and c["scheduler_params"] == {"factor": 0.1, "mode": "max"}
    m = tc["mobilenet1"]
    assert m["batch_size"] == 128 and m["optimizer"] is optim.RMSprop and m["model_params"] == {"alpha": 1}
    assert m["optimizer_params"] == {"lr": 0.045, "alpha": 0.9, "eps": 1.0}
    assert m["scheduler"] is optim.lr_scheduler.StepLR and m["scheduler_params"] == {"step_size": 2, "gamma": 0.94}
    assert tc["resnet50"]["model"] is dv.models.ResNet50 and m["model"] is dv.models.MobileNetV1


def test_torch_schedulers_drive_the_trainer_rate():
    """StepLR and ReduceLROnPlateau(mode='max') (train.py:190,206-209,408-413) move a Trainer's rate through SchedulerTarget."""
    from torch import optim

    class FakeTrainer:
        lr = 0.045

        def set_lr(self, v):
            self.lr = float(v)

    t = FakeTrainer()
    o = dv.epochs.SchedulerTarget(t)
    s = optim.lr_scheduler.StepLR(o, step_size=2, gamma=0.94)
    seen = []
    for _ in range(5):
        o.step()
        s.step()
        seen.append(o.sync())
    want = [0.045, 0.045 * 0.94, 0.045 * 0.94, 0.045 * 0.94 ** 2, 0.045 * 0.94 ** 2]
    assert all(abs(a - b) < 1e-12 for a, b in zip(seen, want))
    assert dv.epochs.get_lr(o) == t.lr
    t2 = FakeTrainer()
    t2.lr = 0.1
    o2 = dv.epochs.SchedulerTarget(t2)
    s2 = optim.lr_scheduler.ReduceLROnPlateau(o2, factor=0.1, mode="max", patience=2)
    for acc in (50.0, 60.0, 59.0, 58.0):
        s2.step(acc)
        assert abs(o2.sync() - 0.1) < 1e-12
    s2.step(57.0)                                     # third epoch without a new best: divide by 10
    assert abs(o2.sync() - 0.01) < 1e-12
    sd = s2.state_dict()
    assert sd["best"] == 60.0 and sd["mode"] == "max"


def test_loggers_have_the_reference_layout():
    lg = dv.epochs.initialize_loggers()
    assert set(lg) == {"train_loss", "val_loss", "val_top1_acc", "val_top5_acc"}
    dv.epochs.log_metrics(lg, "val_top1_acc", 12.5, 3)
    assert lg["val_top1_acc"] == {"epochs": [3], "value": [12.5]}
