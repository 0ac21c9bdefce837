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
