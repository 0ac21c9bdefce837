"""CPU tests of the upper face of the boundary: class names, constructor signatures, state_dict keys/shapes and
initial weights are those of the reference model files (fixtures frozen by oracle/make_golden.py)."""
import inspect
import json
import os

import pytest
import torch
import torch.nn as nn

import deep_vision_b200 as dv
from conftest import GOLDEN


@pytest.fixture(scope="module")
def iface():
    return json.load(open(os.path.join(GOLDEN, "reference_interface.json")))


@pytest.mark.parametrize("name", ["ResNet34", "ResNet50", "ResNet152", "MobileNetV1"])
def test_state_dict_matches_reference(iface, name):
    m = getattr(dv.models, name)()
    sd = m.state_dict()
    ref = iface["state_dict"][name]
    assert list(sd.keys()) == [k for k, _ in ref]
    for (k, shape) in ref:
        assert list(sd[k].shape) == shape, k
    assert sum(p.numel() for p in m.parameters()) == iface["param_count"][name]


@pytest.mark.parametrize("name", ["ResNet34", "ResNet50", "ResNet152", "MobileNetV1", "BasicBlock", "BottleneckBlock",
                                  "DepthwiseSeparableConv", "DepthwiseConv", "PointwiseConv"])
def test_constructor_signatures(iface, name):
    sig = str(inspect.signature(getattr(dv.models, name).__init__))
    assert sig == iface["signatures"][name]


def test_modules_are_torch_subclasses():
    m = dv.models.ResNet34()
    convs = [x for x in m.modules() if isinstance(x, nn.Conv2d)]
    bns = [x for x in m.modules() if isinstance(x, nn.BatchNorm2d)]
    assert len(convs) == 21 and len(bns) == 21
    assert all(isinstance(x, dv.nn.Conv2d) for x in convs)
    # reference checkpoints load, and DataParallel-style 'module.' prefixes are the caller's business (README.md:87-96)
    ref_like = {k: torch.zeros_like(v) for k, v in m.state_dict().items()}
    m.load_state_dict(ref_like)
    assert float(m.conv1.weight.abs().sum()) == 0.0


def test_train_eval_toggle_and_bn_defaults():
    m = dv.models.MobileNetV1()
    bn = m.features[1]
    assert bn.eps == 1e-5 and bn.momentum == 0.1 and bn.num_features == 32
    m.eval()
    assert not bn.training
    m.train()
    assert bn.training
