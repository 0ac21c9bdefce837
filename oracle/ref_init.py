"""Initial state_dicts of the reference's four models, built WITHOUT the product package and without /root/reference.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  bench.py's CPU arm (`--impl reference`, `cpu_baseline`) needs a
state_dict to drive oracle/ref_models.py; importing deep_vision_b200 for that would map the product's shared library
into the reference process.  This module lays the tensors out from the reference's constructors instead:

  * key names, shapes and registration order: ResNet/pytorch/models/resnet50.py:12-45,97-145, resnet34.py:12-46,93-128,
    resnet152.py:12-45, MobileNet/pytorch/models/mobilenet_v1.py:24-93,109-156 (checked key by key against
    tests/golden/reference_interface.json, which was frozen from the reference classes);
  * values: `_initialize_weights` of the ResNets (resnet50.py:84-93) — Kaiming-normal fan_out for every convolution,
    BN gamma 1 / beta 0, nn.Linear left at torch's default (uniform +-1/sqrt(fan_in)); MobileNetV1 has no init routine
    (torch defaults: Kaiming-uniform a=sqrt(5) convolutions).

The draws use their own generator, so the VALUES differ from a reference instance built under the same seed (the RNG
consumption order of nn.Module constructors is not reproduced); distributions, shapes and keys are the reference's.
That is what a timing baseline needs; parity tests use the frozen fixtures instead.
"""
import math

import torch

_RESNET = {
    "resnet34": ("basic", [2, 2, 2, 2], 512),
    "resnet50": ("bottleneck", [3, 4, 6, 3], 2048),
    "resnet152": ("bottleneck", [3, 8, 36, 3], 2048),
}
_MOBILENET_PLAN = [(32, 64, 1), (64, 128, 2), (128, 128, 1), (128, 256, 2), (256, 256, 1), (256, 512, 2), (512, 512, 1),
                   (512, 512, 1), (512, 512, 1), (512, 512, 1), (512, 512, 1), (512, 1024, 2), (1024, 1024, 1)]


def _bn(sd, prefix, c):
    sd[prefix + "weight"] = torch.ones(c)
    sd[prefix + "bias"] = torch.zeros(c)
    sd[prefix + "running_mean"] = torch.zeros(c)
    sd[prefix + "running_var"] = torch.ones(c)
    sd[prefix + "num_batches_tracked"] = torch.zeros((), dtype=torch.int64)


def _conv_he(sd, key, k, c, r, g):
    # nn.init.kaiming_normal_(mode="fan_out", nonlinearity="relu"): std = sqrt(2 / (K*R*S))
    sd[key] = torch.randn(k, c, r, r, generator=g) * math.sqrt(2.0 / (k * r * r))


def _conv_default(sd, key, k, c, r, g):
    # nn.Conv2d default: kaiming_uniform_(a=sqrt(5)) -> U(-b, b), b = 1/sqrt(fan_in)
    b = 1.0 / math.sqrt(c * r * r)
    sd[key] = (torch.rand(k, c, r, r, generator=g) * 2 - 1) * b


def _linear(sd, features, g):
    b = 1.0 / math.sqrt(features)
    sd["linear.weight"] = (torch.rand(1000, features, generator=g) * 2 - 1) * b
    sd["linear.bias"] = (torch.rand(1000, generator=g) * 2 - 1) * b


def init_state(name, seed=0):
    """{key: tensor} in the reference's state_dict order for "resnet34" | "resnet50" | "resnet152" | "mobilenet1"."""
    g = torch.Generator().manual_seed(seed)
    sd = {}
    if name == "mobilenet1":
        _conv_default(sd, "features.0.weight", 32, 3, 3, g)
        _bn(sd, "features.1.", 32)
        for i, (cin, cout, _stride) in enumerate(_MOBILENET_PLAN):
            p = f"features.{i + 3}."
            _conv_default(sd, p + "dw.conv.weight", cin, 1, 3, g)
            _bn(sd, p + "dw.bn.", cin)
            _conv_default(sd, p + "pw.conv.weight", cout, cin, 1, g)
            _bn(sd, p + "pw.bn.", cout)
        _linear(sd, 1024, g)
        return sd
    kind, counts, features = _RESNET[name]
    _conv_he(sd, "conv1.weight", 64, 3, 7, g)
    _bn(sd, "bn1.", 64)
    cin = 64
    for si, (stage, n) in enumerate(zip(("conv2x", "conv3x", "conv4x", "conv5x"), counts)):
        width = 64 << si
        cout = width if kind == "basic" else 4 * width
        for b in range(n):
            p = f"{stage}.{b}."
            if kind == "basic":
                _conv_he(sd, p + "conv1.weight", width, cin, 3, g)
                _bn(sd, p + "bn1.", width)
                _conv_he(sd, p + "conv2.weight", width, width, 3, g)
                _bn(sd, p + "bn2.", width)
            else:
                _conv_he(sd, p + "conv1.weight", width, cin, 1, g)
                _bn(sd, p + "bn1.", width)
                _conv_he(sd, p + "conv2.weight", width, width, 3, g)
                _bn(sd, p + "bn2.", width)
                _conv_he(sd, p + "conv3.weight", cout, width, 1, g)
                _bn(sd, p + "bn3.", cout)
            if b == 0:
                _conv_he(sd, p + "projection.0.weight", cout, cin, 1, g)
                _bn(sd, p + "projection.1.", cout)
            cin = cout
    _linear(sd, features, g)
    return sd
