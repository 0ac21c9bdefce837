"""Freeze outputs of the UNMODIFIED reference model files as golden fixtures (tests/golden/*.npz).

Run in the build container, where /root/reference is mounted:   python oracle/make_golden.py
The GPU box has no /root/reference; tests read only the fixtures written here.

For each case the reference module is built under a fixed seed, run in float64 on CPU (train-mode BN) on a seeded
input, and back-propagated from a seeded upstream gradient.  Stored: input, upstream gradient, output, input gradient,
every parameter (the state_dict) and every parameter gradient, plus the updated BN running statistics.
Whole models are stored as seeds + logits (+ a parameter checksum that detects RNG drift between torch builds).
"""
import os
import sys

import numpy as np
import torch

REF = os.environ.get("DV_REFERENCE", "/root/reference")
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")


def _import(path, module):
    sys.path.insert(0, os.path.join(REF, path))
    try:
        if module in sys.modules:
            del sys.modules[module]
        return __import__(module)
    finally:
        sys.path.pop(0)


def block_case(name, ctor, in_shape, seed):
    torch.manual_seed(seed)
    m = ctor().double().train()
    # non-trivial affine parameters and running stats so every term of BN is exercised
    g = torch.Generator().manual_seed(seed + 1)
    for k, v in m.state_dict().items():
        if k.endswith("bn.weight") or k.endswith(("bn1.weight", "bn2.weight", "bn3.weight", "projection.1.weight")):
            v.copy_(0.5 + torch.rand(v.shape, generator=g, dtype=torch.float64))
        elif k.endswith("bn.bias") or k.endswith(("bn1.bias", "bn2.bias", "bn3.bias", "projection.1.bias")):
            v.copy_(0.2 * torch.randn(v.shape, generator=g, dtype=torch.float64))
    state0 = {k: v.detach().clone() for k, v in m.state_dict().items()}
    x = torch.randn(in_shape, generator=g, dtype=torch.float64, requires_grad=True)
    y = m(x)
    dy = torch.randn(y.shape, generator=g, dtype=torch.float64)
    y.backward(dy)
    rec = {"x": x.detach().numpy(), "dy": dy.numpy(), "y": y.detach().numpy(), "dx": x.grad.numpy()}
    for k, v in state0.items():
        rec["state/" + k] = v.numpy()
    for k, p in m.named_parameters():
        rec["grad/" + k] = p.grad.numpy()
    for k, v in m.state_dict().items():
        if "running_" in k:
            rec["after/" + k] = v.numpy()
    # float32 storage keeps the fixtures small; values are exactly representable inputs for every implementation
    rec = {k: (v.astype(np.float32) if v.dtype == np.float64 else v) for k, v in rec.items()}
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **rec)
    print("wrote", name, {k: v.shape for k, v in rec.items() if k in ("x", "y")})


def model_case(name, ctor, in_shape, seed, bn_train):
    torch.manual_seed(seed)
    m = ctor()
    m.train(bn_train)
    g = torch.Generator().manual_seed(seed + 1)
    x = torch.randn(in_shape, generator=g)
    checksum = float(sum(p.double().sum() for p in m.parameters()))
    abssum = float(sum(p.double().abs().sum() for p in m.parameters()))
    with torch.no_grad():
        y32 = m(x)
    m64 = m.double()
    # the fp32 pass above moved the running stats when bn_train; rebuild for a clean fp64 pass
    torch.manual_seed(seed)
    m64 = ctor().double()
    m64.train(bn_train)
    with torch.no_grad():
        y64 = m64(x.double())
    # torch's own bf16 (CPU autocast) on the same reference module: the torch-vs-torch floor reported beside ours
    torch.manual_seed(seed)
    mb16 = ctor()
    mb16.train(bn_train)
    with torch.no_grad(), torch.autocast("cpu", dtype=torch.bfloat16):
        yb16 = mb16(x).float()
    floor_bf16 = float((yb16.double() - y64).norm() / y64.norm())
    np.savez_compressed(os.path.join(OUT, name + ".npz"), seed=seed, in_shape=np.array(in_shape), bn_train=bn_train,
                        logits_fp32=y32.numpy(), logits_fp64=y64.numpy(), param_sum=checksum, param_abs_sum=abssum,
                        torch_version=torch.__version__, torch_bf16_autocast_rel=floor_bf16)
    print("wrote", name, y32.shape, "fp32-vs-fp64 rel", float((y32.double() - y64).norm() / y64.norm()),
          "torch bf16 autocast vs fp64 rel", floor_bf16)


def interface_case(mods):
    """Class constructor signatures, state_dict keys/shapes and parameter counts of the reference classes."""
    import inspect
    import json
    out = {"state_dict": {}, "param_count": {}, "signatures": {}}
    for mod, names in mods:
        for n in names:
            cls = getattr(mod, n)
            out["signatures"][n] = str(inspect.signature(cls.__init__))
            if n in ("ResNet34", "ResNet50", "ResNet152", "MobileNetV1"):
                m = cls()
                out["state_dict"][n] = [[k, list(v.shape)] for k, v in m.state_dict().items()]
                out["param_count"][n] = sum(p.numel() for p in m.parameters())
    json.dump(out, open(os.path.join(OUT, "reference_interface.json"), "w"))
    print("wrote reference_interface.json")


def main():
    os.makedirs(OUT, exist_ok=True)
    r50 = _import("ResNet/pytorch/models", "resnet50")
    r34 = _import("ResNet/pytorch/models", "resnet34")
    r152 = _import("ResNet/pytorch/models", "resnet152")
    mb = _import("MobileNet/pytorch/models", "mobilenet_v1")

    interface_case([(r34, ["ResNet34", "BasicBlock"]), (r50, ["ResNet50", "BottleneckBlock"]), (r152, ["ResNet152"]),
                    (mb, ["MobileNetV1", "DepthwiseSeparableConv", "DepthwiseConv", "PointwiseConv"])])
    if os.environ.get("DV_GOLDEN_INTERFACE_ONLY"):
        return
    if os.environ.get("DV_GOLDEN_MODELS_ONLY"):
        block_case_enabled = False
    else:
        block_case_enabled = True
    block_case_enabled and block_case("block_bottleneck_256_64_256", lambda: r50.BottleneckBlock(256, 64, 256), (2, 256, 8, 8), 11)
    block_case_enabled and block_case("block_bottleneck_128_64_256_s2_proj", lambda: r50.BottleneckBlock(128, 64, 256, stride=2, downsample=True), (2, 128, 8, 8), 12)
    block_case_enabled and block_case("block_bottleneck_64_64_256_s1_proj", lambda: r50.BottleneckBlock(64, 64, 256, stride=1, downsample=True), (3, 64, 7, 7), 13)
    block_case_enabled and block_case("block_basic_64_64", lambda: r34.BasicBlock(64, 64), (2, 64, 8, 8), 21)
    block_case_enabled and block_case("block_basic_64_128_s2_proj", lambda: r34.BasicBlock(64, 128, stride=2, downsample=True), (2, 64, 8, 8), 22)
    block_case_enabled and block_case("block_dwsep_64_128_s2", lambda: mb.DepthwiseSeparableConv(64, 128, dw_stride=2, pw_stride=1), (2, 64, 8, 8), 31)
    block_case_enabled and block_case("block_dwsep_32_64_s1", lambda: mb.DepthwiseSeparableConv(32, 64, dw_stride=1, pw_stride=1), (2, 32, 9, 9), 32)

    # BASELINE.json configs[0]: ResNet-34 fp32 forward, batch 4 of 3x224x224 synthetic, reference model as-is (eval + train BN)
    model_case("model_resnet34_b4_224_eval", r34.ResNet34, (4, 3, 224, 224), 0, bn_train=False)
    # train-mode fixtures: 4 x 128x128 so that the last stage still normalises over 64 samples per channel
    model_case("model_resnet34_b4_128_train", r34.ResNet34, (4, 3, 128, 128), 0, bn_train=True)
    model_case("model_resnet50_b4_128_train", r50.ResNet50, (4, 3, 128, 128), 0, bn_train=True)
    model_case("model_resnet50_b2_64_eval", r50.ResNet50, (2, 3, 64, 64), 0, bn_train=False)
    model_case("model_resnet152_b4_128_train", r152.ResNet152, (4, 3, 128, 128), 0, bn_train=True)
    model_case("model_resnet152_b2_64_eval", r152.ResNet152, (2, 3, 64, 64), 0, bn_train=False)
    model_case("model_mobilenet1_b4_128_train", mb.MobileNetV1, (4, 3, 128, 128), 0, bn_train=True)
    model_case("model_mobilenet1_b2_64_eval", mb.MobileNetV1, (2, 3, 64, 64), 0, bn_train=False)


if __name__ == "__main__":
    main()
