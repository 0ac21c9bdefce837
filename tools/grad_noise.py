#!/usr/bin/env python
"""Run-to-run difference of the gradient, layer by layer (last layer first): the same weights and the same batch, the
gradient computed twice, in three modes.

  default                               fused BN sums (fp32 atomics in the conv epilogue) + atomics / TMA reduce-add in backward
  forward fixed-order, backward default the BatchNorm statistics come from the fixed-order pass, so the FORWARD is bit-reproducible:
                                        the loss and every dy are identical run to run, and the only thing left that may differ is
                                        what the default backward adds with atomics — the split wgrad tiles (TMA reduce-add into dW,
                                        on the side stream).  Those touch dW only, never a dx, so nothing is amplified: conv weights
                                        ~1e-7, everything else exactly 0.  A data race (a gradient slice read or cleared while a
                                        side-stream kernel still adds to it) would show here as a large number at one layer.
  fixed order                           must give 0 everywhere.

In the default mode the difference enters in the FORWARD: a last-bit change of a fused mean moves a few activations across a
bf16 rounding boundary (2^-9 relative), and train-mode BN at random init amplifies from there (second table).

    python tools/grad_noise.py [--model resnet50] [--batch 8] [--size 96] [--out profiles/r02_grad_noise_per_layer.md]
"""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import deep_vision_b200 as dv  # noqa: E402
from deep_vision_b200.train import Trainer  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--model", default="resnet50")
ap.add_argument("--batch", type=int, default=8)
ap.add_argument("--size", type=int, default=96)
ap.add_argument("--out", default="gpurun_out/grad_noise_per_layer.md")
args = ap.parse_args()
ctor = {"resnet50": "ResNet50", "resnet34": "ResNet34", "resnet152": "ResNet152", "mobilenet1": "MobileNetV1"}[args.model]
torch.cuda.set_device(0)
dv.set_compute_dtype(torch.bfloat16)
g = torch.Generator().manual_seed(100)
x = torch.randn(args.batch, 3, args.size, args.size, generator=g).cuda()
y = torch.randint(0, 1000, (args.batch,), generator=g).cuda()


def autograd_grads(net):
    net.zero_grad(set_to_none=True)
    torch.nn.functional.cross_entropy(net(x), y).backward()
    torch.cuda.synchronize()
    return {k: p.grad.detach().clone() for k, p in net.named_parameters()}


def trainer_grads(tr, net, keep):
    """One eager Trainer step (direct gradient writes, side-stream wgrad when enabled) from the saved state; returns flat_g views."""
    tr.flat_p.copy_(keep[0])
    tr.flat_s.copy_(keep[1])
    for b, s in zip(net.buffers(), keep[2:]):
        b.detach().copy_(s)
    tr.refresh_from_parameters()
    tr.steps_done = 0
    tr.step(x, y)
    torch.cuda.synchronize()
    return {k: p.grad.detach().clone() for k, p in net.named_parameters()}


cols = {}
MODES = (("default (fused BN sums + atomics)", False, True), ("forward fixed-order, backward default", False, False),
         ("fixed order", True, False))
for label, det, fuse in MODES:
    dv.set_deterministic(det)
    dv.set_fuse_bn_stats(fuse)
    torch.manual_seed(0)
    net = getattr(dv.models, ctor)().cuda().train()
    a, b = autograd_grads(net), autograd_grads(net)
    cols["autograd: " + label] = {k: float((a[k] - b[k]).norm() / (b[k].norm() + 1e-30)) for k in a}
    del net
    torch.manual_seed(0)
    net = getattr(dv.models, ctor)().cuda().train()
    tr = Trainer(net, optimizer="sgd", lr=0.1, momentum=0.9, weight_decay=1e-4)
    keep = [tr.flat_p.clone(), tr.flat_s.clone()] + [bf.detach().clone() for bf in net.buffers()]
    a, b = trainer_grads(tr, net, keep), trainer_grads(tr, net, keep)
    cols["Trainer: " + label] = {k: float((a[k] - b[k]).norm() / (b[k].norm() + 1e-30)) for k in a}
    names = [k for k, _ in net.named_parameters()]
    del tr, net
dv.set_fuse_bn_stats(True)
dv.set_deterministic(False)

# forward only, default mode: where the run-to-run difference enters
torch.manual_seed(0)
net = getattr(dv.models, ctor)().cuda().train()
outs = [[], []]
blocks = [(n, m) for n, m in net.named_modules() if n.count(".") == 1 and n.split(".")[0] in ("conv2x", "conv3x", "conv4x", "conv5x", "features")]
for run in range(2):
    hs = [m.register_forward_hook(lambda _m, _i, o, r=run: outs[r].append(o.detach().float().clone())) for _, m in blocks]
    with torch.no_grad():
        logits = net(x)
    outs[run].append(logits.float().clone())
    for h in hs:
        h.remove()
fwd = [(n, float((p - q).norm() / (q.norm() + 1e-30))) for (n, _), p, q in zip(blocks + [("logits", None)], outs[0], outs[1])]
del net

labels = list(cols)
lines = [f"# Run-to-run gradient difference per parameter, last layer first ({args.model}, batch {args.batch}, {args.size}x{args.size}, bf16)", "",
         "Relative L2 difference between two computations of the same gradient (same weights, same batch).", "",
         "| parameter | " + " | ".join(labels) + " |", "|---|" + "---|" * len(labels)]
for k in reversed(names):
    lines.append(f"| `{k}` | " + " | ".join(f"{cols[l][k]:.2e}" for l in labels) + " |")
summary = []
for l in labels:
    v = [cols[l][k] for k in names]
    conv = [cols[l][k] for k in names if k.endswith("weight") and ("conv" in k or "projection.0" in k or k.startswith("features.0") or k == "linear.weight")]
    other = [cols[l][k] for k in names if cols[l][k] not in conv]
    summary.append(f"* {l}: max over all parameters {max(v):.2e}; max over dense filters {max(conv):.2e}; "
                   f"parameters that differ at all: {sum(1 for q in v if q > 0)} of {len(v)}")
lines += ["", "## Summary", ""] + summary
lines += ["", "## Forward pass, default mode: block outputs of two runs (train-mode BN, no_grad)", "", "| block | relative L2 difference |", "|---|---|"]
lines += [f"| `{n}` | {v:.2e} |" for n, v in fwd]
os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
open(args.out, "w").write("\n".join(lines) + "\n")
print("\n".join(lines[:10] + ["..."] + summary + [""] + [f"{n}: {v:.2e}" for n, v in fwd]))
