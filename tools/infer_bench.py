#!/usr/bin/env python
"""Inference throughput (net.eval(), no_grad, bf16): folded conv+BN epilogue vs separate eval-mode BN pass, launched eagerly
from Python and replayed from a CUDA graph (deep_vision_b200.inference.CapturedForward), which takes the host's per-launch cost
out of the picture.  Device-resident synthetic batch, CUDA events, L2 not flushed (inputs 38 MB + activations >> L2)."""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import deep_vision_b200 as dv  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--model", default="resnet50")
ap.add_argument("--batch", type=int, default=256)
ap.add_argument("--iters", type=int, default=20)
args = ap.parse_args()
ctor = {"resnet50": "ResNet50", "resnet34": "ResNet34", "resnet152": "ResNet152", "mobilenet_v1": "MobileNetV1"}[args.model]
torch.manual_seed(0)
dv.set_compute_dtype(torch.bfloat16)
net = getattr(dv.models, ctor)().cuda().eval()
x = torch.randn(args.batch, 3, 224, 224, device="cuda")
out = {"model": args.model, "batch": args.batch}
for fold in (True, False):
    dv.set_fold_bn(fold)
    with torch.no_grad():
        for _ in range(3):
            net(x)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.iters):
            net(x)
        e1.record()
        e1.synchronize()
    ms = e0.elapsed_time(e1) / args.iters
    rec = {"ms_per_batch": round(ms, 3), "images_per_s": round(args.batch / ms * 1e3, 1)}
    fwd = dv.inference.CapturedForward(net, x)
    for _ in range(3):
        fwd(x)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.iters):
        fwd(x)
    e1.record()
    e1.synchronize()
    ms = e0.elapsed_time(e1) / args.iters
    rec["graph_ms_per_batch"], rec["graph_images_per_s"] = round(ms, 3), round(args.batch / ms * 1e3, 1)
    out["folded" if fold else "unfused"] = rec
dv.set_fold_bn(True)
print(json.dumps(out))
