#!/usr/bin/env python
"""The reference's OWN stack on the same B200 (BASELINE.md §4 "Secondary", informational): the reference model step
(oracle/ref_models.py = its forward methods, ResNet/pytorch/train.py:431-485) through stock torch on `cuda` — cuDNN
convolutions and batch norm, channels_last, bf16 autocast, torch.optim — at the bench's batch sizes, plus per-layer cuDNN
times for the 19 ResNet-50 convolution shapes next to this repo's kernels (tools/layer_bench.py).

This is NOT the reference arm of bench.py (that one is the CPU path) and none of this repo's kernels run here; it only says
how far the hand-written kernels are from what NVIDIA's libraries reach on the same box.

    python tools/stock_torch_bench.py [--models resnet50,resnet34,resnet152,mobilenet1] [--layers] [--out gpurun_out/stock_torch]
"""
import argparse
import json
import os
import sys

import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_init, ref_models  # noqa: E402

BATCH = {"resnet50": 256, "resnet34": 256, "resnet152": 128, "mobilenet1": 256}
OPT = {"resnet50": "sgd", "resnet34": "sgd", "resnet152": "sgd", "mobilenet1": "rmsprop"}
CL = torch.channels_last


def step_bench(name, steps=20, warmup=5, graph=False):
    dev = torch.device("cuda")
    sd = {}
    for k, v in ref_init.init_state(name, seed=0).items():
        t = v.to(dev)
        if t.dim() == 4:
            t = t.contiguous(memory_format=CL)
        if t.is_floating_point() and not k.endswith(("running_mean", "running_var")):
            t.requires_grad_(True)
        sd[k] = t
    params = [v for v in sd.values() if v.requires_grad]
    opt = (torch.optim.SGD(params, lr=0.1, momentum=0.9, weight_decay=1e-4) if OPT[name] == "sgd"
           else torch.optim.RMSprop(params, lr=0.045, alpha=0.9, eps=1.0, capturable=graph))
    B = BATCH[name]
    x = torch.randn(B, 3, 224, 224, device=dev).contiguous(memory_format=CL)
    y = torch.randint(0, 1000, (B,), device=dev)

    def step():
        with torch.autocast("cuda", dtype=torch.bfloat16):
            logits = ref_models.model_forward(name, x, sd, bn_train=True)
        loss = F.cross_entropy(logits.float(), y)
        opt.zero_grad(set_to_none=False)
        loss.backward()
        opt.step()
        return loss

    run = step
    if graph:
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            for _ in range(3):
                step()
        torch.cuda.current_stream().wait_stream(s)
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr):
            step()
        run = gr.replay
    for _ in range(warmup):
        run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        run()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    return {"model": name, "batch": B, "cuda_graph": graph, "ms_per_step": round(ms, 3), "images_per_s": round(B / ms * 1e3, 1)}


def layer_bench():
    from tools.layer_bench import R50, timed
    rows = []
    for (Cc, K, R, stride, H, count) in R50:
        pad = R // 2
        x = torch.randn(256, Cc, H, H, device="cuda", dtype=torch.bfloat16).contiguous(memory_format=CL)
        w = (torch.randn(K, Cc, R, R, device="cuda") * 0.05).to(torch.bfloat16).contiguous(memory_format=CL)
        yv = F.conv2d(x, w, None, stride, pad)
        dy = torch.randn_like(yv)
        flops = 2.0 * yv.numel() * Cc * R * R
        conv_bwd = torch.ops.aten.convolution_backward
        fns = {"fprop": lambda: F.conv2d(x, w, None, stride, pad),
               "dgrad": lambda: conv_bwd(dy, x, w, None, [stride, stride], [pad, pad], [1, 1], False, [0, 0], 1, [True, False, False]),
               "wgrad": lambda: conv_bwd(dy, x, w, None, [stride, stride], [pad, pad], [1, 1], False, [0, 0], 1, [False, True, False])}
        for name, fn in fns.items():
            ms = timed(fn, iters=5)
            rows.append(dict(layer=f"{Cc}->{K} {R}x{R}/{stride} @{H}", passname=name, count=count, ms=round(ms, 4), tflops=round(flops / ms / 1e9, 1),
                             gflop=round(flops / 1e9, 1)))
        del x, w, yv, dy
    return rows


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--models", default="resnet50,resnet34,resnet152,mobilenet1")
    ap.add_argument("--layers", action="store_true")
    ap.add_argument("--out", default="gpurun_out/stock_torch")
    args = ap.parse_args()
    torch.backends.cudnn.benchmark = True
    out = {"torch": torch.__version__, "cudnn": torch.backends.cudnn.version(), "device": torch.cuda.get_device_name(0), "steps": []}
    for name in [m for m in args.models.split(",") if m]:
        for graph in (False, True):
            try:
                out["steps"].append(step_bench(name, graph=graph))
            except Exception as e:
                out["steps"].append({"model": name, "cuda_graph": graph, "error": str(e)[:200]})
            print(out["steps"][-1], flush=True)
            torch.cuda.empty_cache()
    md = [f"# Stock torch {out['torch']} / cuDNN {out['cudnn']} on {out['device']}: the reference model step, channels_last, bf16 autocast", "",
          "| model | batch | CUDA graph | ms/step | images/s |", "|---|---|---|---|---|"]
    md += [f"| {r['model']} | {r.get('batch', '')} | {r['cuda_graph']} | {r.get('ms_per_step', r.get('error', ''))} | {r.get('images_per_s', '')} |" for r in out["steps"]]
    if args.layers:
        rows = layer_bench()
        out["layers"] = rows
        tot = sum(r["ms"] * r["count"] for r in rows)
        tflop = sum(r["gflop"] * r["count"] for r in rows) / 1e3
        md += ["", "## cuDNN per layer, batch 256, bf16 channels_last (L2 flushed before every launch, cudnn.benchmark on)", "",
               "| layer | pass | ms | TFLOP/s | x count |", "|---|---|---|---|---|"]
        md += [f"| {r['layer']} | {r['passname']} | {r['ms']} | {r['tflops']} | {r['count']} |" for r in rows]
        md += ["", f"Sum over the 52 non-stem convs x 3 passes: **{tot:.2f} ms** = {tflop / tot * 1e3:.0f} TFLOP/s time-weighted."]
    os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
    open(args.out + ".md", "w").write("\n".join(md) + "\n")
    json.dump(out, open(args.out + ".json", "w"))
    print("\n".join(md))


if __name__ == "__main__":
    main()
