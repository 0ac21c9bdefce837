#!/usr/bin/env python
"""In-step kernel timeline of the CUDA-graph training step (CUPTI through torch.profiler, no ncu): per-kernel time with
the caches as they are in the real step, plus the idle time between kernels.  Writes a markdown table.
A number read from this trace is an attribution aid, not a bench value."""
import argparse
import collections
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import deep_vision_b200 as dv  # noqa: E402
from deep_vision_b200.train import TRAINING_CONFIG, Trainer  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--model", default="resnet50")
ap.add_argument("--batch", type=int, default=256)
ap.add_argument("--out", default="gpurun_out/step_trace.md")
ap.add_argument("--launches", default="", help="also write the launch sequence of the last step: start (us), duration (us), kernel")
args = ap.parse_args()
import torch.distributed as dist  # noqa: E402
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:      # under torchrun: every rank trains, rank 0 writes the timeline (NCCL kernels included)
    os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
dv.set_compute_dtype(torch.bfloat16)
torch.manual_seed(0)
ctor = {"resnet50": "ResNet50", "resnet34": "ResNet34", "resnet152": "ResNet152", "mobilenet1": "MobileNetV1"}[args.model]
net = getattr(dv.models, ctor)().cuda()
cfg = dict(TRAINING_CONFIG[args.model])
cfg.pop("batch_size")
trainer = Trainer(net, use_cuda_graph=True, **cfg)
x = torch.randn(args.batch, 3, 224, 224, device="cuda")
y = torch.randint(0, 1000, (args.batch,), device="cuda")
for _ in range(5):
    trainer.step(x, y)
torch.cuda.synchronize()
from torch.profiler import ProfilerActivity, profile  # noqa: E402

STEPS = 3
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(STEPS):
        trainer.step(x, y)
    torch.cuda.synchronize()
if world > 1:
    dist.barrier()
    torch.cuda.synchronize()
    if rank != 0:
        os._exit(0)
ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA and e.time_range.end > e.time_range.start]
ev.sort(key=lambda e: e.time_range.start)
span = (ev[-1].time_range.end - ev[0].time_range.start) / STEPS
busy = sum(e.time_range.end - e.time_range.start for e in ev) / STEPS
gaps = []
for a, b in zip(ev, ev[1:]):
    gaps.append(max(0.0, b.time_range.start - a.time_range.end))
agg = collections.defaultdict(lambda: [0, 0.0])
for e in ev:
    k = e.name.replace("(anonymous namespace)::", "").replace("void ", "").split("(")[0][:90]
    agg[k][0] += 1
    agg[k][1] += e.time_range.end - e.time_range.start
rows = sorted(agg.items(), key=lambda kv: -kv[1][1])
os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
with open(args.out, "w") as f:
    f.write(f"# In-step kernel timeline, {args.model} batch {args.batch} bf16, CUDA graph, {STEPS} steps averaged (CUPTI), rank 0 of {world}\n\n")
    f.write(f"span per step {span / 1e3:.3f} ms, kernels busy {busy / 1e3:.3f} ms, idle between kernels {(span - busy) / 1e3:.3f} ms "
            f"({len(ev) // STEPS} launches per step, median gap {sorted(gaps)[len(gaps) // 2]:.2f} us)\n\n")
    f.write("| kernel | launches/step | us/step | share |\n|---|---|---|---|\n")
    for k, (n, t) in rows:
        f.write(f"| `{k}` | {n / STEPS:.0f} | {t / STEPS:.1f} | {100 * t / STEPS / span:.1f}% |\n")
big = sorted(((g, ev[i].name[:60], ev[i + 1].name[:60]) for i, g in enumerate(gaps)), reverse=True)[:12 * STEPS]
with open(args.out, "a") as f:
    f.write("\nLargest gaps (us, after kernel -> before kernel), all %d steps:\n\n" % STEPS)
    for g, a, b in big:
        f.write(f"* {g:.1f}  `{a}` -> `{b}`\n")
    hist = collections.Counter(min(int(g), 10) for g in gaps)
    f.write("\nGap histogram (us, 10 = 10 or more): " + ", ".join(f"{k}: {v // STEPS}" for k, v in sorted(hist.items())) + "\n")
if args.launches:
    last = ev[len(ev) - len(ev) // STEPS:]
    t0 = last[0].time_range.start
    with open(args.launches, "w") as f:
        for e in last:
            k = e.name.replace("(anonymous namespace)::", "").replace("void ", "").split("(")[0][:90]
            f.write(f"{e.time_range.start - t0:10.1f} {e.time_range.end - e.time_range.start:8.1f}  {k}\n")
print(open(args.out).read()[:3000])
if world > 1:
    sys.stdout.flush()
    os._exit(0)
