#!/usr/bin/env python
"""Data-parallel equivalence on real GPUs: after one Trainer step on W ranks with different batches,
every rank holds the same parameters, and they equal SGD applied to the AVERAGE of the per-shard gradients computed by plain
autograd (per-shard BatchNorm statistics, as nn.DataParallel).  Run under torchrun; rank 0 prints one JSON line.

DP_CHECK_BACKEND=gloo puts every rank on GPU 0 and exchanges through gloo (NCCL refuses two ranks on one device): the same
Trainer / BucketedAllReduce code path — bucket hooks, side-stream weight gradients, 1/world in the optimiser — on a box with
a single GPU.  gloo collectives cannot be captured in a CUDA graph, so only the eager step is checked there."""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import deep_vision_b200 as dv  # noqa: E402
from deep_vision_b200.train import Trainer  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
backend = os.environ.get("DP_CHECK_BACKEND", "nccl")
if backend == "gloo":
    local = 0
    torch.cuda.set_device(0)
    dist.init_process_group("gloo")
else:
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
dv.set_compute_dtype(torch.bfloat16)
# fixed-order reductions: a random-init ResNet-50 on a tiny batch amplifies the run-to-run rounding of fp32 atomics to O(1)
# gradient differences (measured: the SAME plain-autograd gradient computed twice differed by 1.09 relative), which would
# drown what this script checks
dv.set_deterministic(os.environ.get("DP_CHECK_ATOMICS", "0") != "1")
out = {}
for graph in ((False,) if backend == "gloo" else (False, True)):
    torch.manual_seed(0)
    net = dv.models.ResNet50().cuda()
    ref = dv.models.ResNet50().cuda()
    ref.load_state_dict(net.state_dict())
    lr, wd = 0.1, 1e-4
    tr = Trainer(net, optimizer="sgd", lr=lr, momentum=0.9, weight_decay=wd, use_cuda_graph=graph)
    batches = []
    for r in range(world):
        g = torch.Generator().manual_seed(100 + r)   # CPU generator: the same values on every rank
        batches.append((torch.randn(8, 3, 96, 96, generator=g).cuda(), torch.randint(0, 1000, (8,), generator=g).cuda()))
    p0 = tr.flat_p.clone()
    tr.step(*batches[rank])
    torch.cuda.synchronize()
    if not graph:
        ex = tr.exchange
        out["debug"] = {"world": tr.world, "buckets": len(ex.buckets), "pending_after": list(ex._pending), "launched": list(ex.launch_order),
                        "need": list(ex._need)}
        gg = [torch.empty_like(tr.flat_g) for _ in range(world)]
        dist.all_gather(gg, tr.flat_g)
        out["debug"]["flat_g_identical"] = bool(torch.equal(gg[0], gg[1]))
        out["debug"]["flat_g_norms"] = [float(t.norm()) for t in gg]
        pp = [torch.empty_like(p0) for _ in range(world)]
        dist.all_gather(pp, p0)
        out["debug"]["p0_identical"] = bool(torch.equal(pp[0], pp[1]))
    mine = tr.flat_p.clone()
    gathered = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(gathered, mine)
    out[f"ranks_identical_graph{int(graph)}"] = bool(all(torch.equal(gathered[0], t) for t in gathered))
    if not graph:
        # plain autograd on every shard, averaged
        grads = None
        for (x, y) in batches:
            ref.zero_grad(set_to_none=True)
            ref.train()
            loss = torch.nn.functional.cross_entropy(ref(x), y)
            loss.backward()
            gs = torch.cat([p.grad.flatten() for p in ref.parameters()])
            grads = gs if grads is None else grads + gs
        # the trainer's reduced gradient buffer (sum over ranks) in parameter order
        # noise floor: the same plain-autograd gradient computed twice (fp32 atomics + a random-init network)
        again = None
        for (x, y) in batches:
            ref.zero_grad(set_to_none=True)
            torch.nn.functional.cross_entropy(ref(x), y).backward()
            gs = torch.cat([p.grad.flatten() for p in ref.parameters()])
            again = gs if again is None else again + gs
        out["plain_autograd_twice_rel_diff"] = float((again - grads).norm() / grads.norm())
        gt = torch.cat([p.grad.flatten() for p in net.parameters()])
        out["grad_sum_rel_err"] = float((gt - grads).norm() / grads.norm())
        grads /= world
        pref = torch.cat([p.detach().flatten() for p in ref.parameters()])
        want = pref - lr * (grads + wd * pref)
        got = torch.cat([p.detach().flatten() for p in net.parameters()])
        out["update_rel_err"] = float((got - want).norm() / (want - pref).norm())
    del tr, net, ref
    torch.cuda.empty_cache()
if rank == 0:
    print(json.dumps(out), flush=True)
dist.barrier()
torch.cuda.synchronize()
os._exit(0)
