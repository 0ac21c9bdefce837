#!/usr/bin/env python
"""bench.py — headline benchmark of the hot path: ResNet-50 V1 bf16 training step, batch 256/GPU, synthetic 224x224
(BASELINE.json configs[1]; configs[2] under torchrun with --gpus N).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--model resnet50|resnet34|resnet152|mobilenet1] [--batch B]
    python bench.py --impl reference ...     # the reference's PyTorch CPU path (oracle port) on the host cores

Prints ONE JSON line (rank 0).  `value` = images/s of the whole job, inputs resident in HBM; `e2e` = the same step fed
from pinned host memory with the loss read back every step (the reference's train.py:444-470 behaviour);
`roofline` = the dominant kernel class timed with CUDA events in this process; `cpu_baseline` = the oracle port of the
reference step on the host cores (bounded sample).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

MODELS = {
    "resnet50": dict(ctor="ResNet50", batch=256, gflop_per_img=22.90),
    "resnet34": dict(ctor="ResNet34", batch=256, gflop_per_img=10.72),
    "resnet152": dict(ctor="ResNet152", batch=128, gflop_per_img=67.45),
    "mobilenet1": dict(ctor="MobileNetV1", batch=256, gflop_per_img=3.28),
}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=d["hbm_gbs"], tf_burst=d["bf16_tflops"], tf_sustained=d.get("bf16_tflops_sustained", d["bf16_tflops"]),
                    source="measured")
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sustained=1400.0, source="fallback")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled while the timed region runs."""

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def __enter__(self):
        q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
            "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={q}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def __exit__(self, *a):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except subprocess.TimeoutExpired:
                self.proc.kill()

    def summary(self):
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm)}


def cpu_reference_step(model, batch, steps, warmup):
    """The reference's training step on the host cores: oracle/ref_models.py (a torch-functional port of the model
    files; /root/reference does not exist on the GPU box), fp32, all host threads, same synthetic distribution."""
    import torch
    from oracle import ref_models
    import deep_vision_b200 as dv
    from deep_vision_b200.train import TRAINING_CONFIG
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    torch.manual_seed(0)
    net = getattr(dv.models, MODELS[model]["ctor"])()
    sd = ref_models.clone_state(net.state_dict(), dtype=torch.float32, requires_grad=True)
    cfg = TRAINING_CONFIG[model]
    g = torch.Generator().manual_seed(1234)
    images = torch.randn(batch, 3, 224, 224, generator=g)
    labels = torch.randint(0, 1000, (batch,), generator=g)
    opt_state = {}
    kw = dict(optimizer=cfg["optimizer"], lr=cfg["lr"], momentum=cfg.get("momentum", 0.0), weight_decay=cfg["weight_decay"],
              alpha=cfg.get("alpha", 0.9), eps=cfg.get("eps", 1.0))
    for _ in range(warmup):
        ref_models.train_step(model, sd, images, labels, opt_state, **kw)
    t0 = time.perf_counter()
    for _ in range(steps):
        ref_models.train_step(model, sd, images, labels, opt_state, **kw)
    dt = time.perf_counter() - t0
    return batch * steps / dt, dt / steps, cores


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    batch = args.cpu_batch
    ips, sec, cores = cpu_reference_step(args.model, batch, max(1, min(args.steps, 3)), 1)
    line = {
        "impl": "reference", "metric": "train images/sec", "value": round(ips, 3), "unit": "images/s", "n_gpus": args.gpus,
        "steps": max(1, min(args.steps, 3)), "warmup": 1, "ms_per_step": round(sec * 1e3, 2), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{args.model} V1 training step (fwd+CE+bwd+optimizer), 224x224 synthetic, CPU batch {batch} "
                               f"(bounded sample of the batch-{MODELS[args.model]['batch']}/GPU workload)"},
        "cpu_baseline": {"value": round(ips, 3), "unit": "images/s", "cores": cores, "kind": "port",
                         "sample": f"{max(1, min(args.steps, 3))} steps of batch {batch} after 1 warm-up"},
        "e2e": {"value": round(ips, 3), "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def kernel_roofline(torch, dv, pk):
    """Dominant kernel class = the tcgen05 implicit-GEMM convolution.  Timed here on its heaviest ResNet-50 layer shape
    (3x3, 256->256 at 14x14, batch 256: M=50176, N=256, K=2304, 59.2 GFLOP) with CUDA events on the launching stream,
    L2 flushed between launches."""
    import ctypes as C
    from deep_vision_b200.lib import ConvParams, DV_ALGO_TC, DV_BF16, check, lib
    N, H, Cc, K, R = 256, 14, 256, 256, 3
    x = torch.randn(N, H, H, Cc, device="cuda").to(torch.bfloat16)
    w = (torch.randn(K, R, R, Cc, device="cuda") * 0.05).to(torch.bfloat16)
    y = torch.empty(N, H, H, K, device="cuda", dtype=torch.bfloat16)
    p = ConvParams()
    p.x, p.w, p.y = x.data_ptr(), w.data_ptr(), y.data_ptr()
    p.N, p.H, p.W, p.C, p.K, p.R, p.S, p.stride, p.pad, p.dtype, p.algo = N, H, H, Cc, K, R, R, 1, 1, DV_BF16, DV_ALGO_TC
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    for _ in range(3):
        check(lib.dv_conv_fprop(C.byref(p), st), "dv_conv_fprop")
    times = []
    for _ in range(10):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        check(lib.dv_conv_fprop(C.byref(p), st), "dv_conv_fprop")
        e1.record()
        e1.synchronize()
        times.append(e0.elapsed_time(e1))
    ms = sum(times) / len(times)
    flops = 2.0 * N * H * H * K * Cc * R * R
    achieved = flops / (ms * 1e-3) / 1e12
    return {"bound": "tensor", "kernel": "conv_tc_gemm_pair_kernel<6> (cta_group::2) fprop 3x3 256->256 @14x14 B256", "achieved": round(achieved, 1),
            "peak": pk["tf_burst"], "unit": "TFLOP/s", "frac": round(achieved / pk["tf_burst"], 4),
            # dram__bytes_read.sum + dram__bytes_write.sum per launch from `ncu --set full` (profiles/r01_ncu_full_v9_pair_wgrad.csv);
            # the algorithmic minimum is 52.6 MB (x 25.7 + y 25.7 + w 1.2): the output was still in L2 when the capture ended
            "traffic": 26.9e6, "algorithmic_flops": flops,
            "ms_per_launch": round(ms, 4), "peak_source": pk["source"]}


def hbm_roofline(torch, pk):
    """The dominant HBM-bound kernel class of the step (BatchNorm passes = 45 % of it): dv_bn_bwd_apply on a 256-channel
    56x56 tensor at batch 256 (no residual: reads x and dy, writes dx = 6 bytes per element), CUDA events, L2 flushed."""
    import ctypes as C
    from deep_vision_b200.lib import BnBwdParams, DV_BF16, check, lib
    M, Cc = 256 * 56 * 56, 256
    x, dy, dx = (torch.randn(M, Cc, device="cuda").to(torch.bfloat16) for _ in range(3))
    g, b, sm, si, dg, db = (torch.ones(Cc, device="cuda") for _ in range(6))
    q = BnBwdParams()
    q.x, q.dy, q.dx, q.gamma, q.beta = x.data_ptr(), dy.data_ptr(), dx.data_ptr(), g.data_ptr(), b.data_ptr()
    q.save_mean, q.save_invstd, q.dgamma, q.dbeta = sm.data_ptr(), si.data_ptr(), dg.data_ptr(), db.data_ptr()
    q.M, q.C, q.dtype, q.relu = M, Cc, DV_BF16, 1
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    for _ in range(3):
        check(lib.dv_bn_bwd_apply(C.byref(q), st), "dv_bn_bwd_apply")
    times = []
    for _ in range(10):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        check(lib.dv_bn_bwd_apply(C.byref(q), st), "dv_bn_bwd_apply")
        e1.record()
        e1.synchronize()
        times.append(e0.elapsed_time(e1))
    ms = sum(times) / len(times)
    nbytes = 6.0 * M * Cc
    gbs = nbytes / (ms * 1e-3) / 1e9
    return {"bound": "hbm", "kernel": "bn_bwd_apply_kernel<bf16> C=256 @56x56 B256 (ReLU mask recomputed)", "achieved": round(gbs, 1),
            "peak": pk["hbm"], "unit": "GB/s", "frac": round(gbs / pk["hbm"], 4), "traffic": None, "algorithmic_bytes": nbytes,
            "ms_per_launch": round(ms, 4), "peak_source": pk["source"]}


def run_ours(args):
    import torch
    import torch.distributed as dist
    import deep_vision_b200 as dv
    from deep_vision_b200.lib import lib
    from deep_vision_b200.train import TRAINING_CONFIG, Trainer

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(local_rank)
    if world > 1:
        # NCCL announces its version on stdout when NCCL_DEBUG is VERSION or higher; stdout carries exactly one JSON line
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    if lib.dv_device_check() != 0:
        raise RuntimeError(lib.dv_last_error().decode())
    spec = MODELS[args.model]
    batch = args.batch or spec["batch"]
    dv.set_compute_dtype(torch.bfloat16)
    torch.manual_seed(0)  # identical replicas on every rank
    net = getattr(dv.models, spec["ctor"])().cuda()
    cfg = dict(TRAINING_CONFIG[args.model])
    cfg.pop("batch_size")
    trainer = Trainer(net, use_cuda_graph=not args.no_graph, **cfg)
    g = torch.Generator(device="cuda").manual_seed(1234 + rank)
    images = torch.randn(batch, 3, 224, 224, device="cuda", generator=g)
    labels = torch.randint(0, 1000, (batch,), device="cuda", generator=g)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # launches per step (the CUDA graph replays them without passing through the library's counter)
    c0 = lib.dv_launch_count()
    trainer.use_graph, keep = False, trainer.use_graph
    trainer.step(images, labels)
    torch.cuda.synchronize()
    launches_per_step = int(lib.dv_launch_count() - c0)
    trainer.use_graph = keep

    for _ in range(max(args.warmup, 3)):
        trainer.step(images, labels)
    barrier()
    if args.profile:
        for _ in range(args.steps):
            trainer.step(images, labels)
        barrier()
        if rank == 0:
            print(json.dumps({"profile": True, "launches_per_step": launches_per_step, "steps": args.steps}), flush=True)
        if world > 1:
            sys.stdout.flush()
            os._exit(0)
        return
    with ClockSampler(local_rank) as clk:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.steps):
            loss = trainer.step(images, labels)
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
    t = torch.tensor([ms], device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    loss_val = float(loss.item())

    # end to end: every step's batch starts in pinned host memory and the loss is read back to the host every step
    # (train.py:444-446,470).  HostFeeder copies batch i+1 on a side stream while step i computes.
    from deep_vision_b200.train import HostFeeder
    himg = [torch.randn(batch, 3, 224, 224).pin_memory() for _ in range(2)]
    hlab = [torch.randint(0, 1000, (batch,)).pin_memory() for _ in range(2)]
    feeder = HostFeeder(trainer)
    k2 = max(3, min(args.steps, 20))

    def e2e_loop(n):
        feeder.prefetch(himg[0], hlab[0])
        last = 0.0
        for i in range(n):
            dimg, dlab = feeder.take()
            if i + 1 < n:
                feeder.prefetch(himg[(i + 1) & 1], hlab[(i + 1) & 1])
            loss_t = trainer.step(dimg, dlab)
            feeder.release()
            last = loss_t.item()                     # device -> host read of the step's result, every step
        return last

    e2e_loop(2)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    e2e_loop(k2)
    e1.record()
    barrier()
    t = torch.tensor([e0.elapsed_time(e1)], device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_e2e = float(t.item())

    if rank == 0:
        pk = peaks()
        ips = batch * world * args.steps / (ms * 1e-3)
        ips_e2e = batch * world * k2 / (ms_e2e * 1e-3)
        line = {
            "metric": "train images/sec", "value": round(ips, 1), "unit": "images/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": round(ms / args.steps, 3), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": f"{args.model} V1 bf16 training step (fwd+CE+bwd+grad allreduce+{cfg['optimizer']}), batch {batch}/GPU, "
                                   f"3x224x224 synthetic, random-init weights", "global_batch": batch * world,
                       "parallelism": f"dp{world}", "cuda_graph": not args.no_graph,
                       "l2": "per-step working set (>10 GB of activations) exceeds the 126 MB L2; no explicit flush"},
            "e2e": {"value": round(ips_e2e, 1), "unit": "images/s", "h2d_bytes_per_step": int(himg[0].numel() * 4 + hlab[0].numel() * 8),
                    "d2h_bytes_per_step": 4, "steps": k2},
            "gpu_launches": launches_per_step * args.steps, "launches_per_step": launches_per_step,
            "clocks": clk.summary(), "loss": round(loss_val, 4),
            "conv_tflops": round(spec["gflop_per_img"] * ips / 1e3, 1),
            "conv_frac_of_peak": round(spec["gflop_per_img"] * ips / world / 1e3 / pk["tf_sustained"], 4),
        }
        if world == 1 and not args.no_roofline:
            try:
                line["roofline"] = kernel_roofline(torch, dv, pk)
                line["roofline_hbm"] = hbm_roofline(torch, pk)
            except Exception as e:  # the headline number stands on its own
                line["roofline"] = {"error": str(e)[:200]}
        if world == 1 and not args.no_cpu_baseline:
            nsteps = 2
            ips_cpu, sec, cores = cpu_reference_step(args.model, args.cpu_batch, nsteps, 1)
            line["cpu_baseline"] = {"value": round(ips_cpu, 3), "unit": "images/s", "cores": cores, "kind": "port",
                                    "sample": f"{nsteps} steps of batch {args.cpu_batch} (fp32, oracle/ref_models.py) after 1 warm-up"}
        print(json.dumps(line), flush=True)
    if world > 1:
        # Tearing down a process group whose collectives were captured in a CUDA graph can block for minutes; the numbers
        # are out, every rank is past the barrier, so leave without the teardown.
        sys.stdout.flush()
        sys.stderr.flush()
        dist.barrier()
        torch.cuda.synchronize()
        os._exit(0)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--model", default="resnet50", choices=sorted(MODELS))
    ap.add_argument("--batch", type=int, default=0)
    ap.add_argument("--cpu-batch", type=int, default=16)
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--no-roofline", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--profile", action="store_true", help="warm up, run --steps eager steps, print a short line, exit (for ncu)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
