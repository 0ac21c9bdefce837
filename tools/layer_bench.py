#!/usr/bin/env python
"""Per-layer roofline table of the hot-path kernels on one B200: every distinct ResNet-50 conv shape at batch 256
 x {fprop, dgrad, wgrad}, and the BN / depth-wise passes, timed with CUDA events through the C ABI.
L2 is flushed before every timed launch.  Writes a markdown table (profiles/) and a JSON file."""
import argparse
import ctypes as C
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from deep_vision_b200.lib import (BnBwdParams, BnParams, ConvParams, DwConvParams, DV_ALGO_AUTO, DV_BF16, DV_CONV_DGRAD,  # noqa: E402
                                  DV_CONV_FPROP, DV_CONV_WGRAD, check, lib)

PEAK_TF, PEAK_GB = 1686.9, 6445.0
try:
    _pk = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))
    PEAK_TF, PEAK_GB = _pk["bf16_tflops"], _pk["hbm_gbs"]
except Exception:
    pass

R50 = [  # C, K, R, stride, H(in), count  — SURVEY.md App. A
    (64, 64, 1, 1, 56, 1), (64, 256, 1, 1, 56, 4), (256, 64, 1, 1, 56, 2), (64, 64, 3, 1, 56, 3),
    (256, 128, 1, 2, 56, 1), (256, 512, 1, 2, 56, 1), (128, 128, 3, 1, 28, 4), (128, 512, 1, 1, 28, 4), (512, 128, 1, 1, 28, 3),
    (512, 256, 1, 2, 28, 1), (512, 1024, 1, 2, 28, 1), (256, 256, 3, 1, 14, 6), (256, 1024, 1, 1, 14, 6), (1024, 256, 1, 1, 14, 5),
    (1024, 512, 1, 2, 14, 1), (1024, 2048, 1, 2, 14, 1), (512, 512, 3, 1, 7, 3), (512, 2048, 1, 1, 7, 3), (2048, 512, 1, 1, 7, 2),
]


def stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


_flush = None


def timed(fn, iters=8, warm=2):
    global _flush
    if _flush is None:
        _flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(iters):
        _flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        e1.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]


def conv_rows(batch, iters=8):
    rows = []
    for (Cc, K, R, stride, H, count) in R50:
        pad = R // 2
        P = (H + 2 * pad - R) // stride + 1
        x = torch.randn(batch, H, H, Cc, device="cuda").to(torch.bfloat16)
        w = (torch.randn(K, R, R, Cc, device="cuda") * 0.05).to(torch.bfloat16)
        wt = torch.empty(Cc, R, R, K, device="cuda", dtype=torch.bfloat16)
        check(lib.dv_filter_prep(w.float().data_ptr(), None, wt.data_ptr(), K, R, R, Cc, DV_BF16, stream()), "prep")
        y = torch.empty(batch, P, P, K, device="cuda", dtype=torch.bfloat16)
        dy = torch.randn(batch, P, P, K, device="cuda").to(torch.bfloat16)
        dx = torch.empty_like(x)
        dw = torch.zeros(K, R, R, Cc, device="cuda")
        flops = 2.0 * batch * P * P * K * Cc * R * R
        # algorithmic bytes: what the pass has to touch.  A 1x1 stride-s convolution reads only every s-th pixel in each direction
        # (fprop, wgrad); its dgrad still has to WRITE all of dx (zeros between the strided pixels).
        x_touched = x.numel() if not (R == 1 and stride > 1) else batch * P * P * Cc
        bytes_by_pass = {"fprop": 2.0 * (x_touched + y.numel()) + 2.0 * w.numel(),
                         "dgrad": 2.0 * (x.numel() + y.numel()) + 2.0 * w.numel(),
                         "wgrad": 2.0 * (x_touched + y.numel()) + 4.0 * w.numel()}
        for name, pass_id, fn in (("fprop", DV_CONV_FPROP, lib.dv_conv_fprop), ("dgrad", DV_CONV_DGRAD, lib.dv_conv_dgrad),
                                  ("wgrad", DV_CONV_WGRAD, lib.dv_conv_wgrad)):
            p = ConvParams()
            p.N, p.H, p.W, p.C, p.K, p.R, p.S, p.stride, p.pad, p.dtype, p.algo = batch, H, H, Cc, K, R, R, stride, pad, DV_BF16, DV_ALGO_AUTO
            if pass_id == DV_CONV_FPROP:
                p.x, p.w, p.y = x.data_ptr(), w.data_ptr(), y.data_ptr()
                if os.environ.get("LB_STATS"):
                    sums = torch.zeros(2 * K, device="cuda")
                    p.stat_sums = sums.data_ptr()
            elif pass_id == DV_CONV_DGRAD:
                p.x, p.w, p.y = dx.data_ptr(), w.data_ptr(), dy.data_ptr()      # as in training: KRSC filters read MN-major
                if os.environ.get("LB_WT"):
                    p.w_t = wt.data_ptr()
            else:
                p.x, p.w, p.y = x.data_ptr(), dw.data_ptr(), dy.data_ptr()
                p.accumulate = 0 if os.environ.get("LB_WGRAD_PARTIALS") else 1   # as in training: tiles added into dW
            need = lib.dv_conv_workspace_bytes(C.byref(p), pass_id)
            ws = torch.empty(max(need, 16), dtype=torch.uint8, device="cuda")
            p.workspace, p.workspace_bytes = ws.data_ptr(), need
            algo = lib.dv_conv_algo(C.byref(p), pass_id)
            min_bytes = bytes_by_pass[name]
            ms = timed(lambda: check(fn(C.byref(p), stream()), name), iters=iters)
            tf, gb = flops / ms / 1e9, min_bytes / ms / 1e6
            ai = flops / min_bytes
            bound = "tensor" if ai > PEAK_TF * 1e3 / PEAK_GB else "hbm"
            frac = tf / PEAK_TF if bound == "tensor" else gb / PEAK_GB
            rows.append(dict(layer=f"{Cc}->{K} {R}x{R}/{stride} @{H}", passname=name, algo="tc" if algo == 2 else "simt", count=count,
                             ms=round(ms, 4), tflops=round(tf, 1), gbs=round(gb, 1), ai=round(ai, 1), bound=bound, frac=round(frac, 3),
                             gflop=round(flops / 1e9, 1), mb=round(min_bytes / 1e6, 1)))
        del x, w, wt, y, dy, dx, dw
    return rows


def time_weighted(rows):
    """Whole-step view of the convolution kernels: total FLOPs of all 52 non-stem convs x 3 passes over the sum of their times."""
    ms = sum(r["ms"] * r["count"] for r in rows)
    tflop = sum(r["gflop"] * r["count"] for r in rows) / 1e3
    tf = tflop / ms * 1e3
    return dict(ms=round(ms, 3), tflop=round(tflop, 3), tflops=round(tf, 1), frac_of_burst=round(tf / PEAK_TF, 4))


BN_SHAPES = {"resnet50": [(64, 112), (64, 56), (256, 56), (128, 28), (512, 28), (256, 14), (1024, 14), (512, 7), (2048, 7)],
             "mobilenet1": [(32, 112), (64, 112), (64, 56), (128, 56), (128, 28), (256, 28), (256, 14), (512, 14), (512, 7), (1024, 7)]}


def bn_rows(batch, model="resnet50"):
    rows = []
    for (Cc, H) in BN_SHAPES[model]:
        M = batch * H * H
        x = torch.randn(M, Cc, device="cuda").to(torch.bfloat16)
        y, dy, dx, res, dres = (torch.randn(M, Cc, device="cuda").to(torch.bfloat16) for _ in range(5))
        g, b, rm, rv, sm, si, dg, db = (torch.ones(Cc, device="cuda") for _ in range(8))
        wsz = lib.dv_bn_workspace_bytes(M, Cc)
        ws = torch.empty(wsz, dtype=torch.uint8, device="cuda")
        p = BnParams()
        p.x, p.y, p.residual, p.gamma, p.beta = x.data_ptr(), y.data_ptr(), res.data_ptr(), g.data_ptr(), b.data_ptr()
        p.running_mean, p.running_var, p.save_mean, p.save_invstd = rm.data_ptr(), rv.data_ptr(), sm.data_ptr(), si.data_ptr()
        p.workspace, p.workspace_bytes, p.M, p.C, p.dtype, p.relu, p.eval_mode, p.eps, p.momentum = ws.data_ptr(), wsz, M, Cc, DV_BF16, 1, 0, 1e-5, 0.1
        q = BnBwdParams()
        q.x, q.y, q.dy, q.dx, q.dres = x.data_ptr(), y.data_ptr(), dy.data_ptr(), dx.data_ptr(), dres.data_ptr()
        q.gamma, q.save_mean, q.save_invstd, q.dgamma, q.dbeta = g.data_ptr(), sm.data_ptr(), si.data_ptr(), dg.data_ptr(), db.data_ptr()
        q.workspace, q.workspace_bytes, q.M, q.C, q.dtype, q.relu = ws.data_ptr(), wsz, M, Cc, DV_BF16, 1
        E = M * Cc
        # the non-residual layers (two of three BatchNorms in a bottleneck): no identity, ReLU mask recomputed from x
        p2 = BnParams()
        C.memmove(C.byref(p2), C.byref(p), C.sizeof(p))
        p2.residual = None
        q2 = BnBwdParams()
        C.memmove(C.byref(q2), C.byref(q), C.sizeof(q))
        q2.y, q2.dres, q2.beta = None, None, b.data_ptr()
        for name, fn, arg, nbytes in (("stats", lib.dv_bn_stats, p, 2 * E), ("apply+res+relu", lib.dv_bn_apply, p, 6 * E),
                                      ("bwd_reduce", lib.dv_bn_bwd_reduce, q, 6 * E), ("bwd_apply+dres", lib.dv_bn_bwd_apply, q, 10 * E),
                                      ("apply+relu (no residual)", lib.dv_bn_apply, p2, 4 * E),
                                      ("bwd_reduce (mask recomputed)", lib.dv_bn_bwd_reduce, q2, 4 * E),
                                      ("bwd_apply (mask recomputed)", lib.dv_bn_bwd_apply, q2, 6 * E)):
            ms = timed(lambda: check(fn(C.byref(arg), stream()), name))
            gb = nbytes / ms / 1e6
            rows.append(dict(layer=f"BN C={Cc} @{H}", passname=name, ms=round(ms, 4), gbs=round(gb, 1), frac=round(gb / PEAK_GB, 3),
                             mb=round(nbytes / 1e6, 1)))
    return rows


def dw_rows(batch):
    rows = []
    for (Cc, H, stride) in [(32, 112, 1), (64, 112, 2), (128, 56, 1), (128, 56, 2), (256, 28, 1), (256, 28, 2), (512, 14, 1), (512, 14, 2), (1024, 7, 1)]:
        P = (H - 1) // stride + 1
        x = torch.randn(batch, H, H, Cc, device="cuda").to(torch.bfloat16)
        y = torch.randn(batch, P, P, Cc, device="cuda").to(torch.bfloat16)
        w = torch.randn(Cc, 3, 3, device="cuda")
        p = DwConvParams()
        p.x, p.w, p.y, p.N, p.H, p.W, p.C, p.stride, p.dtype = x.data_ptr(), w.data_ptr(), y.data_ptr(), batch, H, H, Cc, stride, DV_BF16
        need = lib.dv_dwconv_workspace_bytes(C.byref(p))
        ws = torch.empty(need, dtype=torch.uint8, device="cuda")
        p.workspace, p.workspace_bytes = ws.data_ptr(), need
        nbytes = 2 * (x.numel() + y.numel())
        for name, fn in (("fprop", lib.dv_dwconv_fprop), ("dgrad", lib.dv_dwconv_dgrad), ("wgrad", lib.dv_dwconv_wgrad)):
            ms = timed(lambda: check(fn(C.byref(p), stream()), name))
            gb = nbytes / ms / 1e6
            rows.append(dict(layer=f"dw3x3 C={Cc} /{stride} @{H}", passname=name, ms=round(ms, 4), gbs=round(gb, 1), frac=round(gb / PEAK_GB, 3),
                             mb=round(nbytes / 1e6, 1)))
    return rows


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--out", default="gpurun_out/layer_bench")
    ap.add_argument("--what", default="conv,bn,dw")
    ap.add_argument("--bn-model", default="resnet50", choices=sorted(BN_SHAPES), help="whose BatchNorm tensor shapes the bn table uses")
    args = ap.parse_args()
    out = {}
    md = [f"# Per-layer roofline, batch {args.batch}, bf16 (peaks: {PEAK_TF} TFLOP/s burst, {PEAK_GB} GB/s measured)", ""]
    if "conv" in args.what:
        rows = conv_rows(args.batch)
        out["conv"] = rows
        md += ["## ResNet-50 dense convolutions", "", "| layer | pass | algo | ms | TFLOP/s | GB/s (touched bytes) | AI | bound | frac of roof | x count |",
               "|---|---|---|---|---|---|---|---|---|---|"]
        tot = 0.0
        for r in rows:
            md.append(f"| {r['layer']} | {r['passname']} | {r['algo']} | {r['ms']} | {r['tflops']} | {r['gbs']} | {r['ai']} | {r['bound']} | {r['frac']} | {r['count']} |")
            tot += r["ms"] * r["count"]
        tw = time_weighted(rows)
        md += ["", f"Sum over the 52 non-stem convs x 3 passes: **{tot:.2f} ms** per step = {tw['tflops']} TFLOP/s time-weighted = "
                   f"**{tw['frac_of_burst']} of the burst peak** ({tw['tflop']} TFLOP).", ""]
    if "bn" in args.what:
        rows = bn_rows(args.batch, args.bn_model)
        out["bn"] = rows
        md += ["## Batch norm passes", "", "| tensor | pass | ms | GB/s | frac of HBM roof |", "|---|---|---|---|---|"]
        md += [f"| {r['layer']} | {r['passname']} | {r['ms']} | {r['gbs']} | {r['frac']} |" for r in rows]
        md.append("")
    if "dw" in args.what:
        rows = dw_rows(args.batch)
        out["dw"] = rows
        md += ["## MobileNet depth-wise 3x3", "", "| tensor | pass | ms | GB/s | frac of HBM roof |", "|---|---|---|---|---|"]
        md += [f"| {r['layer']} | {r['passname']} | {r['ms']} | {r['gbs']} | {r['frac']} |" for r in rows]
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    open(args.out + ".md", "w").write("\n".join(md) + "\n")
    json.dump(out, open(args.out + ".json", "w"))
    print("\n".join(md))


if __name__ == "__main__":
    main()
