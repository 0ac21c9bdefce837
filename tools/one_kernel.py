#!/usr/bin/env python
"""Launch ONE hot kernel a few times (for `ncu --set full`): the tcgen05 conv fprop on the roofline shape of bench.py
(3x3, 256->256, 14x14, batch 256), or --which wgrad / bn_apply / bn_bwd."""
import argparse
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from deep_vision_b200.lib import BnBwdParams, BnParams, ConvParams, DwConvParams, DV_ALGO_TC, DV_BF16, DV_CONV_WGRAD, check, lib  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--which", default="fprop")
ap.add_argument("--iters", type=int, default=5)
ap.add_argument("--shape", default="256,14,256,256,3", help="N,H,C,K,R (stride 1, pad R//2)")
ap.add_argument("--jobs", default="", help="several kernels in one process: 'fprop:256,14,256,256,3;wgrad_acc:256,56,64,256,1'")
args = ap.parse_args()
st = C.c_void_p(torch.cuda.current_stream().cuda_stream)


def run_dw(which, shape, iters):
    N, H, Cc, stride = (int(v) for v in shape.split(","))
    P = (H - 1) // stride + 1
    x = torch.randn(N, H, H, Cc, device="cuda").to(torch.bfloat16)
    y = torch.randn(N, P, P, Cc, device="cuda").to(torch.bfloat16)
    w = torch.randn(Cc, 3, 3, device="cuda")
    p = DwConvParams()
    p.x, p.w, p.y, p.N, p.H, p.W, p.C, p.stride, p.dtype = x.data_ptr(), w.data_ptr(), y.data_ptr(), N, H, H, Cc, stride, DV_BF16
    need = lib.dv_dwconv_workspace_bytes(C.byref(p))
    ws = torch.empty(max(need, 16), dtype=torch.uint8, device="cuda")
    p.workspace, p.workspace_bytes = ws.data_ptr(), need
    fn = {"dw_fprop": lib.dv_dwconv_fprop, "dw_dgrad": lib.dv_dwconv_dgrad, "dw_wgrad": lib.dv_dwconv_wgrad}[which]
    for _ in range(iters):
        check(fn(C.byref(p), st), which)
    torch.cuda.synchronize()
    print("ok", which)


def run_job(which, shape, iters):
    if which.startswith("dw_"):
        return run_dw(which, shape, iters)
    N, H, Cc, K, R = (int(v) for v in shape.split(","))
    x = torch.randn(N, H, H, Cc, device="cuda").to(torch.bfloat16)
    w = (torch.randn(K, R, R, Cc, device="cuda") * 0.05).to(torch.bfloat16)
    y = torch.empty(N, H, H, K, device="cuda", dtype=torch.bfloat16)
    p = ConvParams()
    p.N, p.H, p.W, p.C, p.K, p.R, p.S, p.stride, p.pad, p.dtype, p.algo = N, H, H, Cc, K, R, R, 1, R // 2, DV_BF16, DV_ALGO_TC
    if which in ("fprop", "fprop_stats"):
        p.x, p.w, p.y = x.data_ptr(), w.data_ptr(), y.data_ptr()
        if which == "fprop_stats":    # with the following BatchNorm's statistics from the epilogue (what the training step launches)
            sums = torch.zeros(2 * K, device="cuda")
            p.stat_sums = sums.data_ptr()
        fn = lambda: check(lib.dv_conv_fprop(C.byref(p), st), "fprop")
    elif which == "dgrad_s2":         # 1x1 stride-2 dgrad (Bottleneck conv1 / projection of a down-sampling block): strided row scatter
        dx = torch.empty_like(x)
        dy2 = torch.randn(N, H // 2, H // 2, K, device="cuda").to(torch.bfloat16)
        p.stride, p.pad = 2, 0
        p.x, p.w, p.y = dx.data_ptr(), w.data_ptr(), dy2.data_ptr()
        keep = dy2
        fn = lambda: check(lib.dv_conv_dgrad(C.byref(p), st), "dgrad_s2")
    elif which in ("dgrad", "dgrad_bnb"):
        dx = torch.empty_like(x)
        p.x, p.w, p.y = dx.data_ptr(), w.data_ptr(), y.normal_().data_ptr()
        if which == "dgrad_bnb":      # with the BatchNorm-backward sums of the producing layer in the epilogue
            x0 = torch.randn_like(x)
            v = [torch.ones(Cc, device="cuda") for _ in range(4)] + [torch.zeros(Cc, device="cuda") for _ in range(2)]
            p.bnb_x = x0.data_ptr()
            p.bnb_gamma, p.bnb_beta, p.bnb_mean, p.bnb_invstd, p.bnb_dgamma, p.bnb_dbeta = [t.data_ptr() for t in v]
            keep = (x0, v)
        fn = lambda: check(lib.dv_conv_dgrad(C.byref(p), st), "dgrad")
    elif which in ("wgrad", "wgrad_acc"):
        dw = torch.zeros(K, R, R, Cc, device="cuda")
        p.accumulate = 1 if which == "wgrad_acc" else 0
        p.x, p.w, p.y = x.data_ptr(), dw.data_ptr(), y.normal_().data_ptr()
        need = lib.dv_conv_workspace_bytes(C.byref(p), DV_CONV_WGRAD)
        ws = torch.empty(need, dtype=torch.uint8, device="cuda")
        p.workspace, p.workspace_bytes = ws.data_ptr(), need
        fn = lambda: check(lib.dv_conv_wgrad(C.byref(p), st), "wgrad")
    else:
        M, Cb = 256 * 56 * 56, 256
        xb = torch.randn(M, Cb, device="cuda").to(torch.bfloat16)
        yb = torch.empty_like(xb)
        v = [torch.ones(Cb, device="cuda") for _ in range(6)]
        q = BnParams()
        q.x, q.y, q.gamma, q.beta, q.running_mean, q.running_var, q.save_mean, q.save_invstd = [t.data_ptr() for t in [xb, yb] + v]
        wsz = lib.dv_bn_workspace_bytes(M, Cb)
        ws = torch.empty(wsz, dtype=torch.uint8, device="cuda")
        q.workspace, q.workspace_bytes, q.M, q.C, q.dtype, q.relu, q.eps, q.momentum = ws.data_ptr(), wsz, M, Cb, DV_BF16, 1, 1e-5, 0.1
        if which == "bn_bwd_apply":     # the shape and variant bench.py's roofline_hbm times: no residual, mask recomputed
            dyb, dxb = torch.randn_like(xb), torch.empty_like(xb)
            b = BnBwdParams()
            b.x, b.dy, b.dx, b.gamma, b.beta = xb.data_ptr(), dyb.data_ptr(), dxb.data_ptr(), v[0].data_ptr(), v[1].data_ptr()
            b.save_mean, b.save_invstd, b.dgamma, b.dbeta = v[2].data_ptr(), v[3].data_ptr(), v[4].data_ptr(), v[5].data_ptr()
            b.M, b.C, b.dtype, b.relu = M, Cb, DV_BF16, 1
            keep = (dyb, dxb)
            fn = lambda: check(lib.dv_bn_bwd_apply(C.byref(b), st), "bwd_apply")
        else:
            fn = (lambda: check(lib.dv_bn_apply(C.byref(q), st), "apply")) if which == "bn_apply" else (lambda: check(lib.dv_bn_stats(C.byref(q), st), "stats"))
    for _ in range(iters):
        fn()
    torch.cuda.synchronize()
    print("ok", which)


if args.jobs:
    for job in args.jobs.split(";"):
        run_job(*job.split(":"), args.iters)
else:
    run_job(args.which, args.shape, args.iters)
