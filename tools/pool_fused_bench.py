"""Times dv_bn_relu_maxpool_{fwd,bwd} against the separate launches they replace, at the ResNet stem's size
(N x 112 x 112 x 64 bf16).  CUDA events on the launching stream, L2 flushed between repetitions by the tensors' own size."""
import ctypes as C
import sys

import torch

sys.path.insert(0, ".")
from deep_vision_b200 import lib as L  # noqa: E402


def timed(fn, reps=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(reps + 1)]
    ev[0].record()
    for i in range(reps):
        fn()
        ev[i + 1].record()
    torch.cuda.synchronize()
    return min(ev[i].elapsed_time(ev[i + 1]) for i in range(reps)) * 1e3


def main():
    N, H, W, Cc = int(sys.argv[1]) if len(sys.argv) > 1 else 256, 112, 112, 64
    P, Q = H // 2, W // 2
    st = torch.cuda.current_stream().cuda_stream
    x = torch.randn(N, H, W, Cc, device="cuda").bfloat16()
    y = torch.empty_like(x)
    pool = torch.empty(N, P, Q, Cc, device="cuda", dtype=torch.bfloat16)
    idx = torch.empty(N, P, Q, Cc, device="cuda", dtype=torch.uint8)
    dpool = torch.randn(N, P, Q, Cc, device="cuda").bfloat16()
    dy, dx = torch.empty_like(x), torch.empty_like(x)
    g, b = torch.rand(Cc, device="cuda") + 0.5, torch.randn(Cc, device="cuda")
    sm, si, dg, db = (torch.empty(Cc, device="cuda") for _ in range(4))
    M = N * H * W
    wsz = L.lib.dv_bn_workspace_bytes(M, Cc)
    ws = torch.empty(wsz, dtype=torch.uint8, device="cuda")
    p = L.BnParams()
    p.x, p.y, p.gamma, p.beta, p.save_mean, p.save_invstd = x.data_ptr(), y.data_ptr(), g.data_ptr(), b.data_ptr(), sm.data_ptr(), si.data_ptr()
    p.workspace, p.workspace_bytes, p.M, p.C, p.dtype, p.relu, p.eps, p.momentum = ws.data_ptr(), wsz, M, Cc, L.DV_BF16, 1, 1e-5, 0.1
    L.check(L.lib.dv_bn_stats(C.byref(p), st), "stats")
    r = L.BnBwdParams()
    r.x, r.dy, r.dx, r.gamma, r.beta = x.data_ptr(), dy.data_ptr(), dx.data_ptr(), g.data_ptr(), b.data_ptr()
    r.save_mean, r.save_invstd, r.dgamma, r.dbeta = sm.data_ptr(), si.data_ptr(), dg.data_ptr(), db.data_ptr()
    r.workspace, r.workspace_bytes, r.M, r.C, r.dtype, r.relu = ws.data_ptr(), wsz, M, Cc, L.DV_BF16, 1
    q = L.BnPoolParams()
    q.x, q.y, q.idx, q.dy, q.dx = x.data_ptr(), pool.data_ptr(), idx.data_ptr(), dpool.data_ptr(), dx.data_ptr()
    q.gamma, q.beta, q.save_mean, q.save_invstd, q.dgamma, q.dbeta = g.data_ptr(), b.data_ptr(), sm.data_ptr(), si.data_ptr(), dg.data_ptr(), db.data_ptr()
    q.N, q.H, q.W, q.C, q.dtype = N, H, W, Cc, L.DV_BF16
    w2 = L.lib.dv_bn_pool_workspace_bytes(C.byref(q))
    ws2 = torch.empty(w2, dtype=torch.uint8, device="cuda")
    q.workspace, q.workspace_bytes = ws2.data_ptr(), w2
    E = x.numel() * 2
    rows = [
        ("bn_apply", lambda: L.lib.dv_bn_apply(C.byref(p), st), 2 * E),
        ("maxpool_fwd", lambda: L.lib.dv_maxpool_fwd(y.data_ptr(), pool.data_ptr(), idx.data_ptr(), N, H, W, Cc, L.DV_BF16, st), E + E // 4 + E // 8),
        ("FUSED fwd", lambda: L.lib.dv_bn_relu_maxpool_fwd(C.byref(q), st), E + E // 4 + E // 8),
        ("maxpool_bwd", lambda: L.lib.dv_maxpool_bwd(dpool.data_ptr(), idx.data_ptr(), dy.data_ptr(), N, H, W, Cc, L.DV_BF16, st), E + E // 4 + E // 8),
        ("bn_bwd_reduce", lambda: L.lib.dv_bn_bwd_reduce(C.byref(r), st), 2 * E),
        ("bn_bwd_apply", lambda: L.lib.dv_bn_bwd_apply(C.byref(r), st), 3 * E),
        ("FUSED bwd (3 launches)", lambda: L.lib.dv_bn_relu_maxpool_bwd(C.byref(q), st), 3 * E + 2 * (E // 4 + E // 8)),
    ]
    print(f"| launch | us | algorithmic MB | GB/s |\n|---|---|---|---|")
    for name, fn, nbytes in rows:
        t = timed(fn)
        print(f"| {name} | {t:.1f} | {nbytes / 1e6:.0f} | {nbytes / t / 1e3:.0f} |")


main()
