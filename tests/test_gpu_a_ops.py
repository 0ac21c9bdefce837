"""GPU parity tests, per op, THROUGH THE C ABI (ctypes on include/dv.h) against the CPU oracle (oracle/ref_ops.py).

Bars (BASELINE.json north_star): fp32 storage <= 1e-4 relative L2 on activations and gradients; bf16 storage
<= 2e-2 relative L2 (inputs are pre-rounded to bf16 so both sides see identical operands; the error left is
accumulation order and the final bf16 rounding).
"""
import ctypes as C

import numpy as np
import pytest
import torch

from helpers import bf16_round, nchw_to_nhwc, nhwc_to_nchw, rel_l2, stream_ptr
from oracle import ref_ops

pytestmark = pytest.mark.gpu

TOL = {torch.float32: 1e-4, torch.bfloat16: 2e-2}


def _lib():
    from deep_vision_b200 import lib
    return lib


def _dv(dtype):
    lib = _lib()
    return lib.DV_F32 if dtype == torch.float32 else lib.DV_BF16


def _dev(a, dtype):
    return torch.from_numpy(np.ascontiguousarray(a)).to(device="cuda", dtype=dtype).contiguous()


def _rand(shape, seed, dtype, scale=1.0):
    g = np.random.default_rng(seed)
    a = g.standard_normal(shape) * scale
    return bf16_round(a) if dtype == torch.bfloat16 else a.astype(np.float32).astype(np.float64)


def run_conv(pass_id, x_nchw, w_oihw, dy_nchw, stride, pad, dtype, algo, use_wt=False):
    """Calls dv_conv_{fprop,dgrad,wgrad}; numpy NCHW/OIHW in, numpy NCHW/OIHW out (float64)."""
    lib = _lib()
    N, Cc, H, W = x_nchw.shape
    K, _, R, S = w_oihw.shape
    P, Q = (H + 2 * pad - R) // stride + 1, (W + 2 * pad - S) // stride + 1
    x = _dev(np.transpose(x_nchw, (0, 2, 3, 1)), dtype)
    w = _dev(np.transpose(w_oihw, (0, 2, 3, 1)), dtype)
    p = lib.ConvParams()
    p.N, p.H, p.W, p.C, p.K, p.R, p.S, p.stride, p.pad = N, H, W, Cc, K, R, S, stride, pad
    p.dtype, p.algo = _dv(dtype), algo
    keep = [x, w]
    if pass_id == lib.DV_CONV_FPROP:
        y = torch.full((N, P, Q, K), float("nan"), device="cuda", dtype=dtype)
        p.x, p.w, p.y = x.data_ptr(), w.data_ptr(), y.data_ptr()
        fn, out = lib.lib.dv_conv_fprop, y
    elif pass_id == lib.DV_CONV_DGRAD:
        dy = _dev(np.transpose(dy_nchw, (0, 2, 3, 1)), dtype)
        dx = torch.full((N, H, W, Cc), float("nan"), device="cuda", dtype=dtype)
        p.x, p.w, p.y = dx.data_ptr(), w.data_ptr(), dy.data_ptr()
        if dtype == torch.bfloat16 and use_wt:
            wf = _dev(np.transpose(w_oihw, (0, 2, 3, 1)), torch.float32)
            wt = torch.empty((Cc, R, S, K), device="cuda", dtype=dtype)
            lib.check(lib.lib.dv_filter_prep(wf.data_ptr(), None, wt.data_ptr(), K, R, S, Cc, _dv(dtype), stream_ptr()), "dv_filter_prep")
            p.w_t = wt.data_ptr()
            keep += [wf, wt]
        keep.append(dy)
        fn, out = lib.lib.dv_conv_dgrad, dx
    else:
        dy = _dev(np.transpose(dy_nchw, (0, 2, 3, 1)), dtype)
        dw = torch.full((K, R, S, Cc), float("nan"), device="cuda", dtype=torch.float32)
        p.x, p.w, p.y = x.data_ptr(), dw.data_ptr(), dy.data_ptr()
        keep.append(dy)
        fn, out = lib.lib.dv_conv_wgrad, dw
    need = lib.lib.dv_conv_workspace_bytes(C.byref(p), pass_id)
    if need:
        ws = torch.empty(need, dtype=torch.uint8, device="cuda")
        p.workspace, p.workspace_bytes = ws.data_ptr(), need
        keep.append(ws)
    lib.check(fn(C.byref(p), stream_ptr()), "conv")
    torch.cuda.synchronize()
    o = out.double().cpu().numpy()
    return np.transpose(o, (0, 3, 1, 2))


CONV_CASES = [
    # N, C, H, W, K, R, stride, pad
    (2, 3, 20, 20, 16, 7, 2, 3),      # ResNet stem shape class (C=3, 7x7/2)
    (2, 3, 16, 16, 32, 3, 2, 1),      # MobileNet stem (3x3/2)
    (2, 16, 9, 9, 24, 3, 1, 1),
    (3, 64, 7, 7, 64, 3, 1, 1),       # ragged M (147 rows)
    (2, 64, 8, 8, 128, 3, 2, 1),      # ResNet-34 3x3 stride 2
    (2, 128, 8, 8, 64, 1, 1, 0),
    (2, 64, 8, 8, 256, 1, 2, 0),      # strided 1x1 projection
    (1, 8, 5, 6, 8, 3, 1, 1),         # H != W
]


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16], ids=["fp32", "bf16"])
@pytest.mark.parametrize("case", CONV_CASES, ids=lambda c: "x".join(map(str, c)))
def test_conv_simt_matches_oracle(case, dtype):
    lib = _lib()
    N, Cc, H, W, K, R, stride, pad = case
    x = _rand((N, Cc, H, W), 1, dtype)
    w = _rand((K, Cc, R, R), 2, dtype, scale=0.2)
    y_ref = ref_ops.conv2d(x, w, stride, pad)
    dy = _rand(y_ref.shape, 3, dtype)
    dx_ref, dw_ref = ref_ops.conv2d_backward(x, w, dy, stride, pad)
    tol = TOL[dtype]
    assert rel_l2(run_conv(lib.DV_CONV_FPROP, x, w, None, stride, pad, dtype, lib.DV_ALGO_SIMT), y_ref) < tol
    assert rel_l2(run_conv(lib.DV_CONV_DGRAD, x, w, dy, stride, pad, dtype, lib.DV_ALGO_SIMT), dx_ref) < tol
    assert rel_l2(run_conv(lib.DV_CONV_WGRAD, x, w, dy, stride, pad, dtype, lib.DV_ALGO_SIMT), dw_ref) < 1e-4  # fp32 output


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16], ids=["fp32", "bf16"])
@pytest.mark.parametrize("shape,residual,relu", [((4, 64, 7, 7), False, True), ((2, 32, 9, 5), True, True),
                                                 ((3, 256, 4, 4), True, False), ((2, 2048, 2, 2), False, False),
                                                 ((64, 8, 16, 16), False, True)])
def test_bn_forward_backward_matches_oracle(shape, residual, relu, dtype):
    lib = _lib()
    N, Cc, H, W = shape
    M = N * H * W
    x = _rand(shape, 5, dtype, scale=2.0) + (0.0 if dtype == torch.bfloat16 else 3.0)  # fp32: large mean vs std
    if dtype == torch.bfloat16:
        x = bf16_round(x + 1.0)
    res = _rand(shape, 6, dtype) if residual else None
    gamma = np.random.default_rng(7).uniform(0.5, 1.5, Cc)
    beta = np.random.default_rng(8).normal(0, 0.3, Cc)
    rm0, rv0 = np.random.default_rng(9).normal(0, 1, Cc), np.random.default_rng(10).uniform(0.5, 2, Cc)
    y_ref, rm_ref, rv_ref, mean, invstd = ref_ops.batch_norm_train(x, gamma, beta, rm0, rv0)
    if residual:
        y_ref = y_ref + res
    pre = y_ref.copy()
    if relu:
        y_ref = ref_ops.relu(y_ref)

    xd = _dev(np.transpose(x, (0, 2, 3, 1)), dtype)
    rd = _dev(np.transpose(res, (0, 2, 3, 1)), dtype) if residual else None
    yd = torch.full((N, H, W, Cc), float("nan"), device="cuda", dtype=dtype)
    f32 = lambda a: torch.from_numpy(np.asarray(a, np.float32)).cuda()
    g_d, b_d, rm_d, rv_d = f32(gamma), f32(beta), f32(rm0), f32(rv0)
    sm, si = torch.empty(Cc, device="cuda"), torch.empty(Cc, device="cuda")
    wsz = lib.lib.dv_bn_workspace_bytes(M, Cc)
    ws = torch.empty(wsz, dtype=torch.uint8, device="cuda")
    p = lib.BnParams()
    p.x, p.y, p.residual = xd.data_ptr(), yd.data_ptr(), (rd.data_ptr() if residual else None)
    p.gamma, p.beta, p.running_mean, p.running_var = g_d.data_ptr(), b_d.data_ptr(), rm_d.data_ptr(), rv_d.data_ptr()
    p.save_mean, p.save_invstd, p.workspace, p.workspace_bytes = sm.data_ptr(), si.data_ptr(), ws.data_ptr(), wsz
    p.M, p.C, p.dtype, p.relu, p.eval_mode, p.eps, p.momentum = M, Cc, _dv(dtype), int(relu), 0, 1e-5, 0.1
    lib.check(lib.lib.dv_bn_stats(C.byref(p), stream_ptr()), "dv_bn_stats")
    lib.check(lib.lib.dv_bn_apply(C.byref(p), stream_ptr()), "dv_bn_apply")
    torch.cuda.synchronize()
    tol = TOL[dtype]
    assert rel_l2(sm.cpu().numpy(), mean) < 1e-5
    assert rel_l2(si.cpu().numpy(), invstd) < 1e-4
    assert rel_l2(rm_d.cpu().numpy(), rm_ref) < 1e-5 and rel_l2(rv_d.cpu().numpy(), rv_ref) < 1e-4
    y_gpu = np.transpose(yd.double().cpu().numpy(), (0, 3, 1, 2))
    assert rel_l2(y_gpu, y_ref) < tol

    # backward: the mask comes from the GPU's own stored output (what the kernel sees), statistics from the oracle
    dy = _rand(shape, 11, dtype)
    mask = (y_gpu > 0) if relu else np.ones_like(y_gpu, bool)
    dz = np.where(mask, dy, 0.0)
    dx_ref, dg_ref, db_ref = ref_ops.batch_norm_backward(x, gamma, mean, invstd, dz)
    dyd = _dev(np.transpose(dy, (0, 2, 3, 1)), dtype)
    dxd = torch.full((N, H, W, Cc), float("nan"), device="cuda", dtype=dtype)
    drd = torch.full((N, H, W, Cc), float("nan"), device="cuda", dtype=dtype) if residual else None
    dg, db = torch.empty(Cc, device="cuda"), torch.empty(Cc, device="cuda")
    q = lib.BnBwdParams()
    q.x, q.y, q.dy, q.dx = xd.data_ptr(), yd.data_ptr(), dyd.data_ptr(), dxd.data_ptr()
    q.dres = drd.data_ptr() if residual else None
    q.gamma, q.save_mean, q.save_invstd, q.dgamma, q.dbeta = g_d.data_ptr(), sm.data_ptr(), si.data_ptr(), dg.data_ptr(), db.data_ptr()
    q.workspace, q.workspace_bytes, q.M, q.C, q.dtype, q.relu = ws.data_ptr(), wsz, M, Cc, _dv(dtype), int(relu)
    lib.check(lib.lib.dv_bn_bwd_reduce(C.byref(q), stream_ptr()), "dv_bn_bwd_reduce")
    lib.check(lib.lib.dv_bn_bwd_apply(C.byref(q), stream_ptr()), "dv_bn_bwd_apply")
    torch.cuda.synchronize()
    assert rel_l2(dg.cpu().numpy(), dg_ref) < 1e-4 and rel_l2(db.cpu().numpy(), db_ref) < 1e-4
    assert rel_l2(np.transpose(dxd.double().cpu().numpy(), (0, 3, 1, 2)), dx_ref) < tol
    if residual:
        assert rel_l2(np.transpose(drd.double().cpu().numpy(), (0, 3, 1, 2)), dz) < tol
    if relu and not residual:
        # mask recomputed from x instead of read from y: must reproduce the stored-mask results bit for bit
        dx1, dg1, db1 = dxd.clone(), dg.clone(), db.clone()
        dxd.fill_(float("nan"))
        q.y, q.beta = None, b_d.data_ptr()
        lib.check(lib.lib.dv_bn_bwd_reduce(C.byref(q), stream_ptr()), "dv_bn_bwd_reduce(recompute)")
        lib.check(lib.lib.dv_bn_bwd_apply(C.byref(q), stream_ptr()), "dv_bn_bwd_apply(recompute)")
        torch.cuda.synchronize()
        assert torch.equal(dxd, dx1) and torch.equal(dg, dg1) and torch.equal(db, db1)
        q.y = yd.data_ptr()
    if relu and dtype == torch.bfloat16:
        # one-bit-per-element mask written by the forward: equals (y > 0) and reproduces the stored-mask backward bit for bit
        mk = torch.full((M, Cc // 8), 0xAA, device="cuda", dtype=torch.uint8)
        p.relu_mask = mk.data_ptr()
        lib.check(lib.lib.dv_bn_apply(C.byref(p), stream_ptr()), "dv_bn_apply(mask)")
        torch.cuda.synchronize()
        bits = np.unpackbits(mk.cpu().numpy().reshape(N, H, W, Cc // 8, 1), axis=-1, bitorder="little").reshape(N, H, W, Cc)
        assert np.array_equal(bits.astype(bool), yd.float().cpu().numpy() > 0)
        dx1, dg1, db1 = dxd.clone(), dg.clone(), db.clone()
        dr1 = drd.clone() if residual else None
        dxd.fill_(float("nan"))
        q.y, q.relu_mask = None, mk.data_ptr()
        lib.check(lib.lib.dv_bn_bwd_reduce(C.byref(q), stream_ptr()), "dv_bn_bwd_reduce(bit mask)")
        lib.check(lib.lib.dv_bn_bwd_apply(C.byref(q), stream_ptr()), "dv_bn_bwd_apply(bit mask)")
        torch.cuda.synchronize()
        assert torch.equal(dxd, dx1) and torch.equal(dg, dg1) and torch.equal(db, db1)
        if residual:
            assert torch.equal(drd, dr1)
        q.y, q.relu_mask, p.relu_mask = yd.data_ptr(), None, None

    # eval mode uses the running statistics
    p.eval_mode = 1
    lib.check(lib.lib.dv_bn_apply(C.byref(p), stream_ptr()), "dv_bn_apply(eval)")
    torch.cuda.synchronize()
    ye = ref_ops.batch_norm_eval(x, gamma, beta, rm_d.double().cpu().numpy(), rv_d.double().cpu().numpy())
    if residual:
        ye = ye + res
    if relu:
        ye = ref_ops.relu(ye)
    assert rel_l2(np.transpose(yd.double().cpu().numpy(), (0, 3, 1, 2)), ye) < tol


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16], ids=["fp32", "bf16"])
@pytest.mark.parametrize("case", [(2, 32, 10, 10, 1), (2, 64, 9, 11, 2), (1, 1024, 7, 7, 1), (3, 128, 8, 8, 2), (2, 2048, 4, 4, 1),
                                  # W >= 14: the TMA-staged kernels (halo tiles through shared memory); ragged tiles in both directions
                                  (2, 64, 20, 30, 1), (2, 32, 17, 14, 1), (1, 128, 33, 16, 1), (2, 96, 12, 20, 1), (3, 256, 28, 28, 1),
                                  (2, 64, 20, 30, 2), (2, 32, 17, 15, 2), (1, 128, 28, 28, 2)],
                         ids=lambda c: "x".join(map(str, c)))
def test_dwconv_matches_oracle(case, dtype):
    lib = _lib()
    N, Cc, H, W, stride = case
    x = _rand((N, Cc, H, W), 21, dtype)
    w = np.random.default_rng(22).standard_normal((Cc, 1, 3, 3)).astype(np.float32).astype(np.float64) * 0.3
    y_ref = ref_ops.conv2d(x, w, stride, 1, groups=Cc)
    dy = _rand(y_ref.shape, 23, dtype)
    dx_ref, dw_ref = ref_ops.conv2d_backward(x, w, dy, stride, 1, groups=Cc)
    P, Q = y_ref.shape[2:]
    xd = _dev(np.transpose(x, (0, 2, 3, 1)), dtype)
    wd = torch.from_numpy(w.astype(np.float32)).cuda().contiguous()
    yd = torch.full((N, P, Q, Cc), float("nan"), device="cuda", dtype=dtype)
    p = lib.DwConvParams()
    p.x, p.w, p.y, p.N, p.H, p.W, p.C, p.stride, p.dtype = xd.data_ptr(), wd.data_ptr(), yd.data_ptr(), N, H, W, Cc, stride, _dv(dtype)
    fused = lib.lib.dv_dwconv_fuses_stats(C.byref(p)) == 1      # TMA-staged stride-1 kernel: BatchNorm sums from its registers
    sums = torch.full((2 * Cc,), 1.5, device="cuda")
    if fused:
        p.stat_sums = sums.data_ptr()
    lib.check(lib.lib.dv_dwconv_fprop(C.byref(p), stream_ptr()), "dv_dwconv_fprop")
    torch.cuda.synchronize()
    tol = TOL[dtype]
    assert rel_l2(np.transpose(yd.double().cpu().numpy(), (0, 3, 1, 2)), y_ref) < tol
    if fused:
        yb = yd.double().reshape(-1, Cc)
        assert rel_l2((sums[:Cc].double() - 1.5).cpu().numpy(), yb.sum(0).cpu().numpy()) < 1e-4       # ADDED to what was there
        assert rel_l2((sums[Cc:].double() - 1.5).cpu().numpy(), (yb * yb).sum(0).cpu().numpy()) < 1e-4
    else:
        p.stat_sums = sums.data_ptr()
        assert lib.lib.dv_dwconv_fprop(C.byref(p), stream_ptr()) == lib.DV_ERR_UNSUPPORTED                 # never silently ignored
    p.stat_sums = None
    dyd = _dev(np.transpose(dy, (0, 2, 3, 1)), dtype)
    dxd = torch.full((N, H, W, Cc), float("nan"), device="cuda", dtype=dtype)
    p.x, p.y = dxd.data_ptr(), dyd.data_ptr()
    lib.check(lib.lib.dv_dwconv_dgrad(C.byref(p), stream_ptr()), "dv_dwconv_dgrad")
    dwd = torch.full((Cc, 1, 3, 3), float("nan"), device="cuda")
    p.x, p.w = xd.data_ptr(), dwd.data_ptr()
    need = lib.lib.dv_dwconv_workspace_bytes(C.byref(p))
    ws = torch.empty(need, dtype=torch.uint8, device="cuda")
    p.workspace, p.workspace_bytes = ws.data_ptr(), need
    lib.check(lib.lib.dv_dwconv_wgrad(C.byref(p), stream_ptr()), "dv_dwconv_wgrad")
    torch.cuda.synchronize()
    assert rel_l2(np.transpose(dxd.double().cpu().numpy(), (0, 3, 1, 2)), dx_ref) < tol
    assert rel_l2(dwd.double().cpu().numpy(), dw_ref) < 1e-4


@pytest.mark.parametrize("hw", [(12, 8), (7, 7), (1, 5), (2, 2)], ids=lambda v: f"{v[0]}x{v[1]}")
def test_maxpool_even_odd_and_tiny_extents(hw):
    """The kernels work on pairs of outputs (forward) and 2x2 input cells (backward): cover even, odd and 1-wide extents."""
    lib = _lib()
    dtype = torch.bfloat16
    N, Cc = 2, 16
    H, W = hw
    x = ref_ops.relu(_rand((N, Cc, H, W), 41, dtype))
    y_ref, tap = ref_ops.max_pool_3x3_s2_p1(x)
    P, Q = y_ref.shape[2:]
    xd = _dev(np.transpose(x, (0, 2, 3, 1)), dtype)
    yd = torch.full((N, P, Q, Cc), float("nan"), device="cuda", dtype=dtype)
    idx = torch.full((N, P, Q, Cc), 99, device="cuda", dtype=torch.uint8)
    lib.check(lib.lib.dv_maxpool_fwd(xd.data_ptr(), yd.data_ptr(), idx.data_ptr(), N, H, W, Cc, _dv(dtype), stream_ptr()), "maxpool_fwd")
    dy = _rand(y_ref.shape, 42, dtype)
    dyd = _dev(np.transpose(dy, (0, 2, 3, 1)), dtype)
    dxd = torch.full((N, H, W, Cc), float("nan"), device="cuda", dtype=dtype)
    lib.check(lib.lib.dv_maxpool_bwd(dyd.data_ptr(), idx.data_ptr(), dxd.data_ptr(), N, H, W, Cc, _dv(dtype), stream_ptr()), "maxpool_bwd")
    torch.cuda.synchronize()
    assert np.array_equal(np.transpose(yd.double().cpu().numpy(), (0, 3, 1, 2)), y_ref)
    assert np.array_equal(np.transpose(idx.cpu().numpy(), (0, 3, 1, 2)), tap)
    assert rel_l2(np.transpose(dxd.double().cpu().numpy(), (0, 3, 1, 2)), ref_ops.max_pool_backward(dy, tap, H, W)) < TOL[dtype]


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16], ids=["fp32", "bf16"])
def test_pools_layout_cast_match_oracle(dtype):
    lib = _lib()
    N, Cc, H, W = 3, 64, 11, 9
    x = ref_ops.relu(_rand((N, Cc, H, W), 31, dtype))  # post-ReLU input: many exact ties at 0
    y_ref, tap = ref_ops.max_pool_3x3_s2_p1(x)
    P, Q = y_ref.shape[2:]
    xd = _dev(np.transpose(x, (0, 2, 3, 1)), dtype)
    yd = torch.empty((N, P, Q, Cc), device="cuda", dtype=dtype)
    idx = torch.empty((N, P, Q, Cc), device="cuda", dtype=torch.uint8)
    lib.check(lib.lib.dv_maxpool_fwd(xd.data_ptr(), yd.data_ptr(), idx.data_ptr(), N, H, W, Cc, _dv(dtype), stream_ptr()), "maxpool_fwd")
    dy = _rand(y_ref.shape, 32, dtype)
    dyd = _dev(np.transpose(dy, (0, 2, 3, 1)), dtype)
    dxd = torch.empty((N, H, W, Cc), device="cuda", dtype=dtype)
    lib.check(lib.lib.dv_maxpool_bwd(dyd.data_ptr(), idx.data_ptr(), dxd.data_ptr(), N, H, W, Cc, _dv(dtype), stream_ptr()), "maxpool_bwd")
    torch.cuda.synchronize()
    assert np.array_equal(np.transpose(yd.double().cpu().numpy(), (0, 3, 1, 2)), y_ref)          # exact: a max of inputs
    assert np.array_equal(np.transpose(idx.cpu().numpy(), (0, 3, 1, 2)), tap)                     # same tie-breaking
    assert rel_l2(np.transpose(dxd.double().cpu().numpy(), (0, 3, 1, 2)), ref_ops.max_pool_backward(dy, tap, H, W)) < TOL[dtype]

    ad = torch.empty((N, Cc), device="cuda", dtype=dtype)
    lib.check(lib.lib.dv_avgpool_fwd(xd.data_ptr(), ad.data_ptr(), N, H * W, Cc, _dv(dtype), stream_ptr()), "avgpool_fwd")
    g = _rand((N, Cc), 33, dtype)
    gd = _dev(g, dtype)
    gx = torch.empty((N, H, W, Cc), device="cuda", dtype=dtype)
    lib.check(lib.lib.dv_avgpool_bwd(gd.data_ptr(), gx.data_ptr(), N, H * W, Cc, _dv(dtype), stream_ptr()), "avgpool_bwd")
    torch.cuda.synchronize()
    assert rel_l2(ad.double().cpu().numpy(), ref_ops.global_avg_pool(x)[:, :, 0, 0]) < TOL[dtype]
    assert rel_l2(np.transpose(gx.double().cpu().numpy(), (0, 3, 1, 2)), np.broadcast_to(g[:, :, None, None] / (H * W), x.shape)) < TOL[dtype]

    img = np.random.default_rng(34).standard_normal((2, 3, 10, 12)).astype(np.float32)
    imd = torch.from_numpy(img).cuda()
    out = torch.empty((2, 10, 12, 3), device="cuda", dtype=dtype)
    lib.check(lib.lib.dv_nchw_to_nhwc(imd.data_ptr(), out.data_ptr(), 2, 3, 10, 12, _dv(dtype), stream_ptr()), "nchw_to_nhwc")
    torch.cuda.synchronize()
    expect = np.transpose(img, (0, 2, 3, 1)).astype(np.float64)
    if dtype == torch.bfloat16:
        expect = bf16_round(expect)
    assert np.array_equal(out.double().cpu().numpy(), expect)


def test_optimisers_and_bias_match_oracle():
    lib = _lib()
    rng = np.random.default_rng(41)
    n = 10007
    p0, g1, g2 = (rng.standard_normal(n).astype(np.float32) for _ in range(3))
    pd, buf = torch.from_numpy(p0.copy()).cuda(), torch.zeros(n, device="cuda")
    pn, bn = p0.astype(np.float64), None
    for step, g in enumerate((g1, g2)):
        gd = torch.from_numpy(g).cuda()
        lib.check(lib.lib.dv_sgd_step(pd.data_ptr(), gd.data_ptr(), buf.data_ptr(), n, 0.1, 0.9, 1e-4, 1.0, int(step == 0), stream_ptr()), "sgd")
        pn, bn = ref_ops.sgd_step(pn, g.astype(np.float64), bn, 0.1, 0.9, 1e-4, first=(step == 0))
    torch.cuda.synchronize()
    assert rel_l2(pd.cpu().numpy(), pn) < 1e-6
    pd, sq = torch.from_numpy(p0.copy()).cuda(), torch.zeros(n, device="cuda")
    pn, sn = p0.astype(np.float64), np.zeros(n)
    for g in (g1, g2):
        gd = torch.from_numpy(g).cuda()
        lib.check(lib.lib.dv_rmsprop_step(pd.data_ptr(), gd.data_ptr(), sq.data_ptr(), n, 0.045, 0.9, 1.0, 0.0, 1.0, stream_ptr()), "rmsprop")
        pn, sn = ref_ops.rmsprop_step(pn, g.astype(np.float64), sn, 0.045, 0.9, 1.0)
    torch.cuda.synchronize()
    assert rel_l2(pd.cpu().numpy(), pn) < 1e-6
    y = rng.standard_normal((5, 1000)).astype(np.float32)
    b = rng.standard_normal(1000).astype(np.float32)
    yd, bd = torch.from_numpy(y.copy()).cuda(), torch.from_numpy(b).cuda()
    lib.check(lib.lib.dv_bias_add(yd.data_ptr(), bd.data_ptr(), 5, 1000, stream_ptr()), "bias_add")
    db = torch.empty(1000, device="cuda")
    lib.check(lib.lib.dv_bias_grad(yd.data_ptr(), db.data_ptr(), 5, 1000, stream_ptr()), "bias_grad")
    torch.cuda.synchronize()
    assert rel_l2(yd.cpu().numpy(), y + b) < 1e-7
    assert rel_l2(db.cpu().numpy(), (y + b).sum(0)) < 1e-5


@pytest.mark.parametrize("case", [(2, 3, 20, 20, 7, 2, 3), (2, 3, 16, 16, 3, 2, 1), (1, 8, 6, 7, 3, 1, 1)], ids=lambda c: "x".join(map(str, c)))
def test_im2col_pad_unpad_match_oracle(case):
    """dv_im2col builds exactly the patch matrix whose product with the padded filter rows is the convolution."""
    lib = _lib()
    N, Cc, H, W, R, stride, pad = case
    dtype = torch.bfloat16
    x = _rand((N, Cc, H, W), 51, dtype)
    K = 16
    w = _rand((K, Cc, R, R), 52, dtype, scale=0.2)
    P, Q = (H + 2 * pad - R) // stride + 1, (W + 2 * pad - R) // stride + 1
    kp = -(-(R * R * Cc) // 64) * 64
    xd = _dev(np.transpose(x, (0, 2, 3, 1)), dtype)
    pat = torch.full((N * P * Q, kp), float("nan"), device="cuda", dtype=dtype)
    lib.check(lib.lib.dv_im2col(xd.data_ptr(), pat.data_ptr(), N, H, W, Cc, R, R, stride, pad, kp, _dv(dtype), stream_ptr()), "im2col")
    wf = torch.from_numpy(np.ascontiguousarray(np.transpose(w, (0, 2, 3, 1))).astype(np.float32)).cuda()
    wp = torch.full((K, kp), float("nan"), device="cuda", dtype=dtype)
    lib.check(lib.lib.dv_pad_rows(wf.data_ptr(), wp.data_ptr(), K, R * R * Cc, kp, _dv(dtype), stream_ptr()), "pad_rows")
    torch.cuda.synchronize()
    y = (pat.double() @ wp.double().t()).cpu().numpy().reshape(N, P, Q, K).transpose(0, 3, 1, 2)
    assert rel_l2(y, ref_ops.conv2d(x, w, stride, pad)) < 1e-12       # exact operands, fp64 product
    assert float(pat[:, R * R * Cc:].abs().sum()) == 0.0 and float(wp[:, R * R * Cc:].abs().sum()) == 0.0
    src = torch.randn(K, kp, device="cuda")
    dst = torch.empty(K, R * R * Cc, device="cuda")
    lib.check(lib.lib.dv_unpad_rows(src.data_ptr(), dst.data_ptr(), K, R * R * Cc, kp, stream_ptr()), "unpad_rows")
    torch.cuda.synchronize()
    assert torch.equal(dst, src[:, :R * R * Cc])


def test_cross_entropy_matches_oracle():
    """dv_cross_entropy: mean loss and d(loss)/d(logits) in one launch (nn.CrossEntropyLoss, train.py:358,452)."""
    lib = _lib()
    rng = np.random.default_rng(61)
    B, Cn = 37, 1000
    logits = (rng.standard_normal((B, Cn)) * 4).astype(np.float32)
    labels = rng.integers(0, Cn, B)
    loss_ref, d_ref = ref_ops.cross_entropy(logits.astype(np.float64), labels)
    ld = torch.from_numpy(logits).cuda()
    lab = torch.from_numpy(labels.astype(np.int64)).cuda()
    loss = torch.full((), float("nan"), device="cuda")
    d = torch.full((B, Cn), float("nan"), device="cuda")
    lib.check(lib.lib.dv_cross_entropy(ld.data_ptr(), lab.data_ptr(), loss.data_ptr(), d.data_ptr(), B, Cn, stream_ptr()), "ce")
    torch.cuda.synchronize()
    assert abs(float(loss) - loss_ref) < 1e-5 * abs(loss_ref)
    assert rel_l2(d.cpu().numpy(), d_ref) < 1e-5
    s = torch.tensor([0.25], device="cuda")
    lib.check(lib.lib.dv_scale_by(d.data_ptr(), s.data_ptr(), B * Cn, stream_ptr()), "scale")
    torch.cuda.synchronize()
    assert rel_l2(d.cpu().numpy(), 0.25 * d_ref) < 1e-5


@pytest.mark.parametrize("hw", [(9, 10), (18, 21)], ids=["strip_kernel", "tma_kernel"])
@pytest.mark.parametrize("stride", [1, 2])
def test_dwconv_folded_batchnorm_epilogue(stride, hw):
    """dv_dwconv_params.ep_shift / ep_relu with taps scaled by dv_bn_fold == oracle dwconv -> eval BN -> ReLU."""
    lib = _lib()
    dtype = torch.bfloat16
    N, Cc, (H, W) = 3, 64, hw
    x = _rand((N, Cc, H, W), 51, dtype)
    w = _rand((Cc, 1, 3, 3), 52, torch.float32, scale=0.3)
    rng = np.random.default_rng(53)
    gamma, beta, rm, rv = rng.uniform(0.5, 1.5, Cc), rng.normal(0, 0.3, Cc), rng.normal(0, 0.5, Cc), rng.uniform(0.5, 2.0, Cc)
    ref = ref_ops.relu(ref_ops.batch_norm_eval(ref_ops.conv2d(x, w, stride, 1, groups=Cc), gamma, beta, rm, rv))
    f32 = lambda a: torch.from_numpy(np.ascontiguousarray(np.asarray(a, np.float32))).cuda()
    g_d, b_d, rm_d, rv_d, w_d = f32(gamma), f32(beta), f32(rm), f32(rv), f32(w)
    scale, shift, w_s = torch.empty(Cc, device="cuda"), torch.empty(Cc, device="cuda"), torch.empty_like(w_d)
    lib.check(lib.lib.dv_bn_fold(g_d.data_ptr(), b_d.data_ptr(), rm_d.data_ptr(), rv_d.data_ptr(), 1e-5, scale.data_ptr(), shift.data_ptr(),
                                 Cc, w_d.data_ptr(), w_s.data_ptr(), stream_ptr()), "dv_bn_fold")
    xd = _dev(np.transpose(x, (0, 2, 3, 1)), dtype)
    P, Q = ref.shape[2:]
    yd = torch.full((N, P, Q, Cc), float("nan"), device="cuda", dtype=dtype)
    p = lib.DwConvParams()
    p.x, p.w, p.y = xd.data_ptr(), w_s.data_ptr(), yd.data_ptr()
    p.N, p.H, p.W, p.C, p.stride, p.dtype = N, H, W, Cc, stride, _dv(dtype)
    p.ep_shift, p.ep_relu = shift.data_ptr(), 1
    lib.check(lib.lib.dv_dwconv_fprop(C.byref(p), stream_ptr()), "dv_dwconv_fprop(folded)")
    torch.cuda.synchronize()
    assert rel_l2(w_s.cpu().numpy(), w * (gamma / np.sqrt(rv + 1e-5))[:, None, None, None]) < 1e-6
    assert rel_l2(np.transpose(yd.double().cpu().numpy(), (0, 3, 1, 2)), ref) < 1e-2
    assert lib.lib.dv_dwconv_dgrad(C.byref(p), stream_ptr()) == lib.DV_ERR_UNSUPPORTED


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16], ids=["fp32", "bf16"])
@pytest.mark.parametrize("shape", [(2, 64, 12, 16), (3, 64, 7, 9), (1, 32, 5, 2), (2, 128, 1, 6)], ids=lambda s: "x".join(map(str, s)))
def test_bn_relu_maxpool_fused_equals_the_three_separate_ops(shape, dtype):
    """dv_bn_relu_maxpool_{fwd,bwd} (the ResNet stem's bn1 -> relu -> maxpool, resnet50.py:49-52): y, the argmax and dx bit for
    bit what dv_bn_apply(relu) -> dv_maxpool_fwd and dv_maxpool_bwd -> dv_bn_bwd_reduce -> dv_bn_bwd_apply give; dgamma/dbeta to
    summation order; and the separate ops are the ones checked against the oracle above."""
    lib = _lib()
    N, Cc, H, W = shape
    M = N * H * W
    P, Q = (H - 1) // 2 + 1, (W - 1) // 2 + 1
    x = _rand((N, H, W, Cc), 31, dtype, scale=1.5)
    if dtype == torch.bfloat16:
        x = np.round(x * 4) / 4          # coarse values: ties inside pooling windows, so the argmax rule is exercised
    xd = _dev(x, dtype)
    f32 = lambda a: torch.from_numpy(np.asarray(a, np.float32)).cuda()
    g_d = f32(np.random.default_rng(32).uniform(-1.5, 1.5, Cc))     # negative gammas flip the order inside a window
    b_d = f32(np.random.default_rng(33).normal(0, 0.3, Cc))
    sm, si = torch.empty(Cc, device="cuda"), torch.empty(Cc, device="cuda")
    wsz = lib.lib.dv_bn_workspace_bytes(M, Cc)
    ws = torch.empty(wsz, dtype=torch.uint8, device="cuda")
    yd = torch.full((N, H, W, Cc), float("nan"), device="cuda", dtype=dtype)
    p = lib.BnParams()
    p.x, p.y, p.gamma, p.beta = xd.data_ptr(), yd.data_ptr(), g_d.data_ptr(), b_d.data_ptr()
    p.save_mean, p.save_invstd, p.workspace, p.workspace_bytes = sm.data_ptr(), si.data_ptr(), ws.data_ptr(), wsz
    p.M, p.C, p.dtype, p.relu, p.eval_mode, p.eps, p.momentum = M, Cc, _dv(dtype), 1, 0, 1e-5, 0.1
    lib.check(lib.lib.dv_bn_stats(C.byref(p), stream_ptr()), "dv_bn_stats")
    lib.check(lib.lib.dv_bn_apply(C.byref(p), stream_ptr()), "dv_bn_apply")
    pool = torch.full((N, P, Q, Cc), float("nan"), device="cuda", dtype=dtype)
    idx = torch.full((N, P, Q, Cc), 0xEE, device="cuda", dtype=torch.uint8)
    lib.check(lib.lib.dv_maxpool_fwd(yd.data_ptr(), pool.data_ptr(), idx.data_ptr(), N, H, W, Cc, _dv(dtype), stream_ptr()), "maxpool_fwd")

    pool_f = torch.full_like(pool, float("nan"))
    idx_f = torch.full_like(idx, 0xEE)
    q = lib.BnPoolParams()
    q.x, q.y, q.idx = xd.data_ptr(), pool_f.data_ptr(), idx_f.data_ptr()
    q.gamma, q.beta, q.save_mean, q.save_invstd = g_d.data_ptr(), b_d.data_ptr(), sm.data_ptr(), si.data_ptr()
    q.N, q.H, q.W, q.C, q.dtype = N, H, W, Cc, _dv(dtype)
    lib.check(lib.lib.dv_bn_relu_maxpool_fwd(C.byref(q), stream_ptr()), "dv_bn_relu_maxpool_fwd")
    torch.cuda.synchronize()
    assert torch.equal(pool_f, pool)
    assert torch.equal(idx_f, idx)

    dpool = _dev(_rand((N, P, Q, Cc), 34, dtype), dtype)
    dyd = torch.full((N, H, W, Cc), float("nan"), device="cuda", dtype=dtype)
    lib.check(lib.lib.dv_maxpool_bwd(dpool.data_ptr(), idx.data_ptr(), dyd.data_ptr(), N, H, W, Cc, _dv(dtype), stream_ptr()), "maxpool_bwd")
    dxd = torch.full((N, H, W, Cc), float("nan"), device="cuda", dtype=dtype)
    dg, db = torch.empty(Cc, device="cuda"), torch.empty(Cc, device="cuda")
    r = lib.BnBwdParams()
    r.x, r.y, r.dy, r.dx = xd.data_ptr(), yd.data_ptr(), dyd.data_ptr(), dxd.data_ptr()
    r.gamma, r.save_mean, r.save_invstd, r.dgamma, r.dbeta = g_d.data_ptr(), sm.data_ptr(), si.data_ptr(), dg.data_ptr(), db.data_ptr()
    r.workspace, r.workspace_bytes, r.M, r.C, r.dtype, r.relu = ws.data_ptr(), wsz, M, Cc, _dv(dtype), 1
    lib.check(lib.lib.dv_bn_bwd_reduce(C.byref(r), stream_ptr()), "dv_bn_bwd_reduce")
    lib.check(lib.lib.dv_bn_bwd_apply(C.byref(r), stream_ptr()), "dv_bn_bwd_apply")

    dx_f = torch.full_like(dxd, float("nan"))
    dg_f, db_f = torch.full_like(dg, float("nan")), torch.full_like(db, float("nan"))
    q.dy, q.dx, q.dgamma, q.dbeta = dpool.data_ptr(), dx_f.data_ptr(), dg_f.data_ptr(), db_f.data_ptr()
    wsz2 = lib.lib.dv_bn_pool_workspace_bytes(C.byref(q))
    ws2 = torch.empty(max(wsz2, 1), dtype=torch.uint8, device="cuda")
    q.workspace, q.workspace_bytes = ws2.data_ptr(), wsz2
    lib.check(lib.lib.dv_bn_relu_maxpool_bwd(C.byref(q), stream_ptr()), "dv_bn_relu_maxpool_bwd")
    torch.cuda.synchronize()
    assert rel_l2(dg_f.cpu().numpy(), dg.cpu().numpy()) < 1e-5 and rel_l2(db_f.cpu().numpy(), db.cpu().numpy()) < 1e-5
    if torch.equal(dg_f, dg) and torch.equal(db_f, db):
        assert torch.equal(dx_f, dxd)
    else:   # sums differ in the last bits by summation order; dx then differs by roundings of those
        assert rel_l2(dx_f.double().cpu().numpy(), dxd.double().cpu().numpy()) < (1e-6 if dtype == torch.float32 else 2e-3)
    # a workspace that is too small is refused
    q.workspace_bytes = max(wsz2 - 1, 0)
    if wsz2:
        assert lib.lib.dv_bn_relu_maxpool_bwd(C.byref(q), stream_ptr()) != 0
    q.C = 24   # not a channel count the kernels take
    assert lib.lib.dv_bn_relu_maxpool_fwd(C.byref(q), stream_ptr()) != 0


@pytest.mark.parametrize("case", [(2, 64, 18, 21), (3, 32, 14, 14), (5, 128, 9, 30), (40, 512, 14, 14)], ids=lambda c: "x".join(map(str, c)))
def test_dwconv_dgrad_fused_bn_backward_sums(case):
    """dv_dwconv_params.bnb_*: the stride-1 depth-wise dgrad adds dbeta = sum dz and dgamma = sum dz*xhat of the BatchNorm that
    produced the convolution's input (dz = dx masked by bn(x0) > 0).  dx is unchanged by the option (bit for bit); the sums equal
    fp64 sums over the kernel's own bf16 dx and dv_bn_bwd_reduce's on the same tensors; they are ADDED to what was there; paths
    that cannot honour the option reject it."""
    lib = _lib()
    N, Cc, H, W = case
    g = torch.Generator(device="cuda").manual_seed(17)
    dy = torch.randn((N, H, W, Cc), generator=g, device="cuda").bfloat16()
    w = (torch.randn((Cc, 1, 3, 3), generator=g, device="cuda") * 0.3).contiguous()
    x0 = (torch.randn((N, H, W, Cc), generator=g, device="cuda") * 1.5 + 0.3).bfloat16()
    gamma = torch.rand(Cc, generator=g, device="cuda") + 0.5
    beta = torch.randn(Cc, generator=g, device="cuda") * 0.3
    xf = x0.float().reshape(-1, Cc)
    mean = xf.mean(0)
    invstd = 1.0 / torch.sqrt(xf.var(0, unbiased=False) + 1e-5)

    def dgrad(with_sums):
        dx = torch.full((N, H, W, Cc), float("nan"), device="cuda", dtype=torch.bfloat16)
        dgamma, dbeta = torch.full((Cc,), 2.0, device="cuda"), torch.full((Cc,), -3.0, device="cuda")
        p = lib.DwConvParams()
        p.x, p.w, p.y, p.N, p.H, p.W, p.C, p.stride, p.dtype = dx.data_ptr(), w.data_ptr(), dy.data_ptr(), N, H, W, Cc, 1, lib.DV_BF16
        assert lib.lib.dv_dwconv_fuses_bn_bwd(C.byref(p)) == 1
        if with_sums:
            p.bnb_x, p.bnb_gamma, p.bnb_beta, p.bnb_mean, p.bnb_invstd = x0.data_ptr(), gamma.data_ptr(), beta.data_ptr(), mean.data_ptr(), invstd.data_ptr()
            p.bnb_dgamma, p.bnb_dbeta = dgamma.data_ptr(), dbeta.data_ptr()
        lib.check(lib.lib.dv_dwconv_dgrad(C.byref(p), stream_ptr()), "dv_dwconv_dgrad")
        torch.cuda.synchronize()
        return dx, dgamma, dbeta, p

    dx_plain, _, _, _ = dgrad(False)
    dx, dgamma, dbeta, p = dgrad(True)
    assert torch.equal(dx.view(torch.int16), dx_plain.view(torch.int16))
    sc = (invstd * gamma).double()
    pre = x0.double().reshape(-1, Cc) * sc + (beta.double() - mean.double() * sc)
    dz = dx.double().reshape(-1, Cc) * (pre > 0)
    xhat = (x0.double().reshape(-1, Cc) - mean.double()) * invstd.double()
    want_b, want_g = dz.sum(0), (dz * xhat).sum(0)
    scale_b, scale_g = dz.abs().sum(0) + 1e-6, (dz * xhat).abs().sum(0) + 1e-6
    assert float(((dbeta.double() + 3.0 - want_b).abs() / scale_b).max()) < 1e-4
    assert float(((dgamma.double() - 2.0 - want_g).abs() / scale_g).max()) < 1e-4
    q = lib.BnBwdParams()
    dg2, db2 = torch.empty(Cc, device="cuda"), torch.empty(Cc, device="cuda")
    M = N * H * W
    ws = torch.empty(lib.lib.dv_bn_workspace_bytes(M, Cc), dtype=torch.uint8, device="cuda")
    q.x, q.dy, q.gamma, q.beta, q.save_mean, q.save_invstd = x0.data_ptr(), dx.data_ptr(), gamma.data_ptr(), beta.data_ptr(), mean.data_ptr(), invstd.data_ptr()
    q.dgamma, q.dbeta, q.workspace, q.workspace_bytes = dg2.data_ptr(), db2.data_ptr(), ws.data_ptr(), ws.numel()
    q.M, q.C, q.dtype, q.relu = M, Cc, lib.DV_BF16, 1
    lib.check(lib.lib.dv_bn_bwd_reduce(C.byref(q), stream_ptr()), "bwd_reduce")
    torch.cuda.synchronize()
    assert float(((dbeta + 3.0 - db2).double().abs() / scale_b).max()) < 1e-4
    assert float(((dgamma - 2.0 - dg2).double().abs() / scale_g).max()) < 1e-4
    # not honoured elsewhere: forward, weight gradient, stride 2
    assert lib.lib.dv_dwconv_fprop(C.byref(p), stream_ptr()) == lib.DV_ERR_UNSUPPORTED
    p.stride = 2
    assert lib.lib.dv_dwconv_fuses_bn_bwd(C.byref(p)) == 0
