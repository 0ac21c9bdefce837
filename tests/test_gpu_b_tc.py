"""GPU parity tests of the tcgen05/TMEM/TMA convolution kernels (DV_ALGO_TC, bf16) through the C ABI:
against the CPU oracle at sizes numpy finishes in seconds, and against the independent CUDA-core kernel plus
size-independent properties (linearity, zero-padding border rows) at BASELINE.json's full sizes."""
import numpy as np
import pytest
import torch

from helpers import bf16_round, rel_l2
from oracle import ref_ops
from test_gpu_a_ops import _lib, _rand, run_conv

pytestmark = pytest.mark.gpu
BF = torch.bfloat16

TC_CASES = [
    # N, C, H, W, K, R, stride, pad          what it exercises
    (2, 64, 8, 8, 64, 1, 1, 0),       # 1x1 GEMM, BN=64, M=128 exactly
    (3, 128, 7, 7, 256, 1, 1, 0),     # ragged M (147), BN=256, two K blocks
    (2, 256, 8, 8, 128, 1, 1, 0),     # BN=128, four K blocks (ring wraps at 3 stages)
    (2, 64, 8, 8, 64, 3, 1, 1),       # 3x3 im2col, padding rows
    (3, 64, 7, 7, 64, 3, 1, 1),       # 3x3, 7x7 image, M-tile spans images
    (2, 128, 14, 14, 128, 3, 1, 1),   # 3x3, 18 K blocks
    (2, 64, 8, 8, 128, 1, 2, 0),      # strided 1x1 (Bottleneck conv1 / projection)
    (2, 64, 9, 9, 128, 3, 2, 1),      # 3x3 stride 2 (ResNet-34), odd size
    (2, 512, 4, 4, 1000, 1, 1, 0),    # K not a multiple of the N tile (FC-like)
    (5, 64, 6, 5, 192, 3, 1, 1),      # H != W, K=192
    (2, 32, 8, 8, 64, 1, 1, 0),       # C=32: MobileNet's first point-wise conv; TMA zero-fills the half-empty K block
    (3, 96, 5, 5, 32, 1, 1, 0),       # C=96 (1.5 K blocks), K=32 (tile wider than the tensor)
    (2, 192, 6, 6, 64, 1, 1, 0),      # patch-matrix shape class of the 7x7 stem (Kp=192)
    (151, 64, 16, 8, 256, 1, 1, 0),   # CTA-pair kernel (M >= 18944, K = 256): 151 M tiles = an odd, half-empty last pair
    (149, 64, 13, 10, 256, 1, 1, 0),  # CTA-pair kernel, ragged last tile (M = 19370)
    (38, 64, 24, 24, 256, 3, 1, 1),   # CTA-pair kernel with im2col loads (M = 21888 = 171 tiles)
    (3, 64, 30, 56, 64, 3, 1, 1),     # halo-tile kernel (3x3, 64 -> 64, 24 <= W <= 62): ResNet's conv2x row width
    (2, 64, 7, 62, 64, 3, 1, 1),      # halo-tile kernel: widest row the 64-pixel box takes, odd height (last tile = one row)
    (150, 64, 4, 25, 64, 3, 1, 1),    # halo-tile kernel: more tiles than SMs (ring and accumulators wrap), second warp column mostly empty
]


@pytest.mark.parametrize("case", TC_CASES, ids=lambda c: "x".join(map(str, c)))
def test_tc_fprop_matches_oracle(case):
    lib = _lib()
    N, Cc, H, W, K, R, stride, pad = case
    x = _rand((N, Cc, H, W), 1, BF)
    w = _rand((K, Cc, R, R), 2, BF, scale=0.1)
    y_ref = ref_ops.conv2d(x, w, stride, pad)
    y = run_conv(lib.DV_CONV_FPROP, x, w, None, stride, pad, BF, lib.DV_ALGO_TC)
    assert np.isfinite(y).all()
    assert rel_l2(y, y_ref) < 2e-2
    # tighter: identical operands, so only the final bf16 rounding of y separates the two (2^-9 relative per element)
    assert np.max(np.abs(y - y_ref) / (np.abs(y_ref) + 1e-2)) < 2e-2


DGRAD_CASES = [c for c in TC_CASES if (c[4] % 64 == 0 or (c[5] == 1 and c[4] % 8 == 0))] + [
    (2, 64, 8, 8, 128, 3, 2, 1),      # 3x3 stride 2, even size: four parity classes of equal extent
    (3, 128, 10, 7, 64, 3, 2, 1),     # H != W, even/odd mix: the odd classes are one row / column shorter
    (2, 64, 2, 3, 64, 3, 2, 1),       # tiny: a class with a single row
]


@pytest.mark.parametrize("use_wt", [False, True], ids=["krsc_mn_major", "transposed_copy"])
@pytest.mark.parametrize("case", DGRAD_CASES, ids=lambda c: "x".join(map(str, c)))
def test_tc_dgrad_matches_oracle(case, use_wt):
    lib = _lib()
    N, Cc, H, W, K, R, stride, pad = case
    x = _rand((N, Cc, H, W), 1, BF)
    w = _rand((K, Cc, R, R), 2, BF, scale=0.1)
    P, Q = (H + 2 * pad - R) // stride + 1, (W + 2 * pad - R) // stride + 1
    dy = _rand((N, K, P, Q), 3, BF)
    dx_ref, _ = ref_ops.conv2d_backward(x, w, dy, stride, pad)
    dx = run_conv(lib.DV_CONV_DGRAD, x, w, dy, stride, pad, BF, lib.DV_ALGO_TC, use_wt=use_wt)
    assert np.isfinite(dx).all()
    assert rel_l2(dx, dx_ref) < 2e-2


WGRAD_CASES = TC_CASES + [(4, 64, 12, 12, 64, 3, 1, 1), (8, 256, 8, 8, 64, 1, 1, 0),
                                                         (2, 512, 6, 6, 128, 3, 1, 1)]


@pytest.mark.parametrize("case", WGRAD_CASES, ids=lambda c: "x".join(map(str, c)))
def test_tc_wgrad_matches_oracle(case):
    lib = _lib()
    N, Cc, H, W, K, R, stride, pad = case
    x = _rand((N, Cc, H, W), 1, BF)
    w = _rand((K, Cc, R, R), 2, BF, scale=0.1)
    P, Q = (H + 2 * pad - R) // stride + 1, (W + 2 * pad - R) // stride + 1
    dy = _rand((N, K, P, Q), 3, BF)
    _, dw_ref = ref_ops.conv2d_backward(x, w, dy, stride, pad)
    dw = run_conv(lib.DV_CONV_WGRAD, x, w, dy, stride, pad, BF, lib.DV_ALGO_TC)
    assert np.isfinite(dw).all()
    assert rel_l2(dw, dw_ref) < 1e-4      # bf16 products are exact in fp32; only summation order differs
    dw2 = run_conv(lib.DV_CONV_WGRAD, x, w, dy, stride, pad, BF, lib.DV_ALGO_TC)
    assert np.array_equal(dw, dw2)         # fixed-order split reduction: bit-reproducible


@pytest.mark.parametrize("case", WGRAD_CASES, ids=lambda c: "x".join(map(str, c)))
def test_tc_wgrad_accumulate_adds_into_dw(case):
    """dv_conv_params.accumulate = 1 (what the trainer uses): the tiles are ADDED into dW — through TMA reduce-add of fp32 tiles,
    clipped at the tensor's K / C edges — on top of what is already there; twice in a row adds twice."""
    import ctypes as C
    from helpers import stream_ptr
    lib = _lib()
    N, Cc, H, W, K, R, stride, pad = case
    x = _rand((N, Cc, H, W), 1, BF)
    w = _rand((K, Cc, R, R), 2, BF, scale=0.1)
    P, Q = (H + 2 * pad - R) // stride + 1, (W + 2 * pad - R) // stride + 1
    dy = _rand((N, K, P, Q), 3, BF)
    _, dw_ref = ref_ops.conv2d_backward(x, w, dy, stride, pad)
    prior = np.random.default_rng(4).standard_normal((K, R, R, Cc)).astype(np.float32)
    xd = torch.from_numpy(np.ascontiguousarray(x.transpose(0, 2, 3, 1))).cuda().to(BF)
    dyd = torch.from_numpy(np.ascontiguousarray(dy.transpose(0, 2, 3, 1))).cuda().to(BF)
    guard = torch.full((K * R * R * Cc + 64,), 7.0, device="cuda")          # dW sits inside a larger buffer: nothing outside may change
    dwd = guard[32:32 + K * R * R * Cc].view(K, R, R, Cc)
    dwd.copy_(torch.from_numpy(prior))
    p = lib.ConvParams()
    p.x, p.w, p.y = xd.data_ptr(), dwd.data_ptr(), dyd.data_ptr()
    p.N, p.H, p.W, p.C, p.K, p.R, p.S, p.stride, p.pad, p.dtype, p.algo, p.accumulate = N, H, W, Cc, K, R, R, stride, pad, lib.DV_BF16, lib.DV_ALGO_TC, 1
    for times in (1, 2):
        lib.check(lib.lib.dv_conv_wgrad(C.byref(p), stream_ptr()), "wgrad accumulate")
        torch.cuda.synchronize()
        got = dwd.double().cpu().numpy().transpose(0, 3, 1, 2)
        want = prior.astype(np.float64).transpose(0, 3, 1, 2) + times * dw_ref
        assert rel_l2(got, want) < 1e-4, times
    assert float(guard[:32].min()) == 7.0 and float(guard[:32].max()) == 7.0
    assert float(guard[32 + K * R * R * Cc:].min()) == 7.0 and float(guard[32 + K * R * R * Cc:].max()) == 7.0


def _gpu_conv(pass_id, xd, wd, dyd, stride, pad, algo, wt=None):
    """device tensors in (NHWC / KRSC, bf16) -> device tensor out; for the full-size checks (no host round trip)."""
    import ctypes as C
    from helpers import stream_ptr
    lib = _lib()
    N, H, W, Cc = xd.shape
    K, R, S, _ = wd.shape
    P, Q = (H + 2 * pad - R) // stride + 1, (W + 2 * pad - S) // stride + 1
    p = lib.ConvParams()
    p.N, p.H, p.W, p.C, p.K, p.R, p.S, p.stride, p.pad, p.dtype, p.algo = N, H, W, Cc, K, R, S, stride, pad, lib.DV_BF16, algo
    if pass_id == lib.DV_CONV_FPROP:
        out = torch.empty((N, P, Q, K), device="cuda", dtype=BF)
        p.x, p.w, p.y = xd.data_ptr(), wd.data_ptr(), out.data_ptr()
        fn = lib.lib.dv_conv_fprop
    elif pass_id == lib.DV_CONV_DGRAD:
        out = torch.empty((N, H, W, Cc), device="cuda", dtype=BF)
        p.x, p.w, p.y = out.data_ptr(), wd.data_ptr(), dyd.data_ptr()
        p.w_t = wt.data_ptr() if wt is not None else None
        fn = lib.lib.dv_conv_dgrad
    else:
        out = torch.empty((K, R, S, Cc), device="cuda", dtype=torch.float32)
        p.x, p.w, p.y = xd.data_ptr(), out.data_ptr(), dyd.data_ptr()
        fn = lib.lib.dv_conv_wgrad
    need = lib.lib.dv_conv_workspace_bytes(C.byref(p), pass_id)
    ws = torch.empty(max(need, 16), dtype=torch.uint8, device="cuda")
    p.workspace, p.workspace_bytes = ws.data_ptr(), need
    lib.check(fn(C.byref(p), stream_ptr()), "conv")
    torch.cuda.synchronize()
    return out


FULL_CASES = [
    # ResNet-50 layer shapes at batch 256; N, C, H, K, R, stride, pad
    (256, 64, 56, 64, 3, 1, 1),
    (256, 64, 56, 256, 1, 1, 0),
    (256, 256, 56, 128, 1, 2, 0),
    (256, 256, 14, 256, 3, 1, 1),
    (256, 512, 7, 512, 3, 1, 1),
    (256, 2048, 7, 512, 1, 1, 0),
    (256, 64, 56, 128, 3, 2, 1),      # ResNet-34's first down-sampling 3x3 (resnet34.py:96-103): dgrad = four parity classes
]


@pytest.mark.parametrize("case", FULL_CASES, ids=lambda c: "x".join(map(str, c)))
def test_tc_full_size_against_cuda_core_kernel_and_properties(case):
    """At BASELINE sizes the numpy oracle is too slow, so: (1) the tensor-core result equals the independent
    CUDA-core kernel's on the same operands; (2) linearity conv(a*x1 + x2) == a*conv(x1)+conv(x2) with a = 2 (exact in
    bf16 up to output rounding); (3) sampled output pixels equal the oracle's dot products."""
    import ctypes as C
    from helpers import stream_ptr
    lib = _lib()
    N, Cc, H, K, R, stride, pad = case
    g = torch.Generator(device="cuda").manual_seed(5)
    xd = torch.randn((N, H, H, Cc), generator=g, device="cuda").to(BF)
    wd = (torch.randn((K, R, R, Cc), generator=g, device="cuda") * 0.05).to(BF)
    P = (H + 2 * pad - R) // stride + 1
    dyd = torch.randn((N, P, P, K), generator=g, device="cuda").to(BF)
    wf = wd.float().contiguous()
    wt = torch.empty((Cc, R, R, K), device="cuda", dtype=BF)
    lib.check(lib.lib.dv_filter_prep(wf.data_ptr(), None, wt.data_ptr(), K, R, R, Cc, lib.DV_BF16, stream_ptr()), "filter_prep")

    y_tc = _gpu_conv(lib.DV_CONV_FPROP, xd, wd, None, stride, pad, lib.DV_ALGO_TC)
    y_si = _gpu_conv(lib.DV_CONV_FPROP, xd, wd, None, stride, pad, lib.DV_ALGO_SIMT)
    err = float((y_tc.float() - y_si.float()).norm() / y_si.float().norm())
    assert err < 4e-3, f"fprop TC vs SIMT {err}"
    # sampled pixels against the oracle
    xs = xd[:2].double().cpu().numpy().transpose(0, 3, 1, 2)
    y_or = ref_ops.conv2d(xs, wd.double().cpu().numpy().transpose(0, 3, 1, 2), stride, pad)
    assert rel_l2(y_tc[:2].double().cpu().numpy().transpose(0, 3, 1, 2), y_or) < 2e-2
    # linearity
    x2 = torch.randn((N, H, H, Cc), generator=g, device="cuda").to(BF)
    lhs = _gpu_conv(lib.DV_CONV_FPROP, (2 * xd.float() + x2.float()).to(BF), wd, None, stride, pad, lib.DV_ALGO_TC).float()
    rhs = 2 * y_tc.float() + _gpu_conv(lib.DV_CONV_FPROP, x2, wd, None, stride, pad, lib.DV_ALGO_TC).float()
    assert float((lhs - rhs).norm() / rhs.norm()) < 2e-2

    dx_tc = _gpu_conv(lib.DV_CONV_DGRAD, xd, wd, dyd, stride, pad, lib.DV_ALGO_TC, None)
    dx_tc2 = _gpu_conv(lib.DV_CONV_DGRAD, xd, wd, dyd, stride, pad, lib.DV_ALGO_TC, wt)
    assert float((dx_tc.float() - dx_tc2.float()).abs().max()) <= 1e-2 * float(dx_tc2.float().abs().max())
    dx_si = _gpu_conv(lib.DV_CONV_DGRAD, xd, wd, dyd, stride, pad, lib.DV_ALGO_SIMT)
    err = float((dx_tc.float() - dx_si.float()).norm() / dx_si.float().norm())
    assert err < 4e-3, f"dgrad TC vs SIMT {err}"

    dw_tc = _gpu_conv(lib.DV_CONV_WGRAD, xd, wd, dyd, stride, pad, lib.DV_ALGO_TC)
    dw_si = _gpu_conv(lib.DV_CONV_WGRAD, xd, wd, dyd, stride, pad, lib.DV_ALGO_SIMT)
    err = float((dw_tc - dw_si).norm() / dw_si.norm())
    assert err < 1e-4, f"wgrad TC vs SIMT {err}"

    # dgrad and wgrad of the FULL-SIZE launch against the oracle.  dgrad is per image: the first and the last image of the
    # batch (both ends of the tile walk) are checked.  wgrad sums over the batch: with dy zeroed outside those two images the
    # same full-size launch computes exactly their partial sum, which numpy finishes in seconds.
    sel = [0, N - 1]
    nchw = lambda t: t.double().cpu().numpy().transpose(0, 3, 1, 2)
    dx_or, dw_or = ref_ops.conv2d_backward(nchw(xd[sel]), nchw(wd), nchw(dyd[sel]), stride, pad)
    assert rel_l2(nchw(dx_tc[sel]), dx_or) < 2e-2, "full-size dgrad vs oracle"
    dy2 = torch.zeros_like(dyd)
    dy2[sel] = dyd[sel]
    dw2 = _gpu_conv(lib.DV_CONV_WGRAD, xd, wd, dy2, stride, pad, lib.DV_ALGO_TC)
    assert rel_l2(nchw(dw2), dw_or) < 1e-4, "full-size wgrad (two-image partial) vs oracle"


@pytest.mark.parametrize("size", [30, 64], ids=["patch_matrix_path", "direct_path"])
@pytest.mark.parametrize("geom", [(7, 2, 3, 64), (3, 2, 1, 32)], ids=["resnet_stem_7x7", "mobilenet_stem_3x3"])
def test_stem_conv_module_matches_oracle(geom, size):
    """The C=3 stems through the product's nn.Conv2d: 64x64 images take the direct path (dv_stem_*, output width a multiple of
    16), 30x30 ones the patch matrix + tcgen05 GEMM; wgrad through the tensor cores either way."""
    import deep_vision_b200 as dv
    from deep_vision_b200.lib import lib as _l
    R, stride, pad, K = geom
    dv.set_compute_dtype(torch.bfloat16)
    torch.manual_seed(3)
    conv = dv.nn.Conv2d(3, K, R, stride=stride, padding=pad, bias=False).cuda()
    x = torch.randn(4, 3, size, size)
    xr = bf16_round(x.numpy())
    wr = bf16_round(conv.weight.detach().cpu().numpy())
    y_ref = ref_ops.conv2d(xr, wr, stride, pad)
    dy = _rand(y_ref.shape, 9, BF)
    _, dw_ref = ref_ops.conv2d_backward(xr, wr, dy, stride, pad)
    y = conv(x.cuda())
    y.backward(torch.from_numpy(dy).cuda().to(BF))
    torch.cuda.synchronize()
    assert rel_l2(y.detach().float().cpu().numpy(), y_ref) < 2e-2
    assert rel_l2(conv.weight.grad.cpu().numpy(), dw_ref) < 1e-4


def test_tc_fp32_logits_and_dgrad_accumulate():
    """Two epilogue variants of the tensor-core GEMM: fp32 output (the classifier's logits, nn.Linear at resnet50.py:44)
    and read-modify-write accumulation of dgrad (second gradient of a residual block's input)."""
    import ctypes as C
    from helpers import stream_ptr
    lib = _lib()
    # fp32 logits: x [B,1,1,F] * w [1000,F]
    B, Fdim, K = 48, 512, 1000
    x = _rand((B, Fdim, 1, 1), 1, BF)
    w = _rand((K, Fdim, 1, 1), 2, BF, scale=0.05)
    xd = torch.from_numpy(x.reshape(B, Fdim)).cuda().to(BF)
    wd = torch.from_numpy(w.reshape(K, Fdim)).cuda().to(BF)
    y = torch.full((B, K), float("nan"), device="cuda")
    p = lib.ConvParams()
    p.x, p.w, p.y = xd.data_ptr(), wd.data_ptr(), y.data_ptr()
    p.N, p.H, p.W, p.C, p.K, p.R, p.S, p.stride, p.pad, p.dtype, p.algo, p.y_f32 = B, 1, 1, Fdim, K, 1, 1, 1, 0, lib.DV_BF16, lib.DV_ALGO_TC, 1
    lib.check(lib.lib.dv_conv_fprop(C.byref(p), stream_ptr()), "fprop f32")
    torch.cuda.synchronize()
    ref = x.reshape(B, Fdim) @ w.reshape(K, Fdim).T
    assert rel_l2(y.double().cpu().numpy(), ref) < 1e-5       # exact bf16 products, fp32 accumulate, no output rounding

    for case, algo in (((2, 64, 8, 8, 128, 1, 1, 0), lib.DV_ALGO_TC), ((2, 64, 8, 8, 128, 1, 2, 0), lib.DV_ALGO_TC),
                       ((2, 64, 8, 8, 64, 3, 1, 1), lib.DV_ALGO_TC), ((2, 64, 9, 9, 128, 3, 2, 1), lib.DV_ALGO_SIMT),
                       ((2, 64, 9, 9, 128, 3, 2, 1), lib.DV_ALGO_TC), ((2, 128, 12, 10, 64, 3, 2, 1), lib.DV_ALGO_TC)):
        N, Cc, H, W, K, R, stride, pad = case
        xs = _rand((N, Cc, H, W), 1, BF)
        ws = _rand((K, Cc, R, R), 2, BF, scale=0.1)
        P, Q = (H + 2 * pad - R) // stride + 1, (W + 2 * pad - R) // stride + 1
        dy = _rand((N, K, P, Q), 3, BF)
        prior = _rand((N, Cc, H, W), 4, BF)
        dx_ref, _ = ref_ops.conv2d_backward(xs, ws, dy, stride, pad)
        wk = torch.from_numpy(np.ascontiguousarray(ws.transpose(0, 2, 3, 1))).cuda().to(BF)
        wt = torch.empty((Cc, R, R, K), device="cuda", dtype=BF)
        lib.check(lib.lib.dv_filter_prep(wk.float().contiguous().data_ptr(), None, wt.data_ptr(), K, R, R, Cc, lib.DV_BF16, stream_ptr()), "prep")
        dyd = torch.from_numpy(np.ascontiguousarray(dy.transpose(0, 2, 3, 1))).cuda().to(BF)
        dxd = torch.from_numpy(np.ascontiguousarray(prior.transpose(0, 2, 3, 1))).cuda().to(BF)
        q = lib.ConvParams()
        q.x, q.w, q.y = dxd.data_ptr(), wk.data_ptr(), dyd.data_ptr()
        q.N, q.H, q.W, q.C, q.K, q.R, q.S, q.stride, q.pad, q.dtype, q.algo, q.accumulate = N, H, W, Cc, K, R, R, stride, pad, lib.DV_BF16, algo, 1
        lib.check(lib.lib.dv_conv_dgrad(C.byref(q), stream_ptr()), "dgrad accumulate")
        torch.cuda.synchronize()
        got = dxd.double().cpu().numpy().transpose(0, 3, 1, 2)
        assert rel_l2(got, dx_ref + prior) < 2e-2, case


@pytest.mark.parametrize("case", [(2, 64, 8, 8, 64, 3, 1, 1), (3, 128, 7, 7, 256, 1, 1, 0), (4, 64, 12, 12, 512, 1, 1, 0),
                                  (2, 256, 9, 9, 1000, 1, 1, 0), (3, 64, 9, 40, 64, 3, 1, 1), (150, 64, 4, 25, 64, 3, 1, 1)],
                         ids=lambda c: "x".join(map(str, c)))
def test_tc_fprop_fused_bn_sums(case):
    """The epilogue's per-channel sums (dv_conv_params.stat_sums) equal the sums of the stored bf16 output, and
    dv_bn_stats_from_sums turns them into the statistics dv_bn_stats computes from that output."""
    import ctypes as C
    from helpers import stream_ptr
    lib = _lib()
    N, Cc, H, W, K, R, stride, pad = case
    x = _rand((N, Cc, H, W), 1, BF)
    w = _rand((K, Cc, R, R), 2, BF, scale=0.1)
    xd = torch.from_numpy(np.ascontiguousarray(x.transpose(0, 2, 3, 1))).cuda().to(BF)
    wd = torch.from_numpy(np.ascontiguousarray(w.transpose(0, 2, 3, 1))).cuda().to(BF)
    P, Q = (H + 2 * pad - R) // stride + 1, (W + 2 * pad - R) // stride + 1
    y = torch.empty((N, P, Q, K), device="cuda", dtype=BF)
    sums = torch.zeros(2 * K, device="cuda")
    p = lib.ConvParams()
    p.x, p.w, p.y, p.stat_sums = xd.data_ptr(), wd.data_ptr(), y.data_ptr(), sums.data_ptr()
    p.N, p.H, p.W, p.C, p.K, p.R, p.S, p.stride, p.pad, p.dtype, p.algo = N, H, W, Cc, K, R, R, stride, pad, lib.DV_BF16, lib.DV_ALGO_TC
    lib.check(lib.lib.dv_conv_fprop(C.byref(p), stream_ptr()), "fprop+sums")
    torch.cuda.synchronize()
    yd = y.double().reshape(-1, K)
    assert rel_l2(sums[:K].cpu().numpy(), yd.sum(0).cpu().numpy()) < 1e-5
    assert rel_l2(sums[K:].cpu().numpy(), (yd * yd).sum(0).cpu().numpy()) < 1e-5
    M = N * P * Q
    f = lambda v: torch.full((K,), float(v), device="cuda")
    g_, b_, rm, rv, mean, inv = f(1), f(0), f(0), f(1), f(float("nan")), f(float("nan"))
    q = lib.BnParams()
    q.x, q.gamma, q.beta, q.running_mean, q.running_var = y.data_ptr(), g_.data_ptr(), b_.data_ptr(), rm.data_ptr(), rv.data_ptr()
    q.save_mean, q.save_invstd, q.M, q.C, q.dtype, q.eps, q.momentum = mean.data_ptr(), inv.data_ptr(), M, K, lib.DV_BF16, 1e-5, 0.1
    lib.check(lib.lib.dv_bn_stats_from_sums(C.byref(q), sums.data_ptr(), stream_ptr()), "from_sums")
    torch.cuda.synchronize()
    assert float(sums.abs().sum()) == 0.0                                   # cleared for the next step
    assert rel_l2(mean.cpu().numpy(), yd.mean(0).cpu().numpy()) < 1e-4
    assert rel_l2(inv.cpu().numpy(), (1.0 / torch.sqrt(yd.var(0, unbiased=False) + 1e-5)).cpu().numpy()) < 1e-4
    assert rel_l2(rm.cpu().numpy(), (0.1 * yd.mean(0)).cpu().numpy()) < 1e-4


FOLD_CASES = [
    # N, C, H, W, K, R, stride, pad, residual, relu
    (2, 64, 8, 8, 64, 3, 1, 1, False, True),      # BN=64 tile, conv2 of a bottleneck
    (3, 128, 7, 7, 256, 1, 1, 0, True, True),     # conv3 + identity + ReLU, ragged M
    (2, 64, 8, 8, 128, 1, 2, 0, False, False),    # projection shortcut: strided 1x1, BN only
    (2, 512, 4, 4, 1000, 1, 1, 0, True, True),    # K not a multiple of the N tile
    (2, 192, 6, 6, 64, 1, 1, 0, False, True),     # patch-matrix shape class of the stem
    (2, 64, 160, 128, 256, 1, 1, 0, True, True),  # M = 40960 >= 148 x 128 x 2: the CTA-pair kernel's epilogue
]


@pytest.mark.parametrize("case", FOLD_CASES, ids=lambda c: "x".join(map(str, c)))
def test_tc_fprop_folded_batchnorm_epilogue(case):
    """dv_conv_params.ep_*: y = relu(conv * scale + shift + identity) against the oracle's conv -> eval BN -> add -> ReLU,
    with scale/shift made by dv_bn_fold."""
    import ctypes as C
    from helpers import stream_ptr
    lib = _lib()
    N, Cc, H, W, K, R, stride, pad, residual, relu = case
    x = _rand((N, Cc, H, W), 1, BF)
    w = _rand((K, Cc, R, R), 2, BF, scale=0.1)
    rng = np.random.default_rng(3)
    gamma, beta = rng.uniform(0.5, 1.5, K), rng.normal(0, 0.3, K)
    rm, rv = rng.normal(0, 0.5, K), rng.uniform(0.5, 2.0, K)
    P, Q = (H + 2 * pad - R) // stride + 1, (W + 2 * pad - R) // stride + 1
    res = _rand((N, K, P, Q), 4, BF) if residual else None
    ref = ref_ops.batch_norm_eval(ref_ops.conv2d(x, w, stride, pad), gamma, beta, rm, rv)
    if residual:
        ref = ref + res
    if relu:
        ref = ref_ops.relu(ref)
    f32 = lambda a: torch.from_numpy(np.asarray(a, np.float32)).cuda()
    g_d, b_d, rm_d, rv_d = f32(gamma), f32(beta), f32(rm), f32(rv)
    scale, shift = torch.empty(K, device="cuda"), torch.empty(K, device="cuda")
    lib.check(lib.lib.dv_bn_fold(g_d.data_ptr(), b_d.data_ptr(), rm_d.data_ptr(), rv_d.data_ptr(), 1e-5, scale.data_ptr(), shift.data_ptr(),
                                 K, None, None, stream_ptr()), "dv_bn_fold")
    torch.cuda.synchronize()
    assert rel_l2(scale.cpu().numpy(), gamma / np.sqrt(rv + 1e-5)) < 1e-6
    assert rel_l2(shift.cpu().numpy(), beta - rm * gamma / np.sqrt(rv + 1e-5)) < 1e-5
    xd = torch.from_numpy(np.ascontiguousarray(x.transpose(0, 2, 3, 1))).cuda().to(BF)
    wd = torch.from_numpy(np.ascontiguousarray(w.transpose(0, 2, 3, 1))).cuda().to(BF)
    rd = torch.from_numpy(np.ascontiguousarray(res.transpose(0, 2, 3, 1))).cuda().to(BF) if residual else None
    y = torch.full((N, P, Q, K), float("nan"), device="cuda", dtype=BF)
    p = lib.ConvParams()
    p.x, p.w, p.y = xd.data_ptr(), wd.data_ptr(), y.data_ptr()
    p.N, p.H, p.W, p.C, p.K, p.R, p.S, p.stride, p.pad, p.dtype, p.algo = N, H, W, Cc, K, R, R, stride, pad, lib.DV_BF16, lib.DV_ALGO_TC
    p.ep_scale, p.ep_shift, p.ep_relu = scale.data_ptr(), shift.data_ptr(), int(relu)
    p.ep_residual = rd.data_ptr() if residual else None
    lib.check(lib.lib.dv_conv_fprop(C.byref(p), stream_ptr()), "fprop(folded)")
    torch.cuda.synchronize()
    out = y.double().cpu().numpy().transpose(0, 3, 1, 2)
    assert np.isfinite(out).all()
    assert rel_l2(out, ref) < 1e-2   # one bf16 rounding of the final value
    # the epilogue is refused, not ignored, where it does not exist
    p.algo = lib.DV_ALGO_SIMT
    assert lib.lib.dv_conv_fprop(C.byref(p), stream_ptr()) == lib.DV_ERR_UNSUPPORTED
    p.algo = lib.DV_ALGO_TC
    assert lib.lib.dv_conv_dgrad(C.byref(p), stream_ptr()) == lib.DV_ERR_UNSUPPORTED


BNB_CASES = [
    # N, C, H, W, K, R, pad        dgrad output = [N,H,W,C]: the tensor whose BatchNorm sums are fused
    (2, 64, 8, 8, 64, 3, 1),          # 64-wide tile, 3x3, M = 128
    (3, 128, 7, 7, 256, 1, 0),        # 128-wide tile, ragged M (147)
    (2, 256, 8, 8, 128, 1, 0),        # 256-wide tile, persistent kernel
    (5, 64, 6, 5, 192, 3, 1),         # H != W, three reduction blocks
    (3, 96, 5, 5, 32, 1, 0),          # C = 96: the second 64-column group is half empty
    (2, 512, 4, 4, 64, 1, 0),         # two 256-wide n tiles per CTA walk
    (151, 64, 16, 8, 256, 1, 0),      # many M tiles per CTA: running sums across tiles, one flush
    (38, 256, 24, 24, 64, 1, 0),      # CTA-pair kernel (M = 21888, 256 columns)
    (37, 256, 24, 23, 128, 3, 1),     # CTA-pair kernel, 3x3, odd number of M tiles (the peer's last tile is empty)
    (3, 64, 9, 40, 64, 3, 1),         # the halo-tile geometry (its plain dgrad takes the halo kernel, the one with sums the im2col kernel), odd height
    (150, 64, 4, 25, 64, 3, 1),       # the same, more tiles than SMs
]


@pytest.mark.parametrize("case", BNB_CASES, ids=lambda c: "x".join(map(str, c)))
def test_tc_dgrad_fused_bn_backward_sums(case):
    """dv_conv_params.bnb_*: the dgrad epilogue adds dbeta = sum dz and dgamma = sum dz*xhat of the BatchNorm that produced the
    convolution's input (dz = dx masked by relu(bn(x0)) > 0).  Checked three ways: dx is unchanged by the option (bit for bit);
    the sums equal fp64 sums over the kernel's own bf16 dx (what a separate reduce pass would read); and they equal
    dv_bn_bwd_reduce's on the same tensors.  The accumulators are pre-loaded, so the test also shows that the sums are ADDED."""
    import ctypes as C
    from helpers import stream_ptr
    lib = _lib()
    N, Cc, H, W, K, R, pad = case
    g = torch.Generator(device="cuda").manual_seed(11)
    dy = torch.randn((N, H + 2 * pad - R + 1, W + 2 * pad - R + 1, K), generator=g, device="cuda").to(BF)
    w = (torch.randn((K, R, R, Cc), generator=g, device="cuda") * 0.1).to(BF)
    x0 = (torch.randn((N, H, W, Cc), generator=g, device="cuda") * 1.5 + 0.3).to(BF)
    gamma = torch.rand(Cc, generator=g, device="cuda") + 0.5
    beta = torch.randn(Cc, generator=g, device="cuda") * 0.3
    xf = x0.float().reshape(-1, Cc)
    mean = xf.mean(0)
    invstd = 1.0 / torch.sqrt(xf.var(0, unbiased=False) + 1e-5)

    def dgrad(with_sums):
        dx = torch.full((N, H, W, Cc), float("nan"), device="cuda", dtype=BF)
        dgamma = torch.full((Cc,), 2.0, device="cuda")
        dbeta = torch.full((Cc,), -3.0, device="cuda")
        p = lib.ConvParams()
        p.x, p.w, p.y = dx.data_ptr(), w.data_ptr(), dy.data_ptr()
        p.N, p.H, p.W, p.C, p.K, p.R, p.S, p.stride, p.pad, p.dtype, p.algo = N, H, W, Cc, K, R, R, 1, pad, lib.DV_BF16, lib.DV_ALGO_TC
        if with_sums:
            p.bnb_x, p.bnb_gamma, p.bnb_beta, p.bnb_mean, p.bnb_invstd = x0.data_ptr(), gamma.data_ptr(), beta.data_ptr(), mean.data_ptr(), invstd.data_ptr()
            p.bnb_dgamma, p.bnb_dbeta = dgamma.data_ptr(), dbeta.data_ptr()
        lib.check(lib.lib.dv_conv_dgrad(C.byref(p), stream_ptr()), "dgrad")
        torch.cuda.synchronize()
        return dx, dgamma, dbeta

    dx_plain, _, _ = dgrad(False)
    dx, dgamma, dbeta = dgrad(True)
    assert torch.equal(dx.view(torch.int16), dx_plain.view(torch.int16))
    # fp64 sums over the kernel's dx; elements whose pre-activation is within rounding of zero may fall on either side
    sc = (invstd * gamma).double()
    pre = x0.double().reshape(-1, Cc) * sc + (beta.double() - mean.double() * sc)
    dz = dx.double().reshape(-1, Cc) * (pre > 0)
    xhat = (x0.double().reshape(-1, Cc) - mean.double()) * invstd.double()
    want_b, want_g = dz.sum(0), (dz * xhat).sum(0)
    scale_b = dz.abs().sum(0) + 1e-6          # the sums cancel heavily: compare against the size of what was summed
    scale_g = (dz * xhat).abs().sum(0) + 1e-6
    assert float(((dbeta.double() + 3.0 - want_b).abs() / scale_b).max()) < 1e-4
    assert float(((dgamma.double() - 2.0 - want_g).abs() / scale_g).max()) < 1e-4
    # and the stand-alone reduce pass on the same tensors
    q = lib.BnBwdParams()
    dg2, db2 = torch.empty(Cc, device="cuda"), torch.empty(Cc, device="cuda")
    M = N * H * W
    ws = torch.empty(lib.lib.dv_bn_workspace_bytes(M, Cc), dtype=torch.uint8, device="cuda")
    q.x, q.dy, q.gamma, q.beta, q.save_mean, q.save_invstd = x0.data_ptr(), dx.data_ptr(), gamma.data_ptr(), beta.data_ptr(), mean.data_ptr(), invstd.data_ptr()
    q.dgamma, q.dbeta, q.workspace, q.workspace_bytes = dg2.data_ptr(), db2.data_ptr(), ws.data_ptr(), ws.numel()
    q.M, q.C, q.dtype, q.relu = M, Cc, lib.DV_BF16, 1
    lib.check(lib.lib.dv_bn_bwd_reduce(C.byref(q), stream_ptr()), "bwd_reduce")
    torch.cuda.synchronize()
    assert float(((dbeta - (-3.0) - db2).double().abs() / scale_b).max()) < 1e-4
    assert float(((dgamma - 2.0 - dg2).double().abs() / scale_g).max()) < 1e-4


def test_tc_dgrad_bn_sums_rejects_what_it_cannot_fuse():
    import ctypes as C
    from helpers import stream_ptr
    lib = _lib()
    t = torch.zeros(2 * 8 * 8 * 64, device="cuda", dtype=BF)
    v = torch.zeros(64, device="cuda")
    p = lib.ConvParams()
    p.x, p.w, p.y = t.data_ptr(), t.data_ptr(), t.data_ptr()
    p.N, p.H, p.W, p.C, p.K, p.R, p.S, p.stride, p.pad, p.dtype, p.algo = 2, 8, 8, 64, 64, 1, 1, 1, 0, lib.DV_BF16, lib.DV_ALGO_TC
    p.bnb_x, p.bnb_gamma, p.bnb_beta, p.bnb_mean, p.bnb_invstd, p.bnb_dgamma, p.bnb_dbeta = [t.data_ptr()] + [v.data_ptr()] * 6
    p.accumulate = 1
    assert lib.lib.dv_conv_dgrad(C.byref(p), stream_ptr()) == lib.DV_ERR_UNSUPPORTED       # accumulate: the final dx never is on chip
    p.accumulate, p.algo = 0, lib.DV_ALGO_SIMT
    assert lib.lib.dv_conv_dgrad(C.byref(p), stream_ptr()) == lib.DV_ERR_UNSUPPORTED       # CUDA-core path has no such epilogue
    p.algo = lib.DV_ALGO_TC
    assert lib.lib.dv_conv_fprop(C.byref(p), stream_ptr()) == lib.DV_ERR_UNSUPPORTED       # dgrad only
    p.bnb_dbeta = None
    assert lib.lib.dv_conv_dgrad(C.byref(p), stream_ptr()) == lib.DV_ERR_ARG


SHORTCUT_CASES = [c for c in BNB_CASES if c[1] % 32 == 0]


@pytest.mark.parametrize("case", SHORTCUT_CASES, ids=lambda c: "x".join(map(str, c)))
def test_tc_dgrad_adds_masked_shortcut_gradient(case):
    """dv_conv_params.ep_residual / ep_residual_mask in dgrad: dx = dgrad(dy) + (mask ? dout : 0), the input gradient of an identity
    residual block (resnet50.py:147-164) in one pass.  With dy = 0 the result must be the masked dout bit for bit (bit layout, row
    and column addressing of the mask); with dout = 0 it must be the plain dgrad bit for bit; with both it must agree with their
    fp32 sum to one bf16 rounding of the accumulator (the plain dgrad rounds before the add, the fused one after)."""
    import ctypes as C
    from helpers import stream_ptr
    lib = _lib()
    N, Cc, H, W, K, R, pad = case
    g = torch.Generator(device="cuda").manual_seed(21)
    dy = torch.randn((N, H + 2 * pad - R + 1, W + 2 * pad - R + 1, K), generator=g, device="cuda").to(BF)
    w = (torch.randn((K, R, R, Cc), generator=g, device="cuda") * 0.1).to(BF)
    dout = torch.randn((N, H, W, Cc), generator=g, device="cuda").to(BF)
    keep = torch.rand((N * H * W, Cc), generator=g, device="cuda") < 0.6
    # bit c%8 of byte c/8, one row of C/8 bytes per pixel: the layout dv_bn_apply writes (relu_mask)
    weights = (2 ** torch.arange(8, device="cuda", dtype=torch.int32)).view(1, 1, 8)
    mask = (keep.view(-1, Cc // 8, 8).to(torch.int32) * weights).sum(-1).to(torch.uint8).contiguous()

    def dgrad(dy_t, res, msk):
        dx = torch.full((N, H, W, Cc), float("nan"), device="cuda", dtype=BF)
        p = lib.ConvParams()
        p.x, p.w, p.y = dx.data_ptr(), w.data_ptr(), dy_t.data_ptr()
        p.N, p.H, p.W, p.C, p.K, p.R, p.S, p.stride, p.pad, p.dtype, p.algo = N, H, W, Cc, K, R, R, 1, pad, lib.DV_BF16, lib.DV_ALGO_TC
        if res is not None:
            p.ep_residual = res.data_ptr()
            p.ep_residual_mask = msk.data_ptr() if msk is not None else None
        lib.check(lib.lib.dv_conv_dgrad(C.byref(p), stream_ptr()), "dgrad")
        torch.cuda.synchronize()
        return dx

    masked = torch.where(keep.view(N, H, W, Cc), dout, torch.zeros_like(dout))
    only_res = dgrad(torch.zeros_like(dy), dout, mask)
    assert torch.equal(only_res.view(torch.int16), (masked + 0.0).view(torch.int16))    # (+0.0: -0 + 0 accumulators give +0)
    unmasked = dgrad(torch.zeros_like(dy), dout, None)
    assert torch.equal(unmasked.view(torch.int16), (dout + 0.0).view(torch.int16))
    plain = dgrad(dy, None, None)
    assert torch.equal(dgrad(dy, torch.zeros_like(dout), mask).view(torch.int16), plain.view(torch.int16))
    both = dgrad(dy, dout, mask)
    want = plain.double() + masked.double()
    err = (both.double() - want).abs()
    # two bf16 roundings apart at most: of the accumulator (plain) and of the sum
    assert bool((err <= 2.0 ** -8 * (plain.double().abs() + want.abs()) + 1e-30).all())
    assert rel_l2(both.double().cpu().numpy(), want.cpu().numpy()) < 4e-3


def test_tc_dgrad_shortcut_epilogue_rejects_what_it_cannot_fuse():
    import ctypes as C
    from helpers import stream_ptr
    lib = _lib()
    t = torch.zeros(2 * 8 * 8 * 64, device="cuda", dtype=BF)
    m = torch.zeros(2 * 8 * 8 * 8, device="cuda", dtype=torch.uint8)
    p = lib.ConvParams()
    p.x, p.w, p.y = t.data_ptr(), t.data_ptr(), t.data_ptr()
    p.N, p.H, p.W, p.C, p.K, p.R, p.S, p.stride, p.pad, p.dtype, p.algo = 2, 8, 8, 64, 64, 1, 1, 1, 0, lib.DV_BF16, lib.DV_ALGO_TC
    p.ep_residual, p.ep_residual_mask = t.data_ptr(), m.data_ptr()
    assert lib.lib.dv_conv_dgrad(C.byref(p), stream_ptr()) == lib.DV_OK
    p.accumulate = 1
    assert lib.lib.dv_conv_dgrad(C.byref(p), stream_ptr()) == lib.DV_ERR_UNSUPPORTED       # either accumulate or add in the epilogue
    p.accumulate, p.algo = 0, lib.DV_ALGO_SIMT
    assert lib.lib.dv_conv_dgrad(C.byref(p), stream_ptr()) == lib.DV_ERR_UNSUPPORTED       # CUDA-core path has no such epilogue
    p.algo, p.ep_residual = lib.DV_ALGO_TC, None
    assert lib.lib.dv_conv_dgrad(C.byref(p), stream_ptr()) == lib.DV_ERR_ARG               # a mask without the tensor it masks
    p.ep_residual, p.stride, p.H, p.W = t.data_ptr(), 2, 16, 16
    assert lib.lib.dv_conv_dgrad(C.byref(p), stream_ptr()) == lib.DV_ERR_UNSUPPORTED       # stride 1 only
    torch.cuda.synchronize()


STEM_CASES = [
    # N, H, W, C, K, R, pad          stride 2
    (2, 64, 64, 3, 64, 7, 3),       # ResNet stem geometry, Q = 32
    (3, 96, 64, 3, 64, 7, 3),       # H != W (P = 48, Q = 32)
    (2, 64, 96, 3, 32, 3, 1),       # MobileNet stem: 3x3, K = 32, Q = 48
    (1, 32, 224, 4, 64, 7, 3),      # full output width Q = 112, C = 4
    (5, 20, 32, 3, 16, 5, 2),       # Q = 16, odd-sized everything else
]


@pytest.mark.parametrize("case", STEM_CASES, ids=lambda c: "x".join(map(str, c)))
def test_stem_direct_convolution_matches_oracle(case):
    """dv_stem_*: the C <= 4 stride-2 stems on the tensor cores straight from the padded NHWC4 image (sliding-window tensor map),
    fprop with fused BatchNorm sums and wgrad in both reduction modes, against the numpy oracle."""
    import ctypes as C
    from helpers import stream_ptr
    lib = _lib()
    N, H, W, Cc, K, R, pad = case
    x = _rand((N, Cc, H, W), 1, BF)
    w = _rand((K, Cc, R, R), 2, BF, scale=0.1)
    y_ref = ref_ops.conv2d(x, w, 2, pad)
    dy = _rand(y_ref.shape, 3, BF)
    _, dw_ref = ref_ops.conv2d_backward(x, w, dy, 2, pad)
    P, Q = y_ref.shape[2], y_ref.shape[3]
    xd = torch.from_numpy(x.transpose(0, 2, 3, 1).copy()).cuda().to(BF)
    wd = torch.from_numpy(w.transpose(0, 2, 3, 1).copy()).cuda().float().contiguous()        # fp32 KRSC master filters
    dyd = torch.from_numpy(dy.transpose(0, 2, 3, 1).copy()).cuda().to(BF)
    p = lib.StemParams()
    p.N, p.H, p.W, p.C, p.K, p.R, p.S, p.stride, p.pad = N, H, W, Cc, K, R, R, 2, pad
    hp, wp = C.c_int(0), C.c_int(0)
    assert lib.lib.dv_stem_supported(C.byref(p), C.byref(hp), C.byref(wp)) == 1
    xp = torch.full((N, hp.value, wp.value, 4), float("nan"), device="cuda", dtype=BF)
    wpk = torch.full((64, R, 32), float("nan"), device="cuda", dtype=BF)
    y = torch.full((N, P, Q, K), float("nan"), device="cuda", dtype=BF)
    sums = torch.zeros(2 * K, device="cuda")
    p.x, p.w, p.y, p.stat_sums = xp.data_ptr(), wpk.data_ptr(), y.data_ptr(), sums.data_ptr()
    lib.check(lib.lib.dv_stem_pad_input(C.byref(p), xd.data_ptr(), stream_ptr()), "pad")
    lib.check(lib.lib.dv_stem_pack_filters(C.byref(p), wd.data_ptr(), stream_ptr()), "pack")
    lib.check(lib.lib.dv_stem_fprop(C.byref(p), stream_ptr()), "stem fprop")
    torch.cuda.synchronize()
    got = y.double().cpu().numpy().transpose(0, 3, 1, 2)
    assert np.isfinite(got).all()
    assert rel_l2(got, y_ref) < 2e-2
    assert np.max(np.abs(got - y_ref) / (np.abs(y_ref) + 1e-2)) < 2e-2
    yb = y.double().reshape(-1, K)
    assert rel_l2(sums[:K].double().cpu().numpy(), yb.sum(0).cpu().numpy()) < 1e-4
    assert rel_l2(sums[K:].double().cpu().numpy(), (yb * yb).sum(0).cpu().numpy()) < 1e-4
    for accumulate in (0, 1):
        dw = torch.full((K, R, R, Cc), float("nan"), device="cuda")
        p.y, p.dw, p.accumulate = dyd.data_ptr(), dw.data_ptr(), accumulate
        ws = torch.empty(max(lib.lib.dv_stem_workspace_bytes(C.byref(p)), 16), dtype=torch.uint8, device="cuda")
        p.workspace, p.workspace_bytes = ws.data_ptr(), ws.numel()
        for _ in range(2):      # twice: the result replaces dw, it is not added to it
            lib.check(lib.lib.dv_stem_wgrad(C.byref(p), stream_ptr()), "stem wgrad")
        torch.cuda.synchronize()
        assert rel_l2(dw.double().cpu().numpy().transpose(0, 3, 1, 2), dw_ref) < 1e-4, accumulate


def test_stem_direct_rejects_other_geometries():
    import ctypes as C
    lib = _lib()
    p = lib.StemParams()
    p.N, p.H, p.W, p.C, p.K, p.R, p.S, p.stride, p.pad = 2, 64, 64, 3, 64, 7, 7, 2, 3
    assert lib.lib.dv_stem_supported(C.byref(p), None, None) == 1
    for field, bad in (("stride", 1), ("C", 8), ("K", 128), ("R", 9), ("W", 60)):     # W = 60 -> Q = 30, not a multiple of 16
        q = lib.StemParams.from_buffer_copy(p)
        setattr(q, field, bad)
        assert lib.lib.dv_stem_supported(C.byref(q), None, None) == 0, field


@pytest.mark.parametrize("case", STEM_CASES[:3], ids=lambda c: "x".join(map(str, c)))
def test_stem_pad_input_from_the_loaders_fp32_nchw_batch(case):
    """dv_stem_pad_input_nchw (fp32 [N,C,H,W] -> padded NHWC4 bf16 in one pass) == dv_nchw_to_nhwc followed by dv_stem_pad_input,
    bit for bit; and nn.Conv2d takes that route in training (no nchw_to_nhwc launch) with the same output."""
    import ctypes as C
    from helpers import stream_ptr
    import deep_vision_b200 as dv
    lib = _lib()
    N, H, W, Cc, K, R, pad = case
    g = torch.Generator(device="cuda").manual_seed(5)
    x = torch.randn((N, Cc, H, W), generator=g, device="cuda") * 60 + 20          # pixel-scale values (train.py:324 normalises 0..255)
    p = lib.StemParams()
    p.N, p.H, p.W, p.C, p.K, p.R, p.S, p.stride, p.pad = N, H, W, Cc, K, R, R, 2, pad
    hp, wp = C.c_int(0), C.c_int(0)
    assert lib.lib.dv_stem_supported(C.byref(p), C.byref(hp), C.byref(wp)) == 1
    nhwc = torch.empty((N, H, W, Cc), device="cuda", dtype=BF)
    lib.check(lib.lib.dv_nchw_to_nhwc(x.data_ptr(), nhwc.data_ptr(), N, Cc, H, W, lib.DV_BF16, stream_ptr()), "nchw_to_nhwc")
    a = torch.full((N, hp.value, wp.value, 4), float("nan"), device="cuda", dtype=BF)
    b = torch.full_like(a, float("nan"))
    p.x = a.data_ptr()
    lib.check(lib.lib.dv_stem_pad_input(C.byref(p), nhwc.data_ptr(), stream_ptr()), "pad")
    p.x = b.data_ptr()
    lib.check(lib.lib.dv_stem_pad_input_nchw(C.byref(p), x.data_ptr(), stream_ptr()), "pad_nchw")
    torch.cuda.synchronize()
    assert torch.equal(a.view(torch.int16), b.view(torch.int16))
    # the module: fp32 NCHW in training goes through the one-pass route; a pre-converted input through the two-pass one
    if Cc == 3:
        dv.set_compute_dtype(BF)
        conv = dv.nn.Conv2d(Cc, K, R, stride=2, padding=pad, bias=False).cuda().train()
        n0 = lib.lib.dv_launch_count()
        y1 = conv(x)
        n1 = lib.lib.dv_launch_count()
        y2 = conv(nhwc.permute(0, 3, 1, 2))
        n2 = lib.lib.dv_launch_count()
        torch.cuda.synchronize()
        assert torch.equal(y1, y2)
        assert (n1 - n0) - (n2 - n1) == 1 - 0 + 0 or (n1 - n0) <= (n2 - n1) + 1   # first call also packs the filters once
