"""dv_bn_stats_static — batch-norm statistics for callers that are not allowed to allocate.

Some hosts of this library own the allocator: they reserve their device memory once and forbid cudaMalloc afterwards
(it synchronises the whole device and can fail in the middle of a step), or they are plug-ins that receive input and
output buffers and nothing else.  They cannot give dv_bn_stats the scratch buffer it asks for, so the library offers a
twin whose scratch is a fixed 2.3 MiB array inside the library image (C <= 2048 channels).

The test drives both entry points through the C ABI on the same tensor and requires the same mean, inverse standard
deviation and updated running statistics (to the tolerances of the other BN tests), TWICE IN A ROW with different
data: the second call finds the first call's partials in the shared array and must not be disturbed by them.
The twin's own call receives no workspace at all.  Sizes beyond the built-in scratch must be refused with a message.

This file sorts last on purpose (tests/conftest.py keeps test_zz_* at the end): a `-x` run has then already executed
every other test before reaching it.  It must pass alone in a fresh process and after the rest of the suite.
"""
import ctypes as C

import numpy as np
import pytest
import torch

from helpers import rel_l2, stream_ptr

pytestmark = pytest.mark.gpu


def _params(lib, x, M, Cc, dtype, gamma, beta, rm, rv, mean, invstd, ws=None):
    p = lib.BnParams()
    p.x, p.gamma, p.beta = x.data_ptr(), gamma.data_ptr(), beta.data_ptr()
    p.running_mean, p.running_var = rm.data_ptr(), rv.data_ptr()
    p.save_mean, p.save_invstd = mean.data_ptr(), invstd.data_ptr()
    if ws is not None:
        p.workspace, p.workspace_bytes = ws.data_ptr(), ws.numel()
    p.M, p.C, p.dtype, p.eps, p.momentum = M, Cc, dtype, 1e-5, 0.1
    return p


@pytest.mark.parametrize("tdtype", [torch.bfloat16, torch.float32], ids=["bf16", "fp32"])
def test_static_scratch_bn_stats_matches_workspace_version_twice(tdtype):
    from deep_vision_b200 import lib
    dv_dtype = lib.DV_BF16 if tdtype == torch.bfloat16 else lib.DV_F32
    # first call: a tensor with many more rows than the built-in scratch has partial rows, and the widest C it takes;
    # second call: a different shape, so stale partials of the first call sit beyond and inside the region in use
    for call, (M, Cc, seed) in enumerate([(40000, 2048, 1), (150000, 64, 2)]):
        g = torch.Generator(device="cuda").manual_seed(seed)
        x = (torch.randn(M, Cc, device="cuda", generator=g) * 2.0 + 0.5).to(tdtype)
        gamma, beta = torch.ones(Cc, device="cuda"), torch.zeros(Cc, device="cuda")
        rm0 = torch.randn(Cc, device="cuda", generator=g)
        rv0 = torch.rand(Cc, device="cuda", generator=g) + 0.5
        # reference: the existing entry point with a caller-provided workspace
        rm_a, rv_a = rm0.clone(), rv0.clone()
        mean_a, inv_a = torch.empty(Cc, device="cuda"), torch.empty(Cc, device="cuda")
        ws = torch.empty(lib.lib.dv_bn_workspace_bytes(M, Cc), dtype=torch.uint8, device="cuda")
        pa = _params(lib, x, M, Cc, dv_dtype, gamma, beta, rm_a, rv_a, mean_a, inv_a, ws)
        lib.check(lib.lib.dv_bn_stats(C.byref(pa), stream_ptr()), "dv_bn_stats")
        # the allocation-free twin: no workspace passed
        rm_b, rv_b = rm0.clone(), rv0.clone()
        mean_b, inv_b = torch.full((Cc,), float("nan"), device="cuda"), torch.full((Cc,), float("nan"), device="cuda")
        pb = _params(lib, x, M, Cc, dv_dtype, gamma, beta, rm_b, rv_b, mean_b, inv_b, None)
        lib.check(lib.lib.dv_bn_stats_static(C.byref(pb), stream_ptr()), "dv_bn_stats_static")
        torch.cuda.synchronize()
        for name, a, b, tol in (("mean", mean_a, mean_b, 1e-5), ("invstd", inv_a, inv_b, 1e-4),
                                ("running_mean", rm_a, rm_b, 1e-5), ("running_var", rv_a, rv_b, 1e-4)):
            assert torch.isfinite(b).all(), (call, name)
            assert rel_l2(b.cpu().numpy(), a.cpu().numpy()) < tol, (call, name)
        # and against the data itself (fp64 on the host), as the other BN tests do
        xd = x.double()
        assert rel_l2(mean_b.cpu().numpy(), xd.mean(0).cpu().numpy()) < 1e-5
        assert rel_l2(inv_b.cpu().numpy(), (1.0 / torch.sqrt(xd.var(0, unbiased=False) + 1e-5)).cpu().numpy()) < 1e-4


def test_static_scratch_rejects_sizes_beyond_its_capacity():
    from deep_vision_b200 import lib
    Cc, M = 4096, 64
    x = torch.zeros(M, Cc, device="cuda", dtype=torch.bfloat16)
    v = [torch.zeros(Cc, device="cuda") for _ in range(6)]
    p = _params(lib, x, M, Cc, lib.DV_BF16, *v, None)
    rc = lib.lib.dv_bn_stats_static(C.byref(p), stream_ptr())
    assert rc == 2                                      # DV_ERR_UNSUPPORTED
    msg = lib.lib.dv_last_error().decode()
    assert "2048" in msg and "dv_bn_stats" in msg
    torch.cuda.synchronize()
