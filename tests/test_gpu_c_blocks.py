"""GPU parity of whole blocks (the product's nn.Modules, i.e. autograd Functions over the C ABI) against the golden
fixtures frozen from the reference's own BasicBlock / BottleneckBlock / DepthwiseSeparableConv.

fp32: output, input gradient and every parameter gradient <= 1e-4 relative L2 (north_star).  bf16: forward <= 2e-2;
bf16 gradients are reported by torch's own autocast at 6-11e-2 from fp64 and gated at 1.5e-1 here.
"""
import numpy as np
import pytest
import torch

import deep_vision_b200 as dv
from helpers import golden_state, load_golden, rel_l2

pytestmark = pytest.mark.gpu

CASES = {
    "block_bottleneck_256_64_256": lambda: dv.models.BottleneckBlock(256, 64, 256),
    "block_bottleneck_128_64_256_s2_proj": lambda: dv.models.BottleneckBlock(128, 64, 256, stride=2, downsample=True),
    "block_bottleneck_64_64_256_s1_proj": lambda: dv.models.BottleneckBlock(64, 64, 256, stride=1, downsample=True),
    "block_basic_64_64": lambda: dv.models.BasicBlock(64, 64),
    "block_basic_64_128_s2_proj": lambda: dv.models.BasicBlock(64, 128, stride=2, downsample=True),
    "block_dwsep_64_128_s2": lambda: dv.models.DepthwiseSeparableConv(64, 128, dw_stride=2, pw_stride=1),
    "block_dwsep_32_64_s1": lambda: dv.models.DepthwiseSeparableConv(32, 64, dw_stride=1, pw_stride=1),
}


@pytest.fixture(autouse=True)
def _restore_dtype():
    yield
    dv.set_compute_dtype(torch.bfloat16)
    dv.set_conv_algo(0)


def _run(name, dtype, algo=0):
    rec = load_golden(name)
    dv.set_compute_dtype(dtype)
    dv.set_conv_algo(algo)
    m = CASES[name]()
    m.load_state_dict(golden_state(rec))
    m = m.cuda().train()
    x = torch.from_numpy(rec["x"]).cuda().requires_grad_(True)
    y = m(x)
    assert y.dtype == dtype and y.is_contiguous(memory_format=torch.channels_last)
    y.backward(torch.from_numpy(rec["dy"]).cuda().to(dtype))
    torch.cuda.synchronize()
    out = {"y": rel_l2(y.detach().float().cpu().numpy(), rec["y"]), "dx": rel_l2(x.grad.float().cpu().numpy(), rec["dx"])}
    for k, p in m.named_parameters():
        out["grad/" + k] = rel_l2(p.grad.float().cpu().numpy(), rec["grad/" + k])
    for k, v in m.state_dict().items():
        if "running_" in k:
            out["after/" + k] = rel_l2(v.float().cpu().numpy(), rec["after/" + k])
    return out


@pytest.mark.parametrize("name", sorted(CASES))
def test_block_fp32_matches_reference(name):
    errs = _run(name, torch.float32)
    bad = {k: v for k, v in errs.items() if v > 1e-4}
    assert not bad, bad


@pytest.mark.parametrize("algo", [0, 1], ids=["auto_tc", "simt"])
@pytest.mark.parametrize("name", sorted(CASES))
def test_block_bf16_matches_reference(name, algo):
    errs = _run(name, torch.bfloat16, algo)
    assert errs["y"] < 2e-2, errs
    # running statistics: fp32 arithmetic, but taken over the bf16-rounded conv output (as torch autocast does)
    assert all(v < 2e-2 for k, v in errs.items() if k.startswith("after/")), errs
    assert errs["dx"] < 1.5e-1 and all(v < 1.5e-1 for k, v in errs.items() if k.startswith("grad/")), errs


@pytest.mark.parametrize("name,fused_layers", [("block_bottleneck_256_64_256", 2), ("block_bottleneck_128_64_256_s2_proj", 2),
                                                ("block_basic_64_64", 1), ("block_dwsep_32_64_s1", 1)])
def test_block_bf16_fused_bn_backward_sums_equal_the_reduce_pass(name, fused_layers):
    """With set_fuse_bn_bwd(True) (default) the dgrad of conv2 / conv3 / the point-wise conv produces dgamma, dbeta of the
    BatchNorm in front of it, and that BatchNorm's reduce + finalize launches disappear.  Same gradients as with the separate
    pass (fixed forward statistics, so the two runs see identical activations), two launches fewer per fused layer."""
    from deep_vision_b200.lib import lib
    rec = load_golden(name)
    dv.set_compute_dtype(torch.bfloat16)
    dv.set_fuse_bn_stats(False)        # bit-reproducible forward: both runs back-propagate through the same tensors
    grads, launches = {}, {}
    try:
        for fuse in (False, True):
            dv.set_fuse_bn_bwd(fuse)
            m = CASES[name]()
            m.load_state_dict(golden_state(rec))
            m = m.cuda().train()
            x = torch.from_numpy(rec["x"]).cuda().requires_grad_(True)
            n0 = lib.dv_launch_count()
            y = m(x)
            y.backward(torch.from_numpy(rec["dy"]).cuda().to(torch.bfloat16))
            torch.cuda.synchronize()
            launches[fuse] = lib.dv_launch_count() - n0
            grads[fuse] = {k: p.grad.detach().double().cpu().numpy() for k, p in m.named_parameters()}
            grads[fuse]["dx"] = x.grad.double().cpu().numpy()
            fused = [k for k, mod in m.named_modules() if getattr(mod, "_dv_bwd_sums_fused", False)]
            assert len(fused) == (fused_layers if fuse else 0), fused
    finally:
        dv.set_fuse_bn_bwd(True)
        dv.set_fuse_bn_stats(True)
    assert launches[False] - launches[True] == 2 * fused_layers, launches
    for k in grads[True]:
        assert rel_l2(grads[True][k], grads[False][k]) < 2e-3, k     # bf16 dx downstream of an fp32 sum that moved in its last bits


@pytest.mark.parametrize("name", ["block_bottleneck_256_64_256", "block_basic_64_64"])
def test_block_bf16_shortcut_gradient_in_dgrad_epilogue_equals_the_accumulate_path(name):
    """set_fuse_shortcut_grad(True) (opt-in, DV_FUSE_SHORTCUT=1): in an identity block the last BatchNorm's backward no longer writes the masked
    gradient of the shortcut; conv1's dgrad adds `dout * (out > 0)` in its epilogue (dv_conv_params.ep_residual[_mask]).  Every
    parameter gradient is bit-identical (none depends on dx); dx agrees to one bf16 rounding (the accumulate path rounds the
    dgrad before adding)."""
    rec = load_golden(name)
    dv.set_compute_dtype(torch.bfloat16)
    dv.set_deterministic(True)       # fixed-order reductions: the two runs differ in the shortcut handling only
    out = {}
    try:
        for fuse in (False, True):
            dv.set_fuse_shortcut_grad(fuse)
            m = CASES[name]()
            m.load_state_dict(golden_state(rec))
            m = m.cuda().train()
            x = torch.from_numpy(rec["x"]).cuda().requires_grad_(True)
            y = m(x)
            y.backward(torch.from_numpy(rec["dy"]).cuda().to(torch.bfloat16))
            torch.cuda.synchronize()
            out[fuse] = {k: p.grad.detach().clone() for k, p in m.named_parameters()}
            out[fuse]["dx"] = x.grad.detach().clone()
    finally:
        dv.set_fuse_shortcut_grad(False)
        dv.set_deterministic(False)
    for k in out[True]:
        if k != "dx":
            assert torch.equal(out[True][k], out[False][k]), k
    assert rel_l2(out[True]["dx"].double().cpu().numpy(), out[False]["dx"].double().cpu().numpy()) < 4e-3
    assert rel_l2(out[True]["dx"].float().cpu().numpy(), rec["dx"]) < 1.5e-1
