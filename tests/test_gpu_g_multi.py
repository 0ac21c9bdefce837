"""Tests of the data-parallel step on real GPUs: tools/dp_check.py under torchrun — replicas identical after a step, reduced
gradient == sum of per-shard plain-autograd gradients, update == SGD on their average (reference
semantics train.py:353-355).  Two ranks over NCCL when the box has two GPUs (eager and CUDA graph); on a single GPU the same
two-rank run goes through gloo with both ranks on device 0, so the exchange logic is exercised in every round-end run."""
import json
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu


def test_two_ranks_on_one_gpu_through_gloo_equal_averaged_shard_gradients():
    """(e) on a 1-GPU box: fixed-order mode, so the reduced gradient must equal the sum of the two per-shard plain-autograd
    gradients to rounding, the update must be SGD on their average, and both replicas must end bit-identical."""
    out = _run_dp_check(29535, DP_CHECK_BACKEND="gloo")
    assert out["ranks_identical_graph0"] and out["debug"]["flat_g_identical"] and out["debug"]["p0_identical"]
    assert out["debug"]["world"] == 2 and out["debug"]["pending_after"] == [0] * out["debug"]["buckets"]
    assert out["plain_autograd_twice_rel_diff"] == 0.0
    assert out["grad_sum_rel_err"] < 1e-6 and out["update_rel_err"] < 1e-5


def test_two_ranks_on_one_gpu_default_mode_keeps_replicas_identical():
    """The benchmarked mode (fp32 atomics / TMA reduce-add, weight gradients on the side stream) through the same two-rank run:
    every bucket is exchanged exactly once and after all of its gradients, so both ranks hold bit-identical reduced buffers and
    parameters (each rank's own summation order differs, the all-reduce output does not)."""
    out = _run_dp_check(29536, DP_CHECK_BACKEND="gloo", DP_CHECK_ATOMICS="1")
    assert out["ranks_identical_graph0"] and out["debug"]["flat_g_identical"]
    assert out["debug"]["pending_after"] == [0] * out["debug"]["buckets"]


def _run_dp_check(port, **env):
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(root, "tools", "dp_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=root, env=dict(os.environ, **env))
    lines = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
    assert r.returncode == 0 and lines, r.stderr[-2000:]
    return json.loads(lines[-1])


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_two_rank_default_mode_keeps_replicas_identical():
    """The default configuration (fp32 atomics, weight gradients on the side stream): every bucket is exchanged once, after all of
    its gradients — the reduced buffers and the updated parameters are bit-identical on both ranks, eager and captured."""
    out = _run_dp_check(29534, DP_CHECK_ATOMICS="1")
    assert out["ranks_identical_graph0"] and out["ranks_identical_graph1"] and out["debug"]["flat_g_identical"]
    assert out["debug"]["pending_after"] == [0] * out["debug"]["buckets"]


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_two_rank_step_equals_averaged_shard_gradients():
    out = _run_dp_check(29533)
    assert out["ranks_identical_graph0"] and out["ranks_identical_graph1"]
    assert out["debug"]["pending_after"] == [0] * out["debug"]["buckets"]
    assert out["plain_autograd_twice_rel_diff"] == 0.0          # fixed-order mode is bit-reproducible
    assert out["grad_sum_rel_err"] < 1e-6 and out["update_rel_err"] < 1e-5
