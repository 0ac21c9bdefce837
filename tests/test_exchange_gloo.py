"""CPU tests (gloo, world_size 2) of the data-parallel host logic: flat parameter/gradient buffers, bucket planning,
hook-driven bucketed all-reduce.  The semantics checked are the reference's DataParallel step (train.py:353-355,452):
mean loss over the global batch == average of the per-shard gradients, replicas stay identical."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp
import torch.nn as nn

from deep_vision_b200.train import BucketedAllReduce, FlatParams, plan_buckets


def test_plan_buckets_covers_everything_back_to_front():
    offsets = [0, 100, 150, 400, 1000]
    total = 1200
    b = plan_buckets(offsets, total, 300)
    assert b[0][1] == total and b[-1][0] == 0
    assert all(b[i][0] == b[i + 1][1] for i in range(len(b) - 1))         # contiguous, descending
    assert all(s in offsets for s, _ in b)                                    # cut only at parameter boundaries
    assert all((e - s) >= 300 for s, e in b[:-1])
    assert plan_buckets([0], 10, 1 << 30) == [(0, 10)]
    assert plan_buckets([0, 5], 10, 1) == [(5, 10), (0, 5)]
    # the exposed tail: the range with the first parameters is cut so that what waits for the last gradient is small
    assert plan_buckets(offsets, total, 600) == [(400, 1200), (0, 400)]
    assert plan_buckets(offsets, total, 600, tail_elems=160) == [(400, 1200), (150, 400), (0, 150)]
    assert plan_buckets(offsets, total, 600, tail_elems=50) == [(400, 1200), (0, 400)]   # no parameter boundary within reach


def _tiny():
    torch.manual_seed(0)
    return nn.Sequential(nn.Conv2d(3, 8, 3, padding=1, bias=False), nn.BatchNorm2d(8), nn.ReLU(), nn.Conv2d(8, 4, 1, bias=False),
                         nn.AdaptiveAvgPool2d(1), nn.Flatten(), nn.Linear(4, 5))


def test_flat_params_keeps_values_and_uses_krsc_for_filters():
    m = _tiny()
    before = {k: v.clone() for k, v in m.state_dict().items()}
    flat = FlatParams(m)
    for k, v in m.state_dict().items():
        assert torch.equal(v, before[k]), k
    w = m[0].weight
    assert w.shape == (8, 3, 3, 3) and w.is_contiguous(memory_format=torch.channels_last)
    assert w.grad is not None and w.grad.stride() == w.stride()
    # the views alias the flat buffers
    flat.flat_p.zero_()
    assert float(m[6].weight.abs().sum()) == 0.0
    assert flat.total == sum(p.numel() for p in m.parameters())


def _shard(rank):
    g = torch.Generator().manual_seed(100 + rank)
    return torch.randn(4, 3, 6, 6, generator=g), torch.randint(0, 5, (4,), generator=g)


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        m = _tiny()
        flat = FlatParams(m)
        ex = BucketedAllReduce(flat, bucket_bytes=256, tail_bytes=0)   # tiny buckets -> several collectives (no tail cut: a 1-parameter bucket may overtake)
        x, y = _shard(rank)
        flat.flat_g.zero_()
        ex.reset()
        nn.functional.cross_entropy(m(x), y).backward()
        ex.finish()
        q.put((rank, flat.flat_g.clone() / world, list(ex.launch_order), len(ex.buckets)))
    finally:
        dist.destroy_process_group()


def test_bucketed_allreduce_world2_equals_global_batch_gradient():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    results = sorted([q.get(timeout=120) for _ in procs], key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    (_, g0, order0, nb), (_, g1, order1, _) = results
    assert nb > 1 and sorted(order0) == list(range(nb)) and order0 == order1
    assert order0[0] == 0, "the bucket holding the last layers must be exchanged first"
    assert torch.equal(g0, g1)                     # replicas see the same averaged gradient
    # DataParallel semantics with per-replica BN: average of the per-shard gradients
    expect = None
    for r in range(2):
        m = _tiny()
        x, y = _shard(r)
        nn.functional.cross_entropy(m(x), y).backward()
        g = torch.cat([p.grad.permute(0, 2, 3, 1).reshape(-1) if p.dim() == 4 else p.grad.reshape(-1) for p in m.parameters()])
        expect = g if expect is None else expect + g
    expect = expect / 2
    assert torch.allclose(g0, expect, rtol=1e-5, atol=1e-7)


def test_hooks_count_each_parameter_once_per_step():
    """A parameter whose gradient a kernel writes directly reports to the exchange itself, and torch runs the
    post-accumulate hook for it too (even when backward returned None for it).  The second report must not count: it used to
    release a bucket's all-reduce after HALF of its gradients (found on 2 GPUs with tools/dp_check.py)."""
    m = _tiny()
    flat = FlatParams(m)
    ex = BucketedAllReduce.__new__(BucketedAllReduce)      # host logic only: no process group, collectives stubbed
    ex.flat, ex.pg = flat, None
    ex.buckets = plan_buckets(flat.offsets, flat.total, 64)
    ex._bucket_of, ex._need = {}, []
    for b, (s, e) in enumerate(ex.buckets):
        members = [j for j, o in enumerate(flat.offsets) if s <= o < e]
        ex._need.append(len(members))
        for j in members:
            ex._bucket_of[j] = b
    launched = []
    import deep_vision_b200.train as tr

    class _Work:
        def wait(self):
            pass

    orig = tr.dist.all_reduce
    tr.dist.all_reduce = lambda t, **kw: (launched.append((t.data_ptr(), t.numel())), _Work())[1]
    try:
        ex.reset()
        hooks = [ex._make_hook(j) for j in range(len(flat.params))]
        order = list(reversed(range(len(hooks))))           # backward visits parameters last to first
        for n, j in enumerate(order):
            hooks[j](flat.params[j])
            hooks[j](flat.params[j])                         # the duplicate report
            done = {ex._bucket_of[k] for k in order[:n + 1]}
            # a bucket is out only when ALL of its members have reported once
            for b in ex.launch_order:
                assert all(k in order[:n + 1] for k, bb in ex._bucket_of.items() if bb == b)
            assert set(ex.launch_order) <= done
        assert sorted(ex.launch_order) == list(range(len(ex.buckets))) and ex._pending == [0] * len(ex.buckets)
        assert len(launched) == len(ex.buckets)
    finally:
        tr.dist.all_reduce = orig
