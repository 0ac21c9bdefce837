"""Per-rank training step: the B200 restatement of `train()` in ResNet/pytorch/train.py:431-485 and of what
`nn.DataParallel` (train.py:353-355) does around it.

Reference step:   output = net(image); loss = CrossEntropyLoss()(output, label); optimizer.zero_grad();
                  loss.backward(); optimizer.step(); loss.item()
with SGD(lr .1, momentum .9, weight_decay 1e-4) for the ResNets (train.py:166-184) and RMSprop(lr .045, alpha .9,
eps 1.0) for MobileNet (train.py:185-214); weight decay applies to every parameter (single param group, :363).

Here: one process per GPU.  All parameters live in ONE flat fp32 buffer and all gradients in another, so that
  * the optimiser is a single fused kernel (dv_sgd_step / dv_rmsprop_step) over the flat buffer,
  * gradient exchange is an NCCL all-reduce (sum, then 1/world folded into the optimiser) over contiguous BUCKETS of
    the flat gradient buffer, launched from autograd hooks as soon as the last gradient of a bucket is produced, so
    the transfer over NVLink overlaps the rest of backward,
  * BN statistics stay local to the rank (the reference's DataParallel replicas never synchronise BN either),
  * no per-step parameter broadcast is needed: replicas start identical and apply identical updates.
The mean cross-entropy over the rank's shard followed by gradient averaging equals DataParallel's mean loss over the
gathered global batch (train.py:452).  The whole step can be captured in a CUDA graph (static shapes).
"""
import ctypes as C

import torch
import torch.distributed as dist
import torch.nn.functional as F

from . import config
from . import functional as dvF
from .lib import check, lib

# train.py:166-214 — the reference's optimiser settings per model
TRAINING_CONFIG = {
    "resnet34": dict(optimizer="sgd", lr=0.1, momentum=0.9, weight_decay=1e-4, batch_size=256),
    "resnet50": dict(optimizer="sgd", lr=0.1, momentum=0.9, weight_decay=1e-4, batch_size=256),
    "resnet152": dict(optimizer="sgd", lr=0.1, momentum=0.9, weight_decay=1e-4, batch_size=128),  # no entry in the reference
    "mobilenet1": dict(optimizer="rmsprop", lr=0.045, alpha=0.9, eps=1.0, weight_decay=0.0, batch_size=128),
}


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def plan_buckets(offsets, total, cap_elems, tail_elems=0):
    """Cut [0,total) into contiguous ranges of >= cap_elems elements (the last may be smaller), walking parameters
    from the END of the flat buffer: backward produces the last layers' gradients first, so bucket 0 is ready first.
    tail_elems > 0: the range that holds the FIRST parameters — whose all-reduce can only start after the very last weight
    gradient of the step (the stem's) and is therefore the one transfer nothing overlaps — is cut once more so that this
    exposed piece carries at most about tail_elems elements; the rest of it goes out as soon as its own gradients are in.
    Returns [(start, end)] in launch order."""
    buckets, end, i = [], total, len(offsets) - 1
    while i >= 0:
        start = end
        while i >= 0 and (end - start) < cap_elems:
            start = offsets[i]
            i -= 1
        buckets.append((start, end))
        end = start
    if tail_elems > 0 and buckets:
        s, e = buckets[-1]
        inner = [o for o in offsets if s < o < e and o - s <= tail_elems]
        if inner:
            cut = max(inner)
            buckets[-1:] = [(cut, e), (s, cut)]
    return buckets


def flat_view(flat, off, shape):
    """The slice of a flat buffer that belongs to a parameter of logical `shape`, shaped like the parameter.  4-D filters
    are stored KRSC (channels_last), so their view is a permutation of the [K,R,S,C] block — NOT `.view(K,C,R,S)`.
    Every place that pairs a parameter with flat optimiser state or gradients goes through this (checkpoint.py too)."""
    n = 1
    for d in shape:
        n *= int(d)
    if len(shape) == 4:
        K, Cc, R, S = (int(d) for d in shape)
        return flat[off:off + n].view(K, R, S, Cc).permute(0, 3, 1, 2)
    return flat[off:off + n].view(tuple(shape))


class FlatParams:
    """All trainable parameters of a module as views into ONE flat fp32 buffer (and their gradients into another).
    4-D filters are laid out KRSC (channels_last) so the convolution kernels read the master copy in place."""

    def __init__(self, module):
        self.params = [p for p in module.parameters() if p.requires_grad]
        dev = self.params[0].device
        total = sum(p.numel() for p in self.params)
        self.flat_p = torch.empty(total, dtype=torch.float32, device=dev)
        self.flat_g = torch.zeros(total, dtype=torch.float32, device=dev)
        self.offsets = []
        off = 0
        for p in self.params:
            n = p.numel()
            pv = flat_view(self.flat_p, off, p.shape)
            gv = flat_view(self.flat_g, off, p.shape)
            pv.copy_(p.data)
            p.data = pv
            p.grad = gv
            self.offsets.append(off)
            off += n
        self.total = total


class BucketedAllReduce:
    """Gradient exchange of the data-parallel step: NCCL (or gloo in the CPU tests) all-reduce SUM over buckets of the
    flat gradient buffer, each launched asynchronously from the autograd hook of the bucket's last-arriving
    parameter; `finish()` makes the current stream wait for all of them."""

    def __init__(self, flat, bucket_bytes=25 << 20, process_group=None, tail_bytes=1 << 20):
        self.flat, self.pg = flat, process_group
        self.buckets = plan_buckets(flat.offsets, flat.total, max(1, bucket_bytes // 4), tail_bytes // 4)
        self._bucket_of, self._need = {}, []
        for b, (s, e) in enumerate(self.buckets):
            members = [j for j, o in enumerate(flat.offsets) if s <= o < e]
            self._need.append(len(members))
            for j in members:
                self._bucket_of[j] = b
        self.reset()
        self.hooks = [self._make_hook(j) for j in range(len(flat.params))]
        for p, h in zip(flat.params, self.hooks):
            p.register_post_accumulate_grad_hook(h)   # gradients that arrive through autograd's accumulation

    def reset(self):
        self._pending = list(self._need)
        self._fired = [False] * len(self.flat.params)
        self._works = []
        self.launch_order = []

    def _make_hook(self, j):
        def hook(_param):
            # A parameter whose gradient is written directly by a kernel reports here itself (functional.grad_sink), and
            # torch may run the post-accumulate hook for it as well (it does so even when backward returned None for the
            # parameter): count each parameter once per step.
            if self._fired[j]:
                return
            self._fired[j] = True
            b = self._bucket_of[j]
            self._pending[b] -= 1
            if self._pending[b] == 0:
                s, e = self.buckets[b]
                self.launch_order.append(b)
                side = getattr(self, "side", None)
                if side is not None and side.used:   # weight gradients of this bucket may still be running on the side stream
                    torch.cuda.current_stream().wait_stream(side.stream)
                self._works.append(dist.all_reduce(self.flat.flat_g[s:e], op=dist.ReduceOp.SUM, group=self.pg, async_op=True))
        return hook

    def finish(self):
        for w in self._works:
            w.wait()
        self._works = []


class Trainer:
    def __init__(self, model, optimizer="sgd", lr=0.1, momentum=0.9, weight_decay=1e-4, alpha=0.9, eps=1.0,
                 bucket_bytes=25 << 20, process_group=None, use_cuda_graph=False, **_ignored):
        self.model = model
        self.kind = optimizer
        self.lr, self.momentum, self.weight_decay, self.alpha, self.eps = lr, momentum, weight_decay, alpha, eps
        self.world = dist.get_world_size(process_group) if (dist.is_available() and dist.is_initialized()) else 1
        self.use_graph = use_cuda_graph
        self.steps_done = 0
        self._graph = None
        if next(model.parameters()).device.type != "cuda":
            raise RuntimeError("Trainer: move the model to a CUDA device first; there is no CPU path")
        self.flat = FlatParams(model)
        self.flat_p, self.flat_g = self.flat.flat_p, self.flat.flat_g
        self.flat_s = torch.zeros_like(self.flat_p)  # momentum buffer / squared-gradient average
        config.bump_weight_epoch()
        self.exchange = BucketedAllReduce(self.flat, bucket_bytes, process_group) if self.world > 1 else None
        self._side = dvF.SideWgrad(self.flat_p.device)
        self._install_direct_grads()
        self._flatten_bn_counters()
        self._install_bf16_filters()

    def _install_bf16_filters(self):
        """One flat bf16 mirror of the parameter buffer, rewritten by the optimiser kernel itself; every dense
        convolution (and the classifier) reads its compute-dtype filters as a view of it, so no per-layer cast kernel runs."""
        from . import nn as dvnn
        self.flat_p16 = None
        if config.compute_dtype() != torch.bfloat16:
            return
        self.flat_p16 = self.flat_p.to(torch.bfloat16)
        index = {id(p): j for j, p in enumerate(self.flat.params)}
        for m in self.model.modules():
            if isinstance(m, dvnn.Conv2d) and not m._dv_depthwise and id(m.weight) in index:
                off = self.flat.offsets[index[id(m.weight)]]
                K, Cc, R, S = m.weight.shape
                view = self.flat_p16[off:off + m.weight.numel()].view(K, R, S, Cc).permute(0, 3, 1, 2)
                m._dv_cache.external = (torch.bfloat16, view)
            elif isinstance(m, dvnn.Linear) and id(m.weight) in index:
                off = self.flat.offsets[index[id(m.weight)]]
                K, Fdim = m.weight.shape
                m._dv_cache.external = (torch.bfloat16, self.flat_p16[off:off + K * Fdim].view(K, Fdim, 1, 1))

    def _install_direct_grads(self):
        """Dense-conv filters and BN affine parameters receive their gradient by a direct kernel write into the flat
        gradient buffer (functional.grad_sink) instead of autograd's temporary + accumulate pass; the callback is the
        exchange's bucket hook (a no-op on one GPU).  Each of these parameters is used once per forward."""
        from . import nn as dvnn
        index = {id(p): j for j, p in enumerate(self.flat.params)}
        noop = lambda _p: None
        for m in self.model.modules():
            targets = []
            if isinstance(m, dvnn.Conv2d):     # dense and depth-wise ([C,1,3,3] is laid out [C][3][3] either way)
                targets = [m.weight]
            elif isinstance(m, dvnn.BatchNorm2d) and m.affine:
                targets = [m.weight, m.bias]
            for p in targets:
                if id(p) in index:
                    p._dv_grad_sink = self.exchange.hooks[index[id(p)]] if self.exchange is not None else noop

    def _flatten_bn_counters(self):
        """nn.BatchNorm2d.num_batches_tracked of every layer as views into one int64 vector: one increment per step."""
        from . import nn as dvnn
        bns = [m for m in self.model.modules() if isinstance(m, dvnn.BatchNorm2d) and m.num_batches_tracked is not None]
        self._nbt = None
        if not bns:
            return
        self._nbt = torch.stack([m.num_batches_tracked.detach().clone() for m in bns])
        for i, m in enumerate(bns):
            m.num_batches_tracked.data = self._nbt[i]
            m._dv_external_nbt = True

    # ---- one step -------------------------------------------------------------------------------------------------
    def _step_body(self, images, labels):
        self.flat_g.zero_()
        if self._nbt is not None:
            self._nbt.add_(1)
        if self.exchange is not None:
            self.exchange.reset()
        logits = self.model(images)
        loss = dvF.cross_entropy(logits, labels)
        side = self._side if config.side_wgrad() else None
        dvF.set_side_wgrad(side)
        if self.exchange is not None:
            self.exchange.side = side
        try:
            loss.backward()
        finally:
            dvF.set_side_wgrad(None)
        if side is not None:
            side.join()
        if self.exchange is not None:
            self.exchange.finish()
        self._optimizer_step()
        return loss.detach()

    def _optimizer_step(self):
        n = self.flat_p.numel()
        scale = 1.0 / self.world
        if self.kind == "sgd":
            first = 1 if self.steps_done == 0 and self._graph is None and not self._capturing else 0
            if self.flat_p16 is not None:
                check(lib.dv_sgd_step_bf16(self.flat_p.data_ptr(), self.flat_g.data_ptr(), self.flat_s.data_ptr(), self.flat_p16.data_ptr(),
                                           n, self.lr, self.momentum, self.weight_decay, scale, first, _stream()), "dv_sgd_step_bf16")
            else:
                check(lib.dv_sgd_step(self.flat_p.data_ptr(), self.flat_g.data_ptr(), self.flat_s.data_ptr(), n, self.lr, self.momentum,
                                      self.weight_decay, scale, first, _stream()), "dv_sgd_step")
        elif self.kind == "rmsprop":
            if self.flat_p16 is not None:
                check(lib.dv_rmsprop_step_bf16(self.flat_p.data_ptr(), self.flat_g.data_ptr(), self.flat_s.data_ptr(),
                                               self.flat_p16.data_ptr(), n, self.lr, self.alpha, self.eps, self.weight_decay, scale,
                                               _stream()), "dv_rmsprop_step_bf16")
            else:
                check(lib.dv_rmsprop_step(self.flat_p.data_ptr(), self.flat_g.data_ptr(), self.flat_s.data_ptr(), n, self.lr, self.alpha,
                                          self.eps, self.weight_decay, scale, _stream()), "dv_rmsprop_step")
        else:
            raise ValueError(f"unknown optimizer {self.kind}")
        config.bump_weight_epoch()

    _capturing = False
    _warmed = False

    def set_lr(self, lr):
        """Learning-rate schedulers call this (epochs.SchedulerTarget).  The rate is a by-value kernel argument, so a
        captured step is re-recorded on the next call to `step` (once per change, i.e. at most once per epoch)."""
        lr = float(lr)
        if lr != self.lr:
            self.lr = lr
            self._graph = None

    def step(self, images, labels):
        """images: [B,3,H,W] fp32 CUDA (NCHW, as the reference's loader yields), labels: [B] int64 CUDA.
        Returns the loss as a 0-d CUDA tensor (no host sync; call .item() to reproduce train.py:470)."""
        self.model.train()
        if not self.use_graph:
            loss = self._step_body(images, labels)
            self.steps_done += 1
            return loss
        if self._graph is None:
            self._capture(images, labels)
        if images.shape != self._static_images.shape or images.dtype != self._static_images.dtype:
            # A batch of another shape — the ragged last batch of an epoch: the reference's DataLoader has no drop_last
            # (train.py:229-234; 1,281,167 % 256 = 143) and trains it.  It runs eagerly through the same kernels; the optimiser
            # takes lr by value and the flat state is shared with the captured step, so nothing is re-recorded.
            loss = self._step_body(images, labels)
            self.steps_done += 1
            return loss
        self._static_images.copy_(images, non_blocking=True)
        self._static_labels.copy_(labels, non_blocking=True)
        self._graph.replay()
        self.steps_done += 1
        # the replayed kernels rewrote parameters and running statistics through raw pointers: nothing moved a Tensor._version,
        # so caches keyed on versions (padded stem filters, folded BatchNorm scale/shift of the inference path) must be told
        config.bump_weight_epoch()
        return self._static_loss

    def _capture(self, images, labels):
        """Warm up eagerly on a side stream (allocator, filter caches, NCCL), then record fwd+bwd+exchange+optimiser
        into one CUDA graph.  The two warm-up steps are NOT training steps: parameters, optimiser state, BN running
        statistics and counters are put back afterwards, so the first replay is the one update this batch gets (the
        reference updates once per batch, train.py:457-467).  With the state buffer at zero the recorded update
        buf = momentum*buf + g equals torch's first-step special case buf = g."""
        self._static_images = images.clone()
        self._static_labels = labels.clone()
        if not self._warmed:   # a re-capture (set_lr) records straight away: nothing is executed
            s = torch.cuda.Stream()
            s.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(s):
                keep = [self.flat_p.clone(), self.flat_s.clone()] + [b.detach().clone() for b in self.model.buffers()]
                for _ in range(2):
                    self._step_body(self._static_images, self._static_labels)
                self.flat_p.copy_(keep[0])
                self.flat_s.copy_(keep[1])
                for b, saved in zip(self.model.buffers(), keep[2:]):
                    b.detach().copy_(saved)
                if self.flat_p16 is not None:
                    self.flat_p16.copy_(self.flat_p)
                config.bump_weight_epoch()
            torch.cuda.current_stream().wait_stream(s)
            self._warmed = True
        torch.cuda.synchronize()
        self._graph = torch.cuda.CUDAGraph()
        self._capturing = True
        try:
            with torch.cuda.graph(self._graph):
                self._static_loss = self._step_body(self._static_images, self._static_labels)
        finally:
            self._capturing = False

    def refresh_from_parameters(self):
        """After parameters were written through torch (load_state_dict, manual edits): rebuild the bf16 mirror."""
        if getattr(self, "flat_p16", None) is not None:
            self.flat_p16.copy_(self.flat_p)
        config.bump_weight_epoch()

    def state_dict(self):
        """The optimiser state in torch.optim's own layout, so that the reference's `optimizer.load_state_dict(
        checkpoint['optimizer'])` (train.py:295) accepts a file written here: per parameter index (net.parameters() order)
        the momentum buffer (SGD) or squared-gradient average (RMSprop) as an OIHW fp32 tensor, plus one param_group with the
        hyper-parameters.  `load_checkpoint` reads the same layout back, whoever wrote it."""
        key = {"sgd": "momentum_buffer", "rmsprop": "square_avg"}[self.kind]
        state = {}
        if self.steps_done > 0:
            for idx, (p, off) in enumerate(zip(self.flat.params, self.flat.offsets)):
                entry = {key: flat_view(self.flat_s, off, p.shape).detach().contiguous().cpu().clone()}
                if self.kind == "rmsprop":
                    entry["step"] = torch.tensor(float(self.steps_done))
                state[idx] = entry
        if self.kind == "sgd":
            group = dict(lr=self.lr, momentum=self.momentum, dampening=0, weight_decay=self.weight_decay, nesterov=False,
                         maximize=False, foreach=None, differentiable=False, fused=None)
        else:
            group = dict(lr=self.lr, momentum=0, alpha=self.alpha, eps=self.eps, centered=False, weight_decay=self.weight_decay,
                         capturable=False, foreach=None, maximize=False, differentiable=False)
        group["params"] = list(range(len(self.flat.params)))
        group["dv_steps"] = int(self.steps_done)   # torch carries unknown group keys along untouched
        return {"state": state, "param_groups": [group]}


class HostFeeder:
    """Host-to-device input path of the training loop (train.py:444-446 moves every batch with `.to(device)`):
    pinned host batches are copied on a side stream into one of two device buffers while the previous step computes, so
    the PCIe transfer of batch i+1 hides behind step i.

        feeder = HostFeeder(trainer)
        feeder.prefetch(images_host, labels_host)            # batch 0
        for i in range(steps):
            ready = feeder.take()                            # waits for batch i on the compute stream
            feeder.prefetch(next_images_host, next_labels_host)
            loss = trainer.step(*ready).item()
    """

    def __init__(self, trainer):
        self.trainer = trainer
        self.stream = torch.cuda.Stream()
        self.slots = [None, None]
        self.events = [torch.cuda.Event(), torch.cuda.Event()]
        self.free_events = [None, None]
        self.head = 0   # next slot to fill
        self.tail = 0   # next slot to hand out

    def prefetch(self, images_host, labels_host):
        i = self.head
        if self.slots[i] is None:
            dev = self.trainer.flat_p.device
            self.slots[i] = (torch.empty(images_host.shape, dtype=images_host.dtype, device=dev),
                             torch.empty(labels_host.shape, dtype=labels_host.dtype, device=dev))
        with torch.cuda.stream(self.stream):
            if self.free_events[i] is not None:
                self.stream.wait_event(self.free_events[i])   # the step that read this slot has been enqueued and finished
            self.slots[i][0].copy_(images_host, non_blocking=True)
            self.slots[i][1].copy_(labels_host, non_blocking=True)
            self.events[i].record(self.stream)
        self.head ^= 1

    def take(self):
        i = self.tail
        torch.cuda.current_stream().wait_event(self.events[i])
        self.tail ^= 1
        self._last = i
        return self.slots[i]

    def release(self):
        """Call after the step that consumed the last `take()` has been enqueued."""
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream())
        self.free_events[self._last] = ev
