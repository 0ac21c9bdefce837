"""torch.autograd.Functions over the C ABI (include/dv.h): the replacement for the torch.nn primitives on the
reference's conv-BN-ReLU hot path (ResNet/pytorch/models/*.py, MobileNet/pytorch/models/mobilenet_v1.py).

Tensors are ordinary torch CUDA tensors with logical shape [N,C,H,W] and ``torch.channels_last`` strides, i.e. NHWC in
memory; torch only owns the memory and the autograd graph, every byte of arithmetic is done by libdv_b200.so on the
current CUDA stream.  Nothing here falls back to torch ops or to the CPU.
"""
import ctypes as C
from os import environ as _environ

import torch

from . import config
from .lib import (DV_ALGO_AUTO, DV_ALGO_SIMT, DV_ALGO_TC, DV_BF16, DV_CONV_DGRAD, DV_CONV_FPROP, DV_CONV_WGRAD, DV_F32,
                  BnBwdParams, BnParams, BnPoolParams, ConvParams, DwConvParams, StemParams, check, lib)

_CL = torch.channels_last


def _dv_dtype(t):
    if t == torch.float32:
        return DV_F32
    if t == torch.bfloat16:
        return DV_BF16
    raise TypeError(f"deep_vision_b200 computes in float32 or bfloat16, got {t}")


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _require_cuda(t, what):
    if not t.is_cuda:
        raise RuntimeError(f"{what}: deep_vision_b200 runs on CUDA (sm_100a) tensors only; got a {t.device} tensor. "
                           "There is no CPU path.")


def _nhwc(t):
    """channels_last-contiguous view/copy of a 4-D tensor (a copy only when the layout is wrong)."""
    return t if t.is_contiguous(memory_format=_CL) else t.contiguous(memory_format=_CL)


_workspaces = {}


def _workspace(nbytes, device):
    """Scratch buffer per (device, stream); grown geometrically, never shrunk.  Kernels that share it are ordered by
    the stream they run on."""
    key = (device.index, torch.cuda.current_stream(device).cuda_stream)
    buf = _workspaces.get(key)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(max(int(nbytes), 1 << 20), dtype=torch.uint8, device=device)
        _workspaces[key] = buf
    return buf


# ----------------------------------------------------------------------------------------------------------------
# layout change at the model entrance
# ----------------------------------------------------------------------------------------------------------------
class _NoCtx:
    """Stand-in context for calling an autograd Function's forward as a plain function (nothing is saved for backward)."""


class _ToNHWC(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, dtype):
        ctx.in_dtype = x.dtype
        N, Cc, H, W = x.shape
        if x.dtype == torch.float32 and x.is_contiguous():
            y = torch.empty((N, Cc, H, W), dtype=dtype, device=x.device, memory_format=_CL)
            check(lib.dv_nchw_to_nhwc(x.data_ptr(), y.data_ptr(), N, Cc, H, W, _dv_dtype(dtype), _stream()), "dv_nchw_to_nhwc")
            return y
        if x.dtype == dtype:
            return x.contiguous(memory_format=_CL)
        # rare entrance (e.g. a half-precision or already channels_last image batch): cast, then lay out
        xc = x.contiguous(memory_format=_CL)
        y = torch.empty_like(xc, dtype=dtype)
        check(lib.dv_cast(xc.data_ptr(), _dv_dtype(xc.dtype), y.data_ptr(), _dv_dtype(dtype), xc.numel(), _stream()), "dv_cast")
        return y

    @staticmethod
    def backward(ctx, dy):
        if dy.dtype == ctx.in_dtype:
            return dy, None
        dyc = dy.contiguous(memory_format=_CL)
        out = torch.empty_like(dyc, dtype=ctx.in_dtype)
        check(lib.dv_cast(dyc.data_ptr(), _dv_dtype(dyc.dtype), out.data_ptr(), _dv_dtype(ctx.in_dtype), dyc.numel(), _stream()), "dv_cast")
        return out, None


def to_compute_layout(x, dtype=None):
    """NCHW fp32 batch (what train.py:444-449 feeds the model) -> NHWC compute dtype."""
    dtype = dtype or config.compute_dtype()
    _require_cuda(x, "input")
    if x.dtype == dtype and x.is_contiguous(memory_format=_CL):
        return x
    return _ToNHWC.apply(x, dtype)


# ----------------------------------------------------------------------------------------------------------------
# dense convolution
# ----------------------------------------------------------------------------------------------------------------
class FilterCache:
    """Compute-dtype copies of one fp32 filter tensor, rebuilt when the parameter's version counter moves
    (optimizer step, load_state_dict, init)."""

    def __init__(self):
        self.key = None
        self.w = None
        self.w_t = None
        self.pad_key = None
        self.w_pad = None
        self.external = None   # (dtype, tensor): a copy some other kernel keeps current (the trainer's fused optimiser)
        self.stem_key = None
        self.w_stem = None

    def get_stem(self, weight, sp):
        """[64][R][32] bf16 filters in the packed K layout of the direct stem kernels (dv_stem_pack_filters)."""
        key = (weight.data_ptr(), weight._version, config.weight_epoch(), weight.device)
        if key != self.stem_key:
            wm = weight.detach()
            if not wm.is_contiguous(memory_format=_CL):
                wm = wm.contiguous(memory_format=_CL)
            if self.w_stem is None or self.w_stem.device != weight.device:
                self.w_stem = torch.empty((64, weight.shape[2], 32), dtype=torch.bfloat16, device=weight.device)
            sp.w = self.w_stem.data_ptr()
            check(lib.dv_stem_pack_filters(C.byref(sp), wm.data_ptr(), _stream()), "dv_stem_pack_filters")
            self.stem_key = key
        return self.w_stem

    def get_padded(self, weight, dtype, kp):
        """[K][Kp] compute-dtype filters (rows zero-padded from R*S*C to Kp) for the patch-matrix path of the stems."""
        K, Cc, R, S = weight.shape
        key = (weight.data_ptr(), weight._version, config.weight_epoch(), dtype, weight.device, kp)
        if key != self.pad_key:
            wm = weight.detach()
            if not wm.is_contiguous(memory_format=_CL):
                wm = wm.contiguous(memory_format=_CL)
            self.w_pad = torch.empty((K, kp), dtype=dtype, device=weight.device)
            check(lib.dv_pad_rows(wm.data_ptr(), self.w_pad.data_ptr(), K, R * S * Cc, kp, _dv_dtype(dtype), _stream()), "dv_pad_rows")
            self.pad_key = key
        return self.w_pad

    def get(self, weight, dtype, need_t):
        K, Cc, R, S = weight.shape
        if self.external is not None and self.external[0] == dtype and not need_t:
            return self.external[1], None
        key = (weight.data_ptr(), weight._version, config.weight_epoch(), dtype, weight.device)
        if key != self.key or (need_t and self.w_t is None):
            wm = weight.detach()
            if not wm.is_contiguous(memory_format=_CL):
                wm = wm.contiguous(memory_format=_CL)
            if dtype == torch.float32:
                self.w, self.w_t = wm, None
            else:
                self.w = torch.empty((K, Cc, R, S), dtype=dtype, device=weight.device, memory_format=_CL)
                self.w_t = torch.empty((Cc, K, R, S), dtype=dtype, device=weight.device, memory_format=_CL) if need_t else None
                check(lib.dv_filter_prep(wm.data_ptr(), self.w.data_ptr(), self.w_t.data_ptr() if need_t else None,
                                         K, R, S, Cc, _dv_dtype(dtype), _stream()), "dv_filter_prep")
            self.key = key
        return self.w, self.w_t


def _conv_params(x_ptr, w_ptr, y_ptr, wt_ptr, N, H, W, Cc, K, R, S, stride, pad, dtype, algo, y_f32=0, accumulate=0):
    p = ConvParams()
    p.x, p.w, p.y, p.w_t = x_ptr, w_ptr, y_ptr, wt_ptr
    p.N, p.H, p.W, p.C, p.K, p.R, p.S, p.stride, p.pad = N, H, W, Cc, K, R, S, stride, pad
    p.dtype, p.algo, p.y_f32, p.accumulate = dtype, algo, y_f32, accumulate
    return p


def grad_sink(param):
    """(tensor, callback) when the trainer has registered `param` for DIRECT gradient writes: the backward kernel then
    writes its result straight into the parameter's slice of the flat gradient buffer (no temporary, no accumulate
    pass) and the callback tells the gradient exchange that the slice is final.  None otherwise (plain autograd)."""
    sink = getattr(param, "_dv_grad_sink", None)
    if sink is None or param.grad is None:
        return None
    return param.grad, sink


class StatSums:
    """[2][K] fp32 accumulator a convolution's epilogue adds sum(y), sum(y^2) to, owned by the conv module and handed
    to the BatchNorm that follows through the activation tensor (`y._dv_stat_sums`).  `pending` guards against a
    producer whose sums were never consumed (they are then cleared before the next use)."""

    __slots__ = ("t", "pending")

    def __init__(self, k, device):
        self.t = torch.zeros(2 * k, dtype=torch.float32, device=device)
        self.pending = False


class GradSlot:
    """Hand-over point for the two gradients of a tensor with two consumers (the input of a residual block feeds the
    first convolution AND the shortcut: resnet50.py:147-163).  The shortcut side DEPOSITS its gradient here instead of
    returning it to autograd; the main-path convolution COLLECTS it and lets its dgrad kernel accumulate on top
    (dv_conv_params.accumulate), so the sum costs no extra pass over the tensor.  Autograd runs nodes in reverse creation
    order, and the blocks create the shortcut's nodes after the main path's, so the deposit always comes first; a
    collect without a deposit raises instead of silently dropping a gradient."""

    __slots__ = ("t", "lazy_ok")

    def __init__(self):
        self.t = None
        # set by the collecting convolution at forward time: its dgrad can add `dout * (out > 0)` in its own epilogue
        # (dv_conv_params.ep_residual / ep_residual_mask), so the depositor hands over (dout, mask) instead of a masked copy
        self.lazy_ok = False


class BnBwdHandoff:
    """What the dgrad of a dense convolution needs in order to produce, in its epilogue, the backward sums of the BatchNorm
    whose output is that convolution's input (conv1 -> bn1 -> relu -> conv2, resnet50.py:152-158: conv2's dgrad computes dy of
    bn1 and can sum dz and dz*xhat while the tile is on chip; dv_conv_params.bnb_*).  Filled by _BatchNormAct.forward for a
    train-mode bf16 BatchNorm with ReLU and without residual, attached to its output tensor, claimed by the convolution
    that consumes it.  `claims` counts consumers seen at forward time: with more than one the gradient of the tensor is a
    sum of several dgrads and nothing is fused.  In backward the convolution leaves (dx, dgamma, dbeta) here; the BatchNorm
    uses the sums only if the dy it receives IS that dx (autograd summed nothing else onto it)."""

    __slots__ = ("x", "gamma", "beta", "mean", "invstd", "claims", "result", "owner")

    def __init__(self, owner=None):
        self.x = self.gamma = self.beta = self.mean = self.invstd = self.result = None
        self.claims = 0
        self.owner = owner


def _run_conv(fn, p, pass_id, device, what):
    need = lib.dv_conv_workspace_bytes(C.byref(p), pass_id)
    if need:
        ws = _workspace(need, device)
        p.workspace, p.workspace_bytes = ws.data_ptr(), ws.numel()
    check(fn(C.byref(p), _stream()), what)


class SideWgrad:
    """Weight-gradient kernels on a second stream.  A layer's wgrad depends only on tensors that exist before its dgrad
    starts and nothing in the backward chain waits for it, so the trainer lets it run beside the chain
    dgrad -> BN backward -> dgrad: a tensor-bound 3x3 wgrad then shares the SMs with the HBM-bound BatchNorm passes.
    `fork` makes the side stream wait for the producer of dy; the tensors it reads are kept alive until `join`, which
    the trainer calls once after backward (before the gradient exchange finishes and the optimiser runs)."""

    def __init__(self, device):
        self.stream = torch.cuda.Stream(device=device)
        self.keep = []
        self.used = False

    def fork(self, *tensors):
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream())
        self.stream.wait_event(ev)
        self.keep.extend(tensors)
        self.used = True
        return self.stream

    def join(self):
        if self.used:
            torch.cuda.current_stream().wait_stream(self.stream)
        self.keep.clear()
        self.used = False


_side_wgrad = {"obj": None}


def set_side_wgrad(obj):
    """Installed by the Trainer around loss.backward(); None = weight gradients on the current stream."""
    _side_wgrad["obj"] = obj


class _Conv2d(torch.autograd.Function):
    """y = conv2d(x, weight) without bias — nn.Conv2d at resnet50.py:20-28,101-143; resnet34.py:20-28,96-126;
    mobilenet_v1.py:30,139-145.  Backward = dgrad + wgrad (autograd of train.py:460)."""

    @staticmethod
    def forward(ctx, x, weight, stride, pad, cache, algo, slot=None, role=None, stats=None, handoff=None):
        ctx.slot, ctx.role = slot, role
        ctx.handoff = handoff
        if handoff is not None:
            handoff.claims += 1
        _require_cuda(x, "conv2d")
        # raw: the fp32 NCHW batch itself (nn.Conv2d.forward hands it over for a C <= 4 stem): the direct path below lays it out,
        # rounds it and pads it in one pass; every other path converts it first, as to_compute_layout would have
        raw = x.dtype == torch.float32 and config.compute_dtype() == torch.bfloat16 and x.is_contiguous() and not x.requires_grad
        cdt = torch.bfloat16 if raw else x.dtype
        if not raw:
            x = _nhwc(x)
        N, Cc, H, W = x.shape
        K, Cw, R, S = weight.shape
        if Cw != Cc:
            raise RuntimeError(f"conv2d: input has {Cc} channels, weight expects {Cw}")
        P = (H + 2 * pad - R) // stride + 1
        Q = (W + 2 * pad - S) // stride + 1
        dt = _dv_dtype(cdt)
        y = torch.empty((N, K, P, Q), dtype=cdt, device=x.device, memory_format=_CL)
        # C=3 stems (resnet50.py:20-28, mobilenet_v1.py:30).  Direct path: padded NHWC4 image + sliding-window tensor map
        # (dv_stem_*); the 100 MB image is also what the backward pass keeps.
        ctx.patches = None
        ctx.stem = None
        if (cdt == torch.bfloat16 and algo != DV_ALGO_SIMT and Cc <= 4 and config.stem_direct() and not ctx.needs_input_grad[0]):
            sp = StemParams()
            sp.N, sp.H, sp.W, sp.C, sp.K, sp.R, sp.S, sp.stride, sp.pad = N, H, W, Cc, K, R, S, stride, pad
            hp, wp = C.c_int(0), C.c_int(0)
            if lib.dv_stem_supported(C.byref(sp), C.byref(hp), C.byref(wp)) == 1:
                xp = torch.empty((N, hp.value, wp.value, 4), dtype=cdt, device=x.device)
                sp.x = xp.data_ptr()
                if raw:
                    check(lib.dv_stem_pad_input_nchw(C.byref(sp), x.data_ptr(), _stream()), "dv_stem_pad_input_nchw")
                else:
                    check(lib.dv_stem_pad_input(C.byref(sp), x.data_ptr(), _stream()), "dv_stem_pad_input")
                w_stem = cache.get_stem(weight, sp)
                sp.w, sp.y = w_stem.data_ptr(), y.data_ptr()
                if stats is not None:
                    if stats.pending:
                        stats.t.zero_()
                    sp.stat_sums = stats.t.data_ptr()
                    stats.pending = True
                check(lib.dv_stem_fprop(C.byref(sp), _stream()), "dv_stem_fprop")
                ctx.save_for_backward(xp, weight)
                ctx.stem = (N, H, W, Cc, stride, pad)
                ctx.geom = (stride, pad, algo)
                return y
        if raw:   # no direct path for this geometry: the layout pass after all
            x = _ToNHWC.forward(_NoCtx(), x, cdt)
        if (x.dtype == torch.bfloat16 and algo != DV_ALGO_SIMT and Cc < 16 and R * S * Cc <= 256 and K % 8 == 0
                and not ctx.needs_input_grad[0]):
            kp = -(-(R * S * Cc) // 64) * 64
            M = N * P * Q
            patches = torch.empty((M, kp), dtype=x.dtype, device=x.device)
            check(lib.dv_im2col(x.data_ptr(), patches.data_ptr(), N, H, W, Cc, R, S, stride, pad, kp, dt, _stream()), "dv_im2col")
            w_pad = cache.get_padded(weight, x.dtype, kp)
            p = _conv_params(patches.data_ptr(), w_pad.data_ptr(), y.data_ptr(), None, M, 1, 1, kp, K, 1, 1, 1, 0, dt, DV_ALGO_TC)
            _attach_stats(p, stats, y, DV_CONV_FPROP)
            _run_conv(lib.dv_conv_fprop, p, DV_CONV_FPROP, x.device, "dv_conv_fprop(stem)")
            ctx.save_for_backward(patches, weight)
            ctx.patches = (M, kp, N, Cc, H, W)
            ctx.geom = (stride, pad, algo)
            return y
        w_c, w_t = cache.get(weight, x.dtype, False)  # dgrad reads the KRSC copy MN-major: no transposed filters
        p = _conv_params(x.data_ptr(), w_c.data_ptr(), y.data_ptr(), None, N, H, W, Cc, K, R, S, stride, pad, dt, algo)
        _attach_stats(p, stats, y, DV_CONV_FPROP)
        _run_conv(lib.dv_conv_fprop, p, DV_CONV_FPROP, x.device, "dv_conv_fprop")
        if role == "collect" and slot is not None:
            slot.lazy_ok = bool(x.dtype == torch.bfloat16 and stride == 1 and w_t is None and Cc % 32 == 0 and config.fuse_shortcut_grad()
                                and lib.dv_conv_algo(C.byref(p), DV_CONV_DGRAD) == DV_ALGO_TC)
        ctx.save_for_backward(x, weight)
        ctx.w_c, ctx.w_t = w_c, w_t
        ctx.geom = (stride, pad, algo)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, weight = ctx.saved_tensors
        stride, pad, algo = ctx.geom
        dy = _nhwc(dy)
        if dy.dtype != x.dtype:
            dy = dy.to(x.dtype)
        K, _, R, S = weight.shape
        dt = _dv_dtype(x.dtype)
        dx = dw = None
        if ctx.stem is not None:
            if ctx.needs_input_grad[1]:
                N, H, W, Cc, _, _ = ctx.stem
                sink = grad_sink(weight)
                dw = sink[0] if sink else torch.empty((K, Cc, R, S), dtype=torch.float32, device=x.device, memory_format=_CL)
                sp = StemParams()
                sp.N, sp.H, sp.W, sp.C, sp.K, sp.R, sp.S, sp.stride, sp.pad = N, H, W, Cc, K, R, S, stride, pad
                sp.x, sp.y, sp.dw = x.data_ptr(), dy.data_ptr(), dw.data_ptr()
                sp.accumulate = 1 if config.atomic_wgrad() else 0

                def launch():
                    ws = _workspace(lib.dv_stem_workspace_bytes(C.byref(sp)), x.device)
                    sp.workspace, sp.workspace_bytes = ws.data_ptr(), ws.numel()
                    check(lib.dv_stem_wgrad(C.byref(sp), _stream()), "dv_stem_wgrad")

                side = _side_wgrad["obj"] if (sink and sp.accumulate) else None
                if side is not None:
                    with torch.cuda.stream(side.fork(x, dy)):
                        launch()
                else:
                    launch()
                if sink:
                    sink[1](weight)
                    dw = None
            return None, dw, None, None, None, None, None, None, None, None
        if ctx.patches is not None:
            M, kp, N, Cc, H, W = ctx.patches
            if ctx.needs_input_grad[1]:
                dw_pad = torch.empty((K, kp), dtype=torch.float32, device=x.device)
                p = _conv_params(x.data_ptr(), dw_pad.data_ptr(), dy.data_ptr(), None, M, 1, 1, kp, K, 1, 1, 1, 0, dt, DV_ALGO_TC)
                _run_conv(lib.dv_conv_wgrad, p, DV_CONV_WGRAD, x.device, "dv_conv_wgrad(stem)")
                sink = grad_sink(weight)
                dw = sink[0] if sink else torch.empty((K, Cc, R, S), dtype=torch.float32, device=x.device, memory_format=_CL)
                check(lib.dv_unpad_rows(dw_pad.data_ptr(), dw.data_ptr(), K, R * S * Cc, kp, _stream()), "dv_unpad_rows")
                if sink:
                    sink[1](weight)
                    dw = None
            return None, dw, None, None, None, None, None, None, None, None
        N, Cc, H, W = x.shape
        if ctx.needs_input_grad[0]:
            acc, shortcut = 0, None
            if ctx.role == "collect":
                if ctx.slot.t is None:
                    raise RuntimeError("deep_vision_b200: residual gradient was not deposited before the main path's dgrad")
                dep, ctx.slot.t = ctx.slot.t, None
                if isinstance(dep, tuple):    # (dout, relu mask) of the block's merge: added by this dgrad's epilogue
                    shortcut, dx = dep, torch.empty_like(x, memory_format=_CL)
                else:
                    dx, acc = dep, 1
            else:
                dx = torch.empty_like(x, memory_format=_CL)
            p = _conv_params(dx.data_ptr(), ctx.w_c.data_ptr(), dy.data_ptr(), ctx.w_t.data_ptr() if ctx.w_t is not None else None,
                             N, H, W, Cc, K, R, S, stride, pad, dt, algo, accumulate=acc)
            if shortcut is not None:
                p.ep_residual, p.ep_residual_mask = shortcut[0].data_ptr(), shortcut[1].data_ptr()
            h = ctx.handoff
            fused = None
            if (h is not None and h.x is not None and h.claims == 1 and acc == 0 and stride == 1 and ctx.w_t is None
                    and config.fuse_bn_bwd() and lib.dv_conv_algo(C.byref(p), DV_CONV_DGRAD) == DV_ALGO_TC):
                # this dx is dy of the BatchNorm that produced x: let the epilogue add that layer's dgamma / dbeta
                sg, sb = grad_sink(h.gamma), grad_sink(h.beta)
                dgamma = sg[0] if sg else torch.zeros(Cc, dtype=torch.float32, device=x.device)
                dbeta = sb[0] if sb else torch.zeros(Cc, dtype=torch.float32, device=x.device)
                p.bnb_x, p.bnb_gamma, p.bnb_beta = h.x.data_ptr(), h.gamma.data_ptr(), h.beta.data_ptr()
                p.bnb_mean, p.bnb_invstd = h.mean.data_ptr(), h.invstd.data_ptr()
                p.bnb_dgamma, p.bnb_dbeta = dgamma.data_ptr(), dbeta.data_ptr()
                fused = (dgamma, dbeta)
            _run_conv(lib.dv_conv_dgrad, p, DV_CONV_DGRAD, x.device, "dv_conv_dgrad")
            if fused is not None:
                h.result = (dx, fused[0], fused[1])
            if ctx.role == "deposit":
                ctx.slot.t, dx = dx, None
        if ctx.needs_input_grad[1]:
            sink = grad_sink(weight)
            atomic = 1 if config.atomic_wgrad() else 0
            if sink:
                dw = sink[0]          # slice of the flat gradient buffer, zeroed at the start of the step
            elif atomic:
                dw = torch.empty((K, Cc, R, S), dtype=torch.float32, device=x.device, memory_format=_CL).zero_()
            else:
                dw = torch.empty((K, Cc, R, S), dtype=torch.float32, device=x.device, memory_format=_CL)
            p = _conv_params(x.data_ptr(), dw.data_ptr(), dy.data_ptr(), None, N, H, W, Cc, K, R, S, stride, pad, dt, algo,
                             accumulate=atomic)
            side = _side_wgrad["obj"] if (sink and atomic) else None   # direct write into the flat buffer, no workspace
            if side is not None:
                with torch.cuda.stream(side.fork(x, dy)):
                    _run_conv(lib.dv_conv_wgrad, p, DV_CONV_WGRAD, x.device, "dv_conv_wgrad")
            else:
                _run_conv(lib.dv_conv_wgrad, p, DV_CONV_WGRAD, x.device, "dv_conv_wgrad")
            if sink:
                sink[1](weight)
                dw = None
        return dx, dw, None, None, None, None, None, None, None, None


def _attach_stats(p, stats, y, pass_id):
    """Let the tensor-core epilogue accumulate BatchNorm's sums; tag the output so the BatchNorm finds them."""
    if stats is None or lib.dv_conv_algo(C.byref(p), pass_id) != DV_ALGO_TC:
        return
    if stats.pending:  # produced earlier and never consumed by a BatchNorm
        stats.t.zero_()
    p.stat_sums = stats.t.data_ptr()
    stats.pending = True


def conv2d(x, weight, stride=1, padding=0, cache=None, algo=DV_ALGO_AUTO, slot=None, role=None, stats=None):
    y = _Conv2d.apply(x, weight, int(stride), int(padding), cache or FilterCache(), algo, slot, role, stats,
                      getattr(x, "_dv_bn_handoff", None))
    if stats is not None and stats.pending:
        y._dv_stat_sums = stats  # picked up by the BatchNorm that consumes y
    return y


# ----------------------------------------------------------------------------------------------------------------
# depth-wise 3x3 convolution
# ----------------------------------------------------------------------------------------------------------------
def _dw_params(x_ptr, w_ptr, y_ptr, N, H, W, Cc, stride, dtype):
    p = DwConvParams()
    p.x, p.w, p.y = x_ptr, w_ptr, y_ptr
    p.N, p.H, p.W, p.C, p.stride, p.dtype = N, H, W, Cc, stride, dtype
    return p


# BatchNorm-backward sums from the stride-1 depth-wise dgrad (dv_dwconv_params.bnb_*): correct and tested, but that kernel is bound
# by its fp32 arithmetic, and the extra x0 vector + mask + two sums per output cost what the saved reduce pass gains
# (MobileNetV1 step 7.51 vs 7.48 ms) — opt-in.
_DW_BNB = _environ.get("DV_DW_FUSE_BN_BWD", "0") == "1"
_DW_SIDE = _environ.get("DV_DW_SIDE", "1") == "1"   # experiment knob: depth-wise weight gradients on the side stream


class _DwConv3x3(torch.autograd.Function):
    """nn.Conv2d(C, C, 3, padding=1, stride, groups=C, bias=False) — mobilenet_v1.py:112-122."""

    @staticmethod
    def forward(ctx, x, weight, stride, stats=None, handoff=None):
        ctx.handoff = handoff
        if handoff is not None:
            handoff.claims += 1
        _require_cuda(x, "dwconv3x3")
        x = _nhwc(x)
        N, Cc, H, W = x.shape
        P = (H - 1) // stride + 1
        Q = (W - 1) // stride + 1
        w = weight.detach().contiguous()  # [C,1,3,3] fp32 == [C][3][3]
        y = torch.empty((N, Cc, P, Q), dtype=x.dtype, device=x.device, memory_format=_CL)
        p = _dw_params(x.data_ptr(), w.data_ptr(), y.data_ptr(), N, H, W, Cc, stride, _dv_dtype(x.dtype))
        if stats is not None and lib.dv_dwconv_fuses_stats(C.byref(p)) == 1:
            if stats.pending:
                stats.t.zero_()
            p.stat_sums = stats.t.data_ptr()
            stats.pending = True
        check(lib.dv_dwconv_fprop(C.byref(p), _stream()), "dv_dwconv_fprop")
        ctx.save_for_backward(x, weight)
        ctx.stride = stride
        return y

    @staticmethod
    def backward(ctx, dy):
        x, weight = ctx.saved_tensors
        stride = ctx.stride
        dy = _nhwc(dy)
        if dy.dtype != x.dtype:
            dy = dy.to(x.dtype)
        N, Cc, H, W = x.shape
        dt = _dv_dtype(x.dtype)
        dx = dw = None
        w = weight.detach().contiguous()
        if ctx.needs_input_grad[0]:
            dx = torch.empty_like(x, memory_format=_CL)
            p = _dw_params(dx.data_ptr(), w.data_ptr(), dy.data_ptr(), N, H, W, Cc, stride, dt)
            h, fused = ctx.handoff, None
            if (h is not None and h.x is not None and h.claims == 1 and config.fuse_bn_bwd() and _DW_BNB
                    and lib.dv_dwconv_fuses_bn_bwd(C.byref(p)) == 1):
                # this dx is dy of the BatchNorm that produced x (pw.bn -> relu -> dw): the kernel adds that layer's dgamma / dbeta
                sg, sb = grad_sink(h.gamma), grad_sink(h.beta)
                dgamma = sg[0] if sg else torch.zeros(Cc, dtype=torch.float32, device=x.device)
                dbeta = sb[0] if sb else torch.zeros(Cc, dtype=torch.float32, device=x.device)
                p.bnb_x, p.bnb_gamma, p.bnb_beta = h.x.data_ptr(), h.gamma.data_ptr(), h.beta.data_ptr()
                p.bnb_mean, p.bnb_invstd = h.mean.data_ptr(), h.invstd.data_ptr()
                p.bnb_dgamma, p.bnb_dbeta = dgamma.data_ptr(), dbeta.data_ptr()
                fused = (dgamma, dbeta)
            check(lib.dv_dwconv_dgrad(C.byref(p), _stream()), "dv_dwconv_dgrad")
            if fused is not None:
                h.result = (dx, fused[0], fused[1])
        if ctx.needs_input_grad[1]:
            sink = grad_sink(weight)
            dw = sink[0] if sink else torch.empty_like(w)   # the flat buffer's slice for [C,1,3,3] is [C][3][3] fp32 as well
            p = _dw_params(x.data_ptr(), dw.data_ptr(), dy.data_ptr(), N, H, W, Cc, stride, dt)

            def launch():
                need = lib.dv_dwconv_workspace_bytes(C.byref(p))
                ws = _workspace(need, x.device)     # per stream: the side stream has its own scratch
                p.workspace, p.workspace_bytes = ws.data_ptr(), ws.numel()
                check(lib.dv_dwconv_wgrad(C.byref(p), _stream()), "dv_dwconv_wgrad")

            side = _side_wgrad["obj"] if (sink and _DW_SIDE) else None   # nothing in the backward chain waits for a weight gradient
            if side is not None:
                with torch.cuda.stream(side.fork(x, dy)):
                    launch()
            else:
                launch()
            if sink:
                sink[1](weight)
                dw = None
        return dx, dw, None, None, None


def dwconv3x3(x, weight, stride=1, stats=None):
    y = _DwConv3x3.apply(x, weight, int(stride), stats, getattr(x, "_dv_bn_handoff", None))
    if stats is not None and stats.pending:
        y._dv_stat_sums = stats   # picked up by the BatchNorm that consumes y
    return y


# ----------------------------------------------------------------------------------------------------------------
# batch norm (+ residual) (+ ReLU)
# ----------------------------------------------------------------------------------------------------------------
# Traversal direction of the streaming BatchNorm passes (dv_bn_params.reverse): a pass that starts at the END of a tensor meets
# the bytes its producer (a convolution walking its tiles upwards) wrote last, while they are still in the L2.
import os as _os
_ZIGZAG = {"fwd": _os.environ.get("DV_ZZ_FWD", "1") == "1", "bwd": _os.environ.get("DV_ZZ_BWD", "1") == "1",
           "bwd_fused": _os.environ.get("DV_ZZ_BWD_FUSED", "1") == "1"}


class _BatchNormAct(torch.autograd.Function):
    """y = relu(bn(x) [+ residual]) — nn.BatchNorm2d + nn.ReLU(inplace=True) (+ `x + identity`) at
    resnet50.py:152-164, resnet34.py:135-142, mobilenet_v1.py:129-133,152-156."""

    @staticmethod
    def forward(ctx, x, residual, gamma, beta, running_mean, running_var, training, momentum, eps, relu, res_slot=None,
                sums=None, handoff=None):
        ctx.res_slot = res_slot
        ctx.handoff = None
        _require_cuda(x, "batch_norm")
        x = _nhwc(x)
        N, Cc, H, W = x.shape
        M = N * H * W
        if residual is not None:
            residual = _nhwc(residual)
            if residual.shape != x.shape or residual.dtype != x.dtype:
                raise RuntimeError("batch_norm: residual must match the input's shape and dtype")
        y = torch.empty_like(x, memory_format=_CL)
        p = BnParams()
        p.x, p.y = x.data_ptr(), y.data_ptr()
        p.residual = residual.data_ptr() if residual is not None else None
        p.gamma, p.beta = gamma.data_ptr(), beta.data_ptr()
        p.running_mean = running_mean.data_ptr() if running_mean is not None else None
        p.running_var = running_var.data_ptr() if running_var is not None else None
        p.M, p.C, p.dtype, p.relu = M, Cc, _dv_dtype(x.dtype), int(relu)
        p.eps, p.momentum = eps, momentum
        use_batch_stats = training or running_mean is None
        if use_batch_stats:
            save_mean = torch.empty(Cc, dtype=torch.float32, device=x.device)
            save_invstd = torch.empty(Cc, dtype=torch.float32, device=x.device)
            p.save_mean, p.save_invstd = save_mean.data_ptr(), save_invstd.data_ptr()
            if not training:  # eval without running stats: normalise with batch stats, update nothing
                p.running_mean = p.running_var = None
            if sums is not None and sums.pending and sums.t.numel() == 2 * Cc:
                check(lib.dv_bn_stats_from_sums(C.byref(p), sums.t.data_ptr(), _stream()), "dv_bn_stats_from_sums")
                sums.pending = False
            else:
                ws = _workspace(lib.dv_bn_workspace_bytes(M, Cc), x.device)
                p.workspace, p.workspace_bytes = ws.data_ptr(), ws.numel()
                check(lib.dv_bn_stats(C.byref(p), _stream()), "dv_bn_stats")
            p.eval_mode = 0
        else:
            save_mean = save_invstd = None
            p.eval_mode = 1
        # ReLU after a residual add: the sign cannot be recomputed from x alone, so the forward leaves one bit per element
        # for the backward (bf16) instead of making both backward passes re-read the 16-bit output
        mask = None
        if relu and residual is not None and use_batch_stats and x.dtype == torch.bfloat16 and any(ctx.needs_input_grad[:4]):
            mask = torch.empty((M, Cc // 8), dtype=torch.uint8, device=x.device)
            p.relu_mask = mask.data_ptr()
        p.reverse = int(_ZIGZAG["fwd"] and use_batch_stats)
        check(lib.dv_bn_apply(C.byref(p), _stream()), "dv_bn_apply")
        ctx.batch_stats = use_batch_stats
        ctx.relu = bool(relu)
        ctx.has_res = residual is not None
        # the ReLU mask: without a residual it is recomputed bit-exactly from x in backward (y need not be kept or re-read)
        keep_y = (mask if mask is not None else y) if (relu and residual is not None) else None
        ctx.bit_mask = mask is not None
        if use_batch_stats:
            ctx.save_for_backward(x, keep_y, gamma, save_mean, save_invstd, beta)
            if (handoff is not None and relu and residual is None and x.dtype == torch.bfloat16 and any(ctx.needs_input_grad[:4])):
                handoff.x, handoff.gamma, handoff.beta, handoff.mean, handoff.invstd = x, gamma, beta, save_mean, save_invstd
                ctx.handoff = handoff
        else:
            ctx.save_for_backward(x, keep_y, gamma, running_mean, running_var, beta)
            ctx.eps = eps
        return y

    @staticmethod
    def backward(ctx, dy):
        x, y, gamma, mean, second, beta = ctx.saved_tensors
        if not ctx.batch_stats:
            raise NotImplementedError("deep_vision_b200: backward through eval-mode BatchNorm2d is not on the hot path")
        dy = _nhwc(dy)
        if dy.dtype != x.dtype:
            dy = dy.to(x.dtype)
        N, Cc, H, W = x.shape
        M = N * H * W
        dx = torch.empty_like(x, memory_format=_CL)
        # identity shortcut whose collector adds `dy * mask` in its dgrad epilogue: no masked copy of dy is written here
        lazy = bool(ctx.has_res and ctx.needs_input_grad[1] and ctx.bit_mask and ctx.res_slot is not None and ctx.res_slot.lazy_ok)
        dres = torch.empty_like(x, memory_format=_CL) if (ctx.has_res and ctx.needs_input_grad[1] and not lazy) else None
        sg, sb = grad_sink(gamma), grad_sink(beta)
        atomic = 1 if config.atomic_bn_bwd() else 0
        mk = torch.zeros if atomic else torch.empty
        dgamma = sg[0] if sg else mk(Cc, dtype=torch.float32, device=x.device)
        dbeta = sb[0] if sb else mk(Cc, dtype=torch.float32, device=x.device)
        p = BnBwdParams()
        p.atomic_sums = atomic
        p.x, p.dy, p.dx = x.data_ptr(), dy.data_ptr(), dx.data_ptr()
        if ctx.bit_mask:
            p.relu_mask = y.data_ptr()
        else:
            p.y = y.data_ptr() if y is not None else None
        p.dres = dres.data_ptr() if dres is not None else None
        p.gamma, p.beta, p.save_mean, p.save_invstd = gamma.data_ptr(), beta.data_ptr(), mean.data_ptr(), second.data_ptr()
        p.dgamma, p.dbeta = dgamma.data_ptr(), dbeta.data_ptr()
        p.M, p.C, p.dtype, p.relu = M, Cc, _dv_dtype(x.dtype), int(ctx.relu)
        # sums that the consumer's dgrad epilogue already produced (BnBwdHandoff) — valid only if this dy IS that dgrad's output
        h, pre = ctx.handoff, None
        if h is not None and h.result is not None:
            pre, h.result = h.result, None
            if pre[0].data_ptr() == dy.data_ptr() and pre[0].shape == dy.shape:
                if sg is None:
                    dgamma = pre[1]
                if sb is None:
                    dbeta = pre[2]
                p.dgamma, p.dbeta = pre[1].data_ptr(), pre[2].data_ptr()
            else:                       # autograd added other gradients onto it: the partial sums are void
                if sg:
                    sg[0].zero_()
                if sb:
                    sb[0].zero_()
                pre = None
        if h is not None and h.owner is not None:
            h.owner._dv_bwd_sums_fused = pre is not None
        p.reverse = int(_ZIGZAG["bwd_fused"] if pre is not None else _ZIGZAG["bwd"])
        if pre is None:
            ws = _workspace(lib.dv_bn_workspace_bytes(M, Cc), x.device)
            p.workspace, p.workspace_bytes = ws.data_ptr(), ws.numel()
            check(lib.dv_bn_bwd_reduce(C.byref(p), _stream()), "dv_bn_bwd_reduce")
        check(lib.dv_bn_bwd_apply(C.byref(p), _stream()), "dv_bn_bwd_apply")
        if sg:
            sg[1](gamma)
            dgamma = None
        if sb:
            sb[1](beta)
            dbeta = None
        if lazy:
            ctx.res_slot.t = (dy, y)
        elif ctx.res_slot is not None and dres is not None:
            ctx.res_slot.t, dres = dres, None   # handed to the main path's dgrad, not to autograd
        return dx, dres, dgamma, dbeta, None, None, None, None, None, None, None, None, None


def batch_norm_act(x, residual, gamma, beta, running_mean, running_var, training, momentum, eps, relu, res_slot=None, owner=None):
    sums = getattr(x, "_dv_stat_sums", None)
    handoff = BnBwdHandoff(owner) if (relu and residual is None and torch.is_grad_enabled() and config.fuse_bn_bwd()) else None
    y = _BatchNormAct.apply(x, residual, gamma, beta, running_mean, running_var, bool(training), float(momentum), float(eps),
                            bool(relu), res_slot, sums, handoff)
    if handoff is not None and handoff.x is not None:
        y._dv_bn_handoff = handoff      # claimed by the convolution that consumes y (functional.conv2d / dwconv3x3)
    return y


class _BnReluMaxPool(torch.autograd.Function):
    """maxpool3x3s2(relu(bn(x))) in train mode — bn1 / relu / maxpool of the ResNet stem (resnet50.py:49-52) — without ever
    writing the normalised tensor or the pool's input gradient (dv_bn_relu_maxpool_*)."""

    @staticmethod
    def forward(ctx, x, gamma, beta, running_mean, running_var, momentum, eps, sums):
        _require_cuda(x, "bn_relu_maxpool")
        x = _nhwc(x)
        N, Cc, H, W = x.shape
        M = N * H * W
        save_mean = torch.empty(Cc, dtype=torch.float32, device=x.device)
        save_invstd = torch.empty(Cc, dtype=torch.float32, device=x.device)
        p = BnParams()
        p.x, p.gamma, p.beta = x.data_ptr(), gamma.data_ptr(), beta.data_ptr()
        p.running_mean = running_mean.data_ptr() if running_mean is not None else None
        p.running_var = running_var.data_ptr() if running_var is not None else None
        p.save_mean, p.save_invstd = save_mean.data_ptr(), save_invstd.data_ptr()
        p.M, p.C, p.dtype, p.relu, p.eps, p.momentum = M, Cc, _dv_dtype(x.dtype), 1, eps, momentum
        if sums is not None and sums.pending and sums.t.numel() == 2 * Cc:
            check(lib.dv_bn_stats_from_sums(C.byref(p), sums.t.data_ptr(), _stream()), "dv_bn_stats_from_sums")
            sums.pending = False
        else:
            ws = _workspace(lib.dv_bn_workspace_bytes(M, Cc), x.device)
            p.workspace, p.workspace_bytes = ws.data_ptr(), ws.numel()
            check(lib.dv_bn_stats(C.byref(p), _stream()), "dv_bn_stats")
        P, Q = (H - 1) // 2 + 1, (W - 1) // 2 + 1
        y = torch.empty((N, Cc, P, Q), dtype=x.dtype, device=x.device, memory_format=_CL)
        idx = torch.empty((N, Cc, P, Q), dtype=torch.uint8, device=x.device, memory_format=_CL)
        q = BnPoolParams()
        q.x, q.y, q.idx = x.data_ptr(), y.data_ptr(), idx.data_ptr()
        q.gamma, q.beta, q.save_mean, q.save_invstd = gamma.data_ptr(), beta.data_ptr(), save_mean.data_ptr(), save_invstd.data_ptr()
        q.N, q.H, q.W, q.C, q.dtype = N, H, W, Cc, _dv_dtype(x.dtype)
        check(lib.dv_bn_relu_maxpool_fwd(C.byref(q), _stream()), "dv_bn_relu_maxpool_fwd")
        ctx.save_for_backward(x, idx, gamma, beta, save_mean, save_invstd)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, idx, gamma, beta, mean, invstd = ctx.saved_tensors
        dy = _nhwc(dy)
        if dy.dtype != x.dtype:
            dy = dy.to(x.dtype)
        N, Cc, H, W = x.shape
        dx = torch.empty_like(x, memory_format=_CL)
        sg, sb = grad_sink(gamma), grad_sink(beta)
        dgamma = sg[0] if sg else torch.empty(Cc, dtype=torch.float32, device=x.device)
        dbeta = sb[0] if sb else torch.empty(Cc, dtype=torch.float32, device=x.device)
        q = BnPoolParams()
        q.x, q.idx, q.dy, q.dx = x.data_ptr(), idx.data_ptr(), dy.data_ptr(), dx.data_ptr()
        q.gamma, q.beta, q.save_mean, q.save_invstd = gamma.data_ptr(), beta.data_ptr(), mean.data_ptr(), invstd.data_ptr()
        q.dgamma, q.dbeta = dgamma.data_ptr(), dbeta.data_ptr()
        q.N, q.H, q.W, q.C, q.dtype = N, H, W, Cc, _dv_dtype(x.dtype)
        ws = _workspace(lib.dv_bn_pool_workspace_bytes(C.byref(q)), x.device)
        q.workspace, q.workspace_bytes = ws.data_ptr(), ws.numel()
        check(lib.dv_bn_relu_maxpool_bwd(C.byref(q), _stream()), "dv_bn_relu_maxpool_bwd")
        if sg:
            sg[1](gamma)
            dgamma = None
        if sb:
            sb[1](beta)
            dbeta = None
        return (dx if ctx.needs_input_grad[0] else None), dgamma, dbeta, None, None, None, None, None


def bn_relu_maxpool(x, gamma, beta, running_mean, running_var, momentum, eps):
    sums = getattr(x, "_dv_stat_sums", None)
    return _BnReluMaxPool.apply(x, gamma, beta, running_mean, running_var, float(momentum), float(eps), sums)


class _ReLU(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        _require_cuda(x, "relu")
        xc = x if (x.is_contiguous() or x.is_contiguous(memory_format=_CL)) else x.contiguous()
        y = torch.empty_like(xc)
        check(lib.dv_relu_fwd(xc.data_ptr(), y.data_ptr(), xc.numel(), _dv_dtype(xc.dtype), _stream()), "dv_relu_fwd")
        ctx.save_for_backward(y)
        return y

    @staticmethod
    def backward(ctx, dy):
        (y,) = ctx.saved_tensors
        if dy.stride() != y.stride():
            dy = dy.contiguous(memory_format=_CL) if y.is_contiguous(memory_format=_CL) and y.dim() == 4 else dy.contiguous()
        dx = torch.empty_like(y)
        check(lib.dv_relu_bwd(y.data_ptr(), dy.data_ptr(), dx.data_ptr(), y.numel(), _dv_dtype(y.dtype), _stream()), "dv_relu_bwd")
        return dx


def relu(x):
    return _ReLU.apply(x)


# ----------------------------------------------------------------------------------------------------------------
# stem / tail
# ----------------------------------------------------------------------------------------------------------------
class _MaxPool3x3s2(torch.autograd.Function):
    """nn.MaxPool2d(kernel_size=3, stride=2, padding=1) — resnet50.py:34."""

    @staticmethod
    def forward(ctx, x):
        _require_cuda(x, "max_pool")
        x = _nhwc(x)
        N, Cc, H, W = x.shape
        P, Q = (H - 1) // 2 + 1, (W - 1) // 2 + 1
        y = torch.empty((N, Cc, P, Q), dtype=x.dtype, device=x.device, memory_format=_CL)
        idx = torch.empty((N, Cc, P, Q), dtype=torch.uint8, device=x.device, memory_format=_CL)
        check(lib.dv_maxpool_fwd(x.data_ptr(), y.data_ptr(), idx.data_ptr(), N, H, W, Cc, _dv_dtype(x.dtype), _stream()), "dv_maxpool_fwd")
        ctx.save_for_backward(idx)
        ctx.in_shape = (N, Cc, H, W)
        return y

    @staticmethod
    def backward(ctx, dy):
        (idx,) = ctx.saved_tensors
        N, Cc, H, W = ctx.in_shape
        dy = _nhwc(dy)
        dx = torch.empty((N, Cc, H, W), dtype=dy.dtype, device=dy.device, memory_format=_CL)
        check(lib.dv_maxpool_bwd(dy.data_ptr(), idx.data_ptr(), dx.data_ptr(), N, H, W, Cc, _dv_dtype(dy.dtype), _stream()), "dv_maxpool_bwd")
        return dx


def max_pool3x3s2(x):
    return _MaxPool3x3s2.apply(x)


class _GlobalAvgPool(torch.autograd.Function):
    """nn.AdaptiveAvgPool2d((1, 1)) — resnet50.py:43, mobilenet_v1.py:84."""

    @staticmethod
    def forward(ctx, x):
        _require_cuda(x, "avg_pool")
        x = _nhwc(x)
        N, Cc, H, W = x.shape
        y = torch.empty((N, Cc, 1, 1), dtype=x.dtype, device=x.device, memory_format=_CL)
        check(lib.dv_avgpool_fwd(x.data_ptr(), y.data_ptr(), N, H * W, Cc, _dv_dtype(x.dtype), _stream()), "dv_avgpool_fwd")
        ctx.in_shape = (N, Cc, H, W)
        return y

    @staticmethod
    def backward(ctx, dy):
        N, Cc, H, W = ctx.in_shape
        dy = dy.reshape(N, Cc).contiguous()
        dx = torch.empty((N, Cc, H, W), dtype=dy.dtype, device=dy.device, memory_format=_CL)
        check(lib.dv_avgpool_bwd(dy.data_ptr(), dx.data_ptr(), N, H * W, Cc, _dv_dtype(dy.dtype), _stream()), "dv_avgpool_bwd")
        return dx


def global_avg_pool(x):
    return _GlobalAvgPool.apply(x)


class _Linear(torch.autograd.Function):
    """y = x W^T + b with fp32 logits — nn.Linear at resnet50.py:44,61; mobilenet_v1.py:87,92.
    Runs as a 1x1 convolution on [B,1,1,F] (dv_conv_*), bias through dv_bias_add / dv_bias_grad."""

    @staticmethod
    def forward(ctx, x, weight, bias, cache):
        _require_cuda(x, "linear")
        x = x.contiguous()
        B, F = x.shape
        K = weight.shape[0]
        w_c, w_t = cache.get(weight.view(K, F, 1, 1), x.dtype, False)
        y = torch.empty((B, K), dtype=torch.float32, device=x.device)
        algo = config.conv_algo()
        p = _conv_params(x.data_ptr(), w_c.data_ptr(), y.data_ptr(), None, B, 1, 1, F, K, 1, 1, 1, 0, _dv_dtype(x.dtype), algo, y_f32=1)
        _run_conv(lib.dv_conv_fprop, p, DV_CONV_FPROP, x.device, "dv_conv_fprop(linear)")
        ctx.w_t, ctx.algo = w_t, algo
        if bias is not None:
            check(lib.dv_bias_add(y.data_ptr(), bias.data_ptr(), B, K, _stream()), "dv_bias_add")
        ctx.save_for_backward(x, weight)
        ctx.w_c = w_c
        ctx.has_bias = bias is not None
        return y

    @staticmethod
    def backward(ctx, dy):
        x, weight = ctx.saved_tensors
        B, F = x.shape
        K = weight.shape[0]
        dy = dy.contiguous()
        dt = _dv_dtype(x.dtype)
        if dy.dtype != x.dtype:
            dyc = torch.empty_like(dy, dtype=x.dtype)
            check(lib.dv_cast(dy.data_ptr(), _dv_dtype(dy.dtype), dyc.data_ptr(), dt, dy.numel(), _stream()), "dv_cast")
        else:
            dyc = dy
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty_like(x)
            p = _conv_params(dx.data_ptr(), ctx.w_c.data_ptr(), dyc.data_ptr(), ctx.w_t.data_ptr() if ctx.w_t is not None else None,
                             B, 1, 1, F, K, 1, 1, 1, 0, dt, ctx.algo)
            _run_conv(lib.dv_conv_dgrad, p, DV_CONV_DGRAD, x.device, "dv_conv_dgrad(linear)")
        if ctx.needs_input_grad[1]:
            dw = torch.empty((K, F), dtype=torch.float32, device=x.device)
            p = _conv_params(x.data_ptr(), dw.data_ptr(), dyc.data_ptr(), None, B, 1, 1, F, K, 1, 1, 1, 0, dt, ctx.algo)
            _run_conv(lib.dv_conv_wgrad, p, DV_CONV_WGRAD, x.device, "dv_conv_wgrad(linear)")
        if ctx.has_bias and ctx.needs_input_grad[2]:
            dyf = dy if dy.dtype == torch.float32 else dy.float()
            db = torch.empty(K, dtype=torch.float32, device=x.device)
            check(lib.dv_bias_grad(dyf.data_ptr(), db.data_ptr(), B, K, _stream()), "dv_bias_grad")
        return dx, dw, db, None


def linear(x, weight, bias=None, cache=None):
    return _Linear.apply(x, weight, bias, cache or FilterCache())


class _CrossEntropy(torch.autograd.Function):
    """nn.CrossEntropyLoss() with mean reduction — ResNet/pytorch/train.py:358,452.  Loss and d(loss)/d(logits) come
    out of one launch; backward only applies the upstream scalar."""

    @staticmethod
    def forward(ctx, logits, labels):
        _require_cuda(logits, "cross_entropy")
        if logits.dtype != torch.float32 or logits.dim() != 2 or labels.dtype != torch.int64:
            raise TypeError("cross_entropy expects fp32 [B,C] logits and int64 [B] labels")
        logits = logits.contiguous()
        labels = labels.contiguous()
        B, Cn = logits.shape
        loss = torch.empty((), dtype=torch.float32, device=logits.device)
        dlogits = torch.empty_like(logits)
        check(lib.dv_cross_entropy(logits.data_ptr(), labels.data_ptr(), loss.data_ptr(), dlogits.data_ptr(), B, Cn, _stream()),
              "dv_cross_entropy")
        ctx.save_for_backward(dlogits)
        return loss

    @staticmethod
    def backward(ctx, g):
        (dlogits,) = ctx.saved_tensors
        g = g.contiguous().float()
        check(lib.dv_scale_by(dlogits.data_ptr(), g.data_ptr(), dlogits.numel(), _stream()), "dv_scale_by")
        return dlogits, None


def cross_entropy(logits, labels):
    return _CrossEntropy.apply(logits, labels)
