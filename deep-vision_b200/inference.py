"""Inference path with BatchNorm folded into the convolution's epilogue.

Under ``net.eval()`` + ``torch.no_grad()`` (the reference's ``validate``, ResNet/pytorch/train.py:488-538, and the
notebooks' cell 3) a BatchNorm2d is a per-channel affine map of running statistics, so every
``conv -> bn [-> + identity] [-> relu]`` group of the four models (resnet50.py:152-164, resnet34.py:135-142,
mobilenet_v1.py:129-133,152-156) becomes ONE launch:

    y = relu( conv(x, w) * scale[k] + shift[k] + identity )        scale = gamma / sqrt(running_var + eps)
                                                                    shift = beta - running_mean * scale

* dense convolutions: tcgen05 implicit GEMM, scale/shift/identity/ReLU applied to the fp32 accumulator before the one
  bf16 rounding (`dv_conv_params.ep_*`) — one rounding fewer than conv followed by a separate BN pass;
* depth-wise 3x3: the scale goes into the fp32 taps exactly, the shift initialises the accumulator
  (`dv_dwconv_params.ep_shift`).

The blocks ask for this with ``conv(x, defer=True)``: in folded mode the convolution module returns a `PendingConv`
(input + module, nothing launched) and the BatchNorm2d that receives it launches the fused kernel.  In every other
situation (training, autograd on, fp32 compute, folding switched off, a shape the tensor-core path does not take) the
convolution runs at once and nothing changes.
"""
import ctypes as C

import torch

from . import config
from . import functional as F
from .lib import DV_ALGO_TC, DV_CONV_FPROP, check, lib


class PendingConv:
    """A convolution whose launch is deferred to the BatchNorm2d that consumes it."""
    __slots__ = ("x", "conv")

    def __init__(self, x, conv):
        self.x, self.conv = x, conv

    def materialize(self):
        return self.conv(self.x)


def wanted(conv, x):
    """Whether `conv(x, defer=True)` should defer: inference mode only."""
    return (config.fold_bn() and not conv.training and not torch.is_grad_enabled() and x.is_cuda
            and x.dtype == torch.bfloat16 and config.conv_algo() != 1)


class _Folded:
    """scale / shift (and scaled depth-wise taps) of one BatchNorm2d, rebuilt when any of its tensors changes."""

    def __init__(self):
        self.key = None
        self.scale = self.shift = self.dw = None

    def get(self, bn, dw_weight=None):
        ts = (bn.weight, bn.bias, bn.running_mean, bn.running_var) + ((dw_weight,) if dw_weight is not None else ())
        key = tuple((t.data_ptr(), t._version) for t in ts) + (config.weight_epoch(), bn.eps)
        if key != self.key:
            Cc = bn.num_features
            dev = bn.weight.device
            self.scale = torch.empty(Cc, dtype=torch.float32, device=dev)
            self.shift = torch.empty(Cc, dtype=torch.float32, device=dev)
            self.dw = None
            w_in = None
            if dw_weight is not None:
                w_in = dw_weight.detach().contiguous()
                self.dw = torch.empty_like(w_in)
            check(lib.dv_bn_fold(bn.weight.data_ptr(), bn.bias.data_ptr(), bn.running_mean.data_ptr(), bn.running_var.data_ptr(),
                                 bn.eps, self.scale.data_ptr(), self.shift.data_ptr(), Cc,
                                 w_in.data_ptr() if w_in is not None else None, self.dw.data_ptr() if self.dw is not None else None,
                                 F._stream()), "dv_bn_fold")
            self.key = key
        return self


def conv_bn_act(pending, bn, residual, relu):
    """Launch the deferred convolution with `bn` (eval mode), the residual and the ReLU in its epilogue.
    Returns None when this group cannot be fused (the caller then runs the two passes)."""
    conv, x = pending.conv, F._nhwc(pending.x)
    if bn.training or bn.running_mean is None or not bn.affine or torch.is_grad_enabled():
        return None
    fold = getattr(bn, "_dv_folded", None)
    if fold is None:
        fold = bn._dv_folded = _Folded()
    N, Cc, H, W = x.shape
    K = conv.out_channels
    stride, pad = conv.stride[0], conv.padding[0]
    R, S = conv.kernel_size
    P, Q = (H + 2 * pad - R) // stride + 1, (W + 2 * pad - S) // stride + 1
    dt = F._dv_dtype(x.dtype)
    if residual is not None:
        residual = F._nhwc(residual)
        if residual.shape != (N, K, P, Q) or residual.dtype != x.dtype:
            return None
    if conv._dv_depthwise:
        if residual is not None:
            return None
        fold.get(bn, conv.weight)
        y = torch.empty((N, K, P, Q), dtype=x.dtype, device=x.device, memory_format=F._CL)
        p = F._dw_params(x.data_ptr(), fold.dw.data_ptr(), y.data_ptr(), N, H, W, Cc, stride, dt)
        p.ep_shift, p.ep_relu = fold.shift.data_ptr(), int(relu)
        check(lib.dv_dwconv_fprop(C.byref(p), F._stream()), "dv_dwconv_fprop(folded)")
        return y
    fold.get(bn)
    y = torch.empty((N, K, P, Q), dtype=x.dtype, device=x.device, memory_format=F._CL)
    if Cc < 16 and R * S * Cc <= 256 and K % 8 == 0:   # C=3 stems: patch matrix + 1x1 tensor-core GEMM (functional._Conv2d)
        kp = -(-(R * S * Cc) // 64) * 64
        M = N * P * Q
        patches = torch.empty((M, kp), dtype=x.dtype, device=x.device)
        check(lib.dv_im2col(x.data_ptr(), patches.data_ptr(), N, H, W, Cc, R, S, stride, pad, kp, dt, F._stream()), "dv_im2col")
        w = conv._dv_cache.get_padded(conv.weight, x.dtype, kp)
        p = F._conv_params(patches.data_ptr(), w.data_ptr(), y.data_ptr(), None, M, 1, 1, kp, K, 1, 1, 1, 0, dt, DV_ALGO_TC)
    else:
        w, _ = conv._dv_cache.get(conv.weight, x.dtype, False)
        p = F._conv_params(x.data_ptr(), w.data_ptr(), y.data_ptr(), None, N, H, W, Cc, K, R, S, stride, pad, dt, config.conv_algo())
        if lib.dv_conv_algo(C.byref(p), DV_CONV_FPROP) != DV_ALGO_TC:
            return None
    p.ep_scale, p.ep_shift = fold.scale.data_ptr(), fold.shift.data_ptr()
    p.ep_residual = residual.data_ptr() if residual is not None else None
    p.ep_relu = int(relu)
    F._run_conv(lib.dv_conv_fprop, p, DV_CONV_FPROP, x.device, "dv_conv_fprop(folded)")
    return y


class CapturedForward:
    """The eval-mode forward of a model for ONE input shape, recorded into a CUDA graph and replayed: the serving loop of the
    reference's notebooks (cell 3: net.eval(); net(image)) without the per-launch host cost of ~60-110 Python-level calls.

        fwd = CapturedForward(net, example_batch)       # net.eval(), bf16 compute, weights as they are NOW
        logits = fwd(batch)                             # same shape/dtype as the example; returns a static output buffer

    The graph bakes in device pointers and the folded scale/shift of every BatchNorm2d: capture again after the weights or
    running statistics change (training steps, load_state_dict)."""

    def __init__(self, net, example):
        if net.training:
            raise RuntimeError("CapturedForward records the inference path: call net.eval() first")
        self.net = net
        self.static_in = example.clone()
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side), torch.no_grad():
            for _ in range(2):                      # allocator, filter copies, folded BN tables
                net(self.static_in)
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        with torch.no_grad(), torch.cuda.graph(self.graph):
            self.static_out = net(self.static_in)

    def __call__(self, x):
        if x.shape != self.static_in.shape or x.dtype != self.static_in.dtype:
            raise RuntimeError(f"CapturedForward was recorded for {tuple(self.static_in.shape)} {self.static_in.dtype}, got "
                               f"{tuple(x.shape)} {x.dtype}")
        self.static_in.copy_(x, non_blocking=True)
        self.graph.replay()
        return self.static_out
