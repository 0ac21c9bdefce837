"""Drop-in torch.nn modules: same constructors, attribute names and state_dict layout as the torch classes the
reference instantiates, so `isinstance(m, nn.Conv2d)` in the reference's `_initialize_weights`
(ResNet/pytorch/models/resnet50.py:84-93) and checkpoint loading (ResNet/pytorch/train.py:293-307) keep working,
but `forward` runs the sm_100a kernels of libdv_b200.so through `functional`.
"""
import torch
import torch.nn as nn

from . import config
from . import functional as F
from . import inference


class Conv2d(nn.Conv2d):
    """Dense (groups=1) or depth-wise 3x3 (groups=in=out) convolution without bias."""

    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        self._dv_cache = F.FilterCache()
        self._dv_stats = None
        kh, kw = self.kernel_size
        self._dv_depthwise = self.groups > 1
        if self.bias is not None:
            raise NotImplementedError("deep_vision_b200.nn.Conv2d: the hot path has no biased convolution (resnet50.py:27)")
        if self.dilation != (1, 1) or self.padding_mode != "zeros" or self.stride[0] != self.stride[1] or \
                self.padding[0] != self.padding[1] or isinstance(self.padding, str):
            raise NotImplementedError("deep_vision_b200.nn.Conv2d: square stride/padding, no dilation, zero padding only")
        if self._dv_depthwise and not (self.groups == self.in_channels == self.out_channels and (kh, kw) == (3, 3)
                                       and self.padding == (1, 1)):
            raise NotImplementedError("deep_vision_b200.nn.Conv2d: grouped convolution is supported as depth-wise 3x3 pad 1 only")

    def forward(self, x, grad_slot=None, role=None, defer=False):
        """grad_slot/role: see functional.GradSlot — used by the residual blocks to fuse the shortcut's gradient into
        this convolution's dgrad; plain callers leave them unset.  defer: the caller hands the result straight to a
        BatchNorm2d; in inference mode the launch is then left to that module (inference.PendingConv)."""
        # a C <= 4 stem in training takes the loader's fp32 NCHW batch as it is: its direct path (dv_stem_pad_input_nchw) converts
        # while it pads, which saves the stand-alone layout pass over the batch
        raw_stem = (self.training and self.in_channels <= 4 and not self._dv_depthwise and x.dim() == 4 and x.dtype == torch.float32
                    and x.is_cuda and x.is_contiguous() and not x.requires_grad and config.compute_dtype() == torch.bfloat16
                    and config.stem_direct() and config.stem_raw_input() and config.conv_algo() != 1)
        if not raw_stem:
            x = F.to_compute_layout(x)
        if defer and not raw_stem and inference.wanted(self, x):
            return inference.PendingConv(x, self)
        stats = None
        if self.training and config.fuse_bn_stats() and (x.dtype == torch.bfloat16 or raw_stem) and config.conv_algo() != 1:
            if self._dv_stats is None or self._dv_stats.t.device != x.device:
                self._dv_stats = F.StatSums(self.out_channels, x.device)
            stats = self._dv_stats
        if self._dv_depthwise:
            return F.dwconv3x3(x, self.weight, self.stride[0], stats)
        return F.conv2d(x, self.weight, self.stride[0], self.padding[0], self._dv_cache, config.conv_algo(), grad_slot, role, stats)


class BatchNorm2d(nn.BatchNorm2d):
    """nn.BatchNorm2d with the following ReLU and the residual add folded into the same pass:
    ``bn(x, residual=identity, act=self.relu)`` == ``relu(bn(x) + identity)``."""

    def forward(self, x, residual=None, act=None, res_slot=None):
        if act is not None and not isinstance(act, nn.ReLU):
            raise TypeError("act must be an nn.ReLU module or None")
        if isinstance(x, inference.PendingConv):
            y = inference.conv_bn_act(x, self, residual, act is not None)
            if y is not None:
                return y
            x = x.materialize()
        x = F.to_compute_layout(x)
        eaf, training, rm, rv = self._bookkeeping()
        return F.batch_norm_act(x, residual, self.weight, self.bias, rm, rv, training, eaf, self.eps, act is not None, res_slot, self)

    def _bookkeeping(self):
        """nn.BatchNorm2d.forward's own bookkeeping: the counter, the averaging factor, which statistics are used and updated."""
        if self.momentum is None:
            eaf = 0.0
        else:
            eaf = self.momentum
        if self.training and self.track_running_stats and self.num_batches_tracked is not None:
            if not getattr(self, "_dv_external_nbt", False):  # the trainer bumps all counters with one launch
                self.num_batches_tracked.add_(1)
            if self.momentum is None:
                eaf = 1.0 / float(self.num_batches_tracked)
        training = self.training or (self.running_mean is None and self.running_var is None)
        if not self.affine:
            raise NotImplementedError("deep_vision_b200.nn.BatchNorm2d: affine=False is not on the hot path")
        rm = self.running_mean if (not self.training or self.track_running_stats) else None
        rv = self.running_var if (not self.training or self.track_running_stats) else None
        return eaf, training, rm, rv

    def forward_relu_maxpool(self, x, act, pool):
        """``pool(act(self(x)))`` for the ResNet stem (bn1 -> relu -> maxpool, resnet50.py:49-52).  In train mode the three run as
        one fused pass (functional.bn_relu_maxpool); otherwise — eval mode, folded inference, another pooling geometry, a channel
        count the fused kernels do not take — exactly the separate modules."""
        c = self.num_features
        v = 8 if config.compute_dtype() == torch.bfloat16 else 4
        fusable = (self.training and config.fuse_stem_pool() and self.affine and isinstance(act, nn.ReLU) and isinstance(pool, MaxPool2d)
                   and not isinstance(x, inference.PendingConv) and pool.kernel_size in (3, (3, 3)) and pool.stride in (2, (2, 2))
                   and pool.padding in (1, (1, 1)) and not pool.ceil_mode and not pool.return_indices and c % v == 0 and 256 % (c // v) == 0 and (c // v) <= 256)
        if not fusable:
            return pool(self(x, act=act))
        x = F.to_compute_layout(x)
        eaf, _, rm, rv = self._bookkeeping()
        return F.bn_relu_maxpool(x, self.weight, self.bias, rm, rv, eaf, self.eps)


class ReLU(nn.ReLU):
    """Stand-alone ReLU.  Inside the blocks it is never launched on its own: the module is handed to
    BatchNorm2d.forward(act=...) and fused there."""

    def forward(self, x):
        return F.relu(x)


class MaxPool2d(nn.MaxPool2d):
    def forward(self, x):
        ks = self.kernel_size if isinstance(self.kernel_size, tuple) else (self.kernel_size,) * 2
        st = self.stride if isinstance(self.stride, tuple) else (self.stride,) * 2
        pd = self.padding if isinstance(self.padding, tuple) else (self.padding,) * 2
        if ks != (3, 3) or st != (2, 2) or pd != (1, 1) or self.ceil_mode or self.return_indices:
            raise NotImplementedError("deep_vision_b200.nn.MaxPool2d: only MaxPool2d(3, 2, 1) (resnet50.py:34)")
        return F.max_pool3x3s2(F.to_compute_layout(x))


class AdaptiveAvgPool2d(nn.AdaptiveAvgPool2d):
    def forward(self, x):
        if self.output_size not in ((1, 1), 1):
            raise NotImplementedError("deep_vision_b200.nn.AdaptiveAvgPool2d: only output size (1, 1) (resnet50.py:43)")
        return F.global_avg_pool(F.to_compute_layout(x))


class Linear(nn.Linear):
    """Classifier head; returns fp32 logits whatever the compute dtype."""

    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        self._dv_cache = F.FilterCache()

    def forward(self, x):
        return F.linear(x, self.weight, self.bias, self._dv_cache)
