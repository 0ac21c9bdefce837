"""Shared pieces of the three ResNet files: the stem/tail scaffold and the two residual blocks.

Module registration order is kept identical to the reference (ResNet/pytorch/models/resnet50.py:12-45,97-145;
resnet34.py:12-46,93-128) because it fixes both the state_dict key order and the order in which constructors and
`_initialize_weights` consume the RNG — with the same seed the B200 model and the reference start from the same weights.
"""
import torch.nn as tnn

from .. import functional as F
from .. import nn as dvnn


def _conv(cin, cout, k, stride=1, pad=0):
    return dvnn.Conv2d(cin, cout, kernel_size=k, stride=stride, padding=pad, bias=False)


class ResNetScaffold(tnn.Module):
    """conv1 7x7/2 -> bn1 -> relu -> maxpool 3x3/2 -> conv2x..conv5x -> avgpool -> linear.

    `stage_plan` is a list of four tuples handed to the subclass's `_make_blocks`; `features` is the width that
    reaches the classifier."""

    def __init__(self, stage_plan, features):
        super().__init__()
        self.conv1 = _conv(3, 64, 7, stride=2, pad=3)
        self.bn1 = dvnn.BatchNorm2d(64)
        self.relu = dvnn.ReLU(inplace=True)
        self.maxpool = dvnn.MaxPool2d(kernel_size=3, stride=2, padding=1)
        for name, args in zip(("conv2x", "conv3x", "conv4x", "conv5x"), stage_plan):
            setattr(self, name, self._make_blocks(*args))
        self.avgpool = dvnn.AdaptiveAvgPool2d((1, 1))
        self.linear = dvnn.Linear(features, 1000)
        self._features = features
        self._initialize_weights()

    def forward(self, x):
        x = self.bn1.forward_relu_maxpool(self.conv1(x, defer=True), self.relu, self.maxpool)   # bn1 -> relu -> maxpool
        for stage in (self.conv2x, self.conv3x, self.conv4x, self.conv5x):
            x = stage(x)
        x = self.avgpool(x)
        x = x.view(x.size(0), self._features)
        return self.linear(x)

    def _initialize_weights(self):
        # He-normal (fan_out) for every convolution, BN gamma=1 beta=0, classifier left at torch's default:
        # resnet50.py:84-93 / resnet34.py:80-89.
        for m in self.modules():
            if isinstance(m, tnn.Conv2d):
                tnn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")
            elif isinstance(m, tnn.BatchNorm2d):
                tnn.init.constant_(m.weight, 1)
                tnn.init.constant_(m.bias, 0)


class _Residual(tnn.Module):
    def _shortcut(self, cin, cout, stride, downsample):
        self.projection = None
        self.downsample = downsample
        if downsample:
            self.projection = tnn.Sequential(_conv(cin, cout, 1, stride=stride), dvnn.BatchNorm2d(cout))

    def _shortcut_and_merge(self, x, main, last_bn, slot):
        """relu(last_bn(main) + shortcut(x)).  The shortcut is evaluated AFTER the main path on purpose: autograd then
        runs its backward first, so its gradient w.r.t. x is already in `slot` when the first convolution's dgrad
        accumulates on top of it (functional.GradSlot)."""
        if self.downsample:
            conv, bn = self.projection[0], self.projection[1]
            identity = bn(conv(x, grad_slot=slot, role="deposit" if slot is not None else None, defer=True))
            return last_bn(main, residual=identity, act=self.relu)
        return last_bn(main, residual=x, act=self.relu, res_slot=slot)

    @staticmethod
    def _slot_for(x):
        import torch
        return F.GradSlot() if (torch.is_grad_enabled() and x.requires_grad) else None


class BasicBlock(_Residual):
    """3x3(stride) -> BN -> ReLU -> 3x3 -> BN -> (+shortcut) -> ReLU   (resnet34.py:92-143)."""

    def __init__(self, in_channels, out_channels, stride=1, downsample=False):
        super().__init__()
        self.conv1 = _conv(in_channels, out_channels, 3, stride=stride, pad=1)
        self.bn1 = dvnn.BatchNorm2d(out_channels)
        self.conv2 = _conv(out_channels, out_channels, 3, stride=1, pad=1)
        self.bn2 = dvnn.BatchNorm2d(out_channels)
        self.relu = dvnn.ReLU(inplace=True)
        self._shortcut(in_channels, out_channels, stride, downsample)

    def forward(self, x):
        x = F.to_compute_layout(x)  # no-op inside a network; converts a raw fp32 NCHW tensor fed to a lone block
        slot = self._slot_for(x)
        out = self.bn1(self.conv1(x, grad_slot=slot, role="collect" if slot is not None else None, defer=True), act=self.relu)
        out = self.conv2(out, defer=True)
        return self._shortcut_and_merge(x, out, self.bn2, slot)


class BottleneckBlock(_Residual):
    """1x1(stride) -> BN -> ReLU -> 3x3 -> BN -> ReLU -> 1x1 -> BN -> (+shortcut) -> ReLU   (resnet50.py:96-165).
    The stride sits on the FIRST 1x1 (ResNet V1), not on the 3x3."""

    def __init__(self, in_channels, out1, out2, stride=1, downsample=False):
        super().__init__()
        self.conv1 = _conv(in_channels, out1, 1, stride=stride)
        self.bn1 = dvnn.BatchNorm2d(out1)
        self.conv2 = _conv(out1, out1, 3, stride=1, pad=1)
        self.bn2 = dvnn.BatchNorm2d(out1)
        self.conv3 = _conv(out1, out2, 1)
        self.bn3 = dvnn.BatchNorm2d(out2)
        self.relu = dvnn.ReLU(inplace=True)
        self._shortcut(in_channels, out2, stride, downsample)

    def forward(self, x):
        x = F.to_compute_layout(x)
        slot = self._slot_for(x)
        out = self.bn1(self.conv1(x, grad_slot=slot, role="collect" if slot is not None else None, defer=True), act=self.relu)
        out = self.bn2(self.conv2(out, defer=True), act=self.relu)
        out = self.conv3(out, defer=True)
        return self._shortcut_and_merge(x, out, self.bn3, slot)


def bottleneck_stage(num_blocks, in_channels, out1, out2, stride):
    first = BottleneckBlock(in_channels, out1, out2, stride=stride, downsample=True)
    rest = [BottleneckBlock(out2, out1, out2) for _ in range(1, num_blocks)]
    return tnn.Sequential(first, *rest)
