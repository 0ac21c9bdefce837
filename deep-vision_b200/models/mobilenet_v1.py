"""MobileNetV1 (MobileNet/pytorch/models/mobilenet_v1.py:10-156): 3x3/2 stem, 13 depth-wise-separable blocks,
global average pool, classifier.  No weight-init routine (torch constructor defaults), stem BN hard-wired to 32
channels — both as in the reference."""
import torch.nn as tnn

from .. import nn as dvnn

__all__ = ["MobileNetV1", "DepthwiseSeparableConv", "DepthwiseConv", "PointwiseConv"]

# (output width multiplier of 32, depth-wise stride) for the 13 blocks, Table 1 of the MobileNet paper;
# mobilenet_v1.py:35-81.  The last block keeps stride 1 (the reference notes the paper's typo).
_PLAN = [(2, 1), (4, 2), (4, 1), (8, 2), (8, 1), (16, 2), (16, 1), (16, 1), (16, 1), (16, 1), (16, 1), (32, 2), (32, 1)]


class MobileNetV1(tnn.Module):
    def __init__(self, alpha=1):
        super().__init__()
        self.alpha = alpha
        layers = [dvnn.Conv2d(3, alpha * 32, 3, padding=1, stride=2, bias=False), dvnn.BatchNorm2d(32), dvnn.ReLU(inplace=True)]
        width = 32
        for mult, dw_stride in _PLAN:
            layers.append(DepthwiseSeparableConv(alpha * width, alpha * 32 * mult, dw_stride=dw_stride, pw_stride=1))
            width = 32 * mult
        layers.append(dvnn.AdaptiveAvgPool2d((1, 1)))
        self.features = tnn.Sequential(*layers)
        self.linear = dvnn.Linear(alpha * 1024, 1000)

    def forward(self, x):
        f = self.features
        x = f[1](f[0](x, defer=True), act=f[2])          # conv -> BN+ReLU in one pass
        for i in range(3, len(f)):           # 13 separable blocks, then the pool
            x = f[i](x)
        x = x.view(x.size(0), self.alpha * 1024)
        return self.linear(x)


class DepthwiseSeparableConv(tnn.Module):
    def __init__(self, in_channels, out_channels, dw_stride, pw_stride):
        super().__init__()
        self.dw = DepthwiseConv(in_channels, in_channels, stride=dw_stride)
        self.pw = PointwiseConv(in_channels, out_channels, stride=pw_stride)

    def forward(self, x):
        return self.pw(self.dw(x))


class _ConvBnRelu(tnn.Module):
    def forward(self, x):
        return self.bn(self.conv(x, defer=True), act=self.relu)


class DepthwiseConv(_ConvBnRelu):
    def __init__(self, in_channels, out_channels, stride):
        super().__init__()
        self.conv = dvnn.Conv2d(in_channels, out_channels, 3, padding=1, stride=stride, groups=in_channels, bias=False)
        self.bn = dvnn.BatchNorm2d(out_channels)
        self.relu = dvnn.ReLU(inplace=True)


class PointwiseConv(_ConvBnRelu):
    def __init__(self, in_channels, out_channels, stride):
        super().__init__()
        self.conv = dvnn.Conv2d(in_channels, out_channels, 1, stride=stride, bias=False)
        self.bn = dvnn.BatchNorm2d(out_channels)
        self.relu = dvnn.ReLU(inplace=True)
