"""The four model files of the hot path, constructor- and state_dict-compatible with the reference."""
from .mobilenet_v1 import DepthwiseConv, DepthwiseSeparableConv, MobileNetV1, PointwiseConv
from .resnet34 import BasicBlock, ResNet34
from .resnet50 import BottleneckBlock, ResNet50
from .resnet152 import ResNet152

__all__ = ["ResNet34", "ResNet50", "ResNet152", "MobileNetV1", "BasicBlock", "BottleneckBlock",
           "DepthwiseSeparableConv", "DepthwiseConv", "PointwiseConv"]
