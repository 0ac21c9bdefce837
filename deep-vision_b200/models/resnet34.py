"""ResNet34 as the reference codes it (ResNet/pytorch/models/resnet34.py:8-89): four stages of TWO BasicBlocks each
— the 18-layer depth — with a 1x1 projection shortcut on the first block of every stage, conv2x included."""
from ._resnet import BasicBlock, ResNetScaffold
import torch.nn as tnn

__all__ = ["ResNet34", "BasicBlock"]


class ResNet34(ResNetScaffold):
    def __init__(self):
        super().__init__([(2, 64, 64, 1), (2, 64, 128, 2), (2, 128, 256, 2), (2, 256, 512, 2)], features=512)

    def _make_blocks(self, num_blocks, in_channels, out_channels, stride):
        blocks = [BasicBlock(in_channels, out_channels, stride=stride, downsample=True)]
        blocks += [BasicBlock(out_channels, out_channels) for _ in range(1, num_blocks)]
        return tnn.Sequential(*blocks)
