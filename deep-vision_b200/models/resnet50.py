"""ResNet50 V1 (ResNet/pytorch/models/resnet50.py:8-93): Bottleneck stages of 3/4/6/3 blocks."""
from ._resnet import BottleneckBlock, ResNetScaffold, bottleneck_stage

__all__ = ["ResNet50", "BottleneckBlock"]


class ResNet50(ResNetScaffold):
    def __init__(self):
        super().__init__([(3, 64, 64, 256, 1), (4, 256, 128, 512, 2), (6, 512, 256, 1024, 2), (3, 1024, 512, 2048, 2)],
                         features=2048)

    def _make_blocks(self, num_blocks, in_channels, out1, out2, stride):
        return bottleneck_stage(num_blocks, in_channels, out1, out2, stride)
