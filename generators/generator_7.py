"""generator_7 — drop-in for the reference's generators/generator_7.py (units/get_generators.get_7): three transposed
convolutions z -> 16x16, then three deep-layer-aggregation trees at 16x16, 32x32 and 64x64 joined by transposed
convolutions, a reflect-padded 9x9 convolution to RGB and tanh.  The tree is the one of discriminator_6_wgangp
(ResNet-in-ResNet blocks with depthwise 3x3 / 5x5 / 7x7 bottlenecks and SE gates) with BatchNorm2d after every
convolution of a bottleneck and scalar PReLUs; `ResnetInit`, `BasicBlock` and `Tree` below are that module's classes
with those parts swapped in.  Same module tree, construction order and RNG consumption as the reference file.
generator_8 uses the same classes under another top level."""
import torch.nn as nn

from discriminators import discriminator_6_wgangp as tree
from gan_b200 import layers as L


class SqueezeExcitationBottleneck(tree.SqueezeExcitationBottleneck):
    """reference :5-38 — (1x1, BN, PReLU, depthwise k x k reflect, BN, PReLU, 1x1, BN) times its SE gate."""

    def get_conv_bottleneck(self, last_planes, in_planes, out_planes, dense_depth, kernel_size):
        return nn.Sequential(
            L.Conv2d(last_planes, in_planes, kernel_size=1, stride=1, padding=0),
            L.BatchNorm2d(in_planes),
            L.PReLU(),
            L.Conv2d(in_planes, in_planes, kernel_size=kernel_size, stride=1, padding=(kernel_size - 1) // 2,
                     padding_mode='reflect', groups=in_planes),
            L.BatchNorm2d(in_planes),
            L.PReLU(),
            L.Conv2d(in_planes, out_planes + dense_depth, kernel_size=1, stride=1, padding=0),
            L.BatchNorm2d(out_planes + dense_depth),
        )

    def get_se(self, in_planes, out_planes, dense_depth, feature_size):
        return nn.Sequential(
            L.AvgPool2d(kernel_size=feature_size, padding=0),
            L.Conv2d(out_planes + dense_depth, in_planes, kernel_size=1),
            L.PReLU(),
            L.Conv2d(in_planes, out_planes + dense_depth, kernel_size=1),
            L.Sigmoid(),
        )


class ResnetInit(tree.ResnetInit):
    """reference :41-64"""
    BOTTLENECK = SqueezeExcitationBottleneck


class BasicBlock(tree.BasicBlock):
    """reference :67-128 — `unify` / `shortcut` are (1x1 conv, BatchNorm2d)."""
    RIR = ResnetInit

    @staticmethod
    def _pointwise(in_planes, out_planes):
        return [L.Conv2d(in_planes, out_planes, kernel_size=1, stride=1, padding=0), L.BatchNorm2d(out_planes)]


class Tree(tree.Tree):
    """reference :131-192"""
    BLOCK = BasicBlock


class Generator(nn.Module):
    """reference :195-238"""
    negative_slope = 1e-2

    def get_upsample(self, planes, out_planes, kernel_size, stride, padding):
        return nn.Sequential(
            L.ConvTranspose2d(planes, out_planes, kernel_size=kernel_size, stride=stride, padding=padding),
            L.BatchNorm2d(out_planes),
            L.PReLU(),
        )

    def __init__(self, z_dim):
        super(Generator, self).__init__()
        self.model = nn.Sequential()
        up = lambda cin, cout, k=4, s=2, p=1: self.get_upsample(planes=cin, out_planes=cout, kernel_size=k, stride=s,
                                                                padding=p)
        self.model.add_module("Upsample_1", up(z_dim, 256 * 4, 4, 1, 0))    # 4
        self.model.add_module("Upsample_2", up(256 * 4, 256 * 2))           # 8
        self.model.add_module("Upsample_3", up(256 * 2, 256))               # 16
        last_planes = 256
        plans = [dict(in_planes=64, out_planes=64, dense_depth=8, level=2, block_num=4, feature_size=16),
                 dict(in_planes=32, out_planes=32, dense_depth=4, level=2, block_num=3, feature_size=32),
                 dict(in_planes=16, out_planes=16, dense_depth=4, level=1, block_num=2, feature_size=64)]
        for n, plan in enumerate(plans, start=1):
            t = Tree(last_planes, **plan)
            last_planes = t.get_out_planes()
            self.model.add_module("Tree_%d" % n, t)
            if n < len(plans):
                self.model.add_module("Upsample_%d" % (n + 3), up(last_planes, last_planes))
        self.model.add_module("LastConv", L.Conv2d(last_planes, 3, 9, stride=1, padding=4, padding_mode='reflect'))
        self.model.add_module("LastActivate", L.Tanh())

    def forward(self, x):
        return self.model(x)
