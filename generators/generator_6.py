"""generator_6 — drop-in for the reference's generators/generator_6.py (units/get_generators.get_6): the aggregation
tree of discriminator_6_wgangp with BatchNorm2d + GELU bottlenecks, between a dual-stream stem (transposed convolution
z -> 2x2, two StemBlocks of ResNet-in-ResNet pairs up to 8x8) and a dual-stream LastBlock at 64x64; three trees at 8x8,
16x16 and 32x32 joined by transposed convolutions, 9x9 reflect-padded convolution to RGB, tanh.  GELU is the exact erf
form (csrc/pointwise.cu GANB200_U_GELU and its derivatives).  Same module tree, construction order and RNG consumption
as the reference file."""
import torch.nn as nn

from discriminators import discriminator_6_wgangp as tree
from gan_b200 import layers as L
from gan_b200 import ops


def _bottleneck(last_planes, in_planes, out_planes, kernel_size):
    """1x1, BN, GELU, depthwise k x k (reflect), BN, GELU, 1x1, BN"""
    return nn.Sequential(
        L.Conv2d(last_planes, in_planes, kernel_size=1, stride=1, padding=0), L.BatchNorm2d(in_planes), L.GELU(),
        L.Conv2d(in_planes, in_planes, kernel_size=kernel_size, stride=1, padding=(kernel_size - 1) // 2,
                 padding_mode='reflect', groups=in_planes), L.BatchNorm2d(in_planes), L.GELU(),
        L.Conv2d(in_planes, out_planes, kernel_size=1, stride=1, padding=0), L.BatchNorm2d(out_planes),
    )


def _se(in_planes, out_planes, feature_size):
    return nn.Sequential(
        L.AvgPool2d(kernel_size=feature_size, padding=0),
        L.Conv2d(out_planes, in_planes, kernel_size=1), L.GELU(),
        L.Conv2d(in_planes, out_planes, kernel_size=1), L.Sigmoid(),
    )


class SqueezeExcitationBottleneck(tree.SqueezeExcitationBottleneck):
    """reference :5-38"""

    def get_conv_bottleneck(self, last_planes, in_planes, out_planes, dense_depth, kernel_size):
        return _bottleneck(last_planes, in_planes, out_planes + dense_depth, kernel_size)

    def get_se(self, in_planes, out_planes, dense_depth, feature_size):
        return _se(in_planes, out_planes + dense_depth, feature_size)


class ResnetInit(tree.ResnetInit):
    """reference :41-64"""
    BOTTLENECK = SqueezeExcitationBottleneck


class BasicBlock(tree.BasicBlock):
    """reference :67-125"""
    RIR = ResnetInit

    @staticmethod
    def _pointwise(in_planes, out_planes):
        return [L.Conv2d(in_planes, out_planes, kernel_size=1, stride=1, padding=0), L.BatchNorm2d(out_planes)]


class Tree(tree.Tree):
    """reference :128-189"""
    BLOCK = BasicBlock


class StemSqueezeExcitationBottleneck(nn.Module):
    """reference :192-225 — a bottleneck without a dense tail."""
    negative_slope = 1e-2

    def get_conv_bottleneck(self, last_planes, in_planes, out_planes, kernel_size):
        return _bottleneck(last_planes, in_planes, out_planes, kernel_size)

    def get_se(self, in_planes, out_planes, feature_size):
        return _se(in_planes, out_planes, feature_size)

    def __init__(self, last_planes, in_planes, out_planes, kernel_size, feature_size):
        super().__init__()
        self.res = self.get_conv_bottleneck(last_planes, in_planes, out_planes, kernel_size)
        self.se = self.get_se(in_planes, out_planes, feature_size)

    def forward(self, x):
        out = self.res(x)
        se = self.se(out)
        if se.shape[-2:] == (1, 1):
            return L.gate(out, se)
        return ops.mul(out, se.expand_as(out))


class StemResnetInit(nn.Module):
    """reference :228-258 — the ResNet-in-ResNet pair with an identity (or projected) skip on the residual stream."""
    BOTTLENECK = StemSqueezeExcitationBottleneck

    def __init__(self, last_planes, in_planes, out_planes, kernel_size, feature_size):
        super().__init__()
        mk = lambda: self.BOTTLENECK(last_planes, in_planes, out_planes, kernel_size, feature_size)
        self.residual = mk()
        self.transient = mk()
        self.residual_across = mk()
        self.transient_across = mk()
        self.shortcut = nn.Sequential()
        if last_planes != out_planes:
            self.shortcut = nn.Sequential(L.Conv2d(last_planes, out_planes, kernel_size=1, stride=1, padding=0),
                                          L.BatchNorm2d(out_planes))

    def forward(self, x):
        x_residual, x_transient = x
        res = ops.add(ops.add(self.residual(x_residual), self.transient_across(x_transient)), self.shortcut(x_residual))
        return res, ops.add(self.residual_across(x_residual), self.transient(x_transient))


class _DualStream(nn.Module):
    """StemBlock / LastBlock: two 3x3 + BN + GELU stems, `block_num` ResNet-in-ResNet pairs, concatenation"""
    RIR = StemResnetInit

    def _build(self, last_planes, in_planes, out_planes, block_num, kernel_size, feature_size):
        self.block_num = block_num
        pre = lambda: nn.Sequential(L.Conv2d(last_planes, out_planes, 3, padding=1), L.BatchNorm2d(out_planes), L.GELU())
        self.residual_pre_conv = pre()
        self.transient_pre_conv = pre()
        for i in range(block_num):
            self.__setattr__('rir_%d' % i, self.RIR(out_planes, in_planes, out_planes, kernel_size, feature_size))

    def _streams(self, x):
        pair = (self.residual_pre_conv(x), self.transient_pre_conv(x))
        for i in range(self.block_num):
            pair = self.__getattr__('rir_%d' % i)(pair)
        return ops.cat_channels(list(pair))


class StemBlock(_DualStream):
    """reference :261-300 — ... then ConvTranspose2d x2 + BN + GELU."""
    negative_slope = 1e-2

    def __init__(self, last_planes, in_planes, out_planes, block_num, kernel_size, feature_size):
        super(StemBlock, self).__init__()
        self._build(last_planes, in_planes, out_planes, block_num, kernel_size, feature_size)
        self.upsample = nn.Sequential(L.ConvTranspose2d(out_planes * 2, out_planes, kernel_size=4, stride=2, padding=1),
                                      L.BatchNorm2d(out_planes), L.GELU())

    def forward(self, x):
        return self.upsample(self._streams(x))


class LastSqueezeExcitationBottleneck(StemSqueezeExcitationBottleneck):
    """reference :303-336"""


class LastResnetInit(StemResnetInit):
    """reference :339-369"""
    BOTTLENECK = LastSqueezeExcitationBottleneck


class LastBlock(_DualStream):
    """reference :372-404"""
    negative_slope = 1e-2
    RIR = LastResnetInit

    def __init__(self, last_planes, in_planes, out_planes, block_num, kernel_size, feature_size):
        super(LastBlock, self).__init__()
        self._build(last_planes, in_planes, out_planes, block_num, kernel_size, feature_size)

    def forward(self, x):
        return self._streams(x)


class Generator(nn.Module):
    """reference :407-459"""
    negative_slope = 1e-2

    def get_upsample(self, planes):
        return nn.Sequential(L.ConvTranspose2d(planes, planes, kernel_size=4, stride=2, padding=1),
                             L.BatchNorm2d(planes), L.GELU())

    def __init__(self, z_dim):
        super(Generator, self).__init__()
        self.model = nn.Sequential()
        self.model.add_module("Stem_1", nn.Sequential(L.ConvTranspose2d(z_dim, 128, kernel_size=4, stride=2, padding=1),
                                                      L.BatchNorm2d(128), L.GELU()))
        self.model.add_module("Stem_2", StemBlock(last_planes=128, in_planes=32, out_planes=64, block_num=2,
                                                  kernel_size=3, feature_size=2))
        self.model.add_module("Stem_3", StemBlock(last_planes=64, in_planes=16, out_planes=32, block_num=2,
                                                  kernel_size=3, feature_size=4))
        last_planes = 32
        plans = [dict(in_planes=8, out_planes=16, feature_size=8), dict(in_planes=4, out_planes=8, feature_size=16),
                 dict(in_planes=4, out_planes=4, feature_size=32)]
        for n, plan in enumerate(plans, start=1):
            t = Tree(last_planes, dense_depth=1, level=2, block_num=2, **plan)
            last_planes = t.get_out_planes()
            self.model.add_module("Tree_%d" % n, t)
            self.model.add_module("Upsample_%d" % n, self.get_upsample(planes=last_planes))
        self.model.add_module("Last_block", LastBlock(last_planes=last_planes, in_planes=4, out_planes=4, block_num=2,
                                                      kernel_size=3, feature_size=64))
        self.model.add_module("LastConv", L.Conv2d(4 * 2, 3, 9, stride=1, padding=4, padding_mode='reflect'))
        self.model.add_module("LastActivate", L.Tanh())

    def forward(self, x):
        return self.model(x)
