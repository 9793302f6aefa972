"""generator_11 — drop-in for the reference's generators/generator_11.py (units/get_generators.get_11): generator_10_2
with ProGAN's pixelwise feature-vector normalisation (`Pixnorm`, csrc/pixnorm.cu) in the place of every BatchNorm,
equalised transposed convolutions, 64 base planes and one-block level-1 trees.  Same module tree, construction order and
RNG consumption as the reference file."""
import random

import numpy
import torch
import torch.nn as nn

from gan_b200 import layers as L
from gan_b200 import ops
from generators import generator_10 as g10
from generators.generator_10_2 import EqualizedConv2d


class Pixnorm(torch.nn.Module):
    """reference :9-16 — x / sqrt(1e-8 + mean over channels of x^2)."""

    def __init__(self, in_planes):
        super(Pixnorm, self).__init__()
        self.in_planes = in_planes

    def forward(self, x):
        assert x.shape[1] == self.in_planes
        return ops.pixnorm(x, 1e-8)


class EqualizedConvTranspose2D(torch.nn.Module):
    """reference :92-103 — `convT.weight` ~ N(0,1), `bias` ~ N(0,1) held by the wrapper, y = convT(x * scale) + bias with
    scale = sqrt(2 / in_planes): the scale is the alpha of the transposed-convolution kernel."""

    def __init__(self, in_planes, out_planes, kernel_size, stride=1, padding=0):
        super(EqualizedConvTranspose2D, self).__init__()
        self.convT = L.ConvTranspose2d(in_planes, out_planes, kernel_size, stride, padding)
        self.scale = numpy.sqrt(2) / numpy.sqrt(in_planes)
        self.bias = self.convT.bias
        self.convT.bias = None
        nn.init.normal_(self.convT.weight)
        nn.init.normal_(self.bias)

    def forward(self, x):
        c = self.convT
        assert c.groups == 1 and c.dilation[0] == 1 and c.output_padding[0] == 0
        return ops.conv_transpose2d(x, c.weight, self.bias, c.stride[0], c.padding[0], alpha=float(self.scale))


class SelfAttention(g10.SelfAttention):
    """reference :107-129"""

    def __init__(self, in_planes, embedding_channels):
        nn.Module.__init__(self)
        self.key = EqualizedConv2d(in_planes=in_planes, out_planes=embedding_channels, kernel_size=1)
        self.query = EqualizedConv2d(in_planes=in_planes, out_planes=embedding_channels, kernel_size=1)
        self.value = EqualizedConv2d(in_planes=in_planes, out_planes=embedding_channels, kernel_size=1)
        self.self_att = EqualizedConv2d(in_planes=embedding_channels, out_planes=in_planes, kernel_size=1)
        self.gamma = nn.Parameter(torch.tensor(random.random() * 0.2 + 0.03))
        self.softmax = nn.Softmax(dim=1)


class SEBottleneck(g10.SEBottleneckSelfAttention):
    """reference :133-172"""

    def get_conv_bottleneck(self, last_planes, in_planes, out_planes, dense_depth, kernel_size):
        return nn.Sequential(
            EqualizedConv2d(in_planes=last_planes, out_planes=in_planes, kernel_size=1, stride=1, padding=0),
            Pixnorm(in_planes),
            L.PReLU(),
            EqualizedConv2d(in_planes=in_planes, out_planes=in_planes, kernel_size=kernel_size, stride=1,
                            padding=(kernel_size - 1) // 2, groups=in_planes, padding_mode='reflect'),
            Pixnorm(in_planes),
            L.PReLU(),
            EqualizedConv2d(in_planes=in_planes, out_planes=out_planes + dense_depth, kernel_size=1, stride=1,
                            padding=0),
            Pixnorm(out_planes + dense_depth),
        )

    def get_se(self, in_planes, out_planes, dense_depth, feature_size):
        return nn.Sequential(
            L.AvgPool2d(kernel_size=feature_size, padding=0),
            EqualizedConv2d(in_planes=out_planes + dense_depth, out_planes=in_planes, kernel_size=1),
            L.PReLU(),
            EqualizedConv2d(in_planes=in_planes, out_planes=out_planes + dense_depth, kernel_size=1),
            L.Sigmoid(),
        )


class ResnetInit(g10.ResnetInit):
    """reference :176-199"""
    BOTTLENECK = SEBottleneck


class BasicBlock(g10.BasicBlock):
    """reference :203-257"""
    RIR = ResnetInit
    ATTENTION = SelfAttention

    @staticmethod
    def _pointwise(in_planes, out_planes):
        return [EqualizedConv2d(in_planes=in_planes, out_planes=out_planes, kernel_size=1), Pixnorm(out_planes)]


class Tree(g10.Tree):
    """reference :261-319"""
    BLOCK = BasicBlock


class Generator(g10.Generator):
    """reference :322-405"""
    PLANES = 64
    TREE = Tree
    ATTENTION = SelfAttention

    def get_upsample(self, planes, out_planes, kernel_size, stride, padding):
        return nn.Sequential(
            EqualizedConvTranspose2D(in_planes=planes, out_planes=out_planes, kernel_size=kernel_size, stride=stride,
                                     padding=padding),
            Pixnorm(out_planes),
            L.PReLU(),
        )

    def _stages(self, planes):
        return [(planes * 8, planes * 2, planes // 4, 1, 1, 8), (planes * 4, planes, planes // 8, 1, 1, 16),
                (planes * 2, planes // 2, planes // 16, 1, 1, 32), (planes, planes // 4, planes // 32, 1, 1, 64)]

    def _last_conv(self, last_planes):
        return EqualizedConv2d(in_planes=last_planes, out_planes=3, kernel_size=9, stride=1, padding=4,
                               padding_mode='reflect')
