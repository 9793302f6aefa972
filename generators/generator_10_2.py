"""generator_10_2 — drop-in for the reference's generators/generator_10_2.py (units/get_generators.get_10_2):
generator_10 with every convolution replaced by an equalised-learning-rate one (N(0,1) weights and biases, the input
scaled by sqrt(2 / fan_in) at run time), 32 base planes, two-level trees with one block per level.  The run-time scale is
the `alpha` of the convolution kernels (conv(x * scale) = scale * conv(x)): no scaled copy of the activation is made.
Same module tree, construction order and RNG consumption as the reference file."""
import random

import numpy
import torch
import torch.nn as nn

from gan_b200 import layers as L
from gan_b200 import ops
from generators import generator_10 as g10


class EqualizedConv2d(torch.nn.Module):
    """reference :7-19 — `conv.weight` ~ N(0,1), `bias` ~ N(0,1) held by the wrapper (the inner conv has none),
    y = conv(x * scale) + bias with scale = sqrt(2) / sqrt(k * k * in_planes) (in_planes also for depthwise convs)."""

    def __init__(self, in_planes, out_planes, kernel_size, stride=1, padding=0, groups=1, padding_mode='reflect'):
        super(EqualizedConv2d, self).__init__()
        self.conv = L.Conv2d(in_planes, out_planes, kernel_size, stride, padding, groups=groups,
                             padding_mode=padding_mode)
        self.scale = numpy.sqrt(2) / numpy.sqrt(kernel_size * kernel_size * in_planes)
        self.bias = self.conv.bias
        self.conv.bias = None
        nn.init.normal_(self.conv.weight)
        nn.init.normal_(self.bias)

    def forward(self, x):
        c = self.conv
        pad = c.padding[0]
        if pad:                                    # padding_mode is 'reflect' everywhere in this file
            assert c.padding_mode == 'reflect'
            x, pad = ops.reflect_pad(x, pad), 0
        if c.groups == 1:
            return ops.conv2d(x, c.weight, self.bias, c.stride[0], pad, 0, alpha=float(self.scale))
        assert c.groups == c.in_channels == c.out_channels, "grouped convs: depthwise only"
        return ops.depthwise_conv2d(x, c.weight, self.bias, c.stride[0], pad, alpha=float(self.scale))


class SelfAttention(g10.SelfAttention):
    """reference :22-45"""

    def __init__(self, in_planes, embedding_channels):
        nn.Module.__init__(self)
        self.key = EqualizedConv2d(in_planes=in_planes, out_planes=embedding_channels, kernel_size=1)
        self.query = EqualizedConv2d(in_planes=in_planes, out_planes=embedding_channels, kernel_size=1)
        self.value = EqualizedConv2d(in_planes=in_planes, out_planes=embedding_channels, kernel_size=1)
        self.self_att = EqualizedConv2d(in_planes=embedding_channels, out_planes=in_planes, kernel_size=1)
        self.gamma = nn.Parameter(torch.tensor(random.random() * 0.2 + 0.03))
        self.softmax = nn.Softmax(dim=1)


class SEBottleneck(g10.SEBottleneckSelfAttention):
    """reference :49-89"""

    def get_conv_bottleneck(self, last_planes, in_planes, out_planes, dense_depth, kernel_size):
        return nn.Sequential(
            EqualizedConv2d(in_planes=last_planes, out_planes=in_planes, kernel_size=1, stride=1, padding=0),
            L.BatchNorm2d(in_planes),
            L.PReLU(),
            EqualizedConv2d(in_planes=in_planes, out_planes=in_planes, kernel_size=kernel_size, stride=1,
                            padding=(kernel_size - 1) // 2, groups=in_planes, padding_mode='reflect'),
            L.BatchNorm2d(in_planes),
            L.PReLU(),
            EqualizedConv2d(in_planes=in_planes, out_planes=out_planes + dense_depth, kernel_size=1, stride=1,
                            padding=0),
            L.BatchNorm2d(out_planes + dense_depth),
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
    """reference :93-116"""
    BOTTLENECK = SEBottleneck


class BasicBlock(g10.BasicBlock):
    """reference :120-175"""
    RIR = ResnetInit
    ATTENTION = SelfAttention

    @staticmethod
    def _pointwise(in_planes, out_planes):
        return [EqualizedConv2d(in_planes=in_planes, out_planes=out_planes, kernel_size=1), L.BatchNorm2d(out_planes)]


class Tree(g10.Tree):
    """reference :178-237"""
    BLOCK = BasicBlock


class Generator(g10.Generator):
    """reference :240-322"""
    PLANES = 32
    TREE = Tree
    ATTENTION = SelfAttention

    def _stages(self, planes):
        return [(planes * 8, planes * 2, planes // 4, 2, 1, 8), (planes * 4, planes, planes // 8, 2, 1, 16),
                (planes * 2, planes // 2, planes // 16, 2, 1, 32), (planes, planes // 4, planes // 32, 2, 1, 64)]

    def _last_conv(self, last_planes):
        return EqualizedConv2d(in_planes=last_planes, out_planes=3, kernel_size=9, stride=1, padding=4,
                               padding_mode='reflect')
