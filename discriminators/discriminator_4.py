"""discriminator_4 — drop-in for the reference's discriminators/discriminator_4.py (the discriminator behind
`get_discriminator.get_4`, for the BCE loops): a densely connected ShuffleNet-style stack — a two-conv stem, four stages
of three DenseBlocks (channel split, 3x3 / 5x5 / 7x7 pre-activated branches with squeeze-excitation gates, a 1x1 dense
cut, channel shuffle) each closed by a stride-2 DownBlock, a 4x4 valid conv, Linear(128, 1) and a sigmoid.

Same class names, attribute names and registration order as the reference (state_dict keys `main.0.layers.1.b5x5.se_2.0.weight`,
...).  The reference imports torchsummary at module top (:6); nothing here needs it.  All arithmetic on libganb200: zero-padded
k x k convs and the strided 3x3 on the conv engine, the SE's 1x1 convs on (b,c,1,1) maps on the one-launch linear kernels,
BatchNorm (twice differentiable), LeakyReLU, sigmoid, gate * x + skip as one fused pass, channel slice / concat / shuffle
on the channel kernels."""
import math

import torch.nn as nn

from gan_b200 import layers as L
from gan_b200 import ops


class ShuffleBlock(nn.Module):
    """reference :9-18 — [N,C,H,W] -> [N,g,C/g,H,W] -> [N,C/g,g,H,W] -> [N,C,H,W]."""

    def __init__(self, groups=2):
        super(ShuffleBlock, self).__init__()
        self.groups = groups

    def forward(self, x):
        return ops.channel_shuffle(x, self.groups)


class SplitBlock(nn.Module):
    """reference :20-27."""

    def __init__(self, ratio):
        super(SplitBlock, self).__init__()
        self.ratio = ratio

    def forward(self, x):
        c = int(int(x.size(1)) * self.ratio)
        return ops.slice_channels(x, 0, c), ops.slice_channels(x, c)


def _bn_act(planes):
    return [L.BatchNorm2d(planes), L.LeakyReLU(0.2, inplace=True)]


def _se(planes):
    return nn.Sequential(L.Conv2d(planes, planes // 4, kernel_size=1), L.LeakyReLU(0.2, inplace=True),
                         L.Conv2d(planes // 4, planes, kernel_size=1), L.Sigmoid())


def _squeeze(x):
    """F.avg_pool2d(x, x.size(2)) on a square map: (b,c,h,h) -> (b,c,1,1)."""
    b, c = x.shape[:2]
    return ops.global_avg_pool(x).reshape(b, c, 1, 1)


class BasicModel(nn.Module):
    """reference :30-87 — BN, act, conv (b1); two more pre-activated convs from b1's output: b2 keeps `planes` channels
    (gated, plus the block input), b3 makes `dense_depth` new ones (gated)."""

    def __init__(self, planes, dense_depth, kernel_size):
        super(BasicModel, self).__init__()
        pad = (kernel_size - 1) // 2
        self.preact = nn.Sequential(*_bn_act(planes))
        self.b1 = nn.Sequential(L.Conv2d(planes, planes, kernel_size=kernel_size, stride=1, padding=pad, bias=False))
        self.b2 = nn.Sequential(*_bn_act(planes),
                                L.Conv2d(planes, planes, kernel_size=kernel_size, stride=1, padding=pad, bias=False))
        self.b3 = nn.Sequential(*_bn_act(planes),
                                L.Conv2d(planes, dense_depth, kernel_size=kernel_size, stride=1, padding=pad, bias=False))
        self.se_1 = _se(planes)
        self.se_2 = _se(dense_depth)

    def forward(self, x):
        out = self.b1(self.preact(x))
        out_b2 = self.b2(out)
        out_b3 = self.b3(out)
        out_b2 = L.gated_residual(x, out_b2, self.se_1(_squeeze(out_b2)))   # out_b2 * w_1 + x
        out_b3 = L.gate(out_b3, self.se_2(_squeeze(out_b3)))
        return ops.cat_channels([out_b2, out_b3])


class DenseBlock(nn.Module):
    """reference :90-135 — quarters of the input: the first passes through, the other three go through the 3x3 / 5x5 /
    7x7 BasicModels; their new channels are summed with a 1x1 dense cut of the whole input; shuffle over 4 groups."""

    def __init__(self, last_planes, dense_depth, split_ratio=0.5):
        super(DenseBlock, self).__init__()
        self.last_planes = last_planes // 4
        self.dense_depth = dense_depth
        self.split = SplitBlock(split_ratio)
        self.b3x3 = BasicModel(self.last_planes, dense_depth, 3)
        self.b5x5 = BasicModel(self.last_planes, dense_depth, 5)
        self.b7x7 = BasicModel(self.last_planes, dense_depth, 7)
        self.densecut = nn.Sequential(*_bn_act(last_planes),
                                      L.Conv2d(last_planes, dense_depth, kernel_size=1, stride=1, bias=False))
        self.shuffle = ShuffleBlock(groups=4)

    def forward(self, x):
        x1, x2 = self.split(x)
        x1_1, x1_2 = self.split(x1)
        x2_1, x2_2 = self.split(x2)
        out_12 = self.b3x3(x1_2)
        out_21 = self.b5x5(x2_1)
        out_22 = self.b7x7(x2_2)
        d = self.last_planes
        dense = self.densecut(x)
        for t in (out_12, out_21, out_22):
            dense = ops.add(dense, ops.slice_channels(t, d))
        out = ops.cat_channels([ops.slice_channels(out_12, 0, d), ops.slice_channels(out_21, 0, d),
                                ops.slice_channels(out_22, 0, d), x1_1, dense])
        return self.shuffle(out)


class DownBlock(nn.Module):
    """reference :138-169 — two pre-activated stride-2 branches (3x3 then 1x1 / 1x1 then 3x3), concat, shuffle."""

    def __init__(self, in_channels, out_channels):
        super(DownBlock, self).__init__()
        mid_channels = out_channels // 2
        self.b1 = nn.Sequential(*_bn_act(in_channels),
                                L.Conv2d(in_channels, in_channels, kernel_size=3, stride=2, padding=1, bias=False),
                                *_bn_act(in_channels),
                                L.Conv2d(in_channels, mid_channels, kernel_size=1, bias=False))
        self.b2 = nn.Sequential(*_bn_act(in_channels),
                                L.Conv2d(in_channels, mid_channels, kernel_size=1, bias=False),
                                *_bn_act(mid_channels),
                                L.Conv2d(mid_channels, mid_channels, kernel_size=3, stride=2, padding=1, bias=False))
        self.shuffle = ShuffleBlock(groups=2)

    def forward(self, x):
        return self.shuffle(ops.cat_channels([self.b1(x), self.b2(x)]))


class Block(nn.Module):
    """reference :172-189 — `blocks` DenseBlocks (16 new channels each) and a DownBlock."""

    def __init__(self, planes, out_planes, blocks, is_DownBlock=True):
        super(Block, self).__init__()
        self.is_DownBlock = is_DownBlock
        layers = []
        for i in range(blocks):
            lenth = math.sqrt(planes)
            layers.append(DenseBlock(int(planes), 16))
            planes = planes + 16
        self.layers = nn.Sequential(*layers)
        self.downblock = DownBlock(int(planes), out_planes)

    def forward(self, x):
        out = self.layers(x)
        if self.is_DownBlock:
            out = self.downblock(out)
        return out


class Discriminator(nn.Module):
    """reference :192-231.  (3,64,64) -> (b,1) probabilities."""

    def __init__(self, net_size=1):
        super(Discriminator, self).__init__()
        self.stem = nn.Sequential(
            L.Conv2d(3, 16, kernel_size=3, stride=1, padding=1, bias=False),
            L.LeakyReLU(0.2, inplace=True),
            L.Conv2d(16, 36, kernel_size=3, stride=1, padding=1, bias=False),
            *_bn_act(36),
        )
        self.main = nn.Sequential(Block(36, 40, 3), Block(40, 48, 3), Block(48, 64, 3), Block(64, 80, 3))  # 64 -> 4 px
        self.layer5 = nn.Sequential(*_bn_act(80), L.Conv2d(80, 128, kernel_size=4, stride=1, padding=0, bias=False),
                                    *_bn_act(128))
        self.linear = L.Linear(128, 1)

    def forward(self, x):
        out = self.layer5(self.main(self.stem(x)))
        return ops.sigmoid(self.linear(out.reshape(out.size(0), -1)))
