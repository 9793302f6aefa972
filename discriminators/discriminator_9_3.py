"""discriminator_9_3 — drop-in for the reference's self-attention critic (discriminators/discriminator_9_3.py), on
libganb200 kernels.  discriminator_9 and discriminator_9_2 are the same building blocks at other widths and with more
attention layers; they import the classes below and only state their own plan.

Class names, constructor signatures, attribute names and parameter registration order follow the reference file, so
state_dict keys, optimiser parameter order and the RNG stream consumed at construction are the same.  What differs from
discriminator_9_4: LeakyReLU(0.2) instead of PReLU, the squeeze-excite gate is two 1x1 convolutions on the globally
pooled map (reference :42-56), and SAGAN self-attention layers (reference :58-80) sit between the residual blocks — a
batched matrix product and a softmax, both twice differentiable here (gan_b200.ops.BmmFn / SoftmaxRowsFn) because the
R1 / R2 / WGAN-GP penalties differentiate the critic twice.
"""
from typing import List

import torch
from torch import nn

from gan_b200 import layers as L
from gan_b200 import ops


class EqualizedWeight(L.EqWeightBase):
    """reference :9-17 — parameter `weight`, constant `c`."""
    PARAM, GAIN = "weight", "c"

    def __init__(self, shape: List[int]):
        super().__init__(shape)


class EqualizedLinear(L.EqLinearBase):
    """reference :20-27"""
    WEIGHT = EqualizedWeight

    def __init__(self, in_planes: int, out_planes: int, bias: float = 0.):
        super().__init__()
        self._build(in_planes, out_planes, bias_gain=bias)


class EqualizedConv2d(L.EqConv2dBase):
    """reference :30-40 — explicit replicate padding amount and stride."""
    WEIGHT = EqualizedWeight

    def __init__(self, in_features: int, out_features: int, kernel_size: int, padding: int = 0, stride: int = 1):
        super().__init__()
        self._build(in_features, out_features, kernel_size, padding, stride)


class SEBlock(nn.Module):
    """reference :42-56 — global pool, 1x1 conv, PReLU, 1x1 conv, sigmoid: a (B, C, 1, 1) gate."""

    def __init__(self, in_planes: int, out_planes: int):
        super(SEBlock, self).__init__()
        self.avg_pool = L.AdaptiveAvgPool2d(1)
        self.conv1 = EqualizedConv2d(in_planes, in_planes, kernel_size=1)
        self.activation1 = L.PReLU(in_planes)
        self.conv2 = EqualizedConv2d(in_planes, out_planes, kernel_size=1)
        self.activation2 = L.Sigmoid()

    def forward(self, x: torch.Tensor):
        return self.activation2(self.conv2(self.activation1(self.conv1(self.avg_pool(x)))))


class SelfAttention(nn.Module):
    """reference :58-80 — query / key / value 1x1 convs, attention over all H*W positions, output 1x1 conv, and the
    residual `gamma(o) * o + x` with an SE gate in the place of SAGAN's scalar gamma."""

    def __init__(self, in_planes: int, embedding_channels: int):
        super(SelfAttention, self).__init__()
        self.query = EqualizedConv2d(in_planes, embedding_channels, 1)
        self.key = EqualizedConv2d(in_planes, embedding_channels, 1)
        self.value = EqualizedConv2d(in_planes, embedding_channels, 1)
        self.self_att = EqualizedConv2d(embedding_channels, in_planes, 1)
        self.gamma = SEBlock(in_planes, in_planes)
        self.softmax = nn.Softmax(dim=1)   # kept for the module tree; the attention runs in L.self_attention

    def forward(self, x: torch.Tensor):
        o = self.self_att(L.self_attention(self.query(x), self.key(x), self.value(x)))
        return L.gated_residual(x, o, self.gamma(o))


class MiniBatchStdDev(nn.Module):
    """reference :82-94 — the whole batch tensor viewed as (group_size, -1)."""

    def __init__(self, group_size: int = 4):
        super().__init__()
        self.group_size = group_size

    def forward(self, x: torch.Tensor):
        assert x.shape[0] % self.group_size == 0
        return ops.minibatch_stddev(x, self.group_size, 1e-8)


class Smooth(L.SmoothBase):
    """reference :96-112"""


class DownSample(nn.Module):
    """reference :114-121 — Smooth then bicubic x0.5, applied as one separable stencil."""

    def __init__(self):
        super().__init__()
        self.smooth = Smooth()

    def forward(self, x: torch.Tensor):
        return ops.smooth_bicubic_down2(x)


class DiscriminatorBlock(nn.Module):
    """reference :123-149 — conv, LeakyReLU, conv, LeakyReLU [, Smooth, stride-2 conv, LeakyReLU] -> x * SE(x) + residual."""

    def __init__(self, in_features, out_features, downsample):
        super().__init__()
        self.residual = nn.Sequential()
        self.block = nn.Sequential(
            EqualizedConv2d(in_features, in_features, kernel_size=3, padding=1),
            L.LeakyReLU(0.2, True),
            EqualizedConv2d(in_features, out_features, kernel_size=3, padding=1),
            L.LeakyReLU(0.2, True),
        )
        self.se = SEBlock(out_features, out_features)
        self.down_sample = nn.Sequential()
        if downsample:
            self.residual = nn.Sequential(DownSample(), EqualizedConv2d(in_features, out_features, kernel_size=1))
            self.down_sample = nn.Sequential(
                Smooth(),
                EqualizedConv2d(out_features, out_features, kernel_size=3, padding=1, stride=2),
                L.LeakyReLU(0.2, True),
            )

    def forward(self, x):
        skip = self.residual(x)
        y = self.down_sample(self.block(x))
        return L.gated_residual(skip, y, self.se(y))


# plan entries: ("att", width multiple) | ("std",) | (in multiple, out multiple, downsample); +1 = the stddev channel
FULL_PLAN = [("att", 1), (1, 1, False), (1, 1, False), (1, 2, True),
             ("att", 2), (2, 2, False), (2, 2, False), (2, 4, True),
             ("att", 4), (4, 4, False), (4, 4, False), (4, 8, True),
             ("att", 8), (8, 8, False), (8, 8, False), (8, 16, True),
             ("att", 16), ("std",), ("top", False), ("top", False), ("top", True)]


def build_trunk(features, plan):
    """the `conv` Sequential of the family: stem, then blocks / attention layers / the stddev channel as planned"""
    top = 16 * features + 1
    mods = [EqualizedConv2d(3, features, 3, 1), L.LeakyReLU(0.2, True)]
    for spec in plan:
        if spec[0] == "att":
            mods.append(SelfAttention(spec[1] * features, spec[1] * features))
        elif spec[0] == "std":
            mods.append(MiniBatchStdDev())
        elif spec[0] == "top":
            mods.append(DiscriminatorBlock(top, top, spec[1]))
        else:
            mods.append(DiscriminatorBlock(spec[0] * features, spec[1] * features, spec[2]))
    head = nn.Sequential(EqualizedLinear(2 * 2 * top, 2 * 2 * top), L.LeakyReLU(0.2, True), EqualizedLinear(2 * 2 * top, 1))
    return nn.Sequential(*mods), head


class Discriminator(nn.Module):
    """reference :151-188 — 32 base features, attention after the stem (64x64) and at 16x16, one plain block per level."""
    FEATURES = 32
    PLAN = [("att", 1), (1, 1, False), (1, 2, True),
            (2, 2, False), (2, 4, True),
            ("att", 4), (4, 4, False), (4, 8, True),
            (8, 8, False), (8, 16, True),
            ("std",), ("top", False), ("top", True)]

    def __init__(self):
        super(Discriminator, self).__init__()
        self.conv, self.fc = build_trunk(self.FEATURES, self.PLAN)

    def forward(self, input):
        out = self.conv(input)
        return self.fc(out.view(out.shape[0], -1))
