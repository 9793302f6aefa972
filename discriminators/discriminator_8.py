"""discriminator_8 — drop-in for the reference's discriminators/discriminator_8.py (the StyleGAN2 critic behind
`get_discriminator.get_8(ngpu, device)` = `Discriminator(int(log2(64)))`): 1x1 from-RGB conv, residual blocks of two
zero-padded 3x3 equalised convs + LeakyReLU(0.2) and a Smooth + bicubic x0.5 down-sample on both paths, summed and
scaled by 1/sqrt(2); minibatch-stddev channel, an unpadded 3x3 conv (4x4 -> 2x2), one equalised linear to a raw score.

Same class names, constructor signatures, attribute names and registration order as the reference (state_dict keys:
`from_rgb.0.weight.weight`, `blocks.<i>.residual.1.bias`, ...).  Differences from discriminator_9_4's helpers that are
kept: conv padding is ZERO padding (F.conv2d's `padding`, reference :103) not replicate, conv biases start at 1 and
linear biases at the `bias` constant (no random draw), no stride argument.  All arithmetic on libganb200: the equalised
gain is the conv / linear kernel's alpha, Smooth + bicubic is one separable stencil, everything twice differentiable."""
import math
from typing import List

import torch
from torch import nn

from gan_b200 import layers as L
from gan_b200 import ops


class Smooth(L.SmoothBase):
    """reference :10-43"""


class EqualizedWeight(L.EqWeightBase):
    """reference :107-141 — parameter `weight` ~ N(0,1), constant `c` = 1/sqrt(fan_in)."""
    PARAM, GAIN = "weight", "c"

    def __init__(self, shape: List[int]):
        super().__init__(shape)


class EqualizedLinear(nn.Module):
    """reference :50-74 — linear(x, W*c, bias); bias = ones * `bias`."""

    def __init__(self, in_features: int, out_features: int, bias: float = 0.):
        super().__init__()
        self.weight = EqualizedWeight([out_features, in_features])
        self.bias = nn.Parameter(torch.ones(out_features) * bias)

    def forward(self, x: torch.Tensor):
        return ops.linear(x, self.weight.raw, self.bias, alpha=self.weight.gain)


class EqualizedConv2d(nn.Module):
    """reference :77-104 — conv2d(x, W*c, bias, padding=padding) with zero padding; bias = ones."""

    def __init__(self, in_features: int, out_features: int, kernel_size: int, padding: int = 0):
        super().__init__()
        self.padding = padding
        self.weight = EqualizedWeight([out_features, in_features, kernel_size, kernel_size])
        self.bias = nn.Parameter(torch.ones(out_features))

    def forward(self, x: torch.Tensor):
        return ops.conv2d(x, self.weight.raw, self.bias, 1, self.padding, 0, alpha=self.weight.gain)


class DownSample(nn.Module):
    """reference :236-256 — Smooth, then bicubic to half the size (align_corners=False)."""

    def __init__(self):
        super().__init__()
        self.smooth = Smooth()

    def forward(self, x: torch.Tensor):
        return ops.smooth_bicubic_down2(x)


class DiscriminatorBlock(nn.Module):
    """reference :144-189 — (down(block(x)) + conv1x1(down(x))) / sqrt(2)."""

    def __init__(self, in_features, out_features):
        super().__init__()
        self.residual = nn.Sequential(DownSample(), EqualizedConv2d(in_features, out_features, kernel_size=1))
        self.block = nn.Sequential(
            EqualizedConv2d(in_features, in_features, kernel_size=3, padding=1),
            L.LeakyReLU(0.2, True),
            EqualizedConv2d(in_features, out_features, kernel_size=3, padding=1),
            L.LeakyReLU(0.2, True),
        )
        self.down_sample = DownSample()
        self.scale = 1 / math.sqrt(2)

    def forward(self, x):
        residual = self.residual(x)
        x = self.down_sample(self.block(x))
        return ops.affine(ops.add(x, residual), self.scale, 0.0)


class MiniBatchStdDev(nn.Module):
    """reference :192-233 — the whole batch tensor viewed as (group_size, -1), unbiased variance, eps inside the sqrt."""

    def __init__(self, group_size: int = 4):
        super().__init__()
        self.group_size = group_size

    def forward(self, x: torch.Tensor):
        assert x.shape[0] % self.group_size == 0
        return ops.minibatch_stddev(x, self.group_size, 1e-8)


class Discriminator(nn.Module):
    """reference :259-320."""

    def __init__(self, log_resolution: int, n_features: int = 64, max_features: int = 512):
        super().__init__()
        self.from_rgb = nn.Sequential(EqualizedConv2d(3, n_features, 1), L.LeakyReLU(0.2, True))
        features = [min(max_features, n_features * (2 ** i)) for i in range(log_resolution - 1)]
        n_blocks = len(features) - 1
        self.blocks = nn.Sequential(*[DiscriminatorBlock(features[i], features[i + 1]) for i in range(n_blocks)])
        self.std_dev = MiniBatchStdDev()
        final_features = features[-1] + 1
        self.conv = EqualizedConv2d(final_features, final_features, 3)
        self.final = EqualizedLinear(2 * 2 * final_features, 1)

    def forward(self, x: torch.Tensor):
        x = self.conv(self.std_dev(self.blocks(self.from_rgb(x))))
        return self.final(x.reshape(x.shape[0], -1))
