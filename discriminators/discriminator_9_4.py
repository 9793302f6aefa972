"""discriminator_9_4 — drop-in for the reference's final discriminator, on libganb200 kernels.

Same class names, constructor signatures, attribute names and parameter registration order as
/root/reference/discriminators/discriminator_9_4.py (so state_dict keys, optimiser parameter order and the
RNG stream consumed at construction are identical); every forward() runs on gan_b200.ops (hand-written sm_100a
kernels, twice differentiable for the R1/R2/WGAN-GP penalties) instead of ATen.

Shape trace at 64x64 (reference :163-199): stem 3->64 @64^2, 15 residual blocks
64@64^2 x2 -> 128@32^2 x3 -> 256@16^2 x3 -> 512@8^2 x3 -> 1024@4^2 -> +stddev channel -> 1025@4^2 x2 -> 1025@2^2,
flatten 4100 -> 4100 -> 1 raw logit (no sigmoid).
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
    """reference :20-27 — the bias is N(0,1) * `bias` (0 by default)."""
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


class MiniBatchStdDev(nn.Module):
    """reference :42-54 — the whole batch tensor viewed as (group_size, -1)."""

    def __init__(self, group_size: int = 4):
        super().__init__()
        self.group_size = group_size

    def forward(self, x: torch.Tensor):
        assert x.shape[0] % self.group_size == 0
        return ops.minibatch_stddev(x, self.group_size, 1e-8)


class Smooth(L.SmoothBase):
    """reference :56-72"""


class DownSample(nn.Module):
    """reference :74-81 — Smooth then bicubic x0.5; both are separable linear maps, applied as one stencil."""

    def __init__(self):
        super().__init__()
        self.smooth = Smooth()

    def forward(self, x: torch.Tensor):
        return ops.smooth_bicubic_down2(x)


def _fc_stack(planes, n):
    mods = []
    for _ in range(n):
        mods += [EqualizedLinear(planes, planes), L.PReLU(planes)]
    return nn.Sequential(*mods)


class SEBlock_conv(nn.Module):
    """reference :83-109 — pool to 5x5, two un-padded 3x3 convs (5->3->1), fc, fc, sigmoid gate."""

    def __init__(self, in_planes: int):
        super(SEBlock_conv, self).__init__()
        self.gap_conv = L.AdaptiveAvgPool2d(5)
        mods = []
        for _ in range(2):
            mods += [EqualizedConv2d(in_planes, in_planes, 3), L.PReLU(in_planes)]
        self.convs = nn.Sequential(*mods)
        self.gap_fc = L.AdaptiveAvgPool2d(1)
        self.fcs = _fc_stack(in_planes, 1)
        self.fc_out = EqualizedLinear(in_planes, in_planes)
        self.activation2 = L.Sigmoid()

    def forward(self, x: torch.Tensor):
        b, c, _, side = x.shape
        assert (side >= 8)
        t = self.convs(self.gap_conv(x))
        t = self.gap_fc(t).view(b, c)
        t = self.fc_out(self.fcs(t))
        return self.activation2(t).view(b, c, 1, 1)


class SEBlock_fc(nn.Module):
    """reference :111-128 — global pool, two fc+PReLU, fc, sigmoid gate."""

    def __init__(self, in_planes: int):
        super(SEBlock_fc, self).__init__()
        self.gap = L.AdaptiveAvgPool2d(1)
        self.fcs = _fc_stack(in_planes, 2)
        self.fc_out = EqualizedLinear(in_planes, in_planes)
        self.activation2 = L.Sigmoid()

    def forward(self, x: torch.Tensor):
        b, c = x.shape[:2]
        t = self.gap(x).view(b, c)
        t = self.fc_out(self.fcs(t))
        return self.activation2(t).view(b, c, 1, 1)


class DiscriminatorBlock(nn.Module):
    """reference :131-161 — conv,PReLU,conv,PReLU [,Smooth,stride-2 conv,PReLU] -> x*SE(x) + residual."""

    def __init__(self, in_features, out_features, downsample, image_size):
        super().__init__()
        self.residual = nn.Sequential()
        self.block = nn.Sequential(
            EqualizedConv2d(in_features, in_features, kernel_size=3, padding=1),
            L.PReLU(in_features),
            EqualizedConv2d(in_features, out_features, kernel_size=3, padding=1),
            L.PReLU(out_features),
        )
        self.se = SEBlock_conv(out_features) if image_size > 4 else SEBlock_fc(out_features)
        self.down_sample = nn.Sequential()
        if downsample:
            self.residual = nn.Sequential(DownSample(), EqualizedConv2d(in_features, out_features, kernel_size=1))
            self.down_sample = nn.Sequential(
                Smooth(),
                EqualizedConv2d(out_features, out_features, kernel_size=3, padding=1, stride=2),
                L.PReLU(out_features),
            )

    def forward(self, x):
        skip = self.residual(x)
        y = self.down_sample(self.block(x))
        return L.gated_residual(skip, y, self.se(y))


class Discriminator(nn.Module):
    """reference :163-199"""

    def __init__(self):
        super(Discriminator, self).__init__()
        f = 64
        top = 16 * f + 1
        # (in, out, downsample, image_size) per block; MiniBatchStdDev sits after the 4x4 down block
        plan = [(f, f, False, 64), (f, f, False, 64),
                (f, 2 * f, True, 32), (2 * f, 2 * f, False, 32), (2 * f, 2 * f, False, 32),
                (2 * f, 4 * f, True, 16), (4 * f, 4 * f, False, 16), (4 * f, 4 * f, False, 16),
                (4 * f, 8 * f, True, 8), (8 * f, 8 * f, False, 8), (8 * f, 8 * f, False, 8),
                (8 * f, 16 * f, True, 4),
                None,
                (top, top, False, 4), (top, top, False, 4), (top, top, True, 2)]
        mods = [EqualizedConv2d(3, f, 3, 1), L.PReLU(f)]
        for spec in plan:
            mods.append(MiniBatchStdDev() if spec is None else DiscriminatorBlock(*spec))
        self.conv = nn.Sequential(*mods)
        flat = 2 * 2 * top
        self.fc = nn.Sequential(EqualizedLinear(flat, flat), L.PReLU(flat), EqualizedLinear(flat, 1))

    def forward(self, input):
        out = self.conv(input)
        return self.fc(out.view(out.shape[0], -1))
