"""discriminator_9_small — drop-in for the reference's discriminators/discriminator_9_small.py, on libganb200 kernels: the
critic whose residual blocks merge the skip path and the convolution path with SELECTIVE-KERNEL attention
(`(feas * attention).sum(dim=1)`, reference :196-201) instead of a squeeze-excite gate, one down-sampling block per
resolution.  Same class names, constructor signatures, attribute names and parameter registration order as the reference.

Under the R1 / R2 / WGAN-GP penalties the attention — pooled sum, three padded 3x3 convs on the 5x5 map, fully connected
trunk, two heads, softmax across the heads — is differentiated twice; every op below has that second derivative
(ops.SkMerge2Fn and ops.PoolSum2Fn fall back to their generic twice-differentiable forms under create_graph)."""
from typing import List

import torch
from torch import nn

from gan_b200 import layers as L
from gan_b200 import ops


class EqualizedWeight(L.EqWeightBase):
    """reference :9-17"""
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
    """reference :30-40 — `pad` child module as in the reference; the padding itself is an index clamp in the kernel."""
    WEIGHT = EqualizedWeight

    def __init__(self, in_features: int, out_features: int, kernel_size: int, padding: int = 0, stride: int = 1):
        super().__init__()
        self._build(in_features, out_features, kernel_size, padding, stride)


class MiniBatchStdDev(nn.Module):
    """reference :42-54"""

    def __init__(self, group_size: int = 4):
        super().__init__()
        self.group_size = group_size

    def forward(self, x: torch.Tensor):
        assert x.shape[0] % self.group_size == 0
        return ops.minibatch_stddev(x, self.group_size, 1e-8)


class Smooth(L.SmoothBase):
    """reference :56-72"""


class DownSample(nn.Module):
    """reference :74-81"""

    def __init__(self):
        super().__init__()
        self.smooth = Smooth()

    def forward(self, x: torch.Tensor):
        return ops.smooth_bicubic_down2(x)


def _fc_prelu(planes, n):
    mods = []
    for _ in range(n):
        mods += [EqualizedLinear(planes, planes), L.PReLU(planes)]
    return mods


class _SKHeads(nn.Module):
    """the two attention classes share their heads: M x (fc, PReLU, fc) on the trunk vector, softmax across the heads"""

    def _build_heads(self, planes, m):
        self.M = m
        for i in range(m):
            self.__setattr__('fc_sub_%d' % i, nn.Sequential(*_fc_prelu(planes, 1), EqualizedLinear(planes, planes)))
        self.softmax = nn.Softmax(dim=1)

    def _heads(self, fea_z, b, s, c):
        subs = [self.__getattr__('fc_sub_%d' % i) for i in range(self.M)]
        return L.sk_attention_vectors(fea_z, subs).view(b, s, c, 1, 1)


class SKAttention_conv(_SKHeads):
    """reference :83-128 — sum of the branches, 5x5 pooling, three padded 3x3 convs + PReLU, global pooling, two fc."""

    def __init__(self, planes: int, m: int):
        super(SKAttention_conv, self).__init__()
        self.gap_conv = L.AdaptiveAvgPool2d(5)
        mods = []
        for _ in range(3):
            mods += [EqualizedConv2d(planes, planes, 3, 1), L.PReLU(planes)]
        self.conv_main = nn.Sequential(*mods)
        self.gap_fc = L.AdaptiveAvgPool2d(1)
        self.fc_main = nn.Sequential(*_fc_prelu(planes, 2))
        self._build_heads(planes, m)

    def forward(self, feas):
        branches = L.as_branches(feas)
        b, c, _, side = branches[0].shape
        assert (side >= 8)
        fea_s = self.conv_main(ops.pooled_branch_sum(branches, 5))
        fea_z = self.fc_main(self.gap_fc(fea_s).view(b, c))
        return self._heads(fea_z, b, len(branches), c)


class SKAttention_fc(_SKHeads):
    """reference :131-168 — sum of the branches, global pooling, four fc + PReLU."""

    def __init__(self, planes: int, m: int):
        super(SKAttention_fc, self).__init__()
        self.gap_fc = L.AdaptiveAvgPool2d(1)
        self.fc_main = nn.Sequential(*_fc_prelu(planes, 4))
        self._build_heads(planes, m)

    def forward(self, feas):
        branches = L.as_branches(feas)
        b, c = branches[0].shape[:2]
        fea_z = self.fc_main(self.gap_fc(L.branch_sum(branches)).view(b, c))
        return self._heads(fea_z, b, len(branches), c)


class DiscriminatorBlock(nn.Module):
    """reference :171-201 — the skip path and the conv path are the two branches of a selective-kernel merge."""

    def __init__(self, in_features, out_features, downsample, image_size):
        super().__init__()
        self.residual = nn.Sequential()
        self.block = nn.Sequential(
            EqualizedConv2d(in_features, in_features, kernel_size=3, padding=1),
            L.LeakyReLU(0.2, True),
            EqualizedConv2d(in_features, out_features, kernel_size=3, padding=1),
            L.LeakyReLU(0.2, True),
        )
        self.sk_attention = SKAttention_conv(out_features, 2) if image_size > 4 else SKAttention_fc(out_features, 2)
        self.down_sample = nn.Sequential()
        if downsample:
            self.residual = nn.Sequential(DownSample(), EqualizedConv2d(in_features, out_features, kernel_size=1))
            self.down_sample = nn.Sequential(
                Smooth(),
                EqualizedConv2d(out_features, out_features, kernel_size=3, padding=1, stride=2),
                L.LeakyReLU(0.2, True),
            )

    def forward(self, x):
        branches = [self.residual(x), self.down_sample(self.block(x))]
        return L.sk_merge(branches, self.sk_attention(branches))


class Discriminator(nn.Module):
    """reference :203-239 — 32 base features, one down-sampling block per resolution (64 -> 2), stddev channel at 4x4."""

    def __init__(self):
        super(Discriminator, self).__init__()
        f = 32
        top = 16 * f + 1
        mods = [EqualizedConv2d(3, f, 3, 1), L.LeakyReLU(0.2, True)]
        for cin, cout, size in ((f, 2 * f, 32), (2 * f, 4 * f, 16), (4 * f, 8 * f, 8), (8 * f, 16 * f, 4)):
            mods.append(DiscriminatorBlock(cin, cout, True, size))
        mods += [MiniBatchStdDev(), DiscriminatorBlock(top, top, True, 2)]
        self.conv = nn.Sequential(*mods)
        self.fc = nn.Sequential(EqualizedLinear(2 * 2 * top, 2 * 2 * top), L.LeakyReLU(0.2, True),
                                EqualizedLinear(2 * 2 * top, 1))

    def forward(self, input):
        out = self.conv(input)
        return self.fc(out.view(out.shape[0], -1))
