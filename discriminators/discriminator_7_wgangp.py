"""discriminator_7_wgangp — drop-in for the reference's discriminators/discriminator_7_wgangp.py (the critic behind
`get_discriminator.get_7_wgan`): the ShuffleNetV2 critic WITHOUT spectral norm — ratio 1 by default, stages of 3/7/3/3
stride-1 units after each down-sampling unit, LeakyReLU with torch's default slope 0.01 (the *sn variant uses ratio 2,
4/7/4/4 and slope 0.1).  Same module tree (`pre.0`, `stage2.0.residual.0`, ..., `conv5.0`), same construction order.
1x1 convs on the conv engine, depthwise 3x3 on csrc/dwconv.cu, channel split / concat / shuffle on the channel kernels;
closed under the double backward of the gradient penalty."""
import torch.nn as nn

from gan_b200 import layers as L
from gan_b200 import ops


def channel_split(x, split):
    """reference :16-23."""
    assert x.size(1) == split * 2
    return ops.slice_channels(x, 0, split), ops.slice_channels(x, split, 2 * split)


def channel_shuffle(x, groups):
    """reference :26-40."""
    return ops.channel_shuffle(x, groups)


def _pw(cin, cout):
    return L.Conv2d(cin, cout, 1)


def _dw(c, stride):
    return L.Conv2d(c, c, 3, stride=stride, padding=1, groups=c)


class ShuffleUnit(nn.Module):
    """reference :43-91."""

    def __init__(self, in_channels, out_channels, stride):
        super().__init__()
        self.stride = stride
        self.in_channels = in_channels
        self.out_channels = out_channels
        if stride != 1 or in_channels != out_channels:
            half = int(out_channels / 2)
            self.residual = nn.Sequential(_pw(in_channels, in_channels), L.LeakyReLU(inplace=True),
                                          _dw(in_channels, stride), _pw(in_channels, half), L.LeakyReLU(inplace=True))
            self.shortcut = nn.Sequential(_dw(in_channels, stride), _pw(in_channels, half), L.LeakyReLU(inplace=True))
        else:
            self.shortcut = nn.Sequential()
            half = int(in_channels / 2)
            self.residual = nn.Sequential(_pw(half, half), L.LeakyReLU(inplace=True), _dw(half, stride),
                                          _pw(half, half), L.LeakyReLU(inplace=True))

    def forward(self, x):
        if self.stride == 1 and self.out_channels == self.in_channels:
            shortcut, residual = channel_split(x, int(self.in_channels / 2))
        else:
            shortcut = residual = x
        merged = ops.cat_channels([self.shortcut(shortcut), self.residual(residual)])
        return channel_shuffle(merged, 2)


class Discriminator(nn.Module):
    """reference :94-140."""

    WIDTHS = {0.5: [48, 96, 192, 1024], 1: [116, 232, 464, 928], 1.5: [176, 352, 704, 1024], 2: [244, 488, 976, 2048]}

    def __init__(self, ratio=1):
        super().__init__()
        out_channels = self.WIDTHS[ratio]  # unknown ratio: KeyError (the reference: UnboundLocalError)
        self.pre = nn.Sequential(L.Conv2d(3, 32, 3, stride=1, padding=1))
        self.stage2 = self._make_stage(32, out_channels[0], 3)
        self.stage3 = self._make_stage(out_channels[0], out_channels[1], 7)
        self.stage4 = self._make_stage(out_channels[1], out_channels[2], 3)
        self.stage5 = self._make_stage(out_channels[2], out_channels[3], 3)
        self.conv5 = nn.Sequential(L.Conv2d(out_channels[3], 1, 4, stride=1, padding=0))

    def forward(self, x):
        for stage in (self.pre, self.stage2, self.stage3, self.stage4, self.stage5, self.conv5):
            x = stage(x)
        return x.reshape(x.size(0), -1)

    def _make_stage(self, in_channels, out_channels, repeat):
        units = [ShuffleUnit(in_channels, out_channels, 2)]
        units += [ShuffleUnit(out_channels, out_channels, 1) for _ in range(repeat)]
        return nn.Sequential(*units)
