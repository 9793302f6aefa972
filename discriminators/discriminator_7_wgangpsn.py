"""discriminator_7_wgangpsn — drop-in for the reference's discriminators/discriminator_7_wgangpsn.py (BASELINE config 5,
the critic of the `wgangpsnR2` loop): a ShuffleNetV2 (x2.0) critic in which every one of the 79 convolutions is
spectrally normalised; LeakyReLU(0.1); (3,64,64) -> one score per sample.

Same module tree as the reference (`pre.0`, `stage2.0.residual.0` ..., `conv5.0`, each a parametrised conv with
`bias`, `parametrizations.weight.original`, `parametrizations.weight.0._u/_v`), same construction order.  On
libganb200: 1x1 convs = the conv engine (tcgen05 where eligible), depthwise 3x3 = csrc/dwconv.cu (HBM-bound),
spectral norm = gan_b200/spectral.py (power iteration + W/sigma kernels), channel split / concat / shuffle = the
channel-copy and shuffle kernels; everything closed under the double backward of R1 and the gradient penalty."""
import torch.nn as nn

from gan_b200 import layers as L
from gan_b200 import ops
from gan_b200.spectral import spectral_norm


def channel_split(x, split):
    """reference :17-24 — the two halves along the channel dimension."""
    assert x.size(1) == split * 2
    return ops.slice_channels(x, 0, split), ops.slice_channels(x, split, 2 * split)


def channel_shuffle(x, groups):
    """reference :27-41."""
    return ops.channel_shuffle(x, groups)


def _pw(cin, cout):
    return spectral_norm(L.Conv2d(cin, cout, 1))


def _dw(c, stride):
    return spectral_norm(L.Conv2d(c, c, 3, stride=stride, padding=1, groups=c))


class ShuffleUnit(nn.Module):
    """reference :44-92.  Down-sampling / widening units run both branches on the whole input; the others split the
    channels, pass one half through and transform the other."""

    def __init__(self, in_channels, out_channels, stride):
        super().__init__()
        self.stride = stride
        self.in_channels = in_channels
        self.out_channels = out_channels
        if stride != 1 or in_channels != out_channels:
            half = int(out_channels / 2)
            self.residual = nn.Sequential(_pw(in_channels, in_channels), L.LeakyReLU(0.1, inplace=True),
                                          _dw(in_channels, stride), _pw(in_channels, half),
                                          L.LeakyReLU(0.1, inplace=True))
            self.shortcut = nn.Sequential(_dw(in_channels, stride), _pw(in_channels, half),
                                          L.LeakyReLU(0.1, inplace=True))
        else:
            self.shortcut = nn.Sequential()
            half = int(in_channels / 2)
            self.residual = nn.Sequential(_pw(half, half), L.LeakyReLU(0.1, inplace=True), _dw(half, stride),
                                          _pw(half, half), L.LeakyReLU(0.1, inplace=True))

    def forward(self, x):
        if self.stride == 1 and self.out_channels == self.in_channels:
            shortcut, residual = channel_split(x, int(self.in_channels / 2))
        else:
            shortcut = residual = x
        merged = ops.cat_channels([self.shortcut(shortcut), self.residual(residual)])
        return channel_shuffle(merged, 2)


class Discriminator(nn.Module):
    """reference :95-143."""

    WIDTHS = {0.5: [48, 96, 192, 1024], 1: [116, 232, 464, 928], 1.5: [176, 352, 704, 1024],
              2: [244, 488, 976, 1952], 3: [344, 688, 1376, 2752]}

    def __init__(self, ratio=2):
        super().__init__()
        out_channels = self.WIDTHS[ratio]  # the reference builds (and never raises) a ValueError and then fails
        # with UnboundLocalError on an unknown ratio; here it is a KeyError
        self.pre = nn.Sequential(spectral_norm(L.Conv2d(3, 32, 3, stride=1, padding=1)))
        self.stage2 = self._make_stage(32, out_channels[0], 4)
        self.stage3 = self._make_stage(out_channels[0], out_channels[1], 7)
        self.stage4 = self._make_stage(out_channels[1], out_channels[2], 4)
        self.stage5 = self._make_stage(out_channels[2], out_channels[3], 4)
        self.conv5 = nn.Sequential(spectral_norm(L.Conv2d(out_channels[3], 1, 4, stride=1, padding=0)))

    def forward(self, x):
        for stage in (self.pre, self.stage2, self.stage3, self.stage4, self.stage5, self.conv5):
            x = stage(x)
        return x.reshape(x.size(0), -1)

    def _make_stage(self, in_channels, out_channels, repeat):
        units = [ShuffleUnit(in_channels, out_channels, 2)]
        units += [ShuffleUnit(out_channels, out_channels, 1) for _ in range(repeat)]
        return nn.Sequential(*units)
