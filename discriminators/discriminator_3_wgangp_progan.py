"""discriminator_3_wgangp_progan — drop-in for the reference's discriminators/discriminator_3_wgangp_progan.py (BASELINE
config 4, the critic of the `wgangp` loop): a ProGAN-style critic of equalised convs + scalar PReLUs with the
minibatch-stddev channel before the last 3x3 conv; (nc,64,64) -> one score per sample.

Same module tree and registration order as the reference (`main.<i>.bias`, `main.<i>.conv.weight`, `main.<i>.weight`
for the PReLUs), so state_dicts interchange and a seeded construction draws the same initial weights.  The compute is
libganb200's: `conv(x*scale)+bias` is ONE conv-engine call (the equalised gain is the kernel's alpha, the bias its
epilogue), the stddev channel is the mbstd kernel with G = batch, and all of it is closed under double backward (the
WGAN-GP penalty differentiates the critic twice, Train.py:209-228)."""
import numpy
import torch.nn as nn

from gan_b200 import layers as L
from gan_b200 import ops


class StandardDeviation(nn.Module):
    """reference :7-16 — per-position biased std over the batch (+10e-8 inside the sqrt), mean of that, one extra
    constant channel."""

    def forward(self, in_features):
        return ops.batch_stddev(in_features, 10e-8)


class EqualizedConv2d(nn.Module):
    """reference :18-29.  The conv's own bias Parameter is re-registered on this module and removed from the conv
    (so the key is `<name>.bias`, and it precedes `<name>.conv.weight` in parameters())."""

    def __init__(self, in_planes, out_planes, kernel_size, stride=1, padding=0, groups=1):
        super().__init__()
        self.conv = L.Conv2d(in_planes, out_planes, kernel_size, stride, padding, groups=groups)
        self.scale = numpy.sqrt(2) / numpy.sqrt(kernel_size * kernel_size * in_planes)
        self.bias = self.conv.bias
        self.conv.bias = None
        nn.init.normal_(self.conv.weight)
        nn.init.normal_(self.bias)

    def forward(self, x):
        c = self.conv
        return ops.conv2d(x, c.weight, self.bias, c.stride[0], c.padding[0], 0, alpha=float(self.scale))


class Discriminator(nn.Module):
    """reference :31-71."""

    def __init__(self, ngpu, ndf, nc):
        super().__init__()
        self.ngpu = ngpu
        seq = [EqualizedConv2d(nc, ndf, 1, 1, 0), L.PReLU()]
        width = ndf
        for mult in (1, 2, 4, 8):  # 64 -> 32 -> 16 -> 8 -> 4: widen at full resolution, then a stride-2 conv
            seq += [EqualizedConv2d(width, ndf * mult, 3, 1, 1), L.PReLU(),
                    EqualizedConv2d(ndf * mult, ndf * mult, 3, 2, 1), L.PReLU()]
            width = ndf * mult
        seq += [StandardDeviation(),
                EqualizedConv2d(width + 1, width, 3, 1, 1), L.PReLU(),
                EqualizedConv2d(width, width, 4, 1, 0), L.PReLU(),
                EqualizedConv2d(width, 1, 1, 1, 0)]
        self.main = nn.Sequential(*seq)

    def forward(self, input):
        out = self.main(input)
        return out.reshape(out.shape[0], -1)
