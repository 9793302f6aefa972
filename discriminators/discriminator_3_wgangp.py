"""discriminator_3_wgangp — drop-in for the reference's discriminators/discriminator_3_wgangp.py, the critic that the
reference's main.py (`get_discriminator.get_3_wgan(ngpu, device, ndf=128, nc=3)`, main.py:44) builds: a DCGAN-shaped
stack of bias-free 4x4 stride-2 convolutions with scalar PReLUs, (nc,64,64) -> one score per sample.

Same module tree as the reference (`main.<2i>.weight` convs, `main.<2i+1>.weight` PReLU slopes; reference :15-37), so
state_dicts interchange and a seeded construction draws the same initial weights.  Every layer runs on libganb200
(zero-padded stride-2 convolution on the tcgen05 engine, its dgrad by output phase, PReLU and their double-backward
terms), so the R1 / R2 / WGAN-GP penalties of train/wgangpR2.py differentiate it twice through our kernels."""
import torch.nn as nn

from gan_b200 import layers as L
from gan_b200 import ops


class StandardDeviation(nn.Module):
    """reference :4-13 — defined there, never used by this critic; kept so that the class path exists."""

    def forward(self, in_features):
        return ops.batch_stddev(in_features, 10e-8)


class Discriminator(nn.Module):
    """reference :15-42."""

    def __init__(self, ngpu, ndf, nc):
        super().__init__()
        self.ngpu = ngpu
        widths = [nc, ndf, ndf * 2, ndf * 4, ndf * 8]
        seq = []
        for cin, cout in zip(widths[:-1], widths[1:]):          # 64 -> 32 -> 16 -> 8 -> 4
            seq += [L.Conv2d(cin, cout, 4, 2, 1, bias=False), L.PReLU()]
        seq.append(L.Conv2d(widths[-1], 1, 4, 1, 0, bias=False))  # 4x4 valid conv -> 1x1
        self.main = nn.Sequential(*seq)

    def forward(self, input):
        out = self.main(input)
        return out.reshape(out.shape[0], -1)
