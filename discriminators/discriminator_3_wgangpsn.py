"""discriminator_3_wgangpsn — drop-in for the reference's discriminators/discriminator_3_wgangpsn.py (the critic behind
`get_discriminator.get_3_wgansn`): the DCGAN stack of bias-free 4x4 stride-2 convolutions with every convolution
spectrally normalised and LeakyReLU(0.2) between them; (nc,64,64) -> one score per sample.  Same module tree
(`main.<2i>.parametrizations.weight.original`, `...weight.0._u/_v`; reference :12-29) and the same RNG consumption at
construction.  Spectral norm = gan_b200/spectral.py (power-iteration and W/sigma kernels), convolutions = the conv engine."""
import torch.nn as nn

from gan_b200 import layers as L
from gan_b200.spectral import spectral_norm


class Discriminator(nn.Module):
    """reference discriminator_3_wgangpsn.py:8-35."""

    def __init__(self, ngpu, ndf, nc):
        super().__init__()
        self.ngpu = ngpu
        widths = [nc, ndf, ndf * 2, ndf * 4, ndf * 8]
        seq = []
        for cin, cout in zip(widths[:-1], widths[1:]):
            seq += [spectral_norm(L.Conv2d(cin, cout, 4, 2, 1, bias=False)), L.LeakyReLU(0.2, inplace=True)]
        seq.append(spectral_norm(L.Conv2d(widths[-1], 1, 4, 1, 0, bias=False)))
        self.main = nn.Sequential(*seq)

    def forward(self, input):
        out = self.main(input)
        return out.reshape(out.shape[0], -1)
