"""discriminator_2 — drop-in for the reference's discriminators/discriminator_2.py (the DCGAN discriminator behind
`get_discriminator.get_2`, which also applies `weights_init`): bias-free 4x4 stride-2 convolutions
(nc -> ndf -> 2ndf -> 4ndf -> 8ndf, 64 -> 4 pixels), BatchNorm2d after the 2nd-4th, LeakyReLU(0.2), a 4x4 valid conv to
one logit and a Sigmoid (the BCE loops).  Same flat module tree (`main.0` ... `main.11`, `active`; reference :10-29)."""
import torch.nn as nn

from gan_b200 import layers as L


class Discriminator(nn.Module):
    """reference discriminator_2.py:6-36."""

    def __init__(self, ngpu, ndf, nc):
        super().__init__()
        self.ngpu = ngpu
        seq = [L.Conv2d(nc, ndf, 4, 2, 1, bias=False), L.LeakyReLU(0.2, inplace=True)]
        for m in (1, 2, 4):
            seq += [L.Conv2d(ndf * m, ndf * m * 2, 4, 2, 1, bias=False), L.BatchNorm2d(ndf * m * 2),
                    L.LeakyReLU(0.2, inplace=True)]
        seq.append(L.Conv2d(ndf * 8, 1, 4, 1, 0, bias=False))
        self.main = nn.Sequential(*seq)
        self.active = L.Sigmoid()

    def forward(self, input):
        out = self.main(input)
        return self.active(out.reshape(out.shape[0], -1))
