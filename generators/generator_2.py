"""generator_2 — drop-in for the reference's generators/generator_2.py (the DCGAN generator behind
`get_generators.get_2`, which also applies `weights_init`): five bias-free 4x4 transposed convolutions
(nz -> 8ngf@4x4 -> 4ngf@8x8 -> 2ngf@16x16 -> ngf@32x32 -> nc@64x64), BatchNorm2d + LeakyReLU(0.2) after the first four,
Tanh at the end.  Same flat module tree (`main.0` ... `main.13`, reference :11-40), same construction order, so
state_dicts interchange and a seeded construction draws the same weights.  Every layer on libganb200: the transposed
conv is the conv engine's dgrad, BatchNorm the batch-statistics kernels, LeakyReLU / Tanh the pointwise kernels."""
import torch.nn as nn

from gan_b200 import layers as L


class Generator(nn.Module):
    """reference generator_2.py:8-43."""

    def __init__(self, ngpu, nz, ngf, nc):
        super().__init__()
        self.ngpu = ngpu
        widths = [nz, ngf * 8, ngf * 4, ngf * 2, ngf]
        seq = []
        for i, (cin, cout) in enumerate(zip(widths[:-1], widths[1:])):
            stride, pad = (1, 0) if i == 0 else (2, 1)
            seq += [L.ConvTranspose2d(cin, cout, 4, stride, pad, bias=False), L.BatchNorm2d(cout),
                    L.LeakyReLU(0.2, inplace=True)]
        seq += [L.ConvTranspose2d(ngf, nc, 4, 2, 1, bias=False), L.Tanh()]
        self.main = nn.Sequential(*seq)

    def forward(self, input):
        return self.main(input)
