"""discriminator_5_wgangpsn — drop-in for the reference's discriminators/discriminator_5_wgangpsn.py (the critic behind
`get_discriminator.get_5_wgansn`): the ResNet-in-ResNet critic with EVERY convolution spectrally normalised,
LeakyReLU(0.1), three layers per block, head = one spectrally normalised 4x4 valid conv 512 -> 1.
Blocks: gan_b200/rir.py; spectral norm: gan_b200/spectral.py."""
import torch.nn as nn

from gan_b200 import layers as L
from gan_b200 import rir
from gan_b200.spectral import spectral_norm


class ResnetInit(rir.ResnetInitBase):
    """reference :6-40."""
    SLOPE = 0.1
    wrap = staticmethod(spectral_norm)


class RiRBlock(rir.RiRBlockBase):
    """reference :43-60."""
    LAYER = ResnetInit


class Discriminator(rir.RiRDiscriminatorBase):
    """reference :63-104."""
    BLOCK = RiRBlock
    LAYERS_PER_BLOCK = 3

    def make_head(self):
        return nn.Sequential(spectral_norm(L.Conv2d(512, 1, kernel_size=4, stride=1, padding=0)))
