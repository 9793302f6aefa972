"""discriminator_5_wgangp — drop-in for the reference's discriminators/discriminator_5_wgangp.py (the critic behind
`get_discriminator.get_5_wgan`): the ResNet-in-ResNet critic without normalisation, LeakyReLU(0.2), two layers per
block, head = 4x4 valid conv 512 -> 1024, LeakyReLU, 1x1 conv -> one score.  Blocks: gan_b200/rir.py."""
import torch.nn as nn

from gan_b200 import layers as L
from gan_b200 import rir


class ResnetInit(rir.ResnetInitBase):
    """reference :6-40."""


class RiRBlock(rir.RiRBlockBase):
    """reference :43-60."""
    LAYER = ResnetInit


class Discriminator(rir.RiRDiscriminatorBase):
    """reference :63-106."""
    BLOCK = RiRBlock
    LAYERS_PER_BLOCK = 2

    def make_head(self):
        return nn.Sequential(L.Conv2d(512, 1024, kernel_size=4, stride=1, padding=0), L.LeakyReLU(0.2, inplace=True),
                             L.Conv2d(1024, 1, kernel_size=1, stride=1, padding=0))
