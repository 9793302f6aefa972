"""discriminator_5 — drop-in for the reference's discriminators/discriminator_5.py (the discriminator behind
`get_discriminator.get_5`, for the BCE loops): the ResNet-in-ResNet stack with BatchNorm2d before every LeakyReLU(0.2),
three layers per block, head = 4x4 valid conv 512 -> 1 and a Sigmoid.  Blocks: gan_b200/rir.py."""
import torch.nn as nn

from gan_b200 import layers as L
from gan_b200 import rir


class ResnetInit(rir.ResnetInitBase):
    """reference :6-46."""
    BATCHNORM = True


class RiRBlock(rir.RiRBlockBase):
    """reference :49-66."""
    LAYER = ResnetInit


class Discriminator(rir.RiRDiscriminatorBase):
    """reference :69-115."""
    BLOCK = RiRBlock
    LAYERS_PER_BLOCK = 3
    PRE_BATCHNORM = True

    def make_head(self):
        return nn.Sequential(L.Conv2d(512, 1, kernel_size=4, stride=1, padding=0))

    def __init__(self):
        super().__init__()
        self.LastActivate = L.Sigmoid()

    def forward(self, x):
        return self.LastActivate(self.trunk(x))
