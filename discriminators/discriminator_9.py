"""discriminator_9 — drop-in for the reference's discriminators/discriminator_9.py: the self-attention critic at 64
base features with an attention layer at every resolution (64x64 ... 4x4) and two plain blocks per level.  The building
blocks are those of discriminator_9_3 (same classes in the reference, file for file); the names are re-exported here so
that module paths such as `discriminators.discriminator_9.SelfAttention` resolve as they do in the reference."""
from torch import nn

from discriminators.discriminator_9_3 import (EqualizedWeight, EqualizedLinear, EqualizedConv2d, SEBlock,  # noqa: F401
                                              SelfAttention, MiniBatchStdDev, Smooth, DownSample, DiscriminatorBlock,
                                              FULL_PLAN, build_trunk)


class Discriminator(nn.Module):
    """reference :151-188"""
    FEATURES = 64

    def __init__(self):
        super(Discriminator, self).__init__()
        self.conv, self.fc = build_trunk(self.FEATURES, FULL_PLAN)

    def forward(self, input):
        out = self.conv(input)
        return self.fc(out.view(out.shape[0], -1))
