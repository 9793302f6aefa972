"""discriminator_9_2 — drop-in for the reference's discriminators/discriminator_9_2.py: discriminator_9 at 16 base
features (the only difference between the two reference files, :154)."""
from discriminators.discriminator_9 import (EqualizedWeight, EqualizedLinear, EqualizedConv2d, SEBlock,  # noqa: F401
                                            SelfAttention, MiniBatchStdDev, Smooth, DownSample, DiscriminatorBlock)
from discriminators import discriminator_9 as _d9


class Discriminator(_d9.Discriminator):
    """reference :151-188"""
    FEATURES = 16
