"""generator_13_2 — drop-in for the reference's generators/generator_13_2.py, which is byte for byte its generator_13.py:
the classes are generator_13's."""
from generators.generator_13 import *  # noqa: F401,F403
from generators.generator_13 import (Smooth, EqualizedWeight, EqualizedLinear, EqualizedConv2d, MappingNetwork,  # noqa: F401
                                     SKAttention_conv, SKAttention_fc, SKConvT, Conv2dWeightModulate, StyleConv, SKConv,
                                     StyleBlock, SEBlock_conv, SEBlock_fc, SelfAttention, ResnetInit, BasicBlock, ToRGB,
                                     Tree, GeneratorBlock, GeneratorStart, Generator)
