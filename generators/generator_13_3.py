"""generator_13_3 — drop-in for the reference's generators/generator_13_3.py: generator_13_4 with shallower selective-kernel
attention trunks (two 3x3 convs, one / two linears; the SE blocks keep their depth), a learned per-channel scale instead of the SK merge of a block's skip, 32 base
planes and one block per tree level.  Everything else is generator_13_4's code, re-parametrised through the class hooks."""
import torch
import torch.nn as nn

from gan_b200 import kernels as KN
from gan_b200 import layers as L  # noqa: F401
from gan_b200 import ops
from generators import generator_13_4 as g
from generators.generator_13_4 import EqualizedWeight, EqualizedLinear, EqualizedConv2d, Smooth  # noqa: F401

DEPTH = dict(conv=2, fc=1, fc_only=2)


class SKAttention_conv(g.SKAttention_conv):
    """reference :41-89"""
    DEPTH = DEPTH


class SKAttention_fc(g.SKAttention_fc):
    """reference :92-131"""
    DEPTH = DEPTH


def _pick_sk_attention(planes, m, image_size):
    return g._pick_sk_attention(planes, m, image_size, SKAttention_conv, SKAttention_fc)


class SKConvT(g.SKConvT):
    """reference :152-169"""
    SK_CONV = SKAttention_conv


class SKConv(g.SKConv):
    """reference :172-201"""
    SK = staticmethod(_pick_sk_attention)


class StyleBlock(g.StyleBlock):
    """reference :204-231"""
    SKCONV = SKConv


class ResnetInit(g.ResnetInit):
    """reference :234-261"""
    BLOCK = StyleBlock
    SK = staticmethod(_pick_sk_attention)


class SEBlock_conv(g.SEBlock_conv):
    """reference :264-294 (as deep as generator_13_4's)"""


class SEBlock_fc(g.SEBlock_fc):
    """reference :297-317"""


class SelfAttention(g.SelfAttention):
    """reference :320-349"""
    SE_CONV, SE_FC = SEBlock_conv, SEBlock_fc


class BasicBlock(g.BasicBlock):
    """reference :352-408 — skip + scale * residual head (`scale` ~ U(0.25, 0.35) per channel) in the place of 13_4's SK
    merge."""
    RIR = ResnetInit
    ATTENTION = SelfAttention

    def _first(self, out_planes):
        self.scale = nn.Parameter(torch.nn.init.uniform_(torch.empty(out_planes), a=0.25, b=0.35))

    def _skip_merge(self, out_planes, image_size):
        pass

    def _merge_skip(self, head, res_head):
        return ops.fma_bc(head, self.scale, res_head, KN.V_C)


class ToRGB(g.ToRGB):
    """reference :411-435"""
    SKCONV = SKConv


class Tree(g.Tree):
    """reference :438-509"""
    BLOCK = BasicBlock
    TORGB = ToRGB
    SK = staticmethod(_pick_sk_attention)


class GeneratorBlock(g.GeneratorBlock):
    """reference :512-527"""
    TREE = Tree
    UP = SKConvT


class GeneratorStart(g.GeneratorStart):
    """reference :530-550"""
    TREE = Tree
    TORGB = ToRGB


class Generator(g.Generator):
    """reference :553-576"""
    START, BLOCK = GeneratorStart, GeneratorBlock
    PLAN = [(1, 1), (1, 1), (1, 1), (1, 1), (1, 1)]

    def __init__(self, z_dim, planes=32):
        super(Generator, self).__init__(z_dim, planes)
