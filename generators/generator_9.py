"""generator_9 — drop-in for the reference's generators/generator_9.py (units/get_generators.get_9): generator_8's
densely connected tree generator at 48 base planes with SAGAN self-attention (1x1 key / query / value convolutions, a
softmax over all H*W positions, a learned scalar gate `gamma`) after the root of every tree and after every stage's
concatenation — up to 64x64, i.e. 4 096 x 4 096 attention maps.  Attention = gan_b200.layers.self_attention (batched
matrix product + row softmax kernels, csrc/attn.cu).  Same module tree, construction order and RNG consumption as the
reference file."""
import torch
import torch.nn as nn

from gan_b200 import kernels as KN
from gan_b200 import layers as L
from gan_b200 import ops
from generators import generator_7 as g7


class SelfAttention(nn.Module):
    """reference :7-31 — y = gamma * self_att(value(x) softmax(key(x)^T query(x))) + x."""

    def __init__(self, in_planes, embedding_channels):
        super(SelfAttention, self).__init__()
        self.key = L.Conv2d(in_planes, embedding_channels, 1)
        self.query = L.Conv2d(in_planes, embedding_channels, 1)
        self.value = L.Conv2d(in_planes, embedding_channels, 1)
        self.self_att = L.Conv2d(embedding_channels, in_planes, 1)
        self.gamma = nn.Parameter(torch.tensor(0.1))
        self.softmax = nn.Softmax(dim=1)   # kept for the module tree; the attention runs in L.self_attention

    def forward(self, x):
        o = self.self_att(L.self_attention(self.key(x), self.query(x), self.value(x)))
        return ops.fma_bc(x, self.gamma.reshape(1), o, KN.V_SCALAR)


class SEBottleneckSelfAttention(g7.SqueezeExcitationBottleneck):
    """reference :34-73 — generator_7's bottleneck under another name (there is no attention inside it)."""

    def get_out_planes(self):
        return self.out_planes + self.dense_depth

    def __init__(self, last_planes, in_planes, out_planes, dense_depth, kernel_size, feature_size):
        super(SEBottleneckSelfAttention, self).__init__(last_planes, in_planes, out_planes, dense_depth, kernel_size,
                                                        feature_size)
        self.out_planes = out_planes
        self.dense_depth = dense_depth


class ResnetInit(g7.ResnetInit):
    """reference :76-99"""
    BOTTLENECK = SEBottleneckSelfAttention


class BasicBlock(g7.BasicBlock):
    """reference :102-164"""
    RIR = ResnetInit


class Tree(g7.Tree):
    """reference :167-227 — generator_7's tree followed by a SelfAttention on the root's output (sub-trees included)."""
    BLOCK = BasicBlock

    def __init__(self, last_planes, in_planes, out_planes, dense_depth, level, block_num, feature_size):
        super(Tree, self).__init__(last_planes, in_planes, out_planes, dense_depth, level, block_num, feature_size)
        self.attention = SelfAttention(self.root.get_out_planes(), self.root.get_out_planes())

    def forward(self, x):
        return self.attention(super(Tree, self).forward(x))


class Generator(nn.Module):
    """reference :230-308"""
    negative_slope = 1e-2
    PLANES = 48

    def get_upsample(self, planes, out_planes, kernel_size, stride, padding):
        return nn.Sequential(
            L.ConvTranspose2d(planes, out_planes, kernel_size=kernel_size, stride=stride, padding=padding),
            L.BatchNorm2d(out_planes),
            L.PReLU(),
        )

    def _stages(self, planes):
        """(stage width, tree width, dense depth, level, block_num, feature_size)"""
        return [(planes * 8, planes * 2, planes // 4, 2, 4, 8), (planes * 4, planes, planes // 8, 2, 4, 16),
                (planes * 2, planes // 2, planes // 16, 1, 3, 32), (planes, planes // 4, planes // 32, 1, 2, 64)]

    TREE = Tree
    ATTENTION = SelfAttention

    def _last_conv(self, last_planes):
        return L.Conv2d(last_planes, 3, 9, stride=1, padding=4, padding_mode='reflect')

    def __init__(self, z_dim):
        super(Generator, self).__init__()
        planes = self.PLANES
        up = lambda cin, cout, k=4, s=2, p=1: nn.Sequential(
            self.get_upsample(planes=cin, out_planes=cout, kernel_size=k, stride=s, padding=p))
        self.Upsample_1 = up(z_dim, planes * 16, 4, 1, 0)
        last_planes = planes * 16
        for n, (width, tw, dd, level, block_num, feature_size) in enumerate(self._stages(planes), start=1):
            setattr(self, "Upsample_%d" % (n + 1), up(last_planes, width))
            t = self.TREE(width, in_planes=tw, out_planes=tw, dense_depth=dd, level=level, block_num=block_num,
                          feature_size=feature_size)
            last_planes = t.get_out_planes() + width
            setattr(self, "Tree_%d" % n, nn.Sequential(t))
            setattr(self, "Attention_%d" % n, self.ATTENTION(last_planes, last_planes))
        self.LastConv = self._last_conv(last_planes)
        self.LastActivate = L.Tanh()

    def forward(self, x):
        x = self.Upsample_1(x)
        for n in range(1, 5):
            x = getattr(self, "Upsample_%d" % (n + 1))(x)
            x = ops.cat_channels([x, getattr(self, "Tree_%d" % n)(x)])
            x = getattr(self, "Attention_%d" % n)(x)
        return self.LastActivate(self.LastConv(x))
