"""generator_10 — drop-in for the reference's generators/generator_10.py (units/get_generators.get_10): generator_9 with
lighter blocks and more attention — a block runs ONE ResNet-in-ResNet pair (depthwise 3x3 bottlenecks) and ends in its own
SelfAttention, a tree concatenates the whole output of every stage into its root, every tree is level 1 with two blocks,
64 base planes; the attention gates `gamma` start at random.random() * 0.2 + 0.03 (Python's `random`, as in the
reference).  Same module tree, construction order and RNG consumption as the reference file."""
import random

import torch
import torch.nn as nn

from gan_b200 import layers as L
from gan_b200 import ops
from generators import generator_9 as g9


def _cut(x, c0, c1=None):
    return ops.slice_channels(x, c0, c1)


class SelfAttention(g9.SelfAttention):
    """reference :8-30"""

    def __init__(self, in_planes, embedding_channels):
        super(SelfAttention, self).__init__(in_planes, embedding_channels)
        self.gamma = nn.Parameter(torch.tensor(random.random() * 0.2 + 0.03))


class SEBottleneckSelfAttention(g9.SEBottleneckSelfAttention):
    """reference :32-71"""


class ResnetInit(g9.ResnetInit):
    """reference :74-97"""
    BOTTLENECK = SEBottleneckSelfAttention


class BasicBlock(nn.Module):
    """reference :99-151"""
    RIR = ResnetInit
    ATTENTION = SelfAttention

    @staticmethod
    def _pointwise(in_planes, out_planes):
        return [L.Conv2d(in_planes, out_planes, kernel_size=1, stride=1, padding=0), L.BatchNorm2d(out_planes)]

    def get_out_planes(self):
        if self.is_unify or self.root:
            return 2 * self.out_planes + 2 * self.dense_depth
        return self.last_planes + 1 * self.dense_depth

    def __init__(self, last_planes, in_planes, out_planes, dense_depth, root, feature_size, is_unify):
        super(BasicBlock, self).__init__()
        self.root = root
        self.last_planes = last_planes
        self.out_planes = out_planes
        self.dense_depth = dense_depth
        self.is_unify = is_unify
        self.unify = nn.Sequential()
        stream = last_planes - out_planes
        if is_unify:
            self.unify = nn.Sequential(*self._pointwise(last_planes, 2 * out_planes + dense_depth))
            stream = out_planes + dense_depth
        self.rir_3 = self.RIR(stream, in_planes, out_planes, dense_depth, 3, feature_size)
        self.shortcut = nn.Sequential()
        if root:
            self.shortcut = nn.Sequential(*self._pointwise(last_planes, 2 * out_planes + dense_depth))
        self.attention = self.ATTENTION(self.get_out_planes(), self.get_out_planes())

    def forward(self, x):
        d = self.out_planes
        x = self.unify(x)
        r3, t3 = self.rir_3((ops.cat_channels([_cut(x, 0, d), _cut(x, 2 * d)]), _cut(x, d)))
        x = self.shortcut(x)
        out = ops.cat_channels([ops.add(_cut(x, 0, d), _cut(r3, 0, d)), t3, _cut(x, 2 * d), _cut(r3, d)])
        return self.attention(out)


class Tree(nn.Module):
    """reference :154-213"""
    BLOCK = BasicBlock

    def get_out_planes(self):
        return self.root.get_out_planes()

    def __init__(self, last_planes, in_planes, out_planes, dense_depth, level, block_num, feature_size):
        super(Tree, self).__init__()
        assert (block_num > 0)
        self.level = level
        self.block_num = block_num
        self.out_planes = out_planes
        self.dense_depth = dense_depth
        plain = lambda planes, unify: self.BLOCK(planes, in_planes, out_planes, dense_depth, False, feature_size, unify)
        first = 0
        if level == 1:
            sub_block = plain(last_planes, last_planes < 2 * out_planes)
            last_planes = sub_block.get_out_planes()
            self.root_last_planes = last_planes
            self.__setattr__('block_%d' % 0, sub_block)
            first = 1
        else:
            self.prev_root = plain(last_planes, last_planes < 2 * out_planes)
            self.root_last_planes = self.prev_root.get_out_planes()
            for i in reversed(range(1, level)):
                subtree = type(self)(last_planes, in_planes, out_planes, dense_depth, i, block_num, feature_size)
                last_planes = subtree.get_out_planes()
                self.root_last_planes += last_planes
                self.__setattr__('level_%d' % i, subtree)
        for i in range(first, block_num):
            sub_block = plain(last_planes, False)
            last_planes = sub_block.get_out_planes()
            self.root_last_planes += last_planes
            self.__setattr__('block_%d' % i, sub_block)
        self.root = self.BLOCK(self.root_last_planes, in_planes * block_num, out_planes, dense_depth, True,
                               feature_size, False)

    def forward(self, x):
        xs = [self.prev_root(x)] if self.level > 1 else []
        for i in reversed(range(1, self.level)):
            x = self.__getattr__('level_%d' % i)(x)
            xs.append(x)
        for i in range(self.block_num):
            x = self.__getattr__('block_%d' % i)(x)
            xs.append(x)
        return self.root(ops.cat_channels(xs))


class Generator(g9.Generator):
    """reference :216-296"""
    PLANES = 64
    TREE = Tree
    ATTENTION = SelfAttention

    def _stages(self, planes):
        return [(planes * 8, planes * 2, planes // 4, 1, 2, 8), (planes * 4, planes, planes // 8, 1, 2, 16),
                (planes * 2, planes // 2, planes // 16, 1, 2, 32), (planes, planes // 4, planes // 32, 1, 2, 64)]
