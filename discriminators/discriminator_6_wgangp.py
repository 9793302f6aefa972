"""discriminator_6_wgangp — drop-in for the reference's discriminators/discriminator_6_wgangp.py (the critic behind
units/get_discriminator.get_6_wgan): a deep-layer-aggregation tree of ResNet-in-ResNet blocks.  Every block runs three
pairs of dual-path streams (depthwise 3x3, 5x5 and 7x7 bottlenecks, each gated by a squeeze-excite branch), three trees
at 32x32, 16x16 and 8x8, reflect-padded 5x5 stride-2 convolutions between them; LeakyReLU(0.01); no normalisation, so it
pairs with the gradient-penalty loops.  (3,64,64) -> (B,1,1,1).

Same module tree as the reference (`model.Tree_1.level_1.block_0.rir_5.transient_across.res.2.weight` ...), same
construction order, so state_dict keys, optimiser parameter order and the RNG stream consumed at construction match.
On libganb200: 1x1 convs and the 5x5 stride-2 convs = the conv engine after a reflect-pad gather, depthwise 3x3 / 5x5 /
7x7 = csrc/dwconv.cu on the reflect-padded map, the AvgPool2d(feature_size) squeeze = the separable resampler, channel
slices / concatenations = the channel-copy kernels; all closed under the double backward of the penalties.
"""
import torch.nn as nn

from gan_b200 import layers as L
from gan_b200 import ops

SLOPE = 1e-2


def _act():
    return L.LeakyReLU(negative_slope=SLOPE, inplace=True)


def _cut(x, c0, c1=None):
    return ops.slice_channels(x, c0, c1)


def _sum3(a, b, c):
    return ops.add(ops.add(a, b), c)


class SqueezeExcitationBottleneck(nn.Module):
    """reference :5-35 — 1x1, depthwise k x k (reflect), 1x1 to `out_planes + dense_depth` channels, times an SE gate
    computed from that output (average over feature_size x feature_size windows, two 1x1 convs, sigmoid)."""
    negative_slope = SLOPE

    def get_conv_bottleneck(self, last_planes, in_planes, out_planes, dense_depth, kernel_size):
        return nn.Sequential(
            L.Conv2d(last_planes, in_planes, kernel_size=1, stride=1, padding=0),
            _act(),
            L.Conv2d(in_planes, in_planes, kernel_size=kernel_size, stride=1, padding=(kernel_size - 1) // 2,
                     padding_mode='reflect', groups=in_planes),
            _act(),
            L.Conv2d(in_planes, out_planes + dense_depth, kernel_size=1, stride=1, padding=0),
        )

    def get_se(self, in_planes, out_planes, dense_depth, feature_size):
        return nn.Sequential(
            L.AvgPool2d(kernel_size=feature_size, padding=0),
            L.Conv2d(out_planes + dense_depth, in_planes, kernel_size=1),
            _act(),
            L.Conv2d(in_planes, out_planes + dense_depth, kernel_size=1),
            L.Sigmoid(),
        )

    def __init__(self, last_planes, in_planes, out_planes, dense_depth, kernel_size, feature_size):
        super(SqueezeExcitationBottleneck, self).__init__()
        self.res = self.get_conv_bottleneck(last_planes, in_planes, out_planes, dense_depth, kernel_size)
        self.se = self.get_se(in_planes, out_planes, dense_depth, feature_size)

    def forward(self, x):
        out = self.res(x)
        se = self.se(out)
        if se.shape[-2:] == (1, 1):          # the map is exactly feature_size on a side: a per-(sample, channel) gate
            return L.gate(out, se)
        return ops.mul(out, se.expand_as(out))


class ResnetInit(nn.Module):
    """reference :38-63 — the residual and the transient stream each feed both outputs.
    (BOTTLENECK / RIR / BLOCK / _pointwise below: the tree generators generator_7 / generator_8 are this tree with
    BatchNorm + PReLU bottlenecks; they subclass these classes and swap the parts.)"""
    BOTTLENECK = SqueezeExcitationBottleneck

    def __init__(self, last_planes, in_planes, out_planes, dense_depth, kernel_size, feature_size):
        super(ResnetInit, self).__init__()
        mk = lambda dd: self.BOTTLENECK(last_planes, in_planes, out_planes, dd, kernel_size, feature_size)
        self.residual = mk(dense_depth)
        self.transient = mk(0)
        self.residual_across = mk(0)
        self.transient_across = mk(dense_depth)

    def forward(self, x):
        x_residual, x_transient = x
        return (ops.add(self.residual(x_residual), self.transient_across(x_transient)),
                ops.add(self.residual_across(x_residual), self.transient(x_transient)))


class BasicBlock(nn.Module):
    """reference :66-125 — optional 1x1 `unify`, the three ResnetInit pairs on the same two streams, optional 1x1
    `shortcut` (the aggregation root), and the dense concatenation of the three residual tails."""

    RIR = ResnetInit

    @staticmethod
    def _pointwise(in_planes, out_planes):
        """the layers of `unify` / `shortcut`"""
        return [L.Conv2d(in_planes, out_planes, kernel_size=1, stride=1, padding=0)]

    def get_out_planes(self):
        if self.is_unify or self.root:
            return 2 * self.out_planes + 4 * self.dense_depth
        return self.last_planes + 3 * self.dense_depth

    def __init__(self, last_planes, in_planes, out_planes, dense_depth, root, feature_size, is_unify):
        super(BasicBlock, self).__init__()
        self.root = root
        self.last_planes = last_planes
        self.out_planes = out_planes
        self.dense_depth = dense_depth
        self.is_unify = is_unify
        self.unify = nn.Sequential()
        if is_unify:
            self.unify = nn.Sequential(*self._pointwise(last_planes, 2 * out_planes + dense_depth))
            stream = out_planes + dense_depth
        else:
            stream = last_planes - out_planes
        self.rir_3 = self.RIR(stream, in_planes, out_planes, dense_depth, 3, feature_size)
        self.rir_5 = self.RIR(stream, in_planes, out_planes, dense_depth, 5, feature_size)
        self.rir_7 = self.RIR(stream, in_planes, out_planes, dense_depth, 7, feature_size)
        self.shortcut = nn.Sequential()
        if root:
            self.shortcut = nn.Sequential(*self._pointwise(last_planes, 2 * out_planes + dense_depth))

    def forward(self, x):
        d = self.out_planes
        x = self.unify(x)
        streams = (ops.cat_channels([_cut(x, 0, d), _cut(x, 2 * d)]), _cut(x, d))
        r3, t3 = self.rir_3(streams)
        r5, t5 = self.rir_5(streams)
        r7, t7 = self.rir_7(streams)
        x = self.shortcut(x)
        head = ops.add(_cut(x, 0, d), _sum3(_cut(r3, 0, d), _cut(r5, 0, d), _cut(r7, 0, d)))
        return ops.cat_channels([head, _sum3(t3, t5, t7), _cut(x, 2 * d), _cut(r3, d), _cut(r5, d), _cut(r7, d)])


class Tree(nn.Module):
    """reference :128-189 — deep layer aggregation: sub-trees and blocks in sequence, the first 2*out_planes channels
    of every stage (and the dense tail of the last) concatenated into the root block."""

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
        self.root_last_planes = 2 * out_planes * (block_num - 1)
        first = 0
        if level == 1:
            sub_block = plain(last_planes, last_planes < 2 * out_planes)
            last_planes = sub_block.get_out_planes()
            self.__setattr__('block_%d' % 0, sub_block)
            first = 1
        else:
            self.prev_root = plain(last_planes, last_planes < 2 * out_planes)
            self.root_last_planes += self.prev_root.get_out_planes()
            for i in reversed(range(1, level)):
                subtree = type(self)(last_planes, in_planes, out_planes, dense_depth, i, block_num, feature_size)
                last_planes = subtree.get_out_planes()
                self.root_last_planes += last_planes
                self.__setattr__('level_%d' % i, subtree)
        for i in range(first, block_num):
            sub_block = plain(last_planes, False)
            last_planes = sub_block.get_out_planes()
            self.__setattr__('block_%d' % i, sub_block)
        self.root_last_planes += sub_block.get_out_planes()
        self.root = self.BLOCK(self.root_last_planes, in_planes * block_num, out_planes, dense_depth, True,
                               feature_size, False)

    def forward(self, x):
        d = self.out_planes
        xs = [self.prev_root(x)] if self.level > 1 else []
        for i in reversed(range(1, self.level)):
            x = self.__getattr__('level_%d' % i)(x)
            xs.append(x)
        for i in range(self.block_num):
            x = self.__getattr__('block_%d' % i)(x)
            xs.append(_cut(x, 0, 2 * d))
        xs.append(_cut(x, 2 * d))
        return self.root(ops.cat_channels(xs))


class Discriminator(nn.Module):
    """reference :192-221"""
    negative_slope = SLOPE

    def get_downsample(self, planes, out_planes):
        return nn.Sequential(
            L.Conv2d(planes, out_planes, kernel_size=5, stride=2, padding=2, padding_mode='reflect'),
            _act(),
        )

    def __init__(self):
        super(Discriminator, self).__init__()
        self.model = nn.Sequential()
        self.model.add_module("Downsample_0", self.get_downsample(planes=3, out_planes=16))
        last_planes = 16
        for n, (in_planes, out_planes, feature_size) in enumerate([(6, 12, 32), (12, 16, 16), (16, 24, 8)], start=1):
            tree = Tree(last_planes, in_planes=in_planes, out_planes=out_planes, dense_depth=1, level=2, block_num=2,
                        feature_size=feature_size)
            last_planes = tree.get_out_planes()
            self.model.add_module("Tree_%d" % n, tree)
            self.model.add_module("Downsample_%d" % n, self.get_downsample(planes=last_planes, out_planes=last_planes))
        self.model.add_module("LastConv", L.Conv2d(last_planes, 1, 4, stride=1, padding=0))

    def forward(self, x):
        return self.model(x)
