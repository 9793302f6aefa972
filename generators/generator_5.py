"""generator_5 — drop-in for the reference's generators/generator_5.py (units/get_generators.get_5): the stem of
generator_4 (transposed convolution z -> 2x2, two BatchNorm stem blocks to 8x8) followed by four densely connected
aggregation trees at 8x8, 16x16, 32x32 and 64x64.  A block runs three InstanceNorm bottlenecks (depthwise 3x3 / 5x5 / 7x7,
reflect padding, LeakyReLU 0.2) on the same input, gates each with a squeeze-excite branch, adds their first `out_planes`
channels to the (optionally projected) input and appends their dense tails; a tree concatenates the first `out_planes`
channels of every stage into its root block.  9x9 reflect-padded convolution to RGB, tanh.

Same module tree, construction order and RNG consumption as the reference file."""
import torch.nn as nn

from gan_b200 import layers as L
from gan_b200 import ops


def _inorm(planes):
    return L.InstanceNorm2d(planes, affine=True, track_running_stats=True)


def _cut(x, c0, c1=None):
    return ops.slice_channels(x, c0, c1)


class BasicBlock(nn.Module):
    """reference :5-92"""
    negative_slope = 0.2

    def _act(self):
        return L.LeakyReLU(negative_slope=self.negative_slope, inplace=True)

    def get_conv_bottleneck(self, last_planes, in_planes, out_planes, dense_depth, kernel_size):
        return nn.Sequential(
            L.Conv2d(last_planes, in_planes, kernel_size=1, stride=1, padding=0), _inorm(in_planes), self._act(),
            L.Conv2d(in_planes, in_planes, kernel_size=kernel_size, stride=1, padding=(kernel_size - 1) // 2,
                     padding_mode='reflect', groups=in_planes), _inorm(in_planes), self._act(),
            L.Conv2d(in_planes, out_planes + dense_depth, kernel_size=1, stride=1, padding=0),
            _inorm(out_planes + dense_depth),
        )

    def get_se(self, in_planes, out_planes, dense_depth, feature_size):
        return nn.Sequential(
            L.AvgPool2d(kernel_size=feature_size, padding=0),
            L.Conv2d(out_planes + dense_depth, in_planes, kernel_size=1), self._act(),
            L.Conv2d(in_planes, out_planes + dense_depth, kernel_size=1), L.Sigmoid(),
        )

    def get_out_planes(self):
        if self.is_unify or self.root:
            return self.out_planes + 4 * self.dense_depth
        return self.last_planes + 3 * self.dense_depth

    def __init__(self, last_planes, in_planes, out_planes, dense_depth, root, feature_size, is_unify):
        super(BasicBlock, self).__init__()
        self.root = root
        self.last_planes = last_planes
        self.out_planes = out_planes
        self.dense_depth = dense_depth
        self.is_unify = is_unify
        self.unify = nn.Sequential()
        width = last_planes
        if is_unify:
            width = out_planes + dense_depth
            self.unify = nn.Sequential(L.Conv2d(last_planes, width, kernel_size=1, stride=1, padding=0), _inorm(width))
        for k in (3, 5, 7):
            setattr(self, "model_%d" % k, self.get_conv_bottleneck(width, in_planes, out_planes, dense_depth, k))
        self.shortcut = nn.Sequential()
        if root:
            self.shortcut = nn.Sequential(
                L.Conv2d(last_planes, out_planes + dense_depth, kernel_size=1, stride=1, padding=0),
                _inorm(out_planes + dense_depth))
        for k in (3, 5, 7):
            setattr(self, "se_%d" % k, self.get_se(in_planes, out_planes, dense_depth, feature_size))

    def _branch(self, k, x):
        out = getattr(self, "model_%d" % k)(x)
        se = getattr(self, "se_%d" % k)(out)
        if se.shape[-2:] == (1, 1):
            return L.gate(out, se)
        return ops.mul(out, se.expand_as(out))

    def forward(self, x):
        x = self.unify(x)
        out3, out5, out7 = self._branch(3, x), self._branch(5, x), self._branch(7, x)
        x = self.shortcut(x)
        d = self.out_planes
        head = ops.add(ops.add(ops.add(_cut(x, 0, d), _cut(out3, 0, d)), _cut(out5, 0, d)), _cut(out7, 0, d))
        return ops.cat_channels([head, _cut(x, d), _cut(out3, d), _cut(out5, d), _cut(out7, d)])


class Tree(nn.Module):
    """reference :94-158"""

    def get_out_planes(self):
        return self.root.get_out_planes()

    def __init__(self, last_planes, in_planes, out_planes, dense_depth, level, block_num, feature_size):
        super(Tree, self).__init__()
        self.level = level
        self.block_num = block_num
        self.out_planes = out_planes
        self.dense_depth = dense_depth
        plain = lambda planes, unify: BasicBlock(planes, in_planes, out_planes, dense_depth, False, feature_size, unify)
        self.root_last_planes = out_planes * (block_num - 1)
        first = 0
        if level == 1:
            sub_block = plain(last_planes, last_planes < out_planes)
            last_planes = sub_block.get_out_planes()
            self.__setattr__('block_%d' % 0, sub_block)
            first = 1
        else:
            self.prev_root = plain(last_planes, last_planes < out_planes)
            self.root_last_planes += self.prev_root.get_out_planes()
            for i in reversed(range(1, level)):
                subtree = Tree(last_planes, in_planes, out_planes, dense_depth, i, block_num, feature_size)
                last_planes = subtree.get_out_planes()
                self.root_last_planes += last_planes
                self.__setattr__('level_%d' % i, subtree)
        for i in range(first, block_num):
            sub_block = plain(last_planes, False)
            last_planes = sub_block.get_out_planes()
            self.__setattr__('block_%d' % i, sub_block)
        self.root_last_planes += sub_block.get_out_planes()
        self.root = BasicBlock(self.root_last_planes, in_planes * block_num, out_planes, dense_depth, True,
                               feature_size, False)

    def forward(self, x):
        d = self.out_planes
        xs = [self.prev_root(x)] if self.level > 1 else []
        for i in reversed(range(1, self.level)):
            x = self.__getattr__('level_%d' % i)(x)
            xs.append(x)
        for i in range(self.block_num):
            x = self.__getattr__('block_%d' % i)(x)
            xs.append(_cut(x, 0, d))
        xs.append(_cut(x, d))
        return self.root(ops.cat_channels(xs))


class Stem_block(nn.Module):
    """reference :160-222 — the first bottleneck changes the width (its skip is the 1x1 + BatchNorm `shortcut`), the others
    are residual; full-map SE gates; ConvTranspose2d x2 + BatchNorm + LeakyReLU at the end."""
    negative_slope = 0.2

    def _act(self):
        return L.LeakyReLU(negative_slope=self.negative_slope, inplace=True)

    def get_conv_group(self, in_planes, planes, kernel_size, expansion=1):
        mid = in_planes * expansion
        return nn.Sequential(
            L.Conv2d(in_planes, mid, kernel_size=1, stride=1, padding=0), L.BatchNorm2d(mid), self._act(),
            L.Conv2d(mid, mid, kernel_size=kernel_size, stride=1, padding=(kernel_size - 1) // 2,
                     padding_mode='reflect', groups=mid), L.BatchNorm2d(mid), self._act(),
            L.Conv2d(mid, planes, kernel_size=1, stride=1, padding=0), L.BatchNorm2d(planes),
        )

    def get_se(self, planes):
        return nn.Sequential(
            L.Conv2d(planes, planes, kernel_size=1), self._act(),
            L.Conv2d(planes, planes, kernel_size=1), L.Sigmoid(),
        )

    def __init__(self, in_planes, planes, block_num=3, kernel_size=3):
        super(Stem_block, self).__init__()
        self.block_num = block_num
        self.shortcut = nn.Sequential()
        if in_planes != planes:
            self.shortcut = nn.Sequential(L.Conv2d(in_planes, planes, kernel_size=1, stride=1, padding=0),
                                          L.BatchNorm2d(planes))
        for i in range(block_num):
            self.__setattr__('res_%d' % i, self.get_conv_group(in_planes, planes, kernel_size))
            in_planes = planes
            self.__setattr__('se_%d' % i, self.get_se(planes))
        self.upsample = nn.Sequential(L.ConvTranspose2d(planes, planes, kernel_size=4, stride=2, padding=1),
                                      L.BatchNorm2d(planes), self._act())

    def forward(self, x):
        for i in range(self.block_num):
            out = self.__getattr__('res_%d' % i)(x)
            out = ops.mul(out, self.__getattr__('se_%d' % i)(out))
            x = ops.add(self.shortcut(x) if i == 0 else x, out)
        return self.upsample(x)


class Generator(nn.Module):
    """reference :224-267"""
    negative_slope = 0.2

    def get_upsample(self, planes):
        return nn.Sequential(
            L.ConvTranspose2d(planes, planes, kernel_size=4, stride=2, padding=1), _inorm(planes),
            L.LeakyReLU(negative_slope=self.negative_slope, inplace=True)
        )

    def __init__(self, z_dim):
        super(Generator, self).__init__()
        last_planes = 512
        model = [L.ConvTranspose2d(z_dim, last_planes, kernel_size=4, stride=2, padding=1), L.BatchNorm2d(last_planes),
                 L.LeakyReLU(negative_slope=self.negative_slope, inplace=True)]                          # 2x2
        model += [Stem_block(in_planes=last_planes, planes=last_planes // 2, block_num=3, kernel_size=3)]        # 4x4
        model += [Stem_block(in_planes=last_planes // 2, planes=last_planes // 2, block_num=4, kernel_size=3)]   # 8x8
        last_planes = last_planes // 2
        plans = [dict(in_planes=64, out_planes=128, dense_depth=16, level=3, block_num=6, feature_size=8),
                 dict(in_planes=32, out_planes=64, dense_depth=12, level=3, block_num=6, feature_size=16),
                 dict(in_planes=16, out_planes=32, dense_depth=8, level=2, block_num=5, feature_size=32),
                 dict(in_planes=8, out_planes=16, dense_depth=4, level=2, block_num=4, feature_size=64)]
        for plan in plans[:2]:
            t = Tree(last_planes, **plan)
            last_planes = t.get_out_planes()
            model += [t, self.get_upsample(planes=last_planes)]
        t3 = Tree(last_planes, **plans[2])
        last_planes = t3.get_out_planes()
        t4 = Tree(last_planes, **plans[3])       # built before the transposed convolution between the two (RNG order)
        model += [t3, self.get_upsample(planes=last_planes), t4]
        last_planes = t4.get_out_planes()
        model += [L.Conv2d(last_planes, 3, 9, stride=1, padding=4, padding_mode='reflect'), L.Tanh()]
        self.model = nn.Sequential(*model)

    def forward(self, x):
        return self.model(x)
