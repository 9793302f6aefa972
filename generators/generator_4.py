"""generator_4 — drop-in for the reference's generators/generator_4.py (units/get_generators.get_4): a transposed
convolution z -> 2x2, two BatchNorm stem blocks (residual depthwise bottlenecks with full-map SE gates, each ending in a
x2 transposed convolution) up to 8x8, then four aggregation trees of ShuffleNet-style blocks at 8x8 ... 64x64 joined by
InstanceNorm transposed convolutions, a reflect-padded 7x7 convolution to RGB and tanh.  A block splits its channels in
four: one quarter passes through, the other three run depthwise 3x3 / 5x5 / 7x7 bottlenecks (InstanceNorm2d with affine
parameters and running statistics, LeakyReLU 0.2) with a squeeze-excite residual, and the four are shuffled together.

Same module tree, construction order and RNG consumption as the reference file.  On libganb200: 1x1 convs = the conv
engine, depthwise convs = csrc/dwconv.cu after the reflect-pad gather, InstanceNorm = the instance-norm kernels, channel
split / concat / shuffle = the channel-copy and shuffle kernels."""
import torch.nn as nn

from gan_b200 import layers as L
from gan_b200 import ops


class ShuffleBlock(nn.Module):
    """reference :9-18 — [N, g, C/g, H, W] -> [N, C/g, g, H, W]."""

    def __init__(self, groups):
        super(ShuffleBlock, self).__init__()
        self.groups = groups

    def forward(self, x):
        return ops.channel_shuffle(x, self.groups)


class SplitBlock(nn.Module):
    """reference :20-27 — the first int(C * ratio) channels and the rest."""

    def __init__(self, ratio):
        super(SplitBlock, self).__init__()
        self.ratio = ratio

    def forward(self, x):
        c = int(x.size(1) * self.ratio)
        return ops.slice_channels(x, 0, c), ops.slice_channels(x, c)


def _inorm(planes):
    return L.InstanceNorm2d(planes, affine=True, track_running_stats=True)


def _depthwise(planes, kernel_size):
    return L.Conv2d(planes, planes, kernel_size=kernel_size, stride=1, padding=(kernel_size - 1) // 2,
                    padding_mode='reflect', groups=planes)


class BasicBlock(nn.Module):
    """reference :29-87"""
    negative_slope = 0.2

    def _act(self):
        return L.LeakyReLU(negative_slope=self.negative_slope, inplace=True)

    def get_conv_group(self, planes, kernel_size):
        return nn.Sequential(
            L.Conv2d(planes // 4, planes, kernel_size=1, stride=1, padding=0), _inorm(planes), self._act(),
            _depthwise(planes, kernel_size), _inorm(planes), self._act(),
            L.Conv2d(planes, planes // 4, kernel_size=1, stride=1, padding=0), _inorm(planes // 4),
        )

    def get_se(self, planes):
        return nn.Sequential(
            L.Conv2d(planes // 4, planes, kernel_size=1), self._act(),
            L.Conv2d(planes, planes // 4, kernel_size=1), L.Sigmoid(),
        )

    def __init__(self, in_planes, planes):
        super(BasicBlock, self).__init__()
        self.split = SplitBlock(0.5)
        self.shuffle = ShuffleBlock(groups=4)
        self.unify = nn.Sequential()
        if in_planes != planes:
            self.unify = nn.Sequential(L.Conv2d(in_planes, planes, kernel_size=1, stride=1, padding=0), _inorm(planes),
                                       self._act())
        for k in (3, 5, 7):
            setattr(self, "model_%d" % k, self.get_conv_group(planes, k))
            setattr(self, "se_%d" % k, self.get_se(planes))

    def _branch(self, k, x):
        out = getattr(self, "model_%d" % k)(x)
        assert out.size(2) == out.size(3), "F.avg_pool2d(out, out.size(2)) is a global pool on square maps only"
        b, c = out.shape[:2]
        gate = getattr(self, "se_%d" % k)(ops.global_avg_pool(out).reshape(b, c, 1, 1))
        return L.gated_residual(x, out, gate)

    def forward(self, x):
        x = self.unify(x)
        xl, xr = self.split(x)
        xn, x3 = self.split(xl)
        x5, x7 = self.split(xr)
        out = ops.cat_channels([self._branch(3, x3), self._branch(5, x5), self._branch(7, x7), xn])
        return self.shuffle(out)


class Tree(nn.Module):
    """reference :89-129 — `prev_root`, sub-trees `level_i`, `block_i` in sequence; every stage's output concatenated
    into `root`."""

    def __init__(self, block, in_planes, planes, level=1, block_num=4, origin=True):
        super(Tree, self).__init__()
        self.level = level
        self.block_num = block_num
        self.prev_root = block(in_planes if origin else planes, planes)
        if level == 1:
            self.root = block(block_num * planes, planes)
            self.__setattr__('block_%d' % 0, block(in_planes, planes))
        else:
            self.root = block((level + block_num) * planes, planes)
            for i in reversed(range(1, level)):
                subtree = Tree(block, in_planes if origin else planes, planes, level=i, origin=origin)
                self.__setattr__('level_%d' % i, subtree)
                origin = False
            self.__setattr__('block_%d' % 0, block(planes, planes))
        for i in range(1, block_num):
            self.__setattr__('block_%d' % i, block(planes, planes))

    def forward(self, x):
        xs = [self.prev_root(x)] if self.level > 1 else []
        for i in reversed(range(1, self.level)):
            x = self.__getattr__('level_%d' % i)(x)
            xs.append(x)
        for i in range(self.block_num):
            x = self.__getattr__('block_%d' % i)(x)
            xs.append(x)
        return self.root(ops.cat_channels(xs))


class Stem_block(nn.Module):
    """reference :131-187 — `block_num` residual bottlenecks x + res(x) * se(res(x)) (the gate is a full map: two 1x1
    convs and a sigmoid on the un-pooled features), then ConvTranspose2d x2 + BatchNorm + LeakyReLU."""
    negative_slope = 0.2

    def _act(self):
        return L.LeakyReLU(negative_slope=self.negative_slope, inplace=True)

    def get_conv_group(self, planes, kernel_size):
        return nn.Sequential(
            L.Conv2d(planes, planes, kernel_size=1, stride=1, padding=0), L.BatchNorm2d(planes), self._act(),
            _depthwise(planes, kernel_size), L.BatchNorm2d(planes), self._act(),
            L.Conv2d(planes, planes, kernel_size=1, stride=1, padding=0), L.BatchNorm2d(planes),
        )

    def get_se(self, planes):
        return nn.Sequential(
            L.Conv2d(planes, planes, kernel_size=1), self._act(),
            L.Conv2d(planes, planes, kernel_size=1), L.Sigmoid(),
        )

    def __init__(self, in_planes, planes, block_num=3, kernel_size=3):
        super(Stem_block, self).__init__()
        self.block_num = block_num
        self.unify = nn.Sequential()
        if in_planes != planes:
            self.unify = nn.Sequential(L.Conv2d(in_planes, planes, kernel_size=1, stride=1, padding=0),
                                       L.BatchNorm2d(planes), self._act())
        for i in range(block_num):
            self.__setattr__('res_%d' % i, self.get_conv_group(planes, kernel_size))
            self.__setattr__('se_%d' % i, self.get_se(planes))
        self.upsample = nn.Sequential(L.ConvTranspose2d(planes, planes, kernel_size=4, stride=2, padding=1),
                                      L.BatchNorm2d(planes), self._act())

    def forward(self, x):
        x = self.unify(x)
        for i in range(self.block_num):
            out = self.__getattr__('res_%d' % i)(x)
            x = ops.add(x, ops.mul(out, self.__getattr__('se_%d' % i)(out)))
        return self.upsample(x)


class Generator(nn.Module):
    """reference :189-229"""
    negative_slope = 0.2

    def get_upsample(self, planes):
        return nn.Sequential(
            L.ConvTranspose2d(planes, planes, kernel_size=4, stride=2, padding=1), _inorm(planes),
            L.LeakyReLU(negative_slope=self.negative_slope, inplace=True)
        )

    def __init__(self, z_dim):
        super(Generator, self).__init__()
        in_planes = 512
        model = [L.ConvTranspose2d(z_dim, in_planes, kernel_size=4, stride=2, padding=1), L.BatchNorm2d(in_planes),
                 L.LeakyReLU(negative_slope=self.negative_slope, inplace=True)]                       # 2x2
        model += [Stem_block(in_planes=in_planes, planes=in_planes // 2, block_num=3, kernel_size=3)]         # 4x4
        model += [Stem_block(in_planes=in_planes // 2, planes=in_planes // 2, block_num=4, kernel_size=3)]    # 8x8
        for div in (4, 8, 16):                                                                        # 8, 16, 32
            model += [Tree(BasicBlock, in_planes=in_planes // (div // 2), planes=in_planes // div, level=2, block_num=3),
                      self.get_upsample(planes=in_planes // div)]
        model += [Tree(BasicBlock, in_planes=in_planes // 16, planes=in_planes // 16, level=2, block_num=3)]  # 64x64
        model += [L.Conv2d(in_planes // 16, 3, 7, stride=1, padding=3, padding_mode='reflect'), L.Tanh()]
        self.model = nn.Sequential(*model)

    def forward(self, x):
        return self.model(x)
