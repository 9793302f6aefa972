"""generator_8 — drop-in for the reference's generators/generator_8.py (units/get_generators.get_8): the building blocks
of generator_7 (BatchNorm + PReLU bottleneck trees) under a densely connected top level — at each of 8x8, 16x16, 32x32 and
64x64 a transposed convolution doubles the map, a tree runs on it, and tree input and tree output are concatenated for the
next stage.  Same module tree, construction order and RNG consumption as the reference file."""
import torch.nn as nn

from gan_b200 import layers as L
from gan_b200 import ops
from generators.generator_7 import SqueezeExcitationBottleneck, ResnetInit, BasicBlock, Tree  # noqa: F401 (same classes)


class Generator(nn.Module):
    """reference :195-260"""
    negative_slope = 1e-2

    def get_upsample(self, planes, out_planes, kernel_size, stride, padding):
        return nn.Sequential(
            L.ConvTranspose2d(planes, out_planes, kernel_size=kernel_size, stride=stride, padding=padding),
            L.BatchNorm2d(out_planes),
            L.PReLU(),
        )

    def __init__(self, z_dim):
        super(Generator, self).__init__()
        planes = 64
        up = lambda cin, cout, k=4, s=2, p=1: nn.Sequential(
            self.get_upsample(planes=cin, out_planes=cout, kernel_size=k, stride=s, padding=p))
        self.Upsample_1 = up(z_dim, planes * 16, 4, 1, 0)
        last_planes = planes * 16
        # (stage width, tree width, dense depth, level, block_num, feature_size)
        stages = [(planes * 8, planes * 2, planes // 4, 2, 4, 8), (planes * 4, planes, planes // 8, 2, 4, 16),
                  (planes * 2, planes // 2, planes // 16, 1, 3, 32), (planes, planes // 4, planes // 32, 1, 2, 64)]
        for n, (width, tw, dd, level, block_num, feature_size) in enumerate(stages, start=1):
            setattr(self, "Upsample_%d" % (n + 1), up(last_planes, width))
            t = Tree(width, in_planes=tw, out_planes=tw, dense_depth=dd, level=level, block_num=block_num,
                     feature_size=feature_size)
            last_planes = t.get_out_planes() + width
            setattr(self, "Tree_%d" % n, nn.Sequential(t))
        self.LastConv = L.Conv2d(last_planes, 3, 9, stride=1, padding=4, padding_mode='reflect')
        self.LastActivate = L.Tanh()

    def forward(self, x):
        x = self.Upsample_1(x)
        for n in range(1, 5):
            x = getattr(self, "Upsample_%d" % (n + 1))(x)
            x = ops.cat_channels([x, getattr(self, "Tree_%d" % n)(x)])
        return self.LastActivate(self.LastConv(x))
