"""generator_13_1 — drop-in for the reference's generators/generator_13_1.py (no factory names it): between generator_12
and generator_13.  generator_12's blocks without the noise and without the normalising mapping networks — a StyleConv is
a modulated convolution (style from its own two-layer mapping network) plus a bias — with a self-attention layer inside
every StyleBlock, per-channel attention gates, PReLU after the ResNet-in-ResNet sums, a learned scale on the block skip,
RGB skips merged by `ToRGB(6)` layers at every tree level, and up-sampling that mixes a transposed convolution with the
bicubic + Smooth path through a 3x3 StyleConv.  Same module tree, construction order and RNG consumption as the
reference file."""
from typing import Tuple

import torch
import torch.nn as nn

from gan_b200 import kernels as KN
from gan_b200 import layers as L
from gan_b200 import ops
from generators.generator_12 import Smooth, UpSample_rgb, EqualizedWeight  # noqa: F401 (identical classes)
from generators.generator_13 import EqualizedLinear, MappingNetwork        # noqa: F401 (identical classes)


class Conv2dWeightModulate(nn.Module):
    """reference :69-96"""

    def __init__(self, d_latent: int, in_planes: int, out_planes: int, kernel_size: int, demodulate: bool = True,
                 eps: float = 1e-8):
        super(Conv2dWeightModulate, self).__init__()
        self.to_style = nn.Sequential(MappingNetwork(d_latent, 2), EqualizedLinear(d_latent, in_planes))
        self.out_planes = out_planes
        self.demodulate = demodulate
        self.padding = (kernel_size - 1) // 2
        self.weight = EqualizedWeight([out_planes, in_planes, kernel_size, kernel_size])
        self.eps = eps

    def forward(self, x: torch.Tensor, s: torch.Tensor):
        return ops.modulated_conv2d(x, self.weight.raw, self.to_style(s), self.padding, self.weight.gain,
                                    self.demodulate, self.eps)


class StyleConv(nn.Module):
    """reference :98-107"""

    def __init__(self, d_latent: int, in_planes: int, out_planes: int, kernel_size: int):
        super(StyleConv, self).__init__()
        self.conv = Conv2dWeightModulate(d_latent, in_planes, out_planes, kernel_size=kernel_size)
        self.bias = nn.Parameter(L.normal_(out_planes))

    def forward(self, x: torch.Tensor, w: torch.Tensor):
        return ops.add_bc(self.conv(x, w), self.bias, KN.V_C)


class SelfAttention(nn.Module):
    """reference :109-131 — y = gamma[c] * o + x, gamma ~ U(0.7, 0.8) per channel."""

    def __init__(self, d_latent, in_planes: int, embedding_channels: int):
        super(SelfAttention, self).__init__()
        self.key = StyleConv(d_latent, in_planes, embedding_channels, 1)
        self.query = StyleConv(d_latent, in_planes, embedding_channels, 1)
        self.value = StyleConv(d_latent, in_planes, embedding_channels, 1)
        self.self_att = StyleConv(d_latent, embedding_channels, in_planes, 1)
        self.gamma = nn.Parameter(torch.nn.init.uniform_(torch.empty(in_planes), a=0.7, b=0.8))
        self.softmax = nn.Softmax(dim=1)   # kept for the module tree; the attention runs in L.self_attention

    def forward(self, x: torch.Tensor, w: torch.Tensor):
        o = self.self_att(L.self_attention(self.key(x, w), self.query(x, w), self.value(x, w)), w)
        return ops.fma_bc(x, self.gamma, o, KN.V_C)


class StyleBlock(nn.Module):
    """reference :134-153"""

    def __init__(self, d_latent: int, last_planes: int, in_planes: int, out_planes: int, dense_depth: int,
                 kernel_size: int):
        super(StyleBlock, self).__init__()
        self.conv1 = StyleConv(d_latent, last_planes, in_planes, kernel_size=1)
        self.activation1 = L.PReLU(in_planes)
        self.attention1 = SelfAttention(d_latent, in_planes, in_planes)
        self.conv2 = StyleConv(d_latent, in_planes, in_planes, kernel_size=kernel_size)
        self.activation2 = L.PReLU(in_planes)
        self.conv3 = StyleConv(d_latent, in_planes, out_planes + dense_depth, kernel_size=kernel_size)

    def forward(self, x: torch.Tensor, w: torch.Tensor):
        x = self.attention1(self.activation1(self.conv1(x, w)), w)
        return self.conv3(self.activation2(self.conv2(x, w)), w)


class SEBlock(nn.Module):
    """reference :156-170"""

    def __init__(self, d_latent: int, in_planes: int, out_planes: int, dense_depth: int):
        super(SEBlock, self).__init__()
        self.conv1 = StyleConv(d_latent, out_planes + dense_depth, in_planes, kernel_size=1)
        self.activation1 = L.PReLU(in_planes)
        self.conv2 = StyleConv(d_latent, in_planes, out_planes + dense_depth, kernel_size=1)
        self.activation2 = L.Sigmoid()

    def forward(self, x: torch.Tensor, w: torch.Tensor):
        b, c = x.shape[:2]
        x = ops.global_avg_pool(x).reshape(b, c, 1, 1)
        return self.activation2(self.conv2(self.activation1(self.conv1(x, w)), w))


class SEStyleBlock(nn.Module):
    """reference :173-185"""

    def __init__(self, d_latent: int, last_planes: int, in_planes: int, out_planes: int, dense_depth: int,
                 kernel_size: int):
        super(SEStyleBlock, self).__init__()
        self.out_planes = out_planes
        self.dense_depth = dense_depth
        self.convs = StyleBlock(d_latent, last_planes, in_planes, out_planes, dense_depth, kernel_size)
        self.se = SEBlock(d_latent, in_planes, out_planes, dense_depth)

    def forward(self, x: torch.Tensor, w: torch.Tensor):
        x = self.convs(x, w)
        return L.gate(x, self.se(x, w))


class ResnetInit(nn.Module):
    """reference :188-211"""

    def __init__(self, d_latent, last_planes, in_planes, out_planes, dense_depth, kernel_size):
        super(ResnetInit, self).__init__()
        mk = lambda dd: SEStyleBlock(d_latent, last_planes, in_planes, out_planes, dd, kernel_size)
        self.residual = mk(dense_depth)
        self.transient = mk(0)
        self.residual_across = mk(0)
        self.transient_across = mk(dense_depth)
        self.activation_residual = L.PReLU(out_planes + dense_depth)
        self.activation_transient = L.PReLU(out_planes)

    def forward(self, x: Tuple[torch.Tensor, torch.Tensor], w: torch.Tensor):
        x_residual, x_transient = x
        res = ops.add(self.residual(x_residual, w), self.transient_across(x_transient, w))
        tra = ops.add(self.residual_across(x_residual, w), self.transient(x_transient, w))
        return self.activation_residual(res), self.activation_transient(tra)


class ToRGB(nn.Module):
    """reference :213-224"""

    def __init__(self, d_latent: int, planes: int):
        super().__init__()
        self.attention = SelfAttention(d_latent, planes, planes)
        self.conv = Conv2dWeightModulate(d_latent, planes, 3, kernel_size=1, demodulate=False)
        self.bias = nn.Parameter(L.normal_(3))
        self.activation = L.PReLU(3)

    def forward(self, x: torch.Tensor, w: torch.Tensor):
        return self.activation(ops.add_bc(self.conv(self.attention(x, w), w), self.bias, KN.V_C))


class BasicBlock(nn.Module):
    """reference :226-272"""

    def get_out_planes(self):
        if self.is_unify or self.root:
            return 2 * self.out_planes + 2 * self.dense_depth
        return self.last_planes + 1 * self.dense_depth

    def __init__(self, d_latent: int, last_planes: int, in_planes: int, out_planes: int, dense_depth: int, root: bool,
                 is_unify: bool):
        super(BasicBlock, self).__init__()
        self.root = root
        self.last_planes = last_planes
        self.out_planes = out_planes
        self.dense_depth = dense_depth
        self.scale = nn.Parameter(torch.nn.init.uniform_(torch.empty(out_planes), a=0.3, b=0.4))
        self.is_unify = is_unify
        stream = last_planes - out_planes
        if is_unify:
            self.unify = StyleConv(d_latent, last_planes, 2 * out_planes + dense_depth, kernel_size=1)
            stream = out_planes + dense_depth
        self.rir_3 = ResnetInit(d_latent, stream, in_planes, out_planes, dense_depth, 3)
        if root:
            self.shortcut = StyleConv(d_latent, last_planes, 2 * out_planes + dense_depth, kernel_size=1)

    def forward(self, x: torch.Tensor, w: torch.Tensor):
        d = self.out_planes
        if self.is_unify:
            x = self.unify(x, w)
        x_residual = ops.cat_channels([ops.slice_channels(x, 0, d), ops.slice_channels(x, 2 * d)])
        res3, tra3 = self.rir_3((x_residual, ops.slice_channels(x, d)), w)
        if self.root:
            x = self.shortcut(x, w)
        head = ops.fma_bc(ops.slice_channels(x, 0, d), self.scale, ops.slice_channels(res3, 0, d), KN.V_C)
        return ops.cat_channels([head, tra3, ops.slice_channels(x, 2 * d), ops.slice_channels(res3, d)])


class Tree(nn.Module):
    """reference :275-345 — every sub-tree's RGB is merged into `rgb` by its own ToRGB(6) (`mix_i`), the tree's own by
    `get_rgb`."""

    def get_out_planes(self):
        return self.root.get_out_planes()

    def __init__(self, d_latent: int, last_planes: int, in_planes: int, out_planes: int, dense_depth: int, level: int,
                 block_num: int):
        super(Tree, self).__init__()
        assert (block_num > 0)
        self.level = level
        self.block_num = block_num
        self.out_planes = out_planes
        self.dense_depth = dense_depth
        leaf = lambda last, unify: BasicBlock(d_latent, last, in_planes, out_planes, dense_depth, False, unify)
        width = 2 * out_planes * (block_num - 1)
        first_leaf = 0
        if level == 1:
            blk = leaf(last_planes, last_planes < 2 * out_planes)
            last_planes = blk.get_out_planes()
            self.__setattr__('block_%d' % 0, blk)
            first_leaf = 1
        else:
            self.prev_root = leaf(last_planes, last_planes < 2 * out_planes)
            width += self.prev_root.get_out_planes()
            for i in reversed(range(1, level)):
                sub = Tree(d_latent, last_planes, in_planes, out_planes, dense_depth, i, block_num)
                last_planes = sub.get_out_planes()
                width += last_planes
                self.__setattr__('level_%d' % i, sub)
                self.__setattr__('mix_%d' % i, ToRGB(d_latent, 6))
        for i in range(first_leaf, block_num):
            blk = leaf(last_planes, False)
            last_planes = blk.get_out_planes()
            self.__setattr__('block_%d' % i, blk)
        width += blk.get_out_planes()
        self.root_last_planes = width
        self.root = BasicBlock(d_latent, width, in_planes * block_num, out_planes, dense_depth, True, False)
        self.to_rgb = ToRGB(d_latent, self.get_out_planes())
        self.get_rgb = ToRGB(d_latent, 6)

    def forward(self, x: torch.Tensor, w: torch.Tensor, rgb: torch.Tensor):
        d = self.out_planes
        parts = [self.prev_root(x, w)] if self.level > 1 else []
        for i in reversed(range(1, self.level)):
            x, rgb_new = self.__getattr__('level_%d' % i)(x, w, rgb)
            rgb = self.__getattr__('mix_%d' % i)(ops.cat_channels([rgb, rgb_new]), w)
            parts.append(x)
        for i in range(self.block_num):
            x = self.__getattr__('block_%d' % i)(x, w)
            parts.append(ops.slice_channels(x, 0, 2 * d))
        parts.append(ops.slice_channels(x, 2 * d))
        out = self.root(ops.cat_channels(parts), w)
        return out, self.get_rgb(ops.cat_channels([rgb, self.to_rgb(out, w)]), w)


class UpSample(nn.Module):
    """reference :347-385 — (attention,) ConvTranspose2d + PReLU (+ Smooth), optionally mixed with bicubic x2 (+ Smooth)
    of the input (1x1 StyleConv `unify` if the widths differ) by a 3x3 StyleConv + PReLU."""

    def __init__(self, d_latent: int, planes: int, out_planes: int, kernel_size: int, stride: int, padding: int,
                 use_attention: bool = True, use_bic: bool = True, use_smooth: bool = True):
        super(UpSample, self).__init__()
        self.use_bic = use_bic
        if use_attention:
            self.attention = SelfAttention(d_latent, planes, planes)
        self.use_attention = use_attention
        self.convT = L.ConvTranspose2d(planes, out_planes, kernel_size=kernel_size, stride=stride, padding=padding)
        self.activation_convT = L.PReLU(out_planes)
        if use_bic:
            self.up_sample = L.BicubicUp2()
            self.need_unify = False
            if planes != out_planes:
                self.need_unify = True
                self.unify = StyleConv(d_latent, planes, out_planes, 1)
            self.mix = StyleConv(d_latent, out_planes * 2, out_planes, 3)
            self.activation_mix = L.PReLU(out_planes)
        self.use_smooth = use_smooth
        if use_smooth:
            self.smooth = Smooth()

    def forward(self, x: torch.Tensor, w: torch.Tensor):
        if self.use_attention:
            x = self.attention(x, w)
        out = self.activation_convT(self.convT(x))
        if self.use_smooth:
            out = self.smooth(out)
        if self.use_bic:
            x = ops.bicubic_up2_smooth(x) if self.use_smooth else self.up_sample(x)
            if self.need_unify:
                x = self.unify(x, w)
            out = self.activation_mix(self.mix(ops.cat_channels([x, out]), w))
        return out


class GeneratorBlock(nn.Module):
    """reference :387-407"""

    def get_out_planes(self):
        return self.tree.get_out_planes()

    def __init__(self, d_latent: int, last_planes: int, in_planes: int, out_planes: int, dense_depth: int, level: int,
                 block_num: int, kernel_size: int = 4, stride: int = 2, padding: int = 1):
        super(GeneratorBlock, self).__init__()
        self.upsample = UpSample(d_latent, planes=last_planes, out_planes=last_planes, kernel_size=kernel_size,
                                 stride=stride, padding=padding)
        self.tree = Tree(d_latent, last_planes, in_planes, out_planes, dense_depth, level, block_num)
        self.upsample_rgb = UpSample_rgb()
        self.get_rgb = ToRGB(d_latent, 6)

    def forward(self, x: torch.Tensor, w: torch.Tensor, rgb: torch.Tensor):
        rgb = self.upsample_rgb(rgb)
        x, rgb_new = self.tree(self.upsample(x, w), w, rgb)
        return x, self.get_rgb(ops.cat_channels([rgb, rgb_new]), w)


class Generator(nn.Module):
    """reference :409-439"""

    def __init__(self, z_dim, planes=64):
        super(Generator, self).__init__()
        self.mapping_network = MappingNetwork(z_dim, 8)
        self.upsample1 = UpSample(z_dim, planes=z_dim, out_planes=planes * 8, kernel_size=4, stride=1, padding=0,
                                  use_attention=False, use_bic=False, use_smooth=False)
        self.style1 = SEStyleBlock(z_dim, planes * 8, planes * 8, planes * 8, 0, 3)
        self.activation1 = L.PReLU(planes * 8)
        self.to_rgb1 = ToRGB(z_dim, planes * 8)
        last = planes * 8
        for idx, mult in enumerate([4, 2, 1, 1], start=1):
            blk = GeneratorBlock(z_dim, last, planes * mult, planes * mult, planes // 8, 2, 2)
            last = blk.get_out_planes()
            self.__setattr__('block%d' % idx, blk)

    def forward(self, x: torch.Tensor):
        w = self.mapping_network(torch.squeeze(x))
        x = self.activation1(self.style1(self.upsample1(x, w), w))
        rgb = self.to_rgb1(x, w)
        for idx in range(1, 5):
            x, rgb = self.__getattr__('block%d' % idx)(x, w, rgb)
        return rgb
