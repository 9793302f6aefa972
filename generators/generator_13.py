"""generator_13 — drop-in for the reference's generators/generator_13.py (units/get_generators.get_13; generator_13_2 is
the same file): the first style-based ancestor of generator_13_5.  A mapping network z -> w; every convolution of the
aggregation tree is a modulated convolution whose style comes from ITS OWN two-layer mapping network on w; per-channel
scaled N(0,1) noise and a bias after it; no BatchNorm anywhere; SK merges of the ResNet-in-ResNet streams and of the RGB
skips, a learned per-channel scale on the block skip, and SAGAN self-attention (3x3 convolutions, SE gate) on both streams
of every block.

Class names, constructor signatures, attribute names and construction order follow the reference file (D-style equalised
layers here: `weight.weight`, `weight.c`).  The arithmetic is generator_13_5's: the modulated convolution runs as one
implicit GEMM on the shared weight (style on the input channels inside the gather, demodulation in the epilogue), SK / SE
products are the fused broadcast kernels, attention is the batched matrix product + row softmax of csrc/attn.cu."""
from typing import Tuple, List

import torch
import torch.nn as nn

from gan_b200 import kernels as KN
from gan_b200 import layers as L
from gan_b200 import ops
from gan_b200 import rng


class Smooth(L.SmoothBase):
    """reference :8-24"""


class EqualizedWeight(L.EqWeightBase):
    """reference :27-35 — parameter `weight`, constant `c`."""
    PARAM, GAIN = "weight", "c"

    def __init__(self, shape: List[int]):
        super(EqualizedWeight, self).__init__(shape)


class EqualizedLinear(L.EqLinearBase):
    """reference :38-45"""
    WEIGHT = EqualizedWeight

    def __init__(self, in_planes: int, out_planes: int):
        super(EqualizedLinear, self).__init__()
        self._build(in_planes, out_planes)


class EqualizedConv2d(L.EqConv2dBase):
    """reference :48-57"""
    WEIGHT = EqualizedWeight

    def __init__(self, in_planes: int, out_planes: int, kernel_size: int):
        super().__init__()
        self._build(in_planes, out_planes, kernel_size, (kernel_size - 1) // 2)


class MappingNetwork(nn.Module):
    """reference :60-70 — n x (EqualizedLinear, PReLU)."""

    def __init__(self, planes: int, n_layers: int):
        super(MappingNetwork, self).__init__()
        layers = []
        for i in range(n_layers):
            layers.append(EqualizedLinear(planes, planes))
            layers.append(L.PReLU(planes))
        self.net = nn.Sequential(*layers)

    def forward(self, z: torch.Tensor):
        return self.net(z)


def _conv_act(planes, n):
    mods = []
    for _ in range(n):
        mods += [EqualizedConv2d(planes, planes, 3), L.PReLU(planes)]
    return mods


class _SKAttention(nn.Module):
    def _build_heads(self, planes, m):
        self.M = m
        for i in range(m):
            self.__setattr__('fc_sub_%d' % i, nn.Sequential(MappingNetwork(planes, 2), EqualizedLinear(planes, planes)))
        self.softmax = nn.Softmax(dim=1)

    def _heads(self, fea_z, b, s, c):
        subs = [self.__getattr__('fc_sub_%d' % i) for i in range(self.M)]
        return L.sk_attention_vectors(fea_z, subs).view(b, s, c, 1, 1)


class SKAttention_conv(_SKAttention):
    """reference :73-113"""

    def __init__(self, planes: int, m: int):
        super(SKAttention_conv, self).__init__()
        self.gap_conv = L.AdaptiveAvgPool2d(5)
        self.conv_main = nn.Sequential(*_conv_act(planes, 3))
        self.gap_fc = L.AdaptiveAvgPool2d(1)
        self.fc_main = MappingNetwork(planes, 2)
        self._build_heads(planes, m)

    def forward(self, feas):
        branches = L.as_branches(feas)
        b, c, _, side = branches[0].shape
        assert (side >= 8)
        fea_s = self.conv_main(ops.pooled_branch_sum(branches, 5))
        return self._heads(self.fc_main(self.gap_fc(fea_s).view(b, c)), b, len(branches), c)


class SKAttention_fc(_SKAttention):
    """reference :116-147"""

    def __init__(self, planes: int, m: int):
        super(SKAttention_fc, self).__init__()
        self.gap_fc = L.AdaptiveAvgPool2d(1)
        self.fc_main = MappingNetwork(planes, 4)
        self._build_heads(planes, m)

    def forward(self, feas):
        branches = L.as_branches(feas)
        b, c = branches[0].shape[:2]
        return self._heads(self.fc_main(self.gap_fc(L.branch_sum(branches)).view(b, c)), b, len(branches), c)


def _pick_sk_attention(planes, m, image_size):
    return SKAttention_conv(planes, m) if image_size > 4 else SKAttention_fc(planes, m)


class SKConvT(nn.Module):
    """reference :150-166 — ConvTranspose2d x2 + PReLU and bicubic x2 + Smooth, merged by SK attention."""

    def __init__(self, planes: int):
        super(SKConvT, self).__init__()
        self.convT = L.ConvTranspose2d(planes, planes, kernel_size=4, stride=2, padding=1)
        self.activation_convT = L.PReLU(planes)
        self.up_sample = L.BicubicUp2()
        self.smooth = Smooth()
        self.sk_attention = SKAttention_conv(planes, 2)

    def forward(self, x: torch.Tensor):
        branches = [self.activation_convT(self.convT(x)), ops.bicubic_up2_smooth(x)]
        return L.sk_merge(branches, self.sk_attention(branches))


class Conv2dWeightModulate(nn.Module):
    """reference :169-197 — the style is `to_style(w)`: this layer's own mapping network and a linear to `in_planes`."""

    def __init__(self, d_latent: int, in_planes: int, out_planes: int, kernel_size: int, demodulate: bool = True,
                 eps: float = 1e-8):
        super(Conv2dWeightModulate, self).__init__()
        self.to_style = nn.Sequential(MappingNetwork(d_latent, 2), EqualizedLinear(d_latent, in_planes))
        self.out_planes = out_planes
        self.demodulate = demodulate
        self.padding = (kernel_size - 1) // 2
        self.pad = nn.ReplicationPad2d(self.padding)
        self.weight = EqualizedWeight([out_planes, in_planes, kernel_size, kernel_size])
        self.eps = eps

    def forward(self, x: torch.Tensor, s: torch.Tensor):
        return ops.modulated_conv2d(x, self.weight.raw, self.to_style(s), self.padding, self.weight.gain,
                                    self.demodulate, self.eps)


class StyleConv(nn.Module):
    """reference :200-214 — modulated conv, optional per-channel-scaled noise (one N(0,1) value per output element), bias."""

    def __init__(self, d_latent: int, in_planes: int, out_planes: int, kernel_size: int, use_noise: bool = False):
        super(StyleConv, self).__init__()
        self.conv = Conv2dWeightModulate(d_latent, in_planes, out_planes, kernel_size=kernel_size)
        self.use_noise = use_noise
        if use_noise:
            self.scale_noise = nn.Parameter(torch.nn.init.uniform_(torch.empty(out_planes), a=0.2, b=0.3))
        self.bias = nn.Parameter(L.normal_(out_planes))
        self.noise_layer = 0   # unique per StyleConv (Generator.__init__): keys the counter-based noise on the GPU

    def forward(self, x: torch.Tensor, w: torch.Tensor):
        y = self.conv(x, w)
        if self.use_noise:
            if rng.philox_noise():
                z = KN.backend().noise_fill(tuple(y.shape), rng.current(), self.noise_layer, y.device).to(y.dtype)
            else:   # the reference's own draw: torch.randn of this shape on the same generator
                z = rng.randn(*y.shape, device=y.device).to(y.dtype)
            y = ops.noise_add(y, self.scale_noise, z)
        return ops.add_bc(y, self.bias, KN.V_C)


class SKConv(nn.Module):
    """reference :217-241"""

    def __init__(self, d_latent: int, in_planes: int, out_planes: int, m: int):
        super(SKConv, self).__init__()
        assert (m > 0)
        self.M = m
        for i in range(m):
            self.__setattr__('conv_%d' % i, StyleConv(d_latent, in_planes, out_planes, kernel_size=3 + i * 2))
            self.__setattr__('nonlinear_%d' % i, L.PReLU(out_planes))
        self.sk_attention = SKAttention_conv(out_planes, m)

    def forward(self, x: torch.Tensor, w: torch.Tensor):
        branches = [self.__getattr__('nonlinear_%d' % i)(self.__getattr__('conv_%d' % i)(x, w)) for i in range(self.M)]
        return L.sk_merge(branches, self.sk_attention(branches))


class StyleBlock(nn.Module):
    """reference :244-269"""

    def __init__(self, d_latent: int, last_planes: int, in_planes: int, out_planes: int, dense_depth: int,
                 kernel_size: int, m: int):
        super(StyleBlock, self).__init__()
        assert (m > 0)
        self.conv1 = StyleConv(d_latent, last_planes, in_planes, kernel_size=1)
        self.activation1 = L.PReLU(in_planes)
        self.m = m
        if m == 1:
            self.conv2 = StyleConv(d_latent, in_planes, in_planes, kernel_size=kernel_size)
            self.activation2 = L.PReLU(in_planes)
        else:
            self.skconv = SKConv(d_latent, in_planes, in_planes, m)
        self.conv3 = StyleConv(d_latent, in_planes, out_planes + dense_depth, kernel_size=kernel_size, use_noise=False)
        self.activation3 = L.PReLU(out_planes + dense_depth)

    def forward(self, x: torch.Tensor, w: torch.Tensor):
        x = self.activation1(self.conv1(x, w))
        x = self.activation2(self.conv2(x, w)) if self.m == 1 else self.skconv(x, w)
        return self.activation3(self.conv3(x, w))


class SEBlock_conv(nn.Module):
    """reference :272-296"""

    def __init__(self, in_planes: int):
        super(SEBlock_conv, self).__init__()
        self.gap_conv = L.AdaptiveAvgPool2d(5)
        self.convs = nn.Sequential(*_conv_act(in_planes, 3))
        self.gap_fc = L.AdaptiveAvgPool2d(1)
        self.fcs = MappingNetwork(in_planes, 2)
        self.fc_out = EqualizedLinear(in_planes, in_planes)
        self.activation2 = L.Sigmoid()

    def forward(self, x: torch.Tensor):
        b, c, _, side = x.shape
        assert (side >= 8)
        t = self.gap_fc(self.convs(self.gap_conv(x))).view(b, c)
        return self.activation2(self.fc_out(self.fcs(t))).view(b, c, 1, 1)


class SEBlock_fc(nn.Module):
    """reference :298-311"""

    def __init__(self, in_planes: int):
        super(SEBlock_fc, self).__init__()
        self.gap = L.AdaptiveAvgPool2d(1)
        self.fcs = MappingNetwork(in_planes, 2)
        self.fc_out = EqualizedLinear(in_planes, in_planes)
        self.activation2 = L.Sigmoid()

    def forward(self, x: torch.Tensor):
        b, c = x.shape[:2]
        return self.activation2(self.fc_out(self.fcs(self.gap(x).view(b, c)))).view(b, c, 1, 1)


class SelfAttention(nn.Module):
    """reference :314-341 — plain (un-modulated) 3x3 query / key / value / output convolutions, SE block as gamma."""

    def __init__(self, in_planes: int, embedding_channels: int, image_size: int):
        super(SelfAttention, self).__init__()
        self.query = EqualizedConv2d(in_planes, embedding_channels, 3)
        self.key = EqualizedConv2d(in_planes, embedding_channels, 3)
        self.value = EqualizedConv2d(in_planes, embedding_channels, 3)
        self.self_att = EqualizedConv2d(embedding_channels, in_planes, 3)
        self.gamma = SEBlock_conv(in_planes) if image_size > 4 else SEBlock_fc(in_planes)
        self.softmax = nn.Softmax(dim=1)   # kept for the module tree; the attention runs in L.self_attention

    def forward(self, x: torch.Tensor):
        o = self.self_att(L.self_attention(self.query(x), self.key(x), self.value(x)))
        return L.gated_residual(x, o, self.gamma(o))


class ResnetInit(nn.Module):
    """reference :344-373"""

    def __init__(self, d_latent: int, last_planes: int, in_planes: int, out_planes: int, dense_depth: int,
                 kernel_size: int, m: int, image_size: int):
        super(ResnetInit, self).__init__()
        mk = lambda dd: StyleBlock(d_latent, last_planes, in_planes, out_planes, dd, kernel_size, m)
        self.residual = mk(dense_depth)
        self.transient = mk(0)
        self.residual_across = mk(0)
        self.transient_across = mk(dense_depth)
        self.sk_attention_residual = _pick_sk_attention(out_planes + dense_depth, 2, image_size)
        self.sk_attention_transient = _pick_sk_attention(out_planes, 2, image_size)

    def forward(self, x: Tuple[torch.Tensor, torch.Tensor], w: torch.Tensor):
        x_residual, x_transient = x
        # the reference's call order (it fixes the order of the noise draws): r_r, r_t, t_t, t_r
        r_r = self.residual(x_residual, w)
        r_t = self.residual_across(x_residual, w)
        t_t = self.transient(x_transient, w)
        t_r = self.transient_across(x_transient, w)
        feas_residual, feas_transient = [r_r, t_r], [r_t, t_t]
        return (L.sk_merge(feas_residual, self.sk_attention_residual(feas_residual)),
                L.sk_merge(feas_transient, self.sk_attention_transient(feas_transient)))


class BasicBlock(nn.Module):
    """reference :376-428"""

    def get_out_planes(self):
        if self.is_unify or self.root:
            return 2 * self.out_planes + 2 * self.dense_depth
        return self.last_planes + 1 * self.dense_depth

    def __init__(self, d_latent: int, last_planes: int, in_planes: int, out_planes: int, dense_depth: int, root: bool,
                 is_unify: bool, m: int, image_size: int):
        super(BasicBlock, self).__init__()
        self.root = root
        self.last_planes = last_planes
        self.out_planes = out_planes
        self.dense_depth = dense_depth
        self.scale = nn.Parameter(torch.nn.init.uniform_(torch.empty(out_planes), a=0.25, b=0.35))
        self.is_unify = is_unify
        stream = last_planes - out_planes
        if is_unify:
            self.unify = StyleConv(d_latent, last_planes, 2 * out_planes + dense_depth, kernel_size=1)
            stream = out_planes + dense_depth
        self.rir_3 = ResnetInit(d_latent, stream, in_planes, out_planes, dense_depth, 3, m, image_size)
        if root:
            self.shortcut = StyleConv(d_latent, last_planes, 2 * out_planes + dense_depth, kernel_size=1)
        self.attention_residual = SelfAttention(out_planes + dense_depth, out_planes + dense_depth, image_size)
        self.attention_transient = SelfAttention(out_planes, out_planes, image_size)

    def forward(self, x: torch.Tensor, w: torch.Tensor):
        d = self.out_planes
        if self.is_unify:
            x = self.unify(x, w)
        x_residual = ops.cat_channels([ops.slice_channels(x, 0, d), ops.slice_channels(x, 2 * d)])
        res3, tra3 = self.rir_3((x_residual, ops.slice_channels(x, d)), w)
        if self.root:
            x = self.shortcut(x, w)
        head = ops.fma_bc(ops.slice_channels(x, 0, d), self.scale, ops.slice_channels(res3, 0, d), KN.V_C)
        res3 = self.attention_residual(ops.cat_channels([head, ops.slice_channels(res3, d)]))
        tra3 = self.attention_transient(tra3)
        return ops.cat_channels([ops.slice_channels(res3, 0, d), tra3, ops.slice_channels(x, 2 * d),
                                 ops.slice_channels(res3, d)])


class ToRGB(nn.Module):
    """reference :431-452"""

    def __init__(self, d_latent: int, planes: int, m: int):
        super().__init__()
        assert (m > 0)
        self.m = m
        if m == 1:
            self.pre_conv = StyleConv(d_latent, planes, planes, 3)
            self.pre_activation = L.PReLU(planes)
        else:
            self.skconv = SKConv(d_latent, planes, planes, m)
        self.conv = Conv2dWeightModulate(d_latent, planes, 3, kernel_size=1, demodulate=False)
        self.bias = nn.Parameter(L.normal_(3))
        self.activation = L.PReLU(3)

    def forward(self, x: torch.Tensor, w: torch.Tensor):
        x = self.pre_activation(self.pre_conv(x, w)) if self.m == 1 else self.skconv(x, w)
        return self.activation(ops.add_bc(self.conv(x, w), self.bias, KN.V_C))


class Tree(nn.Module):
    """reference :455-528"""

    def get_out_planes(self):
        return self.root.get_out_planes()

    def __init__(self, d_latent: int, last_planes: int, in_planes: int, out_planes: int, dense_depth: int, level: int,
                 block_num: int, m: int, image_size: int):
        super(Tree, self).__init__()
        assert (block_num > 0)
        self.level = level
        self.block_num = block_num
        self.out_planes = out_planes
        self.dense_depth = dense_depth
        leaf = lambda last, unify: BasicBlock(d_latent, last, in_planes, out_planes, dense_depth, False, unify, m,
                                              image_size)
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
                sub = Tree(d_latent, last_planes, in_planes, out_planes, dense_depth, i, block_num, m, image_size)
                last_planes = sub.get_out_planes()
                width += last_planes
                self.__setattr__('level_%d' % i, sub)
        for i in range(first_leaf, block_num):
            blk = leaf(last_planes, False)
            last_planes = blk.get_out_planes()
            self.__setattr__('block_%d' % i, blk)
        width += blk.get_out_planes()
        self.root_last_planes = width
        self.root = BasicBlock(d_latent, width, in_planes * block_num, out_planes, dense_depth, True, False, m,
                               image_size)
        self.to_rgb = ToRGB(d_latent, self.get_out_planes(), m)
        self.sk_attention = _pick_sk_attention(3, 2, image_size)

    def forward(self, x: torch.Tensor, w: torch.Tensor, rgb: torch.Tensor):
        d = self.out_planes
        parts = [self.prev_root(x, w)] if self.level > 1 else []
        for i in reversed(range(1, self.level)):
            x, rgb = self.__getattr__('level_%d' % i)(x, w, rgb)
            parts.append(x)
        for i in range(self.block_num):
            x = self.__getattr__('block_%d' % i)(x, w)
            parts.append(ops.slice_channels(x, 0, 2 * d))
        parts.append(ops.slice_channels(x, 2 * d))
        out = self.root(ops.cat_channels(parts), w)
        branches = [rgb, self.to_rgb(out, w)]
        return out, L.sk_merge(branches, self.sk_attention(branches))


class GeneratorBlock(nn.Module):
    """reference :531-547"""

    def get_out_planes(self):
        return self.tree.get_out_planes()

    def __init__(self, d_latent: int, last_planes: int, in_planes: int, out_planes: int, dense_depth: int, level: int,
                 block_num: int, m: int, image_size: int):
        super(GeneratorBlock, self).__init__()
        self.upsample = SKConvT(last_planes)
        self.tree = Tree(d_latent, last_planes, in_planes, out_planes, dense_depth, level, block_num, m, image_size)
        self.upsample_rgb = SKConvT(3)

    def forward(self, x: torch.Tensor, w: torch.Tensor, rgb: torch.Tensor):
        rgb = self.upsample_rgb(rgb)
        return self.tree(self.upsample(x), w, rgb)


class GeneratorStart(nn.Module):
    """reference :550-569"""

    def get_out_planes(self):
        return self.tree.get_out_planes()

    def __init__(self, z_dim: int, mapping_layer: int, in_planes: int, out_planes: int, dense_depth: int, level: int,
                 block_num: int, m: int):
        super(GeneratorStart, self).__init__()
        self.mapping_network = MappingNetwork(z_dim, mapping_layer)
        self.convT = L.ConvTranspose2d(z_dim, out_planes, kernel_size=4, stride=1, padding=0)
        self.activation = L.PReLU(out_planes)
        self.to_rgb = ToRGB(z_dim, out_planes, m)
        self.tree = Tree(z_dim, out_planes, in_planes, out_planes // 2, dense_depth, level, block_num, m, 4)

    def forward(self, x: torch.Tensor):
        w = self.mapping_network(torch.squeeze(x))   # all singleton dims, as the reference (a batch of 1 fails there too)
        x = self.activation(self.convT(x))
        rgb = self.to_rgb(x, w)
        x, rgb = self.tree(x, w, rgb)
        return x, w, rgb


class Generator(nn.Module):
    """reference :572-593"""

    def __init__(self, z_dim, planes=56):
        super(Generator, self).__init__()
        dd = planes // 8
        self.block0 = GeneratorStart(z_dim, 8, planes * 8, planes * 8, dd, 1, 2, 1)
        last = self.block0.get_out_planes()
        for idx, (mult, size) in enumerate([(4, 8), (2, 16), (1, 32), (1, 64)], start=1):
            blk = GeneratorBlock(z_dim, last, planes * mult, planes * mult, dd, 2, 2, 2, size)
            last = blk.get_out_planes()
            self.__setattr__('block%d' % idx, blk)
        for i, m in enumerate(mod for mod in self.modules() if isinstance(mod, StyleConv)):
            m.noise_layer = i + 1

    def forward(self, x: torch.Tensor):
        if rng.philox_noise():
            rng.begin_forward(x.device)
        x, w, rgb = self.block0(x)
        for idx in range(1, 5):
            x, rgb = self.__getattr__('block%d' % idx)(x, w, rgb)
        return rgb
