"""generator_12 — drop-in for the reference's generators/generator_12.py (units/get_generators.get_12): the first
StyleGAN2-shaped member of the family.  z is L2-normalised and mapped to w by eight (linear, PReLU) layers; the network
starts from a learned 4x4 constant; every convolution is a `StyleConv`: a modulated convolution whose style is
`to_style(w)` (its own normalising two-layer mapping network and a linear with bias 1), plus one N(0,1) noise plane per
sample scaled by a learned scalar, plus a bias.  ResNet-in-ResNet pairs of SE-gated style blocks inside aggregation
trees, SAGAN self-attention built from StyleConvs in front of every block, ToRGB and transposed convolution, RGB skips
mixed by `ToRGB(6)` layers, tanh.

Class names, constructor signatures, attribute names and construction order follow the reference file; the ORDER OF THE
NOISE DRAWS in forward does too (key, query, value, self_att; residual, residual_across, transient, transient_across ...),
so that the same generator state gives the same image.  The modulated convolution is generator_13_5's shared-weight form
(gan_b200.ops.modulated_conv2d); F.normalize is x * rsqrt(sum x^2 + 1e-24) (equal to x / max(|x|, 1e-12) unless
|x| < 1e-6, which a latent vector is not)."""
from typing import Tuple, List

import torch
import torch.nn as nn

from gan_b200 import kernels as KN
from gan_b200 import layers as L
from gan_b200 import ops
from gan_b200 import rng


def _normalize_rows(z: torch.Tensor):
    """nn.functional.normalize(z, dim=1) for z:(B, D)"""
    return ops.mul_bc(z, ops.rsqrt_eps(ops.sumsq_per_sample(z), 1e-24), KN.V_N)


class Smooth(L.SmoothBase):
    """reference :20-36"""


class UpSample_rgb(nn.Module):
    """reference :39-46 — bicubic x2 then Smooth, applied as one separable stencil."""

    def __init__(self):
        super(UpSample_rgb, self).__init__()
        self.up_sample = L.BicubicUp2()
        self.smooth = Smooth()

    def forward(self, x: torch.Tensor):
        return ops.bicubic_up2_smooth(x)


class EqualizedWeight(L.EqWeightBase):
    """reference :49-57"""
    PARAM, GAIN = "weight", "c"

    def __init__(self, shape: List[int]):
        super(EqualizedWeight, self).__init__(shape)


class EqualizedLinear(L.EqLinearBase):
    """reference :60-67 — bias = N(0,1) * `bias` (0 by default: the draw is made, the value is zero)."""
    WEIGHT = EqualizedWeight

    def __init__(self, in_planes: int, out_planes: int, bias: float = 0.):
        super(EqualizedLinear, self).__init__()
        self._build(in_planes, out_planes, bias_gain=bias)


class MappingNetwork(nn.Module):
    """reference :70-81 — L2-normalise, then n x (EqualizedLinear, PReLU)."""

    def __init__(self, planes: int, n_layers: int):
        super(MappingNetwork, self).__init__()
        layers = []
        for i in range(n_layers):
            layers.append(EqualizedLinear(planes, planes))
            layers.append(L.PReLU(planes))
        self.net = nn.Sequential(*layers)

    def forward(self, z: torch.Tensor):
        return self.net(_normalize_rows(z))


class Conv2dWeightModulate(nn.Module):
    """reference :84-107 — takes the style itself (the caller owns `to_style`)."""

    def __init__(self, in_planes: int, out_planes: int, kernel_size: int,
                 demodulate: bool = True, eps: float = 1e-8):
        super(Conv2dWeightModulate, self).__init__()
        self.out_planes = out_planes
        self.demodulate = demodulate
        self.padding = (kernel_size - 1) // 2
        self.weight = EqualizedWeight([out_planes, in_planes, kernel_size, kernel_size])
        self.eps = eps

    def forward(self, x: torch.Tensor, s: torch.Tensor):
        return ops.modulated_conv2d(x, self.weight.raw, s, self.padding, self.weight.gain, self.demodulate, self.eps)


def _to_style(d_latent, planes):
    return nn.Sequential(MappingNetwork(d_latent, 2), EqualizedLinear(d_latent, planes, bias=1.0))


class StyleConv(nn.Module):
    """reference :110-125 — y = modconv(x, to_style(w)) + scale_noise * N(0,1)[b, 1, h, w] + bias."""

    def __init__(self, d_latent: int, in_planes: int, out_planes: int, kernel_size: int):
        super(StyleConv, self).__init__()
        self.to_style = _to_style(d_latent, in_planes)
        self.conv = Conv2dWeightModulate(in_planes, out_planes, kernel_size=kernel_size)
        self.scale_noise = nn.Parameter(L.normal_(1))
        self.bias = nn.Parameter(L.normal_(out_planes))
        self.noise_layer = 0   # unique per StyleConv (Generator.__init__): keys the counter-based noise on the GPU

    def forward(self, x: torch.Tensor, w: torch.Tensor):
        y = self.conv(x, self.to_style(w))
        b, c, hight, width = y.shape
        if rng.philox_noise():
            z = KN.backend().noise_fill((b, 1, hight, width), rng.current(), self.noise_layer, y.device).to(y.dtype)
        else:   # the reference's own draw: torch.randn of this shape on the same generator
            z = rng.randn(b, 1, hight, width, device=y.device).to(y.dtype)
        # one noise plane per sample, the same for every channel: spread over the channels by a 1x1 convolution from
        # one channel with constant unit weights (the conv engine; its adjoint sums the channels for d scale_noise)
        plane = ops.mul_bc(z, self.scale_noise, KN.V_SCALAR)
        spread = ops.conv2d(plane, torch.ones((c, 1, 1, 1), dtype=y.dtype, device=y.device), None, 1, 0, 0)
        return ops.add_bc(ops.add(y, spread), self.bias, KN.V_C)


class SelfAttention(nn.Module):
    """reference :128-150 — StyleConv key / query / value / output, scalar gate."""

    def __init__(self, d_latent: int, in_planes: int, embedding_channels: int):
        super(SelfAttention, self).__init__()
        self.key = StyleConv(d_latent, in_planes, embedding_channels, 1)
        self.query = StyleConv(d_latent, in_planes, embedding_channels, 1)
        self.value = StyleConv(d_latent, in_planes, embedding_channels, 1)
        self.self_att = StyleConv(d_latent, embedding_channels, in_planes, 1)
        self.gamma = nn.Parameter(torch.nn.init.uniform_(torch.empty(1), a=0.04, b=0.08))
        self.softmax = nn.Softmax(dim=1)   # kept for the module tree; the attention runs in L.self_attention

    def forward(self, x: torch.Tensor, w: torch.Tensor):
        f, g, h = self.key(x, w), self.query(x, w), self.value(x, w)      # (this order: three noise draws)
        o = self.self_att(L.self_attention(f, g, h), w)
        return ops.fma_bc(x, self.gamma, o, KN.V_SCALAR)


class ToRGB(nn.Module):
    """reference :153-169"""

    def __init__(self, d_latent: int, planes: int):
        super().__init__()
        self.to_style = _to_style(d_latent, planes)
        self.attention = SelfAttention(d_latent, planes, planes)
        self.conv = Conv2dWeightModulate(planes, 3, kernel_size=1, demodulate=False)
        self.bias = nn.Parameter(L.normal_(3))
        self.activation = L.PReLU(3)

    def forward(self, x: torch.Tensor, w: torch.Tensor):
        s = self.to_style(w)
        x = self.conv(self.attention(x, w), s)
        return self.activation(ops.add_bc(x, self.bias, KN.V_C))


class UpSample(nn.Module):
    """reference :172-187"""

    def __init__(self, d_latent: int, planes: int, out_planes: int, kernel_size: int, stride: int, padding: int,
                 use_attention: bool = True):
        super(UpSample, self).__init__()
        if use_attention:
            self.attention = SelfAttention(d_latent, planes, planes)
        self.use_attention = use_attention
        self.convT = L.ConvTranspose2d(planes, out_planes, kernel_size=kernel_size, stride=stride, padding=padding)
        self.activation = L.PReLU()

    def forward(self, x: torch.Tensor, w: torch.Tensor):
        if self.use_attention:
            x = self.attention(x, w)
        return self.activation(self.convT(x))


class StyleBlock(nn.Module):
    """reference :190-207"""

    def __init__(self, d_latent: int, last_planes: int, in_planes: int, out_planes: int, dense_depth: int,
                 kernel_size: int):
        super(StyleBlock, self).__init__()
        self.conv1 = StyleConv(d_latent, last_planes, in_planes, kernel_size=1)
        self.activation1 = L.PReLU(in_planes)
        self.conv2 = StyleConv(d_latent, in_planes, in_planes, kernel_size=kernel_size)
        self.activation2 = L.PReLU(in_planes)
        self.conv3 = StyleConv(d_latent, in_planes, out_planes + dense_depth, kernel_size=kernel_size)
        self.activation3 = L.PReLU(out_planes + dense_depth)

    def forward(self, x: torch.Tensor, w: torch.Tensor):
        x = self.activation1(self.conv1(x, w))
        x = self.activation2(self.conv2(x, w))
        return self.activation3(self.conv3(x, w))


class SEBlock(nn.Module):
    """reference :210-224 — global average, two 1x1 StyleConvs (noise included, on the 1x1 map), sigmoid."""

    def __init__(self, d_latent: int, in_planes: int, out_planes: int, dense_depth: int):
        super(SEBlock, self).__init__()
        self.conv1 = StyleConv(d_latent, out_planes + dense_depth, in_planes, kernel_size=1)
        self.activation1 = L.PReLU(in_planes)
        self.conv2 = StyleConv(d_latent, in_planes, out_planes + dense_depth, kernel_size=1)
        self.activation2 = L.Sigmoid()

    def forward(self, x: torch.Tensor, w: torch.Tensor):
        b, c = x.shape[:2]
        x = ops.global_avg_pool(x).reshape(b, c, 1, 1)
        x = self.activation1(self.conv1(x, w))
        return self.activation2(self.conv2(x, w))


class SEStyleBlock(nn.Module):
    """reference :227-238"""

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
    """reference :241-259"""

    def __init__(self, d_latent, last_planes, in_planes, out_planes, dense_depth, kernel_size):
        super(ResnetInit, self).__init__()
        mk = lambda dd: SEStyleBlock(d_latent, last_planes, in_planes, out_planes, dd, kernel_size)
        self.residual = mk(dense_depth)
        self.transient = mk(0)
        self.residual_across = mk(0)
        self.transient_across = mk(dense_depth)

    def forward(self, x: Tuple[torch.Tensor, torch.Tensor], w: torch.Tensor):
        x_residual, x_transient = x
        r_r = self.residual(x_residual, w)              # the reference's call order = the order of the noise draws
        r_t = self.residual_across(x_residual, w)
        t_t = self.transient(x_transient, w)
        t_r = self.transient_across(x_transient, w)
        return ops.add(r_r, t_r), ops.add(r_t, t_t)


class BasicBlock(nn.Module):
    """reference :262-308"""

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
        self.is_unify = is_unify
        width, stream = last_planes, last_planes - out_planes
        if is_unify:
            width, stream = 2 * out_planes + dense_depth, out_planes + dense_depth
            self.unify = StyleConv(d_latent, last_planes, width, kernel_size=1)
        self.attention = SelfAttention(d_latent, width, width)
        self.rir_3 = ResnetInit(d_latent, stream, in_planes, out_planes, dense_depth, 3)
        if root:
            self.shortcut = StyleConv(d_latent, last_planes, 2 * out_planes + dense_depth, kernel_size=1)

    def forward(self, x: torch.Tensor, w: torch.Tensor):
        d = self.out_planes
        if self.is_unify:
            x = self.unify(x, w)
        xa = self.attention(x, w)
        x_residual = ops.cat_channels([ops.slice_channels(xa, 0, d), ops.slice_channels(xa, 2 * d)])
        res3, tra3 = self.rir_3((x_residual, ops.slice_channels(xa, d)), w)
        if self.root:
            x = self.shortcut(x, w)
        return ops.cat_channels([ops.add(ops.slice_channels(x, 0, d), ops.slice_channels(res3, 0, d)), tra3,
                                 ops.slice_channels(x, 2 * d), ops.slice_channels(res3, d)])


class Tree(nn.Module):
    """reference :311-378 — returns the tree's features and ITS OWN ToRGB of them (the caller mixes the RGB skips)."""

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
        for i in range(first_leaf, block_num):
            blk = leaf(last_planes, False)
            last_planes = blk.get_out_planes()
            self.__setattr__('block_%d' % i, blk)
        width += blk.get_out_planes()
        self.root_last_planes = width
        self.root = BasicBlock(d_latent, width, in_planes * block_num, out_planes, dense_depth, True, False)
        self.to_rgb = ToRGB(d_latent, self.get_out_planes())
        self.mix_rgb = ToRGB(d_latent, 6)

    def forward(self, x: torch.Tensor, w: torch.Tensor, rgb: torch.Tensor):
        d = self.out_planes
        parts = [self.prev_root(x, w)] if self.level > 1 else []
        for i in reversed(range(1, self.level)):
            x, rgb_new = self.__getattr__('level_%d' % i)(x, w, rgb)
            rgb = ops.add(rgb, self.mix_rgb(ops.cat_channels([rgb, rgb_new]), w))
            parts.append(x)
        for i in range(self.block_num):
            x = self.__getattr__('block_%d' % i)(x, w)
            parts.append(ops.slice_channels(x, 0, 2 * d))
        parts.append(ops.slice_channels(x, 2 * d))
        out = self.root(ops.cat_channels(parts), w)
        return out, self.to_rgb(out, w)


class Generator(nn.Module):
    """reference :381-442"""

    def __init__(self, z_dim, planes=32):
        super(Generator, self).__init__()
        self.mapping_network = MappingNetwork(z_dim, 8)
        self.upsample1 = UpSample(z_dim, planes=z_dim, out_planes=planes * 16, kernel_size=4, stride=1, padding=0,
                                  use_attention=False)
        self.initial_constant = nn.Parameter(L.normal_((1, planes * 16, 4, 4)))
        self.style1 = SEStyleBlock(z_dim, planes * 16, planes * 8, planes * 16, 0, 3)
        self.activation1 = L.PReLU(planes * 16)
        self.to_rgb1 = ToRGB(z_dim, planes * 16)
        self.upsample2 = UpSample(z_dim, planes=planes * 16, out_planes=planes * 8, kernel_size=4, stride=2, padding=1)
        plans = [(planes * 8, planes * 2, planes * 4, int(planes * 0.25), planes * 4),
                 (planes * 4, planes * 1, planes * 2, int(planes * 0.125), planes * 2),
                 (planes * 2, planes * 1, planes * 1, int(planes * 0.125), planes * 1),
                 (planes * 1, int(planes * 0.5), int(planes * 0.5), int(planes * 0.125), None)]
        for n, (last, inp, out, dd, up_out) in enumerate(plans, start=1):
            tree = Tree(z_dim, last, inp, out, dd, 1, 2)
            setattr(self, "tree%d" % n, tree)
            setattr(self, "mix_rgb%d" % n, ToRGB(z_dim, 6))
            if up_out is not None:
                setattr(self, "upsample%d" % (n + 2), UpSample(z_dim, planes=tree.get_out_planes(), out_planes=up_out,
                                                               kernel_size=4, stride=2, padding=1))
            if n == 3:
                self.upsample_rgb = UpSample_rgb()
        self.activation2 = L.Tanh()
        for i, m in enumerate(mod for mod in self.modules() if isinstance(mod, StyleConv)):
            m.noise_layer = i + 1

    def forward(self, x):
        if rng.philox_noise():
            rng.begin_forward(x.device)
        w = self.mapping_network(torch.squeeze(x))
        g = self.initial_constant.expand(x.shape[0], -1, -1, -1)
        x = self.upsample1(x, w)
        x = self.activation1(self.style1(ops.add(x, g), w))
        rgb = self.to_rgb1(x, w)
        for n in range(1, 5):
            x = getattr(self, "upsample%d" % (n + 1))(x, w)
            rgb = self.upsample_rgb(rgb)
            x, rgb_new = getattr(self, "tree%d" % n)(x, w, rgb)
            rgb = ops.add(rgb, getattr(self, "mix_rgb%d" % n)(ops.cat_channels([rgb, rgb_new]), w))
        return self.activation2(rgb)
