"""generator_13_4 — drop-in for the reference's generators/generator_13_4.py, the direct predecessor of the final
generator_13_5: the same selective-kernel aggregation tree WITHOUT the style path — plain equalised convolutions followed
by BatchNorm, no mapping network — and with two things 13_5 dropped: SAGAN self-attention (3x3 key / query / value
convolutions, SE-gated residual) on both streams of every block, and selective-kernel merges (instead of sums) of the
ResNet-in-ResNet streams, of a block's skip and of the RGB skips.  Deeper attention trunks than 13_5 (three 3x3 convs,
two / four linears, a BatchNorm closing every head).  generator_13_3 subclasses this file.

Class names, constructor signatures, attribute names and construction order follow the reference file; the equalised
layers, Smooth and the SK / SE helpers are generator_13_5's."""
from typing import Tuple

import torch
import torch.nn as nn

from gan_b200 import layers as L
from gan_b200 import ops
from generators.generator_13_5 import (EqualizedWeight, EqualizedLinear, EqualizedConv2d, Smooth,  # noqa: F401
                                       _fc_bn_act, _conv_bn_act)

# depth of the attention trunks: (3x3 convs of SKAttention_conv / SEBlock_conv, linears after them, linears of the *_fc forms)
DEPTH = dict(conv=3, fc=2, fc_only=4)


class _SKAttention(nn.Module):
    """shared tail of the two SK attentions: M heads (fc, BN, PReLU, fc, BN) on fea_z, softmax across heads"""
    DEPTH = DEPTH

    def _build_heads(self, planes, m):
        self.M = m
        for i in range(m):
            self.__setattr__('fc_sub_%d' % i, L.Sequential(*_fc_bn_act(planes, 1), EqualizedLinear(planes, planes),
                                                           L.BatchNorm1d(planes)))
        self.softmax = nn.Softmax(dim=1)

    def _heads(self, fea_z, b, s, c):
        subs = [self.__getattr__('fc_sub_%d' % i) for i in range(self.M)]
        return L.sk_attention_vectors(fea_z, subs).view(b, s, c, 1, 1)


class SKAttention_conv(_SKAttention):
    """reference :41-89"""

    def __init__(self, planes: int, m: int):
        super(SKAttention_conv, self).__init__()
        self.gap_conv = L.AdaptiveAvgPool2d(5)
        self.conv_main = L.Sequential(*_conv_bn_act(planes, self.DEPTH["conv"]))
        self.gap_fc = L.AdaptiveAvgPool2d(1)
        self.fc_main = L.Sequential(*_fc_bn_act(planes, self.DEPTH["fc"]))
        self._build_heads(planes, m)

    def forward(self, feas):
        branches = L.as_branches(feas)
        b, c, _, side = branches[0].shape
        assert (side >= 8)
        fea_s = self.conv_main(ops.pooled_branch_sum(branches, 5))
        fea_z = self.fc_main(self.gap_fc(fea_s).view(b, c))
        return self._heads(fea_z, b, len(branches), c)


class SKAttention_fc(_SKAttention):
    """reference :92-131"""

    def __init__(self, planes: int, m: int):
        super(SKAttention_fc, self).__init__()
        self.gap_fc = L.AdaptiveAvgPool2d(1)
        self.fc_main = L.Sequential(*_fc_bn_act(planes, self.DEPTH["fc_only"]))
        self._build_heads(planes, m)

    def forward(self, feas):
        branches = L.as_branches(feas)
        b, c = branches[0].shape[:2]
        fea_z = self.fc_main(self.gap_fc(L.branch_sum(branches)).view(b, c))
        return self._heads(fea_z, b, len(branches), c)


def _pick_sk_attention(planes, m, image_size, conv=SKAttention_conv, fc=SKAttention_fc):
    return conv(planes, m) if image_size > 4 else fc(planes, m)


class SKConvT(nn.Module):
    """reference :152-169"""
    SK_CONV = SKAttention_conv

    def __init__(self, planes: int):
        super(SKConvT, self).__init__()
        self.convT = L.ConvTranspose2d(planes, planes, kernel_size=4, stride=2, padding=1)
        self.bn = L.BatchNorm2d(planes)
        self.activation_convT = L.PReLU(planes)
        self.up_sample = L.BicubicUp2()
        self.smooth = Smooth()
        self.sk_attention = self.SK_CONV(planes, 2)

    def forward(self, x: torch.Tensor):
        branches = [self.activation_convT(self.bn(self.convT(x))), ops.bicubic_up2_smooth(x)]
        return L.sk_merge(branches, self.sk_attention(branches))


class SKConv(nn.Module):
    """reference :172-201"""
    SK = staticmethod(_pick_sk_attention)

    def __init__(self, in_planes: int, out_planes: int, m: int, image_size: int):
        super(SKConv, self).__init__()
        assert (m > 0)
        self.M = m
        for i in range(m):
            self.__setattr__('conv_%d' % i, EqualizedConv2d(in_planes, out_planes, kernel_size=3 + i * 2))
            self.__setattr__('BatchNorm_%d' % i, L.BatchNorm2d(out_planes))
            self.__setattr__('nonlinear_%d' % i, L.PReLU(out_planes))
        self.sk_attention = self.SK(out_planes, m, image_size)

    def forward(self, x: torch.Tensor):
        branches = []
        for i in range(self.M):
            t = self.__getattr__('BatchNorm_%d' % i)(self.__getattr__('conv_%d' % i)(x))
            branches.append(self.__getattr__('nonlinear_%d' % i)(t))
        return L.sk_merge(branches, self.sk_attention(branches))


class StyleBlock(nn.Module):
    """reference :204-231 — (1x1 conv, BN, PReLU) then (k x k conv, BN, PReLU) or an SKConv; no style input here."""
    SKCONV = SKConv

    def __init__(self, last_planes: int, in_planes: int, out_planes: int, dense_depth: int,
                 kernel_size: int, m: int, image_size: int):
        super(StyleBlock, self).__init__()
        assert (m > 0)
        self.conv1 = EqualizedConv2d(last_planes, in_planes, kernel_size=1)
        self.bn1 = L.BatchNorm2d(in_planes)
        self.activation1 = L.PReLU(in_planes)
        self.m = m
        if m == 1:
            self.conv2 = EqualizedConv2d(in_planes, out_planes + dense_depth, kernel_size=kernel_size)
            self.bn2 = L.BatchNorm2d(out_planes + dense_depth)
            self.activation2 = L.PReLU(out_planes + dense_depth)
        else:
            self.skconv = self.SKCONV(in_planes, out_planes + dense_depth, m, image_size)

    def forward(self, x: torch.Tensor):
        x = self.activation1(self.bn1(self.conv1(x)))
        if self.m == 1:
            return self.activation2(self.bn2(self.conv2(x)))
        return self.skconv(x)


class ResnetInit(nn.Module):
    """reference :234-261 — the two streams and their cross paths, merged pairwise by SK attention."""
    BLOCK = StyleBlock
    SK = staticmethod(_pick_sk_attention)

    def __init__(self, last_planes: int, in_planes: int, out_planes: int, dense_depth: int,
                 kernel_size: int, m: int, image_size: int):
        super(ResnetInit, self).__init__()
        mk = lambda dd: self.BLOCK(last_planes, in_planes, out_planes, dd, kernel_size, m, image_size)
        self.residual = mk(dense_depth)
        self.transient = mk(0)
        self.residual_across = mk(0)
        self.transient_across = mk(dense_depth)
        self.sk_attention_residual = self.SK(out_planes + dense_depth, 2, image_size)
        self.sk_attention_transient = self.SK(out_planes, 2, image_size)

    def forward(self, x: Tuple[torch.Tensor, torch.Tensor]):
        x_residual, x_transient = x
        feas_residual = [self.residual(x_residual), self.transient_across(x_transient)]
        feas_transient = [self.residual_across(x_residual), self.transient(x_transient)]
        return (L.sk_merge(feas_residual, self.sk_attention_residual(feas_residual)),
                L.sk_merge(feas_transient, self.sk_attention_transient(feas_transient)))


class SEBlock_conv(nn.Module):
    """reference :264-294"""
    DEPTH = DEPTH

    def __init__(self, in_planes: int):
        super(SEBlock_conv, self).__init__()
        self.gap_conv = L.AdaptiveAvgPool2d(5)
        self.convs = L.Sequential(*_conv_bn_act(in_planes, self.DEPTH["conv"]))
        self.gap_fc = L.AdaptiveAvgPool2d(1)
        self.fcs = L.Sequential(*_fc_bn_act(in_planes, self.DEPTH["fc"]))
        self.fc_out = EqualizedLinear(in_planes, in_planes)
        self.fc_bn = L.BatchNorm1d(in_planes)
        self.activation2 = L.Sigmoid()

    def forward(self, x: torch.Tensor):
        b, c, _, side = x.shape
        assert (side >= 8)
        t = self.gap_fc(self.convs(self.gap_conv(x))).view(b, c)
        return self.activation2(self.fc_bn(self.fc_out(self.fcs(t)))).view(b, c, 1, 1)


class SEBlock_fc(nn.Module):
    """reference :297-317"""
    DEPTH = DEPTH

    def __init__(self, in_planes: int):
        super(SEBlock_fc, self).__init__()
        self.gap = L.AdaptiveAvgPool2d(1)
        self.fcs = L.Sequential(*_fc_bn_act(in_planes, self.DEPTH["fc_only"]))
        self.fc_out = EqualizedLinear(in_planes, in_planes)
        self.fc_bn = L.BatchNorm1d(in_planes)
        self.activation2 = L.Sigmoid()

    def forward(self, x: torch.Tensor):
        b, c = x.shape[:2]
        return self.activation2(self.fc_bn(self.fc_out(self.fcs(self.gap(x).view(b, c))))).view(b, c, 1, 1)


class SelfAttention(nn.Module):
    """reference :320-349 — 3x3 query / key / value convolutions, attention over all positions, 3x3 output convolution,
    y = gamma(o) * o + x with an SE block as gamma."""
    SE_CONV, SE_FC = SEBlock_conv, SEBlock_fc

    def __init__(self, in_planes: int, embedding_channels: int, image_size: int):
        super(SelfAttention, self).__init__()
        self.query = EqualizedConv2d(in_planes, embedding_channels, 3)
        self.key = EqualizedConv2d(in_planes, embedding_channels, 3)
        self.value = EqualizedConv2d(in_planes, embedding_channels, 3)
        self.self_att = EqualizedConv2d(embedding_channels, in_planes, 3)
        self.gamma = self.SE_CONV(in_planes) if image_size > 4 else self.SE_FC(in_planes)
        self.softmax = nn.Softmax(dim=1)   # kept for the module tree; the attention runs in L.self_attention

    def forward(self, x: torch.Tensor):
        o = self.self_att(L.self_attention(self.query(x), self.key(x), self.value(x)))
        return L.gated_residual(x, o, self.gamma(o))


class BasicBlock(nn.Module):
    """reference :352-413"""
    RIR = ResnetInit
    ATTENTION = SelfAttention
    SK = staticmethod(_pick_sk_attention)

    def get_out_planes(self):
        if self.is_unify or self.root:
            return 2 * self.out_planes + 2 * self.dense_depth
        return self.last_planes + 1 * self.dense_depth

    def _first(self, out_planes):
        """parameters the block owns ahead of its sub-modules (generator_13_3's `scale`)"""

    def _skip_merge(self, out_planes, image_size):
        """how the block's own first `out_planes` channels and the residual stream's are joined: SK attention here"""
        self.sk_attention_residual = self.SK(out_planes, 2, image_size)

    def _merge_skip(self, head, res_head):
        branches = [head, res_head]
        return L.sk_merge(branches, self.sk_attention_residual(branches))

    def __init__(self, last_planes: int, in_planes: int, out_planes: int, dense_depth: int, root: bool,
                 is_unify: bool, m: int, image_size: int):
        super(BasicBlock, self).__init__()
        self.root = root
        self.last_planes = last_planes
        self.out_planes = out_planes
        self.dense_depth = dense_depth
        self._first(out_planes)
        self.is_unify = is_unify
        stream = last_planes - out_planes
        if is_unify:
            self.unify = EqualizedConv2d(last_planes, 2 * out_planes + dense_depth, kernel_size=1)
            self.bn_unify = L.BatchNorm2d(2 * out_planes + dense_depth)
            stream = out_planes + dense_depth
        self.rir_3 = self.RIR(stream, in_planes, out_planes, dense_depth, 3, m, image_size)
        if root:
            self.shortcut = EqualizedConv2d(last_planes, 2 * out_planes + dense_depth, kernel_size=1)
            self.bn_shortcut = L.BatchNorm2d(2 * out_planes + dense_depth)
        self._skip_merge(out_planes, image_size)
        self.attention_residual = self.ATTENTION(out_planes + dense_depth, out_planes + dense_depth, image_size)
        self.attention_transient = self.ATTENTION(out_planes, out_planes, image_size)

    def forward(self, x: torch.Tensor):
        d = self.out_planes
        if self.is_unify:
            x = self.bn_unify(self.unify(x))
        x_residual = ops.cat_channels([ops.slice_channels(x, 0, d), ops.slice_channels(x, 2 * d)])
        res3, tra3 = self.rir_3((x_residual, ops.slice_channels(x, d)))
        if self.root:
            x = self.bn_shortcut(self.shortcut(x))
        head = self._merge_skip(ops.slice_channels(x, 0, d), ops.slice_channels(res3, 0, d))
        res3 = self.attention_residual(ops.cat_channels([head, ops.slice_channels(res3, d)]))
        tra3 = self.attention_transient(tra3)
        return ops.cat_channels([ops.slice_channels(res3, 0, d), tra3, ops.slice_channels(x, 2 * d),
                                 ops.slice_channels(res3, d)])


class ToRGB(nn.Module):
    """reference :416-440"""
    SKCONV = SKConv

    def __init__(self, planes: int, m: int, image_size: int):
        super().__init__()
        assert (m > 0)
        self.m = m
        if m == 1:
            self.pre_conv = EqualizedConv2d(planes, planes, 3)
            self.pre_bn = L.BatchNorm2d(planes)
            self.pre_activation = L.PReLU(planes)
        else:
            self.skconv = self.SKCONV(planes, planes, m, image_size)
        self.conv = EqualizedConv2d(planes, 3, kernel_size=1)
        self.bn = L.BatchNorm2d(3)
        self.activation = L.PReLU(3)

    def forward(self, x: torch.Tensor):
        t = self.pre_activation(self.pre_bn(self.pre_conv(x))) if self.m == 1 else self.skconv(x)
        return self.activation(self.bn(self.conv(t)))


class Tree(nn.Module):
    """reference :443-514 — the aggregation tree; the RGB skip of every (sub-)tree is merged into `rgb` by SK attention."""
    BLOCK = BasicBlock
    TORGB = ToRGB
    SK = staticmethod(_pick_sk_attention)

    def get_out_planes(self):
        return self.root.get_out_planes()

    def __init__(self, last_planes: int, in_planes: int, out_planes: int, dense_depth: int, level: int,
                 block_num: int, m: int, image_size: int):
        super(Tree, self).__init__()
        assert (block_num > 0)
        self.level = level
        self.block_num = block_num
        self.out_planes = out_planes
        self.dense_depth = dense_depth
        leaf = lambda last, unify: self.BLOCK(last, in_planes, out_planes, dense_depth, False, unify, m, image_size)
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
                sub = type(self)(last_planes, in_planes, out_planes, dense_depth, i, block_num, m, image_size)
                last_planes = sub.get_out_planes()
                width += last_planes
                self.__setattr__('level_%d' % i, sub)
        for i in range(first_leaf, block_num):
            blk = leaf(last_planes, False)
            last_planes = blk.get_out_planes()
            self.__setattr__('block_%d' % i, blk)
        width += blk.get_out_planes()
        self.root_last_planes = width
        self.root = self.BLOCK(width, in_planes * block_num, out_planes, dense_depth, True, False, m, image_size)
        self.to_rgb = self.TORGB(self.get_out_planes(), m, image_size)
        self.sk_attention = self.SK(3, 2, image_size)

    def forward(self, x: torch.Tensor, rgb: torch.Tensor):
        d = self.out_planes
        parts = [self.prev_root(x)] if self.level > 1 else []
        for i in reversed(range(1, self.level)):
            x, rgb = self.__getattr__('level_%d' % i)(x, rgb)
            parts.append(x)
        for i in range(self.block_num):
            x = self.__getattr__('block_%d' % i)(x)
            parts.append(ops.slice_channels(x, 0, 2 * d))
        parts.append(ops.slice_channels(x, 2 * d))
        out = self.root(ops.cat_channels(parts))
        branches = [rgb, self.to_rgb(out)]
        return out, L.sk_merge(branches, self.sk_attention(branches))


class GeneratorBlock(nn.Module):
    """reference :517-532"""
    TREE = Tree
    UP = SKConvT

    def get_out_planes(self):
        return self.tree.get_out_planes()

    def __init__(self, last_planes: int, in_planes: int, out_planes: int, dense_depth: int, level: int,
                 block_num: int, m: int, image_size: int):
        super(GeneratorBlock, self).__init__()
        self.upsample = self.UP(last_planes)
        self.tree = self.TREE(last_planes, in_planes, out_planes, dense_depth, level, block_num, m, image_size)
        self.upsample_rgb = self.UP(3)

    def forward(self, x: torch.Tensor, rgb: torch.Tensor):
        rgb = self.upsample_rgb(rgb)
        return self.tree(self.upsample(x), rgb)


class GeneratorStart(nn.Module):
    """reference :535-555"""
    TREE = Tree
    TORGB = ToRGB

    def get_out_planes(self):
        return self.tree.get_out_planes()

    def __init__(self, z_dim: int, in_planes: int, out_planes: int, dense_depth: int, level: int,
                 block_num: int, m: int):
        super(GeneratorStart, self).__init__()
        self.convT = L.ConvTranspose2d(z_dim, out_planes, kernel_size=4, stride=1, padding=0)
        self.bn = L.BatchNorm2d(out_planes)
        self.activation = L.PReLU(out_planes)
        self.to_rgb = self.TORGB(out_planes, m, 4)
        self.tree = self.TREE(out_planes, in_planes, out_planes // 2, dense_depth, level, block_num, m, 4)

    def forward(self, x: torch.Tensor):
        x = self.activation(self.bn(self.convT(x)))
        return self.tree(x, self.to_rgb(x))


class Generator(nn.Module):
    """reference :558-581"""
    START, BLOCK = GeneratorStart, GeneratorBlock
    # (level, block_num) of block0 .. block4
    PLAN = [(1, 2), (1, 2), (1, 2), (2, 2), (2, 2)]

    def __init__(self, z_dim, planes=64):
        super(Generator, self).__init__()
        dd = planes // 8
        self.block0 = self.START(z_dim, planes * 8, planes * 8, dd, *self.PLAN[0], 1)
        last = self.block0.get_out_planes()
        for idx, (mult, size) in enumerate([(4, 8), (2, 16), (1, 32), (1, 64)], start=1):
            blk = self.BLOCK(last, planes * mult, planes * mult, dd, *self.PLAN[idx], 2, size)
            last = blk.get_out_planes()
            self.__setattr__('block%d' % idx, blk)

    def forward(self, x: torch.Tensor):
        x, rgb = self.block0(x)
        for idx in range(1, 5):
            x, rgb = self.__getattr__('block%d' % idx)(x, rgb)
        return rgb
