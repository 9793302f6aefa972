"""generator_13_5 — drop-in for the reference's final generator, on libganb200 kernels.

Class names, constructor signatures, attribute names and construction order follow
/root/reference/generators/generator_13_5.py (state_dict keys, optimiser parameter order and the RNG stream
consumed while building the 11 180 parameter tensors are identical).  The arithmetic is re-derived:

* Conv2dWeightModulate (reference :219-248) never materialises per-sample weights (B*Cout*Cin*k*k, 2.2 GiB per
  image in the reference).  It runs ONE implicit GEMM with the shared weight: the style scales the
  input channels inside the operand gather, sigma_inv scales the output in the epilogue
  (gan_b200.ops.modulated_conv2d).
* replicate padding is an index clamp inside the conv kernels; cat/slice are channel-slice copies; SK/SE
  products and noise injection are fused broadcast kernels; BatchNorm always uses batch statistics because the
  reference never leaves train mode.
* StyleConv.bias is a dead parameter in the reference (:263 computes x + bias and drops it): it is
  kept as a parameter, never used, never receives a gradient.

Channel plan for planes=48 (reference :610-631): 4^2:396 -> 8^2:396 -> 16^2:204 -> 32^2:108 -> 64^2:108 -> rgb.
"""
import math
import os
from typing import Tuple, List

import torch
import torch.nn as nn

from gan_b200 import layers as L
from gan_b200 import ops
from gan_b200 import rng
from gan_b200 import streams
from gan_b200 import kernels as KN


class EqualizedWeight(L.EqWeightBase):
    """reference :8-16 — parameter `weights`, constant `scale`."""
    PARAM, GAIN = "weights", "scale"

    def __init__(self, shape: List[int]):
        super(EqualizedWeight, self).__init__(shape)


class EqualizedLinear(L.EqLinearBase):
    """reference :19-26 — bias ~ N(0,1)."""
    WEIGHT = EqualizedWeight

    def __init__(self, in_planes: int, out_planes: int):
        super(EqualizedLinear, self).__init__()
        self._build(in_planes, out_planes)


class EqualizedConv2d(L.EqConv2dBase):
    """reference :29-38 — 'same' replicate padding, stride 1."""
    WEIGHT = EqualizedWeight

    def __init__(self, in_planes: int, out_planes: int, kernel_size: int):
        super(EqualizedConv2d, self).__init__()
        self._build(in_planes, out_planes, kernel_size, (kernel_size - 1) // 2)


def _fc_bn_act(planes, n):
    mods = []
    for _ in range(n):
        mods += [EqualizedLinear(planes, planes), L.BatchNorm1d(planes), L.PReLU(planes)]
    return mods


def _conv_bn_act(planes, n):
    mods = []
    for _ in range(n):
        mods += [EqualizedConv2d(planes, planes, 3), L.BatchNorm2d(planes), L.PReLU(planes)]
    return mods


class _SKAttention(nn.Module):
    """shared tail of the two SK attentions: M heads (fc,BN,PReLU,fc) on fea_z, softmax across heads."""

    def _build_heads(self, planes, m):
        self.M = m
        for i in range(m):
            self.__setattr__('fc_sub_%d' % i, L.Sequential(*_fc_bn_act(planes, 1), EqualizedLinear(planes, planes)))
        self.softmax = nn.Softmax(dim=1)

    def _heads(self, fea_z, b, s, c):
        subs = [self.__getattr__('fc_sub_%d' % i) for i in range(self.M)]
        return L.sk_attention_vectors(fea_z, subs).view(b, s, c, 1, 1)


class SKAttention_conv(_SKAttention):
    """reference :41-89 — sum of branches -> 5x5 pool -> 2x(conv3x3,BN,PReLU) -> GAP -> fc,BN,PReLU -> heads."""

    def __init__(self, planes: int, m: int):
        super(SKAttention_conv, self).__init__()
        self.gap_conv = L.AdaptiveAvgPool2d(5)
        self.conv_main = L.Sequential(*_conv_bn_act(planes, 2))
        self.gap_fc = L.AdaptiveAvgPool2d(1)
        self.fc_main = L.Sequential(*_fc_bn_act(planes, 1))
        self._build_heads(planes, m)

    def forward(self, feas):
        branches = L.as_branches(feas)
        b, c, _, side = branches[0].shape
        assert (side >= 8)
        fea_s = self.conv_main(ops.pooled_branch_sum(branches, 5))   # gap_conv(sum of the branches), one pass
        fea_z = self.fc_main(self.gap_fc(fea_s).view(b, c))
        return self._heads(fea_z, b, len(branches), c)


class SKAttention_fc(_SKAttention):
    """reference :92-131 — sum of branches -> GAP -> 2x(fc,BN,PReLU) -> heads."""

    def __init__(self, planes: int, m: int):
        super(SKAttention_fc, self).__init__()
        self.gap_fc = L.AdaptiveAvgPool2d(1)
        self.fc_main = L.Sequential(*_fc_bn_act(planes, 2))
        self._build_heads(planes, m)

    def forward(self, feas):
        branches = L.as_branches(feas)
        b, c = branches[0].shape[:2]
        fea_z = self.fc_main(self.gap_fc(L.branch_sum(branches)).view(b, c))
        return self._heads(fea_z, b, len(branches), c)


class Smooth(L.SmoothBase):
    """reference :134-150"""


class SKConvT(nn.Module):
    """reference :153-170 — learned x2 (ConvT k4 s2 p1 + BN + PReLU) and fixed x2 (bicubic then Smooth, applied as
    one separable stencil) merged by SK attention."""

    def __init__(self, planes: int):
        super(SKConvT, self).__init__()
        self.convT = L.ConvTranspose2d(planes, planes, kernel_size=4, stride=2, padding=1)
        self.bn = L.BatchNorm2d(planes)
        self.activation_convT = L.PReLU(planes)
        self.up_sample = L.BicubicUp2()
        self.smooth = Smooth()
        self.sk_attention = SKAttention_conv(planes, 2)

    def forward(self, x: torch.Tensor):
        learned = self.activation_convT(self.bn(self.convT(x)))
        fixed = ops.bicubic_up2_smooth(x)
        branches = [learned, fixed]
        return L.sk_merge(branches, self.sk_attention(branches))


def _pick_sk_attention(planes, m, image_size):
    return SKAttention_conv(planes, m) if image_size > 4 else SKAttention_fc(planes, m)


class SKConv(nn.Module):
    """reference :173-202 — m branches EqualizedConv2d k=3,5,.. + BN + PReLU, SK merge."""

    def __init__(self, in_planes: int, out_planes: int, m: int, image_size: int):
        super(SKConv, self).__init__()
        assert (m > 0)
        self.M = m
        for i in range(m):
            self.__setattr__('conv_%d' % i, EqualizedConv2d(in_planes, out_planes, kernel_size=3 + i * 2))
            self.__setattr__('BatchNorm_%d' % i, L.BatchNorm2d(out_planes))
            self.__setattr__('nonlinear_%d' % i, L.PReLU(out_planes))
        self.sk_attention = _pick_sk_attention(out_planes, m, image_size)

    def forward(self, x: torch.Tensor):
        branches = []
        for i in range(self.M):
            t = self.__getattr__('conv_%d' % i)(x)
            t = self.__getattr__('BatchNorm_%d' % i)(t)
            branches.append(self.__getattr__('nonlinear_%d' % i)(t))
        return L.sk_merge(branches, self.sk_attention(branches))


class MappingNetwork(nn.Module):
    """reference :205-216"""

    def __init__(self, planes: int, n_layers: int):
        super(MappingNetwork, self).__init__()
        self.net = L.Sequential(*_fc_bn_act(planes, n_layers))

    def forward(self, z: torch.Tensor):
        return self.net(z)


class Conv2dWeightModulate(nn.Module):
    """reference :219-248, in shared-weight form (see module docstring)."""

    def __init__(self, d_latent: int, in_planes: int, out_planes: int, kernel_size: int, demodulate: bool = True,
                 eps: float = 1e-8):
        super(Conv2dWeightModulate, self).__init__()
        self.to_style = nn.Sequential(
            MappingNetwork(d_latent, 1),
            EqualizedLinear(d_latent, in_planes),
            L.BatchNorm1d(in_planes)
        )
        self.out_planes = out_planes
        self.demodulate = demodulate
        self.padding = (kernel_size - 1) // 2
        self.pad = nn.ReplicationPad2d(self.padding)
        self.weight = EqualizedWeight([out_planes, in_planes, kernel_size, kernel_size])
        self.eps = eps

    # filled by StyleBank (below); consumed once per forward
    _style = None
    _style_ready = None

    def _take_style(self, s: torch.Tensor):
        style, ready = self._style, self._style_ready
        if style is None:
            return self.to_style(s)
        self._style = self._style_ready = None
        if ready is not None:  # computed on the bank's side stream
            cur = torch.cuda.current_stream(style.device)
            cur.wait_event(ready)
            style.record_stream(cur)
        return style

    def forward(self, x: torch.Tensor, s: torch.Tensor):
        style = self._take_style(s)
        return ops.modulated_conv2d(x, self.weight.raw, style, self.padding, self.weight.gain, self.demodulate,
                                    self.eps)

    def fused(self, x, s, ns, slope, layer):
        """modconv (+ noise scale `ns`) (+ PReLU `slope`) as one fused layer (gan_b200.ops.StyleLayerFn)."""
        style = self._take_style(s)
        return ops.style_layer(x, self.weight.raw, style, ns, slope, self.padding, self.weight.gain, self.demodulate,
                               self.eps, None, rng.current(), layer)


class StyleBank:
    """The 519 per-layer style vectors `to_style(w)` depend on w alone, not on the feature maps, so they are all computed
    right after the mapping network (GeneratorStart.forward) instead of one by one in front of their convolution.
    GANB200_STYLE_BANK selects how:
      "3" (default) grouped launches — gan_b200.group.GroupedStyles: the whole family in 6 launches forward / 11 backward
          (csrc/group.cu) instead of 4 / 8 per layer; same arithmetic per layer, same parameters and running statistics.
          Falls back to "0" when a to_style module is observed by hooks or is not in training mode.
      "1" the same per-layer kernels as "0", only issued up front (bit-identical to "0", tests/test_style_layer.py);
      "2" like "1" on a side stream (measured on B200: no gain, the step is not latency- but throughput-bound there);
      "0" styles computed where the reference computes them (generator_13_5.py:239)."""

    def __init__(self, layers):
        self.layers = list(layers)
        self.stream = None
        self.grouped = None

    def fill(self, w: torch.Tensor):
        mode = os.environ.get("GANB200_STYLE_BANK", "3")
        if mode == "3":
            if self.grouped is None:
                from gan_b200.group import GroupedStyles
                self.grouped = GroupedStyles([m.to_style for m in self.layers])
            if w.dim() == 2 and self.grouped.eligible():
                for m, s in zip(self.layers, self.grouped(w)):
                    m._style = s
            return
        if mode not in ("1", "2"):
            return
        if mode == "2" and w.is_cuda:
            cur = torch.cuda.current_stream(w.device)
            if self.stream is None:
                self.stream = torch.cuda.Stream(device=w.device)
            self.stream.wait_stream(cur)
            w.record_stream(self.stream)
            with torch.cuda.stream(self.stream):
                for m in self.layers:
                    m._style = m.to_style(w)
                    m._style_ready = torch.cuda.Event()
                    m._style_ready.record(self.stream)
            return
        for m in self.layers:
            m._style = m.to_style(w)


class StyleConv(nn.Module):
    """reference :251-266 — modulated conv, the dead bias, optional per-channel-scaled N(0,1) noise."""

    def __init__(self, d_latent: int, in_planes: int, out_planes: int, kernel_size: int, use_noise: bool = False):
        super(StyleConv, self).__init__()
        self.conv = Conv2dWeightModulate(d_latent, in_planes, out_planes, kernel_size=kernel_size)
        self.use_noise = use_noise
        if use_noise:
            self.scale_noise = nn.Parameter(torch.nn.init.uniform_(torch.empty(out_planes), a=0.2, b=0.3))
        self.bias = nn.Parameter(L.normal_(out_planes))
        self.noise_layer = 0  # unique id per StyleConv, assigned by Generator.__init__ (keys the in-kernel noise)

    def forward(self, x: torch.Tensor, w: torch.Tensor, act: nn.Module = None):
        """`act`: the per-channel PReLU the caller applies to the result (fused into the layer when possible)."""
        hw = x.shape[-1] * x.shape[-2]
        if rng.fused_noise() and hw % 4 == 0:
            ns = self.scale_noise if self.use_noise else None
            return self.conv.fused(x, w, ns, act.weight if act is not None else None, self.noise_layer)
        y = self.conv(x, w)
        if self.use_noise:
            if rng.philox_noise():
                z = KN.backend().noise_fill(tuple(y.shape), rng.current(), self.noise_layer, y.device)
            else:
                # the draw stays a torch.randn call of the same shape on the same generator as in the reference
                # (SURVEY F9): identical seeds give identical noise.
                z = rng.randn(*y.shape, device=y.device)
            y = ops.noise_add(y, self.scale_noise, z)
        return act(y) if act is not None else y


class SKStyleConv(nn.Module):
    """reference :269-295"""

    def __init__(self, d_latent: int, in_planes: int, out_planes: int, m: int, image_size: int, use_noise: bool):
        super(SKStyleConv, self).__init__()
        assert (m > 0)
        self.M = m
        for i in range(m):
            self.__setattr__('conv_%d' % i,
                             StyleConv(d_latent, in_planes, out_planes, kernel_size=3 + i * 2, use_noise=use_noise))
            self.__setattr__('nonlinear_%d' % i, L.PReLU(out_planes))
        self.sk_attention = _pick_sk_attention(out_planes, m, image_size)

    def forward(self, x: torch.Tensor, w: torch.Tensor):
        branches = [self.__getattr__('conv_%d' % i)(x, w, self.__getattr__('nonlinear_%d' % i))
                    for i in range(self.M)]
        return L.sk_merge(branches, self.sk_attention(branches))


class StyleBlock(nn.Module):
    """reference :298-322 — 1x1 modconv, PReLU, (kxk modconv+noise | SK of k3,k5), 3x3 modconv."""

    def __init__(self, d_latent: int, last_planes: int, in_planes: int, out_planes: int, dense_depth: int,
                 kernel_size: int, m: int, image_size: int):
        super(StyleBlock, self).__init__()
        assert (m > 0)
        self.conv1 = StyleConv(d_latent, last_planes, in_planes, kernel_size=1)
        self.activation1 = L.PReLU(in_planes)
        self.m = m
        if m == 1:
            self.conv2 = StyleConv(d_latent, in_planes, in_planes, kernel_size, True)
            self.activation2 = L.PReLU(in_planes)
        else:
            self.skconv = SKStyleConv(d_latent, in_planes, in_planes, m, image_size, True)
        self.conv3 = StyleConv(d_latent, in_planes, out_planes + dense_depth, kernel_size=3)

    def forward(self, x: torch.Tensor, w: torch.Tensor, act: nn.Module = None):
        """`act`: the per-channel PReLU the caller applies to the block's output (BasicBlock's activation_unify /
        activation_shortcut): handed to the last StyleConv, whose fused tail applies it in the same pass."""
        t = self.conv1(x, w, self.activation1)
        t = self.conv2(t, w, self.activation2) if self.m == 1 else self.skconv(t, w)
        return self.conv3(t, w, act)


class ResnetInit(nn.Module):
    """reference :325-349 — residual / transient streams with the two cross paths."""

    def __init__(self, d_latent: int, last_planes: int, in_planes: int, out_planes: int, dense_depth: int,
                 kernel_size: int, m: int, image_size: int):
        super(ResnetInit, self).__init__()
        a = (d_latent, last_planes, in_planes, out_planes)
        b = (kernel_size, m, image_size)
        self.residual = StyleBlock(*a, dense_depth, *b)
        self.transient = StyleBlock(*a, 0, *b)
        self.residual_across = StyleBlock(*a, 0, *b)
        self.transient_across = StyleBlock(*a, dense_depth, *b)
        self.activation_residual = L.PReLU(out_planes + dense_depth)
        self.activation_transient = L.PReLU(out_planes)

    def forward(self, x: Tuple[torch.Tensor, torch.Tensor], w: torch.Tensor):
        x_res, x_tra = x
        # four independent StyleBlocks: on four CUDA streams (gan_b200/streams.py) their latency-bound SK-attention
        # tails overlap; same kernels, same results
        r_r, r_t, t_t, t_r = streams.run_parallel(
            [lambda: self.residual(x_res, w), lambda: self.residual_across(x_res, w),
             lambda: self.transient(x_tra, w), lambda: self.transient_across(x_tra, w)], x_res.device)
        return self.activation_residual(ops.add(r_r, t_r)), self.activation_transient(ops.add(r_t, t_t))


class SEBlock_conv(nn.Module):
    """reference :352-382 — like SK's trunk, ending in fc -> BN1d -> sigmoid gate."""

    def __init__(self, in_planes: int):
        super(SEBlock_conv, self).__init__()
        self.gap_conv = L.AdaptiveAvgPool2d(5)
        self.convs = L.Sequential(*_conv_bn_act(in_planes, 2))
        self.gap_fc = L.AdaptiveAvgPool2d(1)
        self.fcs = L.Sequential(*_fc_bn_act(in_planes, 1))
        self.fc_out = EqualizedLinear(in_planes, in_planes)
        self.fc_bn = L.BatchNorm1d(in_planes)
        self.activation2 = L.Sigmoid()

    def forward(self, x: torch.Tensor):
        b, c, _, side = x.shape
        assert (side >= 8)
        t = self.gap_fc(self.convs(self.gap_conv(x))).view(b, c)
        t = self.fc_bn(self.fc_out(self.fcs(t)))
        return self.activation2(t).view(b, c, 1, 1)


class SEBlock_fc(nn.Module):
    """reference :385-405"""

    def __init__(self, in_planes: int):
        super(SEBlock_fc, self).__init__()
        self.gap = L.AdaptiveAvgPool2d(1)
        self.fcs = L.Sequential(*_fc_bn_act(in_planes, 2))
        self.fc_out = EqualizedLinear(in_planes, in_planes)
        self.fc_bn = L.BatchNorm1d(in_planes)
        self.activation2 = L.Sigmoid()

    def forward(self, x: torch.Tensor):
        b, c = x.shape[:2]
        t = self.fc_bn(self.fc_out(self.fcs(self.gap(x).view(b, c))))
        return self.activation2(t).view(b, c, 1, 1)


class BasicBlock(nn.Module):
    """reference :408-467.  Channel layout of a block's input/output: [residual d | transient d | dense ...]."""

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
        self.is_unify = is_unify
        if is_unify:
            self.unify = StyleBlock(d_latent, last_planes, in_planes, 2 * out_planes, dense_depth, 3, m, image_size)
            self.activation_unify = L.PReLU(2 * out_planes + dense_depth)
            rir_in = out_planes + dense_depth
        else:
            rir_in = last_planes - out_planes
        self.rir_3 = ResnetInit(d_latent, rir_in, in_planes, out_planes, dense_depth, 3, m, image_size)
        if root:
            self.shortcut = StyleBlock(d_latent, last_planes, in_planes, 0, dense_depth, 3, m, image_size)
            self.activation_shortcut = L.PReLU(dense_depth)
        self.se_attention_residual = SEBlock_conv(out_planes) if image_size > 4 else SEBlock_fc(out_planes)

    def forward(self, x: torch.Tensor, w: torch.Tensor):
        d = self.out_planes
        if self.is_unify:
            x = self.unify(x, w, self.activation_unify)
        head = ops.slice_channels(x, 0, d)
        dense = ops.slice_channels(x, 2 * d)
        x_res = ops.cat_channels([head, dense])
        x_tra = ops.slice_channels(x, d)
        res3, tra3 = self.rir_3((x_res, x_tra), w)
        res3_head = ops.slice_channels(res3, 0, d)
        feas_residual = L.gated_residual(head, res3_head, self.se_attention_residual(res3_head))
        if self.root:
            mid = self.shortcut(x, w, self.activation_shortcut)
        else:
            mid = dense
        return ops.cat_channels([feas_residual, tra3, mid, ops.slice_channels(res3, d)])


class ToRGB(nn.Module):
    """reference :470-493"""

    def __init__(self, planes: int, m: int, image_size: int):
        super().__init__()
        assert (m > 0)
        self.m = m
        if m == 1:
            self.pre_conv = EqualizedConv2d(planes, planes, 3)
            self.pre_bn = L.BatchNorm2d(planes)
            self.pre_activation = L.PReLU(planes)
        else:
            self.skconv = SKConv(planes, planes, m, image_size)
        self.conv = EqualizedConv2d(planes, 3, kernel_size=5)
        self.bn = L.BatchNorm2d(3)

    def forward(self, x: torch.Tensor):
        t = self.pre_activation(self.pre_bn(self.pre_conv(x))) if self.m == 1 else self.skconv(x)
        return self.bn(self.conv(t))


class Tree(nn.Module):
    """reference :496-564 — DLA-style aggregation tree.  `level_i` sub-trees, `block_i` leaves, a `root` block
    on the concatenation of everything, and a ToRGB skip accumulated into `rgb`."""

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

        def leaf(last, unify):
            return BasicBlock(d_latent, last, in_planes, out_planes, dense_depth, False, unify, m, image_size)

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
        self.to_rgb = ToRGB(self.get_out_planes(), m, image_size)

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
        return out, ops.add(self.to_rgb(out), rgb)


class GeneratorBlock(nn.Module):
    """reference :567-583 — x2 both streams, then a tree."""

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
    """reference :586-607 — mapping network z->w, ConvT z->4x4, first ToRGB and tree."""

    def get_out_planes(self):
        return self.tree.get_out_planes()

    def __init__(self, z_dim: int, mapping_layer: int, in_planes: int, out_planes: int, dense_depth: int, level: int,
                 block_num: int, m: int):
        super(GeneratorStart, self).__init__()
        self.mapping_network = MappingNetwork(z_dim, mapping_layer)
        self.convT = L.ConvTranspose2d(z_dim, out_planes, kernel_size=4, stride=1, padding=0)
        self.bn = L.BatchNorm2d(out_planes)
        self.activation = L.PReLU(out_planes)
        self.to_rgb = ToRGB(out_planes, m, 4)
        self.tree = Tree(z_dim, out_planes, in_planes, out_planes // 2, dense_depth, level, block_num, m, 4)
        self.style_bank = None  # set by Generator.__init__: every modulated conv of the whole generator

    def forward(self, x: torch.Tensor):
        w = self.mapping_network(torch.squeeze(x))  # all singleton dims, as the reference (B=1 fails)
        if self.style_bank is not None:
            self.style_bank.fill(w)
        x = self.activation(self.bn(self.convT(x)))
        rgb = self.to_rgb(x)
        x, rgb = self.tree(x, w, rgb)
        return x, w, rgb


class Generator(nn.Module):
    """reference :610-631"""

    def __init__(self, z_dim, planes=48):
        super(Generator, self).__init__()
        dd = planes // 8
        self.block0 = GeneratorStart(z_dim, 12, planes * 8, planes * 8, dd, 1, 2, 1)
        last = self.block0.get_out_planes()
        for idx, (mult, size) in enumerate([(4, 8), (2, 16), (1, 32), (1, 64)], start=1):
            blk = GeneratorBlock(z_dim, last, planes * mult, planes * mult, dd, 2, 2, 2, size)
            last = blk.get_out_planes()
            self.__setattr__('block%d' % idx, blk)
        for i, m in enumerate(mod for mod in self.modules() if isinstance(mod, StyleConv)):
            m.noise_layer = i + 1
        self.block0.style_bank = StyleBank(mod for mod in self.modules() if isinstance(mod, Conv2dWeightModulate))

    def forward(self, x: torch.Tensor):
        if rng.philox_noise():
            rng.begin_forward(x.device)
        x, w, rgb = self.block0(x)
        for idx in range(1, 5):
            x, rgb = self.__getattr__('block%d' % idx)(x, w, rgb)
        return rgb
