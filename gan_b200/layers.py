"""nn.Module building blocks of the drop-in model files (generators/*.py, discriminators/*.py).

The model files keep the reference's class names, constructor signatures, attribute names (= state_dict
keys, parameter registration order and therefore RNG order at construction) — that is the API.  What a
forward() does is different: every tensor op goes through gan_b200.ops, i.e. through libganb200's CUDA
kernels; nothing here calls ATen math.  Stock torch container / parameter-holder classes are subclassed so
that `isinstance`, `state_dict()`, `.apply(weights_init)` and default initialisers behave as in the reference.

Reference semantics cited per class (paths under /root/reference).
"""
import math

import numpy as np
import os

import torch
import torch.nn as nn

from . import kernels as KN
from . import ops


def normal_(shape):
    return torch.nn.init.normal_(torch.empty(shape), mean=0, std=1)


# --------------------------------------------------------------------------------------------------------
# parameter-holder subclasses of stock modules: same parameters / buffers / init, our kernels in forward()
# --------------------------------------------------------------------------------------------------------
class PReLU(nn.PReLU):
    """nn.PReLU(C) / nn.PReLU() — e.g. discriminators/discriminator_9_4.py:91, generators/generator_13_5.py:50."""

    def forward(self, x):
        return ops.prelu(x, self.weight)


class Sigmoid(nn.Sigmoid):
    def forward(self, x):
        return ops.sigmoid(x)


class Tanh(nn.Tanh):
    def forward(self, x):
        return ops.tanh(x)


class LeakyReLU(nn.LeakyReLU):
    """inplace is accepted and ignored (the reference never reads the pre-activation afterwards)."""

    def forward(self, x):
        return ops.leaky_relu(x, self.negative_slope)


class GELU(nn.GELU):
    """nn.GELU() — generators/generator_6.py:12 (the exact form only)."""

    def forward(self, x):
        assert self.approximate == 'none'
        return ops.gelu(x)


class _BatchNormMixin:
    def forward(self, x):
        self._check_input_dim(x)
        if self.training or not self.track_running_stats:
            n, _, hw = KN._nchw(x)
            if n * hw <= 1:
                raise ValueError(f"Expected more than 1 value per channel when training, got input size {x.size()}")
            if self.momentum is None:
                raise NotImplementedError("BatchNorm(momentum=None) (cumulative moving average) is not implemented; the "
                                          "reference's models all use the default momentum 0.1")
            momentum = self.momentum
            track = self.training and self.track_running_stats
            return ops.batch_norm_train(x, self.weight, self.bias,
                                        self.running_mean if track else None, self.running_var if track else None,
                                        self.eps, momentum, self.num_batches_tracked if track else None)
        return ops.batch_norm_eval(x, self.weight, self.bias, self.running_mean, self.running_var, self.eps)


class Sequential(nn.Sequential):
    """nn.Sequential (same child names, so the same state_dict keys) that runs a training-mode BatchNorm immediately
    followed by PReLU(C) as one fused launch when the per-channel population is small — the (linear | conv, BN, PReLU)
    triples of generators/generator_13_5.py:60-66.  Anything else is executed module by module."""

    def forward(self, x):
        mods = list(self)
        i = 0
        while i < len(mods):
            m = mods[i]
            nxt = mods[i + 1] if i + 1 < len(mods) else None
            if (isinstance(m, _BatchNormMixin) and isinstance(nxt, PReLU) and m.training and m.track_running_stats
                    and m.weight is not None and m.bias is not None and not m._forward_hooks and
                    not nxt._forward_hooks and not nxt._forward_pre_hooks and ops.bn_prelu_fusable(x, nxt.weight)):
                m._check_input_dim(x)
                n, _, hw = KN._nchw(x)
                if n * hw <= 1:
                    raise ValueError(f"Expected more than 1 value per channel when training, got input size {x.size()}")
                x = ops.batch_norm_prelu_train(x, m.weight, m.bias, nxt.weight, m.running_mean, m.running_var, m.eps,
                                               0.0 if m.momentum is None else m.momentum, m.num_batches_tracked)
                i += 2
                continue
            x = m(x)
            i += 1
        return x


class BatchNorm2d(_BatchNormMixin, nn.BatchNorm2d):
    """nn.BatchNorm2d in the generators (generator_13_5.py:49,157,...): batch statistics whenever the module
    is in train mode — which in the reference is always — running stats updated on every call."""


class BatchNorm1d(_BatchNormMixin, nn.BatchNorm1d):
    pass


class AdaptiveAvgPool2d(nn.AdaptiveAvgPool2d):
    def forward(self, x):
        out = self.output_size if isinstance(self.output_size, int) else self.output_size[0]
        if out == 1:
            b, c = x.shape[:2]
            return ops.global_avg_pool(x).reshape(b, c, 1, 1)
        return ops.adaptive_avg_pool(x, out)


class AvgPool2d(nn.AvgPool2d):
    """nn.AvgPool2d(k) with the default stride (= k), no padding — the SE squeeze of the tree generators / critics
    (generators/generator_7.py:19, discriminators/discriminator_6_wgangp.py:18)."""

    def forward(self, x):
        k = self.kernel_size if isinstance(self.kernel_size, int) else self.kernel_size[0]
        assert self.padding in (0, (0, 0)) and self.stride in (k, (k, k))
        return ops.avg_pool(x, k)


class InstanceNorm2d(nn.InstanceNorm2d):
    """nn.InstanceNorm2d (affine / track_running_stats as constructed) — generators/generator_4.py, generator_5.py."""

    def forward(self, x):
        self._check_input_dim(x)
        use_input = self.training or not self.track_running_stats
        track = self.training and self.track_running_stats  # (torch does not advance num_batches_tracked here)
        return ops.instance_norm(x, self.weight, self.bias, self.running_mean if (track or not use_input) else None,
                                 self.running_var if (track or not use_input) else None, self.eps,
                                 0.0 if self.momentum is None else self.momentum, use_input)


class ReflectionPad2d(nn.ReflectionPad2d):
    def forward(self, x):
        p = self.padding if isinstance(self.padding, int) else self.padding[0]
        assert isinstance(self.padding, int) or len(set(self.padding)) == 1, "same padding on all four sides"
        return ops.reflect_pad(x, p)


class BicubicUp2(nn.Upsample):
    """nn.Upsample(scale_factor=2, mode='bicubic', align_corners=False) — generator_13_5.py:160."""

    def __init__(self):
        super().__init__(scale_factor=2, mode='bicubic', align_corners=False)

    def forward(self, x):
        return ops.bicubic_resize(x, 2 * x.shape[-2], 2 * x.shape[-1], scale=0.5)


class ConvTranspose2d(nn.ConvTranspose2d):
    """nn.ConvTranspose2d with square kernel/stride/padding (generator_13_5.py:156,594; generator_2.py)."""

    def forward(self, x):
        assert self.groups == 1 and self.dilation[0] == 1 and self.output_padding[0] == 0
        return ops.conv_transpose2d(x, self.weight, self.bias, self.stride[0], self.padding[0])


class Conv2d(nn.Conv2d):
    """plain nn.Conv2d, zero padding (DCGAN-style discriminators); groups == channels -> the depthwise kernels
    (discriminator_7_wgangpsn.py:55)."""

    def forward(self, x):
        assert self.dilation[0] == 1 and self.padding_mode in ('zeros', 'reflect', 'replicate')
        pad, mode = self.padding[0], 0
        if self.padding_mode == 'reflect' and pad:   # mirror first (a separable gather), then an unpadded conv
            x, pad = ops.reflect_pad(x, pad), 0
        elif self.padding_mode == 'replicate':
            mode = 1                                  # an index clamp inside the conv kernel
        if self.groups == 1:
            return ops.conv2d(x, self.weight, self.bias, self.stride[0], pad, mode)
        assert self.groups == self.in_channels == self.out_channels, "grouped convs: depthwise only"
        assert mode == 0, "depthwise: zero or reflect padding"
        return ops.depthwise_conv2d(x, self.weight, self.bias, self.stride[0], pad)


class Linear(nn.Linear):
    def forward(self, x):
        return ops.linear(x, self.weight, self.bias)


# --------------------------------------------------------------------------------------------------------
# equalised-learning-rate layers.  G and D spell the attributes differently:
#   D: EqualizedWeight.weight / .c       (discriminator_9_4.py:9-17)
#   G: EqualizedWeight.weights / .scale  (generator_13_5.py:8-16)
# The gain is never multiplied into a copy of the weight: it is the `alpha` of the GEMM/conv kernel.
# --------------------------------------------------------------------------------------------------------
class EqWeightBase(nn.Module):
    PARAM, GAIN = "weight", "c"

    def __init__(self, shape):
        super().__init__()
        self.shape = shape
        setattr(self, self.GAIN, 1 / math.sqrt(np.prod(shape[1:])))
        setattr(self, self.PARAM, nn.Parameter(normal_(shape)))

    @property
    def raw(self):
        return getattr(self, self.PARAM)

    @property
    def gain(self):
        return float(getattr(self, self.GAIN))

    def forward(self):
        return ops.affine(self.raw, self.gain, 0.0)


class EqLinearBase(nn.Module):
    """linear(x, W*c, bias) — discriminator_9_4.py:20-27, generator_13_5.py:19-26."""
    WEIGHT = EqWeightBase

    def _build(self, in_planes, out_planes, bias_gain=None):
        self.weight = self.WEIGHT([out_planes, in_planes])
        b = normal_(out_planes)
        if bias_gain is not None:
            b = b * bias_gain
        self.bias = nn.Parameter(b)

    def forward(self, x):
        return ops.linear(x, self.weight.raw, self.bias, alpha=self.weight.gain)


class EqConv2dBase(nn.Module):
    """ReplicationPad2d(p) -> conv2d(x, W*c, bias, stride) — discriminator_9_4.py:30-40, generator_13_5.py:29-38.
    The padding is an index clamp inside the conv kernel; `self.pad` stays as an (unused) child so that the
    module tree has the reference's names."""
    WEIGHT = EqWeightBase

    def _build(self, in_planes, out_planes, kernel_size, padding, stride=1):
        self.stride = stride
        self.padding = padding
        self.weight = self.WEIGHT([out_planes, in_planes, kernel_size, kernel_size])
        self.bias = nn.Parameter(normal_(out_planes))
        self.pad = nn.ReplicationPad2d(padding)

    def forward(self, x):
        return ops.conv2d(x, self.weight.raw, self.bias, self.stride, self.padding, 1, alpha=self.weight.gain)


class SmoothBase(nn.Module):
    """[[1,2,1],[2,4,2],[1,2,1]]/16 blur with replicate padding — discriminator_9_4.py:56-72.
    `kernel` is a frozen Parameter exactly as in the reference (it shows up in parameters())."""

    def __init__(self):
        super().__init__()
        k = torch.tensor([[[[1, 2, 1], [2, 4, 2], [1, 2, 1]]]], dtype=torch.float)
        self.kernel = nn.Parameter(torch.div(k, k.sum()), requires_grad=False)
        self.pad = nn.ReplicationPad2d(1)

    def forward(self, x):
        return ops.smooth(x)


# --------------------------------------------------------------------------------------------------------
# helpers shared by SK / SE attention
# --------------------------------------------------------------------------------------------------------
def as_branches(feas):
    """SK modules take the branch stack either as a list of (b,c,h,w) tensors (our callers) or as the
    reference's (b,s,c,h,w) tensor."""
    if isinstance(feas, (list, tuple)):
        return list(feas)
    return [feas[:, i] for i in range(feas.shape[1])]


def branch_sum(branches):
    acc = branches[0]
    for t in branches[1:]:
        acc = ops.add(acc, t)
    return acc


def sk_attention_vectors(fea_z, subs):
    """[fc_sub_i(fea_z)] stacked on dim 1 and soft-maxed over it: (b, M, c)."""
    b, c = fea_z.shape
    vecs = [fc(fea_z).reshape(b, 1, c) for fc in subs]
    return ops.softmax_m(ops.cat_channels(vecs))


def sk_merge(branches, att):
    """(feas * attention).sum(dim=1) — generator_13_5.py:169,201,294.  att: (b,M,c,1,1) or (b,M,c)."""
    b, m = att.shape[:2]
    att = att.reshape(b, m, -1)
    if (m == 2 and len(branches) == 2 and branches[0].dim() >= 3 and branches[0].shape == branches[1].shape
            and branches[0].dtype == torch.float32 and os.environ.get("GANB200_SK_MERGE_FUSED", "1") != "0"):
        return ops.sk_merge2(branches[0], branches[1], att)      # one pass each way instead of mul_bc + fma_bc
    out = None
    for i, fea in enumerate(branches):
        a_i = ops.slice_channels(att, i, i + 1).reshape(b, -1)
        out = ops.mul_bc(fea, a_i, KN.V_NC) if out is None else ops.fma_bc(out, a_i, fea, KN.V_NC)
    return out


def self_attention(f, g, h):
    """SAGAN attention of discriminators/discriminator_9_3.py:70-77 on (B, C', H, W) query / key / value maps:
    s = f^T g (B, N, N), beta = softmax(s, dim=1), v = h beta, returned as (B, C', H, W).
    softmax over dim 1 of s is the row softmax of s^T = g^T f, and h beta = h (beta^T)^T — so the attention map is only
    ever held transposed and the two products read their operands through strides."""
    b, c, hh, ww = h.shape
    n = hh * ww
    f3, g3, h3 = f.reshape(b, -1, n), g.reshape(b, -1, n), h.reshape(b, -1, n)
    beta_t = ops.softmax_rows(ops.bmm(g3, f3, True, False))      # [b, j, i] = softmax_i(s[b, i, j])
    return ops.bmm(h3, beta_t, False, True).reshape(b, c, hh, ww)


def gate(x, g):
    """x * g with g:(b,c,1,1) — the SE product (discriminator_9_4.py:160, generator_13_5.py:457)."""
    b, c = x.shape[:2]
    return ops.mul_bc(x, g.reshape(b, c), KN.V_NC)


def gated_residual(skip, x, g):
    """skip + x * g in one pass (the SE product followed by the residual add: discriminator_9_4.py:160-161,
    generator_13_5.py:457-458)."""
    b, c = x.shape[:2]
    return ops.fma_bc(skip, g.reshape(b, c), x, KN.V_NC)


class BCELoss(nn.BCELoss):
    """torch.nn.BCELoss(mean) — train/gan.py:21, ganR2.py:17.  Raises the reference's RuntimeError when an input
    lies outside [0,1] (D9_4 emits raw logits)."""

    def forward(self, input, target):
        assert self.weight is None and self.reduction == "mean"
        if input.shape != target.shape:
            raise ValueError(f"Using a target size ({target.size()}) that is different to the input size "
                             f"({input.size()}) is deprecated. Please ensure they have the same size.")
        return ops.bce_loss(input, target)
