"""generator_14 — drop-in for the reference's generators/generator_14.py (BASELINE config 5, the generator of the
`wgangpsnR2` loop): ConvT stem at 4x4, then four stages of SKConvT (learned x2 + bicubic x2 merged by selective-kernel
attention) followed by two SKConv (3x3 and 5x5 branches merged the same way), and a 5x5 equalised ToRGB conv.
(256,1,1) latent -> (3,64,64) image, planes=32.

It is built from the same pieces as the final generator (generators/generator_13_5.py of this repo: equalised
layers with the gain folded into the conv kernel's alpha, fused BatchNorm, fused bicubic+smooth stencil, SK merge), with
this file's deeper attention trunks (3 convs + 2 fcs, or 4 fcs below 8x8) and BatchNorm-terminated attention heads.
Same module tree / attribute names / construction order as the reference, so state_dicts interchange."""
import torch
import torch.nn as nn

from gan_b200 import layers as L
from gan_b200 import ops
from generators.generator_13_5 import (EqualizedConv2d, EqualizedLinear, EqualizedWeight, Smooth,  # noqa: F401
                                       _conv_bn_act, _fc_bn_act)


class _Attention(nn.Module):
    """heads of reference :60-71 / :105-116: fc,BN,PReLU,fc,BN per branch, softmax across the branches."""

    def _build_heads(self, planes, m):
        self.M = m
        for i in range(m):
            head = nn.Sequential(*_fc_bn_act(planes, 1), EqualizedLinear(planes, planes), L.BatchNorm1d(planes))
            self.__setattr__('fc_sub_%d' % i, head)
        self.softmax = nn.Softmax(dim=1)

    def _heads(self, fea_z, b, s, c):
        subs = [self.__getattr__('fc_sub_%d' % i) for i in range(self.M)]
        return L.sk_attention_vectors(fea_z, subs).view(b, s, c, 1, 1)


class SKAttention_conv(_Attention):
    """reference :41-89 — branch sum -> 5x5 pool -> 3x(conv3x3,BN,PReLU) -> GAP -> 2x(fc,BN,PReLU) -> heads."""

    def __init__(self, planes: int, m: int):
        super().__init__()
        self.gap_conv = L.AdaptiveAvgPool2d(5)
        self.conv_main = nn.Sequential(*_conv_bn_act(planes, 3))
        self.gap_fc = L.AdaptiveAvgPool2d(1)
        self.fc_main = nn.Sequential(*_fc_bn_act(planes, 2))
        self._build_heads(planes, m)

    def forward(self, feas):
        branches = L.as_branches(feas)
        b, c, _, side = branches[0].shape
        assert (side >= 8)
        fea_s = self.conv_main(self.gap_conv(L.branch_sum(branches)))
        fea_z = self.fc_main(self.gap_fc(fea_s).view(b, c))
        return self._heads(fea_z, b, len(branches), c)


class SKAttention_fc(_Attention):
    """reference :92-131 — branch sum -> GAP -> 4x(fc,BN,PReLU) -> heads."""

    def __init__(self, planes: int, m: int):
        super().__init__()
        self.gap_fc = L.AdaptiveAvgPool2d(1)
        self.fc_main = nn.Sequential(*_fc_bn_act(planes, 4))
        self._build_heads(planes, m)

    def forward(self, feas):
        branches = L.as_branches(feas)
        b, c = branches[0].shape[:2]
        fea_z = self.fc_main(self.gap_fc(L.branch_sum(branches)).view(b, c))
        return self._heads(fea_z, b, len(branches), c)


class SKConvT(nn.Module):
    """reference :153-170."""

    def __init__(self, planes: int):
        super().__init__()
        self.convT = L.ConvTranspose2d(planes, planes, kernel_size=4, stride=2, padding=1)
        self.bn = L.BatchNorm2d(planes)
        self.activation_convT = L.PReLU(planes)
        self.up_sample = L.BicubicUp2()
        self.smooth = Smooth()
        self.sk_attention = SKAttention_conv(planes, 2)

    def forward(self, x: torch.Tensor):
        branches = [self.activation_convT(self.bn(self.convT(x))), ops.bicubic_up2_smooth(x)]
        return L.sk_merge(branches, self.sk_attention(branches))


class SKConv(nn.Module):
    """reference :173-202."""

    def __init__(self, in_planes: int, out_planes: int, m: int, image_size: int):
        super().__init__()
        assert (m > 0)
        self.M = m
        for i in range(m):
            self.__setattr__('conv_%d' % i, EqualizedConv2d(in_planes, out_planes, kernel_size=3 + i * 2))
            self.__setattr__('BatchNorm_%d' % i, L.BatchNorm2d(out_planes))
            self.__setattr__('nonlinear_%d' % i, L.PReLU(out_planes))
        self.sk_attention = SKAttention_conv(out_planes, m) if image_size > 4 else SKAttention_fc(out_planes, m)

    def forward(self, x: torch.Tensor):
        branches = []
        for i in range(self.M):
            t = self.__getattr__('BatchNorm_%d' % i)(self.__getattr__('conv_%d' % i)(x))
            branches.append(self.__getattr__('nonlinear_%d' % i)(t))
        return L.sk_merge(branches, self.sk_attention(branches))


class _SEBlock(nn.Module):
    def _tail(self, in_planes):
        self.fc_out = EqualizedLinear(in_planes, in_planes)
        self.fc_bn = L.BatchNorm1d(in_planes)
        self.activation2 = L.Sigmoid()

    def _gate(self, t, b, c):
        return self.activation2(self.fc_bn(self.fc_out(self.fcs(t)))).view(b, c, 1, 1)


class SEBlock_conv(_SEBlock):
    """reference :205-235 (not instantiated by this file's Generator)."""

    def __init__(self, in_planes: int):
        super().__init__()
        self.gap_conv = L.AdaptiveAvgPool2d(5)
        self.convs = nn.Sequential(*_conv_bn_act(in_planes, 3))
        self.gap_fc = L.AdaptiveAvgPool2d(1)
        self.fcs = nn.Sequential(*_fc_bn_act(in_planes, 2))
        self._tail(in_planes)

    def forward(self, x: torch.Tensor):
        b, c, _, side = x.shape
        assert (side >= 8)
        return self._gate(self.gap_fc(self.convs(self.gap_conv(x))).view(b, c), b, c)


class SEBlock_fc(_SEBlock):
    """reference :238-260 (not instantiated by this file's Generator)."""

    def __init__(self, in_planes: int):
        super().__init__()
        self.gap = L.AdaptiveAvgPool2d(1)
        self.fcs = nn.Sequential(*_fc_bn_act(in_planes, 4))
        self._tail(in_planes)

    def forward(self, x: torch.Tensor):
        b, c = x.shape[:2]
        return self._gate(self.gap(x).view(b, c), b, c)


class GeneratorBlock(nn.Module):
    """reference :263-278 — x2, then two SK convs."""

    def get_out_planes(self):
        return self.tree.get_out_planes()  # as the reference: there is no `tree` here, calling it raises AttributeError

    def __init__(self, in_planes: int, out_planes: int, m: int, image_size: int):
        super().__init__()
        self.upsample = SKConvT(in_planes)
        self.convs_1 = SKConv(in_planes, in_planes, m, image_size)
        self.convs_2 = SKConv(in_planes, out_planes, m, image_size)

    def forward(self, x: torch.Tensor):
        return self.convs_2(self.convs_1(self.upsample(x)))


class GeneratorStart(nn.Module):
    """reference :281-307 — ConvT z -> 4x4, two conv+BN+PReLU."""

    def get_out_planes(self):
        return self.tree.get_out_planes()

    def __init__(self, z_dim: int, out_planes: int):
        super().__init__()
        self.convT = L.ConvTranspose2d(z_dim, out_planes, kernel_size=4, stride=1, padding=0)
        self.bn = L.BatchNorm2d(out_planes)
        self.activation = L.PReLU(out_planes)
        self.convs_1 = nn.Sequential(EqualizedConv2d(out_planes, out_planes, 3), L.BatchNorm2d(out_planes),
                                     L.PReLU(out_planes))
        self.convs_2 = nn.Sequential(EqualizedConv2d(out_planes, out_planes, 3), L.BatchNorm2d(out_planes),
                                     L.PReLU(out_planes))

    def forward(self, x: torch.Tensor):
        return self.convs_2(self.convs_1(self.activation(self.bn(self.convT(x)))))


class Generator(nn.Module):
    """reference :310-328."""

    def __init__(self, z_dim, planes=32):
        super().__init__()
        self.block0 = GeneratorStart(z_dim, planes * 8)
        widths = [planes * 8, planes * 4, planes * 2, planes * 2, planes * 1]
        for idx in range(1, 5):
            self.__setattr__('block%d' % idx, GeneratorBlock(widths[idx - 1], widths[idx], 2, 4 << idx))
        self.to_rgb = EqualizedConv2d(planes * 1, 3, 5)

    def forward(self, x: torch.Tensor):
        for idx in range(5):
            x = self.__getattr__('block%d' % idx)(x)
        return self.to_rgb(x)
