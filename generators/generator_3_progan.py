"""generator_3_progan — drop-in for the reference's generators/generator_3_progan.py (BASELINE config 4, the generator
of the `wgangp` loop): five ConvTranspose2d + BatchNorm2d + scalar-PReLU blocks (nz -> 8ngf@4x4 -> ... -> 3@64x64) and a
Tanh.  Same module tree (`main.<i>.<0|1|2>`), same construction order (so the same seeded initial weights), every
tensor op on libganb200 kernels: the transposed conv is the conv engine's dgrad (tcgen05 or fp32), BatchNorm the fused
batch-statistics kernels, PReLU / Tanh the pointwise kernels.

The two Equalized* classes exist in the reference file (generator_3_progan.py:5-31) but its Generator does not
instantiate them (the EqualizedConvTranspose2D line is commented out, :37); they are mirrored for completeness."""
import numpy
import torch.nn as nn

from gan_b200 import layers as L
from gan_b200 import ops


class EqualizedConv2d(nn.Module):
    """reference generator_3_progan.py:5-17 — conv(x*scale)+bias; the gain is the conv kernel's alpha."""

    def __init__(self, in_planes, out_planes, kernel_size, stride=1, padding=0, groups=1, padding_mode='reflect'):
        super().__init__()
        self.conv = L.Conv2d(in_planes, out_planes, kernel_size, stride, padding, groups=groups,
                             padding_mode=padding_mode)
        self.scale = numpy.sqrt(2) / numpy.sqrt(kernel_size * kernel_size * in_planes)
        self.bias = self.conv.bias
        self.conv.bias = None
        nn.init.normal_(self.conv.weight)
        nn.init.normal_(self.bias)

    def forward(self, x):
        if self.conv.padding[0] != 0 and self.conv.padding_mode != 'zeros':
            raise NotImplementedError("reflect padding is not on the reference's training path (class unused there)")
        return ops.conv2d(x, self.conv.weight, self.bias, self.conv.stride[0], self.conv.padding[0], 0,
                          alpha=float(self.scale))


class EqualizedConvTranspose2D(nn.Module):
    """reference generator_3_progan.py:20-31 — convT(x*scale)+bias."""

    def __init__(self, in_planes, out_planes, kernel_size, stride=1, padding=0):
        super().__init__()
        self.convT = L.ConvTranspose2d(in_planes, out_planes, kernel_size, stride, padding)
        self.scale = numpy.sqrt(2) / numpy.sqrt(in_planes)
        self.bias = self.convT.bias
        self.convT.bias = None
        nn.init.normal_(self.convT.weight)
        nn.init.normal_(self.bias)

    def forward(self, x):
        y = ops.conv_transpose2d(ops.affine(x, float(self.scale), 0.0), self.convT.weight, None,
                                 self.convT.stride[0], self.convT.padding[0])
        return ops.add_bc(y, self.bias)


class Generator(nn.Module):
    """reference generator_3_progan.py:33-57.  (nz,1,1) latent -> (3,64,64) image."""

    # (in multiplier, out multiplier, stride, padding) of the 4x4 transposed convs; None = nz / 3 image channels
    PLAN = ((None, 8, 1, 0), (8, 4, 2, 1), (4, 2, 2, 1), (2, 1, 2, 1), (1, None, 2, 1))

    def get_upsample(self, planes, out_planes, kernel_size, stride, padding):
        block = nn.Sequential()
        block.add_module("0", L.ConvTranspose2d(planes, out_planes, kernel_size=kernel_size, stride=stride,
                                                padding=padding))
        block.add_module("1", L.BatchNorm2d(out_planes))
        block.add_module("2", L.PReLU())  # one shared slope, as in the reference (nn.PReLU())
        return block

    def __init__(self, ngpu, nz, ngf, nc):
        super().__init__()
        self.ngpu = ngpu
        stages = []
        for cin, cout, stride, padding in self.PLAN:
            stages.append(self.get_upsample(nz if cin is None else ngf * cin, 3 if cout is None else ngf * cout,
                                            4, stride, padding))
        self.main = nn.Sequential(*stages, L.Tanh())

    def forward(self, input):
        return self.main(input)
