"""ResNet-in-ResNet critic blocks shared by the drop-ins for the reference's discriminator_5 family
(discriminators/discriminator_5.py, discriminator_5_wgangp.py, discriminator_5_wgangpsn.py — the critics behind
`get_discriminator.get_5 / get_5_wgan / get_5_wgansn`).  The three files differ in: BatchNorm before the LeakyReLUs
(_5 only), spectral norm around every conv (_wgangpsn), the LeakyReLU slope (0.2 / 0.2 / 0.1), layers per block
(3 / 2 / 3) and the head.  Each drop-in file subclasses these bases under the reference's class names (pickled
checkpoints name classes by path) and keeps the reference's attribute names and registration order, so state_dicts
interchange and a seeded construction draws the same weights.

All arithmetic on libganb200: 3x3 zero-padded convs (stride 1 / 2) and the 1x1 strided shortcut on the conv engine,
sums and LeakyReLU on the pointwise kernels, BatchNorm on the batch-statistics kernels; closed under the double
backward of R1 / the gradient penalty."""
import torch.nn as nn

from . import layers as L
from . import ops


class ResnetInitBase(nn.Module):
    """Two streams, four 3x3 convs (r->r, t->t, r->t, t->r), a shortcut on the residual stream.
    reference discriminator_5_wgangp.py:6-40 (the _5 / _wgangpsn files: same lines +-BN / spectral_norm)."""

    SLOPE = 0.2
    BATCHNORM = False          # _5: BatchNorm2d before each LeakyReLU; the Sequentials are then named *_bn_LeakyReLU

    @staticmethod
    def wrap(conv):            # _wgangpsn: spectral_norm(conv)
        return conv

    def __init__(self, in_channel, out_channel, stride):
        super().__init__()
        mk = lambda: self.wrap(L.Conv2d(in_channel, out_channel, 3, padding=1, stride=stride))
        self.residual_stream_conv = mk()
        self.transient_stream_conv = mk()
        self.residual_stream_conv_across = mk()
        self.transient_stream_conv_across = mk()
        act = lambda: nn.Sequential(*([L.BatchNorm2d(out_channel)] if self.BATCHNORM else []),
                                    L.LeakyReLU(self.SLOPE, inplace=True))
        stem = "bn_LeakyReLU" if self.BATCHNORM else "LeakyReLU"
        setattr(self, "residual_" + stem, act())
        setattr(self, "transient_" + stem, act())
        self._act_names = ("residual_" + stem, "transient_" + stem)
        self.short_cut = nn.Sequential()
        if in_channel != out_channel or stride != 1:
            self.short_cut = nn.Sequential(self.wrap(L.Conv2d(in_channel, out_channel, kernel_size=1, stride=stride)))

    def forward(self, x):
        x_residual, x_transient = x
        r_r = self.residual_stream_conv(x_residual)
        r_t = self.residual_stream_conv_across(x_residual)
        shortcut = self.short_cut(x_residual)
        t_t = self.transient_stream_conv(x_transient)
        t_r = self.transient_stream_conv_across(x_transient)
        act_r, act_t = (getattr(self, n) for n in self._act_names)
        return act_r(ops.add(ops.add(r_r, t_r), shortcut)), act_t(ops.add(r_t, t_t))


class RiRBlockBase(nn.Module):
    """`layer_num` ResnetInit layers, the first with `stride`; children are named 'generalized layers<i>' as in the
    reference (discriminator_5_wgangp.py:43-60)."""

    LAYER = ResnetInitBase

    def __init__(self, in_channel, out_channel, layer_num, stride, layer=None):
        super().__init__()
        self.resnetinit = self._make_layers(in_channel, out_channel, layer_num, stride)

    def forward(self, x):
        x_residual, x_transient = self.resnetinit(x)
        return (x_residual, x_transient)

    def _make_layers(self, in_channel, out_channel, layer_num, stride, layer=None):
        layers = nn.Sequential()
        for index, s in enumerate([stride] + [1] * (layer_num - 1)):
            layers.add_module("generalized layers{}".format(index), self.LAYER(in_channel, out_channel, s))
            in_channel = out_channel
        return layers


class RiRDiscriminatorBase(nn.Module):
    """(3,64,64) -> two 64-channel streams -> eight RiR blocks (64 px ... 4 px, 64 -> 256 channels) -> concat -> head.
    reference discriminator_5_wgangp.py:63-106."""

    BLOCK = RiRBlockBase
    LAYERS_PER_BLOCK = 2
    PRE_BATCHNORM = False
    # (in multiplier, out multiplier, stride) of rir1..rir8
    PLAN = ((1, 1, 1), (1, 1, 1), (1, 2, 2), (2, 2, 2), (2, 2, 1), (2, 4, 2), (4, 4, 2), (4, 4, 1))

    def make_head(self):
        raise NotImplementedError

    def __init__(self):
        super().__init__()
        base = 64
        wrap, slope = self.BLOCK.LAYER.wrap, self.BLOCK.LAYER.SLOPE
        pre = lambda: nn.Sequential(wrap(L.Conv2d(3, base, 3, padding=1)),
                                    *([L.BatchNorm2d(base)] if self.PRE_BATCHNORM else []),
                                    L.LeakyReLU(slope, inplace=True))
        self.residual_pre_conv = pre()
        self.transient_pre_conv = pre()
        for i, (cin, cout, stride) in enumerate(self.PLAN):
            setattr(self, "rir%d" % (i + 1), self.BLOCK(base * cin, base * cout, self.LAYERS_PER_BLOCK, stride))
        self.conv1 = self.make_head()

    def trunk(self, x):
        streams = (self.residual_pre_conv(x), self.transient_pre_conv(x))
        for i in range(len(self.PLAN)):
            streams = getattr(self, "rir%d" % (i + 1))(streams)
        h = self.conv1(ops.cat_channels(list(streams)))
        return h.reshape(h.size(0), -1)

    def forward(self, x):
        return self.trunk(x)
