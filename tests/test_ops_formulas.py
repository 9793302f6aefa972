"""CPU, float64: the autograd formulas of gan_b200/ops.py (first and second order) against stock torch ops,
with the kernels replaced by their CPU restatement (oracle/kernels_ref.py)."""
import pytest
import torch
import torch.nn.functional as F
from torch.autograd import gradcheck, gradgradcheck

from gan_b200 import ops
from gan_b200 import kernels as KN

torch.manual_seed(0)
D = torch.float64


def r(*shape, grad=True):
    return torch.randn(*shape, dtype=D, requires_grad=grad)


@pytest.mark.parametrize("stride,pad,pad_mode,k", [(1, 1, 1, 3), (2, 1, 1, 3), (1, 0, 0, 3), (1, 2, 1, 5), (2, 1, 0, 4), (1, 0, 0, 1)])
def test_conv2d_matches_torch_and_is_twice_differentiable(ref_backend, stride, pad, pad_mode, k):
    x, w, b = r(2, 3, 6, 6), r(4, 3, k, k), r(4)
    y = ops.conv2d(x, w, b, stride, pad, pad_mode, alpha=0.7)
    xp = F.pad(x, (pad,) * 4, mode="replicate" if pad_mode else "constant") if pad else x
    ref = F.conv2d(xp, w * 0.7, b, stride)
    assert torch.allclose(y, ref, atol=1e-12)
    f = lambda x, w, b: ops.conv2d(x, w, b, stride, pad, pad_mode, alpha=0.7)
    assert gradcheck(f, (x, w, b))
    assert gradgradcheck(f, (x, w, b))


def test_conv_transpose_matches_torch(ref_backend):
    x, w, b = r(2, 3, 4, 4), r(3, 5, 4, 4), r(5)
    y = ops.conv_transpose2d(x, w, b, stride=2, pad=1)
    assert torch.allclose(y, F.conv_transpose2d(x, w, b, stride=2, padding=1), atol=1e-12)
    z, w1 = r(2, 6, 1, 1), r(6, 4, 4, 4)
    assert torch.allclose(ops.conv_transpose2d(z, w1, None, 1, 0), F.conv_transpose2d(z, w1), atol=1e-12)
    assert gradcheck(lambda x, w, b: ops.conv_transpose2d(x, w, b, 2, 1), (x, w, b))


def test_linear(ref_backend):
    x, w, b = r(5, 7), r(3, 7), r(3)
    assert torch.allclose(ops.linear(x, w, b, alpha=0.3), F.linear(x, w * 0.3, b), atol=1e-12)
    f = lambda x, w, b: ops.linear(x, w, b, 0.3)
    assert gradcheck(f, (x, w, b)) and gradgradcheck(f, (x, w, b))


@pytest.mark.parametrize("an", [1, 3])
def test_prelu(ref_backend, an):
    x, a = r(2, 3, 4, 4), (torch.rand(an, dtype=D) * 0.5).requires_grad_()
    assert torch.allclose(ops.prelu(x, a), F.prelu(x, a))
    assert gradcheck(ops.prelu, (x, a))
    # second order through the input-gradient chain (what R1 needs): d/d(a), d/d(gy)
    def chain(x, a, gy):
        y = ops.prelu(x, a)
        (gx,) = torch.autograd.grad(y, x, gy, create_graph=True)
        return gx
    gy = r(2, 3, 4, 4)
    assert gradcheck(chain, (x, a, gy))


def test_resample_ops_match_torch(ref_backend):
    x = r(2, 3, 8, 8)
    k = torch.tensor([[1., 2, 1], [2, 4, 2], [1, 2, 1]], dtype=D) / 16
    sm = F.conv2d(F.pad(x.reshape(-1, 1, 8, 8), (1,) * 4, mode="replicate"), k[None, None]).reshape(x.shape)
    assert torch.allclose(ops.smooth(x), sm, atol=1e-12)
    down = F.interpolate(sm, (4, 4), mode="bicubic", align_corners=False)
    assert torch.allclose(ops.smooth_bicubic_down2(x), down, atol=1e-12)
    up = F.interpolate(x, scale_factor=2, mode="bicubic", align_corners=False)
    assert torch.allclose(ops.bicubic_resize(x, 16, 16, 0.5), up, atol=1e-12)
    sm_up = F.conv2d(F.pad(up.reshape(-1, 1, 16, 16), (1,) * 4, mode="replicate"), k[None, None]).reshape(up.shape)
    assert torch.allclose(ops.bicubic_up2_smooth(x), sm_up, atol=1e-12)
    for hw in (8, 16, 64, 7):
        t = r(1, 2, hw, hw)
        assert torch.allclose(ops.adaptive_avg_pool(t, 5), F.adaptive_avg_pool2d(t, 5), atol=1e-12)
    assert torch.allclose(ops.global_avg_pool(x), x.mean((2, 3)), atol=1e-12)
    assert gradcheck(ops.smooth_bicubic_down2, (x,)) and gradgradcheck(ops.smooth_bicubic_down2, (x,))


def test_pointwise_and_broadcast(ref_backend):
    x, v, vc = r(2, 3, 4, 4), r(2, 3), r(3)
    assert torch.allclose(ops.mul_bc(x, v), x * v[:, :, None, None])
    assert torch.allclose(ops.add_bc(x, vc), x + vc[None, :, None, None])
    assert gradcheck(ops.mul_bc, (x, v)) and gradgradcheck(ops.mul_bc, (x, v))
    assert gradcheck(ops.sigmoid, (x,)) and gradgradcheck(ops.sigmoid, (x,))
    assert gradcheck(ops.tanh, (x,)) and gradgradcheck(ops.tanh, (x,))
    assert gradcheck(lambda t: ops.leaky_relu(t, 0.2), (x,)) and gradgradcheck(lambda t: ops.leaky_relu(t, 0.2), (x,))
    assert torch.allclose(ops.gelu(x), torch.nn.functional.gelu(x), atol=1e-12)    # exact erf form (generator_6)
    assert gradcheck(ops.gelu, (x,)) and gradgradcheck(ops.gelu, (x,))
    xp = (torch.rand(2, 3, dtype=D) + 0.5).requires_grad_()
    assert gradcheck(ops.sqrt, (xp,)) and gradgradcheck(ops.sqrt, (xp,))
    assert gradcheck(lambda t: ops.rsqrt_eps(t, 1e-8), (xp,))
    assert gradcheck(ops.square, (x,)) and gradgradcheck(ops.square, (x,))
    assert gradcheck(ops.mean, (x,)) and gradcheck(ops.sumsq_per_sample, (x,))
    z = r(2, 3, 4, 4, grad=False)
    assert torch.allclose(ops.noise_add(x, vc, z), x + vc[None, :, None, None] * z)
    assert gradcheck(lambda x, vc: ops.noise_add(x, vc, z), (x, vc))
    # x + v*z with a differentiable z: the selective-kernel merge and the gated residuals (twice: R1 runs through D's)
    zz = r(2, 3, 4, 4)
    assert torch.allclose(ops.fma_bc(x, v, zz), x + v[:, :, None, None] * zz)
    assert gradcheck(ops.fma_bc, (x, v, zz)) and gradgradcheck(ops.fma_bc, (x, v, zz))
    from gan_b200 import layers as L
    assert torch.allclose(L.gated_residual(x, zz, v.reshape(2, 3, 1, 1)), x + zz * v[:, :, None, None])
    att = torch.softmax(r(2, 2, 3, grad=False), 1)
    assert torch.allclose(L.sk_merge([x, zz], att), x * att[:, 0, :, None, None] + zz * att[:, 1, :, None, None])


def test_minibatch_stddev(ref_backend):
    x = r(8, 3, 2, 2)
    g = x.view(4, -1)
    std = torch.sqrt(g.var(dim=0) + 1e-8).mean().view(1, 1, 1, 1).expand(8, -1, 2, 2)
    assert torch.allclose(ops.minibatch_stddev(x), torch.cat([x, std], 1), atol=1e-12)
    assert gradcheck(ops.minibatch_stddev, (x,)) and gradgradcheck(ops.minibatch_stddev, (x,))


def test_bce(ref_backend):
    p = torch.rand(6, 1, dtype=D).clamp(0.05, 0.95).requires_grad_()
    t = torch.rand(6, 1, dtype=D)
    assert torch.allclose(ops.bce_loss(p, t), F.binary_cross_entropy(p, t))
    assert gradcheck(lambda p: ops.bce_loss(p, t), (p,))
    with pytest.raises(RuntimeError, match="between 0 and 1"):
        ops.bce_loss(p.detach() + 1.0, t)


def test_batchnorm_and_softmax(ref_backend):
    for shape in [(6, 4), (3, 4, 5, 5)]:
        x, g, b = r(*shape), r(4), r(4)
        rm, rv = torch.zeros(4, dtype=D), torch.ones(4, dtype=D)
        rm2, rv2 = rm.clone(), rv.clone()
        y = ops.batch_norm_train(x, g, b, rm, rv, 1e-5, 0.1)
        ref = F.batch_norm(x, rm2, rv2, g, b, True, 0.1, 1e-5)
        assert torch.allclose(y, ref, atol=1e-10) and torch.allclose(rm, rm2) and torch.allclose(rv, rv2)
        assert gradcheck(lambda x, g, b: ops.batch_norm_train(x, g, b, None, None, 1e-5, 0.1), (x, g, b))
        # twice differentiable (a BatchNorm critic under an R1 / gradient penalty): the create_graph backward re-derives
        # the layer from differentiable pieces; its first derivative must equal the fused backward's
        assert gradgradcheck(lambda x, g, b: ops.batch_norm_train(x, g, b, None, None, 1e-5, 0.1), (x, g, b))
        gy = r(*shape, grad=False)
        fused = torch.autograd.grad(ops.batch_norm_train(x, g, b, None, None, 1e-5, 0.1), (x, g, b), gy)
        graph = torch.autograd.grad(ops.batch_norm_train(x, g, b, None, None, 1e-5, 0.1), (x, g, b), gy, create_graph=True)
        assert all(torch.allclose(u, v, atol=1e-10) for u, v in zip(fused, graph))
        assert torch.allclose(ops.batch_norm_eval(x, g, b, rm, rv, 1e-5), F.batch_norm(x, rm, rv, g, b, False, 0.1, 1e-5), atol=1e-10)
    a = r(3, 2, 5)
    assert torch.allclose(ops.softmax_m(a), torch.softmax(a, 1))
    assert gradcheck(ops.softmax_m, (a,))


@pytest.mark.parametrize("shape", [(16, 12), (6, 12, 3, 3)])
def test_fused_batchnorm_prelu_sequential_matches_torch(ref_backend, shape):
    """gan_b200.layers.Sequential runs (BatchNorm, PReLU(C)) as one fused op each way: output, input / gamma / beta /
    slope gradients, running statistics and num_batches_tracked against torch's own modules, fp64; same state_dict."""
    from gan_b200 import layers as L
    torch.manual_seed(0)
    one_d = len(shape) == 2
    seq = L.Sequential((L.BatchNorm1d if one_d else L.BatchNorm2d)(12), L.PReLU(12)).double()
    want = torch.nn.Sequential((torch.nn.BatchNorm1d if one_d else torch.nn.BatchNorm2d)(12), torch.nn.PReLU(12)).double()
    with torch.no_grad():
        seq[0].weight.normal_()
        seq[0].bias.normal_()
        seq[1].weight.uniform_(0.1, 0.4)
    assert list(seq.state_dict().keys()) == list(want.state_dict().keys())
    want.load_state_dict(seq.state_dict())
    x = torch.randn(*shape, dtype=torch.float64, requires_grad=True)
    x2 = x.detach().clone().requires_grad_()
    y, y2 = seq(x), want(x2)
    assert type(y.grad_fn).__name__ == "BnPReluTrainFnBackward"   # the fused path ran
    gy = torch.randn_like(y)
    y.backward(gy)
    y2.backward(gy)
    assert torch.allclose(y, y2, atol=1e-12) and torch.allclose(x.grad, x2.grad, atol=1e-12)
    for a, b in zip(seq.parameters(), want.parameters()):
        assert torch.allclose(a.grad, b.grad, atol=1e-12)
    for a, b in zip(seq.buffers(), want.buffers()):
        assert torch.allclose(a.double(), b.double(), atol=1e-12)
    # eval mode and a scalar-slope PReLU fall back to module-by-module execution
    seq.eval()
    want.eval()
    assert torch.allclose(seq(x.detach()), want(x.detach()), atol=1e-12)
    seq.train()
    loose = L.Sequential((L.BatchNorm1d if one_d else L.BatchNorm2d)(12), L.PReLU()).double()
    assert type(loose(x).grad_fn).__name__ != "BnPReluTrainFnBackward"


@pytest.mark.parametrize("k", [1, 3, 5])
def test_modulated_conv_equals_per_sample_weight_form(ref_backend, k):
    B, Ci, Co, H = 3, 4, 5, 6
    x, w, s = r(B, Ci, H, H), r(Co, Ci, k, k), r(B, Ci)
    alpha, eps, pad = 0.37, 1e-8, (k - 1) // 2

    def reference_form(x, w, s):  # generators/generator_13_5.py:234-248 restated
        ww = (w * alpha)[None] * s[:, None, :, None, None]
        ww = ww * torch.rsqrt((ww ** 2).sum((2, 3, 4), keepdim=True) + eps)
        xx = F.pad(x.reshape(1, -1, H, H), (pad,) * 4, mode="replicate") if pad else x.reshape(1, -1, H, H)
        return F.conv2d(xx, ww.reshape(B * Co, Ci, k, k), groups=B).reshape(B, Co, H, H)

    y = ops.modulated_conv2d(x, w, s, pad, alpha, True, eps)
    assert torch.allclose(y, reference_form(x, w, s), atol=1e-11)
    gy = torch.randn_like(y)
    g1 = torch.autograd.grad(y, (x, w, s), gy)
    g2 = torch.autograd.grad(reference_form(x, w, s), (x, w, s), gy)
    for a, b in zip(g1, g2):
        assert torch.allclose(a, b, atol=1e-10)


# ---------------------------------------------------------------------------------------------------------------------
# ops of the other configs (cfg4 / cfg5)
# ---------------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("stride,k", [(1, 3), (2, 3), (1, 5), (2, 5), (1, 7)])
def test_depthwise_conv_is_twice_differentiable(ref_backend, stride, k):
    import torch.nn.functional as F
    from gan_b200 import ops
    torch.manual_seed(0)
    x = torch.randn(2, 6, 9, 8, dtype=torch.float64, requires_grad=True)
    w = torch.randn(6, 1, k, k, dtype=torch.float64, requires_grad=True)
    b = torch.randn(6, dtype=torch.float64, requires_grad=True)

    def penalty(conv):
        y = conv(x, w, b)
        (gx,) = torch.autograd.grad(y.tanh().sum(), x, create_graph=True)
        return y.square().sum() + gx.square().sum()

    want = torch.autograd.grad(penalty(lambda x, w, b: F.conv2d(x, w, b, stride, k // 2, groups=6)), (x, w, b))
    got = torch.autograd.grad(penalty(lambda x, w, b: ops.depthwise_conv2d(x, w, b, stride, k // 2)), (x, w, b))
    for a, e in zip(got, want):
        assert torch.allclose(a, e, rtol=1e-9, atol=1e-9)


def test_channel_shuffle_and_batch_stddev(ref_backend):
    from gan_b200 import ops
    torch.manual_seed(0)
    x = torch.randn(5, 12, 3, 4, dtype=torch.float64, requires_grad=True)
    want = x.view(5, 2, 6, 3, 4).transpose(1, 2).contiguous().view(5, 12, 3, 4)
    got = ops.channel_shuffle(x, 2)
    assert torch.equal(got, want)
    wgt = torch.randn_like(x)
    assert torch.equal(torch.autograd.grad((got * wgt).sum(), x)[0], torch.autograd.grad((want * wgt).sum(), x)[0])

    def ref_std(t):  # discriminators/discriminator_3_wgangp_progan.py:9-16 (out of place)
        o = t - t.mean(dim=0, keepdim=True)
        o = torch.sqrt(o.pow(2.0).mean(dim=0) + 10e-8).mean().view(1, 1, 1, 1)
        return torch.cat([t, o.repeat(t.shape[0], 1, t.shape[2], t.shape[3])], 1)

    def pen(fn):
        y = fn(x)
        (gx,) = torch.autograd.grad((y * y[:, -1:]).sum(), x, create_graph=True)
        return y.sum() + gx.square().sum()

    assert torch.allclose(ops.batch_stddev(x), ref_std(x), rtol=1e-10, atol=1e-12)
    a, = torch.autograd.grad(pen(ops.batch_stddev), x)
    e, = torch.autograd.grad(pen(ref_std), x)
    assert torch.allclose(a, e, rtol=1e-7, atol=1e-9)


def test_spectral_norm_module_matches_torch(ref_backend):
    import copy
    from torch.nn.utils.parametrizations import spectral_norm as torch_sn
    from gan_b200 import layers as L
    from gan_b200.spectral import spectral_norm as my_sn
    for ctor in (lambda m: m.Conv2d(6, 10, 3, padding=1), lambda m: m.Conv2d(8, 8, 3, padding=1, groups=8)):
        torch.manual_seed(15)
        base = ctor(torch.nn)
        mine = ctor(L)
        mine.load_state_dict(copy.deepcopy(base.state_dict()))
        torch.manual_seed(16)
        a = torch_sn(base)
        torch.manual_seed(16)
        b = my_sn(mine)
        assert list(a.state_dict()) == list(b.state_dict())
        assert [n for n, _ in a.named_parameters()] == [n for n, _ in b.named_parameters()]
        x = torch.randn(4, base.in_channels, 8, 8, requires_grad=True)
        for _ in range(3):
            ya, yb = a(x), b(x)
        assert torch.allclose(ya, yb, atol=1e-5)
        for k, va in a.state_dict().items():
            assert torch.allclose(b.state_dict()[k], va, atol=1e-6), k
        # R1-style double backward through the normalised conv
        for m, y in ((a, ya), (b, yb)):
            (gx,) = torch.autograd.grad(y.tanh().sum(), x, create_graph=True)
            (y.square().sum() + gx.square().sum()).backward()
        assert torch.allclose(a.parametrizations.weight.original.grad, b.parametrizations.weight.original.grad,
                              rtol=1e-4, atol=1e-5)
        assert torch.allclose(a.bias.grad, b.bias.grad, rtol=1e-4, atol=1e-5)


def test_second_wave_layers_match_torch(ref_backend):
    """Building blocks of the models outside the BASELINE configs, composed of existing kernels:
    reflect padding (also as Conv2d(padding_mode='reflect')), AvgPool2d, InstanceNorm2d with affine parameters and
    running statistics, and the twice-differentiable softmax over branches — against torch in float64."""
    from gan_b200 import layers as L
    x = r(2, 3, 6, 8)
    assert torch.equal(ops.reflect_pad(x, 2), F.pad(x, (2, 2, 2, 2), mode="reflect"))
    assert gradcheck(lambda t: ops.reflect_pad(t, 2), (x,)) and gradgradcheck(lambda t: ops.reflect_pad(t, 2), (x,))
    assert torch.allclose(ops.avg_pool(x, 2), F.avg_pool2d(x, 2)) and gradcheck(lambda t: ops.avg_pool(t, 2), (x,))
    torch.manual_seed(0)
    for mode in ("reflect", "replicate", "zeros"):
        ours, ref = L.Conv2d(3, 4, 5, padding=2, padding_mode=mode).double(), None
        ref = torch.nn.Conv2d(3, 4, 5, padding=2, padding_mode=mode).double()
        ref.load_state_dict(ours.state_dict())
        assert torch.allclose(ours(x), ref(x), atol=1e-10), mode
    pool = L.AvgPool2d(kernel_size=2)
    assert torch.allclose(pool(x), F.avg_pool2d(x, 2))
    # InstanceNorm2d: affine + running statistics, train and eval
    ours, ref = L.InstanceNorm2d(3, affine=True, track_running_stats=True).double(), None
    ref = torch.nn.InstanceNorm2d(3, affine=True, track_running_stats=True).double()
    with torch.no_grad():
        ours.weight.copy_(torch.rand(3) + 0.5); ours.bias.copy_(torch.randn(3))
    ref.load_state_dict(ours.state_dict())
    for _ in range(2):
        assert torch.allclose(ours(x), ref(x), atol=1e-10)
    assert torch.allclose(ours.running_mean, ref.running_mean) and torch.allclose(ours.running_var, ref.running_var)
    ours.eval(); ref.eval()
    assert torch.allclose(ours(x), ref(x), atol=1e-10)
    plain = L.InstanceNorm2d(3).double()  # the defaults: no affine, no running statistics
    assert torch.allclose(plain(x), F.instance_norm(x), atol=1e-10)
    g, b = r(3), r(3)
    f = lambda x, g, b: ops.instance_norm(x, g, b, None, None, 1e-5, 0.1)
    assert gradcheck(f, (x, g, b)) and gradgradcheck(f, (x, g, b))
    # softmax over the branch dimension, differentiated twice
    a = r(3, 2, 5)
    assert gradgradcheck(ops.softmax_m, (a,))
    gy = r(3, 2, 5, grad=False)
    one = torch.autograd.grad(ops.softmax_m(a), a, gy)[0]
    two = torch.autograd.grad(ops.softmax_m(a), a, gy, create_graph=True)[0]
    assert torch.allclose(one, two, atol=1e-12)
