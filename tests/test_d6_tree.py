"""discriminator_6_wgangp (the deep-layer-aggregation critic behind get_6_wgan) against the
UNMODIFIED reference class built from the same seed: identical state_dict (names, order, values), and D(x), the input
gradient and a weight gradient taken through the double backward, on the CPU restatement of the kernels (depthwise
3x3 / 5x5 / 7x7 on reflect-padded maps, windowed average pooling, channel slices and concatenations).  The full
R1 + R2 + GP iteration against a recorded golden is tests/test_other_configs.py, "tree6".  Skipped when the reference
tree is not present."""
import importlib

import pytest
import torch

from oracle import reference as R

pytestmark = pytest.mark.skipif(not R.available(), reason="reference tree not present")


def rel(a, b):
    return float((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-30))


def test_tree_critic_equals_reference(ref_backend):
    ref_mod = R.load("discriminators.discriminator_6_wgangp")
    our_mod = importlib.import_module("discriminators.discriminator_6_wgangp")
    torch.manual_seed(11)
    D_ref = ref_mod.Discriminator()
    torch.manual_seed(11)
    D = our_mod.Discriminator()
    sd_ref, sd = D_ref.state_dict(), D.state_dict()
    assert list(sd.keys()) == list(sd_ref.keys()) and len(sd) == 2554
    assert [n for n, _ in D.named_parameters()] == [n for n, _ in D_ref.named_parameters()]
    for k, v in sd_ref.items():
        assert torch.equal(sd[k], v), k
    for cls in ("SqueezeExcitationBottleneck", "ResnetInit", "BasicBlock", "Tree", "Discriminator"):
        assert hasattr(our_mod, cls), cls     # whole-module checkpoints of the reference name these classes
    x = torch.randn(2, 3, 64, 64, generator=torch.Generator().manual_seed(12))

    def run(net, dtype):
        net = net.to(dtype)
        xi = x.to(dtype).clone().requires_grad_()
        y = net(xi)
        assert y.shape == (2, 1, 1, 1)
        (gx,) = torch.autograd.grad(y.sum(), xi, create_graph=True)
        (gw,) = torch.autograd.grad(gx.pow(2).sum(), net.model.Downsample_0[0].weight)   # through the double backward
        return y.detach(), gx.detach(), gw

    for a, b in zip(run(D, torch.float64), run(D_ref, torch.float64)):   # (float32: the "tree6" golden iteration)
        assert rel(a, b) < 1e-12


# (generator_5 / 8 / 9 / 10: structure, initialisation and a whole iteration are compared in tests/test_other_configs.py)
@pytest.mark.parametrize("name", ["generator_4", "generator_7", "generator_10_2"])
def test_tree_generators_equal_reference(ref_backend, name):
    """generator_7 / generator_8: the same tree with BatchNorm + scalar-PReLU bottlenecks (they subclass the critic's
    classes); generator_4 / generator_5: InstanceNorm trees on BatchNorm stems; generator_9 / 10 / 10_2: self-attention
    trees (10_2 with run-time-scaled equalised convolutions).  Identical state_dict; for generator_10_2
    also G(z) and first-order gradients in float64 (a full iteration of each against a recorded golden:
    tests/test_other_configs.py, "tree4" ... "tree10_2"; 10_2 is too ill-conditioned in float32 for that to say much)."""
    import random
    ref_mod = R.load("generators." + name)
    our_mod = importlib.import_module("generators." + name)
    torch.manual_seed(11)
    random.seed(11)              # generator_10 / 10_2 draw their attention gates from Python's `random`
    G_ref = ref_mod.Generator(64)
    torch.manual_seed(11)
    random.seed(11)
    G = our_mod.Generator(64)
    sd_ref, sd = G_ref.state_dict(), G.state_dict()
    assert list(sd.keys()) == list(sd_ref.keys())
    assert [n for n, _ in G.named_parameters()] == [n for n, _ in G_ref.named_parameters()]
    for k, v in sd_ref.items():
        assert torch.equal(sd[k], v), k
    for cls in [n for n in dir(ref_mod) if isinstance(getattr(ref_mod, n), type) and
                issubclass(getattr(ref_mod, n), torch.nn.Module) and getattr(ref_mod, n).__module__ == ref_mod.__name__]:
        assert hasattr(our_mod, cls), cls     # whole-module checkpoints of the reference name these classes
    if name != "generator_10_2":
        return
    z = torch.randn(2, 64, 1, 1, generator=torch.Generator().manual_seed(12), dtype=torch.float64)

    def run(net):
        net = net.double()
        zi = z.clone().requires_grad_()
        y = net(zi)
        assert y.shape == (2, 3, 64, 64)
        first = next(net.parameters())
        gz, gw = torch.autograd.grad((y * torch.linspace(-1, 1, y.numel(), dtype=y.dtype).reshape(y.shape)).sum(), [zi, first])
        return y.detach(), gz, gw

    for a, b in zip(run(G), run(G_ref)):
        assert rel(a, b) < 1e-8


@pytest.mark.parametrize("name", ["generator_13_3", "generator_13_4"])
def test_final_generator_predecessors_equal_reference(ref_backend, name):
    """generator_13_3 / generator_13_4 (no style path, SK merges and SAGAN attention in every block) at planes=8, float64.
    13_3 agrees end to end (output and gradients).  13_4 at initialisation is a chaotic map — a 1e-14 difference after
    block0 grows 30 .. 1 000-fold per resolution level in float64, in the reference as here (its float32 run differs from
    its own float64 run by O(1)) — so it is compared level by level while the difference is still rounding."""
    ref_mod = R.load("generators." + name)
    our_mod = importlib.import_module("generators." + name)
    torch.manual_seed(11)
    G_ref = ref_mod.Generator(32, 8).double()
    torch.manual_seed(11)
    G = our_mod.Generator(32, 8).double()
    sd_ref, sd = G_ref.state_dict(), G.state_dict()
    assert list(sd.keys()) == list(sd_ref.keys())
    assert [n for n, _ in G.named_parameters()] == [n for n, _ in G_ref.named_parameters()]
    for k, v in sd_ref.items():
        assert torch.equal(sd[k], v), k
    for cls in [n for n in dir(ref_mod) if isinstance(getattr(ref_mod, n), type) and
                issubclass(getattr(ref_mod, n), torch.nn.Module) and getattr(ref_mod, n).__module__ == ref_mod.__name__]:
        assert hasattr(our_mod, cls), cls
    z = torch.randn(16, 32, 1, 1, dtype=torch.float64, generator=torch.Generator().manual_seed(12))
    if name == "generator_13_3":
        def run(net):
            zi = z[:8].clone().requires_grad_()
            y = net(zi)
            first = next(net.parameters())
            return (y.detach(),) + torch.autograd.grad((y * torch.linspace(-1, 1, y.numel(), dtype=y.dtype).reshape(y.shape)).sum(),
                                                       [zi, first])
        for a, b in zip(run(G), run(G_ref)):
            assert rel(a, b) < 1e-8
        return

    def levels(net):      # 4x4 .. 32x32 (the 64x64 level costs as much as the other four together and says nothing new)
        out = []
        with torch.no_grad():
            x, rgb = net.block0(z)
            out.append((x, rgb))
            for i in range(1, 4):
                x, rgb = getattr(net, "block%d" % i)(x, rgb)
                out.append((x, rgb))
        return out
    lv, lv_ref = levels(G), levels(G_ref)
    for i, bound in enumerate([1e-11, 1e-9, 1e-7, 1e-4]):       # measured 2e-14, 1e-12, 9e-11, 6e-8 (then 2e-4 at 64x64)
        for a, b in zip(lv[i], lv_ref[i]):
            assert rel(a, b) < bound, (i, rel(a, b))


def test_generator_13_2_is_generator_13():
    """the reference's generator_13_2.py is byte for byte its generator_13.py; here it re-exports generator_13's classes"""
    import filecmp
    import os
    ref_dir = os.path.dirname(R.load("generators.generator_13").__file__)
    assert filecmp.cmp(os.path.join(ref_dir, "generator_13.py"), os.path.join(ref_dir, "generator_13_2.py"), shallow=False)
    a, b = importlib.import_module("generators.generator_13"), importlib.import_module("generators.generator_13_2")
    for cls in ("Generator", "Tree", "BasicBlock", "StyleConv", "Conv2dWeightModulate", "SelfAttention"):
        assert getattr(a, cls) is getattr(b, cls)


def test_generator_13_1_gradients_equal_reference_in_float64(ref_backend):
    """generator_13_1 at planes=8: G(z) and the gradient of EVERY parameter against the reference in float64.  The bound
    is loose for a float64 test on purpose: the key / query paths in front of the saturated attention softmaxes agree to
    ~2e-9 where everything else agrees to 1e-13 — the conditioning that makes their float32 update norms meaningless
    (tests/test_other_configs.py, NOISE_PARAMS["tree13_1"]).  Key biases have an analytically zero gradient."""
    ref_mod = R.load("generators.generator_13_1")
    our_mod = importlib.import_module("generators.generator_13_1")
    torch.manual_seed(11)
    G_ref = ref_mod.Generator(32, 8).double()
    torch.manual_seed(11)
    G = our_mod.Generator(32, 8).double()
    sd_ref, sd = G_ref.state_dict(), G.state_dict()
    assert list(sd.keys()) == list(sd_ref.keys())
    for k, v in sd_ref.items():
        assert torch.equal(sd[k], v), k
    z = torch.randn(2, 32, 1, 1, dtype=torch.float64, generator=torch.Generator().manual_seed(12))

    def run(net):
        y = net(z)
        ps = [p for p in net.parameters() if p.requires_grad]
        return y.detach(), torch.autograd.grad((y * torch.linspace(-1, 1, y.numel(), dtype=y.dtype).reshape(y.shape)).sum(),
                                               ps, allow_unused=True)
    (y, g), (y_ref, g_ref) = run(G), run(G_ref)
    assert rel(y, y_ref) < 1e-12
    names = [n for n, p in G.named_parameters() if p.requires_grad]
    worst = {"logits": 0.0, "rest": 0.0}
    for n, a, b in zip(names, g, g_ref):
        assert (a is None) == (b is None), n
        if a is None or n.endswith("key.bias"):
            continue
        kind = "logits" if (".key." in n or ".query." in n) else "rest"
        worst[kind] = max(worst[kind], rel(a, b))
    assert worst["rest"] < 1e-9 and worst["logits"] < 1e-5, worst
