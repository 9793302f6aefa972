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

    for a, b in zip(run(D, torch.float64), run(D_ref, torch.float64)):
        assert rel(a, b) < 1e-12
    for a, b in zip(run(D, torch.float32), run(D_ref, torch.float32)):
        assert rel(a, b) < 1e-5


# (generator_5 / 8 / 9 / 10: structure, initialisation and a whole iteration are compared in tests/test_other_configs.py)
@pytest.mark.parametrize("name", ["generator_4", "generator_7", "generator_10_2"])
def test_tree_generators_equal_reference(ref_backend, name):
    """generator_7 / generator_8: the same tree with BatchNorm + scalar-PReLU bottlenecks (they subclass the critic's
    classes); generator_4 / generator_5: InstanceNorm trees on BatchNorm stems; generator_9 / 10 / 10_2: self-attention
    trees (10_2 with run-time-scaled equalised convolutions).  Identical state_dict; for generator_7 and generator_10_2
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
    if name not in ("generator_7", "generator_10_2"):
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
        assert rel(a, b) < (1e-11 if name == "generator_7" else 1e-8)
