"""Bug-for-bug behaviours of the reference that a clean rewrite would silently change (SURVEY Appendix B / F-list), checked
on the drop-in modules on CPU.  (B1/B3/B4/B9 are covered by the golden tests: dead StyleConv biases without gradient,
frozen Smooth kernels among the parameters, running statistics moving in the no-grad forward.)"""
import pytest
import torch


def small_pair():
    from discriminators import discriminator_3_wgangp_progan as dmod
    from generators import generator_3_progan as gmod
    return gmod.Generator(1, 16, 4, 3), dmod.Discriminator(1, 4, 3)


def test_fixed_noise_counts_and_rng_position(ref_backend):
    """B12: 64 / 8 / 16 preview samples depending on the loop, drawn from the global torch stream in __init__."""
    import importlib
    want = {"gan": 64, "ganR2": 64, "wgan": 64, "wgansn": 64, "wgansnR2": 64, "wgangp": 64, "wgangpsnR2": 64,
            "wgangpR2": 8, "wganlazygp": 16, "wganlazygpR2": 16}
    for name, n in want.items():
        G, D = small_pair()
        torch.manual_seed(3)
        T = importlib.import_module("train." + name).Train([(torch.zeros(4, 3, 64, 64), torch.zeros(4))],
                                                           torch.device("cpu"), 1, 16, G, "g", D, "d")
        torch.manual_seed(3)
        assert tuple(T.fixed_noise.shape) == (n, 16, 1, 1) and torch.equal(T.fixed_noise, torch.randn(n, 16, 1, 1)), name
        # optimiser overrides: AdamW(betas 0.5, 0.999) in wgan / wgangp, Adam(0.5|0.0, 0.99) elsewhere
        g = T.optimizer_G.param_groups[0]
        d = T.optimizer_D.param_groups[0]
        if name in ("wgan", "wgangp"):
            assert g["decoupled"] and g["betas"] == (0.5, 0.999) and d["betas"] == (0.5, 0.999) and g["weight_decay"] == 1e-2
        else:
            assert not g["decoupled"] and g["betas"] == (0.5, 0.99) and d["betas"] == (0.0, 0.99)
        assert (g["lr"], d["lr"]) == (0.0001, 0.0004)


def test_batch_of_one_fails_in_batchnorm1d_like_the_reference(ref_backend):
    """B16: GeneratorStart squeezes ALL singleton dimensions; with one sample the mapping network receives a 1-D tensor,
    the equalised linear accepts it and BatchNorm1d raises torch's ValueError."""
    from generators import generator_13_5 as g
    m = g.MappingNetwork(256, 1)
    with pytest.raises(ValueError, match="expected 2D or 3D input"):
        m(torch.squeeze(torch.randn(1, 256, 1, 1)))
    assert m(torch.squeeze(torch.randn(2, 256, 1, 1))).shape == (2, 256)


def test_minibatch_stddev_needs_a_multiple_of_four(ref_backend):
    """F5 / B6: `assert x.shape[0] % group_size == 0` — the critic refuses batches that are not a multiple of 4."""
    from discriminators import discriminator_9_4 as d
    layer = d.MiniBatchStdDev()
    with pytest.raises(AssertionError):
        layer(torch.randn(6, 8, 4, 4))
    y = layer(torch.randn(8, 8, 4, 4))
    assert y.shape == (8, 9, 4, 4) and float(y[:, 8].std()) == 0.0  # one scalar for the whole batch


def test_weights_init_crashes_on_equalized_layers_like_the_reference():
    """B13: `weights_init` matches class names containing 'Conv'; EqualizedConv2d's `.weight` is a module, so applying
    the DCGAN init to the final discriminator fails in the reference (it is only ever applied to G2 / D2)."""
    from discriminators import discriminator_9_4 as d
    from units import Weight_Initialization
    layer = d.EqualizedConv2d(4, 4, 3, 1)
    with pytest.raises(AttributeError):
        layer.apply(Weight_Initialization.weights_init)


def test_attribute_spellings_differ_between_g_and_d(ref_backend):
    """B2: generator EqualizedWeight: `.weights` / `.scale`; discriminator EqualizedWeight: `.weight` / `.c`."""
    from discriminators import discriminator_9_4 as d
    from generators import generator_13_5 as g
    gw, dw = g.EqualizedWeight([4, 3]), d.EqualizedWeight([4, 3])
    assert list(gw.state_dict()) == ["weights"] and hasattr(gw, "scale") and not hasattr(gw, "c")
    assert list(dw.state_dict()) == ["weight"] and hasattr(dw, "c") and not hasattr(dw, "scale")
    assert torch.allclose(gw(), gw.weights * gw.scale) and torch.allclose(dw(), dw.weight * dw.c)
