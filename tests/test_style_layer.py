"""CPU, float64: the fused StyleConv layer (gan_b200.ops.StyleLayerFn: modconv + in-kernel noise + PReLU, 2 saved
activations) against the unfused composition (modulated_conv2d -> noise_add -> prelu) fed the same noise, forward
and every gradient; then the whole generator_13_5 in fused vs unfused mode."""
import os
import random

import numpy as np
import pytest
import torch

from gan_b200 import kernels as KN
from gan_b200 import ops, rng

D = torch.float64


@pytest.mark.parametrize("k,noise,act", [(3, True, True), (5, True, True), (1, False, True), (3, False, False),
                                          (3, True, False)])
def test_fused_style_layer_matches_unfused(ref_backend, k, noise, act):
    torch.manual_seed(0)
    B, C, Ko, H = 3, 5, 6, 4
    x = torch.randn(B, C, H, H, dtype=D, requires_grad=True)
    w = torch.randn(Ko, C, k, k, dtype=D, requires_grad=True)
    s = (torch.randn(B, C, dtype=D) + 1.5).requires_grad_()
    ns = (torch.rand(Ko, dtype=D) * 0.3).requires_grad_() if noise else None
    slope = (torch.rand(Ko, dtype=D) * 0.5).requires_grad_() if act else None
    state = torch.tensor([1234, 7], dtype=torch.int64)
    alpha, pad = 0.31, (k - 1) // 2
    out = ops.style_layer(x, w, s, ns, slope, pad, alpha, True, 1e-8, None, state, 9)
    y = ops.modulated_conv2d(x, w, s, pad, alpha, True, 1e-8)
    if noise:
        z = ref_backend.noise_fill(tuple(y.shape), state, 9, y.device).to(D)
        y = ops.noise_add(y, ns, z)
    if act:
        y = ops.prelu(y, slope)
    assert torch.allclose(out, y, atol=1e-12)
    g = torch.randn_like(out)
    ins = [t for t in (x, w, s, ns, slope) if t is not None]
    got = torch.autograd.grad(out, ins, g)
    want = torch.autograd.grad(y, ins, g)
    for a, b in zip(got, want):
        assert torch.allclose(a, b, atol=1e-10), (a - b).abs().max()


def test_generator_fused_equals_unfused_noise_path(ref_backend):
    """generator_13_5 forward + all parameter gradients: fused layers (philox) == unfused ops fed the same Philox noise."""
    from generators import generator_13_5
    res = {}
    for mode in ("philox", "philox_unfused"):
        prev = rng.noise_mode(mode)
        rng._states.clear()
        os.environ["GANB200_STYLE_BANK"] = "1" if mode == "philox" else "0"  # the fused leg also runs banked styles
        try:
            torch.manual_seed(0); np.random.seed(0); random.seed(0)
            G = generator_13_5.Generator(256)
            torch.manual_seed(1)
            z = torch.randn(4, 256, 1, 1)
            img = G(z)
            img.square().mean().backward()
            res[mode] = (img.detach(), {n: p.grad for n, p in G.named_parameters()})
        finally:
            rng.noise_mode(prev)
            os.environ.pop("GANB200_STYLE_BANK", None)
    a, b = res["philox"], res["philox_unfused"]
    assert float((a[0] - b[0]).norm() / b[0].norm()) < 1e-4
    worst, num, den = 0.0, 0.0, 0.0
    for n, gb in b[1].items():
        ga = a[1][n]
        assert (ga is None) == (gb is None), n
        if gb is None:
            continue
        num += float((ga - gb).norm() ** 2)
        den += float(gb.norm() ** 2)
        # biases feeding a BatchNorm have an analytically zero gradient: what is left is rounding noise whose norm
        # (2e-4 for block0.mapping_network.net.6.bias) depends on the noise realisation -> floor at 1e-3
        if gb.norm() > 1e-3:
            worst = max(worst, float((ga - gb).norm() / gb.norm()))
    assert worst < 5e-2, worst  # fp32, ~100 layers and B=4 batch norms: see tests/test_golden_final_pair.py
    assert (num / den) ** 0.5 < 5e-3, (num / den) ** 0.5  # all 8 766 gradient tensors together (measured 4.5e-4)


def test_style_bank_is_the_same_computation(ref_backend, monkeypatch):
    """GANB200_STYLE_BANK=1: every to_style(w) evaluated right after the mapping network instead of in front of its
    convolution — the image is bit-identical (same kernels on the same operands).  Gradients with the bank on are covered
    by test_generator_fused_equals_unfused_noise_path, whose fused leg runs banked."""
    from generators import generator_13_5
    torch.manual_seed(0); np.random.seed(0); random.seed(0)
    G = generator_13_5.Generator(256)
    assert len(G.block0.style_bank.layers) == 519
    z = torch.randn(2, 256, 1, 1)
    out = {}
    for bank in ("0", "1"):
        monkeypatch.setenv("GANB200_STYLE_BANK", bank)
        rng._states.clear()
        torch.manual_seed(1)
        with torch.no_grad():
            out[bank] = G(z)  # train mode: batch statistics, so the second pass sees the same arithmetic
        assert all(m._style is None for m in G.block0.style_bank.layers)  # every banked style was consumed
    assert torch.equal(out["0"], out["1"])


def _mini_bank(dtype, seed=0):
    """a few to_style MLPs of the generator (ragged widths), in train mode"""
    from generators.generator_13_5 import Conv2dWeightModulate
    torch.manual_seed(seed)
    mods = [Conv2dWeightModulate(32, c, 8, 3).to_style for c in (3, 48, 33, 70)]
    for m in mods:
        m.to(dtype)
        for p in m.parameters():
            p.data.add_(0.1 * torch.randn_like(p))  # not the symmetric initial state (gamma 1, beta 0, slope .25)
    return mods


def test_grouped_style_bank_equals_layer_by_layer_fp64(ref_backend):
    """gan_b200.group.GroupedStyles (6 grouped launches forward, 11 backward) against evaluating every to_style module by
    itself: styles, every parameter gradient, d/dw and the BatchNorm running statistics, in float64."""
    from gan_b200.group import GroupedStyles
    B = 6
    w = torch.randn(B, 32, dtype=torch.float64, requires_grad=True)
    gs = [torch.randn(B, c, dtype=torch.float64) for c in (3, 48, 33, 70)]
    res = {}
    for leg in ("layers", "grouped"):
        mods = _mini_bank(torch.float64)
        w_ = w.detach().clone().requires_grad_()
        if leg == "grouped":
            bank = GroupedStyles(mods)
            assert bank.eligible()
            outs = bank(w_)
        else:
            outs = [m(w_) for m in mods]
        torch.autograd.backward(list(outs), gs)
        res[leg] = ([o.detach() for o in outs], w_.grad, [p.grad for m in mods for p in m.parameters()],
                    [b.clone() for m in mods for b in m.buffers()])
    a, b = res["grouped"], res["layers"]
    for u, v in zip(a[0], b[0]):
        assert torch.allclose(u, v, rtol=1e-11, atol=1e-12)
    assert torch.allclose(a[1], b[1], rtol=1e-10, atol=1e-12)
    assert len(a[2]) == len(b[2]) == 4 * 9
    for u, v in zip(a[2], b[2]):
        assert u.shape == v.shape and torch.allclose(u, v, rtol=1e-9, atol=1e-11), (u - v).abs().max()
    for u, v in zip(a[3], b[3]):
        assert torch.allclose(u.double(), v.double(), rtol=1e-11, atol=1e-12)


def test_grouped_style_bank_steps_aside_for_hooks_and_eval_mode(ref_backend):
    from gan_b200.group import GroupedStyles
    mods = _mini_bank(torch.float32)
    assert GroupedStyles(mods).eligible()
    h = mods[1][2].register_forward_hook(lambda *a: None)
    assert not GroupedStyles(mods).eligible()   # somebody observes a per-layer activation
    h.remove()
    mods[2].eval()
    assert not GroupedStyles(mods).eligible()   # eval-mode BatchNorm uses the running statistics
