"""Per-layer parity of the final pair on the bf16 tcgen05 engine (BASELINE.json north_star: "per-layer activations ...
rel. L2 <= 1e-2 for bf16 tensor-core paths").

End to end the bf16 engine accumulates its operand rounding through 40 (D) resp. ~150 (G) convolutions and, at small
batches, through BatchNorm over 4-8 samples; the whole-network goldens (tests/test_golden_final_pair.py) therefore carry
looser bounds.  Here every block is measured by itself: the CPU restatement of the step (oracle/kernels_ref.py, pinned to
the unmodified reference by the CPU golden tests) runs the whole network once and records the input and output of every
`DiscriminatorBlock` / stem / head of discriminator_9_4 and of every `StyleBlock`, `ToRGB` and `SKConvT` of generator_13_5;
the GPU copy of each block is then fed the RECORDED input and compared with the recorded output — fp32 engine <= 2e-4,
bf16 engine <= 1e-2 per block.  Noise: the reference's torch.randn path on the CPU generator, re-seeded per block so that
both sides draw the same values."""
import copy
import os
import random
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

pytestmark = pytest.mark.gpu
B = 8


def rel(a, b):
    a, b = a.detach().cpu().double(), b.detach().cpu().double()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def _record(net, types, run):
    """-> [(name, module, args, output)] for every module of `types`, with the CPU generator seeded per module"""
    recs, hooks = [], []
    index = {}

    def pre(mod, args):
        torch.manual_seed(1000 + index[mod])

    def post(mod, args, out):
        recs.append((index[mod], tuple(a.detach().clone() if torch.is_tensor(a) else a for a in args),
                     out.detach().clone() if torch.is_tensor(out) else tuple(o.detach().clone() for o in out)))

    names = {}
    for i, (n, m) in enumerate(net.named_modules()):
        if isinstance(m, types):
            index[m] = i
            names[i] = n
            hooks.append(m.register_forward_pre_hook(pre))
            hooks.append(m.register_forward_hook(post))
    with torch.no_grad():
        run()
    for h in hooks:
        h.remove()
    return recs, names


def _blockwise(net_cpu, types, run_cpu, precision, tol, monkeypatch, stride=1):
    from gan_b200 import kernels, rng
    from oracle.kernels_ref import RefBackend
    monkeypatch.setenv("GANB200_PRECISION", precision)
    monkeypatch.setenv("GANB200_STYLE_BANK", "0")   # per-layer styles: every StyleBlock is self-contained
    prev_mode, prev_host = rng.noise_mode("torch"), rng.host_stream(True)
    prev = kernels.set_backend(RefBackend())
    try:
        recs, names = _record(net_cpu, types, run_cpu)
    finally:
        kernels.set_backend(prev)
    worst = {}
    try:
        net_gpu = copy.deepcopy(net_cpu).cuda()
        gpu_mods = dict(net_gpu.named_modules())
        cpu_name = {id(m): n for n, m in net_cpu.named_modules()}

        def to_gpu(a):   # tensors move; a module argument (the activation a block hands to its last layer) is mapped
            if torch.is_tensor(a):
                return a.cuda()
            return gpu_mods[cpu_name[id(a)]] if isinstance(a, torch.nn.Module) else a

        for idx, args, want in recs[::stride]:
            mod = gpu_mods[names[idx]]
            torch.manual_seed(1000 + idx)
            with torch.no_grad():
                got = mod(*[to_gpu(a) for a in args])
            if torch.is_tensor(want):
                worst[names[idx]] = rel(got, want)
            else:
                worst[names[idx]] = max(rel(g, w) for g, w in zip(got, want))
    finally:
        rng.noise_mode(prev_mode)
        rng.host_stream(prev_host)
    bad = {k: v for k, v in worst.items() if not v <= tol}
    assert recs and not bad, (len(recs), sorted(bad.items(), key=lambda kv: -kv[1])[:8])
    return worst


@pytest.mark.parametrize("precision,tol", [("fp32", 2e-4), ("auto", 1e-2)])
def test_discriminator_9_4_blockwise(monkeypatch, precision, tol):
    from discriminators import discriminator_9_4 as dm
    from gan_b200 import layers as L
    torch.manual_seed(0); np.random.seed(0); random.seed(0)
    D = dm.Discriminator()
    x = torch.randn(B, 3, 64, 64, generator=torch.Generator().manual_seed(5))
    types = (dm.DiscriminatorBlock, dm.MiniBatchStdDev, dm.EqualizedLinear)
    worst = _blockwise(D, types, lambda: D(x), precision, tol, monkeypatch)
    assert len(worst) >= 15 + 1 + 2
    print(precision, "D9_4 worst block", max(worst.items(), key=lambda kv: kv[1]))


@pytest.mark.parametrize("precision,tol", [("fp32", 2e-4), ("auto", 1e-2)])
def test_generator_13_5_blockwise(monkeypatch, precision, tol):
    from generators import generator_13_5 as gm
    torch.manual_seed(0); np.random.seed(0); random.seed(0)
    G = gm.Generator(256)
    z = torch.randn(B, 256, 1, 1, generator=torch.Generator().manual_seed(5))
    types = (gm.StyleBlock, gm.ToRGB, gm.SKConvT)
    worst = _blockwise(G, types, lambda: G(z), precision, tol, monkeypatch)
    assert len(worst) >= 133 + 10 + 8
    print(precision, "G13_5 worst block", max(worst.items(), key=lambda kv: kv[1]))
