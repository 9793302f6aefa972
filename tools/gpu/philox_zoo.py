"""generator_12 / generator_13 with the counter-based (philox) noise path on the GPU: forward + backward, finite, and two
forwards from the same generator state give the same image.  Usage: python tools/gpu/philox_zoo.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from gan_b200 import rng  # noqa: E402
from generators import generator_12, generator_13  # noqa: E402

for mod, kw in ((generator_12, dict(planes=8)), (generator_13, dict(planes=8))):
    torch.manual_seed(0)
    G = mod.Generator(32, **kw).cuda()
    z = torch.randn(4, 32, 1, 1, device="cuda")
    assert rng.philox_noise()
    y = G(z)
    y.square().mean().backward()
    grads = [p.grad for p in G.parameters() if p.grad is not None]
    ok = bool(torch.isfinite(y).all()) and all(bool(torch.isfinite(g).all()) for g in grads)
    print(mod.__name__, tuple(y.shape), "finite" if ok else "NOT FINITE", "grads", len(grads))
    assert ok
