"""diagnostic: does a captured G forward (fused Philox noise) equal the eager one?"""
import os, sys, random
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
from gan_b200 import rng
from generators import generator_13_5
dev = torch.device("cuda:0")
def seed(s):
    torch.manual_seed(s); np.random.seed(s); random.seed(s)
seed(0)
G = generator_13_5.Generator(256).to(dev)
z = torch.randn(8, 256, 1, 1, generator=torch.Generator().manual_seed(3)).to(dev)
rel = lambda a, b: float((a - b).norm() / b.norm())
for mode in ("philox", "torch"):
    rng.noise_mode(mode)
    rng._states.clear(); seed(1)
    with torch.no_grad():
        e1 = G(z).clone(); e2 = G(z).clone()
    rng._states.clear(); seed(1)
    with torch.no_grad():
        s = torch.cuda.Stream(); s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            G(z); G(z)
        torch.cuda.current_stream().wait_stream(s); torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, capture_error_mode="thread_local"):
            out = G(z)
    for k, v in rng._states.items():
        v[1:].zero_()
    seed(1)
    g.replay(); r1 = out.clone(); g.replay(); r2 = out.clone()
    torch.cuda.synchronize()
    print(mode, "eager1 vs replay1", rel(r1, e1), "eager2 vs replay2", rel(r2, e2), "eager1 vs eager2", rel(e2, e1), flush=True)
    # BN running stats drift between the legs does not enter (train-mode BN uses batch statistics)
