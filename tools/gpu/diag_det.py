"""diagnostic: which module's output / gradient is the first that differs between two identical runs?"""
import os, sys, random
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
from gan_b200 import rng
from generators import generator_13_5
from discriminators import discriminator_9_4
dev = torch.device("cuda:0")
B = int(os.environ.get("B", "8"))
def seed(s):
    torch.manual_seed(s); np.random.seed(s); random.seed(s)
def cks(t):
    t = t.detach().contiguous()
    return int(t.view(torch.int32).to(torch.int64).sum().item()) if t.dtype == torch.float32 else int(t.sum().item())

def instrument(net, log):
    hs = []
    for name, m in net.named_modules():
        def fwd(mod, inp, out, name=name):
            outs = out if isinstance(out, (tuple, list)) else (out,)
            for i, o in enumerate(outs):
                if torch.is_tensor(o):
                    log.append(("F", name, i, cks(o)))
                    if o.requires_grad:
                        o.register_hook(lambda g, name=name, i=i: log.append(("B", name, i, cks(g))))
        hs.append(m.register_forward_hook(fwd))
    return hs

def run(which):
    seed(0)
    G = generator_13_5.Generator(256).to(dev)
    D = discriminator_9_4.Discriminator().to(dev)
    rng._states.clear(); seed(1)
    log = []
    z = torch.randn(B, 256, 1, 1, device=dev)
    x = torch.randn(B, 3, 64, 64, device=dev, requires_grad=True)
    if which == "G":
        instrument(G, log)
        y = G(z)
        y.square().mean().backward()
        grads = {n: cks(p.grad) for n, p in G.named_parameters() if p.grad is not None}
    else:
        instrument(D, log)
        out = D(x)
        (g,) = torch.autograd.grad(out.sum(), x, create_graph=True)
        pen = g.square().sum() / B
        (out.mean() + pen).backward()
        grads = {n: cks(p.grad) for n, p in D.named_parameters() if p.grad is not None}
    torch.cuda.synchronize()
    return log, grads

for which in ("D", "G"):
    a_log, a_gr = run(which)
    for trial in range(3):
        b_log, b_gr = run(which)
        assert len(a_log) == len(b_log), (len(a_log), len(b_log))
        bad = [(u, v) for u, v in zip(a_log, b_log) if u != v]
        badg = [n for n in a_gr if a_gr[n] != b_gr[n]]
        print(which, "trial", trial, "records", len(a_log), "mismatching records", len(bad), "mismatching grads", len(badg), flush=True)
        for u, v in bad[:6]:
            print("   first mismatches:", u[:3], flush=True)
        for n in badg[:6]:
            print("   grad:", n, flush=True)
