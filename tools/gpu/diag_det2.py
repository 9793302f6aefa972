"""diagnostic: two eager ganR2 legs from the same seeds — where do they part?"""
import os, sys, random
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
from gan_b200 import rng
import test_graphed_step as TG
dev = torch.device("cuda:0")
B = TG.B
def cks(t):
    t = t.detach().contiguous()
    return int(t.view(torch.int32).to(torch.int64).sum().item()) if t.dtype == torch.float32 else int(t.sum().item())

def leg(images):
    rng._states.clear()
    T = TG._trainer(dev)
    TG._seed(1)
    log = []
    x = images.clone().requires_grad_()
    for it in range(3):
        d = T.discriminator_trainstep(x, B)
        log.append((it, "d_losses", tuple(float(v) for v in d)))
        for n, p in T.discriminator.named_parameters():
            log.append((it, "D.grad." + n, cks(p.grad) if p.grad is not None else None))
            log.append((it, "D.param." + n, cks(p)))
        log.append((it, "x.grad", cks(x.grad) if x.grad is not None else None))
        _, g = T.generator_trainstep(B)
        x.grad = None
        log.append((it, "g_loss", float(g)))
        for n, p in T.generator.named_parameters():
            log.append((it, "G.grad." + n, cks(p.grad) if p.grad is not None else None))
            log.append((it, "G.param." + n, cks(p)))
        for n, b in T.generator.named_buffers():
            log.append((it, "G.buf." + n, cks(b)))
    del T
    torch.cuda.empty_cache()
    return log

images = torch.randn(B, 3, 64, 64, generator=torch.Generator().manual_seed(3)).to(dev)
a = leg(images)
for trial in range(2):
    b = leg(images)
    bad = [(u, v) for u, v in zip(a, b) if u != v]
    print("trial", trial, "records", len(a), "mismatches", len(bad), flush=True)
    for u, v in bad[:12]:
        print("   ", u[:2], u[2] if not isinstance(u[2], int) else "", v[2] if not isinstance(v[2], int) else "", flush=True)
