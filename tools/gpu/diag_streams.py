import os, sys, random
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
from gan_b200 import rng
from generators import generator_13_5
dev = torch.device("cuda:0")
torch.manual_seed(0)
G = generator_13_5.Generator(256).to(dev)
z = torch.randn(8, 256, 1, 1, device=dev)
rel = lambda a, b: float((a - b).norm() / b.norm())
def run(grad):
    rng._states.clear(); torch.manual_seed(1)
    if grad:
        G.zero_grad()
        out = G(z); out.square().mean().backward()
        return out.detach().clone(), {n: p.grad.clone() for n, p in G.named_parameters() if p.grad is not None}
    with torch.no_grad():
        return G(z).clone(), None
for grad in (False, True):
    os.environ["GANB200_BRANCH_STREAMS"] = "0"; a = run(grad); a2 = run(grad)
    os.environ["GANB200_BRANCH_STREAMS"] = "1"; b = run(grad)
    torch.cuda.synchronize()
    print("grad", grad, "out: serial vs serial", rel(a2[0], a[0]), "streams vs serial", rel(b[0], a[0]), flush=True)
    if grad:
        num = sum(float((b[1][n] - a[1][n]).norm() ** 2) for n in a[1]); den = sum(float(a[1][n].norm() ** 2) for n in a[1])
        num2 = sum(float((a2[1][n] - a[1][n]).norm() ** 2) for n in a[1])
        print("   grads agg: serial vs serial", (num2 / den) ** .5, "streams vs serial", (num / den) ** .5, flush=True)
# capture tests
os.environ["GANB200_BRANCH_STREAMS"] = "1"
for grad in (False, True):
    try:
        s = torch.cuda.Stream(); s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            run(grad)
        torch.cuda.current_stream().wait_stream(s); torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, capture_error_mode="thread_local"):
            if grad:
                G.zero_grad(); out = G(z); out.square().mean().backward()
            else:
                with torch.no_grad():
                    out = G(z)
        g.replay(); torch.cuda.synchronize()
        print("capture grad", grad, "ok")
    except Exception as e:
        print("capture grad", grad, "FAILED", str(e)[:300])
        torch.cuda.synchronize()
