"""diagnostic: grouped style bank on CUDA vs oracle and vs per-layer kernels, per-tensor errors"""
import copy, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
from gan_b200 import kernels as KN
from gan_b200.group import GroupedStyles
from oracle.kernels_ref import RefBackend
from test_gpu_kernels import _bank_modules, rel
cu, ref = KN.CudaBackend(), RefBackend()
names = ["W1", "b1", "g1", "be1", "a1", "W2", "b2", "g2", "be2"]
for B in (4, 32):
    cpu_mods = _bank_modules("cpu")
    gpu_mods = [copy.deepcopy(m).cuda() for m in cpu_mods]
    lay_mods = [copy.deepcopy(m).cuda() for m in cpu_mods]
    torch.manual_seed(3)
    w = torch.randn(B, 64)
    gs = [torch.randn(B, m[1].weight.raw.shape[0]) for m in cpu_mods]
    out = {}
    for leg, mods, dev in (("ref", cpu_mods, "cpu"), ("cuda", gpu_mods, "cuda"), ("layers", lay_mods, "cuda")):
        prev = KN.set_backend(ref if leg == "ref" else cu)
        w_ = w.detach().clone().to(dev).requires_grad_()
        outs = GroupedStyles(mods)(w_) if leg != "layers" else [m(w_) for m in mods]
        torch.autograd.backward(list(outs), [t.to(dev) for t in gs])
        out[leg] = ([o.detach().cpu() for o in outs], w_.grad.cpu(), [p.grad.cpu() for m in mods for p in m.parameters()],
                    [b.detach().cpu() for m in mods for b in m.buffers()])
        KN.set_backend(prev)
    for other in ("ref", "layers"):
        a, b = out["cuda"], out[other]
        print(f"B={B} cuda-grouped vs {other}: outs", [f"{rel(u, v):.1e}" for u, v in zip(a[0], b[0])], "gw", f"{rel(a[1], b[1]):.1e}")
        for i, (u, v) in enumerate(zip(a[2], b[2])):
            print(f"   layer {i // 9} {names[i % 9]:4s} rel {rel(u, v):.2e} |ref| {v.norm().item():.3e} |diff| {(u - v).norm().item():.3e}")
        print("   buffers", [f"{rel(u.double(), v.double()):.1e}" for u, v in zip(a[3], b[3])])
