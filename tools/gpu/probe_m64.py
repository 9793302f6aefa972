"""where does an M = 64 tcgen05.mma (cta_group::1) put its 64 accumulator rows in TMEM?"""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from gan_b200 import _lib
lib = _lib.load()
torch.manual_seed(0)
rows, n = 256, 64
a = torch.randn(rows, 64, device="cuda").to(torch.bfloat16)
b = torch.randn(n, 64, device="cuda").to(torch.bfloat16)
out = torch.empty(128, n, device="cuda")
st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
want = a[:128].float() @ b.float().t()
for m in (128, 64):
    out.fill_(-7.0)
    _lib.check(lib.ganb200_tc_shift_probe_m(ctypes.c_void_p(a.data_ptr()), ctypes.c_void_p(b.data_ptr()),
                                            ctypes.c_void_p(out.data_ptr()), rows, n, 0, 0, m, st))
    torch.cuda.synchronize()
    # for each TMEM lane: which expected row does it hold?
    d = (out[:, None, :] - want[None, :, :]).abs().amax(dim=2)    # [lane][row]
    best = d.argmin(dim=1)
    ok = d.min(dim=1).values < 1e-3
    print(f"M={m}: lanes holding a result row:", [(int(l), int(best[l])) for l in range(128) if ok[l]])
