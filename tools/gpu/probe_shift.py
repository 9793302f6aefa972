"""tcgen05 SWIZZLE_128B operand read from a start address shifted by single 128-byte rows: which descriptor form works?"""
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
for mode in (0, 1):
    res = []
    for shift in (0, 8, 1, 2, 3, 4, 5, 7, 9, 13, 68, 69, 127):
        out.zero_()
        _lib.check(lib.ganb200_tc_shift_probe(ctypes.c_void_p(a.data_ptr()), ctypes.c_void_p(b.data_ptr()),
                                              ctypes.c_void_p(out.data_ptr()), rows, n, shift, mode, st))
        torch.cuda.synchronize()
        want = a[shift:shift + 128].float() @ b.float().t()
        res.append((shift, float((out - want).abs().max() / want.abs().max())))
    print("base_offset_mode", mode, " ".join(f"{s}:{e:.1e}" for s, e in res), flush=True)
