"""Timeline of the flat convolution kernel's pipeline roles in CTA 0 (ganb200_conv_flat_trace): where does a tile's time go?
python tools/flat_trace.py [C K R HW]"""
import ctypes
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gan_b200 import kernels as KN  # noqa: E402

C, K, R, HW = (int(v) for v in sys.argv[1:5]) if len(sys.argv) > 4 else (48, 48, 5, 64)
B = 256
cu = KN.CudaBackend()
cu.lib.ganb200_conv_flat_mode(2)
x = torch.randn(B, C, HW, HW, device="cuda")
w = torch.randn(K, C, R, R, device="cuda")
for _ in range(3):
    cu.conv_fprop(x, w, None, 1, R // 2, 1, 0.1, engine=2)
torch.cuda.synchronize()
buf = torch.zeros(3 * 32 * 4, dtype=torch.int64, device="cuda")
cu.lib.ganb200_conv_flat_trace(ctypes.c_void_p(buf.data_ptr()))
cu.conv_fprop(x, w, None, 1, R // 2, 1, 0.1, engine=2)
torch.cuda.synchronize()
cu.lib.ganb200_conv_flat_trace(None)
t = buf.cpu().view(3, 32, 4)
t0 = int(t[t > 0].min())
clk = 1.9e3   # cycles per us (nominal)
print(f"C{C} K{K} k{R} {HW}x{HW} B{B}: clock64 stamps of CTA 0, us from the first stamp (at {clk / 1e3:.1f} GHz)")
print("tile | producer: wait-free-stage .. issued | MMA: wait-acc .. got-acc .. got-A .. issued-all | epilogue: wait .. got .. done")
for i in range(14):
    f = lambda r, k: (int(t[r, i, k]) - t0) / clk if int(t[r, i, k]) else float("nan")
    print(f"{i:4d} | {f(0,0):8.2f} {f(0,1):8.2f} | {f(1,0):8.2f} {f(1,1):8.2f} {f(1,2):8.2f} {f(1,3):8.2f} | {f(2,0):8.2f} {f(2,1):8.2f} {f(2,2):8.2f}")
