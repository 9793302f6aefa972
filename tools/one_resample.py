"""A few launches of one resampling operator (for ncu --set full).  python tools/one_resample.py smooth3|down|up|avg"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gan_b200 import kernels as KN
from gan_b200 import resample as RS
cu = KN.CudaBackend()
which = sys.argv[1] if len(sys.argv) > 1 else "smooth3"
op, C = {"smooth3": (RS.smooth3(64), 64), "down": (RS.smooth_then_bicubic_down2(64), 64),
         "up": (RS.bicubic_up2_then_smooth(32), 108), "avg": (RS.adaptive_avg(64, 5), 96)}[which]
x = torch.randn(256, C, op.n_in, op.n_in, device="cuda")
for _ in range(3):
    cu.resample2d(x, op, op)
torch.cuda.synchronize()
print("ok")
