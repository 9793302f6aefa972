"""Runs a few launches of one convolution shape on the tcgen05 engine (for ncu --set full).
python tools/one_conv.py C K R HW [halo_mode] [op]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gan_b200 import kernels as KN
cu = KN.CudaBackend()
C, K, R, HW = (int(v) for v in sys.argv[1:5])
mode = int(sys.argv[5]) if len(sys.argv) > 5 else 1
op = sys.argv[6] if len(sys.argv) > 6 else "fprop"
N = int(os.environ.get("B", "256"))
cu.lib.ganb200_conv_halo_mode(mode)
x = torch.randn(N, C, HW, HW, device="cuda"); w = torch.randn(K, C, R, R, device="cuda")
dy = torch.randn(N, K, HW, HW, device="cuda")
for _ in range(3):
    if op == "fprop":
        cu.conv_fprop(x, w, None, 1, R // 2, 1, 1.0, engine=2)
    elif op == "dgrad":
        cu.conv_dgrad(dy, w, (HW, HW), 1, R // 2, 1, 1.0, engine=2)
    else:
        cu.conv_wgrad(x, dy, (R, R), 1, R // 2, 1, 1.0, engine=2)
torch.cuda.synchronize()
print("ok")
