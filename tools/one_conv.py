"""Runs a few launches of the dominant convolution shapes (for ncu --set full)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gan_b200 import kernels as KN
cu = KN.CudaBackend()
N = 256
for (C, K, R, HW) in [(48, 48, 3, 64), (96, 96, 5, 64), (512, 512, 3, 8)]:
    x = torch.randn(N, C, HW, HW, device="cuda"); w = torch.randn(K, C, R, R, device="cuda")
    dy = torch.randn(N, K, HW, HW, device="cuda")
    for _ in range(3):
        cu.conv_fprop(x, w, None, 1, R // 2, 1, 1.0, engine=2)
        cu.conv_dgrad(dy, w, (HW, HW), 1, R // 2, 1, 1.0, engine=2)
        cu.conv_wgrad(x, dy, (R, R), 1, R // 2, 1, 1.0, engine=2)
torch.cuda.synchronize()
print("ok")
