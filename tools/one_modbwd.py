"""One fused modulated-conv backward (96->96 5x5 @64x64, N=256) for ncu."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gan_b200 import kernels as KN
cu = KN.CudaBackend()
N, C, K, R, HW = 256, 96, 96, 5, 64
x = torch.randn(N, C, HW, HW, device="cuda"); w = torch.randn(K, C, R, R, device="cuda")
s = torch.rand(N, C, device="cuda") + 0.5; sinv = torch.rand(N, K, device="cuda") + 0.5
gu = torch.randn(N, K, HW, HW, device="cuda")
for _ in range(2):
    out = cu.modconv_bwd(x, w, s, sinv, gu, R // 2, 0.02)
torch.cuda.synchronize()
print("ok", out[0].shape)
