"""Fused modulated-conv backward (dgrad + fold/scale/reduce + wgrad) per shape, timed as a CUDA graph; and the one-launch
BatchNorm against the three-launch path on the 5x5 attention maps."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gan_b200 import kernels as KN
cu = KN.CudaBackend()
N = int(os.environ.get("B", "256"))


def timed(f, reps=5):
    f(); torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(reps):
            f()
    g.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e3


for C, K, R, HW in [(96, 96, 5, 64), (48, 48, 5, 64), (48, 48, 3, 64), (96, 96, 3, 64), (192, 192, 3, 32), (96, 96, 5, 32)]:
    x = torch.randn(N, C, HW, HW, device="cuda"); w = torch.randn(K, C, R, R, device="cuda")
    s = torch.rand(N, C, device="cuda") + 0.5; sinv = torch.rand(N, K, device="cuda") + 0.5
    gu = torch.randn(N, K, HW, HW, device="cuda")
    us = timed(lambda: cu.modconv_bwd(x, w, s, sinv, gu, R // 2, 0.02))
    print(f"modconv_bwd C{C} K{K} k{R} {HW}x{HW}: {us:8.1f} us", flush=True)

for C, HW in [(48, 5), (96, 5), (192, 5), (96, 3)]:
    x = torch.randn(N, C, HW, HW, device="cuda"); gy = torch.randn_like(x)
    gamma, beta = torch.rand(C, device="cuda") + 0.5, torch.randn(C, device="cuda")
    rm, rv = torch.zeros(C, device="cuda"), torch.ones(C, device="cuda")

    def fused():
        y, m, inv = cu.bn_small_fwd(x, gamma, beta, rm, rv, 1e-5, 0.1)
        cu.bn_small_bwd(gy, x, m, inv, gamma)

    def split():
        m, v = cu.bn_stats(x)
        y, inv = cu.bn_apply(x, m, v, gamma, beta, rm, rv, 1e-5, 0.1, N * HW * HW)
        s1, s2 = cu.bn_bwd_reduce(gy, x, m, inv)
        cu.bn_bwd_apply(gy, x, m, inv, gamma, s1, s2, N * HW * HW)

    print(f"batch norm fwd+bwd C{C} {HW}x{HW} N{N}: one launch each {timed(fused):7.1f} us, split {timed(split):7.1f} us", flush=True)
