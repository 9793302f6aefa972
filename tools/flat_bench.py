"""Forward convolutions of the final pair on the tcgen05 engine: flat kernel vs halo / per-tap kernels (same packed
operands; CUDA graph of REPS calls each).  python tools/flat_bench.py [B]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gan_b200 import kernels as KN  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
REPS = 5
SHAPES = [(96, 64, 96, 5, 2), (48, 64, 48, 5, 2), (48, 64, 48, 3, 1), (96, 64, 96, 3, 1), (64, 64, 64, 3, 1),
          (128, 32, 128, 3, 1), (96, 32, 96, 5, 2), (192, 32, 192, 3, 1), (390, 64, 96, 1, 0), (108, 64, 108, 5, 2),
          (48, 64, 54, 3, 1), (108, 64, 3, 5, 2)]


def timed(fn):
    fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(REPS):
            fn()
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / REPS


cu = KN.CudaBackend()
print(f"B={B}; fprop, ms per call (pack + kernel) and TFLOP/s")
print(f"{'shape':28s} {'old ms':>8s} {'flat ms':>8s} {'speedup':>8s} {'flat TF/s':>10s} {'identical':>10s}")
for C, HW, K, R, pad in SHAPES:
    x = torch.randn(B, C, HW, HW, device="cuda")
    w = torch.randn(K, C, R, R, device="cuda")
    flops = 2.0 * B * K * HW * HW * C * R * R
    fn = lambda: cu.conv_fprop(x, w, None, 1, pad, 1, 0.1, engine=2)
    res = {}
    for mode in (0, 1):
        cu.lib.ganb200_conv_flat_mode(2 * mode)
        out = fn()
        res[mode] = (timed(fn), out)
    cu.lib.ganb200_conv_flat_mode(1)
    t0, t1 = res[0][0], res[1][0]
    print(f"C{C:4d} {HW:2d}x{HW:<2d} K{K:4d} k{R}          {t0:8.3f} {t1:8.3f} {t0 / t1:8.2f} {flops / t1 / 1e9:10.1f} "
          f"{str(bool(torch.equal(res[0][1], res[1][1]))):>10s}", flush=True)
