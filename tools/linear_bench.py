"""Device time of the nn.Linear-shaped calls (1x1 filter on a 1x1 map) per route.  python tools/linear_bench.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gan_b200 import kernels as KN
cu = KN.CudaBackend()
N = 256
REPS = 10
for C, K in [(256, 256), (48, 48), (96, 96), (192, 192), (256, 48), (384, 384), (512, 512)]:
    x = torch.randn(N, C, 1, 1, device="cuda"); w = torch.randn(K, C, 1, 1, device="cuda")
    dy = torch.randn(N, K, 1, 1, device="cuda"); b = torch.randn(K, device="cuda")
    row = []
    for eng, name in ((0, "linear"), (2, "tc")):
        for op, fn in (("fprop", lambda: cu.conv_fprop(x, w, b, 1, 0, 0, 0.1, engine=eng)),
                       ("dgrad", lambda: cu.conv_dgrad(dy, w, (1, 1), 1, 0, 0, 0.1, engine=eng)),
                       ("wgrad", lambda: cu.conv_wgrad(x, dy, (1, 1), 1, 0, 0, 0.1, engine=eng))):
            fn(); torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                for _ in range(REPS):
                    fn()
            g.replay(); torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
            row.append(f"{name}/{op} {e0.elapsed_time(e1) / REPS * 1e3:6.1f} us")
    print(f"N={N} C={C:4d} K={K:4d}: " + "  ".join(row), flush=True)
