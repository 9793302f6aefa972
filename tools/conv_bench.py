"""Times the tcgen05 conv engine on the dominant shapes of the final pair, halo kernel vs per-tap kernel.
Each call is captured in a CUDA graph of REPS repetitions so that host launch overhead does not enter.
Run on the GPU box:  python tools/conv_bench.py [B]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gan_b200 import kernels as KN  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
HALO = int(sys.argv[2]) if len(sys.argv) > 2 else 1   # 2: halo kernel also where the heuristic prefers per-tap
REPS = 5
# C, HW, K, R, stride, pad, pad_mode
SHAPES = [(96, 64, 96, 5, 1, 2, 1), (48, 64, 48, 5, 1, 2, 1), (48, 64, 48, 3, 1, 1, 1), (96, 64, 96, 3, 1, 1, 1),
          (64, 64, 64, 3, 1, 1, 1), (128, 32, 128, 3, 1, 1, 1), (256, 16, 256, 3, 1, 1, 1), (96, 32, 96, 5, 1, 2, 1),
          (390, 64, 96, 1, 1, 0, 1), (108, 64, 108, 5, 1, 2, 1), (256, 32, 256, 3, 2, 1, 1), (512, 8, 512, 3, 1, 1, 1)]


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


def main():
    cu = KN.CudaBackend()
    dev = torch.device("cuda")
    print(f"B={B}; ms per call (pack + kernel [+ fold]) and TFLOP/s")
    print(f"{'shape':34s} {'op':6s} {'per-tap ms':>10s} {'halo ms':>9s} {'speedup':>8s} {'halo TF/s':>10s} {'max|diff|/rms':>14s}")
    for C, HW, K, R, stride, pad, pm in SHAPES:
        x = torch.randn(B, C, HW, HW, device=dev)
        w = torch.randn(K, C, R, R, device=dev)
        P, Q = KN.out_hw(HW, HW, R, R, stride, pad)
        dy = torch.randn(B, K, P, Q, device=dev)
        flops = 2.0 * B * K * P * Q * C * R * R
        for op, fn in (("fprop", lambda: cu.conv_fprop(x, w, None, stride, pad, pm, 0.1, engine=2)),
                       ("dgrad", lambda: cu.conv_dgrad(dy, w, (HW, HW), stride, pad, pm, 0.1, engine=2)),
                       ("wgrad", lambda: cu.conv_wgrad(x, dy, (R, R), stride, pad, pm, 0.1, engine=2))):
            res = {}
            for mode in (0, 1):
                cu.lib.ganb200_conv_halo_mode(HALO if mode else 0)
                out = fn()
                res[mode] = (timed(fn), out)
            cu.lib.ganb200_conv_halo_mode(1)
            diff = ((res[0][1] - res[1][1]).abs().max() / res[0][1].pow(2).mean().sqrt()).item()
            t0, t1 = res[0][0], res[1][0]
            print(f"C{C:4d} {HW:2d}x{HW:<2d} K{K:4d} R{R} s{stride} p{pad} pm{pm}      {op:6s} {t0:10.3f} {t1:9.3f} "
                  f"{t0 / t1:8.2f} {flops / t1 / 1e9:10.1f} {diff:14.2e}", flush=True)
        del x, w, dy


if __name__ == "__main__":
    main()
