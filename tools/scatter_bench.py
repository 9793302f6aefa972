"""dgrad / fused modconv backward: scatter epilogue (mode 1) vs padded domain + fold pass (mode 0), per shape.
python tools/scatter_bench.py [B]"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gan_b200 import kernels as KN

B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
REPS = 5
# C, HW, K, R, stride
SHAPES = [(96, 64, 96, 5, 1), (48, 64, 48, 5, 1), (48, 64, 48, 3, 1), (96, 64, 96, 3, 1), (390, 64, 96, 1, 1),
          (96, 32, 96, 5, 1), (192, 16, 192, 5, 1), (192, 16, 192, 3, 1), (384, 8, 384, 5, 1),
          (64, 64, 64, 3, 1), (128, 32, 128, 3, 1), (256, 16, 256, 3, 1), (512, 8, 512, 3, 1), (1025, 4, 1025, 3, 1),
          (128, 64, 128, 3, 2)]


def timed(fn):
    fn(); torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(REPS):
            fn()
    g.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / REPS


cu = KN.CudaBackend()
print(f"B={B}  ms per call: mode0 (fold pass) / mode1 (scatter epilogue)")
for C, HW, K, R, stride in SHAPES:
    pad = R // 2
    P, Q = KN.out_hw(HW, HW, R, R, stride, pad)
    x = torch.randn(B, C, HW, HW, device="cuda"); w = torch.randn(K, C, R, R, device="cuda")
    dy = torch.randn(B, K, P, Q, device="cuda")
    s = torch.rand(B, C, device="cuda") + 0.5; sinv = torch.rand(B, K, device="cuda") + 0.5
    res = {}
    for mode in (0, 1):
        cu.lib.ganb200_conv_scatter_mode(mode)
        t_d = timed(lambda: cu.conv_dgrad(dy, w, (HW, HW), stride, pad, 1, 0.1, engine=2))
        t_m = None
        if stride == 1:
            t_m = timed(lambda: cu.modconv_bwd(x, w, s, sinv, dy, pad, 0.1, True, True, False, engine=2))
        res[mode] = (t_d, t_m)
    cu.lib.ganb200_conv_scatter_mode(1)
    m = f"modconv_bwd(gx,gs) {res[0][1]:.3f} / {res[1][1]:.3f}" if res[0][1] is not None else ""
    print(f"C{C:4d} {HW:2d}x{HW:<2d} K{K:4d} R{R} s{stride}: dgrad {res[0][0]:.3f} / {res[1][0]:.3f}   {m}", flush=True)
    del x, w, dy
