"""pack_act (fp32 NCHW -> bf16 NHWC padded) bandwidth; GANB200_PACK_V2=0 selects the first-generation kernel."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gan_b200 import kernels as KN
cu = KN.CudaBackend()
for N, C, HW, pad in [(256, 96, 64, 2), (256, 48, 64, 1), (256, 64, 64, 1), (256, 390, 64, 0), (256, 128, 32, 1),
                      (256, 192, 16, 2), (256, 256, 16, 1)]:
    x = torch.randn(N, C, HW, HW, device="cuda"); s = torch.rand(N, C, device="cuda")
    f = lambda: cu.packed_act(x, s, pad, False, transient=True)
    f(); torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(5):
            f()
    g.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    cp = (C + 7) // 8 * 8
    byts = N * C * HW * HW * 4 + N * (HW + 2 * pad) ** 2 * cp * 2
    print(f"N{N} C{C} {HW}x{HW} pad{pad}: {ms*1e3:7.1f} us  {byts/ms/1e9:6.2f} TB/s", flush=True)
