"""Times the separable resampling operators at the final pair's sizes (B = 256): us per call and GB/s of (in + out)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gan_b200 import kernels as KN
from gan_b200 import resample as RS

cu = KN.CudaBackend()
B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
CASES = [("adaptive_avg 64->5", RS.adaptive_avg(64, 5), 96), ("adaptive_avg^T 5->64", RS.adaptive_avg(64, 5).T, 96),
         ("adaptive_avg 32->5", RS.adaptive_avg(32, 5), 192), ("adaptive_avg^T 5->32", RS.adaptive_avg(32, 5).T, 192),
         ("smooth3 64", RS.smooth3(64), 64), ("smooth3 32", RS.smooth3(32), 128),
         ("smooth+bicubic down 64->32", RS.smooth_then_bicubic_down2(64), 64),
         ("(smooth+bicubic down)^T 32->64", RS.smooth_then_bicubic_down2(64).T, 64),
         ("bicubic up+smooth 32->64", RS.bicubic_up2_then_smooth(32), 108),
         ("(bicubic up+smooth)^T 64->32", RS.bicubic_up2_then_smooth(32).T, 108),
         ("replicate_pad 64+2", RS.replicate_pad(64, 2), 96)]
for name, op, C in CASES:
    x = torch.randn(B, C, op.n_in, op.n_in, device="cuda")
    fn = lambda: cu.resample2d(x, op, op)
    y = fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(5):
            fn()
    g.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / 5 * 1e3
    gb = (x.numel() + y.numel()) * 4 / 1e9
    print(f"{name:34s} C={C:4d} {us:8.1f} us  {gb / us * 1e6:7.0f} GB/s", flush=True)
