import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from gan_b200 import kernels as KN
cu = KN.CudaBackend()
def t(f):
    f(); torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(5): f()
    g.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / 5
for N, C, H, W, pad in [(256, 96, 64, 64, 2), (256, 96, 66, 64, 2), (256, 96, 62, 64, 2), (256, 97, 64, 64, 2), (250, 96, 64, 64, 2)]:
    x = torch.randn(N, C, H, W, device="cuda"); s = torch.rand(N, C, device="cuda")
    ms = t(lambda: cu.packed_act(x, s, pad, False, transient=True))
    cp = (C + 7) // 8 * 8
    byts = N * C * H * W * 4 + N * (H + 2 * pad) * (W + 2 * pad) * cp * 2
    print(f"pack N{N} C{C} {H}x{W}: {ms*1e3:7.1f} us {byts/ms/1e9:5.2f} TB/s")
    y = torch.randn(N, C, H, W, device="cuda"); a = torch.rand(C, device="cuda")
    ms = t(lambda: cu.prelu_fwd(x, a)); print(f"   prelu_fwd: {ms*1e3:7.1f} us {2*x.numel()*4/ms/1e9:5.2f} TB/s")
    ms = t(lambda: cu.ew_binary(KN.B_ADD, x, y)); print(f"   add     : {ms*1e3:7.1f} us {3*x.numel()*4/ms/1e9:5.2f} TB/s")
    ms = t(lambda: cu.ew_bcast(KN.B_MUL, x, s, KN.V_NC)); print(f"   mul_bc  : {ms*1e3:7.1f} us {2*x.numel()*4/ms/1e9:5.2f} TB/s")
    ms = t(lambda: x.clone()); print(f"   torch clone: {ms*1e3:7.1f} us {2*x.numel()*4/ms/1e9:5.2f} TB/s")
