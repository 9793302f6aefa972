"""GPU: tcgen05 conv engine vs the CPU oracle on a list of shapes; prints rel-L2 per (shape, op) and timings."""
import ctypes, sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gan_b200 import kernels as KN
from oracle.kernels_ref import RefBackend

cu, ref = KN.CudaBackend(), RefBackend()
CASES = [(4, 128, 18, 18, 128, 3, 1, 1, 2), (4, 24, 16, 16, 32, 4, 1, 0, 2), (4, 108, 16, 16, 3, 5, 2, 1), (4, 3, 16, 16, 64, 3, 1, 1),
         (4, 64, 10, 10, 64, 3, 1, 1, 2), (4, 40, 7, 9, 24, 3, 1, 0, 2), (4, 64, 16, 16, 64, 3, 1, 1), (2, 64, 64, 64, 64, 3, 1, 1), (8, 128, 8, 8, 256, 3, 1, 1), (8, 1025, 4, 4, 1025, 3, 1, 1),
         (4, 48, 16, 16, 54, 5, 2, 1), (4, 390, 8, 8, 96, 1, 0, 1), (4, 64, 5, 5, 64, 3, 0, 0), (3, 24, 9, 11, 40, 3, 1, 0),
         (256, 256, 1, 1, 512, 1, 0, 0), (4, 96, 32, 32, 96, 5, 2, 1)]
which = sys.argv[1:] or ["fprop", "dgrad", "wgrad"]
rel = lambda a, b: ((a.cpu().double() - b.double()).norm() / b.double().norm()).item()
for case in CASES:
    N, C, H, W, K, R, pad, pm = case[:8]
    ST = case[8] if len(case) > 8 else 1
    torch.manual_seed(0)
    x, w, b = torch.randn(N, C, H, W), torch.randn(K, C, R, R), torch.randn(K)
    isc, osc = torch.rand(N, C) + 0.5, torch.rand(N, K) + 0.5
    P, Q = KN.out_hw(H, W, R, R, ST, pad)
    dy = torch.randn(N, K, P, Q)
    d = KN.ConvDesc(N, C, H, W, K, R, R, ST, pad, pm, P, Q, 0.1, 0)
    out = []
    g = lambda t: t.cuda()
    for op, name in ((KN.FPROP, "fprop"), (KN.DGRAD, "dgrad"), (KN.WGRAD, "wgrad")):
        if name not in which or not cu.lib.ganb200_conv2d_tc_eligible(ctypes.byref(d), op):
            out.append(f"{name}: -")
            continue
        try:
            if op == KN.FPROP:
                got = cu.conv_fprop(g(x), g(w), g(b), ST, pad, pm, 0.1, g(isc), g(osc), engine=2)
                want = ref.conv_fprop(x, w, b, ST, pad, pm, 0.1, isc, osc)
            elif op == KN.DGRAD:
                got = cu.conv_dgrad(g(dy), g(w), (H, W), ST, pad, pm, 0.1, g(isc), g(osc), engine=2)
                want = ref.conv_dgrad(dy, w, (H, W), ST, pad, pm, 0.1, isc, osc)
            else:
                got = cu.conv_wgrad(g(x), g(dy), (R, R), ST, pad, pm, 0.1, g(isc), g(osc), engine=2)
                want = ref.conv_wgrad(x, dy, (R, R), ST, pad, pm, 0.1, isc, osc)
            torch.cuda.synchronize()
            out.append(f"{name}: {rel(got, want):.2e}")
        except Exception as e:
            out.append(f"{name}: ERR {str(e)[:80]}")
    print(case, " | ".join(out), flush=True)

# timing at bench-like sizes: engine 1 vs 2
TIMING = [(256, 128, 64, 128, 3, 2), (256, 1024, 8, 1024, 3, 2), (256, 108, 64, 3, 5, 1), (256, 3, 64, 64, 3, 1)]
for (N, C, HW, K, R, ST) in TIMING:
    pad = R // 2
    P, Q = KN.out_hw(HW, HW, R, R, ST, pad)
    x, w = torch.randn(N, C, HW, HW, device="cuda"), torch.randn(K, C, R, R, device="cuda")
    dy = torch.randn(N, K, P, Q, device="cuda")
    fl = 2.0 * N * P * Q * C * K * R * R
    res = []
    for name in which:
        f = {"fprop": lambda: cu.conv_fprop(x, w, None, ST, pad, 1, 1.0, engine=2),
             "dgrad": lambda: cu.conv_dgrad(dy, w, (HW, HW), ST, pad, 1, 1.0, engine=2),
             "wgrad": lambda: cu.conv_wgrad(x, dy, (R, R), ST, pad, 1, 1.0, engine=2)}[name]
        for _ in range(2): f()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5): f()
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 5
        res.append(f"{name}/tc: {ms:.3f} ms {fl / ms / 1e9:.1f} TF/s")
    print((N, C, HW, K, R, ST), " | ".join(res), flush=True)
sys.exit(0)
for (N, C, HW, K, R) in [(64, 64, 64, 64, 3), (64, 128, 32, 128, 3), (64, 256, 16, 256, 3), (64, 512, 8, 512, 3), (64, 1025, 4, 1025, 3), (64, 96, 64, 96, 5), (64, 48, 64, 48, 5)]:
    x, w = torch.randn(N, C, HW, HW, device="cuda"), torch.randn(K, C, R, R, device="cuda")
    dy = torch.randn(N, K, HW, HW, device="cuda")
    fl = 2.0 * N * HW * HW * C * K * R * R
    res = []
    for name in which:
        for eng in (1, 2):
            d = KN.ConvDesc(N, C, HW, HW, K, R, R, 1, R // 2, 1, HW, HW, 1.0, 0)
            opid = {"fprop": KN.FPROP, "dgrad": KN.DGRAD, "wgrad": KN.WGRAD}[name]
            if eng == 2 and not cu.lib.ganb200_conv2d_tc_eligible(ctypes.byref(d), opid):
                continue
            f = {"fprop": lambda: cu.conv_fprop(x, w, None, 1, R // 2, 1, 1.0, engine=eng),
                 "dgrad": lambda: cu.conv_dgrad(dy, w, (HW, HW), 1, R // 2, 1, 1.0, engine=eng),
                 "wgrad": lambda: cu.conv_wgrad(x, dy, (R, R), 1, R // 2, 1, 1.0, engine=eng)}[name]
            for _ in range(2): f()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(5): f()
            e1.record(); torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / 5
            res.append(f"{name}/e{eng}: {ms:.3f} ms {fl / ms / 1e9:.1f} TF/s")
    print((N, C, HW, K, R), " | ".join(res), flush=True)
