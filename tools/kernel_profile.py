"""Per-kernel time of ONE eager training iteration (D step + G step) of the final pair, from CUPTI activity records
(torch.profiler): warm caches, kernels not serialised by a profiler replay.  Complements the ncu launch list.
python tools/kernel_profile.py [batch] [loop] > profiles/...txt"""
import collections
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
loop = sys.argv[2] if len(sys.argv) > 2 else "ganR2"
dev = torch.device("cuda:0")
T = bench.build_trainer(dev, loop, B)
images = torch.randn(B, 3, 64, 64, device=dev).requires_grad_()
for _ in range(2):
    bench.one_iteration(T, images, B)
torch.cuda.synchronize()
from torch.profiler import ProfilerActivity, profile  # noqa: E402

from gan_b200 import kernels as KN  # noqa: E402
KN.OP_CENSUS = {}
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    bench.one_iteration(T, images, B)
    torch.cuda.synchronize()
agg = collections.defaultdict(lambda: [0.0, 0])
for ev in prof.events():
    if ev.device_type == torch.autograd.DeviceType.CUDA:
        a = agg[ev.name]
        a[0] += ev.device_time if hasattr(ev, "device_time") else ev.cuda_time
        a[1] += 1
tot = sum(v[0] for v in agg.values())
print(f"# one eager iteration, batch {B}, loop {loop}: {sum(v[1] for v in agg.values())} device activities, "
      f"{tot / 1e3:.1f} ms of device time (CUPTI)")
print(f"{'share':>7s} {'ms':>9s} {'count':>7s} {'avg us':>8s}  name")
for name, (us, n) in sorted(agg.items(), key=lambda kv: -kv[1][0]):
    if us / tot < 0.0005:
        continue
    print(f"{100 * us / tot:6.2f}% {us / 1e3:9.2f} {n:7d} {us / n:8.1f}  {name[:110]}")
print("\n# resampling calls of the iteration (operator, planes, in, out, taps): calls")
for k, v in sorted(KN.OP_CENSUS.items(), key=lambda kv: -kv[1] * kv[0][2] * kv[0][3] * kv[0][4]):
    print(f"{v:5d}  {k[1]:40s} planes {k[2]:7d}  {k[3]}x{k[4]} -> {k[5]}x{k[6]}  T={k[7]}")
