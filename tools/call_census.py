"""Kernel-call census of ONE training iteration of the final pair on the CPU restatement backend: every backend method
call with its tensor shapes, aggregated, scaled to batch B (shapes scale linearly with N).  Planning tool: shows which
operand tensors the convolution engine packs, how many bytes each kernel class moves, launches per class.
python tools/call_census.py [loop] [B_report]"""
import collections
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from gan_b200 import kernels as KN  # noqa: E402
from oracle.kernels_ref import RefBackend  # noqa: E402

loop = sys.argv[1] if len(sys.argv) > 1 else "ganR2"
BR = int(sys.argv[2]) if len(sys.argv) > 2 else 256
B = 4


class Tracer:
    def __init__(self, inner):
        self.inner = inner
        self.calls = collections.Counter()
        self.phase = "?"

    def __getattr__(self, name):
        f = getattr(self.inner, name)
        if not callable(f):
            return f

        def wrap(*a, **k):
            shapes = tuple(tuple(t.shape) if torch.is_tensor(t) else (t if isinstance(t, (int, float, tuple, type(None))) else "?")
                           for t in a)
            self.calls[(self.phase, name, shapes, tuple(sorted((kk, tuple(v.shape) if torch.is_tensor(v) else v) for kk, v in k.items() if kk in ("iscale", "oscale", "engine"))))] += 1
            return f(*a, **k)
        return wrap


tr = Tracer(RefBackend())
KN.set_backend(tr)
T = bench.build_trainer(torch.device("cpu"), loop, B)
images = torch.randn(B, 3, 64, 64).requires_grad_()
tr.phase = "D"
T.discriminator_trainstep(images, B)
tr.phase = "G"
T.generator_trainstep(B)
import pickle
pickle.dump(dict(tr.calls), open("/tmp/call_census.pkl", "wb"))
by = collections.Counter()
for (ph, name, shapes, kw), n in tr.calls.items():
    by[(ph, name)] += n
for k, v in sorted(by.items(), key=lambda kv: -kv[1]):
    print(v, k)
print("total", sum(by.values()))
