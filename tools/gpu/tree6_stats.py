"""tree6 (generator_2 + discriminator_6_wgangp under R1 + R2 + GP) on the GPU against the recorded reference golden:
distribution of the update-norm differences (most of this critic's gradients sit at Adam's eps, tests/test_other_configs.py
TOL_CPU["tree6"]) and the time of one iteration.  Usage: python tools/gpu/tree6_stats.py [fp32|auto]"""
import importlib
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
os.environ["GANB200_PRECISION"] = sys.argv[1] if len(sys.argv) > 1 else "fp32"
from gan_b200 import rng  # noqa: E402
from oracle.make_golden_configs import CONFIGS, replay  # noqa: E402

cfg = CONFIGS["tree6"]
gold = torch.load(os.path.join(ROOT, "tests", "golden", "tree6.pt"), weights_only=False)
mods = [importlib.import_module(cfg[k]) for k in ("gen", "dis", "train")]
prev = rng.host_stream(True)
for it in range(2):
    torch.cuda.synchronize()
    t0 = time.time()
    got = replay(cfg, *mods, device=torch.device("cuda"), x=gold["x"], z=gold["z"])
    torch.cuda.synchronize()
    print("replay %d: %.2f s" % (it, time.time() - t0))
rng.host_stream(prev)
print("precision", os.environ["GANB200_PRECISION"])
print("d losses", got["step_d_losses"], "\n   golden", gold["step_d_losses"])
print("g loss", got["step_g_loss"], "golden", gold["step_g_loss"])
for k in ("d_out", "g_out"):
    a, b = torch.as_tensor(got[k]).double().cpu(), torch.as_tensor(gold[k]).double()
    print(k, "rel", float((a - b).norm() / b.norm()))
for k in ("step_d_updates", "step_g_updates"):
    names = list(gold[k].keys())
    ab = np.array([abs(got[k][n] - gold[k][n]) for n in names])
    re = np.array([abs(got[k][n] - gold[k][n]) / max(abs(gold[k][n]), 1e-6) for n in names])
    print(k, len(names), "rel median %.2e p90 %.2e p99 %.2e max %.2e | n>5e-3: %d | abs max %.2e" %
          (np.median(re), np.percentile(re, 90), np.percentile(re, 99), re.max(), int((re > 5e-3).sum()), ab.max()))
    for i in np.argsort(-ab)[:5]:
        print("   %-64s got %.4e want %.4e" % (names[i], got[k][names[i]], gold[k][names[i]]))
