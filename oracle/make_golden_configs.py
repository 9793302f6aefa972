"""TEST INFRASTRUCTURE — golden vectors of the other BASELINE configs and of the pair the
reference's main.py builds as committed ("main"), produced by running
the UNMODIFIED reference (/root/reference through oracle/reference.py) on CPU fp32 with fixed seeds.

    python -m oracle.make_golden_configs [cfg1 cfg4 cfg5 main]     (build container only)

writes tests/golden/<cfg>.pt.  Recipe per config (the same the tests replay with the drop-in modules):
  seed 0 -> construct G then D (weights = pure function of the seed; not stored);
  x, z from torch.Generator(123);  D(x) and G(z) in train mode, recorded in full;
  fresh models (seed 0 again), Train(...) of the config's loop, seed 1, one discriminator_trainstep + one
  generator_trainstep: every returned loss, the norm of every parameter's UPDATE (after - before) and of every buffer.
"""
import os
import random
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

# name -> (generator module, ctor, discriminator module, ctor, train module, batch, image shape)
CONFIGS = {
    "cfg1": dict(gen="generators.generator_1", g_args=lambda: (256, (3, 64, 64)), g_kw={},
                 dis="discriminators.discriminator_1", d_args=lambda: ((3, 64, 64),), d_kw={},
                 train="train.gan", B=8, names=("G1", "D1")),
    "cfg4": dict(gen="generators.generator_3_progan", g_args=lambda: (1, 256, 256, 3), g_kw={},
                 dis="discriminators.discriminator_3_wgangp_progan", d_args=lambda: (1, 64, 3), d_kw={},
                 train="train.wgangp", B=4, names=("G3_PROGAN", "D3_WGAN_PROGAN")),
    # what the reference's main.py builds as committed (main.py:44,60,77): generator_1 + the DCGAN-shaped critic
    # behind get_discriminator.get_3_wgan(ndf=128) + train/wgangpR2.py
    "main": dict(gen="generators.generator_1", g_args=lambda: (256, (3, 64, 64)), g_kw={},
                 dis="discriminators.discriminator_3_wgangp", d_args=lambda: (1, 128, 3), d_kw={},
                 train="train.wgangpR2", B=4, names=("G13", "D3_WGAN")),
    "cfg5": dict(gen="generators.generator_14", g_args=lambda: (256,), g_kw={},
                 dis="discriminators.discriminator_7_wgangpsn", d_args=lambda: (), d_kw={},
                 train="train.wgangpsnR2", B=8, names=("G14", "D7_WGANSN")),
    # SURVEY §8(f) rank 3 — pairs outside the BASELINE configs whose layers are all in the kernel set: the DCGAN pair of
    # get_2 (with the factories' weights_init), the spectrally normalised DCGAN critic, the plain ShuffleNetV2 critic and
    # the three ResNet-in-ResNet critics, each under a loop the reference pairs that kind of critic with.
    "dcgan": dict(gen="generators.generator_2", g_args=lambda: (1, 256, 64, 3), g_kw={},
                  dis="discriminators.discriminator_2", d_args=lambda: (1, 64, 3), d_kw={},
                  train="train.gan", B=8, names=("G2", "D2"), weights_init=True),
    "sn3": dict(gen="generators.generator_2", g_args=lambda: (1, 256, 64, 3), g_kw={},
                dis="discriminators.discriminator_3_wgangpsn", d_args=lambda: (1, 64, 3), d_kw={},
                train="train.wgansnR2", B=4, names=("G2", "D3_WGANSN")),
    "shuffle": dict(gen="generators.generator_2", g_args=lambda: (1, 256, 64, 3), g_kw={},
                    dis="discriminators.discriminator_7_wgangp", d_args=lambda: (), d_kw={},
                    train="train.wgangpR2", B=4, names=("G2", "D7_WGAN")),
    "rir_gp": dict(gen="generators.generator_2", g_args=lambda: (1, 256, 64, 3), g_kw={},
                   dis="discriminators.discriminator_5_wgangp", d_args=lambda: (), d_kw={},
                   train="train.wgangp", B=4, names=("G2", "D5_WGAN")),
    "rir_sn": dict(gen="generators.generator_2", g_args=lambda: (1, 256, 64, 3), g_kw={},
                   dis="discriminators.discriminator_5_wgangpsn", d_args=lambda: (), d_kw={},
                   train="train.wgansn", B=4, names=("G2", "D5_WGANSN")),
    "rir_bce": dict(gen="generators.generator_2", g_args=lambda: (1, 256, 64, 3), g_kw={},
                    dis="discriminators.discriminator_5", d_args=lambda: (), d_kw={},
                    train="train.gan", B=4, names=("G2", "D5")),
    # the dense ShuffleNet-style discriminator of get_4 under the BCE loop with R1 + R2
    "dense4": dict(gen="generators.generator_2", g_args=lambda: (1, 256, 64, 3), g_kw={},
                   dis="discriminators.discriminator_4", d_args=lambda: (), d_kw={},
                   train="train.ganR2", B=4, names=("G2", "D4")),
    # the StyleGAN2 critic of get_8 (log_resolution 6) under R1 + R2 + GP
    "style_d": dict(gen="generators.generator_2", g_args=lambda: (1, 256, 64, 3), g_kw={},
                    dis="discriminators.discriminator_8", d_args=lambda: (6,), d_kw={},
                    train="train.wgangpR2", B=4, names=("G2", "D8")),
    # the remaining loops of SURVEY §8(a) a1 / a7: the lazy-penalty variants (idx 0: penalties on and scaled by 5, idx 1: off)
    # and weight clipping (train/wgan.py:47-48; the StyleGAN2 critic's N(0,1) weights make the clamp bite)
    "lazy": dict(gen="generators.generator_2", g_args=lambda: (1, 256, 64, 3), g_kw={},
                 dis="discriminators.discriminator_3_wgangp", d_args=lambda: (1, 64, 3), d_kw={},
                 train="train.wganlazygp", B=4, names=("G2", "D3_WGAN"), d_idx=(0, 1)),
    "lazy_r2": dict(gen="generators.generator_2", g_args=lambda: (1, 256, 64, 3), g_kw={},
                    dis="discriminators.discriminator_8", d_args=lambda: (6,), d_kw={},
                    train="train.wganlazygpR2", B=4, names=("G2", "D8"), d_idx=(0, 1)),
    "clip": dict(gen="generators.generator_2", g_args=lambda: (1, 256, 64, 3), g_kw={},
                 dis="discriminators.discriminator_8", d_args=lambda: (6,), d_kw={},
                 train="train.wgan", B=4, names=("G2", "D8")),
    # the self-attention critic (SAGAN attention after the stem at 64x64 and at 16x16) under R1 + R2 + GP: torch.bmm and
    # nn.Softmax(dim=1) differentiated twice
    "attn9_3": dict(gen="generators.generator_2", g_args=lambda: (1, 256, 64, 3), g_kw={},
                    dis="discriminators.discriminator_9_3", d_args=lambda: (), d_kw={},
                    train="train.wgangpR2", B=4, names=("G2", "D9_3")),
    # the selective-kernel critic (skip and conv paths merged by SK attention) under R1 + R2 + GP: pooled branch sum,
    # attention trunk, softmax across the two heads and the merge, all differentiated twice
    "sk9_small": dict(gen="generators.generator_2", g_args=lambda: (1, 256, 64, 3), g_kw={},
                      dis="discriminators.discriminator_9_small", d_args=lambda: (), d_kw={},
                      train="train.wgangpR2", B=4, names=("G2", "D9_SMALL")),
    # the deep-layer-aggregation critic of get_6_wgan (depthwise 3x3 / 5x5 / 7x7 bottlenecks with reflect padding, windowed
    # SE squeeze, 5x5 stride-2 reflect convs) under R1 + R2 + GP: 1 276 convolutions differentiated twice
    "tree6": dict(gen="generators.generator_2", g_args=lambda: (1, 256, 64, 3), g_kw={},
                  dis="discriminators.discriminator_6_wgangp", d_args=lambda: (), d_kw={},
                  train="train.wgangpR2", B=4, names=("G2", "D6_WGAN")),
    # the tree generators of get_7 / get_8 (the same aggregation tree with BatchNorm + scalar-PReLU bottlenecks, 25 M and
    # 64 M parameters) against DCGAN-shaped critics: WGAN-GP for one, BCE for the other
    "tree7": dict(gen="generators.generator_7", g_args=lambda: (256,), g_kw={},
                  dis="discriminators.discriminator_3_wgangp", d_args=lambda: (1, 64, 3), d_kw={},
                  train="train.wgangp", B=4, names=("G7", "D3_WGAN")),
    "tree8": dict(gen="generators.generator_8", g_args=lambda: (256,), g_kw={},
                  dis="discriminators.discriminator_2", d_args=lambda: (1, 64, 3), d_kw={},
                  train="train.gan", B=4, names=("G8", "D2")),
    # the InstanceNorm tree generators of get_4 / get_5 (ShuffleNet-style split blocks / densely connected blocks, BatchNorm
    # stems at 2x2 .. 8x8) under the spectral-norm loop with R1 + R2 and the BCE loop with R1 + R2
    "tree4": dict(gen="generators.generator_4", g_args=lambda: (256,), g_kw={},
                  dis="discriminators.discriminator_3_wgangpsn", d_args=lambda: (1, 64, 3), d_kw={},
                  train="train.wgansnR2", B=4, names=("G4", "D3_WGANSN")),
    "tree5": dict(gen="generators.generator_5", g_args=lambda: (256,), g_kw={},
                  dis="discriminators.discriminator_2", d_args=lambda: (1, 64, 3), d_kw={},
                  train="train.ganR2", B=4, names=("G5", "D2")),
    # the self-attention tree generators of get_9 / get_10 / get_10_2 (SAGAN attention up to 64x64: 4 096 x 4 096 maps)
    "tree9": dict(gen="generators.generator_9", g_args=lambda: (256,), g_kw={},
                  dis="discriminators.discriminator_3_wgangp", d_args=lambda: (1, 64, 3), d_kw={},
                  train="train.wgangpR2", B=4, names=("G9", "D3_WGAN")),
    "tree10": dict(gen="generators.generator_10", g_args=lambda: (256,), g_kw={},
                   dis="discriminators.discriminator_3_wgangpsn", d_args=lambda: (1, 64, 3), d_kw={},
                   train="train.wgansn", B=4, names=("G10", "D3_WGANSN")),
    "tree10_2": dict(gen="generators.generator_10_2", g_args=lambda: (256,), g_kw={},
                     dis="discriminators.discriminator_2", d_args=lambda: (1, 64, 3), d_kw={},
                     train="train.gan", B=4, names=("G10_2", "D2")),
    # the GELU tree generator of get_6 and the Pixnorm / equalised self-attention generator of get_11
    "tree6g": dict(gen="generators.generator_6", g_args=lambda: (256,), g_kw={},
                   dis="discriminators.discriminator_3_wgangp", d_args=lambda: (1, 64, 3), d_kw={},
                   train="train.wgangp", B=4, names=("G6", "D3_WGAN")),
    "tree11": dict(gen="generators.generator_11", g_args=lambda: (256,), g_kw={},
                   dis="discriminators.discriminator_3_wgangp", d_args=lambda: (1, 64, 3), d_kw={},
                   train="train.wganlazygp", B=4, names=("G11", "D3_WGAN"), d_idx=(0, 1)),
    # the two predecessors of the final generator (no style path; SK merges everywhere, SAGAN attention in every block), at a
    # quarter / half of their default width and a batch of 8
    "tree13_3": dict(gen="generators.generator_13_3", g_args=lambda: (256,), g_kw={"planes": 16},
                     dis="discriminators.discriminator_3_wgangp", d_args=lambda: (1, 64, 3), d_kw={},
                     train="train.wgangp", B=8, names=("G13_3", "D3_WGAN"), slim=True),
    "tree13_4": dict(gen="generators.generator_13_4", g_args=lambda: (256,), g_kw={"planes": 16},
                     dis="discriminators.discriminator_3_wgangp", d_args=lambda: (1, 64, 3), d_kw={},
                     train="train.wgangp", B=8, names=("G13_4", "D3_WGAN"), slim=True),
    # the first style-based ancestor (a mapping network per modulated convolution, no BatchNorm), at planes 16
    "tree13": dict(gen="generators.generator_13", g_args=lambda: (256,), g_kw={"planes": 16},
                   dis="discriminators.discriminator_3_wgangpsn", d_args=lambda: (1, 64, 3), d_kw={},
                   train="train.wgansnR2", B=4, names=("G13", "D3_WGANSN")),
    # the StyleGAN2-shaped generator of get_12 (noise drawn inside every StyleConv: the replay draws the same stream), planes 16
    "tree12": dict(gen="generators.generator_12", g_args=lambda: (256,), g_kw={"planes": 16},
                   dis="discriminators.discriminator_2", d_args=lambda: (1, 64, 3), d_kw={},
                   train="train.ganR2", B=4, names=("G12", "D2")),
    # generator_13_1 (no factory in the reference): attention inside every style block, ToRGB(6) merges of the RGB skips
    "tree13_1": dict(gen="generators.generator_13_1", g_args=lambda: (256,), g_kw={"planes": 8},
                     dis="discriminators.discriminator_3_wgangp", d_args=lambda: (1, 64, 3), d_kw={},
                     train="train.wgangp", B=2, names=("G13_1", "D3_WGAN")),
    # the same BatchNorm critic under R1 + R2: the penalties differentiate 27 BatchNorms twice
    "rir_r1": dict(gen="generators.generator_2", g_args=lambda: (1, 256, 64, 3), g_kw={},
                   dis="discriminators.discriminator_5", d_args=lambda: (), d_kw={},
                   train="train.ganR2", B=4, names=("G2", "D5")),
}


def seed_all(s):
    torch.manual_seed(s)
    np.random.seed(s)
    random.seed(s)


def build(cfg, genmod, dismod, init_fn=None):
    seed_all(0)
    G = genmod.Generator(*cfg["g_args"](), **cfg["g_kw"])
    if cfg.get("weights_init"):  # units/get_generators.py:18-23, get_discriminator.py:20-25 (get_2 only)
        G.apply(init_fn)
    D = dismod.Discriminator(*cfg["d_args"](), **cfg["d_kw"])
    if cfg.get("weights_init"):
        D.apply(init_fn)
    return G, D


def snapshot(model):
    return {n: p.detach().clone() for n, p in model.named_parameters()}


def update_norms(model, before):
    return {n: float((p.detach().double() - before[n].double()).norm()) for n, p in model.named_parameters()}


def buffer_norms(model):
    return {n: float(b.detach().double().norm()) for n, b in model.named_buffers()}


def replay(cfg, genmod, dismod, trainmod, device=torch.device("cpu"), x=None, z=None, init_fn=None):
    """Runs the recipe with whatever modules it is given (the reference here, the drop-in in the tests)."""
    B = cfg["B"]
    out = {"B": B}
    if x is None:
        gx = torch.Generator().manual_seed(123)
        x = torch.randn(B, 3, 64, 64, generator=gx).clamp_(-1, 1)
        z = torch.randn(B, 256, 1, 1, generator=gx)
    out["x"], out["z"] = x, z
    xd, zd = x.to(device), z.to(device)
    G, D = build(cfg, genmod, dismod, init_fn)
    G, D = G.to(device), D.to(device)
    out["param_names_g"] = [n for n, _ in G.named_parameters()]
    out["param_names_d"] = [n for n, _ in D.named_parameters()]
    out["state_keys_g"] = list(G.state_dict().keys())
    out["state_keys_d"] = list(D.state_dict().keys())
    out["init_norms_g"] = {n: float(p.detach().double().norm()) for n, p in G.named_parameters()}
    out["init_norms_d"] = {n: float(p.detach().double().norm()) for n, p in D.named_parameters()}
    with torch.no_grad():
        out["d_out"] = D(xd).detach().cpu().clone()
        out["g_out"] = G(zd).detach().cpu().clone()
    # one full iteration on fresh models
    G, D = build(cfg, genmod, dismod, init_fn)
    G, D = G.to(device), D.to(device)
    T = trainmod.Train([(xd, torch.zeros(B))], device, 1, 256, G, cfg["names"][0], D, cfg["names"][1])
    g0, d0 = snapshot(G), snapshot(D)
    seed_all(1)
    xi = xd.clone().requires_grad_()
    if cfg.get("d_idx"):  # lazy loops: one critic step per idx value, losses concatenated
        d_ret = []
        for idx in cfg["d_idx"]:
            xi.grad = None
            d_ret += list(T.discriminator_trainstep(xi, B, idx))
    else:
        d_ret = T.discriminator_trainstep(xi, B)
    out["step_d_losses"] = [float(v.detach()) for v in d_ret]
    gen_imgs, g_loss = T.generator_trainstep(B)
    out["step_g_loss"] = float(g_loss.detach())
    out["step_gen_imgs_norm"] = float(gen_imgs.detach().double().norm())
    out["step_d_updates"] = update_norms(D, d0)
    out["step_g_updates"] = update_norms(G, g0)
    out["step_d_buffers"] = buffer_norms(D)
    out["step_g_buffers"] = buffer_norms(G)
    return out


def main(argv):
    from oracle import reference as R
    for name in (argv or ["cfg1", "cfg4"]):
        cfg = CONFIGS[name]
        init_fn = R.load("units.Weight_Initialization").weights_init if cfg.get("weights_init") else None
        out = replay(cfg, R.load(cfg["gen"]), R.load(cfg["dis"]), R.load(cfg["train"]), init_fn=init_fn)
        out["torch"] = torch.__version__
        if cfg.get("slim"):   # chaotic in float32 (tests compare structure, initialisation and D(x) only): no update tables
            for k in [k for k in out if k.startswith("step_") and isinstance(out[k], dict)]:
                out[k] = {}
        path = os.path.join(ROOT, "tests", "golden", name + ".pt")
        torch.save(out, path)
        print("wrote", path, os.path.getsize(path), "bytes;", "d losses", out["step_d_losses"], "g loss",
              out["step_g_loss"])


if __name__ == "__main__":
    main(sys.argv[1:])
