"""TEST INFRASTRUCTURE — generates tests/golden/final_pair_b4.pt by running the UNMODIFIED reference
(/root/reference, imported through oracle/reference.py) on CPU fp32 with fixed seeds.

    python -m oracle.make_golden            (only in the build container; the GPU box has no /root/reference)

Recorded (SURVEY §8(c)/(d) recipe: seeds 0 before construction, inputs drawn from seeded generators):
  * discriminator_9_4: logits of a seeded (4,3,64,64) batch; R1 penalty value and gradient norms of selected
    parameters after `real_loss.backward(); r1.backward()` (the create_graph double backward);
  * generator_13_5: an 8x8-strided subsample and the norm of G(z) for a seeded z with torch.manual_seed(1) set right
    before the forward (fixes the 253 in-forward noise draws);
  * one full `train/wgangpR2.py` iteration (discriminator_trainstep + generator_trainstep) from seed 1: the five D
    losses, the G loss, gradient norms of selected parameters after each step and the norm of every tracked
    parameter after the two Adam updates.
The weights themselves are not stored (515 M parameters): they are a pure function of torch.manual_seed(0) on the
CPU generator, which both the reference and the drop-in consume in the same order (parameter registration order).
"""
import os
import random
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

D_TRACK = ["conv.0.weight.weight", "conv.0.bias", "conv.1.weight", "conv.3.block.0.weight.weight",
           "conv.8.se.conv.0.weight.weight", "conv.16.block.1.weight", "fc.0.weight.weight", "fc.2.bias"]
G_TRACK = ["block0.mapping.net.0.weight.weights", "block0.convT.weight", "block0.bn.weight",
           "block4.tree.root.init_block.residual_across.conv1.conv.weight.weights",
           "block4.tree.to_rgb.to_rgb.weight.weights", "block4.upsample_x.convT.bias"]


def seed_all(s):
    torch.manual_seed(s)
    np.random.seed(s)
    random.seed(s)


def pick(names, wanted):
    """resolve the wanted names against what the model really has (keeps the script robust to naming)."""
    names = list(names)
    out = [n for n in wanted if n in names]
    if len(out) < 4:  # fall back to a deterministic spread over all parameters
        step = max(1, len(names) // 12)
        out = names[::step][:12]
    return out


def gnorms(model, names):
    d = dict(model.named_parameters())
    return {n: (float(d[n].grad.double().norm()) if d[n].grad is not None else None) for n in names}


def pnorms(model, names):
    d = dict(model.named_parameters())
    return {n: float(d[n].detach().double().norm()) for n in names}


def build(genmod, dismod):
    seed_all(0)
    G = genmod.Generator(256)
    D = dismod.Discriminator()
    return G, D


def record(genmod, dismod, trainmod, B=4):
    out = {"B": B}
    G, D = build(genmod, dismod)
    d_names = pick([n for n, _ in D.named_parameters()], D_TRACK)
    g_names = pick([n for n, p in G.named_parameters() if p.requires_grad], G_TRACK)
    out["d_names"], out["g_names"] = d_names, g_names
    gx = torch.Generator().manual_seed(123)
    x = torch.randn(B, 3, 64, 64, generator=gx)
    z = torch.randn(B, 256, 1, 1, generator=gx)
    out["x"], out["z"] = x, z
    # --- D forward, R1 double backward
    xr = x.clone().requires_grad_()
    pred = D(xr)
    out["d_logits"] = pred.detach().clone()
    (-pred.mean()).backward(retain_graph=True)
    g = torch.autograd.grad(pred.sum(), xr, create_graph=True, retain_graph=True)[0]
    r1 = g.pow(2).view(B, -1).sum(1).mean()
    r1.backward()
    out["d_r1"] = float(r1)
    out["d_r1_gradnorms"] = gnorms(D, d_names)
    D.zero_grad()
    # --- G forward
    seed_all(1)
    with torch.no_grad():
        img = G(z)
    out["g_out_sub"] = img[:, :, ::8, ::8].clone()
    out["g_out_norm"] = float(img.double().norm())
    # --- one full wgangpR2 iteration on fresh models
    G, D = build(genmod, dismod)
    T = trainmod.Train([(x, torch.zeros(B))], torch.device("cpu"), 1, 256, G, "G13_5", D, "D9_4")
    seed_all(1)
    xi = x.clone().requires_grad_()
    d_out = T.discriminator_trainstep(xi, B)
    out["step_d_losses"] = [float(v) for v in d_out]
    out["step_d_gradnorms"] = gnorms(D, d_names)
    gen_imgs, g_loss = T.generator_trainstep(B)
    out["step_g_loss"] = float(g_loss)
    out["step_g_gradnorms"] = gnorms(G, g_names)
    out["step_gen_imgs_norm"] = float(gen_imgs.detach().double().norm())
    out["step_d_pnorms"] = pnorms(D, d_names)
    out["step_g_pnorms"] = pnorms(G, g_names)
    out["g_dead_grads"] = sum(1 for p in G.parameters() if p.requires_grad and p.grad is None)
    return out


def main():
    from oracle import reference as R
    genmod = R.load("generators.generator_13_5")
    dismod = R.load("discriminators.discriminator_9_4")
    trainmod = R.load("train.wgangpR2")
    out = record(genmod, dismod, trainmod)
    out["torch"] = torch.__version__
    path = os.path.join(ROOT, "tests", "golden", "final_pair_b4.pt")
    torch.save(out, path)
    print("wrote", path, os.path.getsize(path), "bytes")
    for k, v in out.items():
        if not torch.is_tensor(v):
            print(k, v)


if __name__ == "__main__":
    main()
