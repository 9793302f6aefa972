"""The other BASELINE configs against golden vectors recorded from the UNMODIFIED
reference (oracle/make_golden_configs.py -> tests/golden/cfg*.pt): identical parameter names and registration order,
identical seeded initial weights, D(x) and G(z), and one full iteration of the config's own loop (train/gan.py,
train/wgangp.py, train/wgangpsnR2.py): every loss, the norm of every parameter's update and of every buffer.

CPU: drop-in modules on the CPU restatement of the kernels.  GPU: the same through libganb200, fp32 engine and the
bf16 tcgen05 engine (tolerances stated per test)."""
import importlib
import os
import re

import pytest
import torch

from oracle.make_golden_configs import CONFIGS, replay

# Parameters whose gradient is analytically zero (a bias added right before a BatchNorm): what reaches Adam is fp32
# rounding noise (~1e-9), which Adam's normalisation turns into an O(lr) update of arbitrary sign whatever the
# implementation (the CPU port differs from the reference by up to 40 % there and by < 2e-3 everywhere else).  They
# are excluded from the update-norm comparison.
NOISE_PARAMS = {"cfg1": None, "cfg4": re.compile(r"^main\.\d\.0\.bias$"), "cfg5": None, "main": None,
                "rir_bce": re.compile(r"(stream_conv|stream_conv_across|short_cut\.0|pre_conv\.0)\.bias$")}
NOISE_PARAMS["rir_r1"] = NOISE_PARAMS["rir_bce"]
# style_d: the last block's two biases feed the same sum right before the head; the real (-1/B) and fake (+1/B) Wasserstein
# terms cancel on them up to the LeakyReLU masks and what is left (~1e-9 per element, gradient penalty included) is at
# Adam's eps: update norms 0.0013 / 0.0024 here / in the reference against 0.009 for a full sign step.
NOISE_PARAMS["style_d"] = re.compile(r"^blocks\.3\.(residual\.1|block\.2)\.bias$")
NOISE_PARAMS["lazy_r2"] = NOISE_PARAMS["clip"] = NOISE_PARAMS["style_d"]  # the same critic
# attn9_3: the query bias of a self-attention layer adds b^T g_j to every s[i, j] — constant along the softmax dimension, so
# its gradient is analytically zero (reference: discriminators/discriminator_9_3.py:74-75)
NOISE_PARAMS["attn9_3"] = re.compile(r"\.query\.bias$")
# tree7 / tree8: every convolution (and transposed convolution) that feeds a BatchNorm — `res.0/3/6`, `unify.0`, `shortcut.0`,
# `Upsample_N[.0].0` — has a bias with an analytically zero gradient (838 / 1 055 of the generators' tensors)
NOISE_PARAMS["tree7"] = NOISE_PARAMS["tree8"] = re.compile(r"(res\.[036]|unify\.0|shortcut\.0|Upsample_\d(\.0)?\.0)\.bias$")
# tree4 / tree5: the same (a conv or transposed conv feeding an InstanceNorm or BatchNorm), and the first conv of every block's
# SE branch: its input is the spatial mean of an InstanceNorm output, i.e. the norm's bias, 0 at init — the weight gradient
# is (delta x rounding noise)
NOISE_PARAMS["tree4"] = NOISE_PARAMS["tree5"] = re.compile(
    r"((model_\d\.[036]|res_\d\.[036]|unify\.0|shortcut\.0|upsample\.0|model\.(0|6|8|10)(\.0)?)\.bias|(block_\d|root)\.se_[357]\.[01]\.weight)$")
# tree9 / tree10: as tree7 / tree8, and the key bias of every self-attention layer: it adds b^T g_j to every s[i, j], constant
# along the softmax dimension (reference: generators/generator_9.py:20-25)
NOISE_PARAMS["tree9"] = NOISE_PARAMS["tree10"] = re.compile(
    r"((res\.[036]|unify\.0|shortcut\.0|Upsample_\d\.0\.0)\.bias|\.key\.bias)$")
# tree6g: the convolutions in front of a BatchNorm of generator_6; tree11 (Pixnorm: nothing is mean-subtracted): key biases only
NOISE_PARAMS["tree6g"] = re.compile(
    r"(res\.[036]|unify\.0|shortcut\.0|Stem_1\.0|upsample\.0|Upsample_\d\.0|(residual|transient)_pre_conv\.0)\.bias$")
NOISE_PARAMS["tree11"] = re.compile(r"\.key\.bias$")
# tree13: here f = query (generators/generator_13.py:331-335), so the query biases are the softmax-invariant ones
NOISE_PARAMS["tree13"] = re.compile(r"\.query\.bias$")
NOISE_PARAMS["tree12"] = re.compile(r"\.key\.bias$")      # f = key here (generators/generator_12.py:139-143)
# tree13_1: its feature maps reach |x| ~ 1e3 at initialisation, so the 4 096-way attention softmaxes are saturated and the
# gradient of everything in front of the logits (the key / query StyleConvs with their mapping networks: 2 840 tensors) is
# a difference of nearly equal numbers — in float64 those gradients agree with the reference's to 2e-9 against 1e-13 for
# the rest (tests/test_d6_tree.py), a 1e7 amplification that leaves nothing in float32, in either implementation
NOISE_PARAMS["tree13_1"] = re.compile(r"\.(key|query)\.")
NOISE_IN_D = ("rir_bce", "rir_r1", "style_d", "lazy_r2", "clip", "attn9_3")
ZOO = ["dcgan", "sn3", "shuffle", "rir_gp", "rir_sn", "rir_bce", "rir_r1", "style_d", "dense4", "lazy", "lazy_r2", "clip", "attn9_3", "sk9_small", "tree4", "tree5", "tree6", "tree7", "tree8", "tree9", "tree10", "tree10_2", "tree6g", "tree11", "tree13_3", "tree13_4", "tree13", "tree12", "tree13_1"]  # SURVEY §8(f) rank 3 pairs (make_golden_configs.py)

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def rel(a, b):
    a, b = torch.as_tensor(a, dtype=torch.float64), torch.as_tensor(b, dtype=torch.float64)
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def compare(name, device, tol_out, tol_loss, tol_upd, floor=1e-6, chain=True, tol_gloss=None):
    from gan_b200 import rng
    cfg = CONFIGS[name]
    gold = torch.load(os.path.join(GOLDEN, name + ".pt"), weights_only=False)
    mods = [importlib.import_module(cfg[k]) for k in ("gen", "dis", "train")]
    init_fn = importlib.import_module("units.Weight_Initialization").weights_init if cfg.get("weights_init") else None
    prev = rng.host_stream(True)
    try:
        got = replay(cfg, *mods, device=device, x=gold["x"], z=gold["z"], init_fn=init_fn)
    finally:
        rng.host_stream(prev)
    bad = []
    for k in ("param_names_g", "param_names_d", "state_keys_g", "state_keys_d"):
        if got[k] != gold[k]:
            bad.append((k, [a for a in gold[k] if a not in got[k]][:5], [a for a in got[k] if a not in gold[k]][:5]))
    for k in ("init_norms_g", "init_norms_d"):  # same seed, same construction order -> same weights
        for n, w in gold[k].items():
            if abs(got[k].get(n, float("nan")) - w) > 1e-5 * max(abs(w), 1e-3):
                bad.append((k, n, got[k].get(n), w))
    for k in ("d_out", "g_out") if chain else ("d_out",):
        e = rel(got[k], gold[k])
        if not e < tol_out:
            bad.append((k, e))
    if not chain:  # ill-conditioned chain (see the caller): structure, init, critic output and a finite step only
        import math
        assert all(math.isfinite(v) for v in got["step_d_losses"] + [got["step_g_loss"], got["step_gen_imgs_norm"]])
        assert not bad, bad[:20]
        return got
    for g, w in zip(got["step_d_losses"], gold["step_d_losses"]):
        if not abs(g - w) <= tol_loss * max(abs(w), 1e-2):
            bad.append(("d loss", g, w))
    g, w = got["step_g_loss"], gold["step_g_loss"]  # evaluated AFTER the critic's Adam step
    if not abs(g - w) <= (tol_loss if tol_gloss is None else tol_gloss) * max(abs(w), 1e-2):
        bad.append(("g loss", g, w))
    e = abs(got["step_gen_imgs_norm"] - gold["step_gen_imgs_norm"]) / gold["step_gen_imgs_norm"]
    if not e < tol_out:
        bad.append(("step_gen_imgs_norm", e))
    for k in ("step_d_updates", "step_g_updates", "step_d_buffers", "step_g_buffers"):
        for n, w in gold[k].items():
            g = got[k][n]
            noise = NOISE_PARAMS.get(name)
            if noise is not None and k == ("step_d_updates" if name in NOISE_IN_D else "step_g_updates") and noise.search(n):
                continue
            if not abs(g - w) <= tol_upd * max(abs(w), floor) + floor:
                bad.append((k, n, g, w))
    assert not bad, bad[:20]
    return got


# The critics of these pairs start nearly constant (default init, |D(x)| ~ 1e-2, R1 ~ 1e-12): many of their gradients
# are ~1e-9, the size of Adam's eps, so the first update of those tensors follows fp32 rounding (update norms differ by up
# to 3.5e-3 from the reference's, all D-step losses — computed before any update — agree to 1e-6) and the generator loss,
# evaluated on the UPDATED critic, inherits it (measured 2.7 % / 6e-4 / 1.4e-4).
# rir_bce (27 BatchNorms over a batch of 4): the reference's own fp32 weight gradients are 2.3e-4 off its fp64 run (this
# port: 5e-5 off the same fp64 run), the real and fake passes then partly cancel (3e-3 on the accumulated gradient), and
# ~4 600 of 26 M critic weights get the opposite Adam sign: update NORMS agree to 1e-4, the generator loss to 2.8e-3.
# Conv biases feeding a BatchNorm have an analytically zero gradient (NOISE_PARAMS).
TOL_CPU = {"shuffle": dict(tol_gloss=5e-2, tol_upd=5e-3), "rir_gp": dict(tol_gloss=2e-3), "rir_sn": dict(tol_gloss=1e-3),
           "rir_bce": dict(tol_gloss=1e-2), "rir_r1": dict(tol_gloss=1e-2, tol_loss=5e-4),  # R2 through 27 twice-differentiated BatchNorms: 1.06e-4
           # dense4: 139 BatchNorms over a batch of 4, differentiated twice: R2 2.9e-4, generator loss 5e-4, 705 of 707 update
           # norms within 6e-4 (a 4-pixel BatchNorm weight 1.7e-2, an SE weight 3.3e-3); outputs, BCE losses and R1 at 1e-6
           "dense4": dict(tol_loss=1e-3, tol_gloss=2e-3, tol_upd=2.5e-2),
           # tree6: 1 276 un-normalised convolutions at PyTorch's default init leave most of the critic's gradients at Adam's
           # eps (only 509 of 2 554 tensors move by > 0.8 lr per element; those agree to 4e-4).  For the others the update norm
           # follows the ABSOLUTE fp32 rounding of a ~1e-9 gradient: the unmodified reference run with 1 thread differs from
           # its own default-thread golden by median 1.7e-6, 223 tensors > 2e-3, max 0.30, largest absolute difference 3.8e-5;
           # the CPU port: median 1.6e-6, 234 > 2e-3, max 0.34, abs 4.1e-5 (fp64: forward / double backward equal to 1e-15,
           # tests/test_d6_tree.py) -> an absolute floor of 6e-5 on the update norms of this config
           "tree6": dict(floor=6e-5, tol_gloss=1e-3),
           # tree7 / tree8 (BatchNorm over a batch of 4 in ~2 500 places): outputs 9e-6, critic updates 1e-5; the generators'
           # update norms, zero-gradient biases excluded: 194 / 209 tensors > 2e-3 with the largest absolute difference
           # 9e-6 / 1.4e-5 — the unmodified reference, 1 thread against its default-thread golden: 192 / 209 > 2e-3, abs
           # 3.0e-5 / 4.8e-5.  tree8's generator loss is read through discriminator_2's four BatchNorms: 1.2e-4
           "tree7": dict(floor=6e-5, tol_gloss=1e-3), "tree8": dict(floor=6e-5, tol_gloss=1e-3),
           # tree4 / tree5 (BatchNorm over 4 samples at 2x2, ~1 300 / 2 400 InstanceNorms): the unmodified reference, 1 thread
           # against its default-thread golden: G(z) 3.2e-5 / 1.4e-4, generator loss 2e-6 / 2.6e-4, R2 term of tree5 3.6e-5,
           # update norms (zero-gradient tensors excluded) 164 / 368 > 2e-3 with abs 4.5e-5 / 1.2e-4.  The CPU port: G(z)
           # 3.6e-5 / 1.6e-4, generator loss 3e-6 / 1.1e-4, R2 term 1.0e-4, 68 / 148 > 2e-3 with abs 2.9e-5 / 7.2e-5
           # tree9 / tree10: CPU port G(z) 7e-6 / 3e-6, update norms 242 / 23 tensors > 2e-3 with abs 9.5e-6 / 1.4e-5
           "tree9": dict(floor=6e-5, tol_gloss=1e-3), "tree10": dict(floor=6e-5, tol_gloss=1e-3),
           # tree6g: G(z) 7e-5 (7 000 tensors of 4 .. 64 channels, BatchNorm at 2x2), 111 update norms > 2e-3 with abs 2.2e-5;
           # tree11 (Pixnorm, well conditioned): G(z) 3e-6, every update norm within 4.2e-4
           "tree6g": dict(floor=6e-5, tol_gloss=1e-3), "tree11": dict(floor=1e-6),
           # tree13 (no BatchNorm: well conditioned): G(z) 9e-7, all 13 394 update norms within 1e-3
           "tree13": dict(floor=1e-6),
           # tree12 (a noise plane drawn inside each of 579 StyleConvs; the replay draws the same stream): G(z) 6e-7, update
           # norms 64 of 4 396 > 2e-3 with abs 5.2e-5
           "tree12": dict(floor=6e-5),
           # tree13_1: G(z) 2e-6; of the 9 502 update norms outside the key / query paths one exceeds 2e-3 (abs 6e-6)
           "tree13_1": dict(floor=1e-5),
           # tree10_2 is ill-conditioned in fp32 by construction (N(0,1) weights in front of 4 096-way softmaxes): the
           # reference's own fp32 G(z) is 2.7e-2 away from its float64 run and 1.5e-2 from itself at another thread count,
           # the port 2.5e-2 from float64 (and 7e-11 from the reference in float64, tests/test_d6_tree.py) -> structure,
           # initialisation, the critic's output and a finite step only
           "tree10_2": dict(chain=False, floor=6e-5),
           # tree13_3 / tree13_4 (planes 16, batch 8): chaotic at initialisation — BatchNorm1d over 8 samples inside hundreds of
           # attention heads; the port's float32 G(z) is 3e-2 / 1.1 from the reference's, whose own float64 run amplifies a
           # 1e-14 difference to 2e-4 over the five levels (tests/test_d6_tree.py compares them level by level in float64)
           "tree13_3": dict(chain=False, floor=6e-5), "tree13_4": dict(chain=False, floor=6e-5),
           "tree4": dict(floor=6e-5, tol_gloss=1e-3), "tree5": dict(floor=1.5e-4, tol_out=3e-4, tol_loss=5e-4, tol_gloss=1e-3)}


# (tree10_2 / tree13_4 — structure-only comparisons of chaotic generators — run on the GPU only; on the CPU the float64 tests
# of tests/test_d6_tree.py say more about them in less time)
@pytest.mark.parametrize("name", ["cfg1", "cfg4", "cfg5", "main"] + [n for n in ZOO if n not in ("tree10_2", "tree13_4")])
def test_config_cpu_port_matches_reference_golden(ref_backend, name):
    compare(name, torch.device("cpu"), **{**dict(tol_out=1e-4, tol_loss=1e-4, tol_upd=2e-3), **TOL_CPU.get(name, {})})


# cfg5 on the GPU.  The critic is spectrally normalised 79 times over: its outputs are ~5e-3 and its gradients ~1e-9, the
# size of Adam's eps, so an update norm follows the ABSOLUTE rounding error of the gradient (measured up to 3 % between
# the fp32 CUDA kernels and the reference's CPU run; < 2e-3 for the CPU port which shares the reference's GEMM order).
# generator_14 over a batch of 8 amplifies a 1e-6 per-layer difference to 4e-5 at the output (40x: BatchNorm1d over 8
# samples in every attention head), and the bf16 engine's 2e-3 per conv to ~0.2: end-to-end output parity is meaningless
# there, so the bf16 engine is checked block by block on identical inputs instead (test_cfg5_blockwise_tcgen05).
TOL_FP32 = {"cfg1": dict(tol_out=2e-4, tol_loss=2e-4, tol_upd=5e-3), "cfg4": dict(tol_out=2e-4, tol_loss=2e-4, tol_upd=5e-3),
            "main": dict(tol_out=2e-4, tol_loss=1e-3, tol_upd=5e-3),
            "cfg5": dict(tol_out=1e-3, tol_loss=1e-3, tol_upd=5e-2)}
# cfg4's third critic loss is the WGAN-GP term mean((|g|-1)^2) with |g| ~ 0.45 at init: d(term)/term = 2|g|/(|g|-1) * d|g|/|g|
# ~ -1.6x the relative error of the gradient norm, itself the product of a 12-conv double backward in bf16 (measured
# 2.7 % on the term = 1.6 % on |g|; the fp32 engine matches to 2e-4, and the engine is exact on bf16-rounded operands,
# test_conv_family_tcgen05_exact_on_rounded_operands) -> 5e-2 on the losses of this config.
TOL_BF16 = {"cfg1": dict(tol_out=1e-2, tol_loss=2e-2, tol_upd=2e-2), "cfg4": dict(tol_out=1e-2, tol_loss=5e-2, tol_upd=2e-2),
            "cfg5": dict(tol_out=5e-2, tol_loss=None, tol_upd=None, chain=False)}


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["cfg1", "cfg4", "cfg5", "main"])
def test_config_cuda_fp32_matches_reference_golden(monkeypatch, name):
    monkeypatch.setenv("GANB200_PRECISION", "fp32")
    compare(name, torch.device("cuda"), **TOL_FP32[name])


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["cfg1", "cfg4", "cfg5"])
def test_config_cuda_tcgen05_matches_reference_golden(monkeypatch, name):
    """bf16 operands, fp32 accumulation: north_star's 1e-2 class on outputs and losses.  Adam's first update is
    lr*sign(g)-like (|update| ~ lr per element whatever the gradient scale), so the update norms stay tight."""
    monkeypatch.setenv("GANB200_PRECISION", "auto")
    compare(name, torch.device("cuda"), **TOL_BF16[name])


@pytest.mark.gpu
@pytest.mark.parametrize("name", ZOO)
def test_zoo_cuda_fp32_matches_reference_golden(monkeypatch, name):
    monkeypatch.setenv("GANB200_PRECISION", "fp32")
    # Adam's first update is lr * g / (|g| + eps): its norm follows the ABSOLUTE rounding error of tiny gradients.  Measured
    # on B200 against the reference's CPU run: rir_sn 1.9e-2 (a conv bias feeding a BatchNorm: analytically zero gradient),
    # dense4 1.7e-2 (one 7x7-branch weight with a 1e-9 gradient), shuffle 2e-2; everything else within 5e-3.
    loose = {"shuffle": 2e-2, "rir_sn": 2.5e-2, "dense4": 2.5e-2}
    tol = {**dict(tol_out=2e-4, tol_loss=1e-3, tol_upd=5e-3, tol_gloss=5e-2), **({"tol_upd": loose[name]} if name in loose else {})}
    if name.startswith("tree"):   # gradients at Adam's eps: the absolute floor measured on the reference itself (TOL_CPU), doubled
        tol["floor"] = max(2 * TOL_CPU[name]["floor"], 1e-5)
    if name == "tree5":
        tol["tol_out"] = 5e-4
    if TOL_CPU.get(name, {}).get("chain") is False:
        tol["chain"] = False
    compare(name, torch.device("cuda"), **tol)


@pytest.mark.gpu
@pytest.mark.parametrize("precision,tol", [("fp32", 1e-4), ("auto", 2e-2)])
def test_cfg5_blockwise(monkeypatch, precision, tol):
    """generator_14 and discriminator_7_wgangpsn stage by stage: every block / stage runs on the GPU kernels and on the
    CPU restatement from the SAME input (the CPU chain's), so conditioning of the whole chain does not enter."""
    import copy
    from gan_b200 import kernels
    from oracle.kernels_ref import RefBackend
    from discriminators import discriminator_7_wgangpsn as dmod
    from generators import generator_14 as gmod
    monkeypatch.setenv("GANB200_PRECISION", precision)
    torch.manual_seed(0)
    G, D = gmod.Generator(256), dmod.Discriminator()
    Gc, Dc = copy.deepcopy(G).cuda(), copy.deepcopy(D).cuda()
    gen = torch.Generator().manual_seed(5)
    z = torch.randn(8, 256, 1, 1, generator=gen)
    x = torch.randn(8, 3, 64, 64, generator=gen).clamp_(-1, 1)
    g_stages = [(n, getattr(G, n), getattr(Gc, n)) for n in ("block0", "block1", "block2", "block3", "block4", "to_rgb")]
    d_stages = [(n, getattr(D, n), getattr(Dc, n)) for n in ("pre", "stage2", "stage3", "stage4", "stage5", "conv5")]
    worst = {}
    for stages, t in ((g_stages, z), (d_stages, x)):
        for n, cpu_mod, gpu_mod in stages:
            prev = kernels.set_backend(RefBackend())
            try:
                with torch.no_grad():
                    want = cpu_mod(t)
            finally:
                kernels.set_backend(prev)
            with torch.no_grad():
                got = gpu_mod(t.cuda())
            worst[n] = rel(got.cpu(), want)
            t = want
    assert all(v < tol for v in worst.values()), worst
