"""Checkpoint interop: files written by the UNMODIFIED reference's Units.save_ckpt / save_record
(tests/golden/ref_ckpt_small.pth, ref_record_small.pth — oracle/make_golden_ckpt.py) load through the drop-in's
Units.load_*_ckpt into NATIVE modules (no stock torch.nn layer left in the tree) that reproduce the reference's
outputs; the drop-in's own files round-trip; the opt-in flags restore optimiser state and keep the optimiser bound."""
import os
import shutil

import pytest
import torch

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def rel(a, b):
    a, b = a.double(), b.double()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def small_pair():
    from discriminators import discriminator_3_wgangp_progan as dmod
    from generators import generator_3_progan as gmod
    torch.manual_seed(123)  # NOT the seed of the file: the weights must come from the checkpoint
    return gmod.Generator(1, 16, 4, 3), dmod.Discriminator(1, 4, 3)


def make_units(z):
    from train import trainunits
    G, D = small_pair()
    return trainunits.Units(G, "G3_PROGAN", D, "D3_WGAN_PROGAN", z, epoch_len=10)


@pytest.fixture
def workdir(tmp_path, monkeypatch):
    monkeypatch.chdir(tmp_path)
    os.makedirs("checkpoint")
    io = torch.load(os.path.join(GOLDEN, "ref_ckpt_small_io.pt"))
    shutil.copy(os.path.join(GOLDEN, "ref_ckpt_small.pth"), os.path.join("checkpoint", io["stem"] + "_ckpt.pth"))
    return io


def test_reference_pickle_unpickles_into_a_hybrid(workdir):
    """Why adoption is needed: the reference's file names stock torch layers; loaded raw they would run ATen."""
    from gan_b200 import checkpoint as C
    raw = torch.load(os.path.join("checkpoint", workdir["stem"] + "_ckpt.pth"), weights_only=False)
    G, D = small_pair()
    foreign = C.foreign_layers(raw["generator"], G)
    assert ("main.0.0", "torch.nn.modules.conv.ConvTranspose2d", "gan_b200.layers.ConvTranspose2d") in foreign
    assert not C.is_native(raw["discriminator"], D) and C.is_native(G, G)
    assert raw["generator_name"] == "G3_PROGAN" and raw["epoch"] == 3 and raw["i"] == 3 and raw["method"] == "wgangp"


def test_load_reference_checkpoint_native_and_exact(ref_backend, workdir):
    from gan_b200 import checkpoint as C
    io = workdir
    u = make_units(io["z"].clone())
    g0, d0 = u.generator, u.discriminator
    name = io["stem"] + "_ckpt"
    u.load_generator_ckpt(name)
    u.load_discriminator_ckpt(name)
    assert u.epoch == 3 and u.i == 3 and u.record["epoch"] == 3 and u.record["i"] == 3  # reference :103-105
    # the reference replaces the objects (its optimisers stay bound to the old ones): same here by default
    assert u.generator is not g0 and u.discriminator is not d0
    assert u.optimizer_G.param_groups[0]["params"][0] is next(g0.parameters())
    assert C.is_native(u.generator, g0) and C.is_native(u.discriminator, d0)
    assert list(u.generator.state_dict()) == io["keys_g"] and list(u.discriminator.state_dict()) == io["keys_d"]
    with torch.no_grad():
        assert rel(u.discriminator(io["x"]), io["d_out"]) < 1e-5
        assert rel(u.generator(io["z"]), io["g_out"]) < 1e-5


def test_inplace_flag_keeps_the_optimiser_bound(ref_backend, workdir, monkeypatch):
    monkeypatch.setenv("GANB200_CKPT_INPLACE", "1")
    io = workdir
    u = make_units(io["z"].clone())
    g0, d0 = u.generator, u.discriminator
    u.load_generator_ckpt(io["stem"] + "_ckpt")
    u.load_discriminator_ckpt(io["stem"] + "_ckpt")
    assert u.generator is g0 and u.discriminator is d0
    with torch.no_grad():
        assert rel(u.discriminator(io["x"]), io["d_out"]) < 1e-5


def test_mismatched_pair_is_refused_with_the_reason(ref_backend, workdir):
    from generators import generator_3_progan as gmod
    from train import trainunits
    _, D = small_pair()
    u = trainunits.Units(gmod.Generator(1, 16, 8, 3), "G3_PROGAN", D, "D3_WGAN_PROGAN", workdir["z"], epoch_len=10)
    with pytest.raises(RuntimeError, match="does not fit the module built for it"):
        u.load_generator_ckpt(workdir["stem"] + "_ckpt")


def test_missing_file_prints_and_keeps_the_modules(ref_backend, workdir, capsys):
    u = make_units(workdir["z"].clone())
    g0 = u.generator
    u.load_generator_ckpt("no such file")
    assert "not find ckpt" in capsys.readouterr().out and u.generator is g0  # reference :108-109


def test_reference_record_file_reads(workdir):
    rec = torch.load(os.path.join(GOLDEN, "ref_record_small.pth"), weights_only=False)
    assert rec["record"]["d_loss"] == [0.25] and rec["record"]["g_loss"] == [-0.5]
    assert (rec["epoch"], rec["i"], rec["method"]) == (3, 3, "wgangp")


def test_own_checkpoint_round_trip_with_optimiser_state(ref_backend, workdir, monkeypatch):
    """save_ckpt -> load_*_ckpt of this package's own file: same name scheme as the reference (:73-76), native modules
    load as they are, and with both flags a resumed run continues bit-for-bit like the uninterrupted one."""
    from gan_b200 import ops
    monkeypatch.setenv("GANB200_CKPT_OPTIMIZER", "1")
    monkeypatch.setenv("GANB200_CKPT_INPLACE", "1")
    io = workdir

    def one_step(u):
        u.optimizer_G.zero_grad()
        ops.mean(ops.square(u.generator(io["z"]))).backward()
        u.optimizer_G.step()

    a = make_units(io["z"].clone())
    one_step(a)
    a.make_record([("g_loss", 1.0)])
    a.save_ckpt("wgangp", 2, 13)
    a.save_record("wgangp", 2, 13)
    assert os.path.isfile(os.path.join("checkpoint", io["stem"] + "_ckpt.pth"))
    assert os.path.isfile(os.path.join("checkpoint records", io["stem"] + "_record.pth"))
    assert a.record == {"epoch": 0, "i": 0}  # cleared and re-seeded, reference :92-93
    saved = torch.load(os.path.join("checkpoint", io["stem"] + "_ckpt.pth"), weights_only=False)
    assert {"generator", "generator_name", "discriminator", "discriminator_name", "method", "epoch", "i"} <= set(saved)
    assert list(saved["generator_state"]) == io["keys_g"] and "optimizer_G" in saved

    b = make_units(io["z"].clone())
    with torch.no_grad():  # scramble: everything must come back from the file
        for p in b.generator.parameters():
            p.add_(1.0)
    b.load_generator_ckpt(io["stem"] + "_ckpt")
    b.load_discriminator_ckpt(io["stem"] + "_ckpt")
    for (n, p), (_, q) in zip(a.generator.state_dict().items(), b.generator.state_dict().items()):
        assert torch.equal(p, q), n
    one_step(a)
    one_step(b)  # second Adam step: needs the restored moments and step count to match
    for (n, p), (_, q) in zip(a.generator.named_parameters(), b.generator.named_parameters()):
        assert torch.equal(p, q), n


@pytest.mark.gpu
def test_load_reference_checkpoint_on_gpu(workdir, monkeypatch):
    """The adopted modules run on libganb200 (fp32 engine here: 2e-4 against the reference's CPU outputs)."""
    from gan_b200 import checkpoint as C
    monkeypatch.setenv("GANB200_PRECISION", "fp32")
    io = workdir
    u = make_units(io["z"].clone())
    g0 = u.generator
    u.load_generator_ckpt(io["stem"] + "_ckpt")
    u.load_discriminator_ckpt(io["stem"] + "_ckpt")
    assert C.is_native(u.generator, g0)
    G, D = u.generator.cuda(), u.discriminator.cuda()
    with torch.no_grad():
        assert rel(D(io["x"].cuda()).cpu(), io["d_out"]) < 2e-4
        assert rel(G(io["z"].cuda()).cpu(), io["g_out"]) < 2e-4


def test_every_reference_class_path_exists_in_the_drop_in():
    """A whole-module pickle resolves classes by import path: every class the reference defines in the files of the
    BASELINE configs must exist under the same name here (build container only: reads the reference's sources)."""
    import ast
    import importlib
    from oracle import reference
    if not reference.available():
        pytest.skip("reference tree not present")
    files = [("generators", m) for m in ("generator_1", "generator_3_progan", "generator_13_5", "generator_14")]
    files += [("discriminators", m) for m in ("discriminator_1", "discriminator_3_wgangp", "discriminator_3_wgangp_progan",
                                               "discriminator_7_wgangpsn", "discriminator_9_4")]
    for pkg, mod in files:
        with open(os.path.join(reference.ROOT, pkg, mod + ".py")) as f:
            classes = [n.name for n in ast.parse(f.read()).body if isinstance(n, ast.ClassDef)]
        ours = importlib.import_module(pkg + "." + mod)
        assert [c for c in classes if not isinstance(getattr(ours, c, None), type)] == [], (pkg, mod)


def test_train_loop_side_effects_and_resume(ref_backend, tmp_path, monkeypatch):
    """Train.train() of the drop-in loops (reference train/wgangp.py:73-96, wgan.py:51-75, wganlazygp.py:72-103): image
    grid every 10 iterations, a checkpoint per epoch under the reference's file name, records every 30 iterations for
    the lazy loops, wgan's NameError after its first iteration — and a second run resuming from the file."""
    from train import wgan, wganlazygp, wgangp
    monkeypatch.chdir(tmp_path)
    os.makedirs("generated_images")
    gen = torch.Generator().manual_seed(0)
    batches = [(torch.randn(4, 3, 64, 64, generator=gen).clamp_(-1, 1), torch.zeros(4)) for _ in range(2)]
    G, D = small_pair()
    T = wgangp.Train(batches, torch.device("cpu"), 1, 16, G, "G3_PROGAN", D, "D3_WGAN_PROGAN")
    before = [p.detach().clone() for p in G.parameters()]
    T.train()
    assert any(not torch.equal(a, b) for a, b in zip(before, G.parameters()))
    assert os.path.isfile("generated_images/0-0.png") or os.path.isfile("generated_images/0-0.pt")
    name = "G3_PROGAN D3_WGAN_PROGAN WGANGP epoch_1 i_0_ckpt"
    assert os.path.isfile(os.path.join("checkpoint", name + ".pth"))
    G2, D2 = small_pair()
    T2 = wgangp.Train(batches, torch.device("cpu"), 1, 16, G2, "G3_PROGAN", D2, "D3_WGAN_PROGAN")
    T2.load_generator_ckpt(name)
    T2.load_discriminator_ckpt(name)
    assert T2.epoch == 1 and T2.i == 0
    for a, b in zip(T2.generator.state_dict().values(), G.state_dict().values()):
        assert torch.equal(a, b)
    # lazy loop: records at i % 30 == 0 go to memory and to <file>.txt
    G3, D3 = small_pair()
    T3 = wganlazygp.Train(batches, torch.device("cpu"), 1, 16, G3, "g", D3, "d")
    T3.train()
    assert len(T3.record["Generator loss"]) == 1 and os.path.isfile("wganlazygp.txt")
    # wgan: the reference's loop reads an undefined `gp` right after the first iteration
    G4, D4 = small_pair()
    T4 = wgan.Train(batches, torch.device("cpu"), 1, 16, G4, "g", D4, "d")
    with pytest.raises(NameError):
        T4.train()
    assert all(float(p.detach().abs().max()) <= 1.0 for p in D4.parameters())  # the critic was clipped before the crash
