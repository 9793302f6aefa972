"""TEST INFRASTRUCTURE — a checkpoint written by the UNMODIFIED reference (checkpoint interop).

    python oracle/make_golden_ckpt.py          (build container only; run as a script, not with -m)

The reference saves whole-module pickles (train/trainunits.py:58-83): the file names classes by import path
(`generators.generator_3_progan.Generator`, `torch.nn.modules.conv.ConvTranspose2d`, ...).  To get the reference's own
paths into the pickle this script puts /root/reference FIRST on sys.path and takes this repo OFF it (the reference's
packages have no __init__.py — namespace packages — and would lose to this repo's regular packages wherever those sit
on the path), builds a small config-4 pair (generator_3_progan nz=16 ngf=4, discriminator_3_wgangp_progan ndf=4 — the
classes are size-agnostic), and calls the reference's `Units.save_ckpt` / `Units.save_record`.  Outputs:

  tests/golden/ref_ckpt_small.pth      the reference's checkpoint file, byte for byte (~300 KB)
  tests/golden/ref_record_small.pth    the reference's record file
  tests/golden/ref_ckpt_small_io.pt    plain tensors: x, z, D(x), G(z) of the pickled modules (train mode), state_dict
                                       key lists — what a loader must reproduce
"""
import importlib.util
import os
import shutil
import sys
import tempfile

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.environ.get("GAN_REFERENCE_ROOT", "/root/reference")


def main():
    sys.path[:] = [REF] + [p for p in sys.path if os.path.abspath(p or ".") not in (REPO, os.path.join(REPO, "oracle"))]
    sys.dont_write_bytecode = True
    spec = importlib.util.spec_from_file_location("_oracle_reference", os.path.join(REPO, "oracle", "reference.py"))
    ref = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref)
    ref.install_stubs()  # matplotlib / tkinter / torchsummary
    import torch
    from generators import generator_3_progan as gmod
    from discriminators import discriminator_3_wgangp_progan as dmod
    from train import trainunits
    assert os.path.abspath(gmod.__file__).startswith(REF) and os.path.abspath(trainunits.__file__).startswith(REF)

    torch.manual_seed(0)
    G = gmod.Generator(1, 16, 4, 3)
    D = dmod.Discriminator(1, 4, 3)
    gen = torch.Generator().manual_seed(7)
    x = torch.randn(4, 3, 64, 64, generator=gen).clamp_(-1, 1)
    z = torch.randn(4, 16, 1, 1, generator=gen)
    units = trainunits.Units(G, "G3_PROGAN", D, "D3_WGAN_PROGAN", z.clone(), epoch_len=10)
    units.make_record([("d_loss", 0.25), ("g_loss", -0.5)])
    golden = os.path.join(REPO, "tests", "golden")
    cwd = os.getcwd()
    tmp = tempfile.mkdtemp()
    try:
        os.chdir(tmp)
        units.save_ckpt("wgangp", 2, 13)      # -> epoch 3, i 3 with epoch_len 10
        units.save_record("wgangp", 2, 13)
        stem = "G3_PROGAN D3_WGAN_PROGAN wgangp epoch_3 i_3"
        shutil.copy(os.path.join("checkpoint", stem + "_ckpt.pth"), os.path.join(golden, "ref_ckpt_small.pth"))
        shutil.copy(os.path.join("checkpoint records", stem + "_record.pth"),
                    os.path.join(golden, "ref_record_small.pth"))
    finally:
        os.chdir(cwd)
        shutil.rmtree(tmp)
    with torch.no_grad():  # after saving: the forward passes move BatchNorm's running statistics
        io = {"x": x, "z": z, "stem": stem, "keys_g": list(G.state_dict()), "keys_d": list(D.state_dict()),
              "d_out": D(x).clone(), "g_out": G(z).clone()}
    torch.save(io, os.path.join(golden, "ref_ckpt_small_io.pt"))
    for f in ("ref_ckpt_small.pth", "ref_record_small.pth", "ref_ckpt_small_io.pt"):
        print(f, os.path.getsize(os.path.join(golden, f)), "bytes")


if __name__ == "__main__":
    main()
