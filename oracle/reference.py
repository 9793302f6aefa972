"""TEST INFRASTRUCTURE — loads the UNMODIFIED reference (xxxxx23124/-GAN-) from /root/reference as the oracle.

Only tests/, oracle/make_golden.py and bench.py's `--impl reference` / cpu_baseline leg may import this.
/root/reference exists only in the build container; on the GPU box the tree is read from oracle/_ref/, the byte-for-byte
copy that oracle/build_ref.py makes at build time (git-ignored, shipped like a built .so); when neither exists
`available()` is False and callers fall back to the committed fixtures under tests/golden/ (parity) or to the restated
loops (CPU baseline, kind "port").

The reference imports matplotlib, tkinter (through tqdm.tk) and torchsummary at module top
(train/trainunits.py:7, train/ganR2.py:2, discriminators/discriminator_4.py:6); none is installed here, so
permissive stubs are registered in sys.modules first.  No reference file is edited or copied.

Because the drop-in keeps the reference's package names (generators, discriminators, train, units), the two
trees cannot both live in sys.modules under those names: `load(name)` imports a reference module from its
file under the private prefix `_gan_ref.` and resolves the reference's own absolute imports to that prefix.
"""
import importlib.abc
import importlib.util
import os
import sys
import types

_COPY = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")   # written by oracle/build_ref.py (git-ignored)
ROOT = os.environ.get("GAN_REFERENCE_ROOT") or (
    "/root/reference" if os.path.isdir("/root/reference/generators") else _COPY)
PREFIX = "_gan_ref"
_PKGS = ("generators", "discriminators", "train", "units")


def available():
    return os.path.isdir(os.path.join(ROOT, "generators"))


class _Permissive(types.ModuleType):
    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)

        class _Dummy:
            def __init__(self, *a, **k):
                pass

            def __call__(self, *a, **k):
                return self

            def __getattr__(self, n):
                return _Dummy()

        _Dummy.__name__ = name
        setattr(self, name, _Dummy)
        return _Dummy


def install_stubs():
    """matplotlib(.pyplot), tkinter(.ttk), torchsummary, torchviz: only if the real module is missing."""
    for name in ("matplotlib", "matplotlib.pyplot", "tkinter", "tkinter.ttk", "torchsummary", "torchviz"):
        if name in sys.modules:
            continue
        try:
            importlib.import_module(name)
        except Exception:
            m = _Permissive(name)
            m.__path__ = []
            sys.modules[name] = m
            if "." in name:
                setattr(sys.modules[name.split(".")[0]], name.split(".")[1], m)
    plt = sys.modules["matplotlib.pyplot"]
    if isinstance(plt, _Permissive):
        for fn in ("ion", "clf", "imshow", "figure", "savefig", "show", "pause"):
            setattr(plt, fn, lambda *a, **k: None)


class _RefFinder(importlib.abc.MetaPathFinder):
    """Resolves `_gan_ref.<pkg>.<mod>` to files of the reference tree."""

    def find_spec(self, fullname, path=None, target=None):
        if fullname == PREFIX:
            spec = importlib.util.spec_from_loader(fullname, loader=None, is_package=True)
            spec.submodule_search_locations = [ROOT]
            return spec
        if not fullname.startswith(PREFIX + "."):
            return None
        rel = fullname[len(PREFIX) + 1:].split(".")
        base = os.path.join(ROOT, *rel)
        if os.path.isdir(base):
            spec = importlib.util.spec_from_loader(fullname, loader=None, is_package=True)
            spec.submodule_search_locations = [base]
            return spec
        if os.path.isfile(base + ".py"):
            return importlib.util.spec_from_file_location(fullname, base + ".py")
        return None


_finder = None


def load(name):
    """load('discriminators.discriminator_9_4') -> the reference module (cached as _gan_ref.<name>)."""
    global _finder
    if not available():
        raise RuntimeError(f"reference tree not found at {ROOT}")
    install_stubs()
    sys.dont_write_bytecode = True
    if _finder is None:
        _finder = _RefFinder()
        sys.meta_path.insert(0, _finder)
    full = f"{PREFIX}.{name}"
    if full in sys.modules:
        return sys.modules[full]
    # The reference uses absolute imports of its sibling packages (e.g. `from train import trainunits`,
    # `from generators import generator_1`).  While a reference module executes, alias those top-level names
    # to the private copies, then restore whatever the drop-in had registered.
    saved = {k: v for k, v in sys.modules.items() if k.split(".")[0] in _PKGS}
    for k in saved:
        del sys.modules[k]
    for pkg in _PKGS:
        sys.modules[pkg] = importlib.import_module(f"{PREFIX}.{pkg}")
    alias_hook = _AliasFinder()
    sys.meta_path.insert(0, alias_hook)
    try:
        mod = importlib.import_module(full)
    finally:
        sys.meta_path.remove(alias_hook)
        for k in [k for k in sys.modules if k.split(".")[0] in _PKGS]:
            del sys.modules[k]
        sys.modules.update(saved)
    return mod


class _AliasFinder(importlib.abc.MetaPathFinder):
    """While a reference module is being imported: `generators.generator_1` -> `_gan_ref.generators.generator_1`."""

    def find_spec(self, fullname, path=None, target=None):
        parts = fullname.split(".")
        if parts[0] in _PKGS and len(parts) > 1:
            real = importlib.import_module(f"{PREFIX}.{fullname}")
            sys.modules[fullname] = real
            return real.__spec__
        return None
