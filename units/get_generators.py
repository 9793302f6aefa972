"""units.get_generators — drop-in for the reference's units/get_generators.py:12-111: `get_<name>(...)` builds a
generator and moves it to `device`, same names and argument lists.  Generators are imported lazily (the reference
imports fourteen model files at module import); every factory of the reference builds — the one model file without a
drop-in is generators/generator_13_1.py, which no factory names (`_module` raises NotImplementedError for it).

`get_13_5` and `get_14` are additions: the reference has no factory for its two newest generators (its main.py builds
them by hand)."""
import importlib

_CARRIED = ("generator_13_1", "generator_1", "generator_2", "generator_3_progan", "generator_4", "generator_5", "generator_6", "generator_7", "generator_8", "generator_9", "generator_10", "generator_10_2", "generator_11", "generator_12", "generator_13", "generator_13_2", "generator_13_3", "generator_13_4", "generator_13_5", "generator_14")


def _module(name):
    if name not in _CARRIED:
        raise NotImplementedError(f"generators.{name} is outside this drop-in's scope (carried: {', '.join(_CARRIED)})")
    return importlib.import_module("generators." + name)


def get_1(device, z_dim, target_image_size):
    return _module("generator_1").Generator(z_dim=z_dim, target_image_size=target_image_size).to(device)


def get_2(ngpu, device, nz, ngf, nc):
    """reference :17-22 — the only generator factory that applies the DCGAN `weights_init`."""
    from units import Weight_Initialization
    generator = _module("generator_2").Generator(ngpu=ngpu, nz=nz, ngf=ngf, nc=nc).to(device)
    generator.apply(Weight_Initialization.weights_init)
    return generator


def get_3_progan(ngpu, device, nz, ngf, nc):
    return _module("generator_3_progan").Generator(ngpu=ngpu, nz=nz, ngf=ngf, nc=nc).to(device)


def _z_factory(name):
    def get(ngpu, device, z_dim=128):
        return _module(name).Generator(z_dim=z_dim).to(device)
    get.__name__ = "get_" + name.split("_", 1)[1]
    get.__doc__ = f"generators/{name}.py Generator(z_dim) on `device` (ngpu is ignored, as in the reference)."
    return get


for _n in ("4", "5", "6", "7", "8", "9", "10", "10_2", "11", "12", "13", "13_5", "14"):
    globals()["get_" + _n] = _z_factory("generator_" + _n)
del _n
