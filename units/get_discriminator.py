"""units.get_discriminator — drop-in for the reference's units/get_discriminator.py:15-90: `get_<name>(...)` builds a
discriminator / critic and moves it to `device`, same names and argument lists.  Lazy imports; every discriminator file
of the reference has a drop-in here.

`get_9_4` is an addition (the reference's main.py builds discriminator_9_4 by hand).  discriminator_9_2 / _9_3 have no
factory in the reference either; they are importable modules (`discriminators.discriminator_9_3.Discriminator()`)."""
import importlib
import math

_CARRIED = ("discriminator_1", "discriminator_2", "discriminator_3_wgangp", "discriminator_3_wgangp_progan",
            "discriminator_3_wgangpsn", "discriminator_4", "discriminator_5", "discriminator_5_wgangp", "discriminator_5_wgangpsn",
            "discriminator_6_wgangp", "discriminator_7_wgangp", "discriminator_7_wgangpsn", "discriminator_8", "discriminator_9",
            "discriminator_9_2", "discriminator_9_3", "discriminator_9_4", "discriminator_9_small")


def _module(name):
    if name not in _CARRIED:
        raise NotImplementedError(
            f"discriminators.{name} is outside this drop-in's scope (carried: {', '.join(_CARRIED)})")
    return importlib.import_module("discriminators." + name)


def get_1(device, image_size):
    return _module("discriminator_1").Discriminator(image_size=image_size).to(device)


def _dcgan_factory(fn_name, name):
    def get(ngpu, device, ndf, nc):
        return _module(name).Discriminator(ngpu=ngpu, ndf=ndf, nc=nc).to(device)
    get.__name__ = fn_name
    return get


def _plain_factory(fn_name, name, *args):
    def get(ngpu, device):
        return _module(name).Discriminator(*args).to(device)
    get.__name__ = fn_name
    return get


def get_2(ngpu, device, ndf, nc):
    """reference :20-25 — the only discriminator factory that applies the DCGAN `weights_init`."""
    from units import Weight_Initialization
    discriminator = _module("discriminator_2").Discriminator(ngpu=ngpu, ndf=ndf, nc=nc).to(device)
    discriminator.apply(Weight_Initialization.weights_init)
    return discriminator


get_3_wgan = _dcgan_factory("get_3_wgan", "discriminator_3_wgangp")
get_3_wgan_progan = _dcgan_factory("get_3_wgan_progan", "discriminator_3_wgangp_progan")
get_3_wgansn = _dcgan_factory("get_3_wgansn", "discriminator_3_wgangpsn")
get_4 = _plain_factory("get_4", "discriminator_4")
get_5 = _plain_factory("get_5", "discriminator_5")
get_5_wgan = _plain_factory("get_5_wgan", "discriminator_5_wgangp")
get_5_wgansn = _plain_factory("get_5_wgansn", "discriminator_5_wgangpsn")
get_6_wgan = _plain_factory("get_6_wgan", "discriminator_6_wgangp")
get_7_wgan = _plain_factory("get_7_wgan", "discriminator_7_wgangp")
get_7_wgansn = _plain_factory("get_7_wgansn", "discriminator_7_wgangpsn")
get_8 = _plain_factory("get_8", "discriminator_8", int(math.log2(64)))
get_9 = _plain_factory("get_9", "discriminator_9")
get_9_4 = _plain_factory("get_9_4", "discriminator_9_4")
