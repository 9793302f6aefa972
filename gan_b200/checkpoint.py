"""Checkpoint interop with the reference (reference train/trainunits.py:58-131).

The reference stores WHOLE-MODULE pickles: `{'generator': <nn.Module>, 'discriminator': <nn.Module>, names, 'method',
'epoch', 'i'}`.  The pickle names every class by import path.  Because the drop-in keeps the reference's paths
(`generators.generator_13_5.Generator`, every helper class), a file written by the reference unpickles here — but into
a HYBRID: the outer classes are this repo's, the leaves are the stock `torch.nn` layers the reference had built
(`nn.ConvTranspose2d`, `nn.BatchNorm2d`, `nn.PReLU`, ...), whose forward is ATen, not libganb200.  Running that object
would be a silent library fallback, so a loaded module is never used as it comes out of the file:

  adopt(loaded, template)   -> a module with `template`'s (native) layer tree and `loaded`'s parameters and buffers.
                               `template` is the module the caller already holds (built by units/get_*.py with the
                               same constructor arguments, as main.py does before it calls load_*_ckpt).
  is_native(loaded, template) -> the two module trees have the same class at every name.

A file written by this repo's `save_ckpt` holds native modules and loads as is (it needs this package importable;
the reference can read it wherever the drop-in is installed).  `save_ckpt` also stores both state_dicts and, on
request, both optimiser states under extra keys the reference's loader ignores; the tensors alias the modules'
storages inside one `torch.save`, so the state_dicts add no bytes to the file.
"""
import copy

import torch.nn as nn

def _cls(m):
    """Import path of a module's class; torch's dynamically created `Parametrized<X>` classes count as <X>."""
    t = type(m)
    if t.__module__ == "torch.nn.utils.parametrize" and t.__name__.startswith("Parametrized") and t.__bases__:
        t = t.__bases__[0]
    return t.__module__ + "." + t.__qualname__


def foreign_layers(loaded, template):
    """(name, class in the file, class in the native tree) wherever the two module trees differ.  Empty when `loaded`
    was saved by this package; for a file written by the reference it lists the stock `torch.nn` leaves
    (torch.nn.modules.conv.ConvTranspose2d vs gan_b200.layers.ConvTranspose2d, ...)."""
    have = {n: _cls(m) for n, m in loaded.named_modules()}
    want = {n: _cls(m) for n, m in template.named_modules()}
    diff = [(n, have.get(n), c) for n, c in want.items() if have.get(n) != c]
    diff += [(n, c, None) for n, c in have.items() if n not in want]
    return diff


def is_native(loaded, template):
    return not foreign_layers(loaded, template)


def adopt(loaded, template):
    """The module to use after `torch.load`: `loaded` itself when it is native, else a deep copy of `template` carrying
    `loaded`'s state (same keys and shapes required: the error names what differs).  `template` is left untouched —
    the reference REPLACES `self.generator` on load (trainunits.py:101), it does not update it in place."""
    if not isinstance(loaded, nn.Module):
        raise TypeError(f"checkpoint entry is a {type(loaded).__name__}, not a module")
    if is_native(loaded, template):
        return loaded
    state = loaded.state_dict()
    want = template.state_dict()
    missing = [k for k in want if k not in state]
    extra = [k for k in state if k not in want]
    shapes = [(k, tuple(state[k].shape), tuple(want[k].shape)) for k in want
              if k in state and tuple(state[k].shape) != tuple(want[k].shape)]
    if missing or extra or shapes:
        raise RuntimeError(
            "reference checkpoint does not fit the module built for it "
            f"({type(template).__module__}.{type(template).__name__}): missing {missing[:5]}, unexpected {extra[:5]}, "
            f"shape (file, built) {shapes[:5]} — build the pair with the constructor arguments the checkpoint was "
            "trained with (units/get_generators.py / get_discriminator.py) before loading")
    fresh = copy.deepcopy(template)
    fresh.load_state_dict(state, strict=True)
    dev = next(iter(state.values())).device if state else None
    if dev is not None:
        fresh.to(dev)  # torch.load restores tensors on the device they were saved from; so does the reference
    fresh.train(loaded.training)
    return fresh


def load_into(loaded, module):
    """Opt-in fix of the reference's stale-optimiser behaviour: copy `loaded`'s state INTO `module` (whose parameters
    the optimiser already holds) instead of replacing the object."""
    module.load_state_dict(loaded.state_dict(), strict=True)
    return module
