"""Random draws of the training step (label noise, z, GP eps, StyleConv noise — SURVEY F9 / §3.2 'RNG order').

Default: the same torch.rand / torch.randn calls, same shapes, same order as the reference, on the tensor's own
device generator.  Parity mode (`GANB200_PARITY=1` or `host_stream(True)`): the draw is made on the CPU generator
and copied to the device, so that a CUDA run consumes exactly the random stream of the reference's CPU run and can
be compared with the golden vectors under tests/golden/ (CPU and CUDA torch.randn streams differ)."""
import os

import torch

_HOST = [os.environ.get("GANB200_PARITY", "0") == "1"]


def host_stream(on=True):
    prev = _HOST[0]
    _HOST[0] = bool(on)
    return prev


def randn(*shape, device=None):
    if _HOST[0] and device is not None and torch.device(device).type == "cuda":
        return torch.randn(*shape).to(device)
    return torch.randn(*shape, device=device)


def rand(*shape, device=None):
    if _HOST[0] and device is not None and torch.device(device).type == "cuda":
        return torch.rand(*shape).to(device)
    return torch.rand(*shape, device=device)
