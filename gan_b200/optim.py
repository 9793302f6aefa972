"""Fused multi-tensor Adam / AdamW over libganb200 (replaces torch.optim.Adam in train/trainunits.py:18-19 and
torch.optim.AdamW in train/wgangp.py:17-18 of the reference).

torch.optim semantics kept:
  * a parameter whose .grad is None is skipped entirely — no state is created, its step count does not
    advance (G13_5's 519 dead StyleConv biases and the frozen Smooth kernels never get state);
  * per-parameter 1-based step counts; bias correction from them; eps added outside the sqrt;
  * Adam: L2 weight decay folded into the gradient; AdamW: decoupled decay (default 0.01).
One kernel launch updates every live tensor of a param group that shares a step count (normally all of them):
the host builds a table of (p, g, m, v, n) device pointers and a CTA -> (tensor, chunk) map, cached while the
pointers stay the same (they do inside a replayed CUDA graph; with zero_grad(set_to_none=True) in eager mode the
gradient tensors are new every iteration and the table — a few hundred KB for G13_5 — is rebuilt).
"""
import torch

from . import kernels as KN

CHUNK = 32768  # elements per CTA (a multiple of 4); 128 KiB per operand


class FusedAdam(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, decoupled=False,
                 grad_scale=1.0, capturable=False):
        defaults = dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, decoupled=decoupled)
        super().__init__(params, defaults)
        self.grad_scale = grad_scale
        # capturable: the step count lives on the device so that step() can sit inside a CUDA graph
        self.capturable = capturable
        self._cache = {}
        self._step_dev = {}
        self.pre_step = None  # data parallel: gradient all-reduce (gan_b200.dist.GradSync)

    # -- table ----------------------------------------------------------------------------------------
    def _table(self, key, entries, device):
        """entries: list of (p, g, m, v) tensors."""
        sig = tuple((p.data_ptr(), g.data_ptr()) for p, g, _, _ in entries)
        hit = self._cache.get(key)
        if hit is not None and hit[0] == sig:
            return hit[1:]
        n = len(entries)
        tab_host = torch.empty(n * 5, dtype=torch.int64, pin_memory=device.type == "cuda")
        blocks = []
        vals = []
        for i, (p, g, m, v) in enumerate(entries):
            numel = p.numel()
            vals += [p.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr(), numel]
            for c in range((numel + CHUNK - 1) // CHUNK):
                blocks.append((i, c))
        tab_host.copy_(torch.tensor(vals, dtype=torch.int64))
        blk_host = torch.tensor(blocks, dtype=torch.int32).reshape(-1, 2)
        if device.type == "cuda":
            blk_host = blk_host.pin_memory()
        tab = tab_host.to(device, non_blocking=True)
        blk = blk_host.to(device, non_blocking=True)
        self._cache[key] = (sig, tab, blk, len(blocks), (tab_host, blk_host))
        return tab, blk, len(blocks), None

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        k = KN.backend()
        if self.pre_step is not None:
            self.pre_step()
        for gi, group in enumerate(self.param_groups):
            b1, b2 = group["betas"]
            by_step = {}
            for p in group["params"]:
                if p.grad is None:
                    continue
                if p.grad.is_sparse:
                    raise RuntimeError("FusedAdam does not support sparse gradients")
                st = self.state[p]
                if len(st) == 0:
                    st["step"] = 0
                    st["exp_avg"] = torch.zeros_like(p, memory_format=torch.contiguous_format)
                    st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.contiguous_format)
                st["step"] += 1
                g = p.grad if p.grad.is_contiguous() else p.grad.contiguous()
                by_step.setdefault(st["step"], []).append((p, g, st["exp_avg"], st["exp_avg_sq"]))
            for step, entries in by_step.items():
                device = entries[0][0].device
                if hasattr(k, "adam_reference"):  # CPU restatement installed by tests
                    for p, g, m, v in entries:
                        k.adam_reference(p, g, m, v, group["lr"], b1, b2, group["eps"], group["weight_decay"],
                                         group["decoupled"], step, self.grad_scale)
                    continue
                tab, blk, n_blocks, _ = self._table((gi, 0 if len(by_step) == 1 else step), entries, device)
                step_dev = None
                if self.capturable:
                    step_dev = self._step_dev.get(gi)
                    if step_dev is None:
                        step_dev = torch.full((1,), step - 1, dtype=torch.int64, device=device)
                        self._step_dev[gi] = step_dev
                    k.counter_add(step_dev, 1)
                k.adam_blocks(tab, blk, n_blocks, CHUNK, group["lr"], b1, b2, group["eps"], group["weight_decay"],
                              group["decoupled"], step, self.grad_scale, step_dev=step_dev)
        return loss


def Adam(params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, **kw):
    """torch.optim.Adam signature (the arguments the reference uses)."""
    return FusedAdam(params, lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, decoupled=False, **kw)


def AdamW(params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-2, **kw):
    """torch.optim.AdamW signature; default decoupled decay 0.01."""
    return FusedAdam(params, lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, decoupled=True, **kw)
