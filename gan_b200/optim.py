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
        """entries: list of (p, g, m, v) tensors -> (device table, device block map, n_blocks).
        The pinned staging buffers and the device tables are allocated once per key and re-filled when a pointer
        changes, so that no host allocation happens while a CUDA graph is being captured."""
        sig = tuple((p.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr(), p.numel()) for p, g, m, v in entries)
        hit = self._cache.get(key)
        if hit is not None and hit["sig"] == sig:
            return hit["tab"], hit["blk"], hit["n_blocks"]
        vals, blocks = [], []
        for i, (p, g, m, v) in enumerate(entries):
            numel = p.numel()
            vals += [p.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr(), numel]
            blocks += [(i, c) for c in range((numel + CHUNK - 1) // CHUNK)]
        cuda = device.type == "cuda"
        rebuilt = hit is None or hit["tab_host"].numel() != len(vals) or hit["blk_host"].numel() != 2 * len(blocks)
        if rebuilt:
            hit = {"tab_host": torch.empty(len(vals), dtype=torch.int64, pin_memory=cuda),
                   "blk_host": torch.empty((len(blocks), 2), dtype=torch.int32, pin_memory=cuda),
                   "tab": torch.empty(len(vals), dtype=torch.int64, device=device),
                   "blk": torch.empty((len(blocks), 2), dtype=torch.int32, device=device), "event": None}
        if rebuilt or hit.get("blocks") != blocks:
            # the CTA -> (tensor, chunk) map follows the SIZES of the entries: re-uploaded whenever they change, also when the
            # entry and block counts happen to stay the same (a different set of live parameters)
            hit["blk_host"].copy_(torch.tensor(blocks, dtype=torch.int32).reshape(-1, 2))
            hit["blk"].copy_(hit["blk_host"], non_blocking=True)
            hit["blocks"] = blocks
        capturing = cuda and torch.cuda.is_current_stream_capturing()
        if hit["event"] is not None and not capturing:  # (a capture starts after a device synchronize)
            hit["event"].synchronize()  # the previous async upload has read the staging buffer
        hit["tab_host"].copy_(torch.tensor(vals, dtype=torch.int64))
        hit["tab"].copy_(hit["tab_host"], non_blocking=True)
        if cuda and not capturing:
            hit["event"] = torch.cuda.Event()
            hit["event"].record()
        hit["sig"], hit["n_blocks"] = sig, len(blocks)
        self._cache[key] = hit
        return hit["tab"], hit["blk"], len(blocks)

    def load_state_dict(self, state_dict):
        """torch.optim semantics; the pointer tables and the device-side step counters are rebuilt on the next step."""
        super().load_state_dict(state_dict)
        self._cache.clear()
        self._step_dev.clear()

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        k = KN.backend()
        KN.bump_param_epoch()   # the kernel below rewrites parameters behind torch's back: cached packed weights are stale
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
                if not hasattr(k, "adam_reference") and (p.dtype != torch.float32 or not p.is_contiguous()):
                    raise RuntimeError(f"FusedAdam updates contiguous float32 parameters, got {p.dtype} "
                                       f"{'contiguous' if p.is_contiguous() else 'strided'} {tuple(p.shape)}")
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
                tab, blk, n_blocks = self._table((gi, 0 if len(by_step) == 1 else step), entries, device)
                step_dev = None
                if self.capturable:
                    dkey = (gi, 0 if len(by_step) == 1 else step)   # one device counter per (group, step-count bucket)
                    step_dev = self._step_dev.get(dkey)
                    if step_dev is None:
                        step_dev = torch.full((1,), step - 1, dtype=torch.int64, device=device)
                        self._step_dev[dkey] = step_dev
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
