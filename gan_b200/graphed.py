"""Whole-step CUDA graphs for the training iteration.

One iteration of the final pair is tens of thousands of kernel launches (the reference issues
~31 000 ATen calls per G13_5 forward alone); issued one by one from Python the step is bound by the host, not by the
GPU.  `GraphedStep` runs a few eager warm-up iterations (optimizer state, kernel attributes, BCELoss range check — the
eager iterations raise exactly where the reference raises), then captures `discriminator_trainstep` and
`generator_trainstep` — forward, the create_graph double backward, backward and the fused Adam update — into two CUDA
graphs that share one memory pool, and replays them.  Every launch inside is a libganb200 kernel on the capturing
stream; nothing synchronises, workspaces and tensor maps are fixed at capture time, the Adam step count lives on the
device.

Construction has no side effect on the model: parameters, buffers (BatchNorm running statistics, spectral-norm u/v),
Adam moments / step counts and the Philox noise counter are snapshotted first and restored after the capture, so the
first replay is iteration 1 from the state the trainer was handed over in — `N` replays equal `N` eager iterations
(tests/test_graphed_step.py).

BCELoss range check (reference: torch.nn.BCELoss raises "all elements of input should be between 0 and 1"; train/gan.py,
ganR2.py).  A captured graph cannot read a flag back, so inside the graph the BCE kernel sets a sticky device flag
(ops.BceFn.sticky_flag); it is polled every `poll_every` replays and by `check_bce_range()`, which raises the same
RuntimeError — late by at most `poll_every` iterations, never silently.

Drop-in use:  T = train.ganR2.Train(...);  S = GraphedStep(T, batch, images=first_batch);
S.discriminator_trainstep(images, b);  S.generator_trainstep(b)  — same return values as the Train methods (static
output tensors, overwritten by the next replay).  Lazy-penalty recipes (wganlazygp*) alternate two different step
bodies and are not graphed.
"""
import torch

from . import kernels as KN
from . import ops
from . import rng


class StateSnapshot:
    """Device-side copy of everything one training iteration mutates (both networks' parameters and buffers, both
    optimisers' moments and host / device step counters, the Philox noise counter); `restore()` copies it back in place,
    so CUDA graphs captured over those tensors stay valid."""

    def __init__(self, trainer):
        T = self.T = trainer
        self.live, self.saved = [], []
        for net in (T.generator, T.discriminator):
            for t in list(net.parameters()) + list(net.buffers()):
                self.live.append(t)
                self.saved.append(t.detach().clone())
        self.opt = []
        for opt in (T.optimizer_G, T.optimizer_D):
            st = {p: {k: (v.detach().clone() if torch.is_tensor(v) else v) for k, v in s.items()}
                  for p, s in opt.state.items()}
            dev = {k: v.clone() for k, v in getattr(opt, "_step_dev", {}).items()}
            self.opt.append((opt, st, dev))
        self.rng = {k: v.clone() for k, v in rng._states.items()}

    @torch.no_grad()
    def restore(self):
        KN.bump_param_epoch()
        torch._foreach_copy_([t.detach() for t in self.live], self.saved)
        for opt, st, dev in self.opt:
            for p, s in opt.state.items():
                old = st.get(p)
                for k, v in s.items():
                    if torch.is_tensor(v):
                        if old is not None and k in old:
                            v.copy_(old[k])
                        else:
                            v.zero_()       # state created after the snapshot (first step): back to "never stepped"
                    elif k == "step":
                        s[k] = old[k] if old is not None else 0
            for k, v in getattr(opt, "_step_dev", {}).items():
                if k in dev:
                    v.copy_(dev[k])
                else:
                    v.zero_()
        for k, v in rng._states.items():
            if k in self.rng:
                v.copy_(self.rng[k])
            else:
                v[1:].zero_()               # noise counter created after the snapshot: back to forward 0


_WARMUP_STREAMS = {}


def _warmup_stream(dev):
    """One warm-up stream per device for every GraphedStep of the process (csrc/det_reduce.cuh hands each stream that
    launches a split reduction its own scratch region, and there are 16)."""
    key = torch.device(dev).index or 0
    if key not in _WARMUP_STREAMS:
        _WARMUP_STREAMS[key] = torch.cuda.Stream(device=dev)
    return _WARMUP_STREAMS[key]


class GraphedStep:
    def __init__(self, trainer, batch, warmup=3, images=None, poll_every=16):
        T = self.T = trainer
        rc = T.recipe
        if rc.lazy != 1:
            raise ValueError("lazy-penalty recipes are not graph-capturable (two step bodies)")
        dev = T.device
        self.batch = batch
        self.poll_every = poll_every
        self._replays = 0
        needs_grad = bool(rc.r1)
        if images is not None:
            if images.shape[0] != batch:
                raise ValueError(f"images has batch {images.shape[0]}, GraphedStep was asked for {batch}")
            self.images = images.detach().to(dev, copy=True).contiguous()
        else:
            # shape only: the warm-up below trains on zeros, which is harmless because the state is restored afterwards
            self.images = torch.zeros(batch, 3, 64, 64, device=dev)
        if needs_grad:
            self.images.requires_grad_()
        self._flag = ops.BceFn.sticky_flag(dev) if rc.bce else None
        self.snapshot = StateSnapshot(T)
        for opt in (T.optimizer_G, T.optimizer_D):
            opt.capturable = True
        side = _warmup_stream(dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            for _ in range(max(warmup, 1)):
                T.discriminator_trainstep(self.images, batch)
                T.generator_trainstep(batch)
                self.images.grad = None
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        # the eager warm-up's cached blocks are of no use to the graphs' private pool: hand them back first
        torch.cuda.empty_cache()
        # each graph must re-pack every weight it uses at its own start (a replay may follow ANY other replay): start the
        # captures with the packed-weight cache cold
        KN.bump_param_epoch()
        self.graph_d = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph_d, capture_error_mode="thread_local"):
            self.d_out = T.discriminator_trainstep(self.images, batch)
            self.images.grad = None
        KN.bump_param_epoch()
        self.graph_g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph_g, pool=self.graph_d.pool(), capture_error_mode="thread_local"):
            self.g_out = T.generator_trainstep(batch)
        # the capture bumped the host-side step counts without running anything; the warm-up really trained: undo both
        self.snapshot.restore()
        if self._flag is not None:
            self._flag.zero_()
        torch.cuda.synchronize(dev)

    def discriminator_trainstep(self, images, b_size, idx=0):
        assert b_size == self.batch
        if images is not self.images:
            with torch.no_grad():
                self.images.copy_(images, non_blocking=True)
        self.graph_d.replay()
        KN.bump_param_epoch()   # the replay moved the parameters: packed weights cached by eager code are stale
        self._advance(self.T.optimizer_D)
        return self.d_out

    def generator_trainstep(self, b_size):
        assert b_size == self.batch
        self.graph_g.replay()
        KN.bump_param_epoch()
        self._advance(self.T.optimizer_G)
        self._replays += 1
        if self.poll_every and self._replays % self.poll_every == 0:
            self.check_bce_range()
        return self.g_out

    def check_bce_range(self):
        """Raises the reference's BCELoss error if any replay since the last check fed it a value outside [0, 1]
        (one 4-byte D2H read: synchronises the stream)."""
        if self._flag is None:
            return
        if int(self._flag.item()) != 0:
            self._flag.zero_()
            raise RuntimeError(ops.BceFn.RANGE_ERROR)

    def restore_initial_state(self):
        """Back to the state the trainer had when this object was built (bench.py: keeps the ganR2 weight patch valid
        for any number of timed iterations)."""
        self.snapshot.restore()

    @staticmethod
    def _advance(opt):
        # keep the host-side torch.optim step counters in line with the device counter the graph bumps
        for st in opt.state.values():
            if "step" in st:
                st["step"] += 1
