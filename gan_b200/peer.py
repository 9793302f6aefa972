"""Small data-parallel exchanges over NVLink peer memory (csrc/peer.cu): one single-CTA launch per exchange instead of
torch.stack -> ncclAllReduce -> elementwise tail.

`PeerExchange(group, device)` allocates this rank's exchange buffer through the library, swaps cudaIpc handles with the
other ranks of the (single-node) group over torch.distributed, maps their buffers and keeps the table of the `world`
buffer addresses on the device.  `stats` / `grads` / `mean` are the three exchanges gan_b200.dist installs:

    stats(mean, var)   local BatchNorm (mean, biased var) of equal shards -> global (mean, biased var)
    grads(s1, s2)      BatchNorm backward partial sums -> sums over ranks
    mean(x)            mean over ranks (MiniBatchStdDev scalar and its adjoint)

Vectors longer than the exchange slot (`fits`) stay on NCCL (the grouped style-MLP families: 1 MB per exchange, a handful
per step).  Every rank must issue the same exchanges in the same order on its step stream.  The launches are CUDA-graph
capturable.  `check()` raises if a wait timed out (a peer died) — the kernel never hangs the device.
"""
import ctypes

import torch
import torch.distributed as dist

from . import _lib
from . import kernels as KN

SUM, MEAN, BN_STATS = 0, 1, 2
CAP_FLOATS = 1 << 14          # floats per slot: two vectors of 8 192 channels
CALLS = [0]                   # launches issued by this process (bench.py: peer_calls_per_step)
FUSED = [0]                   # exchanges that ran inside a BatchNorm launch (bench.py: peer_fused_per_step)


class PeerExchange:
    def __init__(self, group=None, device=None, cap_floats=CAP_FLOATS):
        if not (dist.is_available() and dist.is_initialized()):
            raise RuntimeError("PeerExchange needs an initialised process group")
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.device = torch.device(device if device is not None else torch.cuda.current_device())
        self.cap = int(cap_floats)
        self.lib = _lib.load()
        self._opened = []
        self.own = ctypes.c_void_p()
        with torch.cuda.device(self.device):
            # the two collectives below run on every rank whatever happens locally, so that a rank that fails cannot leave
            # the others waiting; the first local error is raised after them
            err, mine = None, None
            try:
                _lib.check(self.lib.ganb200_peer_alloc(self.cap, ctypes.byref(self.own)))
                handle = (ctypes.c_ubyte * 64)()
                _lib.check(self.lib.ganb200_peer_export(self.own, handle))
                mine = bytes(handle)
            except Exception as e:      # noqa: BLE001
                err = e
            handles = [None] * self.world
            dist.all_gather_object(handles, mine, group=group)
            addrs = []
            if err is None and any(h is None for h in handles):
                err = RuntimeError("a rank could not export its exchange buffer")
            if err is None:
                try:
                    for r, h in enumerate(handles):
                        if r == self.rank:
                            addrs.append(self.own.value)
                            continue
                        p = ctypes.c_void_p()
                        raw = (ctypes.c_ubyte * 64).from_buffer_copy(h)
                        _lib.check(self.lib.ganb200_peer_open(raw, ctypes.byref(p)))
                        self._opened.append(p)
                        addrs.append(p.value)
                    self.table = torch.tensor(addrs, dtype=torch.int64).to(self.device)
                    torch.cuda.synchronize(self.device)
                except Exception as e:  # noqa: BLE001
                    err = e
            dist.barrier(group=group)  # nobody signals into a buffer that is not mapped everywhere yet
            if err is not None:
                raise err

    @classmethod
    def connect(cls, group=None, device=None, cap_floats=CAP_FLOATS):
        """-> PeerExchange, or None (on every rank) when cudaIpc peer mappings cannot be set up between the ranks of the
        group (ranks on different nodes, IPC disabled in the container): the caller keeps those exchanges on NCCL.
        Collective: every rank of the group must call it."""
        import sys
        px, why = None, ""
        try:
            px = cls(group, device, cap_floats)
        except Exception as e:      # noqa: BLE001 — any failure here means "no peer path", decided jointly below
            why = f"{type(e).__name__}: {e}"
        ok = torch.tensor([1 if px is not None else 0], dtype=torch.int32,
                          device=torch.device(device if device is not None else torch.cuda.current_device()))
        dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)
        if int(ok.item()) == 1:
            return px
        print(f"[gan_b200.peer] rank {dist.get_rank(group)}: NVLink peer exchange unavailable ({why or 'another rank failed'}); "
              "small exchanges stay on NCCL", file=sys.stderr, flush=True)
        return None

    def count(self):
        """an exchange carried inside another kernel's launch (the fused sync BatchNorm)"""
        FUSED[0] += 1

    def fits(self, n, vectors=1):
        return 0 < n * vectors <= self.cap

    def _launch(self, op, in0, in1, out0, out1):
        st = ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        p = lambda t: ctypes.c_void_p(t.data_ptr()) if t is not None else None
        CALLS[0] += 1
        KN.LAUNCH_COUNT[0] += 1
        _lib.check(self.lib.ganb200_peer_allreduce(ctypes.c_void_p(self.table.data_ptr()), self.rank, self.world, self.cap,
                                                   p(in0), p(in1), p(out0), p(out1), in0.numel(), op, st))

    @staticmethod
    def _f32(t):
        if t.dtype != torch.float32 or not t.is_cuda:
            raise _lib.GanB200Error("peer exchange: float32 CUDA tensors only")
        return t.detach().contiguous()

    def stats(self, mean, var):
        mean, var = self._f32(mean), self._f32(var)
        m, v = torch.empty_like(mean), torch.empty_like(var)
        self._launch(BN_STATS, mean, var, m, v)
        return m, v

    def grads(self, s1, s2):
        s1, s2 = self._f32(s1), self._f32(s2)
        g1, g2 = torch.empty_like(s1), torch.empty_like(s2)
        self._launch(SUM, s1, s2, g1, g2)
        return g1, g2

    def mean(self, x):
        x = self._f32(x)
        y = torch.empty_like(x)
        self._launch(MEAN, x, None, y, None)
        return y

    def sum(self, x):
        x = self._f32(x)
        y = torch.empty_like(x)
        self._launch(SUM, x, None, y, None)
        return y

    def check(self):
        """raises if any exchange since the start gave up waiting for a peer (synchronises the device)"""
        e = ctypes.c_int(0)
        with torch.cuda.device(self.device):
            _lib.check(self.lib.ganb200_peer_error(self.own, ctypes.byref(e)))
        if e.value:
            raise RuntimeError("peer exchange: a rank did not arrive within the time limit")

    def close(self):
        if self.own is None:
            return
        with torch.cuda.device(self.device):
            torch.cuda.synchronize(self.device)
            dist.barrier(group=self.group)     # nobody unmaps while a peer may still read
            for p in self._opened:
                self.lib.ganb200_peer_close(p)
            self._opened = []
            dist.barrier(group=self.group)
            self.lib.ganb200_peer_free(self.own)
        self.own = None
