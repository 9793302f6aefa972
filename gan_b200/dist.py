"""Data-parallel training step: one process per GPU (torchrun), NCCL over NVLink for the exchanges.

The reference is single-device; batch-sharded DP is new.  What must be exchanged so that N ranks with
B/N images each compute the step the reference computes on B images:

  * parameter gradients of G and D: summed over ranks in flat fp32 buckets, the 1/N folded into the fused Adam
    kernel's grad_scale (`GradSync`, installed as the optimiser's pre-step hook);
  * BatchNorm batch statistics in G (`sync-BN`): per-channel (mean, biased var) forward and (sum dy, sum dy*xhat)
    backward are combined across ranks (hooks in gan_b200.ops);
    — each of these small exchanges is ONE launch over NVLink peer memory (gan_b200/peer.py, csrc/peer.cu: stage,
    signal, wait, add in rank order, epilogue); vectors that do not fit its slot go through ncclAllReduce;
  * MiniBatchStdDev's scalar in D: the rank-local mean of sqrt(var) is averaged across ranks, and so is its
    adjoint (twice differentiable: the R1/R2/GP double backward passes through it).  Rank r must hold, for each
    quarter q of the global batch, the r-th slice of that quarter (`shard_quarter_major`) so that the local
    `view(4,-1)` pairs the elements the global view pairs.

Losses are rank-local means; averaging the gradients over ranks makes them global means.
"""
import os

import torch
import torch.distributed as dist
from torch.autograd import Function

from . import ops

PEER = [None]             # gan_b200.peer.PeerExchange of the installed DataParallelStep (small exchanges over NVLink)
BUCKET_BYTES = 256 << 20  # NVSwitch: size buckets for launch latency / overlap, not for link count
CALLS = [0]               # collective calls issued by this process (bench.py: nccl_calls_per_step)


def init_from_env(backend=None):
    """torchrun env (RANK / WORLD_SIZE / LOCAL_RANK / MASTER_*) -> (rank, world, device)."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    use_cuda = torch.cuda.is_available()
    device = torch.device("cuda", local) if use_cuda else torch.device("cpu")
    if use_cuda:
        torch.cuda.set_device(device)
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        kw = {"device_id": device} if use_cuda else {}
        dist.init_process_group(backend or ("nccl" if use_cuda else "gloo"), rank=rank, world_size=world, **kw)
    return rank, world, device


class AllReduceMeanFn(Function):
    """mean over ranks; its adjoint is the same op (linear, self-adjoint up to the 1/N both ways)."""

    @staticmethod
    def forward(ctx, x, group):
        ctx.group = group
        px = PEER[0]
        if px is not None and x.is_cuda and x.dtype == torch.float32 and px.fits(x.numel()):
            return px.mean(x)          # one launch: stage, signal, wait, add in rank order, scale (csrc/peer.cu)
        y = x.detach().clone()
        CALLS[0] += 1
        dist.all_reduce(y, op=dist.ReduceOp.SUM, group=group)
        return ops.affine(y, 1.0 / dist.get_world_size(group), 0.0)

    @staticmethod
    def backward(ctx, g):
        return AllReduceMeanFn.apply(g.contiguous(), ctx.group), None


def shard_quarter_major(x, rank, world, groups=4):
    """rank's slice of a global batch such that the concatenation over ranks of local `view(groups,-1)` rows equals
    the global rows (MiniBatchStdDev, discriminators/discriminator_9_4.py:47-54)."""
    B = x.shape[0]
    assert B % (groups * world) == 0, "global batch must be divisible by 4 * world_size"
    q, l = B // groups, B // (groups * world)
    idx = torch.cat([torch.arange(g * q + rank * l, g * q + (rank + 1) * l) for g in range(groups)])
    return x[idx.to(x.device)]


class GradSync:
    """Sums the gradients of the live parameters over ranks in flat fp32 buckets, overlapped with the backward pass.

    One learning pass (the first optimizer step) runs the exchange after the backward and records, per parameter, how
    many times autograd accumulates into it during one step (a discriminator step calls backward two or three times) and
    in which order the parameters receive their LAST accumulation.  Buckets of BUCKET_BYTES are then laid out in that
    order.  From the second step on a post-accumulate hook counts the accumulations; the moment the last parameter of a
    bucket has its final gradient the bucket is launched on a side stream — ONE multi-tensor copy kernel gathers the
    gradients into the flat buffer (ganb200_multi_tensor_copy), ncclAllReduce sums it in place — while autograd keeps
    running the rest of the backward on the main stream.  `__call__` (the optimiser's pre-step hook) launches whatever is
    left, joins the side stream and points every p.grad at its slice of the reduced flat buffer, which is what the fused
    Adam then reads: no torch.cat, no per-tensor copy back.  Steps whose accumulation counts differ from the learnt ones
    (lazy-penalty recipes alternate two step bodies) are detected and exchanged after the backward instead.
    Under CUDA-graph capture the side-stream work is captured as a fork / join of the graph."""

    def __init__(self, params, group=None, opt=None, overlap=True):
        self.params = [p for p in params]
        self.group = group
        self.opt = opt
        self.overlap = overlap and os.environ.get("GANB200_DP_OVERLAP", "1") != "0"
        self.calls = 0
        self.expected = None           # {param: accumulations per step}
        self.buckets = None            # list of dicts: params, offsets, flat, pending, launched
        self.bucket_of = {}
        self.count = {}
        self.order = {}
        self.seq = 0
        self.stream = None
        self.hooks = [p.register_post_accumulate_grad_hook(self._on_grad) for p in self.params if p.requires_grad]
        self.overlapped_buckets = 0    # buckets launched from inside the backward during the current step
        self.last_overlapped = 0       # ... during the last finished step (tests / bench)

    def remove(self):
        for h in self.hooks:
            h.remove()
        self.hooks = []

    # ---- hook: runs inside autograd, right after a gradient has been accumulated into p.grad ----------------------
    def _on_grad(self, p):
        c = self.count.get(p, 0) + 1
        self.count[p] = c
        self.seq += 1
        self.order[p] = self.seq
        if self.buckets is None or not self.overlap:
            return
        b = self.bucket_of.get(p)
        if b is None:
            return
        if c == self.expected[p]:
            b["pending"] -= 1
            if b["pending"] == 0 and not b["launched"]:
                self._launch(b, overlapped=True)
        elif c > self.expected[p]:
            b["dirty"] = True          # more accumulations than learnt: whatever was sent is stale, redo after backward

    def _launch(self, b, overlapped):
        from . import kernels as KN
        dev = b["flat"].device
        cur = torch.cuda.current_stream(dev) if dev.type == "cuda" else None
        if cur is not None:
            if self.stream is None:
                self.stream = torch.cuda.Stream(device=dev)
            ev = torch.cuda.Event()
            ev.record(cur)
            self.stream.wait_event(ev)
        ctx = torch.cuda.stream(self.stream) if cur is not None else _null()
        with ctx:
            entries = [(b["flat"][o:o + p.numel()].view_as(p), p.grad if p.grad.is_contiguous() else p.grad.contiguous(),
                        None, None) for p, o in zip(b["params"], b["offsets"])]
            if dev.type == "cuda" and self.opt is not None:
                tab, blk, n_blocks = self.opt._table(("gather", id(self), b["index"]),
                                                     [(d, g, d, d) for d, g, _, _ in entries], dev)
                from .optim import CHUNK
                KN.backend().multi_tensor_copy(tab, blk, n_blocks, CHUNK, stream=self.stream)
            else:   # CPU / gloo tests
                for d, g, _, _ in entries:
                    d.copy_(g)
            CALLS[0] += 1
            b["work"] = dist.all_reduce(b["flat"], op=dist.ReduceOp.SUM, group=self.group, async_op=True)
        b["launched"] = True
        b["dirty"] = False
        if overlapped:
            self.overlapped_buckets += 1

    def _learn(self, live):
        """lay the buckets out in the order the parameters received their final gradient"""
        live = sorted(live, key=lambda p: self.order.get(p, 0))
        self.expected = {p: self.count.get(p, 0) for p in live}
        self.buckets, self.bucket_of = [], {}
        cur, nbytes = [], 0
        for p in live:
            cur.append(p)
            nbytes += p.numel() * 4
            if nbytes >= BUCKET_BYTES:
                self._new_bucket(cur)
                cur, nbytes = [], 0
        if cur:
            self._new_bucket(cur)

    def _new_bucket(self, params):
        offs, n = [], 0
        for p in params:
            offs.append(n)
            n += (p.numel() + 3) // 4 * 4       # 16-byte aligned slices: the copy and Adam kernels stay vectorised
        b = {"index": len(self.buckets), "params": list(params), "offsets": offs,
             "flat": torch.zeros(n, dtype=params[0].dtype, device=params[0].device), "pending": len(params),
             "launched": False, "dirty": False, "work": None}
        self.buckets.append(b)
        for p in params:
            self.bucket_of[p] = b

    def __call__(self):
        world = dist.get_world_size(self.group)
        if world == 1:
            return
        live = [p for p in self.params if p.grad is not None]
        if self.buckets is None or set(live) != set(self.expected):
            self._learn(live)          # first step (or the set of live parameters changed): exchange everything now
        for b in self.buckets:
            if b["dirty"] or not b["launched"] or any(self.count.get(p, 0) != self.expected[p] for p in b["params"]):
                if b["launched"] and b["work"] is not None:
                    b["work"].wait()
                self._launch(b, overlapped=False)
        for b in self.buckets:
            b["work"].wait()           # the current stream waits for NCCL's stream
            if self.stream is not None:
                torch.cuda.current_stream(b["flat"].device).wait_stream(self.stream)
            for p, o in zip(b["params"], b["offsets"]):
                p.grad = b["flat"][o:o + p.numel()].view_as(p)
            b["pending"], b["launched"], b["work"] = len(b["params"]), False, None
        self.count.clear()
        self.order.clear()
        self.seq = 0
        self.calls += 1
        self.last_overlapped, self.overlapped_buckets = self.overlapped_buckets, 0


class _null:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False


class DataParallelStep:
    """Installs the three exchanges on a Train object (train/*.py) and removes them again."""

    def __init__(self, trainer, group=None, sync_bn=True, peer=None):
        self.T = trainer
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.sync_bn = sync_bn
        # small exchanges over NVLink peer memory (gan_b200.peer) instead of NCCL: default on CUDA, GANB200_PEER=0 = off
        self.use_peer = (os.environ.get("GANB200_PEER", "1") != "0") if peer is None else bool(peer)
        self.peer = None

    def install(self):
        if self.world == 1:
            return self
        T, world, group = self.T, self.world, self.group
        if self.use_peer and torch.device(T.device).type == "cuda":
            from .peer import PeerExchange
            self.peer = PEER[0] = PeerExchange.connect(group, T.device)
        px = self.peer
        for opt, net in ((T.optimizer_G, T.generator), (T.optimizer_D, T.discriminator)):
            opt.grad_scale = 1.0 / world
            opt.pre_step = GradSync(net.parameters(), group, opt=opt, overlap=self.T.recipe.lazy == 1)
        ops._mbstd_reduce[0] = lambda s: AllReduceMeanFn.apply(s, group)
        if self.sync_bn:
            def stats(mean, var, count):
                # combine (mean, biased var) of equal-sized shards: var_g = E[var] + E[mean^2] - E[mean]^2
                if px is not None and mean.is_cuda and mean.dtype == torch.float32 and px.fits(mean.numel(), 2):
                    m, v = px.stats(mean, var)
                    return m, v, count * world
                pack = torch.stack([mean, var + mean * mean])
                CALLS[0] += 1
                dist.all_reduce(pack, op=dist.ReduceOp.SUM, group=group)
                pack /= world
                m = pack[0].contiguous()
                return m, (pack[1] - m * m).contiguous(), count * world

            def grads(s1, s2):
                if px is not None and s1.is_cuda and s1.dtype == torch.float32 and px.fits(s1.numel(), 2):
                    return px.grads(s1, s2)
                pack = torch.stack([s1, s2])
                CALLS[0] += 1
                dist.all_reduce(pack, op=dist.ReduceOp.SUM, group=group)
                return pack[0].contiguous(), pack[1].contiguous()

            ops._sync_bn_hooks["stats"], ops._sync_bn_hooks["grads"] = stats, grads
            ops._sync_bn_hooks["peer"] = px     # the one-launch BatchNorm kernels carry the exchange themselves
        return self

    def remove(self):
        if self.peer is not None:
            self.peer.check()
            self.peer.close()
            self.peer = PEER[0] = None
        ops._mbstd_reduce[0] = None
        ops._sync_bn_hooks["stats"] = ops._sync_bn_hooks["grads"] = ops._sync_bn_hooks["peer"] = None
        for opt in (self.T.optimizer_G, self.T.optimizer_D):
            opt.grad_scale = 1.0
            if isinstance(opt.pre_step, GradSync):
                opt.pre_step.remove()
            opt.pre_step = None
