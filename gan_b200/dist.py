"""Data-parallel training step: one process per GPU (torchrun), NCCL over NVLink for the exchanges.

The reference is single-device; batch-sharded DP is new.  What must be exchanged so that N ranks with
B/N images each compute the step the reference computes on B images:

  * parameter gradients of G and D: summed over ranks in flat fp32 buckets, the 1/N folded into the fused Adam
    kernel's grad_scale (`GradSync`, installed as the optimiser's pre-step hook);
  * BatchNorm batch statistics in G (`sync-BN`): per-channel (mean, biased var) forward and (sum dy, sum dy*xhat)
    backward are combined across ranks (hooks in gan_b200.ops);
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
    """Sums .grad of the live parameters over ranks in flat buckets.  Called right before optimizer.step()."""

    def __init__(self, params, group=None):
        self.params = list(params)
        self.group = group
        self.calls = 0

    def __call__(self):
        world = dist.get_world_size(self.group)
        if world == 1:
            return
        live = [p for p in self.params if p.grad is not None]
        bucket, nbytes = [], 0
        handles = []
        for p in live:
            bucket.append(p.grad)
            nbytes += p.grad.numel() * 4
            if nbytes >= BUCKET_BYTES:
                handles.append(self._launch(bucket))
                bucket, nbytes = [], 0
        if bucket:
            handles.append(self._launch(bucket))
        for flat, grads, work in handles:
            work.wait()
            off = 0
            for g in grads:
                n = g.numel()
                g.copy_(flat[off:off + n].view_as(g))
                off += n
        self.calls += 1

    def _launch(self, grads):
        flat = torch.cat([g.reshape(-1) for g in grads])
        CALLS[0] += 1
        work = dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.group, async_op=True)
        return flat, grads, work


class DataParallelStep:
    """Installs the three exchanges on a Train object (train/*.py) and removes them again."""

    def __init__(self, trainer, group=None, sync_bn=True):
        self.T = trainer
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.sync_bn = sync_bn

    def install(self):
        if self.world == 1:
            return self
        T, world, group = self.T, self.world, self.group
        for opt, net in ((T.optimizer_G, T.generator), (T.optimizer_D, T.discriminator)):
            opt.grad_scale = 1.0 / world
            opt.pre_step = GradSync(net.parameters(), group)
        ops._mbstd_reduce[0] = lambda s: AllReduceMeanFn.apply(s, group)
        if self.sync_bn:
            def stats(mean, var, count):
                # combine (mean, biased var) of equal-sized shards: var_g = E[var] + E[mean^2] - E[mean]^2
                pack = torch.stack([mean, var + mean * mean])
                CALLS[0] += 1
                dist.all_reduce(pack, op=dist.ReduceOp.SUM, group=group)
                pack /= world
                m = pack[0].contiguous()
                return m, (pack[1] - m * m).contiguous(), count * world

            def grads(s1, s2):
                pack = torch.stack([s1, s2])
                CALLS[0] += 1
                dist.all_reduce(pack, op=dist.ReduceOp.SUM, group=group)
                return pack[0].contiguous(), pack[1].contiguous()

            ops._sync_bn_hooks["stats"], ops._sync_bn_hooks["grads"] = stats, grads
        return self

    def remove(self):
        ops._mbstd_reduce[0] = None
        ops._sync_bn_hooks["stats"] = ops._sync_bn_hooks["grads"] = None
        for opt in (self.T.optimizer_G, self.T.optimizer_D):
            opt.grad_scale = 1.0
            opt.pre_step = None
