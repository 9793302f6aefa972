"""CPU, gloo, world_size 2: the data-parallel exchanges of gan_b200.dist make 2 ranks x B/2 images compute the
gradients and the Adam update a single rank computes on B images: gradient all-reduce (1/N folded into
the fused Adam grad_scale), sync-BN statistics forward and backward (in G, and in the critic — there also through the
R1 / R2 / GP double backward), the global MiniBatchStdDev scalar and its adjoint (including through the double
backward), quarter-major batch sharding."""
import os
import sys

import pytest
import torch
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ITERS = 2


def _build():
    """a miniature G (linear+BN1d+PReLU -> convT -> BN2d) and D (conv+PReLU -> BN2d -> mbstd -> batch-std -> conv -> linear) on our layers"""
    import torch.nn as nn
    from gan_b200 import layers as L
    from gan_b200 import ops

    class EqW(L.EqWeightBase):
        PARAM, GAIN = "weight", "c"

    class EqConv(L.EqConv2dBase):
        WEIGHT = EqW

        def __init__(self, a, b, k, p, s=1):
            super().__init__()
            self._build(a, b, k, p, s)

    class EqLin(L.EqLinearBase):
        WEIGHT = EqW

        def __init__(self, a, b):
            super().__init__()
            self._build(a, b)

    class G(nn.Module):
        def __init__(self):
            super().__init__()
            self.fc, self.bn1, self.act = EqLin(8, 8), L.BatchNorm1d(8), L.PReLU(8)
            self.ct = L.ConvTranspose2d(8, 3, 4, 1, 0)
            self.bn2 = L.BatchNorm2d(3)

        def forward(self, z):
            w = self.act(self.bn1(self.fc(torch.squeeze(z))))
            return self.bn2(self.ct(w.reshape(w.shape[0], 8, 1, 1)))

    class D(nn.Module):
        def __init__(self):
            super().__init__()
            self.c1, self.a1 = EqConv(3, 6, 3, 1), L.PReLU(6)
            self.bn = L.BatchNorm2d(6)  # a BatchNorm critic: sync statistics, also through the penalties' double backward
            self.c2, self.a2 = EqConv(8, 4, 3, 1, 2), L.PReLU(4)
            self.fc = EqLin(4 * 2 * 2, 1)

        def forward(self, x):
            t = self.bn(self.a1(self.c1(x)))
            t = ops.minibatch_stddev(t, 4, 1e-8)
            t = ops.batch_stddev(t)  # the ProGAN per-position batch statistic (discriminator_3_wgangp_progan.py:7-16)
            t = self.a2(self.c2(t))
            return self.fc(t.reshape(t.shape[0], -1))

    return G(), D()


def _step(rank, world, B, out_dir, port):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import torch.distributed as dist
    from gan_b200 import kernels as KN
    from gan_b200 import dist as gdist
    from gan_b200 import rng
    from oracle.kernels_ref import RefBackend
    from train import wgangpR2
    torch.set_default_dtype(torch.float64)
    torch.set_num_threads(1)
    KN.set_backend(RefBackend())
    rng.noise_mode("torch")
    if world > 1:
        dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.manual_seed(0)
    G, D = _build()
    G, D = G.double(), D.double()
    gen = torch.Generator().manual_seed(5)
    images = torch.randn(B, 3, 4, 4, generator=gen, dtype=torch.float64)
    zs = torch.randn(2 * ITERS, B, 8, 1, 1, generator=gen, dtype=torch.float64)
    eps = torch.rand(ITERS, B, generator=gen, dtype=torch.float64)
    if world > 1:
        images = gdist.shard_quarter_major(images, rank, world)
        eps = torch.stack([gdist.shard_quarter_major(eps[i], rank, world) for i in range(ITERS)])
        zs = torch.stack([gdist.shard_quarter_major(zs[i], rank, world) for i in range(2 * ITERS)])
    Bl = images.shape[0]
    T = wgangpR2.Train([(images, torch.zeros(Bl))], torch.device("cpu"), 1, 8, G, "g", D, "d")
    dp = gdist.DataParallelStep(T).install() if world > 1 else None
    # feed the same z / eps to both layouts: patch the draws of the step
    draws = iter([t for i in range(ITERS) for t in (zs[2 * i], eps[i], zs[2 * i + 1])])
    orig_randn, orig_rand = rng.randn, rng.rand
    rng.randn = lambda *s, device=None: next(draws)
    rng.rand = lambda *s, device=None: next(draws)
    overlapped = []
    try:
        # iteration 1 is the learning pass of the bucketed all-reduce (exchange after the backward); from iteration 2 on
        # the buckets are launched from inside the backward (gan_b200.dist.GradSync)
        for _ in range(ITERS):
            d_out = T.discriminator_trainstep(images.clone().requires_grad_(), Bl)
            d_grads = {n: p.grad.clone() for n, p in D.named_parameters() if p.grad is not None}
            _, g_loss = T.generator_trainstep(Bl)
            g_grads = {n: p.grad.clone() for n, p in G.named_parameters() if p.grad is not None}
            if dp is not None:
                overlapped.append((T.optimizer_D.pre_step.last_overlapped, T.optimizer_G.pre_step.last_overlapped))
    finally:
        rng.randn, rng.rand = orig_randn, orig_rand
    res = {"d_losses": [float(v) for v in d_out], "g_loss": float(g_loss),
           "d_params": {n: p.detach().clone() for n, p in D.named_parameters()},
           "g_params": {n: p.detach().clone() for n, p in G.named_parameters()},
           "bn_mean": G.bn2.running_mean.clone(), "bn_var": G.bn2.running_var.clone(),
           "d_grads": d_grads, "g_grads": g_grads, "overlapped": overlapped}
    if dp is not None:
        dp.remove()
        dist.barrier()
        dist.destroy_process_group()
    torch.save(res, os.path.join(out_dir, f"w{world}_r{rank}.pt"))


def _run(world, B, port, out_dir):
    ctx = mp.get_context("spawn")
    procs = [ctx.Process(target=_step, args=(r, world, B, str(out_dir), port)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=300)
        assert p.exitcode == 0, f"rank process exited with {p.exitcode}"
    return {r: torch.load(os.path.join(out_dir, f"w{world}_r{r}.pt"), weights_only=False) for r in range(world)}


@pytest.mark.timeout(600)
def test_two_rank_data_parallel_equals_single_rank_big_batch(tmp_path):
    B = 16
    single = _run(1, B, 29631, tmp_path)[0]
    multi = _run(2, B, 29632, tmp_path)
    for r in (0, 1):
        m = multi[r]
        # parameters after both Adam updates: identical on every rank and equal to the single-rank run
        for key in ("d_params", "g_params"):
            for n, v in single[key].items():
                assert torch.allclose(m[key][n], v, rtol=1e-9, atol=1e-12), (key, n, (m[key][n] - v).abs().max())
        assert m["overlapped"][0] == (0, 0) and m["overlapped"][1][0] >= 1 and m["overlapped"][1][1] >= 1, m["overlapped"]
        assert torch.allclose(m["bn_mean"], single["bn_mean"], atol=1e-12)
        assert torch.allclose(m["bn_var"], single["bn_var"], atol=1e-12)
    # rank-local losses are means over the shard: their average is the global mean (GP/R1/R2 included)
    for i, v in enumerate(single["d_losses"]):
        avg = 0.5 * (multi[0]["d_losses"][i] + multi[1]["d_losses"][i])
        assert abs(avg - v) < 1e-9 * max(1.0, abs(v)), (i, avg, v)
    assert abs(0.5 * (multi[0]["g_loss"] + multi[1]["g_loss"]) - single["g_loss"]) < 1e-9


def test_quarter_major_sharding_pairs_the_same_elements():
    from gan_b200.dist import shard_quarter_major
    x = torch.arange(16.0).reshape(16, 1)
    parts = [shard_quarter_major(x, r, 2) for r in range(2)]
    glob = x.view(4, -1)
    loc = torch.cat([p.view(4, -1) for p in parts], dim=1)
    assert torch.equal(loc.sort(dim=1).values, glob.sort(dim=1).values)
