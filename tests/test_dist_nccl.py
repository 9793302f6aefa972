"""GPU, NCCL, world_size 2 (needs two B200s: `gpurun --gpus 2`; skipped on a one-GPU box): the data-parallel exchanges
of gan_b200.dist on the CUDA kernels make 2 ranks x B/2 images compute what one rank computes on B images — gradient
bucket all-reduce, sync-BN forward / backward (also through the R1 / R2 / GP double backward), the global
MiniBatchStdDev scalar and its adjoint, the ProGAN batch-std plane — in fp32 (engine 1) to rel-L2 <= 1e-5.
The same miniature G / D as tests/test_dist_gloo.py (which proves the algebra in fp64 on CPU)."""
import os
import sys

import pytest
import torch
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

pytestmark = pytest.mark.gpu
ITERS = 2


def _step(rank, world, B, out_dir, port):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank), GANB200_PRECISION="fp32")
    import torch.distributed as dist
    from gan_b200 import dist as gdist
    from gan_b200 import rng
    from test_dist_gloo import _build
    from train import wgangpR2
    rng.noise_mode("torch")
    _, _, dev = gdist.init_from_env("nccl" if world > 1 else None)
    torch.manual_seed(0)
    G, D = _build()
    G, D = G.to(dev), D.to(dev)
    gen = torch.Generator().manual_seed(5)
    images = torch.randn(B, 3, 4, 4, generator=gen)
    zs = torch.randn(2 * ITERS, B, 8, 1, 1, generator=gen)
    eps = torch.rand(ITERS, B, generator=gen)
    if world > 1:
        images = gdist.shard_quarter_major(images, rank, world)
        eps = torch.stack([gdist.shard_quarter_major(eps[i], rank, world) for i in range(ITERS)])
        zs = torch.stack([gdist.shard_quarter_major(zs[i], rank, world) for i in range(2 * ITERS)])
    images, zs, eps = images.to(dev), zs.to(dev), eps.to(dev)
    Bl = images.shape[0]
    T = wgangpR2.Train([(images, torch.zeros(Bl))], dev, 1, 8, G, "g", D, "d")
    dp = gdist.DataParallelStep(T).install() if world > 1 else None
    draws = iter([t.contiguous() for i in range(ITERS) for t in (zs[2 * i], eps[i], zs[2 * i + 1])])
    orig_randn, orig_rand = rng.randn, rng.rand
    rng.randn = lambda *s, device=None: next(draws)
    rng.rand = lambda *s, device=None: next(draws)
    overlapped = []
    try:
        # iteration 1: learning pass of the bucketed all-reduce; iteration 2: buckets launched from inside the backward
        for _ in range(ITERS):
            d_out = T.discriminator_trainstep(images.clone().requires_grad_(), Bl)
            d_grads = {n: p.grad.detach().cpu() / world for n, p in D.named_parameters() if p.grad is not None}
            _, g_loss = T.generator_trainstep(Bl)
            g_grads = {n: p.grad.detach().cpu() / world for n, p in G.named_parameters() if p.grad is not None}
            if dp is not None:
                overlapped.append((T.optimizer_D.pre_step.last_overlapped, T.optimizer_G.pre_step.last_overlapped))
    finally:
        rng.randn, rng.rand = orig_randn, orig_rand
    torch.cuda.synchronize()
    res = {"d_losses": [float(v) for v in d_out], "g_loss": float(g_loss),
           "bn_mean": G.bn2.running_mean.cpu(), "bn_var": G.bn2.running_var.cpu(),
           "d_grads": d_grads, "g_grads": g_grads, "calls": gdist.CALLS[0], "overlapped": overlapped,
           "d_params": {n: p.detach().cpu() for n, p in D.named_parameters()},
           "g_params": {n: p.detach().cpu() for n, p in G.named_parameters()}}
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
        p.join(timeout=600)
        assert p.exitcode == 0, f"rank process exited with {p.exitcode}"
    return {r: torch.load(os.path.join(out_dir, f"w{world}_r{r}.pt"), weights_only=False) for r in range(world)}


@pytest.mark.timeout(1200)
def test_two_rank_nccl_step_equals_single_rank_big_batch(tmp_path):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (gpurun --gpus 2)")
    B = 16
    single = _run(1, B, 29641, tmp_path)[0]
    multi = _run(2, B, 29642, tmp_path)
    tol = 1e-5  # fp32 elementwise / fp32 CUDA-core convolution paths (north_star)
    for r in (0, 1):
        m = multi[r]
        for key in ("d_grads", "g_grads"):
            for n, v in single[key].items():
                if float(v.norm()) < 1e-6:
                    continue  # analytically zero gradients (a bias feeding a BatchNorm): rounding noise only
                rel = float((m[key][n] - v).norm() / v.norm())
                assert rel <= tol, (key, n, rel)
        assert torch.allclose(m["bn_mean"], single["bn_mean"], rtol=1e-4, atol=1e-6)   # running stats: 0.1 * batch mean, fp32 sums in another order
        assert torch.allclose(m["bn_var"], single["bn_var"], rtol=1e-4, atol=1e-6)
        assert m["calls"] > 0
        assert m["overlapped"][0] == (0, 0) and m["overlapped"][1][0] >= 1 and m["overlapped"][1][1] >= 1, m["overlapped"]
        for key in ("d_params", "g_params"):   # after two Adam updates (a sign flip of a ~0 gradient moves an element by 2 lr)
            for n, v in single[key].items():
                assert float((m[key][n] - v).norm()) <= 1e-4 * max(float(v.norm()), 1.0), (key, n)
    for i, v in enumerate(single["d_losses"]):
        avg = 0.5 * (multi[0]["d_losses"][i] + multi[1]["d_losses"][i])
        assert abs(avg - v) <= 1e-5 * max(1.0, abs(v)), (i, avg, v)
    assert abs(0.5 * (multi[0]["g_loss"] + multi[1]["g_loss"]) - single["g_loss"]) <= 1e-5
