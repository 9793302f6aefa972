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
           "peer_calls": __import__("gan_b200.peer", fromlist=["CALLS"]).CALLS[0],
           "peer_fused": __import__("gan_b200.peer", fromlist=["FUSED"]).FUSED[0],
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
        # running stats: 0.1 * batch statistic; the batch mean is a cancelling fp32 sum taken in another order (two shards)
        for key in ("bn_mean", "bn_var"):
            assert torch.allclose(m[key], single[key], rtol=1e-4, atol=1e-5), (key, (m[key] - single[key]).abs().max())
        assert m["calls"] > 0
        assert m["peer_calls"] > 0, "the small exchanges must run over NVLink peer memory (gan_b200.peer), not NCCL"
        assert m["peer_fused"] > 0, "the one-launch BatchNorm kernels must carry the sync-BN exchange themselves"
        assert m["overlapped"][0] == (0, 0) and m["overlapped"][1][0] >= 1 and m["overlapped"][1][1] >= 1, m["overlapped"]
        for key in ("d_params", "g_params"):   # after two Adam updates (a sign flip of a ~0 gradient moves an element by 2 lr)
            for n, v in single[key].items():
                assert float((m[key][n] - v).norm()) <= 1e-4 * max(float(v.norm()), 1.0), (key, n)
    for i, v in enumerate(single["d_losses"]):
        avg = 0.5 * (multi[0]["d_losses"][i] + multi[1]["d_losses"][i])
        assert abs(avg - v) <= 1e-5 * max(1.0, abs(v)), (i, avg, v)
    assert abs(0.5 * (multi[0]["g_loss"] + multi[1]["g_loss"]) - single["g_loss"]) <= 1e-5


def _peer_rank(rank, world, out_dir, port):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    import torch.distributed as dist
    from gan_b200 import dist as gdist
    from gan_b200.peer import PeerExchange
    _, _, dev = gdist.init_from_env("nccl")
    px = PeerExchange(None, dev)        # no NCCL fall-back in the test: a failure to map the peers must fail it
    gen = torch.Generator().manual_seed(100 + rank)
    worst = {"stats": 0.0, "grads": 0.0, "mean": 0.0}

    def nccl_sum(t):
        y = t.clone()
        dist.all_reduce(y)
        return y

    sizes = [1, 7, 48, 519, 1000, 8192]
    for rep in range(40):               # 240 back-to-back exchanges of changing size: the slot / flag protocol under load
        for n in sizes:
            mean = torch.randn(n, generator=gen).to(dev)
            var = torch.rand(n, generator=gen).to(dev) + 0.1
            m, v = px.stats(mean, var)
            m_ref = nccl_sum(mean) / world
            v_ref = nccl_sum(var + mean * mean) / world - m_ref * m_ref
            worst["stats"] = max(worst["stats"], float((m - m_ref).abs().max()), float((v - v_ref).abs().max()))
            g1, g2 = px.grads(mean, var)
            worst["grads"] = max(worst["grads"], float((g1 - nccl_sum(mean)).abs().max()),
                                 float((g2 - nccl_sum(var)).abs().max()))
            worst["mean"] = max(worst["mean"], float((px.mean(var) - nccl_sum(var) / world).abs().max()))
    # captured in a CUDA graph: the sequence number lives on the device, replays keep counting
    x = torch.randn(300, generator=gen).to(dev)
    side = torch.cuda.Stream(device=dev)
    side.wait_stream(torch.cuda.current_stream(dev))
    with torch.cuda.stream(side):
        px.mean(x)
    torch.cuda.current_stream(dev).wait_stream(side)
    torch.cuda.synchronize(dev)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph, capture_error_mode="thread_local"):
        y = px.mean(px.sum(x))
    want = nccl_sum(nccl_sum(x)) / world
    graph_worst = 0.0
    for _ in range(5):
        graph.replay()
        graph_worst = max(graph_worst, float((y - want).abs().max()))
        px.mean(x)                      # eager exchanges between replays
    # all ranks hold the same bits (same order of addition everywhere)
    z = px.sum(torch.randn(4096, generator=gen).to(dev))
    gathered = [torch.empty_like(z) for _ in range(world)]
    dist.all_gather(gathered, z)
    same_bits = all(torch.equal(gathered[0], t) for t in gathered)
    px.check()
    px.close()
    torch.save({"worst": worst, "graph_worst": graph_worst, "same_bits": same_bits}, os.path.join(out_dir, f"peer_r{rank}.pt"))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.timeout(600)
def test_peer_exchange_equals_nccl_all_reduce(tmp_path):
    """gan_b200.peer (one single-CTA launch over cudaIpc peer mappings) against ncclAllReduce + torch arithmetic on the
    same vectors: BatchNorm statistics combine, two-vector sum, mean; under load, inside a replayed CUDA graph, and bit
    for bit the same on every rank."""
    n_gpu = torch.cuda.device_count()
    if n_gpu < 2:
        pytest.skip("needs two GPUs (gpurun --gpus 2)")
    world = min(n_gpu, 8)
    ctx = mp.get_context("spawn")
    procs = [ctx.Process(target=_peer_rank, args=(r, world, str(tmp_path), 29651)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=300)
        assert p.exitcode == 0, f"rank process exited with {p.exitcode}"
    for r in range(world):
        res = torch.load(os.path.join(tmp_path, f"peer_r{r}.pt"), weights_only=False)
        # fp32: the same W addends in another order (NCCL's ring / tree vs rank order) and an fma in the variance
        assert res["worst"]["grads"] <= 1e-5 and res["worst"]["mean"] <= 1e-6 and res["worst"]["stats"] <= 1e-5, res
        assert res["graph_worst"] <= 1e-4, res
        assert res["same_bits"], "every rank must end with the same bits"
