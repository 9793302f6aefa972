"""bench.py — train images/s of one GAN training iteration (discriminator_trainstep + generator_trainstep)
for the final pair generator_13_5 + discriminator_9_4 (BASELINE.json configs[1]: ganR2 penalty structure,
batch 256 per GPU), through the drop-in train/*.py API on libganb200 kernels.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--batch B] [--loop ganR2|wgangpR2] [--impl reference]

Prints ONE JSON line (rank 0).  `value` = images/s with the batch already resident in HBM; `e2e` = the same
through the public step API with the batch in pinned HOST memory (H2D copy of the images and a D2H read of the
losses inside the timed region).  `roofline` is measured live with CUDA events around every launch of the
dominant kernel class (the convolution engine) during the timed steps.  `cpu_baseline` / `--impl reference`
time the CPU restatement (the same step on oracle/kernels_ref.py — the reference's own PyTorch CPU ops) on the
host cores; the unmodified reference tree is not present on the GPU box (it is Python under /root/reference and
cannot travel), so kind = "port".
"""
import argparse
import json
import os
import random
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np
import torch

NZ = 256
G_F = 101.995e9   # generator_13_5 forward FLOPs / image (SURVEY §8(d), 2*MAC conv+matmul)
D_F = 11.922e9    # discriminator_9_4 forward FLOPs / image
FLOPS_PER_IMAGE = {"ganR2": 4 * G_F + 19.16 * D_F, "wgangpR2": 4 * G_F + 25.24 * D_F}


# DRAM bytes (dram__bytes_read.sum + dram__bytes_write.sum) of ONE launch of the engine's dominant kernel shape, from
# the ncu --set full capture summarised in profiles/r01_conv_ncu_full_end.txt (not measured by this script).
NCU_TRAFFIC = {"bytes": 577.24e6, "of": "tc_conv_halo_kernel<128>, fprop 96->96 5x5 @64x64, N=256 (ncu --set full, "
                                        "profiles/r01_conv_ncu_full_end.txt)"}


def seed_all(s):
    torch.manual_seed(s)
    np.random.seed(s)
    random.seed(s)


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm_gbs": d["hbm_gbs"], "tflops": d.get("bf16_tflops_sustained", d["bf16_tflops"]), "src": "measured"}
    return {"hbm_gbs": 6650.0, "tflops": 1400.0, "src": "fallback"}


def build_trainer(device, loop, batch, rank=0):
    """SURVEY §8(d) recipe: seeds 0 before construction; for ganR2 the weight-only patch of D's last linear that
    keeps its logits inside (0,1) so that BCELoss is valid (F2); wgangpR2 runs un-patched."""
    from discriminators import discriminator_9_4
    from generators import generator_13_5
    import importlib
    seed_all(0)
    G = generator_13_5.Generator(NZ).to(device)
    D = discriminator_9_4.Discriminator().to(device)
    if loop == "ganR2":
        with torch.no_grad():
            D.fc[2].weight.weight.mul_(1e-3)
            D.fc[2].bias.fill_(0.5)
    mod = importlib.import_module(f"train.{loop}")
    T = mod.Train([(torch.zeros(batch, 3, 64, 64), torch.zeros(batch))], device, 1, NZ, G, "G13_5", D, "D9_4")
    seed_all(1 + rank)
    return T


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.stop_flag, self.index = [], False, index
        self.th = threading.Thread(target=self.run, daemon=True)

    def run(self):
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                      "-i", str(self.index)], capture_output=True, text=True, timeout=5).stdout
                self.rows.append([c.strip() for c in out.strip().split(",")])
            except Exception:
                pass
            time.sleep(0.2)

    def __enter__(self):
        self.th.start()
        return self

    def __exit__(self, *a):
        self.stop_flag = True
        self.th.join(timeout=6)

    def summary(self):
        sm = [int(r[0]) for r in self.rows if len(r) >= 6 and r[0].isdigit()]
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unavailable"]}
        mx = [int(r[1]) for r in self.rows if len(r) >= 6 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) >= 6 and r[2 + i].lower().startswith("active")
                                                         for r in self.rows)]
        return {"sm_mhz": int(np.median(sm)), "sm_max_mhz": max(mx) if mx else None, "reasons": reasons}


def one_iteration(T, images, batch):
    d_out = T.discriminator_trainstep(images, batch)
    gen_imgs, g_loss = T.generator_trainstep(batch)
    return d_out, g_loss


def cpu_port_rate(loop, batch, seconds_budget, steps=None, warmup=1, threads=None):
    """The same training iteration on the CPU restatement (oracle/kernels_ref.py: stock PyTorch CPU ops, fp32 — what
    the reference executes) on the host cores.  Returns (images/s, steps timed, seconds per step, threads)."""
    from gan_b200 import kernels as KN
    from oracle.kernels_ref import RefBackend
    threads = threads or (os.cpu_count() or 1)
    torch.set_num_threads(threads)
    prev = KN.set_backend(RefBackend())
    try:
        cpu = torch.device("cpu")
        T = build_trainer(cpu, loop, batch)
        images = torch.randn(batch, 3, 64, 64).requires_grad_()
        t0 = time.perf_counter()
        for _ in range(max(warmup, 1)):
            one_iteration(T, images, batch)
        t_w = (time.perf_counter() - t0) / max(warmup, 1)
        n = steps if steps is not None else max(1, int(seconds_budget / max(t_w, 1e-3)))
        if steps is not None and seconds_budget and n * t_w > seconds_budget:
            n = max(1, int(seconds_budget / t_w))
        t0 = time.perf_counter()
        for _ in range(n):
            one_iteration(T, images, batch)
        dt = (time.perf_counter() - t0) / n
    finally:
        KN.set_backend(prev)
    return batch / dt, n, dt, threads


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cpu_batch = args.cpu_batch
    rate, n, dt, threads = cpu_port_rate(args.loop, cpu_batch, args.cpu_seconds, steps=args.steps, warmup=args.warmup)
    sample = (f"{n} full iterations (D step + G step) at batch {cpu_batch} of the same G13_5+D9_4 {args.loop} "
              f"workload, fp32, torch CPU ops, {threads} threads")
    line = {"impl": "reference", "metric": "train images/sec (G13_5+D9_4, 64x64)", "value": rate, "unit": "images/s",
            "n_gpus": args.gpus, "steps": n, "warmup": args.warmup, "ms_per_step": dt * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"generator_13_5+discriminator_9_4 train/{args.loop}.py, batch {args.batch}/GPU "
                                   f"(CPU sample at batch {cpu_batch})", "loop": args.loop},
            "cpu_baseline": {"value": rate, "unit": "images/s", "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": rate, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def replay_census(census, device, reps=3, dump=None):
    """Times every distinct convolution-engine call of one iteration (shape, op, engine) in isolation with CUDA events
    on the launching stream — a CUDA graph of `reps` back-to-back launches, replayed once warm and once timed — and
    weights it by its call count.
    -> {engine: (flops, seconds, calls)} for ONE iteration."""
    from gan_b200 import kernels as KN
    k = KN.backend()
    out = {}
    rows = []
    for key, count in sorted(census.items(), key=lambda kv: -kv[1]):
        op, N, C, H, W, K, R, S, stride, pad, pm, has_i, has_o, has_b, eng = key
        P, Q = KN.out_hw(H, W, R, S, stride, pad)
        e = {"simt": 1, "tc": 2, "linear": 0}[eng]
        x = torch.randn(N, C, H, W, device=device)
        w = torch.randn(K, C, R, S, device=device)
        dy = torch.randn(N, K, P, Q, device=device)
        isc = torch.rand(N, C, device=device) if has_i else None
        osc = torch.rand(N, K, device=device) if has_o else None
        b = torch.randn(K, device=device) if has_b else None
        if op == KN.FPROP:
            f = lambda: k.conv_fprop(x, w, b, stride, pad, pm, 1.0, isc, osc, engine=e)
        elif op == KN.DGRAD:
            f = lambda: k.conv_dgrad(dy, w, (H, W), stride, pad, pm, 1.0, isc, osc, engine=e)
        else:
            f = lambda: k.conv_wgrad(x, dy, (R, S), stride, pad, pm, 1.0, isc, osc, engine=e)
        f()
        torch.cuda.synchronize()
        # `reps` launches captured in one CUDA graph: the timed replay holds device time only (issued from Python the
        # 4-5 us calls of the linear layers were host-bound and understated the engine)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for _ in range(reps):
                f()
        g.replay()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.replay()
        e1.record()
        e1.synchronize()
        sec = e0.elapsed_time(e1) * 1e-3 / reps
        del g
        fl = 2.0 * N * K * P * Q * C * R * S
        a = out.get(eng, (0.0, 0.0, 0))
        out[eng] = (a[0] + fl * count, a[1] + sec * count, a[2] + count)
        rows.append({"op": ["fprop", "dgrad", "wgrad"][op], "N": N, "C": C, "H": H, "W": W, "K": K, "R": R,
                     "stride": stride, "pad": pad, "pad_mode": pm, "engine": eng, "calls": count, "ms_each": sec * 1e3,
                     "ms_total": sec * count * 1e3, "tflops": fl / sec / 1e12})
    rows.sort(key=lambda r: -r["ms_total"])
    if dump:
        os.makedirs(os.path.dirname(dump), exist_ok=True)
        json.dump(rows, open(dump, "w"), indent=0)
    replay_census.top = [r for r in rows if r["engine"] == "tc"][:3]
    return out


def run_native(args):
    from gan_b200 import dist as gdist
    from gan_b200 import kernels as KN
    import torch.distributed as tdist
    rank, world, device = gdist.init_from_env()
    assert device.type == "cuda", "bench.py's own arm needs a B200 (there is no CPU fallback)"
    assert world == args.gpus, f"--gpus {args.gpus} but WORLD_SIZE={world}"
    B = args.batch
    T = build_trainer(device, args.loop, B, rank)
    dp = gdist.DataParallelStep(T).install()
    host_images = torch.randn(B, 3, 64, 64).pin_memory()
    dev_images = host_images.to(device).requires_grad_()

    def barrier():
        if world > 1:
            tdist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=device)
        if world > 1:
            tdist.all_reduce(ms, op=tdist.ReduceOp.MAX)
        return float(ms) / steps

    S = T
    if args.graphs:
        from gan_b200.graphed import GraphedStep
        KN.LAUNCH_COUNT[0] = 0
        KN.CONV_CENSUS = {}
        S = GraphedStep(T, B, warmup=max(args.warmup, 2), images=dev_images)
        launches_per_step = KN.LAUNCH_COUNT[0] // (max(args.warmup, 2) + 1)
        census = {k_: v // (max(args.warmup, 2) + 1) for k_, v in KN.CONV_CENSUS.items()}
        KN.CONV_CENSUS = None
        dev_images = S.images

    def step_resident():
        one_iteration(S, dev_images, B)

    d2h = [0]

    def step_e2e():
        imgs = host_images if args.graphs else host_images.to(device, non_blocking=True).requires_grad_()
        d_out, g_loss = one_iteration(S, imgs, B)
        losses = torch.stack([v.reshape(()) for v in d_out] + [g_loss.reshape(())]).cpu()
        d2h[0] = losses.numel() * 4

    for _ in range(args.warmup):
        step_resident()
    KN.LAUNCH_COUNT[0] = 0
    KN.CONV_TIMER = None if args.graphs else KN.ConvTimer()
    if args.profile_range:
        torch.cuda.profiler.start()   # ncu --profile-from-start off: only the timed steps are profiled
    with ClockSampler(device.index or 0) as clk:
        ms = timed(step_resident, args.steps)
    if args.profile_range:
        torch.cuda.profiler.stop()
    launches = launches_per_step * args.steps if args.graphs else KN.LAUNCH_COUNT[0]
    conv = KN.CONV_TIMER.totals() if KN.CONV_TIMER is not None else {}
    KN.CONV_TIMER = None
    ms_e2e = timed(step_e2e, args.steps)
    if args.graphs:
        conv = {k_: (v[0] * args.steps, v[1] * args.steps, v[2] * args.steps)
                for k_, v in replay_census(census, device, dump=args.census_dump).items()}
    mem_gb = torch.cuda.max_memory_allocated(device) / 2**30
    dp.remove()

    torch.cuda.synchronize()
    if rank != 0:
        # no collective and no communicator teardown after the timed region: the captured CUDA graphs still reference
        # the NCCL communicator, and a rank must never wait for rank 0's CPU-side reporting.  Hard exit, status 0.
        sys.stdout.flush()
        os._exit(0)
    pk = peaks()
    eng = "tc" if "tc" in conv and conv["tc"][1] >= conv.get("simt", (0, 0, 0))[1] * 0.25 else "simt"
    fl, sec, calls = conv.get(eng, (0.0, 1.0, 0))
    achieved = fl / sec / 1e12 if sec > 0 else 0.0
    conv_share = sum(v[1] for v in conv.values()) / (ms * 1e-3 * args.steps)
    roofline = {"bound": "tensor", "kernel": "conv_tc (tcgen05 implicit GEMM)" if eng == "tc" else
                "conv_simt (fp32 CUDA-core implicit GEMM)", "achieved": achieved, "peak": pk["tflops"],
                "unit": "TFLOP/s", "frac": achieved / pk["tflops"], "traffic": NCU_TRAFFIC["bytes"],
                "traffic_of": NCU_TRAFFIC["of"], "top_shapes": getattr(replay_census, "top", []),
                "peak_source": pk["src"],
                "launches": calls, "conv_share_of_step": conv_share,
                "by_engine": {k: {"tflops": v[0] / v[1] / 1e12 if v[1] > 0 else 0.0, "seconds": v[1], "calls": v[2]}
                              for k, v in conv.items()}}
    line = {"metric": "train images/sec (G13_5+D9_4, 64x64)", "value": B * world / (ms * 1e-3), "unit": "images/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16" if KN.default_engine() != 1 else "f32", "data": "synthetic",
            "config": {"workload": f"generator_13_5+discriminator_9_4 train/{args.loop}.py, batch {B}/GPU, "
                                   f"global batch {B * world}, nz {NZ}", "loop": args.loop, "batch_per_gpu": B,
                       "l2": "per-step working set (activations + 2 GB of parameters) exceeds the 126 MB L2",
                       "parallelism": f"dp{world}", "cuda_graphs": bool(args.graphs),
                       "g_step_critic_wgrad": "computed (as the reference)" if os.environ.get("GANB200_G_STEP_D_GRADS") == "1"
                       else "skipped: unread by the reference (SURVEY B9); parameters and losses identical"},
            "model_tflops": FLOPS_PER_IMAGE[args.loop] * B * world / (ms * 1e-3) / 1e12 / world,
            "clocks": clk.summary(), "gpu_launches": launches // max(args.steps, 1), "peak_mem_gib": mem_gb,
            "e2e": {"value": B * world / (ms_e2e * 1e-3), "unit": "images/s",
                    "h2d_bytes_per_step": host_images.numel() * 4, "d2h_bytes_per_step": d2h[0]},
            "roofline": roofline}
    if world == 1 and not args.no_cpu:
        rate, n, dt, threads = cpu_port_rate(args.loop, args.cpu_batch, args.cpu_seconds)
        line["cpu_baseline"] = {"value": rate, "unit": "images/s", "cores": threads, "kind": "port",
                                "sample": f"{n} full iterations at batch {args.cpu_batch} on torch CPU ops (fp32)"}
    print(json.dumps(line), flush=True)
    if world > 1:
        sys.stdout.flush()
        os._exit(0)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--batch", type=int, default=int(os.environ.get("GANB200_BENCH_BATCH", "256")))
    ap.add_argument("--loop", default="ganR2", choices=["ganR2", "wgangpR2"])
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--cpu-batch", type=int, default=4)
    ap.add_argument("--cpu-seconds", type=float, default=25.0)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--profile-range", action="store_true", help="cudaProfilerStart/Stop around the timed steps")
    ap.add_argument("--census-dump", default=os.path.join(ROOT, "gpurun_out", "conv_census.json"))
    ap.add_argument("--graphs", type=int, default=int(os.environ.get("GANB200_GRAPHS", "1")),
                    help="1: replay the two step bodies as CUDA graphs (default); 0: eager launches")
    args = ap.parse_args()
    if args.impl == "reference":
        args.cpu_seconds = 240.0
        run_reference(args)
    else:
        run_native(args)


if __name__ == "__main__":
    main()
