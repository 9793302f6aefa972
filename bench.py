"""bench.py — train images/s of one GAN training iteration (discriminator_trainstep + generator_trainstep)
for the final pair generator_13_5 + discriminator_9_4 (BASELINE.json configs[1]: ganR2 penalty structure,
batch 256 per GPU), through the drop-in train/*.py API on libganb200 kernels.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--batch B] [--loop ganR2|wgangpR2] [--impl reference]

Prints ONE JSON line (rank 0).  `value` = images/s with the batch already resident in HBM; `e2e` = the same
through the public step API with the batch in pinned HOST memory (H2D copy of the images and a D2H read of the
losses inside the timed region).  `roofline` is measured live with CUDA events around every launch of the
dominant kernel class (the convolution engine) during the timed steps.  `cpu_baseline` / `--impl reference`
time the UNMODIFIED reference (its own train/<loop>.py Train.*_trainstep on torch CPU ops, loaded from /root/reference
or from oracle/_ref, the byte-for-byte copy oracle/build_ref.py makes at build time: kind = "reference") on the host
cores; only if neither tree exists does it fall back to the CPU restatement (oracle/kernels_ref.py: kind = "port").

BCELoss validity: the ganR2 loop needs every D logit in [0,1]; the weight-only patch of D's last linear
keeps that true for the first few Adam steps only.  Both arms therefore restore the patched initial state (parameters,
buffers, optimiser state) OUTSIDE the timed intervals — the CPU arm before every iteration, the GPU arm before every
`--restore-every` steps — and the GPU arm reads the BCE range flag after every timed segment: the JSON line carries
"bce_in_range" and the last losses, so the step that was timed is the step the reference defines.
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


def ncu_traffic():
    """DRAM bytes (dram__bytes_read.sum + dram__bytes_write.sum) per launch of the engine's dominant kernel, from the
    newest profiles/r*_conv_traffic.json (written by tools/ncu_traffic.py from one `ncu --set full` capture).  Not
    measurable inside this script (ncu replays kernels): null when no capture has been committed."""
    import glob
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_conv_traffic.json")))
    if not files:
        return None, None
    d = json.load(open(files[-1]))
    return d.get("dram_bytes_per_launch"), f"{d.get('kernel')} {d.get('shape')} ({os.path.relpath(files[-1], ROOT)})"


def physical_cores():
    try:
        import psutil
        n = psutil.cpu_count(logical=False)
        if n:
            return int(n)
    except Exception:
        pass
    return os.cpu_count() or 1


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


class HostSnapshot:
    """state_dict copies of both networks and both optimisers (CPU arm); restore() puts the trainer back to iteration 0."""

    def __init__(self, T):
        import copy
        self.T = T
        self.nets = [copy.deepcopy(m.state_dict()) for m in (T.generator, T.discriminator)]
        self.opts = [copy.deepcopy(o.state_dict()) for o in (T.optimizer_G, T.optimizer_D)]

    def restore(self):
        import copy
        T = self.T
        for m, sd in zip((T.generator, T.discriminator), self.nets):
            m.load_state_dict(sd)
        for o, sd in zip((T.optimizer_G, T.optimizer_D), self.opts):
            o.load_state_dict(copy.deepcopy(sd))


def build_reference_trainer(loop, batch):
    """The unmodified reference: generators/generator_13_5.py + discriminators/discriminator_9_4.py under
    train/<loop>.py, same seeds and the same weight-only patch as build_trainer."""
    from oracle import reference as ref
    cpu = torch.device("cpu")
    seed_all(0)
    G = ref.load("generators.generator_13_5").Generator(NZ).to(cpu)
    D = ref.load("discriminators.discriminator_9_4").Discriminator().to(cpu)
    if loop == "ganR2":
        with torch.no_grad():
            D.fc[2].weight.weight.mul_(1e-3)
            D.fc[2].bias.fill_(0.5)
    T = ref.load(f"train.{loop}").Train([(torch.zeros(batch, 3, 64, 64), torch.zeros(batch))], cpu, 1, NZ, G, "G13_5",
                                        D, "D9_4")
    seed_all(1)
    return T


def cpu_rate(loop, batch, seconds_budget, max_steps=None, min_steps=3, threads=None):
    """One training iteration (D step + G step) of the final pair on the host cores, fp32, torch CPU ops: the reference
    itself when its tree is available (kind "reference"), else the restatement on oracle/kernels_ref.py ("port").
    The patched initial state is restored before every iteration, outside the timed interval, so BCELoss stays valid
    for any number of iterations; one untimed warm-up iteration, then at least `min_steps` timed ones and as many more
    as fit into `seconds_budget` (at most `max_steps`).
    -> dict(rate images/s, steps, sec_per_step, threads, kind)"""
    from gan_b200 import kernels as KN
    from oracle import reference as ref
    threads = threads or physical_cores()
    torch.set_num_threads(threads)
    kind = "reference" if ref.available() else "port"
    prev = None
    if kind == "port":
        from oracle.kernels_ref import RefBackend
        prev = KN.set_backend(RefBackend())
    try:
        T = build_reference_trainer(loop, batch) if kind == "reference" else build_trainer(torch.device("cpu"), loop, batch)
        snap = HostSnapshot(T)
        images = torch.randn(batch, 3, 64, 64).requires_grad_()

        def timed_iteration():
            snap.restore()
            images.grad = None
            t0 = time.perf_counter()
            one_iteration(T, images, batch)
            return time.perf_counter() - t0

        t_w = timed_iteration()                      # warm-up (allocator, thread pool); gives the per-step estimate
        n = max(min_steps, int((seconds_budget - t_w) / max(t_w, 1e-3)))
        if max_steps is not None:
            n = max(1, min(n, max_steps))
        times = [timed_iteration() for _ in range(n)]
        dt = sum(times) / n
    finally:
        if kind == "port":
            KN.set_backend(prev)
    return {"rate": batch / dt, "steps": n, "sec_per_step": dt, "threads": threads, "kind": kind}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cpu_batch = args.cpu_batch
    r = cpu_rate(args.loop, cpu_batch, args.cpu_seconds, max_steps=args.steps)
    what = ("the unmodified reference (train/%s.py Train.discriminator_trainstep + generator_trainstep)" % args.loop
            if r["kind"] == "reference" else "the CPU restatement of the step (oracle/kernels_ref.py)")
    sample = (f"{r['steps']} full iterations (D step + G step) of {what} at batch {cpu_batch} of the same G13_5+D9_4 "
              f"{args.loop} workload, fp32, torch CPU ops, {r['threads']} threads (physical cores), patched initial "
              f"state restored before every iteration outside the timed interval, 1 warm-up iteration")
    line = {"impl": "reference", "metric": "train images/sec (G13_5+D9_4, 64x64)", "value": r["rate"],
            "unit": "images/s", "n_gpus": args.gpus, "steps": r["steps"], "warmup": 1,
            "ms_per_step": r["sec_per_step"] * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"generator_13_5+discriminator_9_4 train/{args.loop}.py, batch {args.batch}/GPU "
                                   f"(CPU sample at batch {cpu_batch})", "loop": args.loop,
                       "requested_steps": args.steps, "requested_warmup": args.warmup,
                       "cpu_seconds_budget": args.cpu_seconds},
            "cpu_baseline": {"value": r["rate"], "unit": "images/s", "cores": r["threads"], "kind": r["kind"],
                             "sample": sample},
            "e2e": {"value": r["rate"], "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
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
    from gan_b200 import peer as gpeer
    import torch.distributed as tdist
    rank, world, device = gdist.init_from_env()
    assert device.type == "cuda", "bench.py's own arm needs a B200 (there is no CPU fallback)"
    assert world == args.gpus, f"--gpus {args.gpus} but WORLD_SIZE={world}"
    B = args.batch

    def barrier():
        if world > 1:
            tdist.barrier()
        torch.cuda.synchronize()

    def measure(loop, steps, warmup, want_census):
        """-> dict with ms (resident), ms_e2e, launches per step, census, losses, bce flag, NCCL calls per step"""
        T = build_trainer(device, loop, B, rank)
        dp = gdist.DataParallelStep(T).install()
        host_images = torch.randn(B, 3, 64, 64).pin_memory()
        dev_images = host_images.to(device).requires_grad_()
        out = {"h2d": host_images.numel() * 4}
        S, restore, check = T, None, (lambda: True)
        if args.graphs:
            from gan_b200.graphed import GraphedStep
            KN.LAUNCH_COUNT[0] = 0
            KN.CONV_CENSUS = {} if want_census else None
            gdist.CALLS[0] = 0
            gpeer.CALLS[0] = gpeer.FUSED[0] = 0
            nw = max(warmup, 2)
            S = GraphedStep(T, B, warmup=nw, images=dev_images, poll_every=0)
            out["launches_per_step"] = KN.LAUNCH_COUNT[0] // (nw + 1)
            out["nccl_calls_per_step"] = gdist.CALLS[0] // (nw + 1)
            out["peer_calls_per_step"] = gpeer.CALLS[0] // (nw + 1)
            out["peer_fused_per_step"] = gpeer.FUSED[0] // (nw + 1)
            if want_census:
                out["census"] = {k_: v // (nw + 1) for k_, v in KN.CONV_CENSUS.items()}
            KN.CONV_CENSUS = None
            dev_images = S.images
            restore = S.restore_initial_state

            def check():
                try:
                    S.check_bce_range()
                    return True
                except RuntimeError:
                    return False
        in_range = [True]
        last = {}

        def step_resident():
            last["d"], last["g"] = one_iteration(S, dev_images, B)

        d2h = [0]

        def step_e2e():
            imgs = host_images if args.graphs else host_images.to(device, non_blocking=True).requires_grad_()
            d_out, g_loss = one_iteration(S, imgs, B)
            losses = torch.stack([v.reshape(()) for v in d_out] + [g_loss.reshape(())]).cpu()
            d2h[0] = losses.numel() * 4
            last["losses"] = [float(v) for v in losses]

        def timed(fn, n):
            """n steps in segments of --restore-every: before each segment the patched initial state is restored
            (untimed); every segment is bracketed by barrier + synchronize and CUDA events on the launching stream."""
            seg = args.restore_every if (restore is not None and args.restore_every > 0) else n
            total, done = 0.0, 0
            while done < n:
                m = min(seg, n - done)
                if restore is not None and args.restore_every > 0:
                    restore()
                barrier()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(m):
                    fn()
                e1.record()
                barrier()
                total += e0.elapsed_time(e1)
                done += m
                in_range[0] = check() and in_range[0]
            ms = torch.tensor([total], device=device)
            if world > 1:
                tdist.all_reduce(ms, op=tdist.ReduceOp.MAX)
            return float(ms) / n

        w_done = 0
        while w_done < warmup:
            m = min(args.restore_every or warmup, warmup - w_done)
            if restore is not None and args.restore_every > 0:
                restore()
            for _ in range(m):
                step_resident()
            w_done += m
        torch.cuda.synchronize()
        in_range[0] = check() and in_range[0]
        KN.LAUNCH_COUNT[0] = 0
        KN.CONV_TIMER = None if args.graphs else KN.ConvTimer()
        if args.profile_range:
            torch.cuda.profiler.start()   # ncu --profile-from-start off: only the timed steps are profiled
        with ClockSampler(device.index or 0) as clk:
            out["ms"] = timed(step_resident, steps)
        if args.profile_range:
            torch.cuda.profiler.stop()
        out["clocks"] = clk.summary()
        if not args.graphs:
            out["launches_per_step"] = KN.LAUNCH_COUNT[0] // max(steps, 1)
        out["conv_timer"] = KN.CONV_TIMER.totals() if KN.CONV_TIMER is not None else {}
        KN.CONV_TIMER = None
        out["ms_e2e"] = timed(step_e2e, steps)
        out["d2h"] = d2h[0]
        out["losses"] = last.get("losses")
        out["bce_in_range"] = bool(in_range[0])
        out.setdefault("nccl_calls_per_step", 0)
        out.setdefault("peer_calls_per_step", 0)
        out.setdefault("peer_fused_per_step", 0)
        out["mem_gb"] = torch.cuda.max_memory_allocated(device) / 2**30
        out["adam"] = adam_roofline(T, device)
        dp.remove()
        del S, T
        return out

    main_run = measure(args.loop, args.steps, args.warmup, True)
    ms, ms_e2e = main_run["ms"], main_run["ms_e2e"]
    conv = main_run["conv_timer"]
    if args.graphs:
        conv = {k_: (v[0] * args.steps, v[1] * args.steps, v[2] * args.steps)
                for k_, v in replay_census(main_run["census"], device, dump=args.census_dump).items()}

    second = None
    if world == 1 and args.second_loop and args.second_loop != args.loop:
        # second headline: the loop main.py:77 really selects for this pair (strict superset: adds WGAN-GP, no BCE)
        try:
            import gc
            gc.collect()
            torch.cuda.empty_cache()
            r2 = measure(args.second_loop, max(2, min(args.steps, 5)), 3, False)
            second = {"loop": args.second_loop, "value": B * world / (r2["ms"] * 1e-3), "unit": "images/s",
                      "ms_per_step": r2["ms"], "e2e": B * world / (r2["ms_e2e"] * 1e-3),
                      "model_tflops": FLOPS_PER_IMAGE[args.second_loop] * B / (r2["ms"] * 1e-3) / 1e12,
                      "steps": max(2, min(args.steps, 5)), "losses": r2["losses"],
                      "gpu_launches": r2["launches_per_step"]}
        except Exception as e:  # never lose the main line to the extra measurement
            second = {"loop": args.second_loop, "error": f"{type(e).__name__}: {e}"[:300]}

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
    traffic, traffic_of = ncu_traffic()
    adam = main_run["adam"]
    roofline = {"bound": "tensor", "kernel": "conv_tc (tcgen05 implicit GEMM)" if eng == "tc" else
                "conv_simt (fp32 CUDA-core implicit GEMM)", "achieved": achieved, "peak": pk["tflops"],
                "unit": "TFLOP/s", "frac": achieved / pk["tflops"], "traffic": traffic,
                "traffic_of": traffic_of, "top_shapes": getattr(replay_census, "top", []),
                "peak_source": pk["src"],
                "launches": calls, "conv_share_of_step": conv_share,
                "by_engine": {k: {"tflops": v[0] / v[1] / 1e12 if v[1] > 0 else 0.0, "seconds": v[1], "calls": v[2]}
                              for k, v in conv.items()},
                "hbm": {"bound": "hbm", "kernel": "multi_tensor_adam_kernel (G + D, 28 B/param)",
                        "achieved": adam["gbs"], "peak": pk["hbm_gbs"], "unit": "GB/s",
                        "frac": adam["gbs"] / pk["hbm_gbs"], "bytes_per_launch_pair": adam["bytes"],
                        "ms": adam["ms"]}}
    line = {"metric": "train images/sec (G13_5+D9_4, 64x64)", "value": B * world / (ms * 1e-3), "unit": "images/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16" if KN.default_engine() != 1 else "f32", "data": "synthetic",
            "config": {"workload": f"generator_13_5+discriminator_9_4 train/{args.loop}.py, batch {B}/GPU, "
                                   f"global batch {B * world}, nz {NZ}", "loop": args.loop, "batch_per_gpu": B,
                       "l2": "per-step working set (activations + 2 GB of parameters) exceeds the 126 MB L2",
                       "parallelism": f"dp{world}", "cuda_graphs": bool(args.graphs),
                       "restore_every": args.restore_every,
                       "g_step_critic_wgrad": "computed (as the reference)" if os.environ.get("GANB200_G_STEP_D_GRADS") == "1"
                       else "skipped: unread by the reference (SURVEY B9); parameters and losses identical"},
            "model_tflops": FLOPS_PER_IMAGE[args.loop] * B * world / (ms * 1e-3) / 1e12 / world,
            "clocks": main_run["clocks"], "gpu_launches": main_run["launches_per_step"],
            "nccl_calls_per_step": main_run["nccl_calls_per_step"],
            "peer_calls_per_step": main_run["peer_calls_per_step"],
            "peer_fused_per_step": main_run["peer_fused_per_step"], "peak_mem_gib": main_run["mem_gb"],
            "bce_in_range": main_run["bce_in_range"], "losses": main_run["losses"],
            "e2e": {"value": B * world / (ms_e2e * 1e-3), "unit": "images/s",
                    "h2d_bytes_per_step": main_run["h2d"], "d2h_bytes_per_step": main_run["d2h"]},
            "roofline": roofline}
    if second is not None:
        line["second_loop"] = second
    if world == 1 and not args.no_cpu:
        r = cpu_rate(args.loop, args.cpu_batch, args.cpu_seconds, max_steps=8)
        line["cpu_baseline"] = {"value": r["rate"], "unit": "images/s", "cores": r["threads"], "kind": r["kind"],
                                "sample": f"{r['steps']} full iterations at batch {args.cpu_batch} of "
                                          f"{'the unmodified reference' if r['kind'] == 'reference' else 'the CPU restatement'}"
                                          f" on torch CPU ops (fp32), initial state restored before each, 1 warm-up"}
    print(json.dumps(line), flush=True)
    if world > 1:
        sys.stdout.flush()
        os._exit(0)


def adam_roofline(T, device):
    """HBM-class roofline entry: the fused multi-tensor Adam of both networks timed alone with CUDA events (28 bytes per
    live parameter: read p, g, m, v; write p, m, v).  Run on throw-away gradients AFTER the timed regions; the moments
    and parameters it touches are not used again."""
    total = 0
    for net in (T.generator, T.discriminator):
        for p in net.parameters():
            if p.grad is not None:
                total += p.numel()
    if total == 0:
        return {"gbs": 0.0, "bytes": 0, "ms": 0.0}
    saved = [(opt, opt.pre_step) for opt in (T.optimizer_G, T.optimizer_D)]
    for opt, _ in saved:
        opt.pre_step = None
        opt.capturable = True
    # device time only: issued eagerly, FusedAdam.step spends ~20 ms on the host walking 10 653 parameters, which is
    # what CUDA events would then measure; captured in a graph the two launches run back to back
    for _ in range(2):
        T.optimizer_G.step(); T.optimizer_D.step()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, capture_error_mode="thread_local"):
        T.optimizer_G.step(); T.optimizer_D.step()
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 5
    e0.record()
    for _ in range(reps):
        g.replay()
    e1.record()
    e1.synchronize()
    del g
    for opt, hook in saved:
        opt.pre_step = hook
    ms = e0.elapsed_time(e1) / reps
    return {"gbs": 28.0 * total / (ms * 1e-3) / 1e9, "bytes": 28 * total, "ms": ms}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--batch", type=int, default=int(os.environ.get("GANB200_BENCH_BATCH", "256")))
    ap.add_argument("--loop", default="ganR2", choices=["ganR2", "wgangpR2"])
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--cpu-batch", type=int, default=4)
    ap.add_argument("--cpu-seconds", type=float, default=None,
                    help="wall-clock budget of the CPU leg (default: 60 s for cpu_baseline, 240 s for --impl reference); "
                         "at least 3 iterations are always timed")
    ap.add_argument("--restore-every", type=int, default=4,
                    help="GPU arm: restore the patched initial state before every N timed steps, outside the timed "
                         "segments (keeps BCELoss inputs in [0,1], SURVEY F2); 0 = never")
    ap.add_argument("--second-loop", default="wgangpR2",
                    help="N=1: also time this loop (fewer steps) and report it under 'second_loop'; '' = off")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--profile-range", action="store_true", help="cudaProfilerStart/Stop around the timed steps")
    ap.add_argument("--census-dump", default=os.path.join(ROOT, "gpurun_out", "conv_census.json"))
    ap.add_argument("--graphs", type=int, default=int(os.environ.get("GANB200_GRAPHS", "1")),
                    help="1: replay the two step bodies as CUDA graphs (default); 0: eager launches")
    args = ap.parse_args()
    if args.cpu_seconds is None:
        args.cpu_seconds = 240.0 if args.impl == "reference" else 60.0
    if args.impl == "reference":
        run_reference(args)
    else:
        run_native(args)


if __name__ == "__main__":
    main()
