"""Fork / join of independent sub-graphs onto CUDA side streams.

generator_13_5's ResnetInit runs four StyleBlocks that do not depend on each other (residual, residual_across,
transient, transient_across: reference generators/generator_13_5.py:341-347).  Each is a chain of ~60 launches of which
only the modulated convolutions fill the GPU; the SK-attention tail between them is dozens of 5-15 us kernels on 5x5 maps
and (B, C) vectors that occupy a few SMs each.  On one stream those tails serialise (~250 ms of a 1.2 s iteration at
B = 256, CUPTI); issued on four streams the tails of the four blocks run side by side, and so do their backward passes
(autograd runs every backward node on the stream its forward ran on and inserts the cross-stream waits itself).  Under
CUDA-graph capture the side streams fork from and join the capturing stream, so the captured graph carries the parallel
branches.

Memory safety with PyTorch's caching allocator (per-stream pools): inputs were produced on the forking stream before the
fork event and are not freed before the join; every tensor a branch returns is `record_stream`-ed on the joining stream,
so its block is not recycled by the side stream while the consumer still reads it.
GANB200_BRANCH_STREAMS=N spreads the branches over N streams.  DEFAULT 1 (off): measured on B200, eager results equal the
single-stream run to the step's own run-to-run noise and forward / backward capture works, but at the benched batch of 256
per GPU the iteration already peaks at 145 of 178 GiB and the per-stream allocator pools plus the frees that the caching
allocator defers until the END of a capture for every cross-stream tensor push the generator-step capture past the HBM
(torch.OutOfMemoryError at 158 GiB, N = 2 and N = 4).  Usable at smaller batches; needs bf16 saved activations first."""
import os

import torch

_pools = {}


def width():
    """number of streams the branches are spread over: GANB200_BRANCH_STREAMS = 1 (off) | 2 | 4 ..."""
    v = os.environ.get("GANB200_BRANCH_STREAMS", "1")
    try:
        return max(1, int(v))
    except ValueError:
        return 1


def _streams(device, n):
    key = (device.type, device.index)
    pool = _pools.setdefault(key, [])
    while len(pool) < n:
        pool.append(torch.cuda.Stream(device=device))
    return pool[:n]


def _tensors(obj):
    if torch.is_tensor(obj):
        yield obj
    elif isinstance(obj, (list, tuple)):
        for o in obj:
            yield from _tensors(o)


def run_parallel(fns, device):
    """[fn() for fn in fns] spread round-robin over `width()` streams (group 0 on the current stream); joined before
    returning.  Per-stream allocator pools cost memory: every extra stream holds its own high-water mark of temporaries."""
    k = min(width(), len(fns))
    if k < 2 or device.type != "cuda":
        return [fn() for fn in fns]
    main = torch.cuda.current_stream(device)
    fork = torch.cuda.Event()
    fork.record(main)
    outs = [None] * len(fns)
    sides = _streams(device, k - 1)
    joins = []
    for g, side in enumerate(sides, start=1):
        side.wait_event(fork)
        with torch.cuda.stream(side):
            for i in range(g, len(fns), k):
                outs[i] = fns[i]()
            done = torch.cuda.Event()
            done.record(side)
        joins.append((g, done))
    for i in range(0, len(fns), k):
        outs[i] = fns[i]()
    for g, done in joins:
        main.wait_event(done)
        for i in range(g, len(fns), k):
            for t in _tensors(outs[i]):
                t.record_stream(main)
    return outs
