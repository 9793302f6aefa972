"""GPU: the execution mode bench.py times — gan_b200.graphed.GraphedStep (two CUDA graphs per iteration, fused Philox
StyleConv noise, generator step with the critic's weights frozen, device-side Adam step counter) — against the same
iterations issued eagerly from the same seeds, on the final pair (generator_13_5 + discriminator_9_4, train/ganR2.py
with bench.py's weight patch).  Also: building a GraphedStep leaves the trainer's state untouched, and the BCELoss
range check that cannot run inside a captured graph trips the sticky device flag and raises the reference's error."""
import os
import random
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

pytestmark = pytest.mark.gpu
B = 8
ITERS = 3


def _seed(s):
    torch.manual_seed(s)
    np.random.seed(s)
    random.seed(s)


def _trainer(dev, patch=True):
    from discriminators import discriminator_9_4
    from generators import generator_13_5
    from train import ganR2
    _seed(0)
    G = generator_13_5.Generator(256).to(dev)
    D = discriminator_9_4.Discriminator().to(dev)
    if patch:  # bench.py / SURVEY §8(d): weights only, keeps D's logits inside (0,1) for the first iterations
        with torch.no_grad():
            D.fc[2].weight.weight.mul_(1e-3)
            D.fc[2].bias.fill_(0.5)
    return ganR2.Train([(torch.zeros(B, 3, 64, 64), torch.zeros(B))], dev, 1, 256, G, "G13_5", D, "D9_4")


def _params(T):
    out = {}
    for tag, net in (("G", T.generator), ("D", T.discriminator)):
        for n, p in net.named_parameters():
            out[f"{tag}.{n}"] = p.detach().clone()
        for n, b in net.named_buffers():
            out[f"{tag}.buf.{n}"] = b.detach().clone()
    return out


@pytest.mark.timeout(1800)
def test_graph_replay_equals_eager_iterations_on_the_final_pair():
    from gan_b200 import rng
    from gan_b200.graphed import GraphedStep
    assert rng.noise_mode() == "philox" and rng.fused_noise(), "the benched mode uses the fused in-kernel noise"
    dev = torch.device("cuda:0")
    images = torch.randn(B, 3, 64, 64, generator=torch.Generator().manual_seed(3)).to(dev)

    # eager leg
    rng._states.clear()
    T = _trainer(dev)
    _seed(1)
    eager_losses = []
    x = images.clone().requires_grad_()
    for _ in range(ITERS):
        d = T.discriminator_trainstep(x, B)
        _, g = T.generator_trainstep(B)
        x.grad = None
        eager_losses.append([float(v) for v in d] + [float(g)])
    eager = _params(T)
    del T
    torch.cuda.empty_cache()

    # graph leg: same seeds; construction (warm-up + capture) must leave the state untouched
    rng._states.clear()
    T = _trainer(dev)
    _seed(1)
    before = _params(T)
    S = GraphedStep(T, B, warmup=2, images=images, poll_every=0)
    after = _params(T)
    for n, v in before.items():
        assert torch.equal(after[n], v), f"GraphedStep construction changed {n}"
    for opt in (T.optimizer_G, T.optimizer_D):
        assert all(st.get("step", 0) == 0 for st in opt.state.values())
    _seed(1)
    graph_losses = []
    for _ in range(ITERS):
        d = S.discriminator_trainstep(S.images, B)
        _, g = S.generator_trainstep(B)
        graph_losses.append([float(v) for v in d] + [float(g)])
    S.check_bce_range()
    graphed = _params(T)

    for a, b in zip(eager_losses, graph_losses):
        for u, v in zip(a, b):
            # same kernels on the same operands; the only order-dependent arithmetic is fp32 atomics (loss / stddev /
            # reduction sums), which Adam's g/sqrt(v) then carries into the next iteration: measured 2e-6 after 2 steps
            assert abs(u - v) <= 2e-5 * max(1.0, abs(u)), (eager_losses, graph_losses)
    worst = ("", 0.0)
    for n, v in eager.items():
        w = graphed[n]
        if not v.is_floating_point():
            assert torch.equal(v, w), n
            continue
        rel = float((v - w).norm() / max(float(v.norm()), 1e-12))
        if rel > worst[1]:
            worst = (n, rel)
    assert worst[1] <= 2e-5, worst
    for opt in (T.optimizer_G, T.optimizer_D):
        assert all(st["step"] == ITERS for st in opt.state.values() if "step" in st)


@pytest.mark.timeout(1800)
def test_graph_replay_trips_the_bce_range_flag_like_the_reference_raises():
    """Un-patched D9_4 logits leave [0,1] at initialisation: eager raises in the warm-up, exactly where the
    reference's torch.nn.BCELoss raises.  With patched weights the capture succeeds; pushing the last bias out of range
    afterwards must trip the sticky flag on replay."""
    from gan_b200.graphed import GraphedStep
    dev = torch.device("cuda:0")
    images = torch.randn(B, 3, 64, 64, generator=torch.Generator().manual_seed(3)).to(dev)
    T = _trainer(dev, patch=False)
    _seed(1)
    with pytest.raises(RuntimeError, match="between 0 and 1"):
        GraphedStep(T, B, warmup=1, images=images)
    del T
    torch.cuda.empty_cache()
    T = _trainer(dev)
    _seed(1)
    S = GraphedStep(T, B, warmup=1, images=images, poll_every=1)
    S.discriminator_trainstep(S.images, B)
    S.generator_trainstep(B)          # polls: in range
    with torch.no_grad():
        T.discriminator.fc[2].bias.fill_(1.5)
    S.discriminator_trainstep(S.images, B)
    with pytest.raises(RuntimeError, match="between 0 and 1"):
        S.generator_trainstep(B)
