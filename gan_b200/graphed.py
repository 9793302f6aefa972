"""Whole-step CUDA graphs for the training iteration.

One iteration of the final pair is ~53 000 kernel launches (the reference issues ~31 000 ATen
calls per G13_5 forward alone); issued one by one from Python the step is bound by the host, not by the GPU.
`GraphedStep` runs a few eager warm-up iterations (optimizer state, kernel attributes, BCELoss range check — the
eager iterations are the ones that can raise), then captures `discriminator_trainstep` and `generator_trainstep` —
forward, the create_graph double backward, backward and the fused Adam update — into two CUDA graphs that share
one memory pool, and replays them.  Every launch inside is a libganb200 kernel on the capturing stream; nothing
synchronises, workspaces and tensor maps are fixed at capture time, the Adam step count lives on the device.

Drop-in use:  T = train.ganR2.Train(...);  S = GraphedStep(T, batch);  S.discriminator_trainstep(images, b);
S.generator_trainstep(b)  — same return values as the Train methods (static output tensors, overwritten by the
next replay).  Lazy-penalty recipes (wganlazygp*) alternate two different step bodies and are not graphed.
"""
import torch


class GraphedStep:
    def __init__(self, trainer, batch, warmup=3, images=None):
        T = self.T = trainer
        rc = T.recipe
        if rc.lazy != 1:
            raise ValueError("lazy-penalty recipes are not graph-capturable (two step bodies)")
        dev = T.device
        self.batch = batch
        needs_grad = bool(rc.r1)
        self.images = torch.zeros(batch, 3, 64, 64, device=dev)
        if images is not None:
            self.images.copy_(images.detach())
        if needs_grad:
            self.images.requires_grad_()
        for opt in (T.optimizer_G, T.optimizer_D):
            opt.capturable = True
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            for _ in range(warmup):
                T.discriminator_trainstep(self.images, batch)
                T.generator_trainstep(batch)
                self.images.grad = None
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        self.graph_d = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph_d, capture_error_mode="thread_local"):
            self.d_out = T.discriminator_trainstep(self.images, batch)
            self.images.grad = None
        self.graph_g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph_g, pool=self.graph_d.pool(), capture_error_mode="thread_local"):
            self.g_out = T.generator_trainstep(batch)

    def discriminator_trainstep(self, images, b_size, idx=0):
        assert b_size == self.batch
        if images is not self.images:
            with torch.no_grad():
                self.images.copy_(images, non_blocking=True)
        self.graph_d.replay()
        self._advance(self.T.optimizer_D)
        return self.d_out

    def generator_trainstep(self, b_size):
        assert b_size == self.batch
        self.graph_g.replay()
        self._advance(self.T.optimizer_G)
        return self.g_out

    @staticmethod
    def _advance(opt):
        # keep the host-side torch.optim step counters in line with the device counter the graph bumps
        for st in opt.state.values():
            if "step" in st:
                st["step"] += 1
