"""The GAN training step, all ten loss/penalty variants of the reference's train/*.py, as one recipe-driven
implementation on gan_b200.ops.

The reference has ten files that repeat the same two functions with small edits; here a `Recipe`
states what differs and `StepMixin` implements `generator_trainstep`, `discriminator_trainstep`,
`compute_grad2`, `gradient_penalty`, `discriminator_loss` once.  Sequencing is the reference's: the same RNG
calls in the same order (label noise, z, GP eps — SURVEY §3.2/F9), every loss term back-propagated separately so
that gradients accumulate in the same order, penalties through `torch.autograd.grad(create_graph=True)` — which
here records a graph made only of libganb200 kernels (conv dgrad-of-dgrad / wgrad-of-dgrad, PReLU mask-scale,
sigmoid', gate products, minibatch-stddev second derivative), see gan_b200/ops.py.

Reference lines: train/gan.py:26-53, ganR2.py:19-63, wgan.py:20-49, wgansn.py:17-39, wgansnR2.py:17-52,
wgangp.py:20-71, wgangpR2.py:17-67, wgangpsnR2.py:22-70, wganlazygp.py:17-70, wganlazygpR2.py:17-77.
"""
import contextlib
import os
from dataclasses import dataclass

import torch

from . import kernels as KN
from . import ops
from . import rng


SEPARATE_BACKWARD = os.environ.get("GANB200_SEPARATE_BACKWARD") == "1"


@dataclass(frozen=True)
class Recipe:
    name: str
    bce: bool = False            # BCE with noisy labels (gan, ganR2) instead of +-mean(D)
    r1: float = 0.0              # weight of mean_b |dD/dx_real|^2 (0 = off)
    r2: float = 0.0              # same on the generated batch
    gp: float = 0.0              # weight of mean_b (|dD/dx_hat| - 1)^2
    lazy: int = 1                # penalties only when idx % lazy == 0, weights multiplied by `lazy`
    clip: float = 0.0            # weight clipping bound (wgan)
    adamw: bool = False          # AdamW(betas 0.5,0.999) override (wgan, wgangp — SURVEY F8)
    fixed_noise: int = 64
    takes_idx: bool = False      # discriminator_trainstep(images, b_size, idx)
    returns: tuple = ("real_loss", "fake_loss")


RECIPES = {
    "gan": Recipe("gan", bce=True),
    "ganR2": Recipe("ganR2", bce=True, r1=1.0, r2=1.0, returns=("real_loss", "fake_loss", "r1", "r2")),
    "wgan": Recipe("wgan", clip=1.0, adamw=True),
    "wgansn": Recipe("wgansn"),
    "wgansnR2": Recipe("wgansnR2", r1=1.0, returns=("real_loss", "fake_loss", "r1")),
    "wgangp": Recipe("wgangp", gp=10.0, adamw=True, returns=("real_loss", "fake_loss", "gp")),
    "wgangpR2": Recipe("wgangpR2", r1=1.0, r2=1.0, gp=10.0, fixed_noise=8,
                       returns=("real_loss", "fake_loss", "gp", "r1", "r2")),
    "wgangpsnR2": Recipe("wgangpsnR2", r1=1.0, gp=10.0, returns=("real_loss", "fake_loss", "r1", "gp")),
    "wganlazygp": Recipe("wganlazygp", gp=10.0, lazy=5, fixed_noise=16, takes_idx=True,
                         returns=("real_loss", "fake_loss", "gp")),
    "wganlazygpR2": Recipe("wganlazygpR2", r1=1.0, r2=1.0, gp=10.0, lazy=5, fixed_noise=16, takes_idx=True,
                           returns=("real_loss", "fake_loss", "gp", "r1", "r2")),
}


def lerp_per_sample(x_real, x_fake, eps):
    """(1 - eps) * x_real + eps * x_fake with eps:(B,) broadcast per sample, no graph (wgangp.py:35-37)."""
    with torch.no_grad():
        one_minus = ops.affine(eps, -1.0, 1.0)
        return ops.add(ops.mul_bc(x_real, one_minus, KN.V_N), ops.mul_bc(x_fake, eps, KN.V_N))


@contextlib.contextmanager
def weights_frozen(module):
    """No parameter of `module` requires grad inside the block: autograd then skips every wgrad / bias / slope reduction of
    it and only propagates to its input.  Used for the critic in the generator step, where the reference's backward also
    fills D.*.grad (ganR2.py:25-26 — SURVEY B9, §8(a) a2: "1 D_f is wasted D-wgrad"): nobody reads those gradients and the
    next discriminator step zeroes them.  GANB200_G_STEP_D_GRADS=1 keeps the reference's accumulation."""
    if os.environ.get("GANB200_G_STEP_D_GRADS") == "1":
        yield
        return
    live = [p for p in module.parameters() if p.requires_grad]
    for p in live:
        p.requires_grad_(False)
    try:
        yield
    finally:
        for p in live:
            p.requires_grad_(True)


class StepMixin:
    """Needs: self.generator, self.discriminator, self.optimizer_G/D, self.device, self.nz, self.recipe
    (and self.criterion for the BCE recipes)."""

    # ---- pieces shared by every variant ------------------------------------------------------------
    def _noisy_labels(self, b_size, base):
        """torch.full + 0.05 * torch.rand: one RNG call of shape (b,1) (gan.py:27-28)."""
        u = rng.rand(b_size, 1, device=self.device)
        return ops.affine(u, 0.05, base)

    def _d_loss(self, pred, target, sign):
        if self.recipe.bce:
            return self.criterion(pred, target)
        return ops.affine(ops.mean(pred), sign, 0.0)

    def compute_grad2(self, d_out, x_in):
        """sum_{c,h,w} (d sum(d_out) / d x_in)^2 per sample, differentiable (ganR2.py:30-39)."""
        batch_size = x_in.size(0)
        grad_dout = torch.autograd.grad(outputs=ops.sum_all(d_out), inputs=x_in, create_graph=True,
                                        retain_graph=True, only_inputs=True)[0]
        assert (grad_dout.size() == x_in.size())
        return ops.sumsq_per_sample(grad_dout).reshape(batch_size)

    def gradient_penalty(self, x_real, x_fake, batch_size, device, center=1.):
        """WGAN-GP on per-sample interpolates (wgangp.py:34-43)."""
        eps = rng.rand(batch_size, device=device)
        x_interp = lerp_per_sample(x_real.detach(), x_fake.detach(), eps).requires_grad_()
        d_out = self.discriminator(x_interp)
        norm = ops.sqrt(self.compute_grad2(d_out, x_interp))
        return ops.mean(ops.square(ops.affine(norm, 1.0, -float(center))))

    def discriminator_loss(self, real_pred, fake_pred):
        """unused helper kept for API parity (wgangp.py:29-32)."""
        return ops.sub(ops.mean(fake_pred), ops.mean(real_pred))

    # ---- generator step ----------------------------------------------------------------------------
    def generator_trainstep(self, b_size):
        rc = self.recipe
        valid_ = self._noisy_labels(b_size, 0.95) if rc.bce else None
        self.optimizer_G.zero_grad()
        z = rng.randn(b_size, self.nz, 1, 1, device=self.device)
        gen_imgs = self.generator(z)
        with weights_frozen(self.discriminator):
            g_loss = self._d_loss(self.discriminator(gen_imgs), valid_, -1.0)
            g_loss.backward()
        self.optimizer_G.step()
        return gen_imgs, g_loss

    # ---- discriminator step ------------------------------------------------------------------------
    def discriminator_trainstep(self, images, b_size, idx=0):
        rc = self.recipe
        valid_ = fake_ = None
        if rc.bce:
            valid_ = self._noisy_labels(b_size, 0.95)
            fake_ = self._noisy_labels(b_size, 0.0)
            z = rng.randn(b_size, self.nz, 1, 1, device=self.device)
            self.optimizer_D.zero_grad()
        else:
            self.optimizer_D.zero_grad()
            z = rng.randn(b_size, self.nz, 1, 1, device=self.device)
        with torch.no_grad():
            gen_imgs = self.generator(z)
        gen_imgs.requires_grad_()
        penal = (idx % rc.lazy == 0)
        scale = float(rc.lazy)
        zero = (lambda: torch.zeros(1, device=self.device))
        out = {}

        def side(x, target, sign, weight, tag):
            pred = self.discriminator(x)
            loss = self._d_loss(pred, target, sign)
            if weight and penal:
                if not x.requires_grad:
                    raise RuntimeError("One of the differentiated Tensors does not require grad")
                if SEPARATE_BACKWARD:
                    loss.backward(retain_graph=True)
                reg = ops.mean(self.compute_grad2(pred, x))
                if weight * scale != 1.0:
                    reg = ops.affine(reg, weight * scale, 0.0)
                if SEPARATE_BACKWARD:
                    reg.backward()
                else:
                    # ONE sweep for the loss and its penalty (ganR2.py:53-56 runs two: real_loss.backward(retain_graph=True)
                    # then r2_reg_r.backward()).  Both sweeps walk the critic's forward graph backwards — a dgrad and a
                    # wgrad per convolution, 2 of the ~7 critic-forwards of work per side — with different seeds; the
                    # backward is linear in the seed, so seeding both roots at once and walking the shared part once
                    # gives the same parameter gradients (sums re-associated: fp32 rounding only) for 2 D_f less per
                    # side.  GANB200_SEPARATE_BACKWARD=1 restores the reference's two calls.
                    torch.autograd.backward([loss, reg])
                out[tag] = reg
            else:
                loss.backward()
                if weight:
                    out[tag] = zero()
            return loss

        out["real_loss"] = side(images, valid_, -1.0, rc.r1, "r1")
        out["fake_loss"] = side(gen_imgs, fake_, +1.0, rc.r2, "r2")
        if rc.gp:
            if penal:
                gp = ops.affine(self.gradient_penalty(images, gen_imgs, b_size, self.device), rc.gp * scale, 0.0)
                gp.backward()
                out["gp"] = gp
            else:
                out["gp"] = zero()
        self.optimizer_D.step()
        if rc.clip:
            k = KN.backend()
            with torch.no_grad():
                for p in self.discriminator.parameters():
                    k.clamp_(p.data, -rc.clip, rc.clip)
        return tuple(out[name] for name in rc.returns)


# ======================================================================================================
# train(): the epoch/batch loop and its per-variant side effects (reference train/<mode>.py `train`)
# ======================================================================================================
@dataclass(frozen=True)
class LoopPlan:
    images_require_grad: bool = True
    show: str = "gen"             # live preview each iteration: "gen", "both" (real+fake), "" (off)
    save_images_every: int = 0    # generator(fixed_noise) grid every N iterations (0 = never)
    record_every: int = 0         # make_record + write_record_to_txt every N iterations
    record_file: str = ""
    ckpt_name: str = ""           # save_ckpt(name, epoch+1, 0) after every epoch
    ckpt_at_end: bool = False
    plot_at_end: int = 0          # draw_plt_record(record_file, N)
    crash_gp: bool = False        # train/wgan.py:69 reads an undefined `gp`


LOOPS = {
    "gan": LoopPlan(images_require_grad=False, show="both"),
    "ganR2": LoopPlan(),
    "wgan": LoopPlan(images_require_grad=False, save_images_every=10, crash_gp=True),
    "wgansn": LoopPlan(save_images_every=10, ckpt_name="WGANSN"),
    "wgansnR2": LoopPlan(save_images_every=10, ckpt_name="WGANSNR2"),
    "wgangp": LoopPlan(images_require_grad=False, save_images_every=10, ckpt_name="WGANGP"),
    "wgangpR2": LoopPlan(show="", save_images_every=20, ckpt_name="WGANGPR2", ckpt_at_end=True),
    "wgangpsnR2": LoopPlan(save_images_every=10, ckpt_name="WGANGPSNR2"),
    "wganlazygp": LoopPlan(images_require_grad=False, show="", save_images_every=30, record_every=30,
                           record_file="wganlazygp", plot_at_end=30),
    "wganlazygpR2": LoopPlan(show="", save_images_every=30, record_every=30, record_file="wganlazygpR2",
                             plot_at_end=30),
}

_RECORD_LABEL = {"real_loss": "Discriminator real loss", "fake_loss": "Discriminator fake loss",
                 "gp": "Gradient penalties", "r1": "Simplified gradient penalties R1",
                 "r2": "Simplified gradient penalties R2"}
_POSTFIX_LABEL = {"gan": {"real_loss": "Real_loss", "fake_loss": "Fake_loss"},
                  "*": {"r1": "r2_reg_r", "r2": "r2_reg_f"},
                  "wgansnR2": {"r1": "r2_reg"}, "wgangpsnR2": {"r1": "r2_reg"}}


def _progress(total):
    """tqdm.tk in the reference; headless boxes have no Tk, so fall back to the console bar."""
    try:
        from tqdm.tk import tqdm as bar
        return bar(total=total)
    except Exception:
        from tqdm import tqdm as bar
        return bar(total=total)


class LoopMixin:
    def train(self):
        rc, plan = self.recipe, LOOPS[self.recipe.name]
        self.load_generator_ckpt('')
        self.load_discriminator_ckpt('')
        proc_bar = _progress(len(self.dataloader))
        print("Starting Training Loop...")
        for epoch in range(self.num_epochs):
            for i, (images, _) in enumerate(self.dataloader):
                images = images.to(self.device)
                if plan.images_require_grad:
                    images.requires_grad_()
                b_size = images.shape[0]
                if rc.takes_idx:
                    d_out = self.discriminator_trainstep(images, b_size, i)
                else:
                    d_out = self.discriminator_trainstep(images, b_size)
                gen_imgs, g_loss = self.generator_trainstep(b_size)
                named = dict(zip(rc.returns, d_out))
                if plan.show == "gen":
                    self.show_images(gen_imgs, b_size)
                elif plan.show == "both":
                    self.show_images(torch.cat([images, gen_imgs]), b_size * 2)
                if plan.record_every and i % plan.record_every == 0:
                    record = [(_RECORD_LABEL[k], v.item()) for k, v in named.items()]
                    record.append(('Generator loss', g_loss.item()))
                    self.make_record(record)
                    self.write_record_to_txt(plan.record_file, record)
                if plan.save_images_every and i % plan.save_images_every == 0:
                    self.save_images("generated_images/", epoch, i)
                if plan.crash_gp:
                    raise NameError("name 'gp' is not defined")
                labels = {**_POSTFIX_LABEL["*"], **_POSTFIX_LABEL.get(rc.name, {})}
                post = {"epoch": f"{epoch}", "Loss_G": f"{g_loss.item():.4f}"}
                post.update({labels.get(k, k): f"{v.item():.4f}" for k, v in named.items()})
                proc_bar.set_postfix(post)
                proc_bar.update(1)
            proc_bar.reset()
            if plan.ckpt_name:
                self.save_ckpt(plan.ckpt_name, epoch + 1, 0)
        if plan.plot_at_end:
            self.draw_plt_record(plan.record_file, plan.plot_at_end)
        proc_bar.close()
        if plan.ckpt_at_end:
            self.save_ckpt(plan.ckpt_name, self.num_epochs, 0)
