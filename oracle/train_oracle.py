"""ORACLE — test infrastructure only, not product code.

CPU restatement (torch fp32 autograd) of ONE iteration of the reference's denoise-GAN training loop,
backend/trainingcode/denoise_gan_code/training.py:
    DenoiseGenerator        :19-74   (= backend/app.py:39-103)
    DenoiseDiscriminator    :77-98
    trainer set-up          :230-242 (MSELoss, BCELoss, Adam lr 1e-4 betas (0.9, 0.999))
    the iteration           :409-427 (D step on real + detached fake, then G step: MSE + 0.001 * BCE)
The per-batch PSNR/SSIM/LPIPS metric pass (:432) and the VGG loss module (constructed at :233 but never used in
`train`) are out of scope.

Pinned: oracle/make_golden_train.py imports the reference's own `DenoiseGenerator` / `DenoiseDiscriminator` classes from
training.py (its absent third-party imports - skimage, lpips, pytorch_msssim, matplotlib, tqdm - stubbed in
sys.modules; none is used by the classes or the iteration), runs the iteration's statements verbatim on seeded
inputs and commits losses / gradient and parameter digests to tests/golden/train_step.npz;
tests/test_train_oracle.py checks this restatement against them on every run.
"""
from __future__ import annotations

from typing import Dict, Tuple

import torch
import torch.nn as nn


class DenoiseGenerator(nn.Module):
    """training.py:19-74 (module names = the state_dict keys of backend/app.py:42-78)."""

    def __init__(self):
        super().__init__()
        def block(ci, co, last_relu=True):
            mods = [nn.Conv2d(ci, co, 3, padding=1), nn.ReLU(), nn.Conv2d(co, co if last_relu else 3, 3, padding=1)]
            if last_relu:
                mods.append(nn.ReLU())
            return nn.Sequential(*mods)
        self.down1 = block(3, 64)
        self.pool1 = nn.MaxPool2d(2, 2)
        self.down2 = block(64, 128)
        self.pool2 = nn.MaxPool2d(2, 2)
        self.bottleneck = block(128, 256)
        self.up2 = nn.ConvTranspose2d(256, 128, 2, stride=2)
        self.upconv2 = block(256, 128)
        self.up1 = nn.ConvTranspose2d(128, 64, 2, stride=2)
        self.upconv1 = block(128, 64, last_relu=False)

    def forward(self, x):
        d1 = self.down1(x)
        d2 = self.down2(self.pool1(d1))
        b = self.bottleneck(self.pool2(d2))
        u2 = self.upconv2(torch.cat([self.up2(b), d2], dim=1))
        u1 = self.upconv1(torch.cat([self.up1(u2), d1], dim=1))
        return torch.tanh(u1)


class DenoiseDiscriminator(nn.Module):
    """training.py:77-98."""

    def __init__(self):
        super().__init__()
        self.model = nn.Sequential(
            nn.Conv2d(3, 64, kernel_size=3, padding=1), nn.LeakyReLU(0.2),
            nn.Conv2d(64, 64, kernel_size=3, stride=2, padding=1), nn.BatchNorm2d(64), nn.LeakyReLU(0.2),
            nn.Conv2d(64, 128, kernel_size=3, padding=1), nn.BatchNorm2d(128), nn.LeakyReLU(0.2),
            nn.Conv2d(128, 128, kernel_size=3, stride=2, padding=1), nn.BatchNorm2d(128), nn.LeakyReLU(0.2),
            nn.AdaptiveAvgPool2d(1), nn.Conv2d(128, 1, kernel_size=1), nn.Sigmoid())

    def forward(self, x):
        return self.model(x).view(-1)


def make_models(seed: int) -> Tuple[DenoiseGenerator, DenoiseDiscriminator]:
    torch.manual_seed(seed)
    return DenoiseGenerator(), DenoiseDiscriminator()


def make_optimizers(g: nn.Module, d: nn.Module, lr: float = 1e-4):
    """training.py:237-238."""
    return (torch.optim.Adam(g.parameters(), lr=lr, betas=(0.9, 0.999)),
            torch.optim.Adam(d.parameters(), lr=lr, betas=(0.9, 0.999)))


def train_iteration(g: nn.Module, d: nn.Module, g_opt, d_opt, noisy: torch.Tensor, clean: torch.Tensor) -> Dict[str, object]:
    """training.py:409-427, statement by statement. Returns losses and clones of both gradient sets."""
    mse, bce = nn.MSELoss(), nn.BCELoss()
    g.train()
    d.train()
    with torch.enable_grad():
        d_opt.zero_grad()
        denoised = g(noisy)
        real_pred = d(clean)
        fake_pred = d(denoised.detach())
        d_loss = bce(real_pred, torch.ones_like(real_pred)) + bce(fake_pred, torch.zeros_like(fake_pred))
        d_loss.backward()
        d_grads = {k: p.grad.detach().clone() for k, p in d.named_parameters()}
        d_opt.step()

        g_opt.zero_grad()
        fake_pred = d(denoised)
        content = mse(denoised, clean)
        adv = bce(fake_pred, torch.ones_like(fake_pred))
        g_loss = content + 0.001 * adv
        g_loss.backward()
        g_grads = {k: p.grad.detach().clone() for k, p in g.named_parameters()}
        g_opt.step()
    return {"d_loss": float(d_loss.detach()), "g_loss": float(g_loss.detach()), "content": float(content.detach()),
            "adv": float(adv.detach()),
            "d_grads": d_grads, "g_grads": g_grads, "denoised": denoised.detach()}


def synthetic_pair(seed: int, B: int, H: int, W: int, sigma: float = 0.1) -> Tuple[torch.Tensor, torch.Tensor]:
    """SURVEY §8(d) synthetic faces: smooth random field in [0,1], Gaussian noise sigma on [0,1], both mapped to [-1,1]."""
    gen = torch.Generator().manual_seed(seed)
    base = torch.rand(B, 3, max(H // 8, 1), max(W // 8, 1), generator=gen)
    clean = torch.nn.functional.interpolate(base, size=(H, W), mode="bicubic", align_corners=False).clamp(0, 1)
    noisy = (clean + sigma * torch.randn(B, 3, H, W, generator=gen)).clamp(0, 1)
    return (noisy - 0.5) / 0.5, (clean - 0.5) / 0.5


def instrumented_gradients(g: DenoiseGenerator, d: DenoiseDiscriminator, noisy: torch.Tensor, clean: torch.Tensor) -> Dict[str, object]:
    """The same two backward passes as `train_iteration` but WITHOUT the optimizer steps in between (so both use the same
    discriminator), with every pre-activation kept: returns losses, parameter gradients and the activation gradients
    named like the CUDA path's debug buffers (NHWC float32). Used to localise a failing layer."""
    import torch.nn.functional as F
    g.train()
    d.train()
    keep: Dict[str, torch.Tensor] = {}

    def k(name, t):
        t.retain_grad()
        keep[name] = t
        return t

    def gen_forward(x):
        z0 = k("dz0", g.down1[0](x)); a0 = F.relu(z0)
        z1 = k("dz1", g.down1[2](a0)); e1 = F.relu(z1)
        p1 = k("gp1", F.max_pool2d(e1, 2, 2))
        z2 = k("dz2", g.down2[0](p1)); a2 = F.relu(z2)
        z3 = k("dz3", g.down2[2](a2)); e2 = F.relu(z3)
        p2 = k("gp2", F.max_pool2d(e2, 2, 2))
        z4 = k("dz4", g.bottleneck[0](p2)); b0 = F.relu(z4)
        z5 = k("dz5", g.bottleneck[2](b0)); b1 = F.relu(z5)
        cat2 = k("gcat2", torch.cat([g.up2(b1), e2], dim=1))
        z7 = k("dz7", g.upconv2[0](cat2)); u2a = F.relu(z7)
        z8 = k("dz8", g.upconv2[2](u2a)); u2b = F.relu(z8)
        cat1 = k("gcat1", torch.cat([g.up1(u2b), e1], dim=1))
        z10 = k("dz10", g.upconv1[0](cat1)); u1a = F.relu(z10)
        z11 = k("gz", g.upconv1[2](u1a))
        return torch.tanh(z11)

    def disc_forward(x, tag=None):
        m = d.model
        z0 = m[0](x); a0 = m[1](z0)
        z1 = m[2](a0); a1 = m[4](m[3](z1))
        z2 = m[5](a1); a2 = m[7](m[6](z2))
        z3 = m[8](a2); a3 = m[10](m[9](z3))
        if tag:
            for n, t in (("d_g0", z0), ("d_g1", z1), ("d_g2", z2), ("d_g3", z3)):
                k(n, t)
            for n, t in (("d_a0", a0), ("d_z1", z1), ("d_a1", a1), ("d_z2", z2), ("d_a2", a2), ("d_z3", z3), ("d_a3", a3)):
                keep["fwd_" + n] = t.detach()
        return m[13](m[12](m[11](a3))).view(-1)

    bce, mse = nn.BCELoss(), nn.MSELoss()
    with torch.enable_grad():
        d.zero_grad(); g.zero_grad()
        y = gen_forward(noisy)
        real_pred = disc_forward(clean)
        fake_pred = disc_forward(y.detach())
        d_loss = bce(real_pred, torch.ones_like(real_pred)) + bce(fake_pred, torch.zeros_like(fake_pred))
        d_loss.backward()
        d_grads = {n: p.grad.detach().clone() for n, p in d.named_parameters()}
        d.zero_grad()
        y_in = y.detach().clone().requires_grad_(True)   # the discriminator's input gradient alone
        adv_only = bce(disc_forward(y_in), torch.ones(y.shape[0]))
        (0.001 * adv_only).backward()
        gx = y_in.grad.detach().clone()
        d.zero_grad()
        fake_pred = disc_forward(y, tag="g")
        content = mse(y, clean)
        adv = bce(fake_pred, torch.ones_like(fake_pred))
        g_loss = content + 0.001 * adv
        g_loss.backward()
        g_grads = {n: p.grad.detach().clone() for n, p in g.named_parameters()}
    nhwc = lambda t: t.permute(0, 2, 3, 1).contiguous()
    acts = {n: nhwc(t.grad) for n, t in keep.items() if not n.startswith("fwd_") and t.grad is not None}
    acts.update({n[4:]: nhwc(t) for n, t in keep.items() if n.startswith("fwd_")})
    acts["gx"] = nhwc(gx)
    acts["y"] = y.detach()
    return {"d_loss": float(d_loss.detach()), "g_loss": float(g_loss.detach()), "content": float(content.detach()),
            "adv": float(adv.detach()),
            "d_grads": d_grads, "g_grads": g_grads, "acts": acts}
