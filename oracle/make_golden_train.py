"""Golden vectors for the training-step oracle, minted from the UNMODIFIED reference classes (build container only).

Imports backend/trainingcode/denoise_gan_code/training.py from /root/reference with its absent third-party imports
stubbed (skimage, lpips, pytorch_msssim, matplotlib, tqdm, sklearn are present or stubbed; none is touched by the
classes or by the iteration), builds the reference's DenoiseGenerator / DenoiseDiscriminator under a fixed seed, runs the
statements of training.py:409-427 on a seeded synthetic batch and writes losses, per-tensor gradient norms, a few raw
gradient tensors and the post-step parameter digests to tests/golden/train_step.npz.
"""
from __future__ import annotations

import importlib
import sys
from pathlib import Path
from unittest import mock

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from oracle.ref_import import REFERENCE_ROOT  # noqa: E402
from oracle.train_oracle import synthetic_pair  # noqa: E402


def load_reference_training():
    for name in ("skimage", "skimage.metrics", "lpips", "pytorch_msssim", "matplotlib", "matplotlib.pyplot", "tqdm"):
        if name not in sys.modules:
            try:
                importlib.import_module(name)
            except Exception:
                sys.modules[name] = mock.MagicMock()
    d = str(REFERENCE_ROOT / "backend" / "trainingcode" / "denoise_gan_code")
    sys.path.insert(0, d)
    try:
        return importlib.import_module("training")
    finally:
        sys.path.remove(d)


def main():
    ref = load_reference_training()
    torch.manual_seed(7)
    g, d = ref.DenoiseGenerator(), ref.DenoiseDiscriminator()
    g_opt = torch.optim.Adam(g.parameters(), lr=1e-4, betas=(0.9, 0.999))
    d_opt = torch.optim.Adam(d.parameters(), lr=1e-4, betas=(0.9, 0.999))
    mse_loss, bce_loss = torch.nn.MSELoss(), torch.nn.BCELoss()
    noisy_img, clean_img = synthetic_pair(11, 4, 32, 32)
    out = {}
    g.train(); d.train()
    for it in range(2):
        # ---- training.py:409-427 verbatim (self.* -> locals) ----
        d_opt.zero_grad()
        denoised_img = g(noisy_img)
        real_pred = d(clean_img)
        fake_pred = d(denoised_img.detach())
        d_loss = bce_loss(real_pred, torch.ones_like(real_pred)) + \
                 bce_loss(fake_pred, torch.zeros_like(fake_pred))
        d_loss.backward()
        d_gn = {k: float(p.grad.norm()) for k, p in d.named_parameters()}
        d_g0 = d.model[0].weight.grad.clone()
        d_opt.step()
        g_opt.zero_grad()
        fake_pred = d(denoised_img)
        content_loss = mse_loss(denoised_img, clean_img)
        adv_loss = bce_loss(fake_pred, torch.ones_like(fake_pred))
        g_loss = content_loss + 0.001 * adv_loss
        g_loss.backward()
        g_gn = {k: float(p.grad.norm()) for k, p in g.named_parameters()}
        g_g11 = g.upconv1[2].weight.grad.clone()
        g_opt.step()
        # ---------------------------------------------------------
        out[f"it{it}_losses"] = np.array([float(d_loss), float(g_loss), float(content_loss), float(adv_loss)], np.float64)
        out[f"it{it}_d_gradnorm"] = np.array([d_gn[k] for k in sorted(d_gn)], np.float64)
        out[f"it{it}_g_gradnorm"] = np.array([g_gn[k] for k in sorted(g_gn)], np.float64)
        out[f"it{it}_d_grad_model0"] = d_g0.numpy()
        out[f"it{it}_g_grad_upconv1_2"] = g_g11.numpy()
    out["d_keys"] = np.array(sorted(k for k, _ in d.named_parameters()))
    out["g_keys"] = np.array(sorted(k for k, _ in g.named_parameters()))
    out["g_param_sums"] = np.array([float(p.double().sum()) for _, p in sorted(g.named_parameters())], np.float64)
    out["d_param_sums"] = np.array([float(p.double().sum()) for _, p in sorted(d.named_parameters())], np.float64)
    out["d_running_mean3"] = d.model[3].running_mean.numpy()
    out["d_running_var9"] = d.model[9].running_var.numpy()
    dst = ROOT / "tests" / "golden" / "train_step.npz"
    np.savez_compressed(dst, **out)
    print("wrote", dst, {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
