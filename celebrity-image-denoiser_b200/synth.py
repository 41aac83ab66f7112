"""Synthetic "face" batches for benchmarks and tests (no dataset access on the build/GPU boxes):
a low-frequency random field per image, clamped to [0,1], plus Gaussian noise sigma=0.1 on the
[0,1] scale (the esrgan_addNoise.py:7-13 convention), truncated to uint8."""
from __future__ import annotations

import torch


def synthetic_noisy_u8(seed: int, B: int, H: int, W: int, sigma: float = 0.1) -> torch.Tensor:
    """CPU uint8 [B,H,W,3]."""
    g = torch.Generator().manual_seed(seed)
    base = torch.rand(B, 3, max(H // 8, 2), max(W // 8, 2), generator=g)
    clean = torch.nn.functional.interpolate(base, size=(H, W), mode="bicubic", align_corners=False).clamp(0, 1)
    noisy = (clean + sigma * torch.randn(B, 3, H, W, generator=g)).clamp(0, 1)
    return (noisy * 255).to(torch.uint8).permute(0, 2, 3, 1).contiguous()


def synthetic_noisy_u8_device(seed: int, B: int, H: int, W: int, device="cuda", var: float = 0.01):
    """The same kind of batch built ON THE DEVICE by this package's own data-prep kernels (what a training-data job
    would run): a random low-frequency u8 field -> Pillow-compatible bicubic up-sampling (cid_resize_bicubic_u8) -> the
    [0,1]-domain Gaussian corruption with sigma = sqrt(var) = 0.1 drawn by Philox4x32-10 (cid_noise_unit_u8,
    esrgan_addNoise.py:7-13). Returns (noisy, clean), CUDA uint8 [B,H,W,3]."""
    from . import noiser

    dev = torch.device(device)
    g = torch.Generator(device=dev).manual_seed(seed)
    base = torch.randint(0, 256, (B, max(H // 8, 2), max(W // 8, 2), 3), dtype=torch.uint8, device=dev, generator=g)
    clean = noiser.resize_bicubic(base, (W, H))
    return noiser.add_gaussian(clean, 0.0, var, seed=seed), clean
