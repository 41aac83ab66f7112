"""Device-side synthetic noiser (test-input builder), same names as the reference's functions.

Reference: add_gaussian_noise / add_salt_pepper_noise / add_speckle_noise / add_poisson_noise / add_uniform_noise
(backend/trainingcode/denoise_gan_code/noise_generation.py:6-39, u8 domain) and add_gaussian / add_salt_pepper /
add_speckle / add_uniform (backend/trainingcode/esrgan_code/esrgan_addNoise.py:7-41, [0,1] domain). Inputs/outputs are CUDA uint8
tensors [..., H, W, C]; RNG is Philox4x32-10 on the device, or the caller's own draws (numpy's
MT19937 stream is not reproducible on a GPU) for bit-exact parity.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import lib as _lib
from .generator import _Engine

_engines = {}


def _engine(device: torch.device) -> _Engine:
    if device.type != "cuda":
        raise RuntimeError("the noiser kernels run only on CUDA tensors; there is no CPU fallback")
    idx = device.index if device.index is not None else torch.cuda.current_device()
    if idx not in _engines:
        _engines[idx] = _Engine(idx)
    return _engines[idx]


def add_gaussian_noise(image: torch.Tensor, mean: float = 0.0, sigma: float = 25.0, *, seed: int = 0,
                       offset: int = 0, noise: Optional[torch.Tensor] = None) -> torch.Tensor:
    """u8(clip(image + N(mean, sigma), 0, 255)), truncating cast. `noise`: optional CUDA float64
    tensor of image's shape holding the draws to use verbatim."""
    eng = _engine(image.device)
    x = image.contiguous()
    if x.dtype != torch.uint8:
        raise RuntimeError("image must be uint8")
    out = torch.empty_like(x)
    nptr = None
    if noise is not None:
        noise = noise.to(device=x.device, dtype=torch.float64).contiguous()
        if noise.numel() != x.numel():
            raise RuntimeError("noise must have the image's number of elements")
        nptr = C.c_void_p(noise.data_ptr())
    stream = torch.cuda.current_stream(x.device).cuda_stream
    _lib.check(eng.lib.cid_noise_gaussian_u8(eng.handle, C.c_void_p(x.data_ptr()), C.c_void_p(out.data_ptr()),
                                             x.numel(), float(mean), float(sigma), nptr, seed, offset,
                                             C.c_void_p(stream)), "cid_noise_gaussian_u8")
    return out


def _elementwise(image: torch.Tensor, kind: int, p0: float, p1: float, seed: int, offset: int,
                 draws: Optional[torch.Tensor]) -> torch.Tensor:
    eng = _engine(image.device)
    x = image.contiguous()
    if x.dtype != torch.uint8:
        raise RuntimeError("image must be uint8")
    out = torch.empty_like(x)
    dptr = None
    if draws is not None:
        draws = draws.to(device=x.device, dtype=torch.float64).contiguous()
        if draws.numel() != x.numel():
            raise RuntimeError("draws must have the image's number of elements")
        dptr = C.c_void_p(draws.data_ptr())
    stream = torch.cuda.current_stream(x.device).cuda_stream
    _lib.check(eng.lib.cid_noise_elementwise_u8(eng.handle, C.c_void_p(x.data_ptr()), C.c_void_p(out.data_ptr()),
                                                x.numel(), kind, float(p0), float(p1), dptr, seed, offset,
                                                C.c_void_p(stream)), "cid_noise_elementwise_u8")
    return out


def add_speckle_noise(image: torch.Tensor, mean: float = 0.0, sigma: float = 0.1, *, seed: int = 0, offset: int = 0,
                      noise: Optional[torch.Tensor] = None) -> torch.Tensor:
    """u8(clip(image + image * N(mean, sigma), 0, 255)) (noise_generation.py:24-28). `noise`: the normal draws."""
    return _elementwise(image, 1, mean, sigma, seed, offset, noise)


def add_uniform_noise(image: torch.Tensor, low: float = 0.0, high: float = 25.0, *, seed: int = 0, offset: int = 0,
                      noise: Optional[torch.Tensor] = None) -> torch.Tensor:
    """u8(clip(image + U(low, high), 0, 255)) (noise_generation.py:35-39). `noise`: the uniform draws."""
    return _elementwise(image, 2, low, high, seed, offset, noise)


def add_poisson_noise(image: torch.Tensor, *, seed: int = 0, offset: int = 0,
                      counts: Optional[torch.Tensor] = None) -> torch.Tensor:
    """np.random.poisson(image).astype(np.uint8) (noise_generation.py:30-33; counts above 255 wrap).
    `counts`: the reference's own Poisson draws."""
    return _elementwise(image, 3, 0.0, 0.0, seed, offset, counts)


def add_salt_pepper_noise(image: torch.Tensor, salt_prob: float = 0.02, pepper_prob: float = 0.02, *, seed: int = 0,
                          offset: int = 0, coords: Optional[torch.Tensor] = None) -> torch.Tensor:
    """image: CUDA uint8 [H,W,C] or [B,H,W,C]. coords: optional CUDA int32 [B, n_salt+n_pepper, 2]
    of (row, col) — salt draws first, then pepper draws."""
    eng = _engine(image.device)
    x = image.contiguous()
    if x.dtype != torch.uint8 or x.dim() not in (3, 4):
        raise RuntimeError("image must be uint8 [H,W,C] or [B,H,W,C]")
    B, H, W, Cc = (1, *x.shape) if x.dim() == 3 else x.shape
    out = torch.empty_like(x)
    cptr, per_image = None, 0
    if coords is not None:
        coords = coords.to(device=x.device, dtype=torch.int32).contiguous()
        if coords.dim() == 2:
            coords = coords[None]
        if coords.dim() != 3 or coords.shape[0] != B or coords.shape[2] != 2:
            raise RuntimeError(f"coords must be int32 [B, n_salt+n_pepper, 2], got {tuple(coords.shape)}")
        cptr, per_image = C.c_void_p(coords.data_ptr()), int(coords.shape[1])
    stream = torch.cuda.current_stream(x.device).cuda_stream
    _lib.check(eng.lib.cid_noise_salt_pepper_u8(eng.handle, C.c_void_p(x.data_ptr()), C.c_void_p(out.data_ptr()),
                                                B, H, W, Cc, float(salt_prob), float(pepper_prob), cptr, per_image,
                                                seed, offset, C.c_void_p(stream)), "cid_noise_salt_pepper_u8")
    return out


# ---- the [0,1]-domain noiser of the ESRGAN data-prep script (esrgan_code/esrgan_addNoise.py), same function names ----
def _unit(image: torch.Tensor, kind: int, p0: float, p1: float, seed: int, offset: int,
          draws: Optional[torch.Tensor]) -> torch.Tensor:
    eng = _engine(image.device)
    x = image.contiguous()
    if x.dtype != torch.uint8:
        raise RuntimeError("image must be uint8 (the script's img = u8 / 255.0 and (noisy * 255).astype(uint8) are fused in)")
    out = torch.empty_like(x)
    dptr = None
    if draws is not None:
        draws = draws.to(device=x.device, dtype=torch.float64).contiguous()
        if draws.numel() != x.numel():
            raise RuntimeError("draws must have the image's number of elements")
        dptr = C.c_void_p(draws.data_ptr())
    stream = torch.cuda.current_stream(x.device).cuda_stream
    _lib.check(eng.lib.cid_noise_unit_u8(eng.handle, C.c_void_p(x.data_ptr()), C.c_void_p(out.data_ptr()), x.numel(), kind,
                                         float(p0), float(p1), dptr, seed, offset, C.c_void_p(stream)), "cid_noise_unit_u8")
    return out


def add_gaussian(image: torch.Tensor, mean: float = 0.0, var: float = 0.01, *, seed: int = 0, offset: int = 0,
                 noise: Optional[torch.Tensor] = None) -> torch.Tensor:
    """u8((clip(image / 255 + N(mean, sqrt(var)), 0, 1)) * 255) (esrgan_addNoise.py:7-13, 66, 79): sigma 0.1 on the [0,1]
    scale, the corruption BASELINE.json configs[0] names. `noise`: the reference's own normal draws (float64)."""
    return _unit(image, 0, mean, var ** 0.5, seed, offset, noise)


def add_speckle(image: torch.Tensor, *, seed: int = 0, offset: int = 0, noise: Optional[torch.Tensor] = None) -> torch.Tensor:
    """clip(img + img * randn, 0, 1) (esrgan_addNoise.py:27-30)."""
    return _unit(image, 1, 0.0, 1.0, seed, offset, noise)


def add_uniform(image: torch.Tensor, low: float = -0.05, high: float = 0.05, *, seed: int = 0, offset: int = 0,
                noise: Optional[torch.Tensor] = None) -> torch.Tensor:
    """clip(img + U(low, high), 0, 1) (esrgan_addNoise.py:38-41)."""
    return _unit(image, 2, low, high, seed, offset, noise)


def add_salt_pepper(image: torch.Tensor, amount: float = 0.004, *, seed: int = 0, offset: int = 0,
                    coords: Optional[torch.Tensor] = None) -> torch.Tensor:
    """esrgan_addNoise.py:15-25: ceil(amount * size * 0.5) single ELEMENTS (row, col, channel) set to 1, then as many to 0.
    image: CUDA uint8 [H,W,C] or [B,H,W,C]; coords: optional CUDA int32 [B, 2*count, 3], salt draws first."""
    eng = _engine(image.device)
    x = image.contiguous()
    if x.dtype != torch.uint8 or x.dim() not in (3, 4):
        raise RuntimeError("image must be uint8 [H,W,C] or [B,H,W,C]")
    B, H, W, Cc = (1, *x.shape) if x.dim() == 3 else x.shape
    out = torch.empty_like(x)
    cptr, per_image = None, 0
    if coords is not None:
        coords = coords.to(device=x.device, dtype=torch.int32).contiguous()
        if coords.dim() == 2:
            coords = coords[None]
        if coords.dim() != 3 or coords.shape[0] != B or coords.shape[2] != 3:
            raise RuntimeError(f"coords must be int32 [B, 2*count, 3], got {tuple(coords.shape)}")
        cptr, per_image = C.c_void_p(coords.data_ptr()), int(coords.shape[1])
    stream = torch.cuda.current_stream(x.device).cuda_stream
    _lib.check(eng.lib.cid_noise_salt_pepper_unit_u8(eng.handle, C.c_void_p(x.data_ptr()), C.c_void_p(out.data_ptr()), B, H, W,
                                                     Cc, float(amount), cptr, per_image, seed, offset, C.c_void_p(stream)),
               "cid_noise_salt_pepper_unit_u8")
    return out


add_gaussian_unit, add_salt_pepper_unit = add_gaussian, add_salt_pepper  # explicit-domain aliases


def resize_bicubic(image: torch.Tensor, size_wh) -> torch.Tensor:
    """Image.resize(size, BICUBIC) (noise_generation.py:61) on CUDA uint8 [H,W,C] or [B,H,W,C];
    bit-identical to Pillow's fixed-point resampler."""
    eng = _engine(image.device)
    x = image.contiguous()
    if x.dtype != torch.uint8 or x.dim() not in (3, 4):
        raise RuntimeError("image must be uint8 [H,W,C] or [B,H,W,C]")
    squeeze = x.dim() == 3
    if squeeze:
        x = x[None]
    B, H, W, Cc = x.shape
    ow, oh = int(size_wh[0]), int(size_wh[1])
    out = torch.empty((B, oh, ow, Cc), dtype=torch.uint8, device=x.device)
    tmp = torch.empty((B, H, ow, Cc), dtype=torch.uint8, device=x.device)
    stream = torch.cuda.current_stream(x.device).cuda_stream
    _lib.check(eng.lib.cid_resize_bicubic_u8(eng.handle, C.c_void_p(x.data_ptr()), C.c_void_p(out.data_ptr()),
                                             C.c_void_p(tmp.data_ptr()), B, H, W, oh, ow, Cc, C.c_void_p(stream)),
               "cid_resize_bicubic_u8")
    return out[0] if squeeze else out


def normalize_to_tensor(image: torch.Tensor, mean: float = 0.5, std: float = 0.5) -> torch.Tensor:
    """ToTensor + Normalize (denoisegan_eval.py:76-79): CUDA uint8 [B,H,W,3] -> fp32 [B,3,H,W]."""
    eng = _engine(image.device)
    x = image.contiguous()
    B, H, W, Cc = x.shape
    if Cc != 3 or x.dtype != torch.uint8:
        raise RuntimeError("image must be uint8 [B,H,W,3]")
    out = torch.empty((B, 3, H, W), dtype=torch.float32, device=x.device)
    stream = torch.cuda.current_stream(x.device).cuda_stream
    _lib.check(eng.lib.cid_normalize_u8_to_f32_nchw(eng.handle, C.c_void_p(x.data_ptr()), C.c_void_p(out.data_ptr()),
                                                    B, H, W, float(mean), float(std), C.c_void_p(stream)),
               "cid_normalize_u8_to_f32_nchw")
    return out
