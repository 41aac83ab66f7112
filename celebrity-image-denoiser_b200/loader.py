"""Model-load entry point and registry hook, same surface as the reference.

Reference: `load_state_safely` (backend/app.py:257-274), the load loop (:319-336), `MODEL_CFG`
(:228-233), `get_padding` (:276-281). Behaviour kept: a checkpoint may be a bare state_dict or a
dict holding it under "generator" / "state_dict" / "G" (the trainer writes "generator",
trainingcode/denoise_gan_code/training.py:360-361); a leading "module." (DataParallel) is
stripped; loading is strict=False; the model is put in eval mode; a failed load is the caller's
to catch (the reference logs a warning and serves random init, :333-336).
"""
from __future__ import annotations

import logging
from typing import Dict, Iterable, Tuple

import torch
import torch.nn as nn

logger = logging.getLogger("gan_api")

MODEL_CFG: Dict[str, Dict] = {  # backend/app.py:228-233
    "denoise": {"normalize": ([0.5] * 3, [0.5] * 3), "activation": "tanh", "pad_divisor": 4, "scale": 1},
    "cgan": {"normalize": ([0.5] * 3, [0.5] * 3), "activation": "tanh", "pad_divisor": 4, "scale": 1},
    "srgan": {"normalize": ([0.5] * 3, [0.5] * 3), "activation": "tanh", "pad_divisor": 4, "scale": 4},
    "esrgan": {"normalize": None, "activation": None, "pad_divisor": 4, "scale": 1},
}


def pick_state(ckpt, key_candidates: Iterable[str] = ("generator", "state_dict", "G")) -> Dict[str, torch.Tensor]:
    """The sub-dict selection + prefix stripping of load_state_safely (backend/app.py:259-271)."""
    state = ckpt
    if isinstance(ckpt, dict):
        for k in key_candidates:
            if k in ckpt and isinstance(ckpt[k], dict):
                state = ckpt[k]
                break
    cleaned = {}
    for k, v in state.items():
        if isinstance(k, str) and k.startswith("module."):
            k = k.replace("module.", "")
        cleaned[k] = v
    return cleaned


def load_state_safely(model: nn.Module, checkpoint_path: str, key_candidates=("generator", "state_dict", "G"),
                      map_location=None) -> None:
    # The trainer's checkpoints also carry optimizer state and python lists
    # (training.py:360-370), so weights_only must be off, as the eval scripts do (denoisegan_eval.py:68).
    ckpt = torch.load(checkpoint_path, map_location=map_location, weights_only=False)
    model.load_state_dict(pick_state(ckpt, key_candidates), strict=False)
    model.eval()
    logger.info(f"Loaded PyTorch weights from {checkpoint_path}")


def get_padding(image_or_size, divisor: int, scale: int = 1) -> Tuple[int, int, int, int]:
    """(left, top, right, bottom) that pads (w, h) up to a multiple of divisor*scale, split with the
    smaller half first. Accepts a PIL image (as the reference does) or a (w, h) tuple."""
    w, h = image_or_size.size if hasattr(image_or_size, "size") and not isinstance(image_or_size, tuple) \
        else image_or_size
    d = divisor * scale
    pw, ph = (d - w % d) % d, (d - h % d) % d
    return (pw // 2, ph // 2, pw - pw // 2, ph - ph // 2)


def patch_app(app_module, device=None, models=("denoise", "esrgan", "srgan")):
    """Drop the B200-native generators in behind the reference's routes: replaces
    `app.PT_MODELS[name]` (backend/app.py:319-324) with the same-named module of this package carrying the
    same weights. The FastAPI routes are untouched (`net(x_pt)` at :393 / :433 now runs the CUDA path).
    Returns the new `PT_MODELS["denoise"]` (or the first swapped model)."""
    from .generator import DenoiseGenerator
    from .sr_generators import ESRGANGenerator, SRGANGenerator

    classes = {"denoise": DenoiseGenerator, "esrgan": ESRGANGenerator, "srgan": SRGANGenerator}
    dev = torch.device(device) if device is not None else torch.device("cuda")
    first = None
    for name in models:
        old = app_module.PT_MODELS[name]
        new = classes[name]()
        new.load_state_dict(old.state_dict(), strict=True)
        new = new.to(dev).eval()
        app_module.PT_MODELS[name] = new
        first = first if first is not None else new
    app_module.DEVICE = dev  # the route moves its input tensor to app.DEVICE (:406)
    return first
