"""ORACLE — test infrastructure only, not product code.

CPU restatement (torch fp32 functional ops) of the reference's ESRGAN and SRGAN generators and of the
esrgan / srgan branches of the /enhance route, used by tests/ as the checker.

Pinned: oracle/make_golden_sr.py imports the unmodified /root/reference/backend/app.py in the build
container, loads seeded state_dicts (from `init_state_dict` below) into `app.ESRGANGenerator` /
`app.SRGANGenerator`, runs them in eval mode and commits inputs + outputs to tests/golden/sr_*.npz;
tests/test_sr_oracle.py checks the functions below against those fixtures on every run (and against the
live modules when /root/reference is present).

  ESRGANGenerator   backend/app.py:203-217 (ResidualBlock :189-200)
  SRGANGenerator    backend/app.py:146-184
  esrgan route      backend/app.py:387-397 (preprocess_image :243-249, postprocess_tensor :251-254), crop :473-479
  srgan route       backend/app.py:383-385, 401-406, 422-436, 471-472 (not cropped, :480-485)
"""
from __future__ import annotations

import math
from typing import Dict, Tuple

import numpy as np
import torch
import torch.nn.functional as F

BN_EPS = 1e-5


def _conv_keys(model: str):
    """(prefix, cout, cin, k) of every Conv2d, (prefix, c) of every BatchNorm2d, PReLU keys — forward order."""
    convs, bns, prelus = [("initial.0", 64, 3, 9)], [], ["initial.1.weight"]
    if model == "esrgan":
        for i in range(8):
            b = f"residuals.{i}.block"
            convs += [(f"{b}.0", 64, 64, 3), (f"{b}.3", 64, 64, 3)]
            bns += [(f"{b}.1", 64), (f"{b}.4", 64)]
            prelus.append(f"{b}.2.weight")
        convs.append(("final", 3, 64, 9))
    elif model == "srgan":
        for i in range(5):
            b = f"res_blocks.{i}"
            convs += [(f"{b}.0", 64, 64, 3), (f"{b}.3", 64, 64, 3)]
            bns += [(f"{b}.1", 64), (f"{b}.4", 64)]
            prelus.append(f"{b}.2.weight")
        convs += [("mid", 64, 64, 3), ("upscale.0", 256, 64, 3), ("upscale.3", 256, 64, 3), ("final", 3, 64, 9)]
        prelus += ["upscale.2.weight", "upscale.5.weight"]
    else:
        raise ValueError(model)
    return convs, bns, prelus


def init_state_dict(model: str, seed: int, gain: float = 1.0) -> Dict[str, torch.Tensor]:
    """Seeded state_dict with the reference's keys/shapes: torch's default conv init distribution
    (U(+-1/sqrt(fan_in)), weights scaled by `gain`), NON-trivial BatchNorm affine + running statistics
    (so the eval-mode fold is exercised) and PReLU slopes around the default 0.25. numpy PCG64, so the
    values are machine independent. The statistics are those of a TRAINED net rather than of a freshly
    constructed one: BatchNorm gains around 0.4 keep each residual branch smaller than the stream it is
    added to, and the ESRGAN output layer is scaled/shifted so the image lands inside [0,1] (the domain
    postprocess_tensor clamps to) — with unit gains the 8-block residual chain amplifies the activations
    x20..50 and the un-squashed ESRGAN output leaves any image range (gain > 1 still probes that regime)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    convs, bns, prelus = _conv_keys(model)
    sd: Dict[str, torch.Tensor] = {}
    for p, co, ci, k in convs:
        bound = 1.0 / math.sqrt(ci * k * k)
        sd[p + ".weight"] = torch.from_numpy((rng.uniform(-bound, bound, (co, ci, k, k)) * gain).astype(np.float32))
        sd[p + ".bias"] = torch.from_numpy(rng.uniform(-bound, bound, (co,)).astype(np.float32))
    for p, c in bns:
        sd[p + ".weight"] = torch.from_numpy((rng.uniform(0.6, 1.4, (c,)) * 0.4).astype(np.float32))
        sd[p + ".bias"] = torch.from_numpy(rng.uniform(-0.2, 0.2, (c,)).astype(np.float32))
        sd[p + ".running_mean"] = torch.from_numpy(rng.uniform(-0.2, 0.2, (c,)).astype(np.float32))
        sd[p + ".running_var"] = torch.from_numpy(rng.uniform(0.05, 1.5, (c,)).astype(np.float32))
        sd[p + ".num_batches_tracked"] = torch.tensor(100, dtype=torch.long)
    for p in prelus:
        sd[p] = torch.from_numpy(rng.uniform(0.1, 0.4, (1,)).astype(np.float32))
    if model == "esrgan":
        sd["final.weight"] = sd["final.weight"] * 0.35
        sd["final.bias"] = sd["final.bias"] + 0.5
    return sd


def _bn(t, sd, p):
    return F.batch_norm(t, sd[p + ".running_mean"], sd[p + ".running_var"], sd[p + ".weight"], sd[p + ".bias"],
                        training=False, eps=BN_EPS)


def _conv(t, sd, p, pad):
    return F.conv2d(t, sd[p + ".weight"], sd[p + ".bias"], padding=pad)


def esrgan_forward(sd: Dict[str, torch.Tensor], x: torch.Tensor, upto: int = 99) -> torch.Tensor:
    """ESRGANGenerator.forward (backend/app.py:214-217), eval mode. `upto` = number of residual blocks after
    which the running activation is returned (test hook)."""
    x1 = F.prelu(_conv(x, sd, "initial.0", 4), sd["initial.1.weight"])
    t = x1
    for i in range(8):
        if upto == i:
            return t
        b = f"residuals.{i}.block"
        u = F.prelu(_bn(_conv(t, sd, b + ".0", 1), sd, b + ".1"), sd[b + ".2.weight"])
        t = t + _bn(_conv(u, sd, b + ".3", 1), sd, b + ".4")       # ResidualBlock.forward :199-200
    return _conv(x1 + t, sd, "final", 4)


def srgan_forward(sd: Dict[str, torch.Tensor], x: torch.Tensor) -> torch.Tensor:
    """SRGANGenerator.forward (backend/app.py:179-184), eval mode."""
    x0 = F.prelu(_conv(x, sd, "initial.0", 4), sd["initial.1.weight"])
    t = x0
    for i in range(5):
        b = f"res_blocks.{i}"
        t = F.prelu(_bn(_conv(t, sd, b + ".0", 1), sd, b + ".1"), sd[b + ".2.weight"])
        t = _bn(_conv(t, sd, b + ".3", 1), sd, b + ".4")            # plain Sequential: no per-block skip
    t = _conv(t, sd, "mid", 1) + x0
    for c, a in (("upscale.0", "upscale.2.weight"), ("upscale.3", "upscale.5.weight")):
        t = F.prelu(F.pixel_shuffle(_conv(t, sd, c, 1), 2), sd[a])
    return torch.tanh(_conv(t, sd, "final", 4))


def forward(model: str, sd, x):
    with torch.no_grad():
        return esrgan_forward(sd, x) if model == "esrgan" else srgan_forward(sd, x)


def get_padding(w: int, h: int, divisor: int, scale: int) -> Tuple[int, int, int, int]:
    """get_padding (backend/app.py:276-281): MODEL_CFG esrgan = (4, 1), srgan = (4, 4)."""
    d = divisor * scale
    pw, ph = (d - w % d) % d, (d - h % d) % d
    return pw // 2, ph // 2, pw - pw // 2, ph - ph // 2


def _to_u8(v01: torch.Tensor) -> np.ndarray:
    """ToPILImage on a float CHW tensor: mul(255).byte() — truncation (torchvision functional.to_pil_image)."""
    return v01.mul(255).byte().permute(0, 2, 3, 1).contiguous().numpy()


def esrgan_enhance_u8(sd, img_u8: np.ndarray) -> np.ndarray:
    """esrgan branch of the route, u8 NHWC [B,H,W,3] in/out: ToTensor only on the UN-padded image (:387-390,
    243-249); net; clamp(0,1) + ToPILImage (:251-254); then `y_pil.crop((l, t, l+w, t+h))` with the padding
    get_padding computed for the original size (:383-384, 473-479) — PIL fills the part outside with black."""
    B, H, W, _ = img_u8.shape
    x = torch.from_numpy(img_u8).permute(0, 3, 1, 2).float().div(255)
    y = forward("esrgan", sd, x).clamp(0, 1)
    q = _to_u8(y)
    l, t, _, _ = get_padding(W, H, 4, 1)
    out = np.zeros_like(q)
    out[:, :H - t, :W - l] = q[:, t:, l:]
    return out


def srgan_enhance_u8(sd, img_u8: np.ndarray) -> np.ndarray:
    """srgan branch: Pad(fill=0) to a multiple of 16 (:383-385), ToTensor, Normalize(.5,.5) (:401-406), net,
    (y*0.5+0.5).clamp(0,1) (:435), ToPILImage (:472); the 4x output is NOT cropped (:480-485)."""
    B, H, W, _ = img_u8.shape
    l, t, r, b = get_padding(W, H, 4, 4)
    xp = np.zeros((B, H + t + b, W + l + r, 3), dtype=np.uint8)
    xp[:, t:t + H, l:l + W] = img_u8
    x = torch.from_numpy(xp).permute(0, 3, 1, 2).float().div(255)
    x = (x - 0.5) / 0.5
    y = (forward("srgan", sd, x) * 0.5 + 0.5).clamp(0, 1)
    return _to_u8(y)


def psnr(a: torch.Tensor, b: torch.Tensor, data_range: float = 2.0) -> float:
    mse = float(((a.double() - b.double()) ** 2).mean())
    return 10.0 * math.log10(data_range ** 2 / max(mse, 1e-20))
