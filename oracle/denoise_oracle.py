"""ORACLE — test infrastructure only, not product code.

CPU restatement (torch fp32 functional ops + numpy) of the reference's denoise hot path, used by
tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs as the
checker and the timed CPU baseline. Nothing in the product package imports this module, and the
product has no CPU path of its own.

Pinned: the reference ships no tests or golden vectors, so this restatement
is pinned against the reference ITSELF: oracle/make_golden.py imports the unmodified
/root/reference/backend/app.py in the build container, runs `app.DenoiseGenerator` /
`app.get_padding` / the torchvision transform chain / noise_generation.py / PIL's resize on
seeded inputs and commits the outputs to tests/golden/*.npz; tests/test_oracle_golden.py checks
every function below against those fixtures on every run.

Each function cites the reference lines it follows (paths relative to /root/reference).
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Tuple

import numpy as np
import torch
import torch.nn.functional as F

# state_dict keys / shapes of DenoiseGenerator (backend/app.py:42-78)
_CONVS = [
    ("down1.0", 3, 64), ("down1.2", 64, 64), ("down2.0", 64, 128), ("down2.2", 128, 128),
    ("bottleneck.0", 128, 256), ("bottleneck.2", 256, 256),
    ("upconv2.0", 256, 128), ("upconv2.2", 128, 128), ("upconv1.0", 128, 64), ("upconv1.2", 64, 3),
]
_CONVTS = [("up2", 256, 128), ("up1", 128, 64)]


def state_dict_shapes() -> Dict[str, Tuple[int, ...]]:
    shapes: Dict[str, Tuple[int, ...]] = {}
    for name, ci, co in _CONVS:
        shapes[name + ".weight"] = (co, ci, 3, 3)
        shapes[name + ".bias"] = (co,)
    for name, ci, co in _CONVTS:
        shapes[name + ".weight"] = (ci, co, 2, 2)
        shapes[name + ".bias"] = (co,)
    return shapes


def init_state_dict(seed: int, gain: float = 1.0) -> Dict[str, torch.Tensor]:
    """Seeded random state_dict with torch's default Conv init distribution
    (kaiming_uniform(a=sqrt(5)) == U(-1/sqrt(fan_in), 1/sqrt(fan_in)) for weight and bias; for
    ConvTranspose2d torch computes fan_in from dim 1, i.e. Cout*kh*kw). Drawn with numpy's PCG64 so
    the values are identical on every machine; `gain` scales the weights (not the biases) to
    drive tanh towards saturation (SURVEY.md §7 "amplified" variant)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    sd: Dict[str, torch.Tensor] = {}
    for key, shape in sorted(state_dict_shapes().items()):
        base = key.rsplit(".", 1)[0]
        wshape = state_dict_shapes()[base + ".weight"]
        fan_in = wshape[1] * wshape[2] * wshape[3]
        bound = 1.0 / math.sqrt(fan_in)
        v = rng.uniform(-bound, bound, size=shape).astype(np.float32)
        if key.endswith(".weight"):
            v = (v * np.float32(gain)).astype(np.float32)
        sd[key] = torch.from_numpy(v)
    return sd


def init_state_dict_trained_like(seed: int, mid_gain: float = 1.25, last_gain: float = 24.0) -> Dict[str, torch.Tensor]:
    """A state_dict with the statistics of a TRAINED denoiser rather than of a random net: moderate gains in the
    body (no chaotic amplification of round-off) and a strong last layer, so the tanh output spans the whole
    [-1, 1] range with a few per cent of the pixels saturated. The north-star tolerances (max-abs 2e-2, +-1 LSB) are
    asserted on it unrelaxed: the reference's shipped .pth is not in the tree."""
    sd = init_state_dict(seed, mid_gain)
    sd["upconv1.2.weight"] = sd["upconv1.2.weight"] * np.float32(last_gain / mid_gain)
    return sd


def forward(sd: Dict[str, torch.Tensor], x: torch.Tensor, intermediates: Optional[dict] = None) -> torch.Tensor:
    """DenoiseGenerator.forward (backend/app.py:80-103), fp32 on CPU.
    x: [B,3,H,W] in [-1,1]. If `intermediates` is a dict it receives the NCHW activations named
    like the CUDA workspace buffers (a0, e1, p1, a2, e2, p2, b0, b1, d2, u2a, u2b, d1, u1a)."""
    def conv(t, name, relu=True):
        t = F.conv2d(t, sd[name + ".weight"], sd[name + ".bias"], padding=1)
        return F.relu(t) if relu else t

    keep = intermediates if intermediates is not None else {}
    a0 = conv(x, "down1.0")                                    # :43-44
    e1 = conv(a0, "down1.2")                                   # :45-46
    p1 = F.max_pool2d(e1, 2, 2)                                # :48, :82
    a2 = conv(p1, "down2.0")                                   # :51-52
    e2 = conv(a2, "down2.2")                                   # :53-54
    p2 = F.max_pool2d(e2, 2, 2)                                # :56, :85
    b0 = conv(p2, "bottleneck.0")                              # :59-60
    b1 = conv(b0, "bottleneck.2")                              # :61-62
    d2 = F.conv_transpose2d(b1, sd["up2.weight"], sd["up2.bias"], stride=2)    # :65, :89
    if d2.shape != e2.shape:                                   # defensive crop :90-92
        e2 = e2[:, :, : d2.shape[2], : d2.shape[3]]
    u2a = conv(torch.cat([d2, e2], dim=1), "upconv2.0")        # :93, :67-68
    u2b = conv(u2a, "upconv2.2")                               # :69-70
    d1 = F.conv_transpose2d(u2b, sd["up1.weight"], sd["up1.bias"], stride=2)   # :73, :96
    if d1.shape != e1.shape:                                   # :97-99
        e1 = e1[:, :, : d1.shape[2], : d1.shape[3]]
    u1a = conv(torch.cat([d1, e1], dim=1), "upconv1.0")        # :100, :75-76
    out = torch.tanh(conv(u1a, "upconv1.2", relu=False))       # :77, :103
    keep.update(a0=a0, e1=e1, p1=p1, a2=a2, e2=e2, p2=p2, b0=b0, b1=b1, d2=d2, u2a=u2a, u2b=u2b, d1=d1, u1a=u1a)
    return out


def _bf16(t: torch.Tensor) -> torch.Tensor:
    return t.to(torch.bfloat16).to(torch.float32)


def forward_bf16_storage(sd: Dict[str, torch.Tensor], x: torch.Tensor) -> torch.Tensor:
    """The same network (backend/app.py:80-103) with the STORAGE precision of the CUDA path modelled on
    the CPU: weights rounded to bf16, every stored activation rounded to bf16, fp32 accumulation and
    fp32 bias/ReLU/tanh, the network input kept as a hi+lo bf16 pair (~16 mantissa bits). It is a
    second checker, not a second reference: tests use it to tell a kernel bug from bf16 round-off
    when amplified weights (gain >= 3) make the net chaotic enough that bf16 storage alone exceeds the
    north-star tolerance against the fp32 reference."""
    w = {k: (_bf16(v) if k.endswith(".weight") else v) for k, v in sd.items()}
    w0 = sd["down1.0.weight"]
    w["down1.0.weight"] = _bf16(w0) + _bf16(w0 - _bf16(w0))  # first layer: weights kept as a hi+lo pair too

    def conv(t, name, relu=True, store=True):
        t = F.conv2d(t, w[name + ".weight"], w[name + ".bias"], padding=1)
        if relu:
            t = F.relu(t)
        return _bf16(t) if store else t

    hi = _bf16(x)
    xin = hi + _bf16(x - hi)
    a0 = conv(xin, "down1.0")
    e1 = conv(a0, "down1.2")
    p1 = F.max_pool2d(e1, 2, 2)
    a2 = conv(p1, "down2.0")
    e2 = conv(a2, "down2.2")
    p2 = F.max_pool2d(e2, 2, 2)
    b0 = conv(p2, "bottleneck.0")
    b1 = conv(b0, "bottleneck.2")
    d2 = _bf16(F.conv_transpose2d(b1, w["up2.weight"], w["up2.bias"], stride=2))
    u2a = conv(torch.cat([d2, e2], dim=1), "upconv2.0")
    u2b = conv(u2a, "upconv2.2")
    d1 = _bf16(F.conv_transpose2d(u2b, w["up1.weight"], w["up1.bias"], stride=2))
    u1a = conv(torch.cat([d1, e1], dim=1), "upconv1.0")
    return torch.tanh(conv(u1a, "upconv1.2", relu=False, store=False))


def get_padding(w: int, h: int, divisor: int = 4, scale: int = 1) -> Tuple[int, int, int, int]:
    """get_padding (backend/app.py:276-281) on (width, height): (left, top, right, bottom)."""
    d = divisor * scale
    pw = (d - w % d) % d
    ph = (d - h % d) % d
    return (pw // 2, ph // 2, pw - pw // 2, ph - ph // 2)


def preprocess_u8(img: np.ndarray, padding: Tuple[int, int, int, int]) -> torch.Tensor:
    """Pad(fill=0) -> ToTensor -> Normalize(0.5, 0.5) -> unsqueeze(0) (backend/app.py:383-385, 401-406).
    img: u8 [H,W,3] or [B,H,W,3]. Returns fp32 [B,3,H',W']."""
    a = img if img.ndim == 4 else img[None]
    l, t, r, b = padding
    a = np.pad(a, ((0, 0), (t, b), (l, r), (0, 0)), mode="constant", constant_values=0)
    x = torch.from_numpy(np.ascontiguousarray(a)).permute(0, 3, 1, 2).contiguous()
    x = x.to(torch.float32).div(255)          # ToTensor
    x = x.sub(0.5).div(0.5)                   # Normalize(mean=.5, std=.5)
    return x


def postprocess_u8(y: torch.Tensor, padding: Tuple[int, int, int, int], size_wh: Tuple[int, int]) -> np.ndarray:
    """(y*0.5+0.5).clamp(0,1) (backend/app.py:435) -> ToPILImage: mul(255) then a TRUNCATING
    uint8 cast (torchvision functional.to_pil_image; :472) -> crop to the original size (:474-480).
    Returns u8 [B,H,W,3]."""
    v = (y * 0.5 + 0.5).clamp(0, 1)
    u = v.mul(255).byte()                     # truncation, as to_pil_image does
    l, t, _, _ = padding
    w, h = size_wh
    u = u[:, :, t:t + h, l:l + w]
    return u.permute(0, 2, 3, 1).contiguous().numpy()


def enhance_u8(sd: Dict[str, torch.Tensor], img: np.ndarray) -> np.ndarray:
    """The denoise branch of POST /enhance between PIL decode and PNG encode
    (backend/app.py:378-385, 400-406, 422-436, 471-480). img: u8 [H,W,3] or [B,H,W,3]."""
    a = img if img.ndim == 4 else img[None]
    h, w = a.shape[1], a.shape[2]
    pad = get_padding(w, h, 4, 1)
    with torch.no_grad():
        y = forward(sd, preprocess_u8(a, pad))
    return postprocess_u8(y, pad, (w, h))


def psnr(a: torch.Tensor, b: torch.Tensor, data_range: float = 2.0) -> float:
    """PSNR with data_range=2.0, the convention of denoise_gan_code/training.py:382."""
    mse = float(((a.double() - b.double()) ** 2).mean())
    return float("inf") if mse == 0 else 10.0 * math.log10(data_range ** 2 / mse)


# --------------------------------------------------------------------------------------------
# noiser (trainingcode/denoise_gan_code/noise_generation.py)
# --------------------------------------------------------------------------------------------
def gaussian_noise_apply(image: np.ndarray, noise: np.ndarray) -> np.ndarray:
    """add_gaussian_noise (:6-10) with the normal draw made explicit:
    np.clip(image + noise, 0, 255).astype(np.uint8), float64, truncating."""
    return np.clip(image + noise.astype(np.float64), 0, 255).astype(np.uint8)


def speckle_noise_apply(image: np.ndarray, noise: np.ndarray) -> np.ndarray:
    """add_speckle_noise (:24-28) with the normal draws made explicit: clip(x + x * noise, 0, 255).astype(u8)."""
    return np.clip(image + image * noise, 0, 255).astype(np.uint8)


def uniform_noise_apply(image: np.ndarray, noise: np.ndarray) -> np.ndarray:
    """add_uniform_noise (:35-39) with the uniform draws made explicit."""
    return np.clip(image + noise, 0, 255).astype(np.uint8)


def poisson_noise_apply(counts: np.ndarray) -> np.ndarray:
    """add_poisson_noise (:30-33): np.random.poisson(image).astype(np.uint8) — the int64 counts WRAP mod 256."""
    return counts.astype(np.uint8)


def salt_pepper_counts(shape_hwc: Tuple[int, int, int], salt_prob: float, pepper_prob: float) -> Tuple[int, int]:
    """num_salt / num_pepper = int(image_array.size * prob) (:15-17) — size counts channels."""
    total = shape_hwc[0] * shape_hwc[1] * shape_hwc[2]
    return int(total * salt_prob), int(total * pepper_prob)


def salt_pepper_apply(image: np.ndarray, salt_rc: np.ndarray, pepper_rc: np.ndarray) -> np.ndarray:
    """add_salt_pepper_noise (:12-22) with the coordinate draws made explicit: salt_rc / pepper_rc
    are int arrays [n,2] of (row, col); salt (255 on all channels) is written first, pepper (0) second."""
    out = image.copy()
    out[salt_rc[:, 0], salt_rc[:, 1], :] = 255
    out[pepper_rc[:, 0], pepper_rc[:, 1], :] = 0
    return out


# --------------------------------------------------------------------------------------------
# [0,1]-domain noiser of the ESRGAN data-prep script (backend/trainingcode/esrgan_code/esrgan_addNoise.py), with the
# random draws made explicit and the script's own conversions (:66 img = u8 / 255.0, :79 (noisy * 255).astype(uint8))
# --------------------------------------------------------------------------------------------
def unit_noise_apply(image_u8: np.ndarray, kind: str, draws: np.ndarray) -> np.ndarray:
    """add_gaussian (:7-13) / add_speckle (:27-30) / add_uniform (:38-41) given the function's own random array."""
    img = image_u8 / 255.0
    noisy = img + img * draws if kind == "speckle" else img + draws
    return (np.clip(noisy, 0, 1) * 255).astype(np.uint8)


def unit_salt_pepper_count(size: int, amount: float = 0.004) -> int:
    """np.ceil(amount * img.size * 0.5) (:17, :21), the same number for salt and for pepper."""
    return int(np.ceil(amount * size * 0.5))


def unit_salt_pepper_apply(image_u8: np.ndarray, salt_idx: np.ndarray, pepper_idx: np.ndarray) -> np.ndarray:
    """add_salt_pepper (:15-25): salt_idx / pepper_idx are int arrays [n,3] of (row, col, channel) ELEMENT coordinates."""
    out = image_u8 / 255.0
    out[salt_idx[:, 0], salt_idx[:, 1], salt_idx[:, 2]] = 1
    out[pepper_idx[:, 0], pepper_idx[:, 1], pepper_idx[:, 2]] = 0
    return (out * 255).astype(np.uint8)


# --------------------------------------------------------------------------------------------
# PIL bicubic resize (Pillow src/libImaging/Resample.c; called at noise_generation.py:61)
# --------------------------------------------------------------------------------------------
_PRECISION_BITS = 32 - 8 - 2


def _bicubic(x: float) -> float:
    a = -0.5
    x = abs(x)
    if x < 1.0:
        return ((a + 2.0) * x - (a + 3.0)) * x * x + 1
    if x < 2.0:
        return (((x - 5) * x + 8) * x - 4) * a
    return 0.0


def _coeffs(in_size: int, out_size: int) -> Tuple[List[Tuple[int, int]], np.ndarray]:
    """precompute_coeffs + normalize_coeffs_8bpc for the full-image box."""
    scale = float(np.float32(in_size) - np.float32(0)) / out_size
    filterscale = max(scale, 1.0)
    support = 2.0 * filterscale
    ksize = int(math.ceil(support)) * 2 + 1
    bounds = []
    kk = np.zeros((out_size, ksize), dtype=np.int64)
    for xx in range(out_size):
        center = (xx + 0.5) * scale
        ss = 1.0 / filterscale
        xmin = max(int(center - support + 0.5), 0)
        xmax = min(int(center + support + 0.5), in_size) - xmin
        k = [_bicubic((x + xmin - center + 0.5) * ss) for x in range(xmax)]
        ww = sum(k)
        if ww != 0.0:
            k = [v / ww for v in k]
        for x, v in enumerate(k):
            s = v * (1 << _PRECISION_BITS)
            kk[xx, x] = int(-0.5 + s) if v < 0 else int(0.5 + s)
        bounds.append((xmin, xmax))
    return bounds, kk


def _resample_axis(a: np.ndarray, out_size: int, axis: int) -> np.ndarray:
    in_size = a.shape[axis]
    bounds, kk = _coeffs(in_size, out_size)
    a = np.moveaxis(a, axis, 0).astype(np.int64)
    out = np.empty((out_size,) + a.shape[1:], dtype=np.uint8)
    for xx, (lo, cnt) in enumerate(bounds):
        w = kk[xx, :cnt].reshape((cnt,) + (1,) * (a.ndim - 1))
        ss = (1 << (_PRECISION_BITS - 1)) + (a[lo:lo + cnt] * w).sum(axis=0)
        out[xx] = np.clip(ss >> _PRECISION_BITS, 0, 255).astype(np.uint8)
    return np.moveaxis(out, 0, axis)


def resize_bicubic_u8(img: np.ndarray, size_wh: Tuple[int, int]) -> np.ndarray:
    """Image.resize(size, BICUBIC) on a u8 [H,W,C] array: horizontal pass, then vertical pass,
    22-bit fixed point with round-half-up and clip8."""
    w, h = size_wh
    tmp = _resample_axis(img, w, axis=1)
    return _resample_axis(tmp, h, axis=0)
