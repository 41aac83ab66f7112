"""ORACLE — test infrastructure only, not product code.

CPU restatement (torch fp32 functional ops) of the Keras cGAN generator that the reference serves
as `POST /enhance?model=cgan&cgan_backend=keras` — the only generator whose TRAINED weights ship
with the reference (backend/weights/cgan_epoch_500.keras).

  architecture   backend/trainingcode/cgan_code/training5barrev.py:24-43  (== config.json inside the archive)
  load           backend/app.py:338-345    tf.keras.models.load_model
  call           backend/app.py:438-467    run_keras_cgan: pad -> /255 -> (x-0.5)/0.5 -> predict -> *0.5+0.5 -> clip
  quantise/crop  backend/app.py:471-480    ToPILImage (truncating x255) + crop

PARITY UNPINNED: TensorFlow / Keras are not installable in the build container or on the GPU boxes,
so this restatement cannot be run against `KERAS_CGAN.predict` itself. It follows the published
Keras/TF semantics (named below) and is anchored by (a) the parameter count and per-layer shapes of
the shipped archive (530,755 = config.json's model), (b) a plausibility check on the shipped trained
weights (tests/test_keras_cgan_oracle.py: a wrong BatchNorm variable order, transposed-conv
orientation or padding rule turns the trained net's output into garbage; the restated net
reproduces its input image to > 20 dB PSNR), and (c) an independent direct-summation restatement
of each layer (`naive_*` below) checked against the torch functional ops on small cases.

Keras/TF semantics restated here (TensorFlow 2.x / Keras 3.10, `channels_last`):
  Conv2D(k4, s2, 'same') on even H,W: total pad = k - s = 2 -> 1 before, 1 after  == torch pad 1;
      kernel variable [kh, kw, Cin, Cout], cross-correlation (no flip).
  Conv2DTranspose(k4, s2, 'same'): tf.nn.conv2d_transpose = gradient of the Conv2D above w.r.t.
      its input: out[2i-1+k] += in[i] * w[k]; output 2H x 2W; kernel variable [kh, kw, Cout, Cin]
      == torch.conv_transpose2d(stride 2, padding 1) with weight.permute(3, 2, 0, 1).
  BatchNormalization inference: gamma * (x - moving_mean) / sqrt(moving_variance + 1e-3) + beta;
      variables in the order gamma, beta, moving_mean, moving_variance (vars/0..3 in the .h5).
  LeakyReLU(negative_slope=0.2); final Conv2D(k3,'same') with tanh.
"""
from __future__ import annotations

from typing import Dict, Tuple

import numpy as np
import torch
import torch.nn.functional as F

BN_EPS = 1e-3
SLOPE = 0.2

# variable name -> shape, in Keras' own layouts (what cid_set_param / KerasCGANGenerator accept)
PARAM_SHAPES: Dict[str, Tuple[int, ...]] = {
    "conv2d.kernel": (4, 4, 3, 64), "conv2d.bias": (64,),
    "conv2d_1.kernel": (4, 4, 64, 128), "conv2d_1.bias": (128,),
    "batch_normalization.gamma": (128,), "batch_normalization.beta": (128,),
    "batch_normalization.moving_mean": (128,), "batch_normalization.moving_variance": (128,),
    "conv2d_transpose.kernel": (4, 4, 128, 128), "conv2d_transpose.bias": (128,),
    "batch_normalization_1.gamma": (128,), "batch_normalization_1.beta": (128,),
    "batch_normalization_1.moving_mean": (128,), "batch_normalization_1.moving_variance": (128,),
    "conv2d_transpose_1.kernel": (4, 4, 64, 128), "conv2d_transpose_1.bias": (64,),
    "batch_normalization_2.gamma": (64,), "batch_normalization_2.beta": (64,),
    "batch_normalization_2.moving_mean": (64,), "batch_normalization_2.moving_variance": (64,),
    "conv2d_2.kernel": (3, 3, 64, 3), "conv2d_2.bias": (3,),
}


def init_params(seed: int, gain: float = 1.0) -> Dict[str, np.ndarray]:
    """Seeded synthetic variables with the archive's names/shapes (glorot-uniform kernels scaled by
    `gain`, small biases, non-trivial BatchNorm statistics so the fold is exercised)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    out: Dict[str, np.ndarray] = {}
    for name, shape in PARAM_SHAPES.items():
        kind = name.rsplit(".", 1)[1]
        if kind == "kernel":
            fan = shape[0] * shape[1] * (shape[2] + shape[3])
            lim = gain * np.sqrt(6.0 / fan)
            v = rng.uniform(-lim, lim, size=shape)
        elif kind == "bias" or kind == "beta":
            v = rng.uniform(-0.1, 0.1, size=shape)
        elif kind == "gamma":
            v = rng.uniform(0.5, 1.5, size=shape)
        elif kind == "moving_mean":
            v = rng.uniform(-0.3, 0.3, size=shape)
        else:  # moving_variance
            v = rng.uniform(0.05, 2.0, size=shape)
        out[name] = v.astype(np.float32)
    return out


def _bn(x, p, name):
    g, b = torch.from_numpy(p[name + ".gamma"]), torch.from_numpy(p[name + ".beta"])
    m, v = torch.from_numpy(p[name + ".moving_mean"]), torch.from_numpy(p[name + ".moving_variance"])
    return F.batch_norm(x, m, v, g, b, training=False, eps=BN_EPS)


def _conv_w(p, name):
    return torch.from_numpy(p[name + ".kernel"]).permute(3, 2, 0, 1).contiguous()  # [Cout,Cin,kh,kw]


def _convt_w(p, name):
    return torch.from_numpy(p[name + ".kernel"]).permute(3, 2, 0, 1).contiguous()  # [Cin,Cout,kh,kw]


def forward_nchw(p: Dict[str, np.ndarray], x: torch.Tensor, upto: int = 5) -> torch.Tensor:
    """x: fp32 NCHW in [-1,1], H and W multiples of 4. `upto` stops after that many layers."""
    t = F.leaky_relu(F.conv2d(x, _conv_w(p, "conv2d"), torch.from_numpy(p["conv2d.bias"]), stride=2, padding=1), SLOPE)
    if upto == 1:
        return t
    t = F.conv2d(t, _conv_w(p, "conv2d_1"), torch.from_numpy(p["conv2d_1.bias"]), stride=2, padding=1)
    t = F.leaky_relu(_bn(t, p, "batch_normalization"), SLOPE)
    if upto == 2:
        return t
    t = F.conv_transpose2d(t, _convt_w(p, "conv2d_transpose"), torch.from_numpy(p["conv2d_transpose.bias"]),
                           stride=2, padding=1)
    t = F.leaky_relu(_bn(t, p, "batch_normalization_1"), SLOPE)
    if upto == 3:
        return t
    t = F.conv_transpose2d(t, _convt_w(p, "conv2d_transpose_1"), torch.from_numpy(p["conv2d_transpose_1.bias"]),
                           stride=2, padding=1)
    t = F.leaky_relu(_bn(t, p, "batch_normalization_2"), SLOPE)
    if upto == 4:
        return t
    t = F.conv2d(t, _conv_w(p, "conv2d_2"), torch.from_numpy(p["conv2d_2.bias"]), stride=1, padding=1)
    return torch.tanh(t)


def predict(p: Dict[str, np.ndarray], x_nhwc: np.ndarray) -> np.ndarray:
    """`KERAS_CGAN.predict(x_np)` (backend/app.py:458): fp32 NHWC [-1,1] in, fp32 NHWC tanh out."""
    x = torch.from_numpy(np.ascontiguousarray(x_nhwc, dtype=np.float32)).permute(0, 3, 1, 2).contiguous()
    with torch.no_grad():
        y = forward_nchw(p, x)
    return y.permute(0, 2, 3, 1).contiguous().numpy()


def get_padding(w: int, h: int, divisor: int = 4) -> Tuple[int, int, int, int]:
    """backend/app.py:276-281 with MODEL_CFG['cgan'] (pad_divisor 4, scale 1)."""
    pw = (divisor - w % divisor) % divisor
    ph = (divisor - h % divisor) % divisor
    return pw // 2, ph // 2, pw - pw // 2, ph - ph // 2


def enhance_u8(p: Dict[str, np.ndarray], img_u8: np.ndarray) -> np.ndarray:
    """The keras-cgan branch of the route for a batch of equally sized images, u8 NHWC in/out:
    Pad(fill=0) (app.py:383-385) -> /255 -> (x-0.5)/0.5 (:442-444) -> predict (:458) -> *0.5+0.5,
    clip (:462-464) -> ToPILImage truncating x255 (:472) -> crop (:474-480)."""
    B, H, W, _ = img_u8.shape
    l, t, r, b = get_padding(W, H)
    xp = np.zeros((B, H + t + b, W + l + r, 3), dtype=np.uint8)
    xp[:, t:t + H, l:l + W] = img_u8
    x = xp.astype(np.float32) / 255.0
    x = (x - 0.5) / 0.5
    y = predict(p, x)
    y = np.clip(y * 0.5 + 0.5, 0.0, 1.0).astype(np.float32)
    q = torch.from_numpy(y).mul(255).byte().numpy()  # torchvision to_pil_image: mul(255).byte() (truncation)
    return np.ascontiguousarray(q[:, t:t + H, l:l + W])


# --------------------------------------------------------------------------------------------
# independent direct-summation restatements (small cases only) used to cross-check the layer
# semantics above without going through torch's conv code
# --------------------------------------------------------------------------------------------
def naive_conv_k4s2_same(x_hwc: np.ndarray, kernel: np.ndarray, bias: np.ndarray) -> np.ndarray:
    """TF Conv2D(k4, s2, 'same') on one HWC image, even H,W: out[y,x] = sum in[2y-1+ky, 2x-1+kx] * k[ky,kx]."""
    H, W, _ = x_hwc.shape
    co = kernel.shape[3]
    xp = np.zeros((H + 2, W + 2, x_hwc.shape[2]), dtype=np.float64)
    xp[1:H + 1, 1:W + 1] = x_hwc
    out = np.zeros((H // 2, W // 2, co), dtype=np.float64)
    for y in range(H // 2):
        for x in range(W // 2):
            win = xp[2 * y:2 * y + 4, 2 * x:2 * x + 4]  # rows 2y-1.. in un-padded coordinates
            out[y, x] = np.tensordot(win, kernel.astype(np.float64), axes=([0, 1, 2], [0, 1, 2])) + bias
    return out


def naive_convt_k4s2_same(x_hwc: np.ndarray, kernel: np.ndarray, bias: np.ndarray) -> np.ndarray:
    """TF Conv2DTranspose(k4, s2, 'same') on one HWC image: scatter form of the conv gradient,
    out[2i-1+ky, 2j-1+kx, co] += in[i,j,ci] * k[ky,kx,co,ci], cropped to 2H x 2W."""
    H, W, _ = x_hwc.shape
    co = kernel.shape[2]
    full = np.zeros((2 * H + 2, 2 * W + 2, co), dtype=np.float64)  # index o+1
    k = kernel.astype(np.float64)
    for i in range(H):
        for j in range(W):
            contrib = np.tensordot(k, x_hwc[i, j].astype(np.float64), axes=([3], [0]))  # [4,4,co]
            full[2 * i:2 * i + 4, 2 * j:2 * j + 4] += contrib
    return full[1:2 * H + 1, 1:2 * W + 1] + bias
