"""Drop-in for the reference's Keras cGAN backend: `KERAS_CGAN = tf.keras.models.load_model(path)`
(backend/app.py:338-345) and `y_np = KERAS_CGAN.predict(x_np, verbose=0)` (:458).

`load_model(path)` reads the `.keras` archive without TensorFlow/h5py (keras_h5.py), checks that its
config.json is the generator of backend/trainingcode/cgan_code/training5barrev.py:24-43 and returns
a `KerasCGANGenerator` whose `predict` takes/returns fp32 NHWC numpy arrays exactly like the Keras
model the route calls — but runs the hand-written sm_100a kernels through the C ABI
(`cid_create_model("keras_cgan")`). `.inputs` has one entry, so the route takes its single-input
branch (:446-458). No CPU path: without a B200 and the built library every call raises.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, Optional, Tuple

import numpy as np
import torch

from . import lib as _lib
from .generator import _Engine
from .keras_h5 import read_keras_archive

# Keras variable name -> shape (Keras layouts: Conv2D kernel [kh,kw,Cin,Cout], Conv2DTranspose [kh,kw,Cout,Cin])
KERAS_CGAN_SHAPES: Dict[str, Tuple[int, ...]] = {
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

_EXPECTED_LAYERS = [  # (class_name, checks) of config.json, in order (training5barrev.py:24-43)
    ("InputLayer", {}),
    ("Conv2D", {"filters": 64, "kernel_size": [4, 4], "strides": [2, 2], "padding": "same"}),
    ("LeakyReLU", {"negative_slope": 0.2}),
    ("Conv2D", {"filters": 128, "kernel_size": [4, 4], "strides": [2, 2], "padding": "same"}),
    ("BatchNormalization", {"epsilon": 0.001}),
    ("LeakyReLU", {"negative_slope": 0.2}),
    ("Conv2DTranspose", {"filters": 128, "kernel_size": [4, 4], "strides": [2, 2], "padding": "same"}),
    ("BatchNormalization", {"epsilon": 0.001}),
    ("LeakyReLU", {"negative_slope": 0.2}),
    ("Conv2DTranspose", {"filters": 64, "kernel_size": [4, 4], "strides": [2, 2], "padding": "same"}),
    ("BatchNormalization", {"epsilon": 0.001}),
    ("LeakyReLU", {"negative_slope": 0.2}),
    ("Conv2D", {"filters": 3, "kernel_size": [3, 3], "strides": [1, 1], "padding": "same", "activation": "tanh"}),
]

_VAR_NAMES = {"conv": ("kernel", "bias"), "batch": ("gamma", "beta", "moving_mean", "moving_variance")}


def check_config(config: dict) -> None:
    """Raises unless config.json describes exactly the generator this backend implements."""
    layers = config.get("config", {}).get("layers", [])
    if len(layers) != len(_EXPECTED_LAYERS):
        raise RuntimeError(f"keras cGAN: expected {len(_EXPECTED_LAYERS)} layers, archive has {len(layers)}")
    for i, (layer, (cls, checks)) in enumerate(zip(layers, _EXPECTED_LAYERS)):
        if layer.get("class_name") != cls:
            raise RuntimeError(f"keras cGAN: layer {i} is {layer.get('class_name')}, expected {cls}")
        cfg = layer.get("config", {})
        for k, v in checks.items():
            got = cfg.get(k)
            if isinstance(v, float):
                ok = got is not None and abs(float(got) - v) < 1e-9
            else:
                ok = (list(got) if isinstance(got, (list, tuple)) else got) == v
            if not ok:
                raise RuntimeError(f"keras cGAN: layer {i} ({cls}) has {k}={got!r}, expected {v!r}")


def variables_from_h5(datasets: Dict[str, np.ndarray]) -> Dict[str, np.ndarray]:
    """`layers/<name>/vars/<i>` datasets of model.weights.h5 -> {"<name>.<variable>": array}.
    Keras 3 stores a layer's variables in creation order: Conv kernel, bias; BatchNormalization
    gamma, beta, moving_mean, moving_variance."""
    out: Dict[str, np.ndarray] = {}
    for path, arr in datasets.items():
        parts = path.split("/")
        if len(parts) != 4 or parts[0] != "layers" or parts[2] != "vars":
            continue
        names = _VAR_NAMES["batch" if parts[1].startswith("batch_normalization") else "conv"]
        out[f"{parts[1]}.{names[int(parts[3])]}"] = np.ascontiguousarray(arr, dtype=np.float32)
    missing = [k for k in KERAS_CGAN_SHAPES if k not in out]
    if missing:
        raise RuntimeError(f"keras cGAN: archive lacks variables {missing}")
    for k, shape in KERAS_CGAN_SHAPES.items():
        if tuple(out[k].shape) != shape:
            raise RuntimeError(f"keras cGAN: {k} has shape {tuple(out[k].shape)}, expected {shape}")
    return {k: out[k] for k in KERAS_CGAN_SHAPES}


class KerasCGANGenerator:
    """The object the reference keeps in `KERAS_CGAN`."""

    LAYER_NAMES = ("conv2d", "conv2d_1", "conv2d_transpose", "conv2d_transpose_1", "conv2d_2")

    def __init__(self, variables: Dict[str, np.ndarray], device="cuda"):
        self.variables = {k: torch.from_numpy(np.ascontiguousarray(variables[k], dtype=np.float32))
                          for k in KERAS_CGAN_SHAPES}
        self.device = torch.device(device)
        self.inputs = [("input_layer", (None, 256, 256, 3))]  # one input -> the route's single-input branch
        self._engine: Optional[_Engine] = None
        self._options: Dict[str, int] = {}

    # ---- plumbing ----------------------------------------------------------------------------
    def _eng(self) -> _Engine:
        if self.device.type != "cuda":
            raise RuntimeError("KerasCGANGenerator (B200-native) runs only on CUDA; there is no CPU fallback")
        if self._engine is None:
            idx = self.device.index if self.device.index is not None else torch.cuda.current_device()
            eng = _Engine(idx, "keras_cgan")
            for k, v in self._options.items():
                eng.set_option(k, v)
            eng.upload(self.variables, KERAS_CGAN_SHAPES)
            self._engine = eng
        return self._engine

    def to(self, device) -> "KerasCGANGenerator":
        device = torch.device(device)
        if device != self.device:
            self.device = device
            self._engine = None
        return self

    def set_option(self, name: str, value: int) -> "KerasCGANGenerator":
        self._options[name] = int(value)
        if self._engine is not None:
            self._engine.set_option(name, value)
        return self

    def kernel_launches(self) -> int:
        return self._engine.launches() if self._engine is not None else 0

    def profile_read(self):
        ms, n = self._eng().profile_read()
        return dict(zip(self.LAYER_NAMES, ms[:5])), n

    # ---- Keras surface: y_np = KERAS_CGAN.predict(x_np, verbose=0)  (backend/app.py:458) ------
    def predict(self, x, verbose=0, **_):
        x_np = np.asarray(x, dtype=np.float32)
        if x_np.ndim != 4 or x_np.shape[3] != 3:
            raise ValueError(f"expected input [B,H,W,3], got {x_np.shape}")
        xd = torch.from_numpy(np.ascontiguousarray(x_np)).to(self.device)
        return self.predict_device(xd).cpu().numpy()

    __call__ = predict

    def predict_device(self, x: torch.Tensor) -> torch.Tensor:
        """x: CUDA fp32 NHWC [B,H,W,3] in [-1,1] -> CUDA fp32 NHWC tanh output."""
        if x.dim() != 4 or x.size(3) != 3:
            raise RuntimeError(f"expected input [B,H,W,3], got {tuple(x.shape)}")
        if x.device.type != "cuda":
            raise RuntimeError("KerasCGANGenerator (B200-native) runs only on CUDA tensors; there is no CPU fallback")
        eng = self._eng()
        if x.dtype != torch.float32 or not x.is_contiguous():
            x = x.float().contiguous()
        B, H, W, _ = x.shape
        y = torch.empty_like(x)
        if B == 0:
            return y
        ws = eng.workspace(B, H, W)
        wp, wn = eng.ws_ptr(ws)
        stream = torch.cuda.current_stream(x.device).cuda_stream
        with torch.cuda.device(x.device):
            _lib.check(eng.lib.cid_forward_f32_nhwc(eng.handle, C.c_void_p(x.data_ptr()), C.c_void_p(y.data_ptr()),
                                                    B, H, W, C.c_void_p(wp), wn, C.c_void_p(stream)),
                       "cid_forward_f32_nhwc")
        return y

    def forward_u8(self, x: torch.Tensor, padding: Tuple[int, int, int, int] = (0, 0, 0, 0),
                   out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """The whole keras branch of the route on device: u8 NHWC [B,H,W,3] -> u8 NHWC (pad, /255,
        (x-0.5)/0.5, generator, *0.5+0.5, clip, truncating x255, crop; backend/app.py:438-480)."""
        if x.dim() != 4 or x.size(3) != 3 or x.dtype != torch.uint8:
            raise RuntimeError(f"expected uint8 input [B,H,W,3], got {x.dtype} {tuple(x.shape)}")
        if x.device.type != "cuda":
            raise RuntimeError("KerasCGANGenerator (B200-native) runs only on CUDA tensors; there is no CPU fallback")
        eng = self._eng()
        x = x.contiguous()
        B, H, W, _ = x.shape
        l, t, r, b = (int(v) for v in padding)
        y = out if out is not None else torch.empty_like(x)
        if B == 0:
            return y
        ws = eng.workspace(B, H + t + b, W + l + r)
        wp, wn = eng.ws_ptr(ws)
        stream = torch.cuda.current_stream(x.device).cuda_stream
        with torch.cuda.device(x.device):
            _lib.check(eng.lib.cid_forward_u8_nhwc(eng.handle, C.c_void_p(x.data_ptr()), C.c_void_p(y.data_ptr()),
                                                   B, H, W, l, t, r, b, C.c_void_p(wp), wn, C.c_void_p(stream)),
                       "cid_forward_u8_nhwc")
        return y

    def debug_activation(self, name: str, B: int, H: int, W: int) -> torch.Tensor:
        """Test hook: intermediate activation c0..c3 of the last forward, float32 NHWC on the CPU."""
        eng = self._eng()
        ws = eng.workspace(B, H, W)
        wp, _ = eng.ws_ptr(ws)
        buf = torch.empty(B * H * W * 128, dtype=torch.bfloat16)
        c, h, w = C.c_int(), C.c_int(), C.c_int()
        _lib.check(eng.lib.cid_debug_copy_activation(eng.handle, name.encode(), C.c_void_p(wp), B, H, W,
                                                     C.c_void_p(buf.data_ptr()), buf.numel() * 2, C.byref(c),
                                                     C.byref(h), C.byref(w)), "cid_debug_copy_activation")
        n = B * h.value * w.value * c.value
        return buf[:n].view(B, h.value, w.value, c.value).float()


def load_model(path, compile: bool = False, device="cuda") -> KerasCGANGenerator:  # noqa: A002 (Keras' own argument name)
    """`tf.keras.models.load_model(CGAN_CKPT_KERAS, compile=False)` (backend/app.py:342)."""
    config, datasets = read_keras_archive(path)
    check_config(config)
    return KerasCGANGenerator(variables_from_h5(datasets), device=device)
