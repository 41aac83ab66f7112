"""Drop-in `ESRGANGenerator` and `SRGANGenerator`: the reference's nn.Module surface over the CUDA C-ABI.

Mirrors backend/app.py:146-184 (SRGANGenerator), :189-217 (ResidualBlock, ESRGANGenerator): same class
names and constructor arguments, the same module tree — hence the same `state_dict` keys, including the
BatchNorm buffers — and the same `forward(x)` contract (fp32 NCHW in; ESRGAN: fp32 NCHW out, no
activation, [0,1] domain; SRGAN: tanh output at 4x the size). `PT_MODELS["esrgan"|"srgan"] = ...`,
`load_state_safely` and `net(x)` (backend/app.py:322-324, 393, 433) work unchanged.

The parameters are ordinary nn.Parameters / buffers, but the modules never compute with them in PyTorch:
`forward` hands them to libcid_b200.so, which folds the eval-mode BatchNorm into the convolutions, packs
bf16 tensor-core layouts and runs the hand-written sm_100a kernels. The route always runs these models in
eval mode (`.eval()` at backend/app.py:391, 423); train-mode BatchNorm is not implemented and raises.
No cuDNN, no CPU path: a CPU tensor or a missing library raises.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, Optional, Tuple

import numpy as np
import torch
import torch.nn as nn

from . import lib as _lib
from .generator import _Engine


class _CudaGenerator(nn.Module):
    MODEL = ""          # cid_create_model name
    OUT_SCALE = 1       # spatial scale of the output
    LAYER_NAMES: Tuple[str, ...] = ()

    def _init_engine_state(self):
        self._engines: Dict[int, _Engine] = {}
        self._options: Dict[str, int] = {}

    def _float_state(self) -> Dict[str, torch.Tensor]:
        return {k: v for k, v in self.state_dict().items() if v.dtype.is_floating_point}

    def _signature(self):
        return tuple((k, v.data_ptr(), v._version) for k, v in self.state_dict(keep_vars=True).items()
                     if v.dtype.is_floating_point)

    def _engine(self, device: torch.device) -> _Engine:
        if device.type != "cuda":
            raise RuntimeError(f"{type(self).__name__} (B200-native) runs only on CUDA tensors; there is no CPU fallback")
        if self.training:
            raise RuntimeError(f"{type(self).__name__} (B200-native) implements eval-mode BatchNorm only; call .eval() "
                               "as the route does (backend/app.py:391, 423)")
        idx = device.index if device.index is not None else torch.cuda.current_device()
        eng = self._engines.get(idx)
        if eng is None:
            eng = _Engine(idx, self.MODEL)
            for k, v in self._options.items():
                eng.set_option(k, v)
            self._engines[idx] = eng
        sig = self._signature()
        if eng.signature != sig:
            state = self._float_state()
            eng.upload(state, {k: tuple(v.shape) for k, v in state.items()})
            eng.signature = sig
        return eng

    def set_option(self, name: str, value: int):
        self._options[name] = int(value)
        for eng in self._engines.values():
            eng.set_option(name, value)
        return self

    def kernel_launches(self) -> int:
        return sum(e.launches() for e in self._engines.values())

    def profile_read(self):
        eng = next(iter(self._engines.values()))
        ms, n = eng.profile_read()
        return dict(zip(self.LAYER_NAMES, ms)), n

    # ---- y = net(x) (backend/app.py:393, 433) ----------------------------------------------------
    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if x.dim() != 4 or x.size(1) != 3:
            raise RuntimeError(f"expected input [B,3,H,W], got {tuple(x.shape)}")
        eng = self._engine(x.device)
        if x.dtype != torch.float32 or not x.is_contiguous():
            x = x.float().contiguous()
        B, _, H, W = x.shape
        s = self.OUT_SCALE
        y = torch.empty((B, 3, H * s, W * s), dtype=torch.float32, device=x.device)
        if B == 0 or H == 0 or W == 0:
            return y
        ws = eng.workspace(B, H, W)
        wp, wn = eng.ws_ptr(ws)
        stream = torch.cuda.current_stream(x.device).cuda_stream
        with torch.cuda.device(x.device):
            _lib.check(eng.lib.cid_forward_f32_nchw(eng.handle, C.c_void_p(x.data_ptr()), C.c_void_p(y.data_ptr()),
                                                    B, H, W, C.c_void_p(wp), wn, C.c_void_p(stream)),
                       "cid_forward_f32_nchw")
        return y

    # ---- fused route path: u8 NHWC image(s) in -> u8 NHWC image(s) out -----------------------------
    def forward_u8(self, x: torch.Tensor, padding: Tuple[int, int, int, int] = (0, 0, 0, 0),
                   out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """x: CUDA uint8 [B,H,W,3] (the un-padded upload); padding = get_padding()'s (l,t,r,b) for this model.
        ESRGAN: returns [B,H,W,3] — the route feeds the un-padded image and crops the result at the padding
        offsets (backend/app.py:387-397, 473-479). SRGAN: returns [B,4(H+t+b),4(W+l+r),3], not cropped (:480-485)."""
        if x.dim() != 4 or x.size(3) != 3 or x.dtype != torch.uint8:
            raise RuntimeError(f"expected uint8 input [B,H,W,3], got {x.dtype} {tuple(x.shape)}")
        eng = self._engine(x.device)
        x = x.contiguous()
        B, H, W, _ = x.shape
        l, t, r, b = (int(v) for v in padding)
        s = self.OUT_SCALE
        shape = (B, H, W, 3) if s == 1 else (B, s * (H + t + b), s * (W + l + r), 3)
        y = out if out is not None else torch.empty(shape, dtype=torch.uint8, device=x.device)
        if tuple(y.shape) != shape or y.dtype != torch.uint8 or not y.is_contiguous() or y.device != x.device:
            raise RuntimeError(f"out must be a contiguous uint8 tensor of shape {shape} on the input's device")
        if B == 0:
            return y
        ws = eng.workspace(B, H, W) if s == 1 else eng.workspace(B, H + t + b, W + l + r)
        wp, wn = eng.ws_ptr(ws)
        stream = torch.cuda.current_stream(x.device).cuda_stream
        with torch.cuda.device(x.device):
            _lib.check(eng.lib.cid_forward_u8_nhwc(eng.handle, C.c_void_p(x.data_ptr()), C.c_void_p(y.data_ptr()),
                                                   B, H, W, l, t, r, b, C.c_void_p(wp), wn, C.c_void_p(stream)),
                       "cid_forward_u8_nhwc")
        return y


class ResidualBlock(nn.Module):
    """backend/app.py:189-200 (parameter container only; the arithmetic runs in the CUDA library)."""

    def __init__(self, channels):
        super().__init__()
        self.block = nn.Sequential(
            nn.Conv2d(channels, channels, 3, 1, 1),
            nn.BatchNorm2d(channels),
            nn.PReLU(),
            nn.Conv2d(channels, channels, 3, 1, 1),
            nn.BatchNorm2d(channels),
        )


class ESRGANGenerator(_CudaGenerator):
    MODEL = "esrgan"
    OUT_SCALE = 1
    LAYER_NAMES = ("initial",) + tuple(f"residuals.{i}.block.{j}" for i in range(8) for j in (0, 3)) + ("final",)

    def __init__(self, num_residuals=8):
        super().__init__()
        if num_residuals != 8:
            raise ValueError("the B200-native ESRGANGenerator implements the reference's default of 8 residual blocks")
        self.initial = nn.Sequential(nn.Conv2d(3, 64, 9, 1, 4), nn.PReLU())
        self.residuals = nn.Sequential(*[ResidualBlock(64) for _ in range(num_residuals)])
        self.final = nn.Conv2d(64, 3, 9, 1, 4)
        self._init_engine_state()


class SRGANGenerator(_CudaGenerator):
    MODEL = "srgan"
    OUT_SCALE = 4
    LAYER_NAMES = ("initial",) + tuple(f"res_blocks.{i}.{j}" for i in range(5) for j in (0, 3)) + (
        "mid", "upscale.0", "upscale.3", "final")

    def __init__(self, scale_factor=4):
        super().__init__()
        if scale_factor != 4:
            raise ValueError("the B200-native SRGANGenerator implements the reference's default scale_factor=4")
        self.initial = nn.Sequential(nn.Conv2d(3, 64, kernel_size=9, padding=4), nn.PReLU())
        res_blocks = []
        for _ in range(5):
            res_blocks.append(nn.Sequential(
                nn.Conv2d(64, 64, kernel_size=3, padding=1),
                nn.BatchNorm2d(64),
                nn.PReLU(),
                nn.Conv2d(64, 64, kernel_size=3, padding=1),
                nn.BatchNorm2d(64),
            ))
        self.res_blocks = nn.Sequential(*res_blocks)
        self.mid = nn.Conv2d(64, 64, kernel_size=3, padding=1)
        upscale = []
        for _ in range(int(np.log2(scale_factor))):
            upscale.append(nn.Conv2d(64, 256, kernel_size=3, padding=1))
            upscale.append(nn.PixelShuffle(2))
            upscale.append(nn.PReLU())
        self.upscale = nn.Sequential(*upscale)
        self.final = nn.Conv2d(64, 3, kernel_size=9, padding=4)
        self._init_engine_state()
