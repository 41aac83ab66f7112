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

from typing import Tuple  # noqa: F401

import numpy as np
import torch
import torch.nn as nn

from .generator import _CudaGenerator


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
    PADS_INPUT = False
    EVAL_ONLY = True
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
    EVAL_ONLY = True
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
