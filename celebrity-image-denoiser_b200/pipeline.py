"""Image-in -> image-out entry points of the denoise path (the reference's inference surface).

Reference: the denoise branch of `enhance_image` (backend/app.py:378-385 pad, :401-406
ToTensor/Normalize/.to(DEVICE), :422-436 run_torch_path, :471-480 ToPILImage + crop). Here the pad,
normalise, de-normalise, clamp, truncating quantise and crop are all fused into the first and last
CUDA kernels (cid_forward_u8_nhwc); the host only moves uint8 pixels.
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import numpy as np
import torch

from .generator import DenoiseGenerator
from .loader import MODEL_CFG, get_padding


def enhance_batch_u8(net: DenoiseGenerator, images: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """images: CUDA uint8 [B,H,W,3] (any H, W >= 1). Returns CUDA uint8 [B,H,W,3]."""
    cfg = MODEL_CFG["denoise"]
    H, W = int(images.shape[1]), int(images.shape[2])
    pad = get_padding((W, H), cfg["pad_divisor"], cfg["scale"])
    return net.forward_u8(images, pad, out=out)


class HostPipeline:
    """End-to-end batches from HOST memory: pinned staging, H2D, generator, D2H, double-buffered
    over `depth` CUDA streams so the copies of one sub-batch overlap the kernels of another."""

    def __init__(self, net: DenoiseGenerator, device: torch.device, sub_batch: int, H: int, W: int, depth: int = 2):
        self.net, self.device, self.sub, self.H, self.W, self.depth = net, torch.device(device), sub_batch, H, W, depth
        self.pad = get_padding((W, H), 4, 1)
        self.streams = [torch.cuda.Stream(self.device) for _ in range(depth)]
        shape = (sub_batch, H, W, 3)
        self.d_in = [torch.empty(shape, dtype=torch.uint8, device=self.device) for _ in range(depth)]
        self.d_out = [torch.empty(shape, dtype=torch.uint8, device=self.device) for _ in range(depth)]
        # one generator clone per stream is unnecessary: the workspace is reused strictly in stream
        # order per slot, so each slot gets its own network handle + workspace instead.
        self.nets = [net] + [self._clone(net) for _ in range(depth - 1)]
        self._open = False

    @staticmethod
    def _clone(net: DenoiseGenerator) -> DenoiseGenerator:
        c = DenoiseGenerator()
        c.load_state_dict(net.state_dict(), strict=True)
        p = next(net.parameters())
        c = c.to(p.device).eval()
        for k, v in net._options.items():
            c.set_option(k, v)
        return c

    def run(self, host_in: torch.Tensor, host_out: torch.Tensor, join: bool = True) -> Tuple[int, int]:
        """host_in/host_out: pinned CPU uint8 [B,H,W,3]. Returns (h2d_bytes, d2h_bytes).
        join=False streams batch after batch without re-joining the slot streams in between (each slot's staging
        buffers are reused strictly in its own stream's order); call finish() before reading the last results."""
        B = host_in.shape[0]
        cur = torch.cuda.current_stream(self.device)
        if join or not self._open:
            for s in self.streams:
                s.wait_stream(cur)
            self._open = True
        i = 0
        for b0 in range(0, B, self.sub):
            nb = min(self.sub, B - b0)
            slot = i % self.depth
            with torch.cuda.stream(self.streams[slot]):
                self.d_in[slot][:nb].copy_(host_in[b0:b0 + nb], non_blocking=True)
                self.nets[slot].forward_u8(self.d_in[slot][:nb], self.pad, out=self.d_out[slot][:nb])
                host_out[b0:b0 + nb].copy_(self.d_out[slot][:nb], non_blocking=True)
            i += 1
        if join:
            self.finish()
        n = B * self.H * self.W * 3
        return n, n

    def finish(self) -> None:
        """Make the caller's current stream wait for everything the slot streams have been given."""
        cur = torch.cuda.current_stream(self.device)
        for s in self.streams:
            cur.wait_stream(s)
        self._open = False


def enhance_batch_host(net: DenoiseGenerator, host_images: torch.Tensor, device="cuda", sub_batch: int = 64,
                       host_out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """host_images: CPU uint8 [B,H,W,3] (pinned for full speed). Returns CPU uint8 [B,H,W,3]."""
    B, H, W, _ = host_images.shape
    if host_out is None:
        host_out = torch.empty_like(host_images).pin_memory()
    pipe = HostPipeline(net, torch.device(device), min(sub_batch, B), H, W)
    pipe.run(host_images, host_out)
    torch.cuda.current_stream(torch.device(device)).synchronize()
    return host_out


def enhance_image(net: DenoiseGenerator, image, device="cuda"):
    """PIL RGB image in -> denoised PIL RGB image out (what POST /enhance?model=denoise returns as
    PNG, backend/app.py:378-491, minus the matplotlib graph)."""
    from PIL import Image

    arr = np.asarray(image.convert("RGB"), dtype=np.uint8)
    x = torch.from_numpy(np.ascontiguousarray(arr))[None].to(device)
    y = enhance_batch_u8(net, x)
    return Image.fromarray(y[0].cpu().numpy())
