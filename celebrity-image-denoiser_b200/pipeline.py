"""Image-in -> image-out entry points of the denoise path (the reference's inference surface).

Reference: the denoise branch of `enhance_image` (backend/app.py:378-385 pad, :401-406
ToTensor/Normalize/.to(DEVICE), :422-436 run_torch_path, :471-480 ToPILImage + crop). Here the pad,
normalise, de-normalise, clamp, truncating quantise and crop are all fused into the first and last
CUDA kernels (cid_forward_u8_nhwc); the host only moves uint8 pixels.
"""
from __future__ import annotations

import os
from typing import Optional, Tuple

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


def set_rank_affinity(local_rank: int, local_world: int, device_index: Optional[int] = None) -> Optional[list]:
    """Give this rank's host threads their own slice of the CPUs next to its GPU, so the pinned staging
    buffers are first-touched on that NUMA node and the launch threads of the ranks of one box do not pile
    up on the same cores. The CPU set comes from NVML's per-GPU affinity mask (the whole allowed set when
    NVML is unavailable); ranks that share a set split it evenly. Returns the chosen CPUs (None if the
    platform has no sched_setaffinity)."""
    if not hasattr(os, "sched_setaffinity"):
        return None
    allowed = sorted(os.sched_getaffinity(0))
    near = allowed
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local_rank if device_index is None else device_index)
        words = (len(os.sched_getaffinity(0)) + 63) // 64 + 8
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        cpus = [w * 64 + b for w, m in enumerate(mask) for b in range(64) if (m >> b) & 1]
        cpus = [c for c in cpus if c in set(allowed)]
        if cpus:
            near = cpus
    except Exception:
        pass
    if local_world > 1 and len(near) >= 2 * local_world:
        per = len(near) // local_world
        near = near[local_rank * per:(local_rank + 1) * per]
    try:
        os.sched_setaffinity(0, near)
    except OSError:
        return None
    return near


class HostPipeline:
    """End-to-end batches from HOST memory as a three-stage pipeline: pinned host uint8 -> H2D (copy
    stream) -> generator (compute stream) -> D2H (second copy stream) -> pinned host uint8, over
    `depth` staging slots. The three stages of different sub-batches overlap (the two copy directions
    use the GPU's two copy engines); one sub-batch's stages are ordered by events. All kernels run on
    ONE compute stream, so a single generator handle / workspace serves every slot and consecutive
    sub-batches keep the programmatic-dependent-launch chain of the layers intact.

    One cudaMemcpyAsync per sub-batch and direction; stage = "all" | "copy" | "compute" selects what
    `run` really does (the last two are the copy-only / compute-only diagnostics of bench.py)."""

    def __init__(self, net: DenoiseGenerator, device: torch.device, sub_batch: int, H: int, W: int, depth: int = 3):
        self.net, self.device, self.sub, self.H, self.W = net, torch.device(device), sub_batch, H, W
        self.depth = max(int(depth), 1)
        self.pad = get_padding((W, H), 4, 1)
        self.s_in = torch.cuda.Stream(self.device)
        self.s_run = torch.cuda.Stream(self.device)
        self.s_out = torch.cuda.Stream(self.device)
        shape = (sub_batch, H, W, 3)
        self.d_in = [torch.empty(shape, dtype=torch.uint8, device=self.device) for _ in range(self.depth)]
        self.d_out = [torch.empty(shape, dtype=torch.uint8, device=self.device) for _ in range(self.depth)]
        mk = lambda: [torch.cuda.Event() for _ in range(self.depth)]  # noqa: E731
        self.in_ready, self.run_done, self.out_done = mk(), mk(), mk()
        self._used = [False] * self.depth
        self._open = False
        self._i = 0

    def run(self, host_in: torch.Tensor, host_out: torch.Tensor, join: bool = True, stage: str = "all") -> Tuple[int, int]:
        """host_in/host_out: pinned CPU uint8 [B,H,W,3]. Returns (h2d_bytes, d2h_bytes).
        join=False streams batch after batch without re-joining the pipeline in between (each slot's staging
        buffers are reused strictly in event order); call finish() before reading the last results."""
        B = host_in.shape[0]
        if B == 0:
            return 0, 0
        do_copy, do_run = stage in ("all", "copy"), stage in ("all", "compute")
        cur = torch.cuda.current_stream(self.device)
        if join or not self._open:
            for s in (self.s_in, self.s_run, self.s_out):
                s.wait_stream(cur)
            self._open = True
        for b0 in range(0, B, self.sub):
            nb = min(self.sub, B - b0)
            k = self._i % self.depth
            self._i += 1
            with torch.cuda.stream(self.s_in):
                if self._used[k]:
                    self.s_in.wait_event(self.run_done[k])   # the kernels that read d_in[k] last time are done
                if do_copy:
                    self.d_in[k][:nb].copy_(host_in[b0:b0 + nb], non_blocking=True)
                self.in_ready[k].record(self.s_in)
            with torch.cuda.stream(self.s_run):
                self.s_run.wait_event(self.in_ready[k])
                if self._used[k]:
                    self.s_run.wait_event(self.out_done[k])  # d_out[k]'s previous result has left the device
                if do_run:
                    self.net.forward_u8(self.d_in[k][:nb], self.pad, out=self.d_out[k][:nb])
                self.run_done[k].record(self.s_run)
            with torch.cuda.stream(self.s_out):
                self.s_out.wait_event(self.run_done[k])
                if do_copy:
                    host_out[b0:b0 + nb].copy_(self.d_out[k][:nb], non_blocking=True)
                self.out_done[k].record(self.s_out)
            self._used[k] = True
        if join:
            self.finish()
        n = B * self.H * self.W * 3
        return (n, n) if do_copy else (0, 0)

    def finish(self) -> None:
        """Make the caller's current stream wait for everything the pipeline has been given."""
        cur = torch.cuda.current_stream(self.device)
        for s in (self.s_in, self.s_run, self.s_out):
            cur.wait_stream(s)
        self._open = False


def enhance_batch_host(net: DenoiseGenerator, host_images: torch.Tensor, device="cuda", sub_batch: int = 64,
                       host_out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """host_images: CPU uint8 [B,H,W,3] (pinned for full speed). Returns CPU uint8 [B,H,W,3]."""
    B, H, W, _ = host_images.shape
    if host_out is None:
        host_out = torch.empty_like(host_images)
        if B:
            host_out = host_out.pin_memory()
    if B == 0:
        return host_out
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
