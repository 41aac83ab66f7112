"""Route-level micro-batching: concurrent uploads of the same size share ONE generator forward.

Reference shape: the upload route builds a batch of one per request (`transform(image).unsqueeze(0)`,
backend/app.py:401-406) and runs `net(x_pt)` once per request (:433); concurrent requests therefore run
one after another. A batch-1 forward at 128x128 keeps the 148 SMs ~6 % busy (bench.py `latency`), so N
concurrent requests cost N x the batch-1 latency although one batch-N forward takes barely longer than a
batch-1 forward. `MicroBatcher` sits between the route and the generator: a request hands over its RGB
uint8 image and gets a future; one worker thread drains the queue, groups the waiting requests by image
size (the fused kernels take one (H, W, padding) per launch), stages each group in pinned host memory,
and runs H2D -> `cid_forward_u8_nhwc` -> D2H once per group. Batches are bit-identical to single-image
forwards (no op of the generator mixes batch elements; tests/test_gpu_parity.py batch-invariance tests),
so batching is invisible in the results.

    batcher = MicroBatcher(PT_MODELS["denoise"], max_batch=32, max_wait_ms=2.0)
    out_u8 = batcher.enhance(np.asarray(image))           # blocking, thread-safe
    out_u8 = await batcher.enhance_async(np.asarray(image))   # inside an `async def` route
"""
from __future__ import annotations

import queue
import threading
import time
from concurrent.futures import Future
from typing import Callable, Dict, List, Optional, Tuple

import numpy as np
import torch


class _Request:
    __slots__ = ("image", "future", "t_in")

    def __init__(self, image: np.ndarray):
        self.image = image
        self.future: Future = Future()
        self.t_in = time.perf_counter()


def group_by_shape(requests: List[_Request], max_batch: int) -> List[List[_Request]]:
    """Arrival-ordered groups of same-(H, W) requests, each at most max_batch long."""
    by_shape: Dict[Tuple[int, int], List[List[_Request]]] = {}
    order: List[List[_Request]] = []
    for r in requests:
        key = (int(r.image.shape[0]), int(r.image.shape[1]))
        groups = by_shape.setdefault(key, [])
        if not groups or len(groups[-1]) >= max_batch:
            groups.append([])
            order.append(groups[-1])
        groups[-1].append(r)
    return order


class MicroBatcher:
    """Collects concurrent single-image requests into batched forwards.

    net: a module with `forward_u8(images_cuda_u8, pad, out=None)` (DenoiseGenerator) - or pass
    `run_batch`, a callable `(pinned CPU uint8 [B,H,W,3]) -> CPU uint8 [B,H,W,3]`, to replace the device
    path (host-logic tests do). max_wait_ms bounds the time the first request of a batch waits for company;
    the worker never waits while the GPU is idle AND nothing else is queued beyond that window."""

    def __init__(self, net=None, device="cuda", max_batch: int = 32, max_wait_ms: float = 2.0,
                 run_batch: Optional[Callable[[torch.Tensor], torch.Tensor]] = None):
        if net is None and run_batch is None:
            raise ValueError("MicroBatcher needs a generator or a run_batch callable")
        self.net, self.device = net, torch.device(device)
        self.max_batch, self.max_wait = max(int(max_batch), 1), max(float(max_wait_ms), 0.0) * 1e-3
        self._run_batch = run_batch or self._run_on_device
        self._q: "queue.Queue[Optional[_Request]]" = queue.Queue()
        self._staging: Dict[Tuple[int, int], Tuple[torch.Tensor, torch.Tensor]] = {}
        self._stream: Optional[torch.cuda.Stream] = None
        self.batches: List[int] = []        # size of every forward that ran (diagnostics / tests)
        self._closed = False
        self._worker = threading.Thread(target=self._loop, name="cid-microbatcher", daemon=True)
        self._worker.start()

    # ------------------------------------------------------------------ public
    def submit(self, image: np.ndarray) -> Future:
        """image: uint8 [H,W,3] RGB. Returns a Future whose result is the denoised uint8 [H,W,3]."""
        if self._closed:
            raise RuntimeError("MicroBatcher is closed")
        image = np.ascontiguousarray(image)
        if image.dtype != np.uint8 or image.ndim != 3 or image.shape[2] != 3 or image.shape[0] < 1 or image.shape[1] < 1:
            raise ValueError(f"expected uint8 [H,W,3], got {image.dtype} {image.shape}")
        r = _Request(image)
        self._q.put(r)
        return r.future

    def enhance(self, image: np.ndarray, timeout: Optional[float] = None) -> np.ndarray:
        return self.submit(image).result(timeout)

    async def enhance_async(self, image: np.ndarray) -> np.ndarray:
        import asyncio
        return await asyncio.wrap_future(self.submit(image))

    def close(self) -> None:
        if not self._closed:
            self._closed = True
            self._q.put(None)
            self._worker.join()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    # ------------------------------------------------------------------ worker
    def _collect(self) -> Optional[List[_Request]]:
        first = self._q.get()
        if first is None:
            return None
        got = [first]
        deadline = first.t_in + self.max_wait
        while len(got) < self.max_batch:
            left = deadline - time.perf_counter()
            try:
                r = self._q.get(timeout=left) if left > 0 else self._q.get_nowait()
            except queue.Empty:
                break
            if r is None:
                self._q.put(None)   # finish this batch, then stop
                break
            got.append(r)
        return got

    def _loop(self) -> None:
        while True:
            got = self._collect()
            if got is None:
                return
            for group in group_by_shape(got, self.max_batch):
                try:
                    H, W = group[0].image.shape[:2]
                    host_in, _ = self._buffers(H, W)
                    nb = len(group)
                    staged = host_in.numpy()          # a view of the (pinned) staging tensor
                    for i, r in enumerate(group):
                        np.copyto(staged[i], r.image)  # (np.asarray(PIL image) is read-only: no torch.from_numpy on it)
                    out = self._run_batch(host_in[:nb])
                    self.batches.append(nb)
                    res = out.numpy()
                    for i, r in enumerate(group):
                        r.future.set_result(res[i].copy())
                except BaseException as e:  # the route turns this into its HTTP 500 (backend/app.py:499-502)
                    for r in group:
                        if not r.future.done():
                            r.future.set_exception(e)

    def _buffers(self, H: int, W: int) -> Tuple[torch.Tensor, torch.Tensor]:
        key = (H, W)
        if key not in self._staging:
            pin = self.device.type == "cuda" and torch.cuda.is_available()
            mk = lambda: torch.empty((self.max_batch, H, W, 3), dtype=torch.uint8, pin_memory=pin)  # noqa: E731
            if len(self._staging) >= 8:   # bound the pinned memory held for rarely seen sizes
                self._staging.pop(next(iter(self._staging)))
            self._staging[key] = (mk(), mk())
        return self._staging[key]

    def _run_on_device(self, host_in: torch.Tensor) -> torch.Tensor:
        from .pipeline import enhance_batch_u8
        nb, H, W, _ = host_in.shape
        _, host_out = self._buffers(H, W)
        if self._stream is None:
            self._stream = torch.cuda.Stream(self.device)
        with torch.cuda.stream(self._stream):
            x = host_in.to(self.device, non_blocking=True)
            y = enhance_batch_u8(self.net, x)
            host_out[:nb].copy_(y, non_blocking=True)
        self._stream.synchronize()
        return host_out[:nb]
