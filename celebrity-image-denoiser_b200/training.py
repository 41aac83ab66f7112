"""Host mirror of the reference's denoise-GAN training step, on the C-ABI training entry points.

Reference: backend/trainingcode/denoise_gan_code/training.py
    DenoiseDiscriminator        :77-98   (same module tree -> same state_dict keys)
    DenoiseGANTrainer.__init__  :230-242 (Adam lr 1e-4, betas (0.9, 0.999); BCELoss + MSELoss)
    DenoiseGANTrainer.train     :403-427 (one iteration = `DenoiseGANTrainStep.step`)

The arithmetic runs in libcid_b200.so (cid_train_*, include/cid_b200.h). PyTorch supplies the device buffers, the
stream and - when torch.distributed is initialised with world size > 1 - the NCCL all-reduce of the flat gradient
sections between each backward pass and its Adam update (BASELINE config #5; the reference itself is single-GPU).
There is no CPU or PyTorch compute path here.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, Optional

import torch
import torch.nn as nn

from . import lib as _lib
from .generator import DenoiseGenerator


class DenoiseDiscriminator(nn.Module):
    """Parameter container with the reference's module tree (training.py:77-98). The forward pass only exists inside
    the training step (train-mode BatchNorm on the tensor-core path); calling the module directly raises."""

    # (kind, arguments) per Sequential slot: the indices - hence the state_dict keys model.0.weight ... model.12.bias -
    # are the reference's (training.py:79-94)
    _SPEC = (("conv", 3, 64, 1), ("lrelu",), ("conv", 64, 64, 2), ("bn", 64), ("lrelu",), ("conv", 64, 128, 1), ("bn", 128),
             ("lrelu",), ("conv", 128, 128, 2), ("bn", 128), ("lrelu",), ("pool",), ("head", 128, 1), ("sigmoid",))

    def __init__(self):
        super().__init__()
        make = {"conv": lambda ci, co, s: nn.Conv2d(ci, co, kernel_size=3, stride=s, padding=1),
                "lrelu": lambda: nn.LeakyReLU(0.2), "bn": nn.BatchNorm2d, "pool": lambda: nn.AdaptiveAvgPool2d(1),
                "head": lambda ci, co: nn.Conv2d(ci, co, kernel_size=1), "sigmoid": nn.Sigmoid}
        self.model = nn.Sequential(*[make[kind](*a) for kind, *a in self._SPEC])

    def forward(self, x):  # pragma: no cover - by design
        raise RuntimeError("DenoiseDiscriminator runs inside DenoiseGANTrainStep (cid_train_*); it has no standalone "
                           "forward and no PyTorch fallback")


def allreduce_mean_(flat: torch.Tensor, group=None) -> torch.Tensor:
    """Average a flat gradient section over the data-parallel ranks, in place (NCCL on GPUs, gloo in the CPU tests).
    A no-op without an initialised process group or with world size 1. One collective per network per step: the whole
    section (7.3 MB generator / 1.0 MB discriminator) is a single bucket - NVSwitch all-reduce cost is launch latency,
    not per-link bandwidth, at this size."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized():
        world = dist.get_world_size(group)
        if world > 1:
            dist.all_reduce(flat, group=group)
            flat.mul_(1.0 / world)
    return flat


def sync_replicas_(tensors, group=None, src: int = 0) -> None:
    """Make every data-parallel rank start from rank `src`'s values (parameters, BatchNorm running statistics), in place.
    Both gradient-exchange paths average gradients and assume the replicas hold identical weights; modules constructed
    per process are randomly initialised per process, so without this step unseeded ranks would silently train
    different weights on the averaged gradients. A no-op without an initialised process group / with world size 1."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) < 2:
        return
    g = group if group is not None else dist.group.WORLD
    root = dist.get_global_rank(g, src) if group is not None else src
    for t in tensors:
        dist.broadcast(t, src=root, group=group)


class DenoiseGANTrainStep:
    """One data-parallel replica of the reference's training iteration.

    `generator` / `discriminator` provide the initial parameters (their state_dict keys are the reference's); the
    training state then lives in ONE flat fp32 device buffer (parameters | gradients | Adam exp_avg | exp_avg_sq |
    BatchNorm running statistics) that the C library updates in place. `export_state_dicts()` copies it back.
    """

    def __init__(self, generator: nn.Module, discriminator: nn.Module, device="cuda", lr: float = 1e-4,
                 betas=(0.9, 0.999), eps: float = 1e-8, process_group=None, peer_allreduce: bool = True):
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("the training step runs on a CUDA device (sm_100a); there is no CPU fallback")
        self.index = self.device.index if self.device.index is not None else torch.cuda.current_device()
        self.lib = _lib.load()
        self.lr, self.betas, self.eps = float(lr), (float(betas[0]), float(betas[1])), float(eps)
        self.group = process_group
        h = C.c_void_p()
        _lib.check(self.lib.cid_train_create(self.index, C.byref(h)), "cid_train_create")
        self.handle = h
        n = C.c_size_t()
        _lib.check(self.lib.cid_train_state_floats(h, C.byref(n)), "cid_train_state_floats")
        self.peer_fused = False
        self._symm = None
        self._peer_error = None
        peers = self._symmetric_state(n.value) if peer_allreduce else None
        self.state = peers[0] if peers else torch.zeros(n.value, dtype=torch.float32, device=self.device)
        _lib.check(self.lib.cid_train_bind_state(h, C.c_void_p(self.state.data_ptr()), n.value), "cid_train_bind_state")
        if peers:
            self._bind_peers(*peers)
        self.names = [self.lib.cid_train_tensor_name(h, i).decode() for i in range(self.lib.cid_train_tensor_count(h))]
        self._ws: Optional[torch.Tensor] = None
        self._ws_key = None
        self.load_state_dicts(generator.state_dict(), discriminator.state_dict())

    # ---- data-parallel peers -----------------------------------------------------------------
    def _symmetric_state(self, n: int):
        """With >1 ranks on one NVLink domain: allocate the state (and the barrier flags) as torch symmetric memory so
        that cid_train_adam can run the gradient all-reduce inside the optimiser kernel (include/cid_b200.h,
        cid_train_set_peers). Returns (state, flags, state_handle, flags_handle) or None; every rank takes the same
        decision (a MIN all-reduce of the outcome), otherwise the NCCL all-reduce path of `step` is used."""
        import torch.distributed as dist
        if not (dist.is_available() and dist.is_initialized()) or os.environ.get("CID_PEER_ALLREDUCE", "1") == "0":
            return None
        group = self.group if self.group is not None else dist.group.WORLD
        world = dist.get_world_size(group)
        if world < 2 or dist.get_backend(group) != "nccl":
            return None
        out, ok = None, 1
        try:
            import torch.distributed._symmetric_memory as symm
            if world > 8:
                raise RuntimeError("more than 8 ranks")
            with torch.cuda.device(self.device):
                state = symm.empty(n, dtype=torch.float32, device=self.device)
                flags = symm.empty(int(self.lib.cid_train_peer_flag_ints()), dtype=torch.int32, device=self.device)
                state.zero_()
                flags.zero_()
                torch.cuda.synchronize(self.device)
                hs = symm.rendezvous(state, group)
                hf = symm.rendezvous(flags, group)
            if len(hs.buffer_ptrs) != world or hs.buffer_ptrs[hs.rank] != state.data_ptr():
                raise RuntimeError("unexpected symmetric-memory pointer table")
            out = (state, flags, hs, hf)
        except Exception as e:  # symmetric memory unavailable (no P2P, old driver, ...): NCCL path
            self._peer_error = repr(e)
            ok = 0
        agree = torch.tensor([ok], dtype=torch.int32, device=self.device)
        dist.all_reduce(agree, op=dist.ReduceOp.MIN, group=group)  # also orders the zeroing above before any peer's first flag write
        return out if int(agree.item()) == 1 else None

    def _bind_peers(self, state, flags, hs, hf) -> None:
        world = hs.world_size
        sp = (C.c_void_p * world)(*[C.c_void_p(int(p)) for p in hs.buffer_ptrs])
        fp = (C.c_void_p * world)(*[C.c_void_p(int(p)) for p in hf.buffer_ptrs])
        _lib.check(self.lib.cid_train_set_peers(self.handle, world, hs.rank, sp, fp), "cid_train_set_peers")
        self._symm = (flags, hs, hf)  # keep the mappings alive
        self.peer_fused = True

    def bind_peers_raw(self, world: int, rank: int, state_ptrs, flag_ptrs, keep_alive=None) -> None:
        """cid_train_set_peers with caller-provided device pointers (rank order). For callers that map the peers' buffers
        themselves (CUDA IPC, several replicas inside one process); `flag_ptrs[r]` = cid_train_peer_flag_ints() zeroed ints."""
        sp = (C.c_void_p * world)(*[C.c_void_p(int(p)) for p in state_ptrs])
        fp = (C.c_void_p * world)(*[C.c_void_p(int(p)) for p in flag_ptrs])
        _lib.check(self.lib.cid_train_set_peers(self.handle, world, rank, sp, fp), "cid_train_set_peers")
        self._symm = keep_alive
        self.peer_fused = world > 1

    # ---- state ------------------------------------------------------------------------------
    def close(self):
        if getattr(self, "handle", None):
            torch.cuda.synchronize(self.device)
            self.lib.cid_train_destroy(self.handle)
            self.handle = None

    def __del__(self):  # best effort
        try:
            self.close()
        except Exception:
            pass

    def set_option(self, name: str, value: int) -> None:
        _lib.check(self.lib.cid_train_set_option(self.handle, name.encode(), int(value)), "cid_train_set_option")

    def view(self, key: str, section: int = 0) -> torch.Tensor:
        """Flat view of one named tensor inside the state buffer (section 0 params, 1 grads, 2 exp_avg, 3 exp_avg_sq)."""
        off, n = C.c_size_t(), C.c_size_t()
        _lib.check(self.lib.cid_train_tensor(self.handle, key.encode(), section, C.byref(off), C.byref(n)), "cid_train_tensor")
        return self.state[off.value:off.value + n.value]

    def section(self, net: str, section: int) -> torch.Tensor:
        off, n = self._section_range(net, section)
        return self.state[off:off + n]

    def load_state_dicts(self, g_sd: Dict[str, torch.Tensor], d_sd: Dict[str, torch.Tensor]) -> None:
        for name in self.names:
            net, key = name.split(".", 1)
            src = (g_sd if net == "generator" else d_sd)[key]
            self.view(name).copy_(src.detach().to(self.device, torch.float32).reshape(-1))
        self._sync_replicas()
        _lib.check(self.lib.cid_train_sync_weights(self.handle, C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)),
                   "cid_train_sync_weights")

    def _sync_replicas(self) -> None:
        """Rank 0's parameters and BatchNorm statistics become everybody's (see sync_replicas_)."""
        g_off, g_n = self._section_range("generator", 0)
        d_off, d_n = self._section_range("discriminator", 0)
        P = g_n + d_n
        assert g_off == 0 and d_off == g_n
        sync_replicas_([self.state[:P], self.state[4 * P:]], self.group)

    def _section_range(self, net: str, section: int):
        off, n = C.c_size_t(), C.c_size_t()
        _lib.check(self.lib.cid_train_section(self.handle, net.encode(), section, C.byref(off), C.byref(n)), "cid_train_section")
        return off.value, n.value

    def export_state_dicts(self, generator: nn.Module, discriminator: nn.Module) -> None:
        """Copy the trained parameters and BatchNorm statistics back into the two modules."""
        for name in self.names:
            net, key = name.split(".", 1)
            dst = (generator if net == "generator" else discriminator).state_dict()[key]
            v = self.view(name)
            dst.copy_(v.reshape(dst.shape).to(dst.dtype) if dst.dim() else v[0].to(dst.dtype))

    def grads(self, net: str) -> Dict[str, torch.Tensor]:
        out = {}
        for name in self.names:
            n, key = name.split(".", 1)
            if n == net and "running_" not in key and "num_batches" not in key:
                out[key] = self.view(name, 1).clone()
        return out

    # ---- one iteration ----------------------------------------------------------------------
    def _workspace(self, B: int, H: int, W: int):
        key = (B, H, W)
        if self._ws_key != key:
            n = C.c_size_t()
            _lib.check(self.lib.cid_train_workspace_bytes(self.handle, B, H, W, C.byref(n)), "cid_train_workspace_bytes")
            self._ws = None
            self._ws = torch.empty(n.value + 1024, dtype=torch.uint8, device=self.device)
            self._ws_key = key
        p = (self._ws.data_ptr() + 1023) & ~1023
        return p, self._ws.numel() - (p - self._ws.data_ptr())

    def _allreduce(self, flat: torch.Tensor) -> None:
        if not self.peer_fused:  # fused: cid_train_adam reads every rank's gradients itself
            allreduce_mean_(flat, self.group)

    def d_backward(self, noisy: torch.Tensor, clean: torch.Tensor) -> None:
        B, H, W, ws, wsn, st = self._args(noisy, clean)
        _lib.check(self.lib.cid_train_d_backward(self.handle, C.c_void_p(noisy.data_ptr()), C.c_void_p(clean.data_ptr()), B, H, W,
                                                 C.c_void_p(ws), wsn, st), "cid_train_d_backward")

    def g_backward(self, noisy: torch.Tensor, clean: torch.Tensor) -> None:
        B, H, W, ws, wsn, st = self._args(noisy, clean)
        _lib.check(self.lib.cid_train_g_backward(self.handle, C.c_void_p(noisy.data_ptr()), C.c_void_p(clean.data_ptr()), B, H, W,
                                                 C.c_void_p(ws), wsn, st), "cid_train_g_backward")

    def adam(self, which: int) -> None:
        st = C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        _lib.check(self.lib.cid_train_adam(self.handle, which, self.lr, self.betas[0], self.betas[1], self.eps, st), "cid_train_adam")

    def _args(self, noisy, clean):
        for t in (noisy, clean):
            if not (t.is_cuda and t.dtype == torch.float32 and t.is_contiguous() and t.dim() == 4 and t.shape[1] == 3):
                raise RuntimeError("noisy / clean must be contiguous float32 CUDA tensors [B,3,H,W]")
        if noisy.shape != clean.shape or noisy.device != clean.device or noisy.device.index != self.index:
            raise RuntimeError("noisy and clean must have the same shape and live on the trainer's device")
        B, _, H, W = noisy.shape
        ws, wsn = self._workspace(B, H, W)
        return B, H, W, ws, wsn, C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def step(self, noisy: torch.Tensor, clean: torch.Tensor, read_losses: bool = True):
        """training.py:409-427: D step then G step. Returns {d_loss, g_loss, content, adv} (synchronises) or None."""
        self.d_backward(noisy, clean)
        self._allreduce(self.section("discriminator", 1))
        self.adam(1)
        self.g_backward(noisy, clean)
        self._allreduce(self.section("generator", 1))
        self.adam(0)
        return self.losses() if read_losses else None

    def losses(self) -> Dict[str, float]:
        out = (C.c_double * 4)()
        _lib.check(self.lib.cid_train_losses(self.handle, out), "cid_train_losses")
        return {"d_loss": out[0], "g_loss": out[1], "content": out[2], "adv": out[3]}

    def kernel_launches(self) -> int:
        return int(self.lib.cid_train_kernel_launches(self.handle))

    def debug_buffer(self, name: str, B: int, H: int, W: int) -> torch.Tensor:
        """Test hook: a named training workspace buffer of the last step as a float32 CPU tensor (NHWC, or NCHW for 'y')."""
        ws, _ = self._workspace(B, H, W)
        c, lv, es = C.c_int(), C.c_int(), C.c_int()
        host = torch.empty(B * H * W * 256 * 2 + 16, dtype=torch.uint8)
        _lib.check(self.lib.cid_train_debug_buffer(self.handle, name.encode(), C.c_void_p(ws), B, H, W, C.c_void_p(host.data_ptr()),
                                                   host.numel(), C.byref(c), C.byref(lv), C.byref(es)), "cid_train_debug_buffer")
        h, w = H >> lv.value, W >> lv.value
        n = B * h * w * c.value
        if es.value == 4:
            t = host[:n * 4].view(torch.float32)
            return t.reshape(B, 3, h, w).clone() if name == "y" else t.reshape(B, h, w, c.value).clone()
        return host[:n * 2].view(torch.bfloat16).reshape(B, h, w, c.value).float()
