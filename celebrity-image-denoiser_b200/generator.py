"""Drop-in `DenoiseGenerator`: the reference's nn.Module surface over the CUDA C-ABI.

Mirrors `DenoiseGenerator` of the reference (backend/app.py:39-103): same class name, no
constructor arguments, the same 24 `state_dict` keys and shapes, `forward(x)` taking fp32 NCHW in
[-1,1] and returning the tanh output as fp32 NCHW on the same device. So
`PT_MODELS["denoise"] = DenoiseGenerator().to(DEVICE)` (backend/app.py:319-320),
`load_state_safely(model, path)` (:257-274) and `net(x_pt)` (:433) work unchanged.

The parameters are ordinary `nn.Parameter`s (so `.to()`, `load_state_dict(strict=False)`,
`state_dict()` behave exactly as in the reference) but the module never computes with them in
PyTorch: `forward` hands them to libcid_b200.so, which packs them to its bf16 tensor-core layout
and runs the hand-written sm_100a kernels. No cuDNN, no CPU path: a CPU tensor or a missing
library raises.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, Optional, Tuple

import torch
import torch.nn as nn

from . import lib as _lib

# (attribute, [(kind, cin, cout), ...]); conv entries are followed by a parameter-less ReLU slot
# so the Sequential indices — and therefore the state_dict keys — are the reference's
# (down1.0, down1.2, ...; the last conv of upconv1 has no ReLU after it: backend/app.py:74-78).
_BLOCKS = (
    ("down1", ((3, 64), (64, 64)), True),
    ("down2", ((64, 128), (128, 128)), True),
    ("bottleneck", ((128, 256), (256, 256)), True),
    ("up2", (256, 128), None),
    ("upconv2", ((256, 128), (128, 128)), True),
    ("up1", (128, 64), None),
    ("upconv1", ((128, 64), (64, 3)), False),
)

STATE_DICT_SHAPES: Dict[str, Tuple[int, ...]] = {}
for _name, _spec, _last_relu in _BLOCKS:
    if _last_relu is None:
        STATE_DICT_SHAPES[f"{_name}.weight"] = (_spec[0], _spec[1], 2, 2)  # ConvTranspose2d: [Cin,Cout,kh,kw]
        STATE_DICT_SHAPES[f"{_name}.bias"] = (_spec[1],)
    else:
        for _i, (_ci, _co) in enumerate(_spec):
            STATE_DICT_SHAPES[f"{_name}.{2 * _i}.weight"] = (_co, _ci, 3, 3)
            STATE_DICT_SHAPES[f"{_name}.{2 * _i}.bias"] = (_co,)


class _Engine:
    """One C-ABI handle (= one device) plus its cached workspace."""

    def __init__(self, device_index: int, model: str = "denoise"):
        self.lib = _lib.load()
        self.device_index = device_index
        self.model = model
        h = C.c_void_p()
        _lib.check(self.lib.cid_create_model(device_index, model.encode(), C.byref(h)), "cid_create_model")
        self.handle = h
        self.signature = None
        self._ws: Optional[torch.Tensor] = None
        self._ws_key = None

    def close(self):
        if self.handle:
            self.lib.cid_destroy(self.handle)
            self.handle = None

    def __del__(self):  # best effort
        try:
            self.close()
        except Exception:
            pass

    def set_option(self, name: str, value: int):
        _lib.check(self.lib.cid_set_option(self.handle, name.encode(), int(value)), "cid_set_option")
        self._ws = None
        self._ws_key = None

    def upload(self, state: Dict[str, torch.Tensor], shapes: Optional[Dict[str, Tuple[int, ...]]] = None):
        # cid_set_param copies with a blocking cudaMemcpy (legacy default stream); parameters written on the caller's
        # current non-blocking side stream (load_state_dict / an optimizer step / .to() under torch.cuda.stream) must
        # have landed before that copy reads them
        torch.cuda.current_stream(self.device_index).synchronize()
        for key, shape in (shapes or STATE_DICT_SHAPES).items():
            t = state[key].detach()
            if t.dtype != torch.float32 or not t.is_contiguous():
                t = t.float().contiguous()
            if tuple(t.shape) != shape:
                raise RuntimeError(f"{key}: shape {tuple(t.shape)} != {shape}")
            dims = (C.c_int64 * len(shape))(*shape)
            _lib.check(self.lib.cid_set_param(self.handle, key.encode(), C.c_void_p(t.data_ptr()), dims, len(shape)),
                       f"cid_set_param({key})")
        _lib.check(self.lib.cid_finalize(self.handle), "cid_finalize")

    def workspace(self, B: int, H: int, W: int) -> torch.Tensor:
        key = (B, H, W)
        if self._ws_key != key:
            n = C.c_size_t()
            _lib.check(self.lib.cid_workspace_bytes(self.handle, B, H, W, C.byref(n)), "cid_workspace_bytes")
            if self._ws is None or self._ws.numel() < n.value:
                self._ws = None  # release before allocating the larger one
                self._ws = torch.empty(n.value + 1024, dtype=torch.uint8, device=f"cuda:{self.device_index}")
            self._ws_key = key
        return self._ws

    def ws_ptr(self, ws: torch.Tensor) -> Tuple[int, int]:
        p = (ws.data_ptr() + 1023) & ~1023
        return p, ws.numel() - (p - ws.data_ptr())

    def launches(self) -> int:
        return int(self.lib.cid_kernel_launches(self.handle))

    def profile_read(self):
        ms = (C.c_double * 24)()
        n = C.c_int64()
        _lib.check(self.lib.cid_profile_read(self.handle, ms, C.byref(n)), "cid_profile_read")
        return list(ms), int(n.value)


class _CudaGenerator(nn.Module):
    """Engine plumbing shared by the generators of PT_MODELS (backend/app.py:319-324): one C-ABI handle per
    device, lazy re-pack of the parameters when they change, `net(x)` and the fused uint8 route path."""

    MODEL = "denoise"     # cid_create_model name
    OUT_SCALE = 1         # spatial scale of the output
    PADS_INPUT = True     # forward_u8 feeds the padded frame (ESRGAN feeds the un-padded image, backend/app.py:387-390)
    EVAL_ONLY = False     # modules with BatchNorm: only the eval-mode (folded) form exists
    LAYER_NAMES: Tuple[str, ...] = ()
    PARAM_SHAPES: Optional[Dict[str, Tuple[int, ...]]] = None  # None: every floating-point state_dict entry

    def _init_engine_state(self):
        self._engines: Dict[int, _Engine] = {}
        self._options: Dict[str, int] = {}
        self._slots = None

    # ---- change detection ------------------------------------------------------------------------
    # forward() re-packs the weights when a parameter was loaded, moved or modified since the last pack. The check
    # walks a cached list of (owner dict, name) slots - tensor identity, storage pointer and version counter - instead
    # of rebuilding state_dict() on every call (that alone cost tens of microseconds on a ~100 us batch-1 forward).
    def _param_slots(self):
        if self._slots is None:
            slots = []
            for prefix, mod in self.named_modules():
                for store in (mod._parameters, mod._buffers):
                    for name, t in store.items():
                        if t is None or not t.dtype.is_floating_point:
                            continue
                        key = f"{prefix}.{name}" if prefix else name
                        if self.PARAM_SHAPES is None or key in self.PARAM_SHAPES:
                            slots.append((key, store, name))
            self._slots = slots
        return self._slots

    def _signature(self):
        sig = []
        for _, store, name in self._param_slots():
            t = store[name]
            sig.append((id(t), t.data_ptr(), t._version))
        return tuple(sig)

    def _float_state(self) -> Dict[str, torch.Tensor]:
        return {key: store[name].detach() for key, store, name in self._param_slots()}

    def _engine(self, device: torch.device) -> _Engine:
        if device.type != "cuda":
            raise RuntimeError(f"{type(self).__name__} (B200-native) runs only on CUDA tensors; there is no CPU fallback")
        if self.EVAL_ONLY and self.training:
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
        if eng.signature != sig:  # parameters were loaded / moved / modified since the last pack
            state = self._float_state()
            eng.upload(state, self.PARAM_SHAPES or {k: tuple(v.shape) for k, v in state.items()})
            eng.signature = sig
        return eng

    def set_option(self, name: str, value: int):
        """Forwarded to cid_set_option (e.g. "chunk")."""
        self._options[name] = int(value)
        for eng in self._engines.values():
            eng.set_option(name, value)
        return self

    def kernel_launches(self) -> int:
        return sum(e.launches() for e in self._engines.values())

    def graph_replays(self) -> int:
        """Passes that ran as one CUDA-graph launch (small batches; cid_set_option "graph")."""
        return sum(int(e.lib.cid_graph_replays(e.handle)) for e in self._engines.values())

    def release_workspace(self) -> None:
        """Drop the cached activation workspace(s); the next forward allocates what it needs."""
        for e in self._engines.values():
            e._ws, e._ws_key = None, None

    def profile_read(self):
        """(per-layer summed ms dict, passes) since the last read; needs set_option("profile", 1)."""
        eng = next(iter(self._engines.values()))
        ms, n = eng.profile_read()
        return dict(zip(self.LAYER_NAMES, ms)), n

    # ---- the reference's call: y = net(x) (backend/app.py:393, 433) --------------------------------
    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if x.dim() != 4 or x.size(1) != 3:
            raise RuntimeError(f"expected input [B,3,H,W], got {tuple(x.shape)}")
        eng = self._engine(x.device)
        if x.dtype != torch.float32 or not x.is_contiguous():
            x = x.float().contiguous()
        B, _, H, W = x.shape
        s = self.OUT_SCALE
        oh, ow = self._out_hw(H, W)
        y = torch.empty((B, 3, oh * s, ow * s), dtype=torch.float32, device=x.device)
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

    @staticmethod
    def _out_hw(H: int, W: int) -> Tuple[int, int]:
        """Spatial size of the network's result before OUT_SCALE (the SR generators keep the input's)."""
        return H, W

    # ---- fused route path: u8 NHWC image(s) in -> u8 NHWC image(s) out -----------------------------
    def forward_u8(self, x: torch.Tensor, padding: Tuple[int, int, int, int] = (0, 0, 0, 0),
                   out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """x: CUDA uint8 NHWC [B,H,W,3] (the un-padded upload); padding = get_padding()'s (l,t,r,b) for this model.
        denoise: returns [B,H,W,3] = pad -> normalise -> generator -> denorm -> clamp -> truncating x255 -> crop, as
        backend/app.py:383-480 does. ESRGAN: returns [B,H,W,3] - the route feeds the un-padded image and crops the
        result at the padding offsets (:387-397, 473-479). SRGAN: returns [B,4(H+t+b),4(W+l+r),3], not cropped (:480-485)."""
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
        ws = eng.workspace(B, H + t + b, W + l + r) if self.PADS_INPUT else eng.workspace(B, H, W)
        wp, wn = eng.ws_ptr(ws)
        stream = torch.cuda.current_stream(x.device).cuda_stream
        with torch.cuda.device(x.device):
            _lib.check(eng.lib.cid_forward_u8_nhwc(eng.handle, C.c_void_p(x.data_ptr()), C.c_void_p(y.data_ptr()),
                                                   B, H, W, l, t, r, b, C.c_void_p(wp), wn, C.c_void_p(stream)),
                       "cid_forward_u8_nhwc")
        return y


class DenoiseGenerator(_CudaGenerator):
    MODEL = "denoise"

    @staticmethod
    def _out_hw(H: int, W: int) -> Tuple[int, int]:
        """Two floor-halving pools and two doubling transposed convs: for sizes that are not multiples of 4 the
        reference crops the skip tensors to the decoder (backend/app.py:89-99) and returns this smaller image."""
        return 4 * (H // 4), 4 * (W // 4)
    PARAM_SHAPES = STATE_DICT_SHAPES
    LAYER_NAMES = ("down1.0", "down1.2", "down2.0", "down2.2", "bottleneck.0", "bottleneck.2", "up2", "upconv2.0",
                   "upconv2.2", "up1", "upconv1.0", "upconv1.2")

    def __init__(self):
        super().__init__()
        for name, spec, last_relu in _BLOCKS:
            if last_relu is None:
                self.add_module(name, nn.ConvTranspose2d(spec[0], spec[1], kernel_size=2, stride=2))
                continue
            mods = []
            for i, (ci, co) in enumerate(spec):
                mods.append(nn.Conv2d(ci, co, kernel_size=3, padding=1))
                if last_relu or i + 1 < len(spec):
                    mods.append(nn.ReLU())
            self.add_module(name, nn.Sequential(*mods))
        self._init_engine_state()

    def debug_activation(self, name: str, x_shape: Tuple[int, int, int, int]) -> torch.Tensor:
        """Test hook: intermediate activation `name` of the last forward with input [B,3,H,W], as a
        float32 NHWC CPU tensor (values are bf16-exact)."""
        B, _, H, W = x_shape
        eng = next(iter(self._engines.values()))
        ws = eng.workspace(B, H, W)
        wp, _ = eng.ws_ptr(ws)
        lvl_max = B * H * W * 256  # generous upper bound (elements)
        buf = torch.empty(lvl_max, dtype=torch.bfloat16)
        c, h, w = C.c_int(), C.c_int(), C.c_int()
        _lib.check(eng.lib.cid_debug_copy_activation(eng.handle, name.encode(), C.c_void_p(wp), B, H, W,
                                                     C.c_void_p(buf.data_ptr()), buf.numel() * 2, C.byref(c),
                                                     C.byref(h), C.byref(w)), "cid_debug_copy_activation")
        n = B * h.value * w.value * c.value
        return buf[:n].view(B, h.value, w.value, c.value).float()
