"""ctypes binding of csrc/libcid_b200.so (the C ABI declared in include/cid_b200.h).

There is deliberately no fallback: if the shared library is missing or a call fails, a
RuntimeError is raised (inside the reference's route that becomes its HTTP 500,
backend/app.py:499-502).
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path

import os

# CID_B200_LIB: another build of the same library (same-box A/B runs of two kernel versions, tools/gpu_session.sh)
_LIB_PATH = Path(os.environ.get("CID_B200_LIB") or Path(__file__).resolve().parent / "csrc" / "libcid_b200.so")
_lib = None

ABI_VERSION = 7

# name -> (restype, argtypes); mirrors include/cid_b200.h one to one
_SIGNATURES = {
    "cid_version": (C.c_int, []),
    "cid_last_error": (C.c_char_p, []),
    "cid_create": (C.c_int, [C.c_int, C.POINTER(C.c_void_p)]),
    "cid_create_model": (C.c_int, [C.c_int, C.c_char_p, C.POINTER(C.c_void_p)]),
    "cid_destroy": (C.c_int, [C.c_void_p]),
    "cid_set_param": (C.c_int, [C.c_void_p, C.c_char_p, C.c_void_p, C.POINTER(C.c_int64), C.c_int]),
    "cid_finalize": (C.c_int, [C.c_void_p]),
    "cid_set_option": (C.c_int, [C.c_void_p, C.c_char_p, C.c_int64]),
    "cid_workspace_bytes": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_size_t)]),
    "cid_forward_f32_nchw": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int,
                                       C.c_void_p, C.c_size_t, C.c_void_p]),
    "cid_forward_f32_nhwc": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int,
                                       C.c_void_p, C.c_size_t, C.c_void_p]),
    "cid_forward_u8_nhwc": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int,
                                      C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.c_void_p]),
    "cid_noise_gaussian_u8": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_float, C.c_float,
                                        C.c_void_p, C.c_uint64, C.c_uint64, C.c_void_p]),
    "cid_noise_elementwise_u8": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_float, C.c_float,
                                           C.c_void_p, C.c_uint64, C.c_uint64, C.c_void_p]),
    "cid_noise_salt_pepper_u8": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int,
                                           C.c_double, C.c_double, C.c_void_p, C.c_int64, C.c_uint64, C.c_uint64,
                                           C.c_void_p]),
    "cid_noise_unit_u8": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_float, C.c_float,
                                    C.c_void_p, C.c_uint64, C.c_uint64, C.c_void_p]),
    "cid_noise_salt_pepper_unit_u8": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int,
                                                C.c_double, C.c_void_p, C.c_int64, C.c_uint64, C.c_uint64, C.c_void_p]),
    "cid_resize_bicubic_u8": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int,
                                        C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "cid_resize_bicubic_u8_static": (C.c_int, [C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int,
                                               C.c_int, C.c_int, C.c_void_p]),
    "cid_normalize_u8_to_f32_nchw": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int,
                                               C.c_float, C.c_float, C.c_void_p]),
    "cid_kernel_launches": (C.c_int64, [C.c_void_p]),
    "cid_graph_replays": (C.c_int64, [C.c_void_p]),
    "cid_profile_read": (C.c_int, [C.c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_int64)]),
    "cid_debug_copy_activation": (C.c_int, [C.c_void_p, C.c_char_p, C.c_void_p, C.c_int, C.c_int, C.c_int,
                                            C.c_void_p, C.c_size_t, C.POINTER(C.c_int), C.POINTER(C.c_int),
                                            C.POINTER(C.c_int)]),
    # training step
    "cid_train_create": (C.c_int, [C.c_int, C.POINTER(C.c_void_p)]),
    "cid_train_destroy": (C.c_int, [C.c_void_p]),
    "cid_train_set_option": (C.c_int, [C.c_void_p, C.c_char_p, C.c_int64]),
    "cid_train_state_floats": (C.c_int, [C.c_void_p, C.POINTER(C.c_size_t)]),
    "cid_train_bind_state": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t]),
    "cid_train_tensor": (C.c_int, [C.c_void_p, C.c_char_p, C.c_int, C.POINTER(C.c_size_t), C.POINTER(C.c_size_t)]),
    "cid_train_tensor_count": (C.c_int, [C.c_void_p]),
    "cid_train_tensor_name": (C.c_char_p, [C.c_void_p, C.c_int]),
    "cid_train_section": (C.c_int, [C.c_void_p, C.c_char_p, C.c_int, C.POINTER(C.c_size_t), C.POINTER(C.c_size_t)]),
    "cid_train_sync_weights": (C.c_int, [C.c_void_p, C.c_void_p]),
    "cid_train_workspace_bytes": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_size_t)]),
    "cid_train_d_backward": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                       C.c_size_t, C.c_void_p]),
    "cid_train_g_backward": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                       C.c_size_t, C.c_void_p]),
    "cid_train_adam": (C.c_int, [C.c_void_p, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float, C.c_void_p]),
    "cid_train_losses": (C.c_int, [C.c_void_p, C.POINTER(C.c_double)]),
    "cid_train_peer_flag_ints": (C.c_int, []),
    "cid_train_set_peers": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p)]),
    "cid_train_kernel_launches": (C.c_int64, [C.c_void_p]),
    "cid_train_debug_buffer": (C.c_int, [C.c_void_p, C.c_char_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                         C.c_size_t, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]),
}

EXPORTED_SYMBOLS = tuple(_SIGNATURES)


def library_path() -> Path:
    return _LIB_PATH


def load():
    """dlopen the C-ABI library (once). Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not _LIB_PATH.exists():
        raise RuntimeError(
            f"{_LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(there is no CPU / PyTorch fallback for the generator path)")
    lib = C.CDLL(str(_LIB_PATH))
    for name, (res, args) in _SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the .so does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    if lib.cid_version() != ABI_VERSION:
        raise RuntimeError(f"libcid_b200.so ABI {lib.cid_version()} != binding ABI {ABI_VERSION}; rebuild")
    _lib = lib
    return lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = load().cid_last_error().decode("utf-8", "replace")
        raise RuntimeError(f"libcid_b200 {what} failed: {msg}")
