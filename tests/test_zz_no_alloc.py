"""cid_resize_bicubic_u8_static: the allocation-free resize entry point (scratch in a static device array) against
cid_resize_bicubic_u8, through the C-ABI. No counterpart in the reference (include/cid_b200.h says why it exists); the
file is named so that pytest collects it last - a failure here must not hide the parity tests under `pytest -x`.

Buffers come from torch BEFORE the calls under test; the entry point itself gets plain device pointers."""
import ctypes as C

import numpy as np
import pytest
import torch

from cid_b200 import lib as _lib
from cid_b200.synth import synthetic_noisy_u8

TABLE_BYTES = 256 << 10   # CID_RESIZE_STATIC_TABLE_BYTES
TMP_BYTES = 8 << 20       # CID_RESIZE_STATIC_TMP_BYTES


def _p(t):
    return C.c_void_p(t.data_ptr())


@pytest.fixture(scope="module")
def dev():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    return torch.device("cuda", 0)


def _existing(lib, handle, x, oh, ow):
    B, ih, iw, c = x.shape
    out = torch.empty((B, oh, ow, c), dtype=torch.uint8, device=x.device)
    tmp = torch.empty((B, ih, ow, c), dtype=torch.uint8, device=x.device)
    _lib.check(lib.cid_resize_bicubic_u8(handle, _p(x), _p(out), _p(tmp), B, ih, iw, oh, ow, c, None),
               "cid_resize_bicubic_u8")
    return out


@pytest.mark.gpu
def test_static_scratch_resize_equals_the_allocating_entry_point(dev):
    lib = _lib.load()
    handle = C.c_void_p()
    _lib.check(lib.cid_create(0, C.byref(handle)), "cid_create")
    try:
        # two calls in a row share the static array: different shapes, tap counts and channel counts, so that stale
        # tables / intermediate bytes of the previous call would show; then the first case again
        cases = [(2, 300, 413, 256, 256, 3), (3, 64, 48, 128, 96, 3), (1, 97, 131, 40, 50, 1), (2, 300, 413, 256, 256, 3),
                 (4, 128, 128, 32, 32, 3)]
        for B, ih, iw, oh, ow, c in cases:
            x = torch.from_numpy(np.ascontiguousarray(
                synthetic_noisy_u8(ih + 3 * ow, B, ih, iw).numpy()[..., :c])).to(dev)
            want = _existing(lib, handle, x, oh, ow)
            got = torch.full((B, oh, ow, c), 0x5A, dtype=torch.uint8, device=dev)
            torch.cuda.synchronize()
            before = torch.cuda.memory_allocated(dev)
            rc = lib.cid_resize_bicubic_u8_static(0, _p(x), _p(got), B, ih, iw, oh, ow, c, None)
            _lib.check(rc, "cid_resize_bicubic_u8_static")
            torch.cuda.synchronize()
            assert torch.cuda.memory_allocated(dev) == before
            assert torch.equal(got, want), (B, ih, iw, oh, ow, c)   # u8: bit-exact, like every resize test
    finally:
        lib.cid_destroy(handle)


@pytest.mark.gpu
def test_static_scratch_resize_rejects_sizes_beyond_the_array(dev):
    lib = _lib.load()
    x = torch.zeros(16, dtype=torch.uint8, device=dev)   # never read: the limits are checked before any launch
    # intermediate image B*in_h*out_w*C beyond CID_RESIZE_STATIC_TMP_BYTES
    B, ih, iw, oh, ow, c = 64, 512, 512, 256, 256, 3
    assert B * ih * ow * c > TMP_BYTES
    assert lib.cid_resize_bicubic_u8_static(0, _p(x), _p(x), B, ih, iw, oh, ow, c, None) != 0
    assert b"exceeds the static limit" in lib.cid_last_error()
    # coefficient tables beyond CID_RESIZE_STATIC_TABLE_BYTES (a very wide kernel: 1000x down-scale along x)
    assert lib.cid_resize_bicubic_u8_static(0, _p(x), _p(x), 1, 4, 64000, 4, 64, 1, None) != 0
    assert b"exceed the static limit" in lib.cid_last_error()
    assert lib.cid_resize_bicubic_u8_static(0, None, _p(x), 1, 4, 4, 4, 4, 3, None) != 0
