"""Minimal pure-Python reader for the HDF5 weight store inside a Keras 3 `.keras` archive.

The reference loads its only shipped trained generator with
`tf.keras.models.load_model("weights/cgan_epoch_500.keras")` (backend/app.py:338-345). Neither
TensorFlow nor h5py exists on the B200 boxes, so the model-load entry point of the Keras backend is
restated here: a `.keras` file is a zip of `config.json`, `metadata.json` and `model.weights.h5`;
the `.h5` written by Keras 3 / h5py defaults is "classic" HDF5 — superblock v0, symbol-table groups
(v1 B-trees + local heaps), v1 object headers, contiguous (or compact) uncompressed
little-endian datasets. Only that subset is parsed; anything else raises.
"""
from __future__ import annotations

import io
import json
import struct
import zipfile
from typing import Dict, Tuple

import numpy as np

_SIG = b"\x89HDF\r\n\x1a\n"
_UNDEF = 0xFFFFFFFFFFFFFFFF


class H5Error(RuntimeError):
    pass


class _H5:
    def __init__(self, data: bytes):
        self.d = data
        if data[:8] != _SIG:
            raise H5Error("not an HDF5 file")
        if data[8] != 0:
            raise H5Error(f"unsupported HDF5 superblock version {data[8]} (only v0)")
        if data[13] != 8 or data[14] != 8:
            raise H5Error("only 8-byte offsets/lengths are supported")
        self.base = self.u64(24)
        # root group symbol-table entry at offset 56
        self.root_header = self.u64(56 + 8)

    def u16(self, o): return struct.unpack_from("<H", self.d, o)[0]
    def u32(self, o): return struct.unpack_from("<I", self.d, o)[0]
    def u64(self, o): return struct.unpack_from("<Q", self.d, o)[0]

    # ---- object headers (version 1) ------------------------------------------------------
    def messages(self, addr: int):
        d = self.d
        if d[addr] != 1:
            raise H5Error(f"unsupported object header version {d[addr]} at {addr}")
        nmsg = self.u16(addr + 2)
        size = self.u32(addr + 8)
        blocks = [(addr + 16, size)]
        out = []
        while blocks and len(out) < nmsg:
            o, left = blocks.pop(0)
            end = o + left
            while o + 8 <= end and len(out) < nmsg:
                mtype, msize = self.u16(o), self.u16(o + 2)
                body = o + 8
                if mtype == 0x10:  # continuation
                    blocks.append((self.u64(body), self.u64(body + 8)))
                out.append((mtype, body, msize))
                o = body + msize
        return out

    # ---- groups: symbol table -> {name: object header address} ---------------------------
    def group_entries(self, addr: int) -> Dict[str, int]:
        for mtype, body, _ in self.messages(addr):
            if mtype == 0x11:
                return self._walk_btree(self.u64(body), self.u64(body + 8))
        return {}

    def is_group(self, addr: int) -> bool:
        return any(m[0] == 0x11 for m in self.messages(addr))

    def _heap_data(self, heap: int) -> int:
        if self.d[heap:heap + 4] != b"HEAP":
            raise H5Error("bad local heap signature")
        return self.u64(heap + 24)

    def _walk_btree(self, node: int, heap: int) -> Dict[str, int]:
        d = self.d
        if d[node:node + 4] != b"TREE":
            raise H5Error("bad B-tree signature")
        level, used = d[node + 5], self.u16(node + 6)
        out: Dict[str, int] = {}
        o = node + 24
        for i in range(used):
            child = self.u64(o + 8 + i * 16)
            if level > 0:
                out.update(self._walk_btree(child, heap))
            else:
                out.update(self._snod(child, heap))
        return out

    def _snod(self, addr: int, heap: int) -> Dict[str, int]:
        d = self.d
        if d[addr:addr + 4] != b"SNOD":
            raise H5Error("bad symbol node signature")
        n = self.u16(addr + 6)
        names = self._heap_data(heap)
        out = {}
        for i in range(n):
            e = addr + 8 + i * 40
            no = names + self.u64(e)
            end = d.index(b"\0", no)
            out[d[no:end].decode()] = self.u64(e + 8)
        return out

    # ---- datasets ------------------------------------------------------------------------
    def dataset(self, addr: int) -> np.ndarray:
        shape = dtype = None
        raw = None
        for mtype, body, msize in self.messages(addr):
            d = self.d
            if mtype == 0x01:
                ver, rank = d[body], d[body + 1]
                o = body + (8 if ver == 1 else 4)
                shape = tuple(self.u64(o + 8 * i) for i in range(rank))
            elif mtype == 0x03:
                cls = d[body] & 0x0F
                bits0 = d[body + 1]
                size = self.u32(body + 4)
                if bits0 & 1:
                    raise H5Error("big-endian datasets are not supported")
                if cls == 1 and size in (2, 4, 8):
                    dtype = np.dtype(f"<f{size}")
                elif cls == 0 and size in (1, 2, 4, 8):
                    dtype = np.dtype(f"<{'i' if bits0 & 8 else 'u'}{size}")
                else:
                    raise H5Error(f"unsupported datatype class {cls} size {size}")
            elif mtype == 0x0B:
                raise H5Error("filtered (compressed) datasets are not supported")
            elif mtype == 0x08:
                ver = d[body]
                if ver != 3:
                    raise H5Error(f"unsupported data layout version {ver}")
                lclass = d[body + 1]
                if lclass == 1:
                    a, n = self.u64(body + 2), self.u64(body + 10)
                    raw = b"" if a == _UNDEF else d[self.base + a:self.base + a + n]
                elif lclass == 0:
                    n = self.u16(body + 2)
                    raw = d[body + 4:body + 4 + n]
                else:
                    raise H5Error("chunked datasets are not supported")
        if shape is None or dtype is None or raw is None:
            raise H5Error("object is not a simple dataset")
        count = int(np.prod(shape)) if shape else 1
        return np.frombuffer(raw, dtype=dtype, count=count).reshape(shape).copy()

    def walk(self, addr: int, prefix: str, out: Dict[str, np.ndarray]):
        for name, child in sorted(self.group_entries(addr).items()):
            path = f"{prefix}/{name}" if prefix else name
            if self.is_group(child):
                self.walk(child, path, out)
            else:
                out[path] = self.dataset(child)


def read_h5_datasets(data: bytes) -> Dict[str, np.ndarray]:
    """All datasets of an HDF5 image as {"group/sub/name": ndarray}."""
    h = _H5(data)
    out: Dict[str, np.ndarray] = {}
    h.walk(h.root_header, "", out)
    return out


def read_keras_archive(path) -> Tuple[dict, Dict[str, np.ndarray]]:
    """(`config.json` as a dict, {dataset path: ndarray}) of a Keras 3 `.keras` file."""
    with zipfile.ZipFile(path) as z:
        config = json.loads(z.read("config.json"))
        weights = read_h5_datasets(z.read("model.weights.h5"))
    return config, weights
