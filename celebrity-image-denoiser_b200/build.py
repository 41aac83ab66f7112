"""Build the C-ABI CUDA library in-tree with nvcc for sm_100a.

The library (csrc/libcid_b200.so) is the product: Python only binds it through ctypes. nvcc
cross-compiles without a GPU, so this runs in the CPU-only build container as well.
"""
from __future__ import annotations

import os
import shutil
import subprocess
from pathlib import Path

CSRC = Path(__file__).resolve().parent / "csrc"
LIB_PATH = CSRC / "libcid_b200.so"
SOURCES = ["cid_api.cu", "noiser.cu", "resize.cu"]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "-shared", "-Xcompiler", "-fPIC",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found; cannot build libcid_b200.so")


def needs_build() -> bool:
    if not LIB_PATH.exists():
        return True
    t = LIB_PATH.stat().st_mtime
    deps = list(CSRC.glob("*.cu")) + list(CSRC.glob("*.cuh")) + list(CSRC.glob("*.h"))
    deps.append(CSRC.parent.parent / "include" / "cid_b200.h")
    return any(d.stat().st_mtime > t for d in deps)


def source_sha() -> str:
    """Hash of the sources libcid_b200.so is built from (nvcc's output is not byte-reproducible, so profiles name the
    build they were captured with by its sources, not by the binary)."""
    import hashlib
    h = hashlib.sha256()
    deps = sorted(list(CSRC.glob("*.cu")) + list(CSRC.glob("*.cuh")) + list(CSRC.glob("*.h")))
    deps.append(CSRC.parent.parent / "include" / "cid_b200.h")
    for d in deps:
        h.update(d.name.encode())
        h.update(d.read_bytes())
    return h.hexdigest()[:16]


def build_library(force: bool = False, verbose: bool = False) -> Path:
    """Compile csrc/*.cu into csrc/libcid_b200.so (no-op if up to date)."""
    if not force and not needs_build():
        return LIB_PATH
    cmd = [_nvcc(), *NVCC_FLAGS, "-o", str(LIB_PATH), *[str(CSRC / s) for s in SOURCES]]
    if verbose:
        cmd.insert(1, "-Xptxas")
        cmd.insert(2, "-v")
        print(" ".join(cmd))
    res = subprocess.run(cmd, cwd=str(CSRC), capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"nvcc failed:\n{res.stdout}\n{res.stderr}")
    if verbose:
        print(res.stderr)
    return LIB_PATH


if __name__ == "__main__":
    print(build_library(force=True, verbose=True))
