"""Import the UNMODIFIED reference (build container only; /root/reference is absent on GPU boxes).

backend/app.py imports matplotlib unconditionally (app.py:10) and it is not installed here, so it
is stubbed in sys.modules (it is used only by the out-of-scope plotting helpers, app.py:294-313).
TensorFlow is optional in the reference (app.py:15-19) and absent here. The reference's weight
files are not in the tree, so its own load loop warns and keeps random init (app.py:333-336).
"""
from __future__ import annotations

import importlib
import os
import sys
from pathlib import Path
from unittest import mock

REFERENCE_ROOT = Path(os.environ.get("CID_REFERENCE_ROOT", "/root/reference"))


def reference_available() -> bool:
    return (REFERENCE_ROOT / "backend" / "app.py").exists()


def load_reference_app():
    """Returns the reference's `app` module (backend/app.py), imported once."""
    if "app" in sys.modules and getattr(sys.modules["app"], "__file__", "").startswith(str(REFERENCE_ROOT)):
        return sys.modules["app"]
    if not reference_available():
        raise RuntimeError(f"reference not found under {REFERENCE_ROOT}")
    if "matplotlib" not in sys.modules:
        mpl, plt = mock.MagicMock(), mock.MagicMock()
        plt.subplots = lambda *a, **k: (mock.MagicMock(), mock.MagicMock())  # `fig, axs = plt.subplots(...)`
        mpl.pyplot = plt
        sys.modules["matplotlib"], sys.modules["matplotlib.pyplot"] = mpl, plt
    backend = str(REFERENCE_ROOT / "backend")
    sys.path.insert(0, backend)
    try:
        return importlib.import_module("app")
    finally:
        sys.path.remove(backend)


def load_reference_noiser():
    """trainingcode/denoise_gan_code/noise_generation.py (numpy + PIL only; has a __main__ guard)."""
    d = str(REFERENCE_ROOT / "backend" / "trainingcode" / "denoise_gan_code")
    sys.path.insert(0, d)
    try:
        return importlib.import_module("noise_generation")
    finally:
        sys.path.remove(d)
