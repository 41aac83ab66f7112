"""Importable alias of the `celebrity-image-denoiser_b200/` package directory.

The package directory carries the reference's name (with hyphens, so it cannot be named in an
`import` statement); this shim puts it on `cid_b200.__path__`, so `import cid_b200.generator`
loads `celebrity-image-denoiser_b200/generator.py`.
"""
from pathlib import Path as _Path

_PKG_DIR = _Path(__file__).resolve().parent.parent / "celebrity-image-denoiser_b200"
__path__.append(str(_PKG_DIR))

PACKAGE_DIR = _PKG_DIR
