"""Importable alias of the `celebrity-image-denoiser_b200/` package directory.

The package directory carries the reference's name (with hyphens, so it cannot be named in an
`import` statement). This shim points `cid_b200.__path__` at it and runs its `__init__.py` in this
module's namespace, so `import cid_b200` / `from cid_b200.generator import ...` load the files in
`celebrity-image-denoiser_b200/`.
"""
from pathlib import Path as _Path

PACKAGE_DIR = _Path(__file__).resolve().parent.parent / "celebrity-image-denoiser_b200"
__path__ = [str(PACKAGE_DIR)]
_init = PACKAGE_DIR / "__init__.py"
exec(compile(_init.read_text(), str(_init), "exec"))
