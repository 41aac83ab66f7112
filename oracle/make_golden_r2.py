"""Round-2 golden vectors minted from the reference itself (run in the build container):

    python -m oracle.make_golden_r2

* noiser_100x100.npz: add_salt_pepper_noise (denoise_gan_code/noise_generation.py) on a batch of
  two 100x100x3 images, where size * 0.02 is an integer (600 draws; a float32 probability gives 599);
* unit_noise_{H}x{W}.npz: the [0,1]-domain functions of esrgan_code/esrgan_addNoise.py (add_gaussian, add_salt_pepper,
  add_speckle, add_uniform) run under np.random.seed with their draws captured, through the script's own u8 -> float
  -> u8 conversions (:66, :79). The module has no __main__ guard: it is imported from an empty working directory, where
  its os.walk("Clean_dataset") loop (:48-80) finds nothing to do.
"""
from __future__ import annotations

import contextlib
import importlib
import io
import os
import sys
import tempfile
from pathlib import Path

import numpy as np

from oracle.make_golden import GOLDEN, synthetic_noisy_u8
from oracle.ref_import import REFERENCE_ROOT, load_reference_noiser


def load_reference_unit_noiser():
    d = str(REFERENCE_ROOT / "backend" / "trainingcode" / "esrgan_code")
    cwd = os.getcwd()
    with tempfile.TemporaryDirectory() as tmp:
        os.chdir(tmp)
        sys.path.insert(0, d)
        try:
            with contextlib.redirect_stdout(io.StringIO()):
                return importlib.import_module("esrgan_addNoise")
        finally:
            sys.path.remove(d)
            os.chdir(cwd)


def main() -> None:
    ng = load_reference_noiser()
    imgs = synthetic_noisy_u8(300, 2, 100, 100)
    outs, salt, pepper = [], [], []
    for i in range(2):
        img = imgs[i]
        np.random.seed(700 + i)
        outs.append(ng.add_salt_pepper_noise(img))
        np.random.seed(700 + i)
        n_salt, n_pepper = int(img.size * 0.02), int(img.size * 0.02)
        assert n_salt == 600
        cs = [np.random.randint(0, d - 1, n_salt) for d in img.shape]
        cp = [np.random.randint(0, d - 1, n_pepper) for d in img.shape]
        salt.append(np.stack([cs[0], cs[1]], 1).astype(np.int32))
        pepper.append(np.stack([cp[0], cp[1]], 1).astype(np.int32))
    np.savez_compressed(GOLDEN / "noiser_100x100.npz", image=imgs, sp_out=np.stack(outs), salt_rc=np.stack(salt),
                        pepper_rc=np.stack(pepper))

    un = load_reference_unit_noiser()
    for i, (H, W) in enumerate([(24, 40), (33, 17)]):
        img_u8 = synthetic_noisy_u8(400 + i, 1, H, W)[0]
        img = np.array(img_u8) / 255.0                       # esrgan_addNoise.py:66
        to_u8 = lambda a: (a * 255).astype(np.uint8)         # :79
        rec = {"image": img_u8}
        np.random.seed(900 + i)
        rec["gaussian_out"] = to_u8(un.add_gaussian(img))
        np.random.seed(900 + i)
        rec["gaussian_draws"] = np.random.normal(0, 0.01 ** 0.5, img.shape)
        np.random.seed(910 + i)
        rec["speckle_out"] = to_u8(un.add_speckle(img))
        np.random.seed(910 + i)
        rec["speckle_draws"] = np.random.randn(*img.shape)
        np.random.seed(920 + i)
        rec["uniform_out"] = to_u8(un.add_uniform(img))
        np.random.seed(920 + i)
        rec["uniform_draws"] = np.random.uniform(-0.05, 0.05, img.shape)
        np.random.seed(930 + i)
        rec["sp_out"] = to_u8(un.add_salt_pepper(img))
        np.random.seed(930 + i)
        n = int(np.ceil(0.004 * img.size * 0.5))
        rec["salt_idx"] = np.stack([np.random.randint(0, d, n) for d in img.shape], 1).astype(np.int32)
        rec["pepper_idx"] = np.stack([np.random.randint(0, d, n) for d in img.shape], 1).astype(np.int32)
        np.savez_compressed(GOLDEN / f"unit_noise_{H}x{W}.npz", **rec)
    print("round-2 golden vectors written to", GOLDEN)


if __name__ == "__main__":
    main()
