"""Build-container script: golden vectors for the speckle / poisson / uniform noise types, from the reference's
own functions (trainingcode/denoise_gan_code/noise_generation.py:24-39) under np.random.seed, with the draws
captured by replaying the same seed."""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from oracle.ref_import import load_reference_noiser  # noqa: E402

GOLDEN = ROOT / "tests" / "golden"

if __name__ == "__main__":
    ng = load_reference_noiser()
    for (H, W) in ((32, 32), (17, 23)):
        img = np.random.default_rng(H * 31 + W).integers(0, 256, (H, W, 3), dtype=np.uint8)
        img[0, :4] = 255  # bright pixels: Poisson counts above 255 wrap in the reference's uint8 cast
        np.random.seed(7); spk = ng.add_speckle_noise(img)
        np.random.seed(7); spk_draws = np.random.normal(0, 0.1, img.shape)
        np.random.seed(8); poi = ng.add_poisson_noise(img)
        np.random.seed(8); poi_draws = np.random.poisson(img)
        np.random.seed(9); uni = ng.add_uniform_noise(img)
        np.random.seed(9); uni_draws = np.random.uniform(0, 25, img.shape)
        np.savez_compressed(GOLDEN / f"noise_more_{H}x{W}.npz", image=img, speckle_draws=spk_draws, speckle_out=spk,
                            poisson_counts=poi_draws.astype(np.int64), poisson_out=poi, uniform_draws=uni_draws,
                            uniform_out=uni)
        print(H, W, "poisson counts > 255:", int((poi_draws > 255).sum()))
