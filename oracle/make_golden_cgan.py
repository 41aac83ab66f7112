"""Build-container script: fixtures for the Keras cGAN path (needs /root/reference).

  tests/golden/keras_cgan_epoch500.npz   the 22 trained variables of the reference's shipped
        backend/weights/cgan_epoch_500.keras (read with the product's own HDF5 reader; fp32, Keras layouts)
        + `face_u8`: the reference's only sample image (backend/trainingcode/srgan_code/input_noise.jpg, 64x64)
          resized to 256x256 with PIL bicubic + `oracle_out_u8`: oracle/keras_cgan_oracle.enhance_u8 of it.
TensorFlow is not installed anywhere, so there is no reference OUTPUT to record: parity for this path is
"unpinned" (see oracle/keras_cgan_oracle.py); the fixture pins the weights, the input and the oracle's own
output (so a later change of torch's CPU kernels cannot silently move the checker).
"""
import sys
from pathlib import Path

import numpy as np
from PIL import Image

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from cid_b200.keras_cgan import check_config, variables_from_h5  # noqa: E402
from cid_b200.keras_h5 import read_keras_archive  # noqa: E402
from oracle import keras_cgan_oracle as ko  # noqa: E402

REF = Path("/root/reference/backend")

if __name__ == "__main__":
    config, datasets = read_keras_archive(REF / "weights" / "cgan_epoch_500.keras")
    check_config(config)
    var = variables_from_h5(datasets)
    assert sum(v.size for v in var.values()) == 530755
    img = Image.open(REF / "trainingcode" / "srgan_code" / "input_noise.jpg").convert("RGB").resize((256, 256), Image.BICUBIC)
    face = np.asarray(img, dtype=np.uint8)[None]
    out = ko.enhance_u8(var, face)
    np.savez_compressed(ROOT / "tests" / "golden" / "keras_cgan_epoch500.npz", face_u8=face, oracle_out_u8=out, **var)
    print("wrote keras_cgan_epoch500.npz", {k: v.shape for k, v in var.items()})
