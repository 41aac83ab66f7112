"""B200-native generator forward for Celebrity-Image-Denoiser (import as `cid_b200`).

Public surface (mirrors the reference's names for the denoise path, backend/app.py):
    DenoiseGenerator      nn.Module with the reference's state_dict keys; forward on CUDA only
    load_state_safely     checkpoint loader with the reference's key-candidate / "module." rules
    get_padding, MODEL_CFG
    enhance_image / enhance_batch_u8 / enhance_batch_host   image-in -> image-out
    patch_app(app)        swap the new generator into the reference's PT_MODELS registry
    MicroBatcher          concurrent single-image requests of the upload route share one batched forward
    keras_cgan.load_model / KerasCGANGenerator.predict   the Keras cGAN backend (backend/app.py:338-345, 458)
    ESRGANGenerator, SRGANGenerator   the other two nn.Modules of PT_MODELS (backend/app.py:146-217)
    DenoiseDiscriminator, DenoiseGANTrainStep   one iteration of trainingcode/denoise_gan_code/training.py:403-427
"""
from .generator import DenoiseGenerator, STATE_DICT_SHAPES  # noqa: F401
from .loader import MODEL_CFG, get_padding, load_state_safely, patch_app  # noqa: F401
from .pipeline import enhance_batch_host, enhance_batch_u8, enhance_image  # noqa: F401
from .batcher import MicroBatcher  # noqa: F401
from . import keras_cgan  # noqa: F401,E402
from .keras_cgan import KerasCGANGenerator  # noqa: F401,E402
from .sr_generators import ESRGANGenerator, SRGANGenerator  # noqa: F401,E402
from .training import DenoiseDiscriminator, DenoiseGANTrainStep  # noqa: F401,E402
