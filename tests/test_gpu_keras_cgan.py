"""GPU parity of the Keras cGAN generator path (cid_create_model("keras_cgan")) against
oracle/keras_cgan_oracle.py, through the C ABI. North-star tolerances: float max-abs <= 2e-2 on
[-1,1], PSNR divergence < 0.1 dB, uint8 within +-1 LSB — asserted on the reference's own TRAINED
weights (tests/golden/keras_cgan_epoch500.npz) and on seeded synthetic variables."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from cid_b200.keras_cgan import KerasCGANGenerator
from oracle import keras_cgan_oracle as ko

MAX_ABS = 2e-2
PSNR_DIV_DB = 0.1
LSB = 1


@pytest.fixture(scope="module")
def dev():
    assert torch.cuda.is_available(), "gpu tests need a CUDA device"
    return torch.device("cuda:0")


@pytest.fixture(scope="module")
def trained(golden_dir):
    z = np.load(golden_dir / "keras_cgan_epoch500.npz")
    return {k: z[k] for k in ko.PARAM_SHAPES}, z["face_u8"]


_nets = {}


def synth_net(dev, seed):
    if seed not in _nets:
        var = ko.init_params(seed)
        _nets[seed] = (KerasCGANGenerator(var, dev), var)
    return _nets[seed]


def psnr_f(a, b):  # data range 2.0, the reference's convention (denoise_gan_code/training.py:382)
    return 10 * np.log10(4.0 / max(float(((a - b) ** 2).mean()), 1e-20))


def test_trained_weights_float_and_u8_parity(dev, trained):
    var, face = trained
    net = KerasCGANGenerator(var, dev)
    x = ((face.astype(np.float32) / 255.0) - 0.5) / 0.5
    y = net.predict(x)
    ref = ko.predict(var, x)
    assert y.shape == ref.shape == (1, 256, 256, 3) and y.dtype == np.float32
    assert float(np.abs(y - ref).max()) <= MAX_ABS
    assert abs(psnr_f(y, x) - psnr_f(ref, x)) < PSNR_DIV_DB
    out = net.forward_u8(torch.from_numpy(face).to(dev), (0, 0, 0, 0)).cpu().numpy()
    ref8 = ko.enhance_u8(var, face)
    assert int(np.abs(out.astype(np.int32) - ref8.astype(np.int32)).max()) <= LSB
    assert net.kernel_launches() == 10  # 5 hand-written kernels per forward, nothing else


def test_trained_weights_layer_by_layer(dev, trained):
    var, face = trained
    net = KerasCGANGenerator(var, dev)
    x = ((face[:, :64, :96].astype(np.float32) / 255.0) - 0.5) / 0.5
    net.predict(x)
    xt = torch.from_numpy(x).permute(0, 3, 1, 2).contiguous()
    for i, name in enumerate(["c0", "c1", "c2", "c3"]):
        ref = ko.forward_nchw(var, xt, upto=i + 1).permute(0, 2, 3, 1)
        got = net.debug_activation(name, 1, 64, 96)
        assert got.shape == ref.shape
        # bf16 storage: half an ulp of the value plus the propagated error of the layers before
        tol = 0.01 * float(ref.abs().max()) + 0.004 * (i + 1)
        assert float((got - ref).abs().max()) <= tol, name


@pytest.mark.parametrize("B,H,W,seed", [(1, 4, 4, 0), (2, 8, 12, 1), (1, 32, 32, 0), (3, 20, 36, 2), (2, 64, 48, 1),
                                         (1, 132, 128, 3), (1, 256, 256, 0), (4, 128, 128, 2)])
def test_synthetic_weights_vs_oracle_shapes(dev, B, H, W, seed):
    net, var = synth_net(dev, seed)
    x = np.random.default_rng(100 + seed).uniform(-1, 1, (B, H, W, 3)).astype(np.float32)
    y = net.predict(x)
    ref = ko.predict(var, x)
    assert float(np.abs(y - ref).max()) <= MAX_ABS
    assert abs(psnr_f(y, x) - psnr_f(ref, x)) < PSNR_DIV_DB


@pytest.mark.parametrize("H,W", [(1, 1), (3, 5), (30, 27), (33, 46), (127, 130)])
def test_route_u8_odd_sizes_pad_and_crop(dev, H, W):
    net, var = synth_net(dev, 1)
    img = np.random.default_rng(H * 131 + W).integers(0, 256, (2, H, W, 3), dtype=np.uint8)
    out = net.forward_u8(torch.from_numpy(img).to(dev), ko.get_padding(W, H)).cpu().numpy()
    ref = ko.enhance_u8(var, img)
    assert out.shape == ref.shape
    assert int(np.abs(out.astype(np.int32) - ref.astype(np.int32)).max()) <= LSB


def test_batch_and_chunk_invariance_bit_exact(dev):
    net, var = synth_net(dev, 2)
    img = torch.from_numpy(np.random.default_rng(9).integers(0, 256, (6, 40, 56, 3), dtype=np.uint8)).to(dev)
    whole = net.forward_u8(img).clone()
    one = torch.cat([net.forward_u8(img[i:i + 1]) for i in range(6)])
    assert torch.equal(whole, one)
    chunked = KerasCGANGenerator(var, dev).set_option("chunk", 4)
    assert torch.equal(whole, chunked.forward_u8(img))


def test_error_behaviour(dev):
    net, var = synth_net(dev, 0)
    with pytest.raises(RuntimeError, match="multiples of 4"):
        net.predict_device(torch.zeros((1, 6, 8, 3), device=dev))
    with pytest.raises(RuntimeError):
        net.predict_device(torch.zeros((1, 8, 8, 4), device=dev))
    bad = dict(var)
    bad["conv2d_1.kernel"] = np.zeros((4, 4, 64, 64), np.float32)
    with pytest.raises(RuntimeError, match="conv2d_1.kernel"):
        KerasCGANGenerator(bad, dev).predict(np.zeros((1, 8, 8, 3), np.float32))
    # the denoise entry point refuses a keras_cgan handle (and vice versa) instead of mis-running it
    import ctypes as C
    eng = net._eng()
    rc = eng.lib.cid_forward_f32_nchw(eng.handle, C.c_void_p(1), C.c_void_p(1), 1, 8, 8, C.c_void_p(1), 0, None)
    assert rc != 0 and b"denoise" in eng.lib.cid_last_error()
