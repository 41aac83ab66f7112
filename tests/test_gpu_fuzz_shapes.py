"""Seeded shape fuzz of all four generators through the C ABI: partial tiles, odd tile counts for the CTA-pair
kernels, batch sizes that do not divide the SM count. Same north-star tolerances as the fixed-shape tests."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from cid_b200 import DenoiseGenerator, ESRGANGenerator, KerasCGANGenerator, SRGANGenerator
from oracle import denoise_oracle as orc
from oracle import keras_cgan_oracle as ko
from oracle import sr_oracle as so

MAX_ABS = 2e-2


@pytest.fixture(scope="module")
def dev():
    assert torch.cuda.is_available()
    return torch.device("cuda:0")


def _shapes(seed, n, mult, hi):
    rng = np.random.default_rng(seed)
    out = []
    for _ in range(n):
        B = int(rng.integers(1, 6))
        H = int(rng.integers(1, hi // mult + 1)) * mult
        W = int(rng.integers(1, hi // mult + 1)) * mult
        out.append((B, H, W))
    return out


def test_denoise_fuzz(dev):
    sd = orc.init_state_dict(9, 2.0)
    net = DenoiseGenerator()
    net.load_state_dict(sd, strict=True)
    net = net.to(dev).eval()
    for i, (B, H, W) in enumerate(_shapes(101, 14, 4, 96)):
        x = torch.rand(B, 3, H, W, generator=torch.Generator().manual_seed(i)) * 2 - 1
        with torch.no_grad():
            ref = orc.forward(sd, x)
        err = float((net(x.to(dev)).cpu() - ref).abs().max())
        assert err <= MAX_ABS, (B, H, W, err)


def test_denoise_route_fuzz_odd_sizes(dev):
    sd = orc.init_state_dict(10, 2.0)
    net = DenoiseGenerator()
    net.load_state_dict(sd, strict=True)
    net = net.to(dev).eval()
    rng = np.random.default_rng(5)
    for _ in range(10):
        B, H, W = int(rng.integers(1, 4)), int(rng.integers(1, 70)), int(rng.integers(1, 70))
        img = rng.integers(0, 256, (B, H, W, 3), dtype=np.uint8)
        out = net.forward_u8(torch.from_numpy(img).to(dev), orc.get_padding(W, H)).cpu().numpy()
        ref = orc.enhance_u8(sd, img)
        assert int(np.abs(out.astype(np.int32) - ref.astype(np.int32)).max()) <= 1, (B, H, W)


@pytest.mark.parametrize("model,cls", [("esrgan", ESRGANGenerator), ("srgan", SRGANGenerator)])
def test_sr_fuzz(dev, model, cls):
    sd = so.init_state_dict(model, 6)
    net = cls()
    net.load_state_dict(sd, strict=True)
    net = net.to(dev).eval()
    for i, (B, H, W) in enumerate(_shapes(202, 10, 1, 56 if model == "esrgan" else 40)):
        x = torch.rand(B, 3, H, W, generator=torch.Generator().manual_seed(50 + i))
        if model == "srgan":
            x = x * 2 - 1
        ref = so.forward(model, sd, x)
        err = float((net(x.to(dev)).cpu() - ref).abs().max())
        assert err <= MAX_ABS, (B, H, W, err)


def test_keras_cgan_fuzz(dev):
    var = ko.init_params(7)
    net = KerasCGANGenerator(var, dev)
    for i, (B, H, W) in enumerate(_shapes(303, 12, 4, 96)):
        x = np.random.default_rng(70 + i).uniform(-1, 1, (B, H, W, 3)).astype(np.float32)
        err = float(np.abs(net.predict(x) - ko.predict(var, x)).max())
        assert err <= MAX_ABS, (B, H, W, err)
