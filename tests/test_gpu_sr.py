"""GPU parity of the ESRGAN / SRGAN generators (cid_create_model("esrgan" | "srgan")) against
oracle/sr_oracle.py and the golden vectors minted from the reference, through the C ABI.
North-star tolerances: float max-abs <= 2e-2, PSNR divergence < 0.1 dB, uint8 within +-1 LSB."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from cid_b200.sr_generators import ESRGANGenerator, SRGANGenerator
from oracle import sr_oracle as so

MAX_ABS = 2e-2
PSNR_DIV_DB = 0.1
LSB = 1
CLS = {"esrgan": ESRGANGenerator, "srgan": SRGANGenerator}


@pytest.fixture(scope="module")
def dev():
    assert torch.cuda.is_available(), "gpu tests need a CUDA device"
    return torch.device("cuda:0")


_nets = {}


def make_net(dev, model, seed, gain=1.0):
    key = (model, seed, gain)
    if key not in _nets:
        sd = so.init_state_dict(model, seed, gain)
        net = CLS[model]()
        net.load_state_dict(sd, strict=True)
        _nets[key] = (net.to(dev).eval(), sd)
    return _nets[key]


def up(x, s):
    return torch.nn.functional.interpolate(x, scale_factor=s, mode="nearest") if s > 1 else x


@pytest.mark.parametrize("model", ["esrgan", "srgan"])
@pytest.mark.parametrize("seed", [0, 1])
def test_forward_vs_reference_golden(dev, golden_dir, model, seed):
    z = np.load(golden_dir / f"sr_{model}_fwd_s{seed}.npz")
    net, _ = make_net(dev, model, int(z["seed"]), float(z["gain"]))
    x = torch.from_numpy(z["x"])
    y = net(x.to(dev)).cpu()
    ref = torch.from_numpy(z["y"])
    assert y.shape == ref.shape
    assert float((y - ref).abs().max()) <= MAX_ABS
    tgt = up(x, net.OUT_SCALE)
    assert abs(so.psnr(y, tgt) - so.psnr(ref, tgt)) < PSNR_DIV_DB


@pytest.mark.parametrize("model", ["esrgan", "srgan"])
@pytest.mark.parametrize("B,H,W,seed", [(1, 1, 1, 0), (2, 5, 7, 1), (1, 16, 16, 0), (3, 20, 36, 2), (1, 64, 48, 1),
                                         (1, 128, 128, 3)])
def test_forward_vs_oracle_shapes(dev, model, B, H, W, seed):
    net, sd = make_net(dev, model, seed)
    x = torch.rand(B, 3, H, W, generator=torch.Generator().manual_seed(7 + seed))
    if model == "srgan":
        x = x * 2 - 1
    y = net(x.to(dev)).cpu()
    ref = so.forward(model, sd, x)
    assert y.shape == ref.shape
    assert float((y - ref).abs().max()) <= MAX_ABS
    tgt = up(x, net.OUT_SCALE)
    assert abs(so.psnr(y, tgt) - so.psnr(ref, tgt)) < PSNR_DIV_DB


@pytest.mark.parametrize("model,name", [("esrgan", "18x23"), ("esrgan", "16x16"), ("srgan", "18x23"), ("srgan", "16x16")])
def test_route_u8_vs_reference_golden(dev, golden_dir, model, name):
    z = np.load(golden_dir / f"sr_{model}_route_{name}.npz")
    net, _ = make_net(dev, model, int(z["seed"]))
    img = z["img"][None]
    H, W = img.shape[1:3]
    pad = so.get_padding(W, H, 4, 1 if model == "esrgan" else 4)
    out = net.forward_u8(torch.from_numpy(img).to(dev), pad).cpu().numpy()[0]
    assert out.shape == z["out"].shape
    assert int(np.abs(out.astype(np.int32) - z["out"].astype(np.int32)).max()) <= LSB


@pytest.mark.parametrize("model", ["esrgan", "srgan"])
@pytest.mark.parametrize("H,W", [(3, 5), (30, 27), (33, 46)])
def test_route_u8_odd_sizes_vs_oracle(dev, model, H, W):
    net, sd = make_net(dev, model, 1)
    img = np.random.default_rng(H * 977 + W).integers(0, 256, (2, H, W, 3), dtype=np.uint8)
    pad = so.get_padding(W, H, 4, 1 if model == "esrgan" else 4)
    out = net.forward_u8(torch.from_numpy(img).to(dev), pad).cpu().numpy()
    ref = (so.esrgan_enhance_u8 if model == "esrgan" else so.srgan_enhance_u8)(sd, img)
    assert out.shape == ref.shape
    assert int(np.abs(out.astype(np.int32) - ref.astype(np.int32)).max()) <= LSB


@pytest.mark.parametrize("model", ["esrgan", "srgan"])
def test_batch_and_chunk_invariance_bit_exact(dev, model):
    net, sd = make_net(dev, model, 2)
    x = torch.rand(5, 3, 24, 40, generator=torch.Generator().manual_seed(3)).to(dev)
    whole = net(x).clone()
    one = torch.cat([net(x[i:i + 1]) for i in range(5)])
    assert torch.equal(whole, one)
    other = CLS[model]()
    other.load_state_dict(sd, strict=True)
    other = other.to(dev).eval().set_option("chunk", 2)
    assert torch.equal(whole, other(x))


def test_esrgan_running_activation_after_first_blocks(dev):
    """Layer-level check through the public surface: truncating the oracle after k residual blocks must agree with
    a generator whose remaining blocks are identities (conv weights 0, BN -> 0) — exercises EPI_ADD's skip path."""
    sd = so.init_state_dict("esrgan", 4)
    for i in range(3, 8):
        b = f"residuals.{i}.block"
        sd[b + ".3.weight"].zero_(); sd[b + ".3.bias"].zero_()
        sd[b + ".4.weight"].zero_(); sd[b + ".4.bias"].zero_()   # BN(conv) == 0 -> block(x) == x
    net = ESRGANGenerator()
    net.load_state_dict(sd, strict=True)
    net = net.to(dev).eval()
    x = torch.rand(1, 3, 32, 32, generator=torch.Generator().manual_seed(11))
    y = net(x.to(dev)).cpu()
    ref = so.forward("esrgan", sd, x)
    assert float((y - ref).abs().max()) <= MAX_ABS


def test_error_behaviour(dev):
    net, _ = make_net(dev, "esrgan", 0)
    with pytest.raises(RuntimeError):
        net(torch.zeros((1, 4, 8, 8), device=dev))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        net(torch.zeros((1, 3, 8, 8)))
    net.train()
    with pytest.raises(RuntimeError, match="eval-mode"):
        net(torch.zeros((1, 3, 8, 8), device=dev))
    net.eval()
