"""The oracle (oracle/denoise_oracle.py) against the golden vectors minted from the reference itself
(oracle/make_golden.py ran /root/reference's unmodified code; fixtures in tests/golden/).
CPU only. This is what pins the oracle."""
import glob
import json
import os

import numpy as np
import pytest
import torch

from oracle import denoise_oracle as orc


def _load(golden_dir, name):
    return np.load(golden_dir / name)


FWD = ["fwd_s0_g1_b2_16x24", "fwd_s1_g2_b1_32x32", "fwd_s2_g3_b3_8x40", "fwd_config1_s0_g1_b4_64x64"]
ROUTE = ["route_s0_g2_30x27", "route_s3_g2_33x46", "route_s3_g1_64x64"]


def test_versions_recorded(golden_dir):
    v = json.loads((golden_dir / "VERSIONS.json").read_text())
    assert {"torch", "torchvision", "numpy", "pillow", "reference"} <= set(v)


def test_state_dict_keys_match_reference_layout():
    shapes = orc.state_dict_shapes()
    assert len(shapes) == 24
    assert shapes["down1.0.weight"] == (64, 3, 3, 3)
    assert shapes["up2.weight"] == (256, 128, 2, 2)      # ConvTranspose2d is [Cin, Cout, kh, kw]
    assert shapes["up1.weight"] == (128, 64, 2, 2)
    assert shapes["upconv1.2.weight"] == (3, 64, 3, 3)
    assert sum(int(np.prod(s)) for s in shapes.values()) == 1_827_587


@pytest.mark.parametrize("name", FWD)
def test_forward_matches_reference_module(golden_dir, name):
    g = _load(golden_dir, name + ".npz")
    wseed = int(g["meta"][0])
    sd = orc.init_state_dict(wseed, float(g["gain"]))
    with torch.no_grad():
        y = orc.forward(sd, torch.from_numpy(g["x"]))
    # same torch ops on the same machine class: fp32 round-off only
    assert float((y - torch.from_numpy(g["y"])).abs().max()) <= 1e-5


@pytest.mark.parametrize("name", ROUTE)
def test_route_matches_reference_fastapi_route(golden_dir, name):
    g = _load(golden_dir, name + ".npz")
    sd = orc.init_state_dict(int(g["meta"][0]), float(g["gain"]))
    h, w = g["image"].shape[:2]
    assert tuple(int(v) for v in g["padding"]) == orc.get_padding(w, h)
    out = orc.enhance_u8(sd, g["image"])[0]
    assert out.shape == g["out"].shape
    # truncating quantise: fp32 round-off may flip a value sitting on an integer boundary
    diff = np.abs(out.astype(np.int32) - g["out"].astype(np.int32))
    assert diff.max() <= 1
    assert (diff != 0).mean() < 1e-3


def test_get_padding_table(golden_dir):
    for w, h, l, t, r, b in _load(golden_dir, "get_padding.npz")["table"]:
        assert orc.get_padding(int(w), int(h)) == (l, t, r, b)
        assert (w + l + r) % 4 == 0 and (h + t + b) % 4 == 0


@pytest.mark.parametrize("name", ["noiser_32x32", "noiser_17x23", "noiser_64x48"])
def test_noiser_bit_exact(golden_dir, name):
    g = _load(golden_dir, name + ".npz")
    img = g["image"]
    assert np.array_equal(orc.gaussian_noise_apply(img, g["gaussian_noise"]), g["gaussian_out"])
    ns, npp = orc.salt_pepper_counts(img.shape, 0.02, 0.02)
    assert ns == len(g["salt_rc"]) and npp == len(g["pepper_rc"])
    assert np.array_equal(orc.salt_pepper_apply(img, g["salt_rc"], g["pepper_rc"]), g["sp_out"])
    # the reference's randint(0, dim-1) never touches the last row / column
    assert g["salt_rc"][:, 0].max() < img.shape[0] - 1 and g["salt_rc"][:, 1].max() < img.shape[1] - 1


@pytest.mark.parametrize("name", ["noise_more_32x32", "noise_more_17x23"])
def test_more_noise_types_bit_exact(golden_dir, name):
    g = _load(golden_dir, name + ".npz")
    img = g["image"]
    assert np.array_equal(orc.speckle_noise_apply(img, g["speckle_draws"]), g["speckle_out"])
    assert np.array_equal(orc.uniform_noise_apply(img, g["uniform_draws"]), g["uniform_out"])
    assert np.array_equal(orc.poisson_noise_apply(g["poisson_counts"]), g["poisson_out"])


def test_resize_bit_exact_vs_pillow(golden_dir):
    files = sorted(glob.glob(str(golden_dir / "resize_*.npz")))
    assert len(files) >= 5
    for f in files:
        g = np.load(f)
        oh, ow = g["out"].shape[:2]
        assert np.array_equal(orc.resize_bicubic_u8(g["image"], (ow, oh)), g["out"]), os.path.basename(f)


def test_oracle_against_live_reference_when_present():
    """In the build container the reference is importable: check one fresh (non-golden) case."""
    from oracle.ref_import import load_reference_app, reference_available

    if not reference_available():
        pytest.skip("/root/reference is not on this machine (GPU box)")
    app = load_reference_app()
    sd = orc.init_state_dict(7, 1.5)
    net = app.DenoiseGenerator()
    net.load_state_dict(sd, strict=True)
    net.eval()
    x = torch.rand(2, 3, 20, 12, generator=torch.Generator().manual_seed(3)) * 2 - 1
    with torch.no_grad():
        assert float((net(x) - orc.forward(sd, x)).abs().max()) <= 1e-6


def test_salt_pepper_integer_product_edge_100x100(golden_dir):
    """100x100x3 * 0.02 is exactly 600 in float64 (599 with a float32 probability); batch of two images."""
    g = _load(golden_dir, "noiser_100x100.npz")
    ns, npp = orc.salt_pepper_counts(g["image"][0].shape, 0.02, 0.02)
    assert (ns, npp) == (600, 600) == (g["salt_rc"].shape[1], g["pepper_rc"].shape[1])
    for i in range(2):
        assert np.array_equal(orc.salt_pepper_apply(g["image"][i], g["salt_rc"][i], g["pepper_rc"][i]), g["sp_out"][i])


@pytest.mark.parametrize("name", ["unit_noise_24x40", "unit_noise_33x17"])
def test_unit_domain_noiser_bit_exact(golden_dir, name):
    """esrgan_addNoise.py:7-41 through the script's u8 -> [0,1] -> u8 conversions, with the reference's own draws."""
    g = _load(golden_dir, name + ".npz")
    img = g["image"]
    for kind in ("gaussian", "speckle", "uniform"):
        assert np.array_equal(orc.unit_noise_apply(img, kind, g[kind + "_draws"]), g[kind + "_out"]), kind
    n = orc.unit_salt_pepper_count(img.size)
    assert n == len(g["salt_idx"]) == len(g["pepper_idx"])
    assert np.array_equal(orc.unit_salt_pepper_apply(img, g["salt_idx"], g["pepper_idx"]), g["sp_out"])
