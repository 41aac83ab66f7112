"""CPU tests: the ESRGAN / SRGAN oracle against golden vectors minted from the reference's own modules
and FastAPI route (oracle/make_golden_sr.py), and the host mirrors' state_dict surface."""
import numpy as np
import pytest
import torch

from cid_b200.sr_generators import ESRGANGenerator, SRGANGenerator
from oracle import sr_oracle as so
from oracle.ref_import import load_reference_app, reference_available


@pytest.mark.parametrize("model", ["esrgan", "srgan"])
@pytest.mark.parametrize("seed", [0, 1])
def test_forward_matches_reference_module(golden_dir, model, seed):
    z = np.load(golden_dir / f"sr_{model}_fwd_s{seed}.npz")
    sd = so.init_state_dict(model, int(z["seed"]), float(z["gain"]))
    y = so.forward(model, sd, torch.from_numpy(z["x"]))
    assert y.shape == z["y"].shape
    assert float((y - torch.from_numpy(z["y"])).abs().max()) <= 2e-5


@pytest.mark.parametrize("model,name", [("esrgan", "18x23"), ("esrgan", "16x16"), ("srgan", "18x23"), ("srgan", "16x16")])
def test_route_matches_reference_fastapi_route(golden_dir, model, name):
    z = np.load(golden_dir / f"sr_{model}_route_{name}.npz")
    sd = so.init_state_dict(model, int(z["seed"]), 1.0)
    fn = so.esrgan_enhance_u8 if model == "esrgan" else so.srgan_enhance_u8
    out = fn(sd, z["img"][None])[0]
    assert out.shape == z["out"].shape
    d = np.abs(out.astype(np.int32) - z["out"].astype(np.int32))
    assert d.max() <= 1 and (d != 0).mean() < 1e-3  # a truncation boundary may flip between CPU builds


def test_esrgan_route_quirk_crop_of_unpadded_output(golden_dir):
    """18x23 is not a multiple of 4: the route crops the UN-padded ESRGAN output at (l,t) = (0,1), so the last
    row of the returned image is PIL's black fill (backend/app.py:473-479)."""
    z = np.load(golden_dir / "sr_esrgan_route_18x23.npz")
    assert so.get_padding(23, 18, 4, 1) == (0, 1, 1, 1)
    assert not z["out"][-1].any()


@pytest.mark.parametrize("cls,model", [(ESRGANGenerator, "esrgan"), (SRGANGenerator, "srgan")])
def test_state_dict_surface(cls, model):
    net = cls()
    sd = so.init_state_dict(model, 0)
    assert set(net.state_dict().keys()) == set(sd.keys())
    for k, v in net.state_dict().items():
        assert tuple(v.shape) == tuple(sd[k].shape), k
    net.load_state_dict(sd, strict=True)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        net.eval()(torch.zeros(1, 3, 8, 8))


@pytest.mark.skipif(not reference_available(), reason="reference only exists in the build container")
@pytest.mark.parametrize("cls,model", [(ESRGANGenerator, "esrgan"), (SRGANGenerator, "srgan")])
def test_live_reference_state_dict_is_interchangeable(cls, model):
    app = load_reference_app()
    ref = getattr(app, cls.__name__)()
    ours = cls()
    ours.load_state_dict(ref.state_dict(), strict=True)
    ref.load_state_dict(ours.state_dict(), strict=True)
    x = torch.rand(1, 3, 12, 10)
    ref.eval()
    with torch.no_grad():
        assert float((ref(x) - so.forward(model, ref.state_dict(), x)).abs().max()) <= 2e-5
