"""Mint golden vectors from the reference itself (run in the build container):

    python -m oracle.make_golden

Imports the unmodified /root/reference code (oracle/ref_import.py), runs it on seeded inputs and
writes small fixtures to tests/golden/. The fixtures travel to the GPU box; the reference does not.
Library versions are recorded in tests/golden/VERSIONS.json.
"""
from __future__ import annotations

import io
import json
from pathlib import Path

import numpy as np
import torch

from oracle import denoise_oracle as orc
from oracle.ref_import import load_reference_app, load_reference_noiser

GOLDEN = Path(__file__).resolve().parent.parent / "tests" / "golden"

# (name, weight seed, weight gain, data seed, B, H, W)
FORWARD_CASES = [
    ("fwd_s0_g1_b2_16x24", 0, 1.0, 11, 2, 16, 24),
    ("fwd_s1_g2_b1_32x32", 1, 2.0, 12, 1, 32, 32),
    ("fwd_s2_g3_b3_8x40", 2, 3.0, 13, 3, 8, 40),
    ("fwd_config1_s0_g1_b4_64x64", 0, 1.0, 1234, 4, 64, 64),   # BASELINE.json configs[0]
]
# (name, weight seed, gain, data seed, H, W) — odd sizes exercise get_padding / crop
ROUTE_CASES = [
    ("route_s0_g2_30x27", 0, 2.0, 21, 30, 27),
    ("route_s3_g2_33x46", 3, 2.0, 22, 33, 46),
    ("route_s3_g1_64x64", 3, 1.0, 23, 64, 64),
]


def synthetic_noisy_u8(seed: int, B: int, H: int, W: int) -> np.ndarray:
    """Smooth random field + N(0, 0.1^2), clipped and truncated to u8."""
    g = torch.Generator().manual_seed(seed)
    base = torch.rand(B, 3, max(H // 8, 2), max(W // 8, 2), generator=g)
    clean = torch.nn.functional.interpolate(base, size=(H, W), mode="bicubic", align_corners=False).clamp(0, 1)
    noisy = (clean + 0.1 * torch.randn(B, 3, H, W, generator=g)).clamp(0, 1)
    return (noisy * 255).to(torch.uint8).permute(0, 2, 3, 1).contiguous().numpy()


def main() -> None:
    GOLDEN.mkdir(parents=True, exist_ok=True)
    app = load_reference_app()
    import PIL
    import torchvision
    from PIL import Image
    versions = {"torch": torch.__version__, "torchvision": torchvision.__version__, "numpy": np.__version__,
                "pillow": PIL.__version__, "reference": "flawyer/Celebrity-Image-Denoiser @ /root/reference"}

    # ---- generator forward: app.DenoiseGenerator with a seeded state_dict ----
    for name, wseed, gain, dseed, B, H, W in FORWARD_CASES:
        sd = orc.init_state_dict(wseed, gain)
        net = app.DenoiseGenerator()
        net.load_state_dict(sd, strict=True)
        net.eval()
        u8 = synthetic_noisy_u8(dseed, B, H, W)
        x = orc.preprocess_u8(u8, (0, 0, 0, 0))
        with torch.no_grad():
            y = net(x)
        np.savez_compressed(GOLDEN / f"{name}.npz", x=x.numpy(), y=y.numpy(),
                            meta=np.array([wseed, dseed, B, H, W], dtype=np.int64), gain=np.float64(gain))
        print(name, "out range", float(y.min()), float(y.max()))

    # ---- full route: POST /enhance?model=denoise through the reference's FastAPI app ----
    from fastapi.testclient import TestClient
    client = TestClient(app.app)
    for name, wseed, gain, dseed, H, W in ROUTE_CASES:
        sd = orc.init_state_dict(wseed, gain)
        app.PT_MODELS["denoise"].load_state_dict(sd, strict=True)
        u8 = synthetic_noisy_u8(dseed, 1, H, W)[0]
        buf = io.BytesIO()
        Image.fromarray(u8).save(buf, format="PNG")
        r = client.post("/enhance?model=denoise", files={"file": ("in.png", buf.getvalue(), "image/png")})
        assert r.status_code == 200, r.text
        import base64
        out = np.array(Image.open(io.BytesIO(base64.b64decode(r.json()["denoised_image_base64"]))).convert("RGB"))
        pad = app.get_padding(Image.fromarray(u8), 4, 1)
        np.savez_compressed(GOLDEN / f"{name}.npz", image=u8, out=out, padding=np.array(pad, dtype=np.int64),
                            meta=np.array([wseed, dseed, H, W], dtype=np.int64), gain=np.float64(gain))
        print(name, "pad", pad, "out mean", out.mean())

    # ---- get_padding table ----
    rows = []
    for w in range(1, 40):
        for h in (1, 2, 3, 4, 5, 17, 30, 64):
            rows.append((w, h) + tuple(app.get_padding(Image.new("RGB", (w, h)), 4, 1)))
    np.savez_compressed(GOLDEN / "get_padding.npz", table=np.array(rows, dtype=np.int64))

    # ---- noiser: reference functions under np.random.seed, with the draws captured ----
    ng = load_reference_noiser()
    for i, (H, W) in enumerate([(32, 32), (17, 23), (64, 48)]):
        img = synthetic_noisy_u8(100 + i, 1, H, W)[0]
        np.random.seed(500 + i)
        ref_g = ng.add_gaussian_noise(img)               # mean 0, sigma 25
        np.random.seed(500 + i)
        noise = np.random.normal(0, 25, img.shape)       # the same draw the call above made
        np.random.seed(600 + i)
        ref_sp = ng.add_salt_pepper_noise(img)           # 2 % + 2 %
        np.random.seed(600 + i)
        n_salt, n_pepper = int(img.size * 0.02), int(img.size * 0.02)
        cs = [np.random.randint(0, d - 1, n_salt) for d in img.shape]
        cp = [np.random.randint(0, d - 1, n_pepper) for d in img.shape]
        np.savez_compressed(GOLDEN / f"noiser_{H}x{W}.npz", image=img, gaussian_noise=noise, gaussian_out=ref_g,
                            salt_rc=np.stack([cs[0], cs[1]], 1).astype(np.int32),
                            pepper_rc=np.stack([cp[0], cp[1]], 1).astype(np.int32), sp_out=ref_sp)

    # ---- PIL bicubic resize (noise_generation.py:61) ----
    for i, ((H, W), (oh, ow)) in enumerate([((37, 53), (32, 32)), ((20, 20), (40, 48)), ((150, 100), (64, 64)),
                                            ((64, 64), (64, 64)), ((48, 31), (16, 64))]):
        img = synthetic_noisy_u8(200 + i, 1, H, W)[0]
        out = np.array(Image.fromarray(img).resize((ow, oh), resample=Image.Resampling.BICUBIC))
        np.savez_compressed(GOLDEN / f"resize_{H}x{W}_to_{oh}x{ow}.npz", image=img, out=out)

    (GOLDEN / "VERSIONS.json").write_text(json.dumps(versions, indent=1) + "\n")
    print("golden vectors written to", GOLDEN)


if __name__ == "__main__":
    main()
