"""Build-container script: golden vectors for the ESRGAN / SRGAN oracle from the UNMODIFIED reference
modules (/root/reference/backend/app.py: ESRGANGenerator, SRGANGenerator, eval mode, seeded state_dicts)
and from the reference's own FastAPI route for model=esrgan / model=srgan."""
import base64
import io
import sys
from pathlib import Path

import numpy as np
import torch
from PIL import Image

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from oracle import sr_oracle as so  # noqa: E402
from oracle.ref_import import load_reference_app  # noqa: E402

GOLD = ROOT / "tests" / "golden"


def main():
    app = load_reference_app()
    from fastapi.testclient import TestClient
    client = TestClient(app.app)
    for model, cls in (("esrgan", app.ESRGANGenerator), ("srgan", app.SRGANGenerator)):
        for seed, gain, (B, H, W) in ((0, 1.0, (2, 16, 24)), (1, 1.5, (1, 21, 13))):
            sd = so.init_state_dict(model, seed, gain)
            net = cls()
            net.load_state_dict(sd, strict=True)
            net.eval()
            x = torch.rand(B, 3, H, W, generator=torch.Generator().manual_seed(100 + seed))
            if model == "srgan":
                x = x * 2 - 1
            with torch.no_grad():
                y = net(x)
            np.savez_compressed(GOLD / f"sr_{model}_fwd_s{seed}.npz", x=x.numpy(), y=y.numpy(), seed=seed, gain=gain)
        # the real route, with the reference's registry entry carrying our seeded weights
        sd = so.init_state_dict(model, 2, 1.0)
        app.PT_MODELS[model].load_state_dict(sd, strict=True)
        for (H, W) in ((18, 23), (16, 16)):
            img = np.random.default_rng(H * 100 + W).integers(0, 256, (H, W, 3), dtype=np.uint8)
            buf = io.BytesIO()
            Image.fromarray(img).save(buf, format="PNG")
            r = client.post(f"/enhance?model={model}", files={"file": ("a.png", buf.getvalue(), "image/png")})
            assert r.status_code == 200, r.text
            out = np.asarray(Image.open(io.BytesIO(base64.b64decode(r.json()["denoised_image_base64"]))).convert("RGB"))
            np.savez_compressed(GOLD / f"sr_{model}_route_{H}x{W}.npz", img=img, out=out, seed=2)
            print(model, "route", img.shape, "->", out.shape)


if __name__ == "__main__":
    main()
