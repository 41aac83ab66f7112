"""GPU bring-up: ESRGAN / SRGAN vs oracle. python tools/sr_debug.py"""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import torch
from cid_b200.sr_generators import ESRGANGenerator, SRGANGenerator
from oracle import sr_oracle as so

dev = torch.device("cuda:0")
for model, cls in (("esrgan", ESRGANGenerator), ("srgan", SRGANGenerator)):
    for seed, gain, shape in ((0, 1.0, (1, 16, 16)), (1, 1.0, (1, 64, 48)), (3, 1.0, (1, 128, 128)), (1, 1.5, (1, 21, 13)), (2, 1.0, (3, 20, 36))):
        sd = so.init_state_dict(model, seed, gain)
        net = cls(); net.load_state_dict(sd, strict=True); net = net.to(dev).eval()
        B, H, W = shape
        x = torch.rand(B, 3, H, W, generator=torch.Generator().manual_seed(7 + seed))
        if model == "srgan":
            x = x * 2 - 1
        y = net(x.to(dev)).cpu()
        ref = so.forward(model, sd, x)
        d = (y - ref).abs()
        idx = np.unravel_index(int(d.argmax()), d.shape)
        print(f"{model} seed {seed} gain {gain} {shape}: max-abs {float(d.max()):.4f} mean {float(d.mean()):.5f} ref range [{float(ref.min()):.2f},{float(ref.max()):.2f}] worst {tuple(int(i) for i in idx)}")
    img = np.random.default_rng(5).integers(0, 256, (1, 16, 16, 3), dtype=np.uint8)
    pad = so.get_padding(16, 16, 4, 1 if model == "esrgan" else 4)
    out = net.forward_u8(torch.from_numpy(img).to(dev), pad).cpu().numpy()
    sdr = sd
    ref = (so.esrgan_enhance_u8 if model == "esrgan" else so.srgan_enhance_u8)(sdr, img)
    dd = np.abs(out.astype(int) - ref.astype(int))
    print(model, "u8 16x16 max diff", dd.max(), "hist", np.bincount(dd.ravel())[:6], "out mean", out.mean(), "ref mean", ref.mean())
