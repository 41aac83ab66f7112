"""GPU bring-up: keras cGAN path vs oracle, layer by layer.  python tools/cgan_debug.py [B] [H] [W] [seed|real]"""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import torch
from cid_b200.keras_cgan import KerasCGANGenerator, KERAS_CGAN_SHAPES
from oracle import keras_cgan_oracle as ko

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1
H = int(sys.argv[2]) if len(sys.argv) > 2 else 32
W = int(sys.argv[3]) if len(sys.argv) > 3 else 32
which = sys.argv[4] if len(sys.argv) > 4 else "0"
if which == "real":
    z = np.load(Path(__file__).resolve().parent.parent / "tests/golden/keras_cgan_epoch500.npz")
    var = {k: z[k] for k in KERAS_CGAN_SHAPES}
else:
    var = ko.init_params(int(which))
rng = np.random.default_rng(5)
x = rng.uniform(-1, 1, (B, H, W, 3)).astype(np.float32)
if which == "real":
    face = z["face_u8"]
    x = np.repeat((face[:, :H, :W].astype(np.float32) / 255.0 - 0.5) / 0.5, B, axis=0)
net = KerasCGANGenerator(var, "cuda")
y = net.predict(x)
xt = torch.from_numpy(x).permute(0, 3, 1, 2).contiguous()
for i, name in enumerate(["c0", "c1", "c2", "c3"]):
    ref = ko.forward_nchw(var, xt, upto=i + 1).permute(0, 2, 3, 1)
    got = net.debug_activation(name, B, H, W)
    d = (got - ref).abs()
    print(name, tuple(got.shape), "max-abs", float(d.max()), "ref max", float(ref.abs().max()),
          "worst idx", np.unravel_index(int(d.argmax()), d.shape))
ref = ko.predict(var, x)
d = np.abs(y - ref)
print("y", y.shape, "max-abs", float(d.max()), "worst", np.unravel_index(int(d.argmax()), d.shape))
u8 = torch.from_numpy(rng.integers(0, 256, (B, H - 1, W - 2, 3), dtype=np.uint8))
if which == "real":
    u8 = torch.from_numpy(np.repeat(z["face_u8"][:, :H - 1, :W - 2], B, axis=0).copy())
pad = ko.get_padding(W - 2, H - 1)
o = net.forward_u8(u8.cuda(), pad).cpu().numpy()
r = ko.enhance_u8(var, u8.numpy())
print("u8 max diff", int(np.abs(o.astype(int) - r.astype(int)).max()), "frac differing", float((o != r).mean()))

if which == "real":
    a, b = o.astype(np.float64), r.astype(np.float64)
    print("psnr(gpu, oracle) dB", 10 * np.log10(255 ** 2 / max(((a - b) ** 2).mean(), 1e-12)),
          "hist of |diff|", np.bincount(np.abs(o.astype(int) - r.astype(int)).ravel())[:6])
