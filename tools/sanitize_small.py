"""Small invocations of every generator for compute-sanitizer (memcheck): python tools/sanitize_small.py"""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import torch
from cid_b200 import DenoiseGenerator, ESRGANGenerator, SRGANGenerator, KerasCGANGenerator
from oracle import denoise_oracle as orc, sr_oracle as so, keras_cgan_oracle as ko

dev = torch.device("cuda:0")
net = DenoiseGenerator(); net.load_state_dict(orc.init_state_dict(0, 1.0)); net = net.to(dev).eval()
x = torch.rand(2, 3, 36, 20, device=dev) * 2 - 1
print("denoise", float(net(x).abs().max()))
img = torch.randint(0, 256, (2, 30, 27, 3), dtype=torch.uint8, device=dev)
print("denoise u8", int(net.forward_u8(img, (0, 1, 1, 1)).sum()))
for model, cls in (("esrgan", ESRGANGenerator), ("srgan", SRGANGenerator)):
    n = cls(); n.load_state_dict(so.init_state_dict(model, 0)); n = n.to(dev).eval()
    print(model, float(n(torch.rand(2, 3, 21, 13, device=dev)).abs().max()))
    pad = so.get_padding(27, 30, 4, 1 if model == "esrgan" else 4)
    print(model, "u8", int(n.forward_u8(img, pad).sum()))
g = KerasCGANGenerator(ko.init_params(0), dev)
print("cgan", float(np.abs(g.predict(np.random.default_rng(0).uniform(-1, 1, (2, 20, 36, 3)).astype(np.float32))).max()))
print("cgan u8", int(g.forward_u8(img, (0, 1, 1, 1)).sum()))
torch.cuda.synchronize()
print("done")
