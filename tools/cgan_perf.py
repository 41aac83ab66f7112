"""Keras cGAN generator at batch B / 256x256: forward time and per-layer times.   python tools/cgan_perf.py [B]"""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from cid_b200.keras_cgan import KerasCGANGenerator
from cid_b200.synth import synthetic_noisy_u8
from oracle import keras_cgan_oracle as ko

B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
S = 256
var = ko.init_params(0)
x = synthetic_noisy_u8(2, B, S, S).cuda()
g = KerasCGANGenerator(var, "cuda")
y = g.forward_u8(x, (0, 0, 0, 0))
for _ in range(3):
    g.forward_u8(x, (0, 0, 0, 0), out=y)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    g.forward_u8(x, (0, 0, 0, 0), out=y)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 10
g.set_option("profile", 1)
g.forward_u8(x, (0, 0, 0, 0), out=y); g.profile_read()
for _ in range(5):
    g.forward_u8(x, (0, 0, 0, 0), out=y)
lm, n = g.profile_read()
print(f"keras cgan {B} x {S}^2: ms/forward {ms:.4f} img/s {B / ms * 1e3:.0f}", {k: round(v / n, 4) for k, v in lm.items()})
