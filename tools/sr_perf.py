"""Per-layer timing of ESRGAN at batch 256 / 128x128. python tools/sr_perf.py [model] [B] [S]"""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from cid_b200.sr_generators import ESRGANGenerator, SRGANGenerator
from cid_b200.synth import synthetic_noisy_u8
from oracle import sr_oracle as so
model = sys.argv[1] if len(sys.argv) > 1 else "esrgan"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 256
S = int(sys.argv[3]) if len(sys.argv) > 3 else 128
net = (ESRGANGenerator if model == "esrgan" else SRGANGenerator)()
net.load_state_dict(so.init_state_dict(model, 0), strict=True)
net = net.cuda().eval()
x = synthetic_noisy_u8(3, B, S, S).cuda()
y = net.forward_u8(x)
for _ in range(3):
    net.forward_u8(x, out=y)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    net.forward_u8(x, out=y)
e1.record(); torch.cuda.synchronize()
print(model, B, S, "UNPROFILED ms/forward", e0.elapsed_time(e1) / 10, "img/s", B / (e0.elapsed_time(e1) / 10) * 1e3)
net.set_option("profile", 1)
for _ in range(3):
    net.forward_u8(x, out=y)
net.profile_read()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    net.forward_u8(x, out=y)
e1.record(); torch.cuda.synchronize()
ms, n = net.profile_read()
print(model, B, S, "ms/forward", e0.elapsed_time(e1) / 10, "img/s", B / (e0.elapsed_time(e1) / 10) * 1e3)
print({k: round(v / n, 4) for k, v in ms.items()})
