"""Two u8 forward passes at a given batch (for ncu captures): python tools/one_pass.py [B] [size] [chunk]"""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from cid_b200.generator import DenoiseGenerator
from cid_b200.synth import synthetic_noisy_u8

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
S = int(sys.argv[2]) if len(sys.argv) > 2 else 128
chunk = int(sys.argv[3]) if len(sys.argv) > 3 else 0
torch.manual_seed(0)
net = DenoiseGenerator().cuda().eval()
if chunk:
    net.set_option("chunk", chunk)
x = synthetic_noisy_u8(1, B, S, S).cuda()
y = torch.empty_like(x)
for _ in range(2):
    net.forward_u8(x, (0, 0, 0, 0), out=y)
torch.cuda.synchronize()
print("ok", int(y.sum()))
