"""Bit-compare the denoise generator's u8 output under two 'variant' masks (cid_set_option): python tools/variant_check.py BASE BITS"""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from cid_b200 import DenoiseGenerator  # noqa: E402

torch.manual_seed(0)
net = DenoiseGenerator().to("cuda:0").eval()
x = torch.randint(0, 256, (8, 72, 104, 3), dtype=torch.uint8, device="cuda:0")
base, bits = int(sys.argv[1]), int(sys.argv[2])
net.set_option("variant", base)
y0 = net.forward_u8(x).clone()
net.set_option("variant", base | bits)
y1 = net.forward_u8(x)
torch.cuda.synchronize()
print("variant", base | bits, "equal", bool(torch.equal(y0, y1)), "maxdiff", int((y0.int() - y1.int()).abs().max()))
