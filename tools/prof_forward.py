"""One batch-256 forward of the denoise generator for an ncu capture (2 warm-up passes = 24 launches, then the 12 launches
to capture):   ncu --set full --clock-control none --import-source on --launch-skip 24 --launch-count 12 \
                   -o gpurun_out/fwd_full python tools/prof_forward.py"""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from cid_b200 import DenoiseGenerator  # noqa: E402
from cid_b200.synth import synthetic_noisy_u8  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
S = int(sys.argv[2]) if len(sys.argv) > 2 else 128
torch.manual_seed(0)
net = DenoiseGenerator().to("cuda:0").eval().set_option("graph", 0)
x = synthetic_noisy_u8(1234, B, S, S).to("cuda:0")
y = torch.empty_like(x)
for _ in range(3):
    net.forward_u8(x, (0, 0, 0, 0), out=y)
torch.cuda.synchronize()
print("launches", net.kernel_launches())
