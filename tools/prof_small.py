"""Batch-1 forwards for an ncu launch list (graph replay off so every kernel is a plain launch):
   ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/b1_launches.csv python tools/prof_small.py 128"""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from cid_b200 import DenoiseGenerator  # noqa: E402
from cid_b200.synth import synthetic_noisy_u8  # noqa: E402

S = int(sys.argv[1]) if len(sys.argv) > 1 else 128
regime = int(sys.argv[2]) if len(sys.argv) > 2 else 0
torch.manual_seed(0)
net = DenoiseGenerator().to("cuda:0").eval().set_option("graph", 0).set_option("regime", regime)
x = synthetic_noisy_u8(1, 1, S, S).to("cuda:0")
y = torch.empty_like(x)
for _ in range(4):
    net.forward_u8(x, (0, 0, 0, 0), out=y)
torch.cuda.synchronize()
