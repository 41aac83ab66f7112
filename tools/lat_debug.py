"""Batch-1 latency over time after idle and after a full-power burst (DVFS state matters: bench.py warms the latency
cases for this reason).   python tools/lat_debug.py"""
import sys, time
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from cid_b200 import DenoiseGenerator
from cid_b200.synth import synthetic_noisy_u8
import pynvml
pynvml.nvmlInit(); h = pynvml.nvmlDeviceGetHandleByIndex(0)
dev = torch.device("cuda:0")
torch.manual_seed(0)
net = DenoiseGenerator().to(dev).eval()
x = synthetic_noisy_u8(77, 1, 128, 128).to(dev); y = torch.empty_like(x)
X = synthetic_noisy_u8(1, 256, 128, 128).to(dev); Y = torch.empty_like(X)

def blocks(tag, n=12, iters=500):
    out = []
    for _ in range(n):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters): net.forward_u8(x, (0, 0, 0, 0), out=y)
        e1.record(); torch.cuda.synchronize()
        out.append((round(e0.elapsed_time(e1) / iters * 1e3, 1), pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)))
    print(tag, out)

for _ in range(10): net.forward_u8(x, (0, 0, 0, 0), out=y)
torch.cuda.synchronize()
blocks("from idle")
for _ in range(40): net.forward_u8(X, (0, 0, 0, 0), out=Y)
torch.cuda.synchronize()
blocks("after a 100 ms full-power burst")
time.sleep(2.0)
blocks("after 2 s idle")
