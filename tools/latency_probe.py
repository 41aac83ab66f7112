"""Small-batch latency of the denoise generator under each execution form (not the driver's bench line; bench.py's `latency`
record is): per (batch, size) the time per forward_u8 call for regime auto / large / split, with and without CUDA-graph
replay, and the per-call host time.   python tools/latency_probe.py > gpurun_out/latency_probe.json"""
import json
import sys
import time
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from cid_b200 import DenoiseGenerator  # noqa: E402
from cid_b200.synth import synthetic_noisy_u8  # noqa: E402

dev = torch.device("cuda:0")
torch.manual_seed(0)
out = []
for B, S in ((1, 128), (1, 256), (1, 512), (4, 128), (32, 128)):
    x = synthetic_noisy_u8(77, min(B, 4), S, S)
    x = x.repeat((B + x.shape[0] - 1) // x.shape[0], 1, 1, 1)[:B].contiguous().to(dev)
    y = torch.empty_like(x)
    for regime in (0, 1, 2):
        for graph in (1, 0):
            net = DenoiseGenerator().to(dev).eval().set_option("regime", regime).set_option("graph", graph)
            f = lambda: net.forward_u8(x, (0, 0, 0, 0), out=y)  # noqa: E731
            for _ in range(10):
                f()
            torch.cuda.synchronize()
            iters = 300 if B == 1 else 50
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0 = time.perf_counter()
            e0.record()
            for _ in range(iters):
                f()
            e1.record()
            host = (time.perf_counter() - t0) / iters * 1e6
            torch.cuda.synchronize()
            us = e0.elapsed_time(e1) / iters * 1e3
            rec = {"batch": B, "size": S, "regime": regime, "graph": graph, "us_per_call": round(us, 2),
                   "host_us_per_call": round(host, 2), "graph_replays": net.graph_replays()}
            out.append(rec)
            print(rec, file=sys.stderr)
print(json.dumps(out))
