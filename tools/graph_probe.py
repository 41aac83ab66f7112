"""Probe: does CUDA-graph replay of the 12-launch forward shorten the batch-1 latency?"""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from cid_b200.generator import DenoiseGenerator
from cid_b200.synth import synthetic_noisy_u8

torch.manual_seed(0)
net = DenoiseGenerator().cuda().eval()
if len(sys.argv) > 1:
    net.set_option("pdl", int(sys.argv[1]))
for B, S in [(1, 128), (1, 256), (1, 512), (4, 128), (256, 128)]:
    x = synthetic_noisy_u8(1, B, S, S).cuda()
    y = torch.empty_like(x)
    for _ in range(3):
        net.forward_u8(x, (0, 0, 0, 0), out=y)
    torch.cuda.synchronize()
    ref = y.clone()
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        net.forward_u8(x, (0, 0, 0, 0), out=y)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            net.forward_u8(x, (0, 0, 0, 0), out=y)
    torch.cuda.synchronize()
    def timeit(fn, n=100):
        for _ in range(5):
            fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n * 1e3
    t_plain = timeit(lambda: net.forward_u8(x, (0, 0, 0, 0), out=y))
    y.zero_()
    t_graph = timeit(g.replay)
    print(f"B={B} {S}x{S}: plain {t_plain:.1f} us, graph {t_graph:.1f} us, same={bool(torch.equal(y, ref))}")
