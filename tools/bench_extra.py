"""Secondary measurements (not the driver's bench line): BASELINE.json configs[3] (high-resolution
latency path, batch 1 and 32 at 256^2 / 512^2) for the denoise generator, and the Keras cGAN generator
at 256^2. CUDA events, device-resident u8 in -> u8 out, >= 5 warm-ups; the oracle port on the host
cores beside each number.   python tools/bench_extra.py > gpurun_out/extra.json"""
import json
import os
import statistics
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import numpy as np
import torch

from cid_b200.generator import DenoiseGenerator
from cid_b200.keras_cgan import KerasCGANGenerator
from cid_b200.sr_generators import ESRGANGenerator, SRGANGenerator
from cid_b200.synth import synthetic_noisy_u8
from oracle import denoise_oracle as orc
from oracle import keras_cgan_oracle as ko
from oracle import sr_oracle as so

DEN_FLOP_PX = 703232
CGAN_LAYER_FLOP_PX = {"conv2d": 2 * 64 * 48 / 4, "conv2d_1": 2 * 128 * 1024 / 16, "conv2d_transpose": 2 * 128 * 512 / 4,
                      "conv2d_transpose_1": 2 * 64 * 512, "conv2d_2": 2 * 3 * 576}
CGAN_FLOP_PX = sum(CGAN_LAYER_FLOP_PX.values())


def time_gpu(fn, iters, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def time_cpu(fn, budget_s=6.0, max_n=9):
    fn()
    ts = []
    t_end = time.perf_counter() + budget_s
    while len(ts) < 2 or (time.perf_counter() < t_end and len(ts) < max_n):
        t0 = time.perf_counter()
        fn()
        ts.append(time.perf_counter() - t0)
    return statistics.median(ts) * 1e3


def main():
    dev = torch.device("cuda:0")
    torch.set_num_threads(os.cpu_count() or 1)
    out = {"cores": torch.get_num_threads(), "denoise": [], "keras_cgan": []}
    torch.manual_seed(0)
    net = DenoiseGenerator().to(dev).eval()
    sd = orc.init_state_dict(0, 1.0)
    for B, S in [(1, 128), (1, 256), (32, 256), (1, 512), (32, 512)]:
        x = synthetic_noisy_u8(1, B, S, S).to(dev)
        y = torch.empty_like(x)
        ms = time_gpu(lambda: net.forward_u8(x, (0, 0, 0, 0), out=y), 50 if B == 1 else 20)
        cb = min(B, 4)
        xc = orc.preprocess_u8(synthetic_noisy_u8(1, cb, S, S).numpy(), (0, 0, 0, 0))
        with torch.no_grad():
            cpu_ms = time_cpu(lambda: orc.forward(sd, xc)) / cb
        out["denoise"].append({"batch": B, "size": S, "ms": ms, "images_per_s": B / ms * 1e3,
                               "tflops": B * S * S * DEN_FLOP_PX / ms / 1e9, "cpu_ms_per_image": cpu_ms,
                               "cpu_batch": cb})
        print(out["denoise"][-1], file=sys.stderr)
    var = ko.init_params(0)
    g = KerasCGANGenerator(var, dev).set_option("profile", 1)
    for B in (1, 32, 256):
        S = 256
        x = synthetic_noisy_u8(2, B, S, S).to(dev)
        y = torch.empty_like(x)
        g.forward_u8(x, (0, 0, 0, 0), out=y)
        g.profile_read()
        iters = 50 if B == 1 else 20
        ms = time_gpu(lambda: g.forward_u8(x, (0, 0, 0, 0), out=y), iters, warm=5)
        layer_ms, passes = g.profile_read()
        layers = {k: {"ms": v / passes, "tflops": CGAN_LAYER_FLOP_PX[k] * B * S * S / (v / passes) / 1e9}
                  for k, v in layer_ms.items()}
        cb = min(B, 4)
        xc = synthetic_noisy_u8(2, cb, S, S).numpy()
        cpu_ms = time_cpu(lambda: ko.enhance_u8(var, xc)) / cb
        out["keras_cgan"].append({"batch": B, "size": S, "ms": ms, "images_per_s": B / ms * 1e3,
                                  "tflops": B * S * S * CGAN_FLOP_PX / ms / 1e9, "layers": layers,
                                  "cpu_ms_per_image": cpu_ms, "cpu_batch": cb})
        print(out["keras_cgan"][-1], file=sys.stderr)
    # ESRGAN / SRGAN: FLOPs = 2*M*N*K over their convolutions
    flop_px = {"esrgan": 2 * (64 * 243 + 16 * 64 * 576 + 3 * 5184),
               "srgan": 2 * (64 * 243 + 11 * 64 * 576 + 256 * 576 + 4 * 256 * 576 + 16 * 3 * 5184)}
    out["sr"] = []
    for model, cls, cases in (("esrgan", ESRGANGenerator, [(1, 128), (256, 128), (32, 256)]),
                              ("srgan", SRGANGenerator, [(1, 128), (64, 128), (8, 256)])):
        sd = so.init_state_dict(model, 0)
        net = cls()
        net.load_state_dict(sd, strict=True)
        net = net.to(dev).eval().set_option("profile", 1)
        for B, S in cases:
            x = synthetic_noisy_u8(3, B, S, S).to(dev)
            y = net.forward_u8(x)
            net.profile_read()
            ms = time_gpu(lambda: net.forward_u8(x, (0, 0, 0, 0), out=y), 50 if B == 1 else 10, warm=3)
            layer_ms, passes = net.profile_read()
            cb = 1
            xc = synthetic_noisy_u8(3, cb, S, S).numpy()
            fn = so.esrgan_enhance_u8 if model == "esrgan" else so.srgan_enhance_u8
            cpu_ms = time_cpu(lambda: fn(sd, xc), budget_s=4.0, max_n=5) / cb
            out["sr"].append({"model": model, "batch": B, "size": S, "ms": ms, "images_per_s": B / ms * 1e3,
                              "tflops": B * S * S * flop_px[model] / ms / 1e9,
                              "layers_ms": {k: v / passes for k, v in layer_ms.items()},
                              "cpu_ms_per_image": cpu_ms})
            print({k: v for k, v in out["sr"][-1].items() if k != "layers_ms"}, file=sys.stderr)
            if B > 1:
                print("   ", {k: round(v / passes, 4) for k, v in layer_ms.items()}, file=sys.stderr)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
