#!/usr/bin/env python
"""Benchmark of the generator-forward hot path: denoised images/sec at 128x128.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One "step" = one pass of the hot path over one batch of synthetic noisy faces (BASELINE.json
configs[1]: batch 256 of 128x128 RGB per GPU, random-init generator weights with the reference's
state_dict layout, bf16 tensor-core arithmetic). Prints ONE JSON line (rank 0).

  value      images/s with the uint8 inputs already resident in HBM (cid_forward_u8_nhwc)
  e2e        images/s through the public API with HOST buffers: pinned uint8 in -> H2D ->
             generator -> D2H -> pinned uint8 out, all inside the timed region
  roofline   dominant kernel (the tcgen05 conv of the slowest layer): algorithmic FLOPs / CUDA-event
             time measured live in this run (per-layer events recorded by the library)
  cpu_baseline  the oracle port of the reference's PyTorch CPU path on this box's host cores
                (rank 0, N=1 only, bounded sample)

--impl reference times the reference's CPU implementation of the path (oracle port: the same
torch CPU ops the reference module runs; /root/reference itself is Python source that cannot
travel to the GPU box) on the same metric/config.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

METRIC = "denoised images/sec at 128x128"
UNIT = "images/s"
FLOP_PER_PIXEL = 703232  # conv + convT multiply-adds x2, per output pixel
# per-layer FLOPs per output-resolution pixel of the whole image (2*M*N*K summed per layer / H*W)
LAYER_FLOP_PER_PIXEL = {
    "down1.0": 2 * 64 * 27, "down1.2": 2 * 64 * 576, "down2.0": 2 * 128 * 576 / 4, "down2.2": 2 * 128 * 1152 / 4,
    "bottleneck.0": 2 * 256 * 1152 / 16, "bottleneck.2": 2 * 256 * 2304 / 16, "up2": 2 * 512 * 256 / 16,
    "upconv2.0": 2 * 128 * 2304 / 4, "upconv2.2": 2 * 128 * 1152 / 4, "up1": 2 * 256 * 128 / 4,
    "upconv1.0": 2 * 64 * 1152, "upconv1.2": 2 * 3 * 576,
}


# dram__bytes_read.sum + dram__bytes_write.sum per launch at batch 256 of 128x128, from the committed
# `ncu --set full` capture profiles/r01i_ncu_full_batch256.md (MB); scaled linearly for other batches/sizes.
NCU_TRAFFIC_MB_B256_128 = {
    "down1.0": 13.41 + 476.60, "down1.2": 537.07 + 616.13, "down2.0": 134.45 + 208.42, "down2.2": 268.86 + 284.56,
    "bottleneck.0": 67.77 + 76.91, "bottleneck.2": 135.45 + 89.74, "up2": 134.51 + 210.47,
    "upconv2.0": 537.52 + 238.86, "upconv2.2": 268.82 + 223.15, "up1": 268.53 + 478.73,
    "upconv1.0": 1074.0 + 506.96, "upconv1.2": 536.94 + 8.01,
}


def load_peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return {"bf16": float(d["bf16_tflops"]), "bf16_sustained": float(d.get("bf16_tflops_sustained", 0.0)),
                "hbm": float(d["hbm_gbs"]), "source": "measured"}
    return {"bf16": 1590.0, "bf16_sustained": 1400.0, "hbm": 6650.0, "source": "fallback"}


class ClockSampler:
    """Samples SM clock / throttle reasons through NVML while the timed region runs."""

    def __init__(self, index: int):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thread = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.dev = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.dev, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _loop(self):
        nv = self.nv
        names = {
            "hw_slowdown": getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8),
            "hw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40),
            "sw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20),
            "sw_power_cap": getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4),
            "hw_power_brake": getattr(nv, "nvmlClocksEventReasonHwPowerBrakeSlowdown", 0x80),
        }
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.dev, nv.NVML_CLOCK_SM))
                try:
                    mask = nv.nvmlDeviceGetCurrentClocksEventReasons(self.dev)
                except Exception:
                    mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.dev)
                for k, bit in names.items():
                    if mask & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            time.sleep(0.004)

    def start(self):
        if self.nv is not None:
            self._thread = threading.Thread(target=self._loop, daemon=True)
            self._thread.start()

    def stop(self):
        self._stop.set()
        if self._thread is not None:
            self._thread.join()
        med = statistics.median(self.samples) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


def cpu_forward_baseline(H: int, W: int, batch: int, budget_s: float, steps=None, warmup: int = 1):
    """Times the oracle port of the reference's CPU path (torch fp32, all host threads)."""
    import torch
    from oracle import denoise_oracle as orc
    from cid_b200.synth import synthetic_noisy_u8

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    sd = orc.init_state_dict(0, 1.0)
    x = orc.preprocess_u8(synthetic_noisy_u8(1234, batch, H, W).numpy(), (0, 0, 0, 0))
    times = []
    with torch.no_grad():
        for _ in range(warmup):
            orc.forward(sd, x)
        t_end = time.perf_counter() + budget_s
        while (steps is None and time.perf_counter() < t_end and len(times) < 50) or \
                (steps is not None and len(times) < steps):
            t0 = time.perf_counter()
            orc.forward(sd, x)
            times.append(time.perf_counter() - t0)
    med = statistics.median(times)
    return {"value": batch / med, "unit": UNIT, "cores": torch.get_num_threads(), "kind": "port",
            "sample": f"{len(times)} timed forwards of batch {batch} @ {H}x{W} fp32 (median), oracle port of "
                      f"backend/app.py:80-103 on torch CPU", "ms_per_forward": med * 1e3}


def run_reference(args, rank: int, world: int):
    if rank != 0:
        return
    H = W = args.size
    batch = 32
    res = cpu_forward_baseline(H, W, batch, budget_s=1e9, steps=max(args.steps, 1), warmup=max(args.warmup, 1))
    line = {
        "impl": "reference", "metric": METRIC, "value": res["value"], "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": res["ms_per_forward"], "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"generator forward, batch {args.batch} of {H}x{W} RGB per GPU "
                               f"(BASELINE.json configs[1]); each reference step is a bounded sample of {batch} images"},
        "cpu_baseline": {k: res[k] for k in ("value", "unit", "cores", "kind", "sample")},
        "e2e": {"value": res["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=256, help="images per GPU per step")
    ap.add_argument("--size", type=int, default=128)
    ap.add_argument("--chunk", type=int, default=-1, help="images per pass through the layers (-1 = library default)")
    ap.add_argument("--sub-batch", type=int, default=128, help="e2e: images per H2D/compute/D2H pipeline slot")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--debug-flags", type=int, default=0, help="bring-up experiments (invalidates results)")
    ap.add_argument("--variant", type=int, default=-1, help="CTA-pair layer bit mask (cid_set_option 'variant'; -1 = library default)")
    ap.add_argument("--layers", action="store_true", help="also print the per-layer table to stderr")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch
    import torch.distributed as dist

    from cid_b200.generator import DenoiseGenerator
    from cid_b200.pipeline import HostPipeline
    from cid_b200.synth import synthetic_noisy_u8

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the generator path has no CPU fallback "
                         "(use --impl reference for the CPU baseline)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # rank 0 prints ONE JSON line on stdout: NCCL's "NCCL version ..." banner (written to stdout whenever
        # NCCL_DEBUG >= VERSION) goes to stderr instead — fd 1 points at stderr while the communicator comes up
        sys.stdout.flush()
        saved_stdout = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=dev)
            dist.barrier()  # the communicator (and its banner) is created lazily by the first collective
            torch.cuda.synchronize(dev)
        finally:
            sys.stdout.flush()
            os.dup2(saved_stdout, 1)
            os.close(saved_stdout)

    B, H, W = args.batch, args.size, args.size
    peaks = load_peaks()

    # random-init generator with the reference's state_dict layout (the shipped .pth is not in the tree)
    torch.manual_seed(0)
    net = DenoiseGenerator().to(dev).eval()
    if args.chunk >= 0:
        net.set_option("chunk", args.chunk)
    net.set_option("profile", 1)
    if args.debug_flags:
        net.set_option("debug_flags", args.debug_flags)
    if args.variant >= 0:
        net.set_option("variant", args.variant)

    host_in = synthetic_noisy_u8(1234 + rank, B, H, W).pin_memory()
    host_out = torch.empty_like(host_in).pin_memory()
    x = host_in.to(dev)
    y = torch.empty_like(x)
    pad = (0, 0, 0, 0)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    # ---------------- device-resident throughput ----------------
    for _ in range(max(args.warmup, 3)):
        net.forward_u8(x, pad, out=y)
    torch.cuda.synchronize(dev)
    net.profile_read()
    launches0 = net.kernel_launches()
    sampler = ClockSampler(local_rank)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    sampler.start()
    e0.record()
    for _ in range(args.steps):
        net.forward_u8(x, pad, out=y)
    e1.record()
    barrier()
    clocks = sampler.stop()
    ms_total = e0.elapsed_time(e1)
    launches = net.kernel_launches() - launches0
    layer_ms, passes = net.profile_read()

    # ---------------- end to end through the public host API ----------------
    pipe = HostPipeline(net, dev, min(args.sub_batch, B), H, W, depth=2)
    for _ in range(3):
        pipe.run(host_in, host_out)
    torch.cuda.synchronize(dev)
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    f0.record()
    for _ in range(args.steps):
        h2d, d2h = pipe.run(host_in, host_out, join=False)  # a serving loop: batch k+1's upload overlaps batch k's tail
    pipe.finish()
    f1.record()
    barrier()
    ms_e2e = f0.elapsed_time(f1)
    checksum = int(host_out.view(-1)[::4099].to(torch.int64).sum())  # the D2H result is really read

    t = torch.tensor([ms_total, ms_e2e], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total, ms_e2e = float(t[0]), float(t[1])

    if rank == 0:
        imgs = B * world * args.steps
        value = imgs / (ms_total * 1e-3)
        e2e_value = imgs / (ms_e2e * 1e-3)
        # dominant kernel = the tensor-core conv layer with the largest share of the step
        conv_layers = [k for k in layer_ms if k != "down1.0"]
        dom = max(conv_layers, key=lambda k: layer_ms[k])
        dom_ms = layer_ms[dom] / max(passes, 1)
        launches_per_pass = passes / max(args.steps, 1)
        imgs_per_pass = B / max(launches_per_pass, 1)
        dom_flops = LAYER_FLOP_PER_PIXEL[dom] * H * W * imgs_per_pass
        achieved = dom_flops / (dom_ms * 1e-3) / 1e12
        roofline = {"bound": "tensor", "kernel": f"conv_umma_kernel[{dom}]", "achieved": achieved,
                    "peak": peaks["bf16"], "unit": "TFLOP/s", "frac": achieved / peaks["bf16"],
                    "traffic": NCU_TRAFFIC_MB_B256_128[dom] * 1e6 * (imgs_per_pass * H * W) / (256 * 128 * 128),
                    "traffic_source": "dram__bytes_read+write per launch, ncu --set full (profiles/r01i_ncu_full_batch256.md), "
                                      "scaled to this launch's pixel count",
                    "peak_source": f"{peaks['source']} burst bf16 (MEASURED_PEAKS.json bf16_tflops)",
                    "flops_per_launch": dom_flops, "ms_per_launch": dom_ms}
        net_tflops = value * FLOP_PER_PIXEL * H * W / 1e12
        step_ms = ms_total / args.steps
        per_layer = {k: {"ms": layer_ms[k] / args.steps,
                         "tflops": LAYER_FLOP_PER_PIXEL[k] * H * W * B / (layer_ms[k] / args.steps * 1e-3) / 1e12,
                         "share": layer_ms[k] / args.steps / step_ms} for k in layer_ms}
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": step_ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": f"generator forward, batch {B} of {H}x{W} RGB per GPU, random-init weights "
                                   f"(BASELINE.json configs[1]); u8 NHWC in -> u8 NHWC out",
                       "batch_per_gpu": B, "height": H, "width": W, "parallelism": f"batch-sharded x{world}, no collective",
                       "l2": "per-step activation working set (~4 GB at batch 256) exceeds the 126 MB L2; no flush needed"},
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": ms_e2e / args.steps, "checksum": checksum},
            "gpu_launches": launches,
            "roofline": roofline,
            "net_tensor": {"achieved_tflops": net_tflops, "frac_of_burst_peak": net_tflops / peaks["bf16"],
                           "frac_of_sustained_peak": (net_tflops / peaks["bf16_sustained"]) if peaks["bf16_sustained"] else None},
            "layers": per_layer,
        }
        if world == 1 and not args.no_cpu_baseline:
            cb = cpu_forward_baseline(H, W, 32, budget_s=15.0)
            line["cpu_baseline"] = {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")}
        if args.layers:
            for k, v in per_layer.items():
                print(f"  {k:14s} {v['ms']:8.4f} ms  {v['tflops']:8.1f} TFLOP/s  {100 * v['share']:5.1f} %", file=sys.stderr)
        print(json.dumps(line), flush=True)

    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
