#!/usr/bin/env python
"""Benchmark of the generator-forward hot path: denoised images/sec at 128x128.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One "step" = one pass of the hot path over one batch of synthetic noisy faces (BASELINE.json
configs[1]: batch 256 of 128x128 RGB per GPU, random-init generator weights with the reference's
state_dict layout, bf16 tensor-core arithmetic). Prints ONE JSON line (rank 0).

  value      images/s with the uint8 inputs already resident in HBM (cid_forward_u8_nhwc), timed with
             the per-layer profiling events OFF (the configuration a user runs: the layers chain
             through programmatic dependent launch)
  e2e        images/s through the public API with HOST buffers: pinned uint8 in -> H2D ->
             generator -> D2H -> pinned uint8 out, all inside the timed region (three-stage
             pipeline); e2e.breakdown times the copies alone and the kernels alone
  roofline   dominant kernel (the tcgen05 conv of the slowest layer): algorithmic FLOPs / CUDA-event
             time measured live in a separate profiled pass (per-layer events recorded by the library);
             traffic = dram bytes of that kernel from the newest profiles/r*_ncu_traffic.json
  sustained  the headline loop repeated for >= 3 s with its own clock samples
  latency    BASELINE configs[3]: batch 1 and 32 at 128^2 / 256^2 / 512^2 (the upload route's shape)
  strong_2048  BASELINE configs[2]: 2048 images split contiguously over the N ranks
  train      BASELINE configs[4]: one G+D+Adam iteration at batch 64 per GPU (gradient all-reduce
             fused into Adam over NVLink peer memory, or NCCL) through cid_train_*
  membound   achieved HBM GB/s of the pre-processing kernels (noiser, resize, normalise)
  cpu_baseline  the oracle port of the reference's PyTorch CPU path on this box's host cores
                (rank 0, N=1 only, bounded sample)

--impl reference times the reference's CPU implementation of the path (oracle port: the same
torch CPU ops the reference module runs; /root/reference itself is Python source that cannot
travel to the GPU box) on the same metric/config.
"""
from __future__ import annotations

import argparse
import hashlib
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
# conv-only FLOPs (2*M*N*K) of one training iteration per 128x128 image: G fwd 11.52 + G bwd 22.99
# + 3 D fwd x 1.265 + 2 D bwd (data + weight gradients) x 2.47 + 1 D bwd (data gradient only) x 1.265
TRAIN_GFLOP_PER_IMAGE_128 = 11.5218 + 22.987 + 3 * 1.265 + 2 * 2.473 + 1.265


def load_peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return {"bf16": float(d["bf16_tflops"]), "bf16_sustained": float(d.get("bf16_tflops_sustained", 0.0)),
                "hbm": float(d["hbm_gbs"]), "source": "measured (MEASURED_PEAKS.json)"}
    return {"bf16": 1590.0, "bf16_sustained": 1400.0, "hbm": 6650.0, "source": "fallback (B200_PROFILING.md)"}


def load_ncu_traffic():
    """Newest profiles/r*_ncu_traffic.json (written by tools/ncu_summary.py --traffic-json from an `ncu --set full`
    capture of one batch-256 forward): per-layer dram__bytes_read + dram__bytes_write and the hash of the library it
    was captured with."""
    files = sorted((ROOT / "profiles").glob("r*_ncu_traffic.json"))
    if not files:
        return None
    d = json.loads(files[-1].read_text())
    d["file"] = f"profiles/{files[-1].name}"
    return d


def library_sha():
    """Hash of the library's SOURCES (nvcc output is not byte-reproducible): names the build a profile belongs to."""
    from cid_b200.build import source_sha
    return source_sha()


class ClockSampler:
    """Samples SM clock / throttle reasons through NVML while the timed region runs."""

    def __init__(self, index: int, period_s: float = 0.002):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self.period = period_s
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
            time.sleep(self.period)

    def start(self):
        if self.nv is not None:
            try:  # first NVML queries of a process can take tens of ms: pay for them before the timed region
                self.nv.nvmlDeviceGetClockInfo(self.dev, self.nv.NVML_CLOCK_SM)
                self.nv.nvmlDeviceGetCurrentClocksEventReasons(self.dev)
            except Exception:
                pass
            self._thread = threading.Thread(target=self._loop, daemon=True)
            self._thread.start()
        return self

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


class Timer:
    """CUDA-event timing of a region on the current stream, with the barrier + synchronize bracket of the contract.
    Every region's milliseconds are collected and reduced (MAX over ranks) in one all-reduce at the end."""

    def __init__(self, torch, dist, dev, world):
        self.torch, self.dist, self.dev, self.world = torch, dist, dev, world
        self.ms = {}

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize(self.dev)

    def region(self, name, fn, iters, warm=0, sync_first=True):
        for _ in range(warm):
            fn()
        e0, e1 = self.torch.cuda.Event(enable_timing=True), self.torch.cuda.Event(enable_timing=True)
        if sync_first:
            self.barrier()
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        self.barrier()
        self.ms[name] = e0.elapsed_time(e1)
        return self.ms[name]

    def reduce_max(self):
        names = sorted(self.ms)
        t = self.torch.tensor([self.ms[n] for n in names], dtype=self.torch.float64, device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        self.ms = {n: float(v) for n, v in zip(names, t.tolist())}
        return self.ms


def bench_membound(torch, dev, tm: Timer):
    """Achieved HBM GB/s of the memory-bound pre-processing kernels (north star: 'achieved HBM GB/s against peak for the
    elementwise and norm kernels'). Inputs are far larger than the 126 MB L2, so every pass streams from DRAM.
    Algorithmic bytes: noiser 1 B read + 1 B written per element; resize input + output bytes (the intermediate of the
    two-pass resampler is extra traffic, not algorithmic); normalise 1 B read + 4 B written per element."""
    from cid_b200 import noiser as nz

    out = []
    g = torch.Generator(device=dev).manual_seed(5)
    big = torch.randint(0, 256, (1024, 512, 512, 3), dtype=torch.uint8, device=dev, generator=g)   # 805 MB
    n = big.numel()
    cases = [
        ("gaussian_u8_kernel (Philox4x32-10 + Box-Muller)", lambda: nz.add_gaussian_noise(big, 0.0, 25.0, seed=1), 2 * n),
        ("elementwise_noise_kernel<uniform> (Philox)", lambda: nz.add_uniform_noise(big, 0.0, 25.0, seed=2), 2 * n),
        ("elementwise_noise_kernel<speckle> (Philox)", lambda: nz.add_speckle_noise(big, 0.0, 0.1, seed=3), 2 * n),
        ("salt_pepper_slab_kernel (one pass: bulk load -> Philox draws in shared memory -> bulk store; 2 % + 2 %)",
         lambda: nz.add_salt_pepper_noise(big, 0.02, 0.02, seed=4), 2 * n),
    ]
    if hasattr(nz, "add_gaussian_unit"):
        cases.append(("unit_noise_kernel<gaussian> ([0,1] domain, sigma 0.1, Philox)",
                      lambda: nz.add_gaussian_unit(big, seed=6), 2 * n))
    for name, fn, nbytes in cases:
        ms = tm.region("mb:" + name, fn, 5, warm=2) / 5
        out.append({"kernel": name, "bytes": nbytes, "ms_key": "mb:" + name, "iters": 5})
        del ms
    del big
    torch.cuda.empty_cache()
    src = torch.randint(0, 256, (256, 512, 512, 3), dtype=torch.uint8, device=dev, generator=g)    # 201 MB
    name = "resize_fused_kernel (PIL bicubic, both passes in one kernel, 512^2 -> 256^2, batch 256)"
    tm.region("mb:" + name, lambda: nz.resize_bicubic(src, (256, 256)), 5, warm=2)
    out.append({"kernel": name, "bytes": src.numel() + 256 * 256 * 256 * 3, "ms_key": "mb:" + name, "iters": 5})
    del src
    img = torch.randint(0, 256, (4096, 128, 128, 3), dtype=torch.uint8, device=dev, generator=g)   # 201 MB in, 805 MB out
    name = "normalize_kernel (ToTensor + Normalize, u8 NHWC -> f32 NCHW, batch 4096 @128^2)"
    tm.region("mb:" + name, lambda: nz.normalize_to_tensor(img), 5, warm=2)
    out.append({"kernel": name, "bytes": img.numel() * 5, "ms_key": "mb:" + name, "iters": 5})
    del img
    torch.cuda.empty_cache()
    return out


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
    ap.add_argument("--depth", type=int, default=3, help="e2e: staging slots of the three-stage pipeline")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--layers", action="store_true", help="also print the per-layer table to stderr")
    ap.add_argument("--opt", action="append", default=[], metavar="NAME=VALUE",
                    help="cid_set_option on the generator (A/B experiments, e.g. --opt wst7=8)")
    ap.add_argument("--only", default="", help="comma list of extra sections to run (default: all): "
                                               "sustained,latency,strong,train,membound")
    ap.add_argument("--sustained-seconds", type=float, default=3.0)
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
    from cid_b200.pipeline import HostPipeline, set_rank_affinity
    from cid_b200.synth import synthetic_noisy_u8

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the generator path has no CPU fallback "
                         "(use --impl reference for the CPU baseline)")
    sections = set(s for s in args.only.split(",") if s) or {"sustained", "latency", "strong", "train", "membound"}
    affinity = set_rank_affinity(local_rank, world) if world > 1 else None
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
    K, Wm = max(args.steps, 1), max(args.warmup, 3)
    peaks = load_peaks()
    tm = Timer(torch, dist, dev, world)

    # random-init generator with the reference's state_dict layout (the shipped .pth is not in the tree)
    torch.manual_seed(0)
    net = DenoiseGenerator().to(dev).eval()
    if args.chunk >= 0:
        net.set_option("chunk", args.chunk)
    for kv in args.opt:
        name, val = kv.split("=", 1)
        net.set_option(name, int(val))

    host_in = synthetic_noisy_u8(1234 + rank, B, H, W).pin_memory()
    host_out = torch.empty_like(host_in).pin_memory()
    x = host_in.to(dev)
    y = torch.empty_like(x)
    pad = (0, 0, 0, 0)
    step = lambda: net.forward_u8(x, pad, out=y)  # noqa: E731

    # ---------------- device-resident throughput (profiling events off) ----------------
    for _ in range(Wm):
        step()
    torch.cuda.synchronize(dev)
    launches0 = net.kernel_launches()
    sampler = ClockSampler(local_rank).start()
    tm.region("value", step, K)
    clocks = sampler.stop()
    launches = net.kernel_launches() - launches0

    # ---------------- end to end through the public host API ----------------
    pipe = HostPipeline(net, dev, min(args.sub_batch, B), H, W, depth=args.depth)
    io_bytes = [0, 0]
    for stage, key in (("all", "e2e"), ("copy", "e2e_copy"), ("compute", "e2e_compute")):
        # every timed region starts from the same board state as `value` did (idle, then W warm-up passes): straight after
        # a full-power region the board is power-capped and the same loop runs 5-10 % slower (power-limited)
        torch.cuda.synchronize(dev)
        time.sleep(0.5)
        for _ in range(max(Wm, 3)):
            pipe.run(host_in, host_out, stage=stage)
        torch.cuda.synchronize(dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        tm.barrier()
        e0.record()
        for _ in range(K):  # a serving loop: batch k+1's upload overlaps batch k's tail
            h2d, d2h = pipe.run(host_in, host_out, join=False, stage=stage)
        pipe.finish()
        e1.record()
        tm.barrier()
        tm.ms[key] = e0.elapsed_time(e1)
        if stage == "all":
            io_bytes = [h2d, d2h]
            checksum = int(host_out.view(-1)[::4099].to(torch.int64).sum())  # the D2H result is really read

    # ---------------- per-layer profiled pass (cudaEventRecord between layers: not the headline) ----------------
    net.set_option("profile", 1)
    for _ in range(2):
        step()
    torch.cuda.synchronize(dev)
    net.profile_read()
    tm.region("profiled", step, K)
    layer_ms, passes = net.profile_read()
    net.set_option("profile", 0)

    # ---------------- BASELINE configs[3]: latency path ----------------
    latency = []
    if "latency" in sections:
        for (lb, ls) in ((1, 128), (32, 128), (1, 256), (32, 256), (1, 512), (32, 512)):
            xi = synthetic_noisy_u8(77, min(lb, 4), ls, ls)
            xi = xi.repeat((lb + xi.shape[0] - 1) // xi.shape[0], 1, 1, 1)[:lb].contiguous().to(dev)
            yo = torch.empty_like(xi)
            iters = 500 if lb == 1 else 30
            key = f"lat:{lb}x{ls}"
            call = lambda: net.forward_u8(xi, pad, out=yo)  # noqa: E731
            # settle the clocks on THIS workload first: right after a full-power phase the first ~50-80 ms of a light
            # workload run at depressed clocks (batch 1 @128: 97 us instead of the steady 82 us, tools/lat_debug.py)
            t_settle = time.perf_counter()
            while time.perf_counter() - t_settle < 0.15:
                for _ in range(20):
                    call()
                torch.cuda.synchronize(dev)
            tm.region(key, call, iters, warm=5)
            latency.append({"batch": lb, "size": ls, "iters": iters, "ms_key": key})
            del xi, yo
        net.release_workspace()
        torch.cuda.empty_cache()

    # ---------------- BASELINE configs[2]: 2048 images split over the ranks ----------------
    strong = None
    if "strong" in sections:
        from cid_b200.sharding import shard_range
        a, b = shard_range(2048, rank, world)
        xs = x.repeat((b - a + B - 1) // B, 1, 1, 1)[: b - a].contiguous()
        ys = torch.empty_like(xs)
        net.set_option("chunk", 256)
        n_strong = 20  # (5 passes were a 15 ms window at 8 GPUs: one slow rank's clock ramp decided the max over ranks)
        tm.region("strong", lambda: net.forward_u8(xs, pad, out=ys), n_strong, warm=3)
        net.set_option("chunk", max(args.chunk, 0))
        strong = {"images": 2048, "images_this_rank": b - a, "iters": n_strong}
        del xs, ys
        torch.cuda.empty_cache()

    # ---------------- sustained: the headline loop for >= 3 s (after the short latency cases: it leaves the GPU hot) ----------------
    sustained = None
    if "sustained" in sections:
        est = max(tm.ms["value"] / K, 1e-3)
        n_sus = max(int(args.sustained_seconds * 1e3 / est) + 1, K)
        sus_sampler = ClockSampler(local_rank, period_s=0.02).start()
        tm.region("sustained", step, n_sus, warm=3)   # (the latency section released the workspace)
        sustained = {"steps": n_sus, "clocks": sus_sampler.stop()}

    # ---------------- memory-bound pre-processing kernels ----------------
    membound = bench_membound(torch, dev, tm) if "membound" in sections else []

    # ---------------- BASELINE configs[4]: training step ----------------
    train = None
    if "train" in sections:
        from cid_b200.training import DenoiseDiscriminator, DenoiseGANTrainStep
        torch.manual_seed(0)
        tr = DenoiseGANTrainStep(DenoiseGenerator(), DenoiseDiscriminator(), device=dev)
        TB = 64
        gen = torch.Generator(device="cpu").manual_seed(4321 + rank)
        clean = torch.rand(TB, 3, H // 8, W // 8, generator=gen)
        clean = torch.nn.functional.interpolate(clean, size=(H, W), mode="bicubic", align_corners=False).clamp(0, 1)
        noisy = (clean + 0.1 * torch.randn(TB, 3, H, W, generator=gen)).clamp(0, 1)
        noisy, clean = ((noisy - 0.5) / 0.5).to(dev).contiguous(), ((clean - 0.5) / 0.5).to(dev).contiguous()
        tstep = lambda: tr.step(noisy, clean, read_losses=False)  # noqa: E731
        for _ in range(3):
            tstep()
        torch.cuda.synchronize(dev)
        l0 = tr.kernel_launches()
        tm.region("train", tstep, 10)
        t_launch = (tr.kernel_launches() - l0) // 10
        losses = tr.losses()
        train = {"batch_per_gpu": TB, "iters": 10, "gpu_launches_per_step": t_launch, "losses": losses,
                 "allreduce": ("fused into Adam over NVLink peer memory (adam_allreduce_kernel)" if tr.peer_fused
                               else "nccl all_reduce" if world > 1 else "none (1 GPU)")}
        tr.close()
        del tr, noisy, clean
        torch.cuda.empty_cache()

    ms = tm.reduce_max()

    if rank == 0:
        ms_total, ms_e2e = ms["value"], ms["e2e"]
        imgs = B * world * K
        value = imgs / (ms_total * 1e-3)
        e2e_value = imgs / (ms_e2e * 1e-3)
        step_ms = ms_total / K
        # dominant kernel = the tensor-core conv layer with the largest share of the (profiled) step
        conv_layers = [k for k in layer_ms if k != "down1.0" and layer_ms[k] > 0]
        dom = max(conv_layers, key=lambda k: layer_ms[k])
        dom_ms = layer_ms[dom] / max(passes, 1)
        launches_per_pass = passes / K
        imgs_per_pass = B / max(launches_per_pass, 1)
        dom_flops = LAYER_FLOP_PER_PIXEL[dom] * H * W * imgs_per_pass
        achieved = dom_flops / (dom_ms * 1e-3) / 1e12
        traffic, traffic_src = None, "no profiles/r*_ncu_traffic.json committed for this build"
        ncu = load_ncu_traffic()
        if ncu and dom in ncu.get("layers", {}):
            lay = ncu["layers"][dom]
            scale = (imgs_per_pass * H * W) / (ncu["batch"] * ncu["height"] * ncu["width"])
            traffic = (lay["dram_read_bytes"] + lay["dram_write_bytes"]) * scale
            same = ncu.get("source_sha") == library_sha()
            traffic_src = (f"dram__bytes_read+write per launch, ncu --set full, {ncu['file']} "
                           f"({'captured with a build of these csrc/ sources' if same else 'captured with a build of OLDER csrc/ sources'})"
                           ", scaled to this launch's pixel count")
        roofline = {"bound": "tensor", "kernel": f"conv_umma_kernel[{dom}]", "achieved": achieved,
                    "peak": peaks["bf16"], "unit": "TFLOP/s", "frac": achieved / peaks["bf16"],
                    "traffic": traffic, "traffic_source": traffic_src,
                    "peak_source": f"{peaks['source']} burst bf16 (bf16_tflops)",
                    "flops_per_launch": dom_flops, "ms_per_launch": dom_ms}
        prof_step_ms = ms["profiled"] / K
        per_layer = {k: {"ms": layer_ms[k] / K,
                         "tflops": (LAYER_FLOP_PER_PIXEL[k] * H * W * B / (layer_ms[k] / K * 1e-3) / 1e12) if layer_ms[k] > 0 else None,
                         "share": layer_ms[k] / K / prof_step_ms} for k in layer_ms}
        per_gpu_tflops = value / world * FLOP_PER_PIXEL * H * W / 1e12
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K,
            "warmup": Wm, "ms_per_step": step_ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": f"generator forward, batch {B} of {H}x{W} RGB per GPU, random-init weights "
                                   f"(BASELINE.json configs[1]); u8 NHWC in -> u8 NHWC out",
                       "batch_per_gpu": B, "height": H, "width": W, "parallelism": f"batch-sharded x{world}, no collective",
                       "l2": "per-step activation working set (~7 GB at batch 256) exceeds the 126 MB L2; no flush needed",
                       "host_affinity": (f"{len(affinity)} CPUs from {affinity[0]}" if affinity else None)},
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": io_bytes[0], "d2h_bytes_per_step": io_bytes[1],
                    "ms_per_step": ms_e2e / K, "checksum": checksum,
                    "pipeline": f"3 stages (H2D / kernels / D2H streams), {args.depth} slots of {min(args.sub_batch, B)} images",
                    "breakdown": {"copy_only_ms_per_step": ms["e2e_copy"] / K, "compute_only_ms_per_step": ms["e2e_compute"] / K,
                                  "note": "same pipeline with the kernels (copy_only) or the copies (compute_only) left out; "
                                          "max over ranks"}},
            "gpu_launches": launches,
            "roofline": roofline,
            "net_tensor": {"achieved_tflops_per_gpu": per_gpu_tflops, "frac_of_burst_peak": per_gpu_tflops / peaks["bf16"],
                           "frac_of_sustained_peak": (per_gpu_tflops / peaks["bf16_sustained"]) if peaks["bf16_sustained"] else None},
            "profiled_pass": {"ms_per_step": prof_step_ms, "note": "per-layer cudaEventRecord between the layers; "
                                                                   "`layers` and `roofline` come from this pass, `value` does not"},
            "layers": per_layer,
        }
        if sustained:
            s_ms = ms["sustained"] / sustained["steps"]
            s_tf = B * FLOP_PER_PIXEL * H * W / (s_ms * 1e-3) / 1e12
            sustained.update({"seconds": ms["sustained"] * 1e-3, "ms_per_step": s_ms, "value": B * world / (s_ms * 1e-3),
                              "unit": UNIT, "tflops_per_gpu": s_tf, "frac_of_burst_peak": s_tf / peaks["bf16"],
                              "frac_of_sustained_peak": (s_tf / peaks["bf16_sustained"]) if peaks["bf16_sustained"] else None})
            line["sustained"] = sustained
        if latency:
            for e in latency:
                t = ms[e.pop("ms_key")] / e["iters"]
                tf = e["batch"] * e["size"] ** 2 * FLOP_PER_PIXEL / (t * 1e-3) / 1e12
                e.update({"us": t * 1e3, "images_per_s": e["batch"] / (t * 1e-3), "tflops": tf,
                          "frac_of_burst_peak": tf / peaks["bf16"]})
            line["latency"] = {"config": "BASELINE.json configs[3]: forward_u8 calls back to back on one stream, u8 in -> u8 out "
                                         "resident in HBM, per call (max over ranks); each case is timed after 150 ms of its own calls, "
                                         "i.e. at the clocks a process serving this shape settles to", "cases": latency}
        if strong:
            t = ms["strong"] / strong["iters"]
            strong.update({"ms": t, "value": 2048 / (t * 1e-3), "unit": UNIT, "scaling": "strong",
                           "config": "BASELINE.json configs[2]: 2048 images of 128x128 split contiguously over the ranks, "
                                     "256 images per pass"})
            line["strong_2048"] = strong
        if train:
            t = ms["train"] / train["iters"]
            gflop = TRAIN_GFLOP_PER_IMAGE_128 * (H * W / 16384.0)
            tf = train["batch_per_gpu"] * gflop / t
            train.update({"ms_per_step": t, "value": world * train["batch_per_gpu"] / (t * 1e-3), "unit": "images/s",
                          "tflops_per_gpu": tf, "frac_of_burst_peak": tf / peaks["bf16"], "gflop_per_image": gflop,
                          "config": f"BASELINE.json configs[4]: denoise GAN G+D+Adam iteration, batch {train['batch_per_gpu']}/GPU "
                                    f"@{H}x{W} (training.py:409-427), weak scaling"})
            line["train"] = train
        if membound:
            for e in membound:
                t = ms[e.pop("ms_key")] / e.pop("iters")
                gbs = e["bytes"] / (t * 1e-3) / 1e9
                e.update({"ms": t, "gbs": gbs, "frac_of_hbm_peak": gbs / peaks["hbm"]})
            line["membound"] = {"peak_gbs": peaks["hbm"], "peak_source": peaks["source"], "kernels": membound}
        if world == 1 and not args.no_cpu_baseline:
            cb = cpu_forward_baseline(H, W, 32, budget_s=15.0)
            line["cpu_baseline"] = {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")}
        if args.layers:
            for k, v in per_layer.items():
                print(f"  {k:14s} {v['ms']:8.4f} ms  {v['tflops'] or 0:8.1f} TFLOP/s  {100 * v['share']:5.1f} %", file=sys.stderr)
        print(json.dumps(line), flush=True)

    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
