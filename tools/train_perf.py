"""Throughput of the training step (BASELINE config #5: batch 64 per GPU at 128x128, NCCL gradient all-reduce when
launched under torchrun). Prints one JSON line (rank 0). Not the headline bench: bench.py measures the inference path.

    python tools/train_perf.py [--B 64] [--H 128] [--W 128] [--steps 20] [--warmup 5] [--phases]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/train_perf.py
"""
import argparse
import json
import os
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from cid_b200 import DenoiseGenerator  # noqa: E402
from cid_b200.training import DenoiseDiscriminator, DenoiseGANTrainStep  # noqa: E402

# conv-only FLOPs (2*M*N*K) of one iteration per 128x128 image: G fwd 11.52 + G bwd 22.99 + 3 D fwd x 1.265
# + 2 D bwd (data + weight gradients) x 2.47 + 1 D bwd (data gradient only) x 1.265, scaled by H*W/16384
GFLOP_PER_IMAGE_128 = 11.5218 + 22.987 + 3 * 1.265 + 2 * 2.473 + 1.265


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--B", type=int, default=64)
    ap.add_argument("--H", type=int, default=128)
    ap.add_argument("--W", type=int, default=128)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--phases", action="store_true", help="also time the four phases separately")
    ap.add_argument("--opt", action="append", default=[], help="name=value for cid_train_set_option (A/B switches)")
    ap.add_argument("--nccl", action="store_true", help="N > 1: NCCL all-reduce + Adam instead of the fused peer-memory kernel")
    ap.add_argument("--check", action="store_true", help="N > 1: run 3 steps on both all-reduce paths and compare the weights")
    a = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        import torch.distributed as dist
        saved = os.dup(1)
        os.dup2(2, 1)  # NCCL's banner must not reach stdout
        dist.init_process_group("nccl", device_id=dev)
        dist.barrier()
        os.dup2(saved, 1)
    torch.manual_seed(0)
    g, d = DenoiseGenerator(), DenoiseDiscriminator()
    tr = DenoiseGANTrainStep(g, d, device=dev, peer_allreduce=not a.nccl)
    for kv in a.opt:
        tr.set_option(kv.split("=")[0], int(kv.split("=")[1]))
    gen = torch.Generator(device="cpu").manual_seed(1234 + rank)
    clean = (torch.rand(a.B, 3, a.H // 8, a.W // 8, generator=gen))
    clean = torch.nn.functional.interpolate(clean, size=(a.H, a.W), mode="bicubic", align_corners=False).clamp(0, 1)
    noisy = (clean + 0.1 * torch.randn(a.B, 3, a.H, a.W, generator=gen)).clamp(0, 1)
    noisy, clean = ((noisy - 0.5) / 0.5).to(dev).contiguous(), ((clean - 0.5) / 0.5).to(dev).contiguous()
    if a.check and world > 1:
        ref = DenoiseGANTrainStep(g, d, device=dev, peer_allreduce=False)
        worst = 0.0
        # the optimiser paths are compared on the SAME rank-local gradients (two backward passes of one trainer differ in
        # the last bits - atomics order in the weight-gradient kernels - and Adam turns the sign of a ~0 gradient into +-lr)
        for it in range(3):
            for which, net, backward in ((1, "discriminator", tr.d_backward), (0, "generator", tr.g_backward)):
                backward(noisy, clean)
                ref.section(net, 1).copy_(tr.section(net, 1))
                tr.adam(which)
                ref._allreduce(ref.section(net, 1))
                ref.adam(which)
                x, y = tr.section(net, 0), ref.section(net, 0)
                worst = max(worst, float((x - y).abs().max()))
                mine = x.clone()
                dist.broadcast(mine, src=0)
                assert torch.equal(mine, x), f"rank {rank}: replicas diverged ({net}, step {it})"
        tr.losses()
        t = torch.tensor([worst], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        if rank == 0:
            print(json.dumps({"check": "fused peer-memory all-reduce + Adam vs NCCL all-reduce + Adam, 3 steps",
                              "fused_path": bool(tr.peer_fused), "fallback_reason": tr._peer_error,
                              "max_abs_weight_diff": float(t.item()), "replicas_bit_identical": True, "n_gpus": world}))
        assert float(t.item()) < 1e-7
        ref.close()
    for _ in range(a.warmup):
        tr.step(noisy, clean, read_losses=False)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    l0 = tr.kernel_launches()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        tr.step(noisy, clean, read_losses=False)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / a.steps
    launches = (tr.kernel_launches() - l0) // a.steps
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    phases = None
    if a.phases and world == 1:
        phases = {}
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(5)]
        acc = [0.0] * 4
        for _ in range(a.steps):
            ev[0].record(); tr.d_backward(noisy, clean)
            ev[1].record(); tr.adam(1)
            ev[2].record(); tr.g_backward(noisy, clean)
            ev[3].record(); tr.adam(0)
            ev[4].record()
            torch.cuda.synchronize()
            for i in range(4):
                acc[i] += ev[i].elapsed_time(ev[i + 1])
        for i, n in enumerate(("d_backward", "adam_d", "g_backward", "adam_g")):
            phases[n] = acc[i] / a.steps
    losses = tr.losses()
    if rank == 0:
        gflop = GFLOP_PER_IMAGE_128 * (a.H * a.W / 16384.0)
        ips = world * a.B / (ms * 1e-3)
        print(json.dumps({"metric": "training images/sec (G+D step, Adam" + ((", gradient all-reduce fused into Adam over NVLink peer memory" if tr.peer_fused else ", NCCL grad all-reduce") if world > 1 else "") + ")",
                          "value": ips, "unit": "images/s", "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
                          "ms_per_step": ms, "tflops_per_gpu": a.B * gflop / ms, "gflop_per_image": gflop,
                          "gpu_launches_per_step": launches, "dtype": "bf16 (fp32 accumulate, fp32 master weights)",
                          "config": {"workload": f"denoise GAN train step, batch {a.B}/GPU @{a.H}x{a.W}", "scaling": "weak"},
                          "phases_ms": phases, "losses": losses}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
