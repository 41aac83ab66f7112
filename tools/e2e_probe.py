"""Names the end-to-end (host-fed) limiter on a multi-GPU box. Under torchrun, every rank measures its pinned-memory
H2D / D2H copy bandwidth alone and with all ranks copying at once, with default page placement and with the staging
buffers bound to the GPU's own NUMA node (set_mempolicy), then the HostPipeline end to end for a few (sub-batch, depth)
choices. Rank 0 prints one JSON document.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 tools/e2e_probe.py
"""
import ctypes
import json
import os
import subprocess
import sys
from pathlib import Path

import torch
import torch.distributed as dist

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from cid_b200 import DenoiseGenerator  # noqa: E402
from cid_b200.pipeline import HostPipeline, set_rank_affinity  # noqa: E402
from cid_b200.synth import synthetic_noisy_u8  # noqa: E402

rank = int(os.environ.get("RANK", "0"))
local = int(os.environ.get("LOCAL_RANK", "0"))
world = int(os.environ.get("WORLD_SIZE", "1"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)


def barrier():
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def sh(cmd):
    try:
        return subprocess.run(cmd, shell=True, capture_output=True, text=True, timeout=20).stdout.strip()
    except Exception as e:  # noqa: BLE001
        return f"<{e}>"


def gpu_numa_node(i):
    try:
        import pynvml
        pynvml.nvmlInit()
        bdf = pynvml.nvmlDeviceGetPciInfo(pynvml.nvmlDeviceGetHandleByIndex(i)).busId
        bdf = (bdf.decode() if isinstance(bdf, bytes) else bdf).lower()
        if len(bdf.split(":")[0]) == 8:
            bdf = bdf[4:]
        return int(Path(f"/sys/bus/pci/devices/{bdf}/numa_node").read_text()), bdf
    except Exception as e:  # noqa: BLE001
        return -1, str(e)


def set_mempolicy(node):
    """MPOL_BIND (2) to one node for this thread's future page allocations; node < 0 restores the default policy."""
    libc = ctypes.CDLL(None, use_errno=True)
    SYS_set_mempolicy = 238  # x86_64
    if node < 0:
        return libc.syscall(SYS_set_mempolicy, 0, None, 0)
    mask = (ctypes.c_ulong * 16)()
    mask[node // 64] = 1 << (node % 64)
    rc = libc.syscall(SYS_set_mempolicy, 2, mask, 16 * 64)
    return rc if rc == 0 else -ctypes.get_errno()


def copy_bw(h_in, h_out, d_in, d_out, mode, iters=20):
    s1, s2 = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    s1.wait_event(e0)
    s2.wait_event(e0)
    for _ in range(iters):
        if mode in ("h2d", "both"):
            with torch.cuda.stream(s1):
                d_in.copy_(h_in, non_blocking=True)
        if mode in ("d2h", "both"):
            with torch.cuda.stream(s2):
                h_out.copy_(d_out, non_blocking=True)
    torch.cuda.current_stream().wait_stream(s1)
    torch.cuda.current_stream().wait_stream(s2)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    nbytes = h_in.numel() * (2 if mode == "both" else 1)
    return nbytes / ms / 1e6  # GB/s


def gather(x):
    if world == 1:
        return [x]
    out = [None] * world
    dist.all_gather_object(out, x)
    return out


report = {"world": world}
node, bdf = gpu_numa_node(local)
info = {"rank": rank, "gpu_numa_node": node, "bdf": bdf, "cpus_allowed": sorted(os.sched_getaffinity(0))[:64]}
aff = set_rank_affinity(local, world) if world > 1 else None
info["affinity_set"] = aff
if rank == 0:
    report["topology"] = {
        "nvidia_smi_topo": sh("nvidia-smi topo -m | head -30"),
        "numa_nodes": sh("ls /sys/devices/system/node | grep node; cat /sys/devices/system/node/node*/cpulist 2>/dev/null"),
        "cpuset_mems": sh("cat /sys/fs/cgroup/cpuset.mems.effective 2>/dev/null || cat /sys/fs/cgroup/cpuset/cpuset.effective_mems 2>/dev/null"),
        "cpuset_cpus": sh("cat /sys/fs/cgroup/cpuset.cpus.effective 2>/dev/null || cat /sys/fs/cgroup/cpuset/cpuset.effective_cpus 2>/dev/null"),
        "lscpu": sh("lscpu | grep -E 'Model name|Socket|NUMA|^CPU\\(s\\)'"),
        "meminfo": sh("grep -E 'MemTotal|MemFree' /proc/meminfo; cat /sys/devices/system/node/node*/meminfo 2>/dev/null | grep -E 'MemTotal|MemFree'"),
    }

N = 64 << 20
d_in = torch.empty(N, dtype=torch.uint8, device=dev)
d_out = torch.empty(N, dtype=torch.uint8, device=dev)
results = {}
for placement in ("default", "gpu_node"):
    rc = 0
    if placement == "gpu_node":
        rc = set_mempolicy(node) if node >= 0 else -1
    h_in = torch.empty(N, dtype=torch.uint8).pin_memory()
    h_out = torch.empty(N, dtype=torch.uint8).pin_memory()
    h_in.fill_(1)
    h_out.fill_(2)
    set_mempolicy(-1)
    res = {"mempolicy_rc": rc}
    for mode in ("h2d", "d2h", "both"):
        # alone: one rank at a time
        alone = None
        for r in range(world):
            barrier()
            if r == rank:
                copy_bw(h_in, h_out, d_in, d_out, mode, 3)
                alone = copy_bw(h_in, h_out, d_in, d_out, mode)
        barrier()
        copy_bw(h_in, h_out, d_in, d_out, mode, 3)
        barrier()
        together = copy_bw(h_in, h_out, d_in, d_out, mode)
        res[mode] = {"alone_GBs": round(alone, 1), "all_ranks_GBs": round(together, 1)}
    results[placement] = res
    del h_in, h_out
info["copy"] = results

# the pipeline itself, a few shapes, with staging placed by the same two policies
torch.manual_seed(0)
net = DenoiseGenerator().to(dev).eval()
B, H, W = 256, 128, 128
pipe_res = []
for placement in ("default", "gpu_node"):
    if placement == "gpu_node" and node >= 0:
        set_mempolicy(node)
    h_in = synthetic_noisy_u8(1234 + rank, B, H, W).pin_memory()
    h_out = torch.empty_like(h_in).pin_memory()
    set_mempolicy(-1)
    for sub, depth in ((128, 3), (64, 4), (64, 8), (32, 8), (256, 3)):
        pipe = HostPipeline(net, dev, sub, H, W, depth=depth)
        for _ in range(5):
            pipe.run(h_in, h_out, join=False)
        pipe.finish()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(30):
            pipe.run(h_in, h_out, join=False)
        pipe.finish()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 30
        t = torch.tensor([ms], device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        pipe_res.append({"placement": placement, "sub_batch": sub, "depth": depth, "ms_per_step_max": round(float(t), 4),
                         "images_per_s_all": round(world * B / float(t) * 1e3)})
    del h_in, h_out
info["pipeline"] = pipe_res
allinfo = gather(info)
if rank == 0:
    report["ranks"] = allinfo
    print(json.dumps(report, indent=1))
if world > 1:
    dist.destroy_process_group()
