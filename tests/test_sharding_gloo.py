"""N>1 host path on CPU: world_size-2 gloo run of the batch-sharding logic (contiguous shards, no
data-path collective; the all_gather is the test/offline reassembly helper only)."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from cid_b200.sharding import gather_shards, shard_range


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n_items, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        full = torch.arange(n_items * 6, dtype=torch.uint8).view(n_items, 2, 1, 3)   # stand-in "images"
        a, b = shard_range(n_items, rank, world)
        local = full[a:b] + 1                  # per-image independent "processing"
        out = gather_shards(local, n_items)
        ok = torch.equal(out, full + 1)
        t = torch.tensor([float(b - a)])
        dist.all_reduce(t)                     # plumbing check: shards cover the batch
        q.put((rank, bool(ok), int(t.item())))
    finally:
        dist.destroy_process_group()


def _run(n_items):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n_items, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    for rank, ok, total in res:
        assert ok and total == n_items


def test_two_rank_shard_and_gather_even():
    _run(8)


def test_two_rank_shard_and_gather_ragged():
    _run(7)
