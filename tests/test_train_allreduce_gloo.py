"""N>1 host logic of the training step on CPU: world_size-2 gloo run of the gradient averaging the trainer performs
between each backward pass and its Adam update (NCCL on the GPU box; BASELINE config #5), checked against the
data-parallel identity: the mean of the per-rank gradients of the ORACLE equals the oracle's gradient on the
concatenated batch for the batch-mean losses (MSE / BCE) - for the generator, which has no BatchNorm."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from cid_b200.training import allreduce_mean_
        from oracle import train_oracle as to
        torch.set_num_threads(2)
        g, _ = to.make_models(5)
        noisy, clean = to.synthetic_pair(9, 4, 16, 16)
        a, b = rank * 2, rank * 2 + 2                      # contiguous shard of the global batch
        with torch.enable_grad():
            loss = torch.nn.functional.mse_loss(g(noisy[a:b]), clean[a:b])
            loss.backward()
        flat = torch.cat([p.grad.reshape(-1) for p in g.parameters()])
        allreduce_mean_(flat)
        g.zero_grad()
        with torch.enable_grad():
            torch.nn.functional.mse_loss(g(noisy), clean).backward()
        full = torch.cat([p.grad.reshape(-1) for p in g.parameters()])
        err = float((flat - full).norm() / full.norm())
        q.put((rank, err, int(flat.numel())))
    finally:
        dist.destroy_process_group()


def test_two_rank_gradient_average_equals_global_batch_gradient():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=300) for _ in procs]
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    for rank, err, n in res:
        assert n == 1_827_587
        assert err < 1e-4, (rank, err)


def test_allreduce_is_a_noop_without_a_process_group():
    from cid_b200.training import allreduce_mean_
    t = torch.arange(5, dtype=torch.float32)
    assert torch.equal(allreduce_mean_(t.clone()), t)
