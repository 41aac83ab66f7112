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


def _sync_worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from cid_b200 import DenoiseGenerator
        from cid_b200.training import DenoiseDiscriminator, sync_replicas_
        torch.manual_seed(100 + rank)                      # every rank starts from DIFFERENT random weights
        g, d = DenoiseGenerator(), DenoiseDiscriminator()
        d.model[3].running_mean.fill_(float(rank + 1))     # and different BatchNorm statistics
        params = torch.cat([p.detach().reshape(-1) for p in list(g.parameters()) + list(d.parameters())]).clone()
        bufs = torch.cat([b.detach().reshape(-1).float() for b in d.buffers() if b.dtype.is_floating_point]).clone()
        before = float(params.sum())
        sync_replicas_([params, bufs])
        gathered = [torch.zeros_like(params) for _ in range(world)]
        dist.all_gather(gathered, params)
        gb = [torch.zeros_like(bufs) for _ in range(world)]
        dist.all_gather(gb, bufs)
        q.put((rank, before, all(torch.equal(gathered[0], x) for x in gathered), all(torch.equal(gb[0], x) for x in gb),
               float(bufs[:64].mean())))
    finally:
        dist.destroy_process_group()


def test_replicas_start_from_rank0_even_with_different_seeds():
    """ranks built with different seeds must not silently train different weights on averaged gradients."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_sync_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=300) for _ in procs)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    assert res[0][1] != res[1][1]                          # they really started apart
    for rank, _, same_params, same_bufs, bn_mean in res:
        assert same_params and same_bufs, rank
        assert bn_mean == 1.0                              # rank 0's statistics


def test_discriminator_state_dict_keys_match_the_reference_layout():
    from cid_b200.training import DenoiseDiscriminator
    keys = list(DenoiseDiscriminator().state_dict())
    assert keys[:2] == ["model.0.weight", "model.0.bias"] and keys[-2:] == ["model.12.weight", "model.12.bias"]
    assert "model.3.running_mean" in keys and "model.9.num_batches_tracked" in keys and len(keys) == 25
