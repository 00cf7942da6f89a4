"""world_size-2 gloo tests (CPU) of the host-side multi-GPU logic: DDP-style gradient bucketing/averaging of fit_b200.train.GradSync
and the batch-shard bookkeeping of the sampler (row offsets cover the global batch exactly once)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _grad_sync_worker(rank, world, port, q):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from fit_b200.train import GradSync
    gs = GradSync()
    g = torch.Generator().manual_seed(100 + rank)
    groups = {"blocks.1": {"a": torch.randn(7, 5, generator=g), "b": torch.randn(11, generator=g)},
              "blocks.0": {"c": torch.randn(3, 4, generator=g)}}
    local = {k: {n: t.clone() for n, t in v.items()} for k, v in groups.items()}
    for tag, gr in groups.items():
        gs.on_grads(tag, gr)
    gs.finish()
    # expected: mean over ranks
    ok = True
    for tag, gr in local.items():
        for n, t in gr.items():
            both = [torch.zeros_like(t) for _ in range(world)]
            dist.all_gather(both, t)
            ok = ok and torch.allclose(groups[tag][n], sum(both) / world, atol=1e-6)
    # no_sync leaves the gradients untouched
    with gs.no_sync():
        t = torch.ones(4) * (rank + 1)
        gs.on_grads("x", {"t": t})
        gs.finish()
        ok = ok and torch.equal(t, torch.ones(4) * (rank + 1))
    # gradient accumulation (accumulate_grad_batches=2, FiT/train.py:288): micro-step 1 under no_sync accumulates locally into p.grad
    # (autograd's job), micro-step 2 is synced -> every rank must end with mean_over_ranks(g1 + g2) in p.grad, identical on all ranks
    class _M(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.w = torch.nn.Parameter(torch.zeros(6, 3))
            self.b = torch.nn.Parameter(torch.zeros(5))
    m = _M()
    gs.attach(m)
    g1 = {"w": torch.randn(6, 3, generator=g), "b": torch.randn(5, generator=g)}
    g2 = {"w": torch.randn(6, 3, generator=g), "b": torch.randn(5, generator=g)}
    want = {}
    for n in g1:
        both = [torch.zeros_like(g1[n]) for _ in range(world)]
        dist.all_gather(both, g1[n] + g2[n])
        want[n] = sum(both) / world

    def autograd_accumulate(grads):          # what autograd does with the tensors backward() returns
        for n, prm in m.named_parameters():
            prm.grad = grads[n].clone() if prm.grad is None else prm.grad + grads[n]
    with gs.no_sync():
        gs.on_grads("all", {n: t.clone() for n, t in g1.items()})
        gs.finish()
        autograd_accumulate(g1)
    new = {n: t.clone() for n, t in g2.items()}
    flat = torch.cat([new["w"].reshape(-1), new["b"].reshape(-1)])           # in-place bucket, as the training engine allocates them
    views = {"w": flat[:18].view(6, 3), "b": flat[18:]}
    gs.on_grads("all", views, flat)
    gs.finish()
    autograd_accumulate(views)
    for n, prm in m.named_parameters():
        ok = ok and torch.allclose(prm.grad, want[n], atol=1e-6)
    q.put((rank, ok))
    dist.destroy_process_group()


def test_grad_sync_averages_across_two_ranks():
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_grad_sync_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
    assert sorted(r for r, _ in res) == [0, 1] and all(ok for _, ok in res)


def test_batch_shards_cover_global_batch_once():
    from fit_b200.sampler import shard_rows
    for total, world in [(256, 1), (256, 2), (256, 8), (100, 8), (7, 4)]:
        spans = [shard_rows(total, world, r) for r in range(world)]
        assert spans[0][0] == 0 and spans[-1][1] == total
        assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
        assert max(e - b for b, e in spans) - min(e - b for b, e in spans) <= 1
