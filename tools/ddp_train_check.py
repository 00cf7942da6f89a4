"""N-GPU check of the data-parallel training step (BASELINE config 4 shape per GPU: FiT-XL/2, 32 samples x 256 tokens, bf16):
torchrun --nproc-per-node N tools/ddp_train_check.py
  1. parity: the NCCL-averaged gradients of the per-rank shards == the mean of the per-shard gradients computed on ONE rank
     (same kernels, so the bar is the all-reduce's fp32 summation order: <= 1e-5 relative);
  2. timing: forward + backward + bucketed all-reduce (GradSync, overlapped on a side stream), CUDA events, max over ranks."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import bench
from fit_b200.models.fit import FiT_models
from fit_b200.rope import precompute_freqs_cis_2d
from fit_b200.train import GradSync, masked_mse_loss

os.environ.setdefault("NCCL_MAX_CTAS", os.environ.get("FITB200_DDP_RESERVE_SMS") or ("16" if int(os.environ.get("WORLD_SIZE", "1")) > 2 else "8"))   # NCCL stays inside the SMs GradSync leaves free
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
name = os.environ.get("MODEL", "FiT-XL/2")
N, T = int(os.environ.get("BATCH", 32)), 256
D, Hd, L, heads = bench.MODELS[name]
d = D // heads
torch.manual_seed(0)                       # identical replicas
with torch.device(dev):
    m = FiT_models[name](precision="bf16")
bench.randomize_zero_init(m)
m.train(); m.y_embedder.dropout_prob = 0.0
pos = torch.tensor(precompute_freqs_cis_2d(d, 16, 16, max_length=256), dtype=torch.float32)[None].expand(N, T, d).contiguous().to(dev)
mask = torch.ones(N, T, dtype=torch.bool, device=dev)


def shard(r):
    g = torch.Generator().manual_seed(100 + r)
    return (torch.randn(N, T, 16, generator=g).to(dev), torch.randn(N, T, 16, generator=g).to(dev),
            torch.randint(0, 1000, (N,), generator=g).to(dev), torch.randint(0, 1000, (N,), generator=g).to(dev))


def step(data, sync=None):
    x, tgt, t, y = data
    for p in m.parameters():
        p.grad = None
    out = m(x, t, y, pos, mask, train=True)
    loss = masked_mse_loss(out, tgt, mask)
    loss.backward()
    if sync is not None:
        sync.finish()
    return float(loss.detach())


sync = GradSync() if world > 1 else None
if sync is not None:
    sync.attach(m)
loss = step(shard(rank), sync)
torch.cuda.synchronize()
mine = {k: p.grad.detach().float().clone() for k, p in m.named_parameters() if p.grad is not None}
if world > 1:
    # reference on every rank: mean over the shards of the locally computed (unsynchronised) gradients
    with sync.no_sync():
        ref = None
        for r in range(world):
            step(shard(r), None)
            g = {k: p.grad.detach().float() / world for k, p in m.named_parameters() if p.grad is not None}
            ref = g if ref is None else {k: ref[k] + g[k] for k in g}
    worst = max(float((mine[k] - ref[k]).norm() / (ref[k].norm() + 1e-30)) for k in ref)
    assert worst < 1e-4, f"rank {rank}: all-reduced gradients differ from the mean of the shard gradients: {worst}"
    if rank == 0:
        print(f"parity ok: {len(ref)} parameter gradients, worst relative L2 difference {worst:.2e} (world {world})")
# timing
data = shard(rank)
for _ in range(2):
    step(data, sync)
if world > 1:
    dist.barrier()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
K = int(os.environ.get("ITERS", 5))
e0.record()
for _ in range(K):
    step(data, sync)
e1.record()
torch.cuda.synchronize()
ms = torch.tensor([e0.elapsed_time(e1) / K], device=dev, dtype=torch.float64)
if world > 1:
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
if rank == 0:
    fl = 3 * N * world * bench.forward_flops(D, Hd, L, T)
    print(f"{name} DDP train step: world {world}, {N} samples/GPU (global {N * world}): {float(ms):.2f} ms/step (fwd + bwd + gradient all-reduce, max over ranks) "
          f"= {N * world / float(ms) * 1e3:.1f} samples/s, {fl / float(ms) / 1e9:.1f} TFLOP/s algorithmic aggregate")
if world > 1:
    dist.destroy_process_group()
