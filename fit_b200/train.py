"""Training-step arithmetic of the hot path (FiT/train.py:45-69, masked_FiT/train.py:40-60): the autograd bridge that makes
``loss.backward()`` run the hand-written backward, the masked MSE, q_sample (= DDIMScheduler.add_noise) and the data-parallel
gradient exchange (DDP-style NCCL all-reduce, bucketed per block and overlapped with the rest of the backward)."""
from __future__ import annotations

from typing import Dict, List, Optional

import os

import torch

from . import ops
from .engine_train import FiTTrainEngine

Tensor = torch.Tensor
BACKWARD_READY = True


def _train_engine(model) -> FiTTrainEngine:
    eng = getattr(model, "_train_engine", None)
    if eng is None or eng.precision != model.precision:
        eng = FiTTrainEngine(model, model.precision)
        object.__setattr__(model, "_train_engine", eng)
    return eng


def attach_optimizer(model, optimizer):
    """``FusedAdamWEMA`` <-> training engine: the optimizer kernel writes the bf16 / packed operand copies of the parameters it updates
    straight into the engine's weight buffers (FiT/train.py:215-217 runs AdamW on fp32 masters and autocast re-casts them every forward;
    here the cast rides in the optimizer pass).  Without it the engine re-fills its buffers whenever a parameter version changes."""
    optimizer.attach_engine(_train_engine(model))
    return optimizer


class _FiTFunction(torch.autograd.Function):
    """One autograd node for the whole network: forward keeps the activations, backward returns every parameter gradient."""

    @staticmethod
    def forward(ctx, model, x, t, y, pos, mask, train, *params):
        eng = _train_engine(model)
        out, saved = eng.forward(x, t, y, pos, mask, train)
        ctx.eng, ctx.saved, ctx.names = eng, saved, [n for n, _ in model.named_parameters()]
        ctx.sync = getattr(model, "_grad_sync", None)
        return out

    @staticmethod
    def backward(ctx, dout):
        sync = ctx.sync
        grads = ctx.eng.backward(ctx.saved, dout.contiguous(), on_grads=None if sync is None else sync.on_grads)
        if sync is not None:
            sync.finish()
        ctx.saved = None
        return (None,) * 7 + tuple(grads.get(n) for n in ctx.names)


def fit_forward_autograd(model, x, t, y, pos, mask, train=True):
    """FiT.forward under autograd (called by fit_b200.models.fit.FiT.forward when gradients are required)."""
    y = model.y_embedder.labels_for(y, model.training)
    return _FiTFunction.apply(model, x, t, y, pos, mask, train, *model.parameters())


class _MaskedMSE(torch.autograd.Function):
    @staticmethod
    def forward(ctx, out, target, mask):
        acc = torch.zeros(2, dtype=torch.float32, device=out.device)
        loss = torch.empty(1, dtype=torch.float32, device=out.device)
        out_c, target_c = out.float().contiguous(), target.float().contiguous()
        ops.masked_mse_fwd(out_c, target_c, mask, acc, loss)
        ctx.save_for_backward(out_c, target_c, mask, acc)
        return loss[0]

    @staticmethod
    def backward(ctx, g):
        out, target, mask, acc = ctx.saved_tensors
        dout = torch.empty_like(out)
        ops.masked_mse_bwd(out, target, mask, acc, g.reshape(1).float().contiguous(), dout)
        return dout, None, None


def masked_mse_loss(out: Tensor, target: Tensor, mask: Tensor) -> Tensor:
    """F.mse_loss(out[mask], target[mask]) of FiT/train.py:63-66 (tokens [N, T, P], mask [N, T] bool) as one fused kernel each way."""
    if not out.is_cuda:
        raise RuntimeError("fit_b200 kernels need CUDA tensors (there is no CPU fallback)")
    return _MaskedMSE.apply(out, target, mask)


masked_token_mse = masked_mse_loss


def add_noise(diffusion, x0: Tensor, noise: Tensor, t: Tensor) -> Tensor:
    """DDIMScheduler(num_train_timesteps=1000).add_noise call site (FiT/train.py:53) = q_sample with the linear schedule."""
    return diffusion.q_sample(x0, t, noise=noise)


# ---------------------------------------------------------------------------------------------------- data parallel
def default_reserve_sms(world: int) -> int:
    """SMs left to NCCL's CTAs next to the persistent GEMM / attention grids (FITB200_DDP_RESERVE_SMS overrides; set NCCL_MAX_CTAS to the same
    number before the process group is created).  Measured on 8 x B200, FiT-XL/2 at 32 samples per GPU (fwd + bwd + all-reduce of 2.7 GB):
    8 SMs 44.9 ms, 16 SMs 40.6 ms, 32 SMs 42.1 ms per step -- with 8 CTAs the 8-rank all-reduce (22.8 ms alone) no longer hides behind the
    25 ms backward; at 2 ranks (16 ms alone) 8 are enough."""
    env = os.environ.get("FITB200_DDP_RESERVE_SMS")
    if env:
        return int(env)
    return 16 if world > 2 else 8


class GradSync:
    """DDP-style gradient averaging for the custom backward.

    The backward hands over each block's parameter gradients as soon as they are final (``on_grads``); they are packed into one
    flat fp32 bucket per block and all-reduced (NCCL over NVLink / NVSwitch on GPUs, gloo in the CPU tests) on a side stream
    while the backward of the earlier blocks keeps the compute stream busy.  ``finish`` joins the side stream, scales by
    1 / world_size and scatters the averaged values back into the gradient tensors autograd returns.
    ``no_sync`` skips the exchange on non-final accumulation micro-steps (accumulate_grad_batches=2, FiT/train.py:288): autograd keeps
    adding those micro-steps' local gradients into ``p.grad``; on the synced micro-step the accumulated ``p.grad`` is folded into the
    bucket BEFORE the collective (and cleared), so every rank ends with mean_over_ranks(g_1 + ... + g_k) like torch DDP."""

    def __init__(self, process_group=None, bucket_bytes: int = 0):
        import torch.distributed as dist
        self.dist, self.pg = dist, process_group
        self.world = dist.get_world_size(process_group) if dist.is_initialized() else 1
        # NCCL averages inside the collective (ReduceOp.AVG); gloo (CPU tests) sums and the bucket is scaled afterwards
        self.avg_in_collective = dist.is_initialized() and dist.get_backend(process_group) == "nccl"
        self.enabled = True
        self._pending: List = []
        self._stream = None
        self._params: Dict[str, torch.nn.Parameter] = {}
        # NCCL's CTAs need SMs of their own next to the persistent GEMM / attention grids: leave `reserve_sms` free while syncing.
        # (Set NCCL_MAX_CTAS <= reserve_sms before the process group is created to hold NCCL to that budget.)
        self.reserve_sms = default_reserve_sms(self.world) if self.world > 1 and self.avg_in_collective else 0
        if self.reserve_sms > 0:
            from . import _lib as L
            n = torch.cuda.get_device_properties(torch.cuda.current_device()).multi_processor_count
            L.check(L.lib().fitb200_set_sm_limit(n - self.reserve_sms), "set_sm_limit")

    def attach(self, model):
        object.__setattr__(model, "_grad_sync", self)
        self._params = dict(model.named_parameters())
        return self

    class _NoSync:
        def __init__(self, gs): self.gs = gs
        def __enter__(self): self.gs.enabled = False
        def __exit__(self, *a): self.gs.enabled = True

    def no_sync(self):
        return GradSync._NoSync(self)

    def on_grads(self, tag: str, grads: Dict[str, Tensor], flat: Optional[Tensor] = None):
        """``flat``: the gradients are views into this one buffer (the training engine allocates each block's gradients that way):
        it is all-reduced in place, with no packing copy before and no scatter copy after."""
        if not self.enabled or self.world == 1:
            return
        # gradients accumulated under no_sync() by earlier micro-steps join the exchange: reduce (accumulated + new), not only the new part
        for name, g in grads.items():
            prm = self._params.get(name)
            if prm is not None and prm.grad is not None:
                g.add_(prm.grad)     # same accumulation autograd itself performs; g is a view into `flat` when the bucket is in place
                prm.grad = None
        tensors = list(grads.values())
        in_place = flat is not None
        if not in_place:
            flat = torch.cat([g.reshape(-1) for g in tensors])   # packing copy (no arithmetic)
        if flat.is_cuda:
            if self._stream is None:
                self._stream = torch.cuda.Stream(device=flat.device)
            ready = torch.cuda.current_stream().record_event()
            self._stream.wait_event(ready)
            with torch.cuda.stream(self._stream):
                work = self.dist.all_reduce(flat, op=self.dist.ReduceOp.AVG if self.avg_in_collective else self.dist.ReduceOp.SUM,
                                            group=self.pg, async_op=True)
            flat.record_stream(self._stream)
        else:
            work = self.dist.all_reduce(flat, op=self.dist.ReduceOp.SUM, group=self.pg, async_op=True)
        self._pending.append((flat, [] if in_place else tensors, work))

    def finish(self):
        if not self._pending:
            return
        for flat, tensors, work in self._pending:
            work.wait()
        if self._stream is not None:
            torch.cuda.current_stream().wait_stream(self._stream)
        inv = 1.0 / self.world
        for flat, tensors, _ in self._pending:
            if not self.avg_in_collective:
                flat.mul_(inv)   # gloo has no AVG: scale the reduced bucket (CPU test path only)
            off = 0
            for g in tensors:
                n = g.numel()
                g.copy_(flat[off:off + n].view_as(g))
                off += n
        self._pending.clear()


# ---------------------------------------------------------------------------------------------------- CUDA-graph training step
class CapturedTrainStep:
    """One training step -- forward, masked MSE, hand-written backward, fused AdamW / EMA (FiT/train.py:45-69,215-217) -- captured once as a
    CUDA graph and replayed: the step is ~1500 launches of 10-60 us kernels, and issuing them from Python (25-30 us per launch) was slower
    than the GPU executes them (41.6 ms step, 38.6 ms of device time).

        step = CapturedTrainStep(model, optimizer)          # optimizer: FusedAdamWEMA over model.parameters()
        loss = step(x_t, t, y, pos, mask, target)           # device fp32 scalar (static tensor: read it before the next call)

    The first call runs eagerly (it builds the engine's weight buffers and the optimizer tables), the second captures and replays; later
    calls copy the batch into the captured input buffers and replay.  A new batch shape captures again.  Activations and gradients live
    in the graph's private memory pool.  With a GradSync attached (data parallel) the step stays eager unless FITB200_TRAIN_GRAPH_DDP=1
    (the NCCL all-reduces are then captured on their side stream with the rest)."""

    def __init__(self, model, optimizer, loss_fn=masked_mse_loss):
        self.model, self.opt, self.loss_fn = model, optimizer, loss_fn
        attach_optimizer(model, optimizer)
        optimizer.capturable = True        # step number / bias corrections on the device: one graph serves every step
        optimizer.sync_device_state()
        self._graphs = {}      # batch signature -> (graph, static inputs, static loss)
        self._warm = set()
        self.captures = 0

    def _eager(self, x, t, y, pos, mask, target):
        for p in self.model.parameters():
            p.grad = None
        out = self.model(x, t, y, pos, mask, train=True)
        loss = self.loss_fn(out, target, mask)
        loss.backward()
        self.opt.step()
        return loss.detach()

    def __call__(self, x, t, y, pos, mask, target):
        sync = getattr(self.model, "_grad_sync", None)
        if sync is not None and sync.world > 1 and os.environ.get("FITB200_TRAIN_GRAPH_DDP", "0") != "1":
            return self._eager(x, t, y, pos, mask, target)
        batch = (x, t, y, pos, mask, target)
        sig = tuple((tuple(b.shape), b.dtype) for b in batch)
        if sig not in self._warm:     # first step of this shape: eager (allocator warm-up, weight buffers, optimizer tables)
            self._warm.add(sig)
            return self._eager(*batch)
        ent = self._graphs.get(sig)
        if ent is None:
            static = [b.clone() for b in batch]
            for p in self.model.parameters():
                p.grad = None          # the captured backward allocates the gradients inside the graph's pool
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                out = self.model(*static[:5], train=True)
                loss = self.loss_fn(out, static[5], static[4])
                loss.backward()
                self.opt.step()
                static_loss = loss.detach()
            self.captures += 1
            ent = (g, static, static_loss)
            self._graphs[sig] = ent
        g, static, static_loss = ent
        for dst, src in zip(static, batch):
            if dst.data_ptr() != src.data_ptr():
                dst.copy_(src, non_blocking=True)
        g.replay()
        # the replay updated the parameters (and the engine's operand copies) behind Python's back
        self.opt.note_external_step()
        return static_loss
