"""CUDA-graph sampling loop: the caller side of the hot path (sample_lightning.py:36-70 of the reference).

One denoising step = [loop index -> model timestep gather] + [FiT forward on 2B rows: B conditional | B null-class]
+ [Philox noise] + [fused CFG + DDPM/DDIM update of the B latents] + [index decrement], captured ONCE as a CUDA
graph and replayed ``num_timesteps`` times: no per-step host-to-device copy (the reference does ~8 per step:
gaussian_diffusion.py:679,882-894; respace.py:125), no per-step allocation, no Python between kernels.

Only the B latents the reference keeps (``samples, _ = samples.chunk(2)``, sample_lightning.py:69) are carried;
the duplicated half of forward_with_cfg's input (fit.py:545-546) is produced by the embed kernel's row
replication, not by a copy.  Noise is keyed by (seed, step, global image index): a batch sharded over N GPUs
(``row_offset``) reproduces the single-GPU images bit for bit, with no collective.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from . import ops
from ._nvtx import rng

Tensor = torch.Tensor


class CFGSampler:
    def __init__(self, model, diffusion, batch: int, latent_hw: Tuple[int, int], cfg_scale: float = 1.5, seed: int = 0,
                 row_offset: int = 0, mode: str = "ddpm", eta: float = 0.0, clip_denoised: bool = False,
                 use_graph: bool = True, device=None):
        if mode not in ("ddpm", "ddim"):
            raise ValueError("mode must be 'ddpm' or 'ddim'")
        self.model, self.diffusion = model, diffusion
        self.B = batch
        self.cfg_scale, self.seed, self.row_offset = float(cfg_scale), int(seed), int(row_offset)
        self.mode, self.eta, self.clip = mode, float(eta), bool(clip_denoised)
        dev = device or next(model.parameters()).device
        if dev.type != "cuda":
            raise RuntimeError("fit_b200: the sampler needs a CUDA device (no CPU fallback)")
        self.device = dev
        H, W = latent_hw
        p, C = model.patch_size, model.in_channels
        self.T = (H // p) * (W // p)
        self.d = model.hidden_size // model.num_heads
        B = batch
        self.x = torch.empty(B, C, H, W, dtype=torch.float32, device=dev)           # the B latents, updated in place
        self.raw = torch.empty(2 * B, model.out_channels, H, W, dtype=torch.float32, device=dev)
        self.noise = torch.empty(B, C, H, W, dtype=torch.float32, device=dev)
        self.y = torch.empty(2 * B, dtype=torch.int64, device=dev)                   # labels | null class
        self.pos = torch.empty(1, self.T, self.d, dtype=torch.float32, device=dev)
        self.pos_rows = 1
        self.mask = torch.ones(2 * B, self.T, dtype=torch.bool, device=dev)
        self.t_idx = torch.empty(2 * B, dtype=torch.int64, device=dev)               # loop index i (respaced)
        self.t_model = torch.empty(2 * B, dtype=torch.int64, device=dev)             # original timestep fed to the model
        self.step_scalar = torch.empty(1, dtype=torch.int64, device=dev)
        self.use_graph = use_graph
        self._graph: Optional[torch.cuda.CUDAGraph] = None
        self._captured = None          # (engine, packed weights, workspace) the graph holds raw pointers into: kept alive, compared per run
        self.captures = 0              # number of graph captures so far (tests assert one per bucket)
        self.launches_per_step = 0
        # the embedding table has a null-class row only when the model was built with class_dropout_prob > 0 (fit.py:94-95): without it the
        # reference's unconditional half raises an index error -- do the same instead of sampling a wrong class
        if model.y_embedder.embedding_table.weight.shape[0] <= model.y_embedder.num_classes:
            raise IndexError("CFG sampling needs the null-class embedding row (class_dropout_prob > 0); this model has none")
        self.row_offset_dev = torch.tensor([int(row_offset)], dtype=torch.int64, device=dev)   # read by the noise kernel at run time

    # ------------------------------------------------------------------ one step (all C-ABI launches)
    def _step(self):
        d, m = self.diffusion, self.model
        tb = d._tables(self.device)
        with rng("sample/model"):
            ops.gather_i64(d._map_tensor(self.device), self.t_idx, self.t_model)
            m.engine.forward(self.x, self.t_model, self.y, self.pos, self.mask, train=False, rows=2 * self.B, out=self.raw)
        with rng("sample/cfg+step"):
            ops.philox_normal(self.noise, self.seed, self.t_idx, self.row_offset_dev)
            ops.cfg_step(self.x, self.raw, self.noise, self.t_idx[: self.B], tb["coef"], cfg_pairs=self.B, cfg_scale=self.cfg_scale,
                         ddim=(self.mode == "ddim"), eta=self.eta, flags=d._flags(self.clip), sample=self.x)
            ops.add_i64(self.t_idx, -1)

    def kernels_per_step(self) -> int:
        """Number of libfit_b200 kernel launches in one step: 7 per block (un-fused: 2 ln_modulate + 4 GEMMs + attention; LN prologue: 2 small
        c-vector GEMMs + 4 GEMMs + attention) + embed / conditioning / final (8) + 4 loop kernels, + row_stats when the residual stream's
        row sums come from the GEMM epilogues (bf16 default), + modulation_pack when the LayerNorm prologue is active."""
        from .engine import final_gemm_ok, ln_prologue_ok, ln_stats_ok
        L = len(self.model.blocks)
        prec, D = self.model.engine.precision, self.model.hidden_size
        fused = ln_prologue_ok(prec, 2 * self.B * self.T, D, self.T) and L > 0
        stats = fused or ln_stats_ok(prec, D)       # row_stats after the embedder; modulation_pack only with the prologue
        # final layer as a prologue GEMM: modulation_pack + c-vector GEMM + GEMM + unpatchify instead of the one CUDA-core kernel
        fin = 3 if stats and final_gemm_ok(prec, 2 * self.B * self.T, D, self.T) else 0
        return 7 * L + 8 + 4 + (1 if fused else 0) + (1 if stats else 0) + fin

    def noise_for_step(self, i: int, row_offset: int = 0, rows: Optional[int] = None) -> Tensor:
        rows = self.B if rows is None else rows
        out = torch.empty(rows, *self.x.shape[1:], dtype=torch.float32, device=self.device)
        self.step_scalar.fill_(i)
        return ops.philox_normal(out, self.seed, self.step_scalar, self.row_offset + row_offset)

    def set_row_offset(self, row_offset: int):
        """Global index of this sampler's first image (the noise key).  Device-resident: changing it does NOT re-capture the graph."""
        self.row_offset = int(row_offset)
        self.row_offset_dev.fill_(self.row_offset)

    # ------------------------------------------------------------------ driver
    def load(self, z: Tensor, y: Tensor, pos: Tensor, mask: Optional[Tensor] = None):
        """Stage one batch: z [B,C,H,W] (host or device), labels [B], RoPE table [1|2B, T, d], optional mask [2B, T]."""
        B = self.B
        self.x.copy_(z, non_blocking=True)
        self.y[:B].copy_(y, non_blocking=True)
        self.y[B:].fill_(self.model.y_embedder.num_classes)   # null class id (sample_lightning.py:40)
        pos = pos if pos.dim() == 3 else pos[None]
        if pos.shape[0] != self.pos.shape[0]:
            self.pos = torch.empty(pos.shape, dtype=torch.float32, device=self.device)
            self._graph = None
        self.pos.copy_(pos, non_blocking=True)
        if mask is not None:
            self.mask.copy_(mask, non_blocking=True)
        else:
            self.mask.fill_(True)
        self.t_idx.fill_(self.diffusion.num_timesteps - 1)

    def run(self, steps: Optional[int] = None):
        """Enqueue `steps` denoising steps (default: all remaining) on the current stream."""
        n = self.diffusion.num_timesteps if steps is None else steps
        with torch.no_grad():
            if not self.use_graph:
                for _ in range(n):
                    self._step()
                return
            eng = self.model.engine
            if self._graph is not None and (eng is not self._captured[0] or eng.weights() is not self._captured[1]):
                # parameters changed (load_state_dict, optimizer step, EMA swap: new packed weights) or set_precision() built a new
                # engine: the captured pointers are stale
                self._graph = None
            if self._graph is None:
                # warm-up outside capture: workspace allocation, weight packing, cudaFuncSetAttribute
                saved = (self.x.clone(), self.t_idx.clone())
                self._step()
                torch.cuda.synchronize(self.device)
                self.x.copy_(saved[0]); self.t_idx.copy_(saved[1])
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g):
                    self._step()
                self._graph = g
                self.captures += 1
                # strong references: an evicted workspace / repacked weight set must outlive the graph that points into it
                self._captured = (eng, eng.weights(), eng.workspace(2 * self.B, self.T, self.device))
                self.x.copy_(saved[0]); self.t_idx.copy_(saved[1])
            for _ in range(n):
                self._graph.replay()

    def sample(self, z: Tensor, y: Tensor, pos: Tensor, mask: Optional[Tensor] = None, out: Optional[Tensor] = None) -> Tensor:
        """Full loop: returns the B final latents (a fresh device tensor, or copied into `out`, e.g. pinned host memory)."""
        self.load(z, y, pos, mask)
        self.run()
        if out is None:
            return self.x.clone()
        out.copy_(self.x, non_blocking=True)
        return out


def shard_rows(total: int, world: int, rank: int) -> Tuple[int, int]:
    """[begin, end) of the images rank `rank` samples when a batch of `total` is sharded over `world` GPUs (no collective:
    each image's trajectory depends only on its own latent, label and noise).  `begin` is the
    ``row_offset`` to give CFGSampler so the noise is keyed by the global image index."""
    base, rem = divmod(total, world)
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


# ---------------------------------------------------------------------------------------------------- batch driver (row f1)
class SamplingDriver:
    """The batch loop of sample_lightning.py:26-70 on top of CFGSampler.

    Per batch the reference draws labels (``torch.randint``) and latents (``torch.randn``), pads the latent tokens to 256
    (``_pad_latent``), builds the RoPE table / mask for the resolution (``_create_pos_embed`` / ``_create_mask``), runs
    ``ddim_sample_loop(model.forward_with_cfg, ...)`` on 2B rows, keeps the first half and un-pads it.  Here the network runs
    on the latent's own token count (padded key tokens are masked out in the reference, so they change nothing for the valid
    ones), the RoPE table of a resolution is built once and cached (``rope.create_pos_embed``, VisionNTK included), and
    there is one CUDA-graph sampler per (batch size, latent height, latent width) bucket.  With ``world`` > 1 every rank takes
    its ``shard_rows`` slice of each batch (no collective); the noise is keyed by the global image index.
    VAE decoding (sample_lightning.py:72-74) is out of scope: the method returns latents."""

    def __init__(self, model, diffusion, cfg_scale: float = 1.5, mode: str = "ddim", eta: float = 0.0, seed: int = 0, max_length: int = 256,
                 rank: int = 0, world: int = 1, max_buckets: int = 4):
        self.model, self.diffusion = model, diffusion
        self.cfg_scale, self.mode, self.eta, self.seed, self.max_length = cfg_scale, mode, eta, seed, max_length
        self.rank, self.world, self.max_buckets = rank, world, max_buckets
        self._samplers = {}
        self._pos = {}

    def _pos_table(self, h: int, w: int) -> Tensor:
        key = (h, w)
        if key not in self._pos:
            from .rope import create_pos_embed
            m = self.model
            table, valid_t = create_pos_embed(h, w, m.patch_size, self.max_length, m.hidden_size // m.num_heads)
            self._pos[key] = table[:, :valid_t].contiguous().pin_memory()
        return self._pos[key]

    def _sampler(self, B: int, h: int, w: int, row_offset: int) -> CFGSampler:
        key = (B, h, w)
        s = self._samplers.get(key)
        if s is None:
            if len(self._samplers) >= self.max_buckets:
                self._samplers.pop(next(iter(self._samplers)))
            s = CFGSampler(self.model, self.diffusion, batch=B, latent_hw=(h, w), cfg_scale=self.cfg_scale, seed=self.seed, mode=self.mode,
                           eta=self.eta)
            self._samplers[key] = s
        s.set_row_offset(row_offset)   # device-resident noise key: one captured graph per (batch, H, W) bucket serves every batch
        return s

    def sample(self, num_samples: int, image_hw: Tuple[int, int], batch_size: int = 100, num_classes: int = 1000,
               generator: Optional[torch.Generator] = None, labels: Optional[Tensor] = None) -> Tuple[Tensor, Tensor]:
        """Returns (latents [n_local, C, H/8, W/8] on the device, labels [n_local]) of this rank's share of ``num_samples`` images."""
        H, W = image_hw[0] // 8, image_hw[1] // 8                   # sample_lightning.py:35
        C = self.model.in_channels
        pos = self._pos_table(H, W)
        outs, labs = [], []
        done = 0
        while done < num_samples:
            cur = min(batch_size, num_samples - done)               # sample_lightning.py:32
            y_all = torch.randint(0, num_classes, (cur,), generator=generator) if labels is None else labels[done:done + cur].cpu()
            z_all = torch.randn(cur, C, H, W, generator=generator)  # host draw (the reference draws on its device generator)
            b0, b1 = shard_rows(cur, self.world, self.rank)
            if b1 > b0:
                s = self._sampler(b1 - b0, H, W, row_offset=done + b0)
                outs.append(s.sample(z_all[b0:b1].pin_memory(), y_all[b0:b1].to(torch.int64).pin_memory(), pos))
                labs.append(y_all[b0:b1])
            done += cur
        dev = next(self.model.parameters()).device
        if not outs:
            return torch.empty(0, C, H, W, device=dev), torch.empty(0, dtype=torch.int64)
        return torch.cat(outs), torch.cat(labs)
