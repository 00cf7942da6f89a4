#!/usr/bin/env python
"""Headline benchmark: FiT-XL/2 256x256 sampling images/sec (250-step DDPM, CFG 1.5, batch 256 per GPU).

    python bench.py [--gpus N] [--steps K] [--warmup W]            # fit_b200 (CUDA) arm
    python bench.py --impl reference [--steps K] [--warmup W]      # the reference algorithm on the host CPU cores

A "step" is ONE denoising step of the whole batch: FiT-XL/2 forward on 2 x batch rows (conditional | null class)
+ the fused CFG + p_sample update.  images/sec = batch / (250 x seconds per step).  With N GPUs every rank samples its
own `--batch` images (weak scaling, no collective on the data path); value = N x batch / (250 x max-over-ranks step time).
One JSON line on stdout (rank 0).  Nothing here reads /root/reference.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

STEPS_PER_IMAGE = 250
METRIC = "FiT-XL/2 sampling images/sec @256^2 (250-step DDPM, CFG 1.5)"


# ------------------------------------------------------------------------------------------------ FLOP model
def forward_flops(D, Hd, L, T, p2c=16):
    """Algorithmic FLOPs of one sample forward with T valid tokens (2 per MAC), SURVEY.md section 8(d)."""
    return L * (T * (8 * D * D + 6 * D * Hd) + 4 * T * T * D + 12 * D * D) + 2 * T * p2c * D + 2 * (256 * D + D * D) + 4 * D * D + 2 * T * D * p2c


MODELS = {"FiT-XL/2": (1152, 3072, 28, 16), "FiT-B/2": (768, 2048, 12, 12), "FiT-S/2": (384, 1024, 12, 6)}


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as fh:
            pk = json.load(fh)
        return {"hbm_gbs": pk["hbm_gbs"], "tf_burst": pk["bf16_tflops"], "tf_sustained": pk.get("bf16_tflops_sustained", pk["bf16_tflops"]),
                "source": "measured (MEASURED_PEAKS.json)"}
    return {"hbm_gbs": 6650.0, "tf_burst": 1590.0, "tf_sustained": 1400.0, "source": "fallback (B200_PROFILING.md)"}


# ------------------------------------------------------------------------------------------------ clocks sampler
class ClockSampler:
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            if len(r) < 6:
                continue
            try:
                sm.append(float(r[0])); mx = float(r[1])
            except ValueError:
                continue
            for nme, v in zip(names, r[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(nme)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------------ CPU arm (oracle port)
def cpu_reference_run(model_name: str, images: int, steps: int, warmup: int):
    """The reference algorithm (oracle restatement of fit.py / gaussian_diffusion.py, pinned to the reference's goldens) on
    the host cores: `images` latents (2 x images rows with CFG), `steps` timed denoising steps.  Returns (img/s, s/step, info)."""
    import torch
    from oracle import diffusion_oracle as do
    from oracle import fit_oracle as fo
    from oracle.pos_embed_oracle import precompute_freqs_cis_2d
    torch.set_num_threads(os.cpu_count() or 1)
    cfg = fo.model_config(model_name)
    sd = fo.random_state_dict(cfg, seed=0)
    d = cfg["hidden_size"] // cfg["num_heads"]
    tb, tmap = do.create(str(STEPS_PER_IMAGE))
    g = torch.Generator().manual_seed(123)
    z = torch.randn(images, 4, 32, 32, generator=g)
    x = torch.cat([z, z], 0)
    y = torch.cat([torch.randint(0, 1000, (images,), generator=g), torch.full((images,), 1000)])
    pos = torch.tensor(precompute_freqs_cis_2d(d, 16, 16, max_length=256), dtype=torch.float32)[None]
    mask = torch.ones(2 * images, 256, dtype=torch.bool)
    tmap_t = torch.tensor(tmap)
    times = []
    with torch.no_grad():
        for k in range(warmup + steps):
            i = STEPS_PER_IMAGE - 1 - k
            t0 = time.perf_counter()
            t = torch.full((2 * images,), i, dtype=torch.long)
            eps = fo.forward_with_cfg(sd, cfg, x, tmap_t[t], y, pos, mask, 1.5)
            x = do.p_sample(tb, x, t, eps, torch.randn(x.shape, generator=g))["sample"]
            if k >= warmup:
                times.append(time.perf_counter() - t0)
    s_step = sum(times) / len(times)
    return images / (STEPS_PER_IMAGE * s_step), s_step, {"cores": torch.get_num_threads(), "images": images, "steps": steps}


def cpu_unmodified_reference_run(model_name: str, images: int, steps: int, warmup: int):
    """Same sample as cpu_reference_run, but through the UNMODIFIED reference modules (FiT.forward_with_cfg + SpacedDiffusion.p_sample,
    sample_lightning.py:36-70) imported from /root/reference (build container only: the GPU box has no /root/reference).
    Returns None when the reference tree is absent."""
    if not os.path.isdir("/root/reference/FiT"):
        return None
    import torch
    from oracle import fit_oracle as fo
    from oracle.make_goldens import import_reference
    torch.set_num_threads(os.cpu_count() or 1)
    FiT, FiT_models, create_diffusion, precompute = import_reference()
    cfg = fo.model_config(model_name)
    model = FiT_models[model_name]()
    model.load_state_dict(fo.random_state_dict(cfg, seed=0))
    model.eval()
    diffusion = create_diffusion(str(STEPS_PER_IMAGE))
    d = cfg["hidden_size"] // cfg["num_heads"]
    g = torch.Generator().manual_seed(123)
    z = torch.randn(images, 4, 32, 32, generator=g)
    x = torch.cat([z, z], 0)
    y = torch.cat([torch.randint(0, 1000, (images,), generator=g), torch.full((images,), 1000)])
    pos = torch.tensor(precompute(d, 16, 16, max_length=256), dtype=torch.float32)[None]
    mask = torch.ones(2 * images, 256, dtype=torch.bool)
    kw = dict(y=y, pos=pos, mask=mask, cfg_scale=1.5)
    times = []
    with torch.no_grad():
        for k in range(warmup + steps):
            i = STEPS_PER_IMAGE - 1 - k
            t0 = time.perf_counter()
            t = torch.full((2 * images,), i, dtype=torch.long)
            x = diffusion.p_sample(model.forward_with_cfg, x, t, clip_denoised=False, model_kwargs=kw)["sample"]
            if k >= warmup:
                times.append(time.perf_counter() - t0)
    s_step = sum(times) / len(times)
    return images / (STEPS_PER_IMAGE * s_step), s_step, {"cores": torch.get_num_threads(), "images": images, "steps": steps}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps = max(1, min(args.steps, 4))
    warm = 1 if args.warmup > 0 else 0
    kind = "reference"
    res = cpu_unmodified_reference_run(args.model, images=2, steps=steps, warmup=warm)
    if res is None:   # GPU box: /root/reference does not exist there -> the pinned restatement
        kind = "port"
        res = cpu_reference_run(args.model, images=2, steps=steps, warmup=warm)
    val, s_step, info = res
    sample = (f"{info['images']} images ({2 * info['images']} rows with CFG) x {steps} of the 250 denoising steps, fp32, "
              f"{info['cores']} threads; img/s = images / (250 x s/step)")
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": "images/sec", "n_gpus": args.gpus, "steps": steps, "warmup": warm,
            "ms_per_step": s_step * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"{args.model} 256x256 (32x32x4 latent, 256 tokens), 250-step DDPM p_sample, CFG 1.5; CPU sample of the same step",
                       "steps_per_image": STEPS_PER_IMAGE},
            "cpu_baseline": {"value": val, "unit": "images/sec", "cores": info["cores"], "kind": kind, "sample": sample},
            "e2e": {"value": val, "unit": "images/sec", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------ CUDA arm
def randomize_zero_init(model, seed=1):
    """adaLN-Zero zero-initialises the modulation and output layers (fit.py:421-430): the random-init network would be the
    identity and output 0.  Re-draw those parameters N(0, 0.02) so the benchmark computes on non-degenerate values."""
    import torch
    g = torch.Generator(device="cpu").manual_seed(seed)
    for name, p in model.named_parameters():
        if "adaLN_modulation" in name or name.startswith("final_layer.linear") or name.endswith(".bias"):
            p.data.copy_(torch.randn(p.shape, generator=g) * 0.02)


def traffic_tag():
    """Newest committed `ncu --set full` summary of the block kernels (profiles/<tag>_ncu_full_block_kernels.csv); its header line names the
    git revision it was captured from."""
    import glob
    tags = sorted(os.path.basename(f).split("_ncu_full")[0] for f in glob.glob(os.path.join(ROOT, "profiles", "*_ncu_full_block_kernels.csv")))
    return tags[-1] if tags else "none"


def ncu_traffic(tag="r1f"):
    """DRAM bytes (read + write) per launch of the block kernels from the committed `ncu --set full` capture of tools/prof_kernels.py at
    the same shape (profiles/<tag>_ncu_full_block_kernels.csv); {} when the summary is absent."""
    import csv
    path = os.path.join(ROOT, "profiles", f"{tag}_ncu_full_block_kernels.csv")
    if not os.path.exists(path):
        return {}
    out, gate_seen = {}, 0
    with open(path) as fh:
        rows = list(csv.reader(fh))
    hdr = rows[0]
    ir, iw = [i for i, h in enumerate(hdr) if h.startswith("dram_read")][0], [i for i, h in enumerate(hdr) if h.startswith("dram_write")][0]
    scale = lambda h: 1e9 if "Gbyte" in h else 1e6 if "Mbyte" in h else 1e3 if "Kbyte" in h else 1.0  # noqa: E731
    for r in rows[1:]:
        name, b = r[0], float(r[ir]) * scale(hdr[ir]) + float(r[iw]) * scale(hdr[iw])
        name = name.replace("void ", "").replace("fitb200::", "")   # ncu prints the namespace for some kernels, not for others
        key = None
        lna = name.rstrip().endswith("1, 1>")   # LN-prologue instantiations
        if name.startswith("gemm_tc_kernel<256, 3"): key = "gemm_tc[ln+swiglu fc1_g|fc1_x]" if lna else "gemm_tc[swiglu fc1_g|fc1_x]"
        elif name.startswith("gemm_tc_kernel<256, 4"): key = "gemm_tc[ln+qkv+rope]" if lna else "gemm_tc[qkv+rope]"
        elif name.startswith("gemm_tc_kernel<256, 2"):
            key = "gemm_tc[proj+gate+resid]" if gate_seen == 0 else "gemm_tc[fc2+gate+resid]"
            gate_seen += 1
        elif name.startswith("attn_tc"): key = "attn_tc[softmax(qk^T)v]"
        elif name.startswith("ln_modulate"): key = "ln_modulate"
        elif name.startswith("ln_apply"): key = "ln_apply"
        if key and key not in out:
            out[key] = b
    return out


def time_kernel(fn, iters, flush):
    """(best, mean) seconds over `iters` launches, each after an L2 flush, CUDA events on the launch stream.  The roofline's `achieved` /
    `frac` use the MEAN launch (`*_best` alongside: MEASURED_PEAKS.json's burst peaks are best-of-10 figures)."""
    import torch
    fn(); fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        flush()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e-3)
    return min(ts), sum(ts) / len(ts)


def kernel_rooflines(model, sampler, peaks, iters=10):
    """Per-kernel roofline, timed alone with CUDA events on the launch stream (L2 flushed between iterations)."""
    import torch
    from fit_b200 import ops
    from fit_b200._lib import EPI_BIAS, EPI_BIAS_GATE_RESID, EPI_BIAS_ROPE, EPI_SWIGLU
    eng = model.engine
    W = eng.weights()
    ws = eng.workspace(2 * sampler.B, sampler.T, sampler.device)
    e = W.blocks[0]
    D, T, M = model.hidden_size, sampler.T, ws.xs.shape[0]
    Hp = W.Hp
    h, dh = model.num_heads, model.hidden_size // model.num_heads
    sh1, sc1, g1, sh2, sc2, g2 = ws.mod_views[0]
    scratch = torch.empty(256 * 1024 * 1024 // 4, dtype=torch.float32, device=sampler.device)  # 256 MB > 126 MB L2
    scratch.fill_(1.0)
    # L2 flush by READING 256 MB (> 126 MB L2): a write flush leaves the L2 full of dirty lines whose write-back is then charged to the
    # timed kernel (ln_modulate: 0.158 ms after a write flush vs 0.126 ms under ncu's own cache control)
    flush = lambda: scratch.sum()  # noqa: E731
    bpe = ws.xs.element_size()
    res = {}

    def add(name, fn, flops, bytes_):
        t_best, t = time_kernel(fn, iters, flush)
        tf, gbs = flops / t / 1e12, bytes_ / t / 1e9
        tensor_bound = flops / (peaks["tf_burst"] * 1e12) >= bytes_ / (peaks["hbm_gbs"] * 1e9)
        peak = peaks["tf_burst"] if tensor_bound else peaks["hbm_gbs"]
        ach = tf if tensor_bound else gbs
        res[name] = {"bound": "tensor" if tensor_bound else "hbm", "ms": t * 1e3, "ms_best": t_best * 1e3,
                     "achieved": ach, "achieved_best": ach * t / t_best, "peak": peak, "unit": "TFLOP/s" if tensor_bound else "GB/s",
                     "frac": ach / peak, "frac_best": ach * t / t_best / peak, "timing": "mean of %d launches (best alongside)" % iters,
                     "algorithmic_bytes": bytes_, "algorithmic_flops": flops, "traffic": None}

    fused = getattr(ws, "modA", None) is not None   # LayerNorm + modulate as the GEMM prologue (FiTEngine.block)
    R2 = 2 * sampler.B
    if fused:
        from fit_b200.engine import LN_EPS
        ops.row_stats(ws.xs, ws.parts)
        add("gemm_tc[ln+swiglu fc1_g|fc1_x]", lambda: ops.gemm(ws.xs, e["fc1_w"], None, ws.g, EPI_SWIGLU, rows_per_sample=T,
                                                              ln=(ws.modA[1, :R2], ws.cf, ws.parts, LN_EPS)),
            2.0 * M * D * 2 * Hp, (M * D + 2 * Hp * D + M * Hp) * 2)
        add("gemm_tc[ln+qkv+rope]", lambda: ops.gemm(ws.xs, e["qkv_w"], None, ws.qkv, EPI_BIAS_ROPE, rope=sampler.pos, rope_cols=2 * D, rows_per_sample=T,
                                                     ln=(ws.modA[0, :R2], ws.cq, ws.parts, LN_EPS)),
            2.0 * M * D * 3 * D, (M * D + 3 * D * D + M * 3 * D) * 2)
        add("gemm_tc[ln vectors qkv]", lambda: ops.gemm(ws.modA[0], e["qkv_w"], e["qkv_b"], ws.cq, EPI_BIAS, bias_row0=R2),
            2.0 * 2 * R2 * D * 3 * D, (2 * R2 * D + 3 * D * D) * 2 + 2 * R2 * 3 * D * 4)
        add("gemm_tc[ln vectors fc1]", lambda: ops.gemm(ws.modA[1], e["fc1_w"], e["fc1_b"], ws.cf, EPI_BIAS, bias_row0=R2),
            2.0 * 2 * R2 * D * 2 * Hp, (2 * R2 * D + 2 * Hp * D) * 2 + 2 * R2 * 2 * Hp * 4)
    else:
        add("gemm_tc[swiglu fc1_g|fc1_x]", lambda: ops.gemm(ws.h, e["fc1_w"], e["fc1_b"], ws.g, EPI_SWIGLU),
            2.0 * M * D * 2 * Hp, (M * D + 2 * Hp * D + M * Hp) * 2)
        add("gemm_tc[qkv+rope]", lambda: ops.gemm(ws.h, e["qkv_w"], e["qkv_b"], ws.qkv, EPI_BIAS_ROPE, rope=sampler.pos, rope_cols=2 * D),
            2.0 * M * D * 3 * D, (M * D + 3 * D * D + M * 3 * D) * 2)
    so = getattr(ws, "parts", None)   # the engine's GATE_RESID GEMMs also write the row sums of their output
    add("gemm_tc[fc2+gate+resid]", lambda: ops.gemm(ws.g, e["fc2_w"], e["fc2_b"], ws.xs, EPI_BIAS_GATE_RESID, gate=g2, ld_gate=W.n_mod,
                                                     rows_per_sample=T, resid=ws.xs, stats_out=so), 2.0 * M * Hp * D, (M * Hp + D * Hp) * 2 + 2 * M * D * bpe)
    add("gemm_tc[proj+gate+resid]", lambda: ops.gemm(ws.att, e["proj_w"], e["proj_b"], ws.xs, EPI_BIAS_GATE_RESID, gate=g1, ld_gate=W.n_mod,
                                                      rows_per_sample=T, resid=ws.xs, stats_out=so), 2.0 * M * D * D, (M * D + D * D) * 2 + 2 * M * D * bpe)
    add("attn_tc[softmax(qk^T)v]", lambda: ops.attention(ws.qkv, None, ws.kv_len, sampler.mask, 2 * sampler.B, T, h, dh, ws.att, prerotated=True),
        4.0 * (2 * sampler.B) * T * T * D, (M * 3 * D + M * D) * 2)
    if not fused:
        if getattr(ws, "parts", None) is not None:   # default bf16 path: row sums from the GATE_RESID epilogues, streaming modulate pass
            from fit_b200.engine import LN_EPS
            ops.row_stats(ws.xs, ws.parts)
            add("ln_apply", lambda: ops.ln_apply(ws.xs, ws.parts, sh1, sc1, W.n_mod, T, ws.h, LN_EPS), 0.0, M * D * (bpe + 2) + ws.parts.numel() * 4)
            add("row_stats", lambda: ops.row_stats(ws.xs, ws.parts), 0.0, M * D * bpe + ws.parts.numel() * 4)
        else:
            add("ln_modulate", lambda: ops.ln_modulate(ws.xs, sh1, sc1, W.n_mod, T, ws.h), 0.0, M * D * (bpe + 2))
    tb = sampler.diffusion._tables(sampler.device)
    nel = sampler.x.numel()
    add("cfg_ddpm_step", lambda: ops.cfg_step(sampler.x, sampler.raw, sampler.noise, sampler.t_idx[: sampler.B], tb["coef"], cfg_pairs=sampler.B,
                                              cfg_scale=1.5, sample=sampler.x), 0.0, nel * 4 * 5)
    # family (3) kernels (north_star: "HBM GB/s of the elementwise and step kernels"): µs-scale, launch-latency bound at this size
    R = 2 * sampler.B
    p_, Cin = model.patch_size, model.in_channels
    add("embed_tokens", lambda: ops.embed_tokens(sampler.x, W.x_w, W.x_b, ws.xs, R, T, p_, Cin), 0.0, sampler.x.numel() * 4 + M * D * bpe)
    fsh, fsc = ws.fin_views
    if getattr(ws, "fin_mod", None) is not None and ws.fin_c is not None:
        # the engine's final layer: tcgen05 GEMM with the LayerNorm + modulate prologue, 16 output columns (one pass over the residual stream)
        from fit_b200.engine import LN_EPS as _EPS
        add("final_layer[ln+gemm N=16]", lambda: ops.gemm(ws.xs, W.fin_w16, None, ws.fin_tok, EPI_BIAS, rows_per_sample=T,
                                                          ln=(ws.fin_mod[0, :R], ws.fin_c, ws.parts, _EPS)), 0.0, M * D * bpe + ws.fin_tok.numel() * 4)
    else:
        add("final_layer", lambda: ops.final_layer(ws.xs, fsh, fsc, W.n_mod, W.fin_w, W.fin_b, sampler.raw, R, T, p_, model.out_channels), 0.0,
            M * D * bpe + sampler.raw.numel() * 4)
    add("philox_normal", lambda: ops.philox_normal(sampler.noise, 1234, sampler.t_idx, sampler.row_offset_dev), 0.0, sampler.noise.numel() * 4)
    tag = traffic_tag()
    for k, b in ncu_traffic(tag).items():
        if k in res:
            res[k]["traffic"] = b
            res[k]["traffic_source"] = f"profiles/{tag}_ncu_full_block_kernels.csv (dram__bytes_read.sum + dram__bytes_write.sum, one launch)"
    return res


# ------------------------------------------------------------------------------------------------ extra workloads (BASELINE configs 3-4)
def strong_scaling_run(model, diffusion, args, world, rank, dev, pos, barrier, all_max):
    """BASELINE config 3 as written: global batch `--batch` SHARDED over the GPUs (batch/world images per GPU, no collective).
    Returns images/sec of the whole job (same metric as the headline line, fixed total work)."""
    import torch
    from fit_b200.sampler import CFGSampler, shard_rows
    b0, b1 = shard_rows(args.batch, world, rank)
    Bs = b1 - b0
    g = torch.Generator().manual_seed(2000 + rank)
    z = torch.randn(Bs, 4, 32, 32, generator=g).pin_memory()
    y = torch.randint(0, 1000, (Bs,), generator=g).pin_memory()
    smp = CFGSampler(model, diffusion, batch=Bs, latent_hw=(32, 32), cfg_scale=1.5, seed=1234, row_offset=b0)
    smp.load(z, y, pos)
    smp.run(max(args.warmup, 1))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    smp.run(args.steps)
    e1.record()
    barrier()
    ms = all_max(e0.elapsed_time(e1)) / args.steps
    del smp
    return {"img_s": args.batch / (STEPS_PER_IMAGE * ms * 1e-3), "ms_per_step": ms, "per_gpu_batch": Bs, "global_batch": args.batch,
            "token_rows_per_gpu": 2 * Bs * 256}


def train_run(model, args, world, rank, dev, barrier, all_max, samples=32, iters=5):
    """BASELINE config 4: FiT-XL/2 bf16 training step, `samples` synthetic latents per GPU (global batch 256 at 8 GPUs), forward + hand-written
    backward + per-block NCCL gradient all-reduce (GradSync, overlapped on a side stream) + fused AdamW/EMA (FiT/train.py:45-69,215-217,283-291).
    Timed with CUDA events, max over ranks.  Also timed: the same step with the exchange switched off (local gradients only) and the
    all-reduce of the same buckets alone, which give the exposed / hidden share of the collective."""
    import torch
    import torch.distributed as dist
    from fit_b200 import _lib as L
    from fit_b200.optim import FusedAdamWEMA
    from fit_b200.train import CapturedTrainStep, GradSync, masked_mse_loss
    from fit_b200.rope import precompute_freqs_cis_2d
    D, Hd, Ld, heads = MODELS[args.model]
    T, d = 256, D // heads
    was_training = model.training
    model.train()
    g = torch.Generator().manual_seed(300 + rank)
    x = torch.randn(samples, T, 16, generator=g).to(dev)
    tgt = torch.randn(samples, T, 16, generator=g).to(dev)
    t = torch.randint(0, 1000, (samples,), generator=g).to(dev)
    y = torch.randint(0, 1000, (samples,), generator=g).to(dev)
    pos = torch.tensor(precompute_freqs_cis_2d(d, 16, 16, max_length=256), dtype=torch.float32)[None].expand(samples, T, d).contiguous().to(dev)
    mask = torch.ones(samples, T, dtype=torch.bool, device=dev)
    sync = GradSync().attach(model) if world > 1 else None
    opt = FusedAdamWEMA(model.parameters(), lr=1e-4, weight_decay=0.0, ema_decay=0.9999)
    # one GPU: the whole step (forward, loss, backward, AdamW/EMA) is ONE replayed CUDA graph; data parallel: issued eagerly, the per-block
    # all-reduces overlap the backward on a side stream (GradSync.finish runs inside the autograd node's backward)
    cstep = CapturedTrainStep(model, opt)

    def step():
        return cstep(x, t, y, pos, mask, tgt)

    def timed(fn, n):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record()
        for _ in range(n):
            fn()
        e1.record()
        barrier()
        return all_max(e0.elapsed_time(e1)) / n

    for _ in range(2):
        loss = step().clone()      # (a captured step returns a static tensor: keep the value of the last warm-up step)
    ms = timed(step, iters)
    res = {"train_step_ms": ms, "samples_per_gpu": samples, "global_batch": samples * world,
           "train_samples_s": samples * world / (ms * 1e-3),
           "train_tflops_per_gpu": 3 * samples * forward_flops(D, Hd, Ld, T) / (ms * 1e-3) / 1e12,
           "train_loss": float(loss.detach()),
           "train_graph": cstep.captures > 0,
           "train_step": ("fwd + bwd + fused AdamW/EMA replayed as one CUDA graph" if cstep.captures > 0 else
                          "fwd + bwd + per-block NCCL all-reduce (AVG, fp32, overlapped) + fused AdamW/EMA, issued eagerly") + "; 3 x forward FLOPs counted"}
    if sync is not None:
        with sync.no_sync():        # always the eager step: a captured graph has the collectives baked in
            ms_local = timed(lambda: cstep._eager(x, t, y, pos, mask, tgt), iters)
        # the same buckets, all-reduced alone (nothing to overlap with)
        names = [n for n, _ in model.named_parameters()]
        sizes = {}
        for n, p in model.named_parameters():
            key = n.split(".")[1] if n.startswith("blocks.") else "rest"
            sizes[key] = sizes.get(key, 0) + p.numel()
        bufs = [torch.zeros(v, dtype=torch.float32, device=dev) for v in sizes.values()]
        def ar_alone():
            for b in bufs:
                dist.all_reduce(b, op=dist.ReduceOp.AVG)
        ar_alone()
        ms_ar = timed(ar_alone, 3)
        exposed = max(ms - ms_local, 0.0)
        res.update({"train_step_local_ms": ms_local, "allreduce_alone_ms": ms_ar, "allreduce_exposed_ms": exposed,
                    "allreduce_overlap_frac": max(0.0, 1.0 - exposed / ms_ar) if ms_ar > 0 else None,
                    "allreduce_bytes": 4 * sum(sizes.values()), "allreduce_buckets": len(bufs), "collective": "ncclAllReduce (AVG, fp32)"})
        object.__setattr__(model, "_grad_sync", None)
        n_sm = torch.cuda.get_device_properties(dev).multi_processor_count
        L.check(L.lib().fitb200_set_sm_limit(n_sm), "set_sm_limit")   # GradSync reserved SMs for NCCL: give them back
    del cstep, opt
    for p in model.parameters():
        p.grad = None
    model.train(was_training)
    object.__setattr__(model, "_train_engine", None)
    torch.cuda.empty_cache()
    return res


def eager_torch_run(model, args, dev, pos, steps=2):
    """The same-box bar: the reference's arithmetic as eager PyTorch (cuBLAS / ATen kernels) on THIS GPU at the headline shape -- the pinned
    restatement oracle/fit_oracle.py + diffusion_oracle.p_sample (checker code used as a timed baseline, like cpu_baseline), fp32 with TF32
    matmuls (torch.set_float32_matmul_precision('high'), FiT/train.py:28) and under bf16 autocast."""
    import torch
    from oracle import diffusion_oracle as do
    from oracle import fit_oracle as fo
    cfg = fo.model_config(args.model)
    sd = {k: v.detach() for k, v in model.state_dict().items()}
    B = args.batch
    tb, tmap = do.create(str(STEPS_PER_IMAGE))
    g = torch.Generator().manual_seed(5)
    z = torch.randn(B, 4, 32, 32, generator=g).to(dev)
    y = torch.cat([torch.randint(0, 1000, (B,), generator=g), torch.full((B,), 1000)]).to(dev)
    mask = torch.ones(2 * B, 256, dtype=torch.bool, device=dev)
    posd = pos.to(dev)
    tmap_t = torch.as_tensor(tmap).to(dev)
    out = {}
    old = torch.get_float32_matmul_precision()
    torch.set_float32_matmul_precision("high")
    try:
        for name, ctx in (("tf32", None), ("bf16_autocast", torch.autocast("cuda", dtype=torch.bfloat16))):
            x = torch.cat([z, z], 0)
            ts = []
            with torch.no_grad():
                for k in range(1 + steps):
                    i = STEPS_PER_IMAGE - 1 - k
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record()
                    t = torch.full((2 * B,), i, dtype=torch.long, device=dev)
                    if ctx is None:
                        eps = fo.forward_with_cfg(sd, cfg, x, tmap_t[t], y, posd, mask, 1.5)
                    else:
                        with ctx:
                            eps = fo.forward_with_cfg(sd, cfg, x, tmap_t[t], y, posd, mask, 1.5)
                    x = do.p_sample(tb, x, t, eps.float(), torch.randn_like(x))["sample"]
                    e1.record()
                    torch.cuda.synchronize()
                    if k >= 1:
                        ts.append(e0.elapsed_time(e1))
            ms = sum(ts) / len(ts)
            out[name] = {"img_s": B / (STEPS_PER_IMAGE * ms * 1e-3), "ms_per_step": ms}
            del x, eps
            torch.cuda.empty_cache()
    finally:
        torch.set_float32_matmul_precision(old)
    return out


def run_cuda_arm(args):
    import torch
    import torch.distributed as dist
    from fit_b200.diffusion import create_diffusion
    from fit_b200.models.fit import FiT_models
    from fit_b200.sampler import CFGSampler
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py: no CUDA device; fit_b200 has no CPU fallback (use --impl reference for the CPU arm)")
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("NCCL_MAX_CTAS", os.environ.get("FITB200_DDP_RESERVE_SMS") or ("16" if int(os.environ.get("WORLD_SIZE", "1")) > 2 else "8"))   # the DDP leg: NCCL stays inside the SMs GradSync leaves free
        # NCCL prints its version banner on STDOUT when the communicator is created (NCCL_DEBUG >= VERSION on the GPU boxes; WARN still
        # prints it).  stdout carries exactly one JSON line, so file descriptor 1 points at stderr while the communicator comes up.
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=dev)
            dist.barrier()
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved, 1)
            os.close(saved)
    peaks = load_peaks()
    D, Hd, L, heads = MODELS[args.model]
    B = args.batch
    torch.manual_seed(0)
    with torch.device(dev):
        model = FiT_models[args.model](precision=args.precision)
    randomize_zero_init(model)
    model = model.eval()
    diffusion = create_diffusion(str(STEPS_PER_IMAGE))
    dh = D // heads
    from fit_b200.rope import precompute_freqs_cis_2d
    pos = torch.tensor(precompute_freqs_cis_2d(dh, 16, 16, max_length=256), dtype=torch.float32)[None]
    g = torch.Generator().manual_seed(1000 + rank)
    z_host = torch.randn(B, 4, 32, 32, generator=g).pin_memory()
    y_host = torch.randint(0, 1000, (B,), generator=g).pin_memory()
    out_host = torch.empty(B, 4, 32, 32).pin_memory()
    sampler = CFGSampler(model, diffusion, batch=B, latent_hw=(32, 32), cfg_scale=1.5, seed=1234, row_offset=rank * B)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- value: device-resident inputs, K graph-replayed steps, CUDA events, max over ranks
    sampler.load(z_host, y_host, pos)
    sampler.run(args.warmup if args.warmup > 0 else 1)
    clocks = ClockSampler(local)
    barrier()
    if rank == 0:
        clocks.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    sampler.run(args.steps)
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    clk = clocks.stop() if rank == 0 else None
    # ---- e2e: the user-level call with HOST buffers: H2D of latents + labels + RoPE table, K steps, D2H of the result
    barrier()
    t0 = time.perf_counter()
    sampler.load(z_host, y_host, pos)
    sampler.run(args.steps)
    out_host.copy_(sampler.x, non_blocking=True)
    barrier()
    e2e_s = time.perf_counter() - t0
    tms = torch.tensor([ms, e2e_s * 1e3], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
    ms, e2e_ms = float(tms[0]), float(tms[1])
    assert bool(torch.isfinite(out_host).all()), "non-finite latents"

    def all_max(v):
        tv = torch.tensor([v], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tv, op=dist.ReduceOp.MAX)
        return float(tv[0])

    extra = {}
    if not args.no_extras:
        # ---- config 3 as written (strong scaling: the global batch sharded) and config 4 (DDP training step) -- all ranks take part
        if world > 1:
            extra["strong"] = strong_scaling_run(model, diffusion, args, world, rank, dev, pos, barrier, all_max)
        else:
            extra["strong"] = {"img_s": world * B / (STEPS_PER_IMAGE * ms / args.steps * 1e-3), "ms_per_step": ms / args.steps, "per_gpu_batch": B,
                               "global_batch": B, "token_rows_per_gpu": 2 * B * 256, "note": "N = 1: identical to the headline run"}
        extra["strong_img_s"] = extra["strong"]["img_s"]
        extra["train"] = train_run(model, args, world, rank, dev, barrier, all_max)
        extra["train_step_ms"] = extra["train"]["train_step_ms"]
        barrier()
    if rank == 0:
        ms_step = ms / args.steps
        value = world * B / (STEPS_PER_IMAGE * ms_step * 1e-3)
        e2e_val = world * B / (STEPS_PER_IMAGE * (e2e_ms / args.steps) * 1e-3)
        flops_img = 2 * STEPS_PER_IMAGE * forward_flops(D, Hd, L, 256)
        kr = kernel_rooflines(model, sampler, peaks)
        dom_key = "gemm_tc[ln+swiglu fc1_g|fc1_x]" if "gemm_tc[ln+swiglu fc1_g|fc1_x]" in kr else "gemm_tc[swiglu fc1_g|fc1_x]"
        dom = kr[dom_key]
        roof = dict(dom, kernel="gemm_tc_kernel<256, SWIGLU%s> (fc1_g|fc1_x, M=%d K=%d N=%d)" % (", LN prologue" if "ln+" in dom_key else "", 2 * B * 256, D, 2 * Hd),
                    peak_source=peaks["source"],
                    whole_step_tflops=value / world * flops_img / 1e12, whole_step_frac_of_sustained=value / world * flops_img / 1e12 / peaks["tf_sustained"],
                    kernels=kr)
        if not args.no_extras:
            eg = eager_torch_run(model, args, dev, pos)
            extra["torch_eager_b200"] = eg
            extra["torch_eager_b200_img_s"] = eg["bf16_autocast"]["img_s"]
            extra["torch_eager_b200_tf32_img_s"] = eg["tf32"]["img_s"]
        cpu = None
        if not args.no_cpu_baseline:
            cv, cs, info = cpu_reference_run(args.model, images=1, steps=2, warmup=1)
            cpu = {"value": cv, "unit": "images/sec", "cores": info["cores"], "kind": "port",
                   "sample": f"1 image (2 rows with CFG) x 2 of the 250 denoising steps of the same {args.model} workload, fp32 oracle port, "
                             f"{cs:.2f} s/step; img/s = 1 / (250 x s/step)"}
        line = {"metric": METRIC, "value": value, "unit": "images/sec", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "bf16" if args.precision == "bf16" else "f32", "data": "synthetic",
                "config": {"workload": f"{args.model} 256x256 (32x32x4 latent, 256 tokens), 250-step DDPM p_sample, CFG 1.5, batch {B} per GPU "
                                       f"({2 * B} rows per step), random-init weights (adaLN-Zero params re-drawn N(0,0.02))",
                           "steps_per_image": STEPS_PER_IMAGE, "per_gpu_batch": B, "global_batch": world * B, "parallelism": f"batch-sharded x{world}, no collective",
                           "l2": "activations per step (>300 MB per tensor) exceed the 126 MB L2; per-kernel timings: best of 10 launches (as the measured peaks), L2 flushed by reading a 256 MB buffer before each",
                           "step_definition": "one denoising step of the whole batch; images/sec = global_batch / (250 x s/step)"},
                "clocks": clk, "gpu_launches": args.steps * sampler.kernels_per_step(),
                "e2e": {"value": e2e_val, "unit": "images/sec", "h2d_bytes_per_step": (z_host.numel() * 4 + y_host.numel() * 8 + pos.numel() * 4) / args.steps,
                        "d2h_bytes_per_step": out_host.numel() * 4 / args.steps,
                        "note": "CFGSampler.load (pinned host latents/labels/RoPE table -> device) + K graph-replayed steps + device -> pinned host copy of the latents"},
                "roofline": roof, "cpu_baseline": cpu, "extra": extra}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    # stdout carries exactly one JSON line: keep NCCL's version banner (NCCL_DEBUG=VERSION on the GPU boxes) off it
    if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
        os.environ["NCCL_DEBUG"] = "WARN"
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="fit_b200", choices=["fit_b200", "reference"])
    ap.add_argument("--model", default="FiT-XL/2", choices=sorted(MODELS))
    ap.add_argument("--batch", type=int, default=256, help="images per GPU")
    ap.add_argument("--precision", default="bf16", choices=["bf16", "fp32"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the strong-scaling, training-step and eager-torch side measurements")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_cuda_arm(args)


if __name__ == "__main__":
    main()
