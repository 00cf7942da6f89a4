"""Fused AdamW + EMA optimizer step.

Mirrors what the reference training loop runs after backward / all-reduce: ``torch.optim.AdamW(self.parameters(), lr=1e-4,
weight_decay=0)`` (FiT/train.py:215-217) wrapped by the EMA optimizer (FiT/ema.py:231-278: ``step`` = optimizer step, then every
``every_n_steps`` the EMA copy is updated with ``ema = ema * decay + p * (1 - decay)``, ema.py:149-154).  Here both are ONE CUDA
kernel over a chunk table of all parameter tensors (``fitb200_adamw_ema_step``): p, g, m, v, ema are read once and p, m, v, ema
written once.  Same constructor arguments and state names (``exp_avg``, ``exp_avg_sq``, ``step``) as torch.optim.AdamW; ``ema_params``,
``swap_model_weights`` and ``swap_ema_weights`` follow FiT/ema.py:280-313.  No CPU fallback.
"""
from __future__ import annotations

import contextlib
import ctypes as C
import math
from typing import Iterable, List, Optional, Tuple

import torch

from . import _lib as L

ADAM_CHUNK = 16384  # FITB200_ADAM_CHUNK


class _AdamTensor(C.Structure):
    """Mirror of fitb200_adam_tensor (include/fit_b200.h)."""
    _fields_ = [("p", C.c_void_p), ("g", C.c_void_p), ("m", C.c_void_p), ("v", C.c_void_p), ("ema", C.c_void_p), ("numel", C.c_int64),
                ("shadow", C.c_void_p), ("shadow_ld", C.c_int64), ("shadow_dtype", C.c_int32), ("row_len", C.c_int32),
                ("pack_off", C.c_int32), ("reserved_", C.c_int32)]


class FusedAdamWEMA:
    def __init__(self, params: Iterable[torch.nn.Parameter], lr: float = 1e-4, betas: Tuple[float, float] = (0.9, 0.999), eps: float = 1e-8,
                 weight_decay: float = 0.0, ema_decay: Optional[float] = 0.9999, every_n_steps: int = 1):
        self.params: List[torch.nn.Parameter] = [p for p in params if p.requires_grad]
        if not self.params:
            raise ValueError("optimizer got an empty parameter list")
        for p in self.params:
            if not p.is_cuda or p.dtype != torch.float32 or not p.is_contiguous():
                raise RuntimeError("fit_b200.optim: parameters must be contiguous fp32 CUDA tensors (no CPU fallback)")
        if ema_decay is not None and not 0.0 <= ema_decay <= 1.0:
            raise ValueError("Decay must be within 0 and 1")   # FiT/ema.py:47-48
        self.lr, self.betas, self.eps, self.weight_decay = lr, betas, eps, weight_decay
        self.ema_decay, self.every_n_steps = ema_decay, every_n_steps
        self.step_count = 0
        self.exp_avg = [torch.zeros_like(p) for p in self.params]
        self.exp_avg_sq = [torch.zeros_like(p) for p in self.params]
        # FiT/ema.py:243-248: the EMA copy starts as a deep copy of the parameters
        self.ema_params = tuple(p.detach().clone() for p in self.params) if ema_decay is not None else ()
        self._table = None
        self._table_key = None
        # capturable: the step number and the hyper-parameters live in device memory and the kernel derives the bias corrections itself, so
        # a CUDA graph that contains step() can be replayed for every step (fit_b200.train.CapturedTrainStep); like torch.optim's capturable
        self.capturable = False
        self._step_dev = None
        self._hyper_dev = None
        self._engine = None            # FiTTrainEngine whose kernel-layout weight buffers this optimizer keeps current
        self._shadows = {}             # parameter index -> (shadow tensor, row_len, shadow_ld, pack_off)

    def attach_engine(self, engine):
        """Let the step kernel write the bf16 / packed copies of the updated parameters into ``engine``'s weight buffers
        (``FiTTrainEngine.weights().shadow_specs``), so the training forward of the next step starts without re-casting anything."""
        W = engine.weights()
        index = {id(p): i for i, p in enumerate(self.params)}
        self._shadows = {}
        for p, shadow, row_len, shadow_ld, pack_off in W.shadow_specs(engine.model):
            if id(p) in index:
                self._shadows[index[id(p)]] = (shadow, row_len, shadow_ld, pack_off)
        self._engine = engine
        self._table = None
        return self

    # ------------------------------------------------------------------ tables
    def _build_table(self, grads: List[torch.Tensor]):
        key = tuple(g.data_ptr() for g in grads) + tuple(p.data_ptr() for p in self.params) + tuple(
            (i, sh[0].data_ptr()) for i, sh in sorted(self._shadows.items()))
        if self._table is not None and key == self._table_key:
            return
        dev = self.params[0].device
        arr = (_AdamTensor * len(self.params))()
        chunks = []
        for i, (p, g) in enumerate(zip(self.params, grads)):
            arr[i].p, arr[i].g, arr[i].m, arr[i].v = p.data_ptr(), g.data_ptr(), self.exp_avg[i].data_ptr(), self.exp_avg_sq[i].data_ptr()
            arr[i].ema = self.ema_params[i].data_ptr() if self.ema_params else 0
            arr[i].numel = p.numel()
            arr[i].shadow, arr[i].row_len, arr[i].shadow_ld, arr[i].pack_off = 0, 1, 1, -1
            if i in self._shadows:
                sh, row_len, shadow_ld, pack_off = self._shadows[i]
                arr[i].shadow, arr[i].shadow_dtype = sh.data_ptr(), L.dtype_code(sh)
                arr[i].row_len, arr[i].shadow_ld, arr[i].pack_off = row_len, shadow_ld, pack_off
            if any(int(x or 0) % 16 for x in (arr[i].p, arr[i].g, arr[i].m, arr[i].v, arr[i].ema)):
                raise RuntimeError("fit_b200.optim: parameter / gradient / state tensors must be 16-byte aligned")
            for first in range(0, p.numel(), ADAM_CHUNK):
                chunks += [i, first]
        # pinned staging + asynchronous copies: legal inside a CUDA-graph capture (CapturedTrainStep); the staging tensors stay alive with the table
        raw = torch.frombuffer(bytearray(bytes(arr)), dtype=torch.uint8).clone().pin_memory()
        chk = torch.tensor(chunks, dtype=torch.int32).pin_memory()
        raw_d, chk_d = torch.empty(raw.shape, dtype=torch.uint8, device=dev), torch.empty(chk.shape, dtype=torch.int32, device=dev)
        raw_d.copy_(raw, non_blocking=True)
        chk_d.copy_(chk, non_blocking=True)
        self._table = (raw_d, chk_d, len(chunks) // 2, raw, chk)
        self._table_key = key

    # ------------------------------------------------------------------ torch.optim-like API
    def zero_grad(self, set_to_none: bool = False):
        for p in self.params:
            if p.grad is not None:
                if set_to_none:
                    p.grad = None
                else:
                    p.grad.zero_()

    @torch.no_grad()
    def step(self):
        grads = []
        for p in self.params:
            if p.grad is None:
                raise RuntimeError("fit_b200.optim: every parameter needs a gradient (the FiT step produces all of them)")
            g = p.grad
            if g.dtype != torch.float32 or not g.is_contiguous():
                raise RuntimeError("fit_b200.optim: gradients must be contiguous fp32")
            grads.append(g)
        self._build_table(grads)
        if self.capturable:
            return self._step_capturable()
        self.step_count += 1
        b1, b2 = self.betas
        bc1 = 1.0 - b1 ** self.step_count
        bc2_sqrt = math.sqrt(1.0 - b2 ** self.step_count)
        # FiT/ema.py:252-257: the EMA moves when (step index before increment) % every_n_steps == 0
        do_ema = self.ema_decay is not None and (self.step_count - 1) % self.every_n_steps == 0
        tab, chunks, n_chunks = self._table[:3]
        L.check(L.lib().fitb200_adamw_ema_step(L.ptr(tab), L.ptr(chunks), n_chunks, self.lr, b1, b2, self.eps, self.weight_decay, bc1, bc2_sqrt,
                                               self.ema_decay if do_ema else -1.0, L.stream_ptr()), "adamw_ema_step")
        self.bump_versions()
        if self._engine is not None:
            self._engine.adopt_parameter_versions()

    def sync_device_state(self):
        """(Re)write the device-side step number and hyper-parameters from the host-side values (after construction, load_state_dict, or an
        lr change).  Not capturable itself: call it outside the graph."""
        dev = self.params[0].device
        if self._step_dev is None:
            self._step_dev = torch.zeros(1, dtype=torch.int64, device=dev)
            self._hyper_dev = torch.zeros(6, dtype=torch.float64, device=dev)
        self._step_dev.fill_(self.step_count)
        self._hyper_dev.copy_(torch.tensor([self.lr, self.betas[0], self.betas[1], self.eps, self.weight_decay,
                                            -1.0 if self.ema_decay is None else self.ema_decay], dtype=torch.float64))

    def _step_capturable(self):
        if self._step_dev is None:
            if torch.cuda.is_current_stream_capturing():
                raise RuntimeError("fit_b200.optim: run one eager step (or sync_device_state()) before capturing a capturable optimizer")
            self.sync_device_state()
        from . import ops
        ops.add_i64(self._step_dev, 1)
        tab, chunks, n_chunks = self._table[:3]
        L.check(L.lib().fitb200_adamw_ema_step_dev(L.ptr(tab), L.ptr(chunks), n_chunks, L.ptr(self._hyper_dev), L.ptr(self._step_dev),
                                                   self.every_n_steps, L.stream_ptr()), "adamw_ema_step_dev")
        if not torch.cuda.is_current_stream_capturing():   # a capture records the launch; the replays are counted by note_external_step
            self.step_count += 1
        self.bump_versions()
        if self._engine is not None:
            self._engine.adopt_parameter_versions()

    def note_external_step(self):
        """A captured CUDA graph replayed this optimizer's step kernel (fit_b200.train.CapturedTrainStep): advance the host-side step count
        and the parameter version counters as ``step`` would have."""
        self.step_count += 1
        self.bump_versions()
        if self._engine is not None:
            self._engine.adopt_parameter_versions()

    def bump_versions(self):
        """The kernel wrote the parameters behind autograd's back: bump their version counters so that the weight caches keyed by
        (data_ptr, _version) (fit_b200.engine._Weights) repack."""
        torch._C._autograd._unsafe_set_version_counter(tuple(self.params), tuple(p._version + 1 for p in self.params))

    # ------------------------------------------------------------------ EMA helpers (FiT/ema.py:280-313)
    def swap_ema_weights(self):
        for p, e in zip(self.params, self.ema_params):
            tmp = p.data.clone()
            p.data.copy_(e)
            e.copy_(tmp)
        self.bump_versions()

    @contextlib.contextmanager
    def swap_ema_weights_ctx(self, enabled: bool = True):
        if enabled:
            self.swap_ema_weights()
        try:
            yield
        finally:
            if enabled:
                self.swap_ema_weights()

    def state_dict(self):
        return {"state": {i: {"step": torch.tensor(float(self.step_count)), "exp_avg": m, "exp_avg_sq": v}
                          for i, (m, v) in enumerate(zip(self.exp_avg, self.exp_avg_sq))},
                "param_groups": [{"lr": self.lr, "betas": self.betas, "eps": self.eps, "weight_decay": self.weight_decay,
                                  "params": list(range(len(self.params)))}],
                "ema": self.ema_params, "current_step": self.step_count, "decay": self.ema_decay, "every_n_steps": self.every_n_steps}

    def load_state_dict(self, sd):
        """Accepts this class's own ``state_dict()``, the reference EMA optimizer's layout ``{"opt": AdamW state, "ema", "current_step",
        "decay", "every_n_steps"}`` (FiT/ema.py:322-347 -- what ``checkpoint.to_lightning_checkpoint`` writes into ``optimizer_states[0]``)
        and a bare ``torch.optim.AdamW`` state dict.  Restores the moments, the step count (``current_step``, else the per-parameter
        ``step`` AdamW keeps) and the EMA copy, so a resumed run continues with the right bias correction."""
        opt = sd["opt"] if "opt" in sd else sd
        steps = []
        for i, st in opt["state"].items():
            self.exp_avg[int(i)].copy_(st["exp_avg"]); self.exp_avg_sq[int(i)].copy_(st["exp_avg_sq"])
            if "step" in st:
                steps.append(int(float(st["step"])))
        if sd.get("current_step") is not None:
            self.step_count = int(sd["current_step"])
        elif steps:
            self.step_count = max(steps)
        else:
            self.step_count = 0
        groups = opt.get("param_groups") or []
        if groups:
            g0 = groups[0]
            self.lr, self.betas = g0.get("lr", self.lr), tuple(g0.get("betas", self.betas))
            self.eps, self.weight_decay = g0.get("eps", self.eps), g0.get("weight_decay", self.weight_decay)
        if sd.get("decay") is not None:
            self.ema_decay = sd["decay"]
        if sd.get("every_n_steps") is not None:
            self.every_n_steps = sd["every_n_steps"]
        if sd.get("ema") is not None and self.ema_params:
            if len(sd["ema"]) != len(self.ema_params):
                raise ValueError(f"EMA state holds {len(sd['ema'])} tensors, the optimizer has {len(self.ema_params)} parameters")
            for e, src in zip(self.ema_params, sd["ema"]):
                e.copy_(src)
        if self._step_dev is not None:
            self.sync_device_state()
