"""Lightning checkpoint compatibility: the on-disk format on the other side of the drop-in boundary.

The reference trains ``FiTModule`` (FiT/train.py:31-40): ``self.model = torch.compile(FiT_models[args.model]())`` and
``save_hyperparameters()``, so a ``.ckpt`` written by its ``ModelCheckpoint`` (train.py:272-277) is a pickled dict with
  state_dict            {"model._orig_mod.<param>": tensor}   ("model.<param>" when the module was not compiled)
  optimizer_states[0]   the EMA optimizer's state (FiT/ema.py:322-337): {"opt": AdamW state, "ema": tuple of tensors in
                        ``parameters()`` order, "current_step", "decay", "every_n_steps"}
  hyper_parameters      {"args": Namespace(model="FiT-XL/2", ...)}
``sample_lightning.py:15`` reads it back with ``FiTModule.load_from_checkpoint``.  These helpers move weights between that format and
``fit_b200.models.fit.FiT`` (same parameter names and shapes), including the EMA weights the reference samples from when the EMA
callback swaps them in (ema.py:280-313).  Pure host logic: no kernels, works on CPU tensors."""
from __future__ import annotations

from argparse import Namespace
from typing import Any, Dict, Optional, Tuple, Union

import torch

_PREFIXES = ("model._orig_mod.", "model.module.", "model.", "_orig_mod.", "")


def _read(ckpt: Union[str, Dict[str, Any]]) -> Dict[str, Any]:
    if isinstance(ckpt, dict):
        return ckpt
    # Lightning pickles an argparse Namespace next to the tensors: weights_only=True would refuse it
    return torch.load(ckpt, map_location="cpu", weights_only=False)


def model_name_from_checkpoint(ckpt: Union[str, Dict[str, Any]]) -> Optional[str]:
    """``args.model`` as saved by ``save_hyperparameters()`` (FiT/train.py:40, 240-ish ``--model``)."""
    hp = _read(ckpt).get("hyper_parameters") or {}
    args = hp.get("args", hp)
    if isinstance(args, Namespace):
        return getattr(args, "model", None)
    if isinstance(args, dict):
        return args.get("model")
    return None


def extract_state_dict(ckpt: Union[str, Dict[str, Any]], model: torch.nn.Module, use_ema: bool = False) -> Dict[str, torch.Tensor]:
    """The FiT parameters of a Lightning checkpoint under this module's names.  ``use_ema``: the EMA weights
    (``optimizer_states[0]["ema"]``, one tensor per parameter in ``parameters()`` order, FiT/ema.py:243-248,328-335)."""
    ck = _read(ckpt)
    sd = ck["state_dict"] if "state_dict" in ck else ck
    want = list(model.state_dict().keys())
    out: Dict[str, torch.Tensor] = {}
    for pre in _PREFIXES:
        if all(pre + k in sd for k in want):
            out = {k: sd[pre + k] for k in want}
            break
    else:
        missing = [k for k in want if not any(p + k in sd for p in _PREFIXES)]
        raise KeyError(f"checkpoint does not hold the FiT parameters (first missing: {missing[:3]}, {len(missing)} of {len(want)})")
    if use_ema:
        states = ck.get("optimizer_states") or []
        if not states or "ema" not in states[0]:
            raise KeyError("checkpoint has no EMA weights (optimizer_states[0]['ema']): it was not trained with the EMA callback")
        ema = list(states[0]["ema"])
        names = [n for n, _ in model.named_parameters()]
        if len(ema) != len(names):
            raise ValueError(f"EMA holds {len(ema)} tensors, the model has {len(names)} parameters")
        for n, e in zip(names, ema):
            if tuple(e.shape) != tuple(out[n].shape):
                raise ValueError(f"EMA tensor for {n} has shape {tuple(e.shape)}, expected {tuple(out[n].shape)}")
            out[n] = e
    return out


def load_lightning_checkpoint(ckpt: Union[str, Dict[str, Any]], model: torch.nn.Module, use_ema: bool = False) -> torch.nn.Module:
    """``model.load_state_dict`` from a reference ``.ckpt`` (strict: every parameter must be present with the reference's shape)."""
    model.load_state_dict(extract_state_dict(ckpt, model, use_ema=use_ema), strict=True)
    return model


def from_checkpoint(ckpt: Union[str, Dict[str, Any]], use_ema: bool = False, device=None, **model_kwargs) -> Tuple[torch.nn.Module, str]:
    """``FiTModule.load_from_checkpoint`` for the hot path only: build ``FiT_models[args.model]`` and load its weights."""
    from .models.fit import FiT_models
    ck = _read(ckpt)
    name = model_name_from_checkpoint(ck)
    if name not in FiT_models:
        raise KeyError(f"checkpoint names model {name!r}; known: {sorted(FiT_models)}")
    model = FiT_models[name](**model_kwargs)
    load_lightning_checkpoint(ck, model, use_ema=use_ema)
    if device is not None:
        model = model.to(device)
    return model.eval(), name


def to_lightning_checkpoint(model: torch.nn.Module, model_name: str, optimizer=None, compiled: bool = True, global_step: int = 0,
                            epoch: int = 0) -> Dict[str, Any]:
    """The dict the reference's ``FiTModule.load_from_checkpoint`` / ``EMAOptimizer.load_state_dict`` expect (write it with
    ``torch.save``).  ``optimizer``: a ``fit_b200.optim.FusedAdamWEMA`` (its AdamW moments and EMA copy go into
    ``optimizer_states[0]``)."""
    pre = "model._orig_mod." if compiled else "model."
    ck: Dict[str, Any] = {
        "epoch": epoch, "global_step": global_step, "pytorch-lightning_version": "2.3.0",
        "state_dict": {pre + k: v.detach().cpu().clone() for k, v in model.state_dict().items()},
        "hyper_parameters": {"args": Namespace(model=model_name)},
    }
    if optimizer is not None:
        sd = optimizer.state_dict()
        ck["optimizer_states"] = [{
            "opt": {"state": {i: {"step": st["step"], "exp_avg": st["exp_avg"].detach().cpu().clone(), "exp_avg_sq": st["exp_avg_sq"].detach().cpu().clone()}
                              for i, st in sd["state"].items()},
                    "param_groups": [dict(g, amsgrad=False, maximize=False, foreach=None, capturable=False, differentiable=False, fused=None)
                                     for g in sd["param_groups"]]},
            "ema": tuple(e.detach().cpu().clone() for e in sd["ema"]), "current_step": sd["current_step"], "decay": sd["decay"],
            "every_n_steps": sd["every_n_steps"],
        }]
    return ck
