"""Lightning .ckpt compatibility (fit_b200.checkpoint): key prefixes of the compiled / plain FiTModule,
EMA weights from optimizer_states[0]['ema'], hyper_parameters['args'].model, and the round trip through torch.save."""
from argparse import Namespace

import pytest
import torch

from fit_b200 import checkpoint as ck
from fit_b200.models.fit import FiT_models


def _model(seed):
    torch.manual_seed(seed)
    m = FiT_models["FiT-S/8"]()
    with torch.no_grad():
        for p in m.parameters():
            p.copy_(torch.randn_like(p) * 0.02)
    return m


@pytest.mark.parametrize("prefix", ["model._orig_mod.", "model."])
def test_load_reference_style_checkpoint(prefix, tmp_path):
    src, dst = _model(0), _model(1)
    ema = tuple(p.detach() + 1.0 for p in src.parameters())
    blob = {"state_dict": {prefix + k: v.clone() for k, v in src.state_dict().items()},
            "optimizer_states": [{"opt": {}, "ema": ema, "current_step": 3, "decay": 0.9999, "every_n_steps": 1}],
            "hyper_parameters": {"args": Namespace(model="FiT-S/8", global_seed=0)}}
    path = tmp_path / "epoch=0.ckpt"
    torch.save(blob, path)
    assert ck.model_name_from_checkpoint(str(path)) == "FiT-S/8"
    ck.load_lightning_checkpoint(str(path), dst)
    for (n, a), (_, b) in zip(src.state_dict().items(), dst.state_dict().items()):
        assert torch.equal(a, b), n
    ck.load_lightning_checkpoint(str(path), dst, use_ema=True)
    for a, b in zip(src.parameters(), dst.parameters()):
        assert torch.equal(a.detach() + 1.0, b.detach())


def test_errors():
    m = _model(0)
    sd = {"model." + k: v for k, v in m.state_dict().items()}
    sd.pop("model.final_layer.linear.weight")
    with pytest.raises(KeyError):
        ck.load_lightning_checkpoint({"state_dict": sd}, m)
    full = {"state_dict": {"model." + k: v for k, v in m.state_dict().items()}}
    with pytest.raises(KeyError):
        ck.load_lightning_checkpoint(full, m, use_ema=True)            # no optimizer_states
    full["optimizer_states"] = [{"ema": tuple(p.detach() for p in list(m.parameters())[:-1])}]
    with pytest.raises(ValueError):
        ck.load_lightning_checkpoint(full, m, use_ema=True)            # wrong tensor count


def test_round_trip_through_reference_format(tmp_path):
    m = _model(2)

    class _Opt:   # the slice of FusedAdamWEMA.state_dict() the writer reads (the optimizer itself needs a GPU)
        def state_dict(self):
            ps = list(m.parameters())
            return {"state": {i: {"step": torch.tensor(5.0), "exp_avg": torch.zeros_like(p), "exp_avg_sq": torch.ones_like(p)} for i, p in enumerate(ps)},
                    "param_groups": [{"lr": 1e-4, "betas": (0.9, 0.999), "eps": 1e-8, "weight_decay": 0.0, "params": list(range(len(ps)))}],
                    "ema": tuple(p.detach() * 2 for p in ps), "current_step": 5, "decay": 0.9999, "every_n_steps": 1}

    blob = ck.to_lightning_checkpoint(m, "FiT-S/8", optimizer=_Opt(), global_step=5)
    assert all(k.startswith("model._orig_mod.") for k in blob["state_dict"])
    path = tmp_path / "x.ckpt"
    torch.save(blob, path)
    back = _model(3)
    ck.load_lightning_checkpoint(str(path), back, use_ema=True)
    for a, b in zip(m.parameters(), back.parameters()):
        assert torch.equal(a.detach() * 2, b.detach())
    # torch.optim.AdamW (the reference's inner optimizer) accepts the written optimizer state
    opt = torch.optim.AdamW(back.parameters(), lr=1e-4, weight_decay=0)
    opt.load_state_dict(torch.load(path, weights_only=False)["optimizer_states"][0]["opt"])
    assert float(opt.state[next(iter(back.parameters()))]["step"]) == 5.0
