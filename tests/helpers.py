"""Shared helpers of the parity tests (the oracle is the checker; the product path is fit_b200)."""
import os

import numpy as np
import torch

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

TINY64 = dict(patch_size=2, in_channels=4, hidden_size=128, depth=2, num_heads=2, mlp_ratio=4.0,
              class_dropout_prob=0.1, num_classes=10, learn_sigma=False)
TINY72 = dict(TINY64, hidden_size=144)
TINY_SHAPES = [(16, 16), (10, 20), (8, 32), (12, 12)]


def rel_l2(a, b):
    a, b = torch.as_tensor(a).double().cpu(), torch.as_tensor(b).double().cpu()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def golden(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"))


def build_model(cfg, sd, precision, device="cuda", train_mode=False):
    """fit_b200 model with the oracle's state_dict (the reference's parameter names)."""
    from fit_b200.models.fit import FiT
    m = FiT(**cfg, precision=precision)
    missing, unexpected = m.load_state_dict(sd, strict=True)
    assert not missing and not unexpected
    m = m.to(device)
    return m.train() if train_mode else m.eval()


def dev(*ts, device="cuda"):
    out = [t.to(device) if isinstance(t, torch.Tensor) else t for t in ts]
    return out if len(out) > 1 else out[0]


def record(key, value):
    """Append a measured parity figure (worst-case error of a test) to gpurun_out/parity_measured.txt so the GPU run leaves the actual
    numbers behind, not only pass/fail against the bound (copied to profiles/ per round)."""
    out = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out")
    try:
        os.makedirs(out, exist_ok=True)
        with open(os.path.join(out, "parity_measured.txt"), "a") as fh:
            fh.write(f"{key}\t{value}\n")
    except OSError:
        pass
