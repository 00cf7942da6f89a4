"""fit_b200 -- B200-native (sm_100a) implementation of the FiT denoising hot path.

    from fit_b200.models.fit import FiT_models          # drop-in for FiT/models/fit.py
    from fit_b200.diffusion import create_diffusion     # drop-in for FiT/diffusion

Everything computes through libfit_b200.so (hand-written CUDA behind the C ABI of include/fit_b200.h);
there is no CPU or PyTorch fallback.
"""
__version__ = "0.1.0"
