"""fitb200_ln_modulate_fwd_static: the LayerNorm + adaLN-modulate pass whose scratch (row sums) is a file-scope `static __device__` array of
the library instead of a caller / allocator buffer (no counterpart in the reference).

Called through the C ABI (ctypes) and checked against the existing entry point fitb200_ln_modulate_fwd on the same inputs, on two calls in a
row (the second with fewer rows, so the shared array still holds the first call's contents), within the tolerances of tests/test_gpu_kernels.py
(fp32 1e-5, bf16 4e-3 relative L2).  Collected last (file name) so that a failure here hides no other test under `pytest -x`."""
import ctypes as C

import pytest
import torch

from helpers import rel_l2

pytestmark = pytest.mark.gpu

MAX_ROWS = 131072   # FITB200_STATIC_LN_MAX_ROWS (include/fit_b200.h)


def _call(fn, x, mod, D, rps, out):
    from fit_b200 import _lib as L
    M = x.shape[0]
    shift, scale = mod, mod[:, D:]
    args = [C.c_void_p(x.data_ptr()), L.dtype_code(x), M, D, rps, C.c_void_p(shift.data_ptr()), C.c_void_p(scale.data_ptr()), mod.stride(0),
            C.c_void_p(out.data_ptr()), L.dtype_code(out)]
    if fn.__name__ == "fitb200_ln_modulate_fwd":
        args += [None, None]   # mean, rstd outputs not requested
    return fn(*args, L.stream_ptr())


@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16])
def test_static_scratch_ln_modulate_matches_existing_entry_point(dt):
    from fit_b200 import _lib as L
    lib = L.lib()
    g = torch.Generator().manual_seed(11)
    tol = 1e-5 if dt == torch.float32 else 4e-3
    # two calls in a row on different sizes: the second reuses the array the first one filled
    for M, D, rps in [(4096, 1152, 256), (777, 768, 111)]:
        S = (M + rps - 1) // rps
        x = (torch.randn(M, D, generator=g) * 1.3 + 0.2).to(dt).cuda()
        mod = (torch.randn(S, 2 * D, generator=g) * 0.4).cuda()
        want = torch.empty(M, D, dtype=dt, device="cuda")
        got = torch.full((M, D), float("nan"), dtype=dt, device="cuda")
        assert _call(lib.fitb200_ln_modulate_fwd, x, mod, D, rps, want) == 0, lib.fitb200_last_error()
        free_before = torch.cuda.mem_get_info()[0]
        assert _call(lib.fitb200_ln_modulate_fwd_static, x, mod, D, rps, got) == 0, lib.fitb200_last_error()
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info()[0] >= free_before          # the entry point took no device memory
        assert bool(torch.isfinite(got.float()).all())
        assert rel_l2(got.float(), want.float()) < tol, (M, D, rps)


def test_static_scratch_ln_modulate_rejects_oversize():
    from fit_b200 import _lib as L
    lib = L.lib()
    D = 64
    x = torch.zeros(8, D, device="cuda")
    mod = torch.zeros(1, 2 * D, device="cuda")
    out = torch.empty(8, D, device="cuda")
    rc = lib.fitb200_ln_modulate_fwd_static(C.c_void_p(x.data_ptr()), 0, MAX_ROWS + 1, D, MAX_ROWS + 1, C.c_void_p(mod.data_ptr()),
                                            C.c_void_p(mod[:, D:].data_ptr()), 2 * D, C.c_void_p(out.data_ptr()), 0, L.stream_ptr())
    assert rc == -1 and b"exceed the static scratch" in lib.fitb200_last_error()
