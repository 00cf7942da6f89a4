"""ctypes binding of libfit_b200.so (the C ABI declared in include/fit_b200.h).

There is NO fallback: if the shared library is missing or an entry point fails, a RuntimeError is
raised.  torch is used only for device memory (``tensor.data_ptr()``) and the current stream.
"""
from __future__ import annotations

import ctypes as C
import os
import re

import torch

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("FITB200_LIB") or os.path.join(_PKG, "libfit_b200.so")  # FITB200_LIB: another build of the same ABI (A/B timing)
HEADER_PATH = os.path.join(os.path.dirname(_PKG), "include", "fit_b200.h")

F32, BF16 = 0, 1
EPI_BIAS, EPI_BIAS_SILU, EPI_BIAS_GATE_RESID, EPI_SWIGLU, EPI_BIAS_ROPE = 0, 1, 2, 3, 4

_lib = None


class GemmArgs(C.Structure):
    """Mirror of fitb200_gemm_args (include/fit_b200.h)."""
    _fields_ = [
        ("M", C.c_int32), ("N", C.c_int32), ("K", C.c_int32),
        ("A", C.c_void_p), ("lda", C.c_int64),
        ("B", C.c_void_p), ("ldb", C.c_int64),
        ("B2", C.c_void_p),
        ("bias", C.c_void_p), ("bias2", C.c_void_p),
        ("out", C.c_void_p), ("ldo", C.c_int64), ("out_dtype", C.c_int32),
        ("epilogue", C.c_int32),
        ("gate", C.c_void_p), ("ld_gate", C.c_int64), ("rows_per_sample", C.c_int32),
        ("resid", C.c_void_p), ("ld_resid", C.c_int64), ("resid_dtype", C.c_int32),
        ("rope_freqs", C.c_void_p), ("rope_rows", C.c_int32), ("rope_tokens", C.c_int32), ("rope_head_dim", C.c_int32),
        ("rope_cols", C.c_int32),
        ("a_mn", C.c_int32), ("b_mn", C.c_int32),
        ("bias_row0", C.c_int32),
        ("ln_g", C.c_void_p), ("ld_ln_g", C.c_int64),
        ("ln_c", C.c_void_p), ("ln_samples", C.c_int32),
        ("ln_parts", C.c_void_p), ("ln_nparts", C.c_int32), ("ln_eps", C.c_float),
        ("stats_out", C.c_void_p),
    ]


def declared_symbols():
    """Every entry point include/fit_b200.h declares (used by the CPU-side export test)."""
    with open(HEADER_PATH) as fh:
        src = fh.read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(fitb200_[a-z0-9_]+)\s*\(", src)))


_SCALARS = {"int": C.c_int, "int32_t": C.c_int32, "int64_t": C.c_int64, "uint64_t": C.c_uint64, "float": C.c_float, "double": C.c_double}


def declared_prototypes():
    """{name: [ctypes argtypes]} parsed from the header, so that every call is type-checked and
    64-bit / float scalars are passed with the right width."""
    with open(HEADER_PATH) as fh:
        src = fh.read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    protos = {}
    for m in re.finditer(r"\b(?:int|const char\*)\s+(fitb200_[a-z0-9_]+)\s*\(([^)]*)\)\s*;", src):
        name, args = m.group(1), m.group(2).strip()
        types = []
        if args and args != "void":
            for a in args.split(","):
                a = a.strip()
                if "*" in a or a.startswith("fitb200_stream_t"):
                    types.append(C.c_void_p)
                else:
                    base = a.replace("const ", "").split()[0]
                    types.append(_SCALARS[base])
        protos[name] = types
    return protos


def lib():
    """The loaded library; raises loudly when it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -m fit_b200.build` "
                "(fit_b200 has no CPU or PyTorch fallback)")
        _lib = C.CDLL(LIB_PATH)
        _lib.fitb200_last_error.restype = C.c_char_p
        protos = declared_prototypes()
        for name in declared_symbols():
            fn = getattr(_lib, name)  # AttributeError if the header and the library disagree
            if name != "fitb200_last_error":
                fn.restype = C.c_int
            fn.argtypes = protos[name]
    return _lib


def check(rc: int, what: str = ""):
    if rc != 0:
        msg = lib().fitb200_last_error().decode(errors="replace")
        raise RuntimeError(f"fit_b200 {what} failed (rc={rc}): {msg}")


def stream_ptr() -> C.c_void_p:
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def ptr(t) -> C.c_void_p:
    return C.c_void_p(0 if t is None else t.data_ptr())


def dtype_code(t: torch.Tensor) -> int:
    if t.dtype == torch.float32:
        return F32
    if t.dtype == torch.bfloat16:
        return BF16
    raise TypeError(f"fit_b200: unsupported dtype {t.dtype}")


def require_cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise RuntimeError("fit_b200 kernels need CUDA tensors (there is no CPU fallback)")
