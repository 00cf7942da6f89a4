"""In-tree build of libfit_b200.so (sm_100a only) with plain nvcc.

``python -m fit_b200.build`` or ``fit_b200.build.build()``.  nvcc cross-compiles without a GPU.
Objects go to ``fit_b200/csrc/_build/`` and the library to ``fit_b200/libfit_b200.so`` (both
git-ignored; the .so travels to the GPU box with the gpurun snapshot).
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
CSRC = os.path.join(PKG, "csrc")
OBJDIR = os.path.join(CSRC, "_build")
LIB = os.path.join(PKG, "libfit_b200.so")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC",
    "--expt-relaxed-constexpr", "-DFITB200_BUILD",
]
if os.environ.get("FITB200_NVCC_EXTRA"):       # A/B builds of one tuning macro (python -m fit_b200.build with FITB200_LIB_OUT=<other .so>)
    NVCC_FLAGS += os.environ["FITB200_NVCC_EXTRA"].split()
    LIB = os.environ.get("FITB200_LIB_OUT", LIB)
    OBJDIR = OBJDIR + "_alt"
if os.environ.get("FITB200_TUNING") == "1":   # measurement-only switches (e.g. FITB200_GEMM_DBG: GEMM mainloop ceiling); never in the product build
    NVCC_FLAGS.append("-DFITB200_TUNING")


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", shutil.which("nvcc")):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found (set NVCC=/path/to/nvcc)")


def sources():
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cu"))


def _digest(path: str) -> str:
    h = hashlib.sha256()
    h.update(" ".join(NVCC_FLAGS).encode())
    deps = [path] + sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h")))
    deps.append(os.path.join(ROOT, "include", "fit_b200.h"))
    for d in deps:
        with open(d, "rb") as fh:
            h.update(fh.read())
    return h.hexdigest()


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile every .cu under csrc/ and link libfit_b200.so.  Returns the library path."""
    os.makedirs(OBJDIR, exist_ok=True)
    nvcc = _nvcc()
    jobs = []
    for src in sources():
        obj = os.path.join(OBJDIR, os.path.basename(src)[:-3] + ".o")
        stamp = obj + ".sha"
        dig = _digest(src)
        fresh = (not force and os.path.exists(obj) and os.path.exists(stamp) and open(stamp).read() == dig)
        jobs.append((src, obj, stamp, dig, fresh))

    def compile_one(job):
        src, obj, stamp, dig, fresh = job
        if fresh:
            return None
        cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", src, "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
        with open(stamp, "w") as fh:
            fh.write(dig)
        return r.stderr if verbose else None

    with ThreadPoolExecutor(max_workers=min(8, len(jobs))) as ex:
        logs = list(ex.map(compile_one, jobs))
    if verbose:
        for lg in logs:
            if lg:
                print(lg)
    objs = [j[1] for j in jobs]
    need_link = force or not os.path.exists(LIB) or any(not j[4] for j in jobs) or any(
        os.path.getmtime(o) > os.path.getmtime(LIB) for o in objs)
    if need_link:
        cmd = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", LIB] + objs
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return LIB


if __name__ == "__main__":
    path = build(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(path)
