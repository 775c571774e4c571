"""Build libpsla_frontend.so in-tree with nvcc for sm_100a (no JIT cache, no torch extension)."""
from __future__ import annotations

import os
import shutil
import subprocess

CSRC = os.path.join(os.path.dirname(os.path.abspath(__file__)), "csrc")
# PSLA_FRONTEND_LIB: load another build of the same sources (kernel experiments, scripts/build_variants.py)
BUILT_LIB = os.path.join(CSRC, "libpsla_frontend.so")
LIB = os.environ.get("PSLA_FRONTEND_LIB") or BUILT_LIB
VARIANT_DIR = os.path.join(CSRC, "variants")
SOURCES = ("psla_frontend.cu", "tables.cc")
HEADERS = ("fft_core.cuh", "fbank_stream.cuh", "stream_kernel.cuh", "noise_roll.cuh", "tables.h", os.path.join("..", "..", "include", "psla_frontend.h"))
NVCC_FLAGS = ("-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "-shared")


def _nvcc() -> str:
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: cannot build libpsla_frontend.so")


def is_stale() -> bool:
    if not os.path.exists(BUILT_LIB):
        return True
    t = os.path.getmtime(BUILT_LIB)
    return any(os.path.getmtime(os.path.join(CSRC, f)) > t for f in SOURCES + HEADERS)


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile the CUDA library if missing or older than its sources; returns its path."""
    if not force and not is_stale():
        return BUILT_LIB
    cmd = [_nvcc(), *NVCC_FLAGS, "-o", BUILT_LIB + ".tmp", *SOURCES]
    if verbose:
        cmd.insert(1, "-Xptxas")
        cmd.insert(2, "-v")
    res = subprocess.run(cmd, cwd=CSRC, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + res.stdout + res.stderr)
    os.replace(BUILT_LIB + ".tmp", BUILT_LIB)
    if verbose:
        print(res.stderr)
    return BUILT_LIB


def build_variant(name: str, defines=(), slim: bool = True, verbose: bool = False, extra_flags=()) -> str:
    """Experimental build csrc/variants/libpsla_<name>.so with extra -D flags.  slim: only the kernels the bench batch uses
    (-DPSLA_SLIM: default clip-mean mode, 128 mel bins, fp32 samples, [B, T, F] output) -- seconds instead of minutes."""
    os.makedirs(VARIANT_DIR, exist_ok=True)
    out = os.path.join(VARIANT_DIR, f"libpsla_{name}.so")
    cmd = [_nvcc(), *NVCC_FLAGS, *(["-DPSLA_SLIM"] if slim else []), *[f"-D{d}" for d in defines], *extra_flags, "-Xptxas", "-v", "-o", out, *SOURCES]
    res = subprocess.run(cmd, cwd=CSRC, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"nvcc failed for variant {name}:\n" + res.stdout + res.stderr)
    if verbose:
        print(res.stderr)
    with open(out + ".ptxas.txt", "w") as f:
        f.write(res.stderr)
    return out


if __name__ == "__main__":
    print(build(force=True, verbose=True))
