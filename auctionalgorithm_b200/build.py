"""Builds libb200lsap.so (the CUDA kernels + C-ABI) in-tree with nvcc for sm_100a.

    python -m auctionalgorithm_b200.build [--force]

nvcc cross-compiles without a GPU; the .so is git-ignored but travels to the GPU box.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

PKG = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG, "csrc")
LIB = os.path.join(PKG, "libb200lsap.so")
SOURCES = ["capi.cu"]
HEADERS = sorted(f for f in os.listdir(CSRC) if f.endswith(".cuh"))  # every header capi.cu may include
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17", "--shared",
              "-Xcompiler", "-fPIC", "-Xcompiler", "-O3"]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: libb200lsap.so cannot be built (there is no CPU fallback)")


def needs_build() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, f) for f in SOURCES + HEADERS]
    deps += [os.path.join(PKG, "..", "include", f) for f in ("b200lsap.h", "b200lsap_gen.h")]
    return any(os.path.exists(d) and os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not needs_build():
        return LIB
    cmd = [_nvcc(), *NVCC_FLAGS, "-o", LIB] + [os.path.join(CSRC, s) for s in SOURCES]
    if os.environ.get("B200LSAP_TIMING"):
        cmd.insert(1, "-DB200LSAP_TIMING")
    if os.environ.get("B200LSAP_VAR"):  # developer experiments (dense_cache.cuh)
        cmd.insert(1, "-DB200LSAP_VAR=" + os.environ["B200LSAP_VAR"])
    for d in os.environ.get("B200LSAP_DEFINES", "").split():  # developer experiments: "-DB200LSAP_SOLO_MAX=3 ..."
        cmd.insert(1, d)
    if os.environ.get("B200LSAP_OUT"):
        cmd[cmd.index("-o") + 1] = os.environ["B200LSAP_OUT"]
    if verbose:
        cmd[1:1] = ["-Xptxas", "-v"]
        print(" ".join(cmd))
    r = subprocess.run(cmd, capture_output=True, text=True)
    if verbose or r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed building libb200lsap.so")
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
