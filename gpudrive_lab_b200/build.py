"""In-tree build of libgpudrive_b200.so (nvcc, sm_100a). The .so is git-ignored but travels to
the GPU box with the repo snapshot."""
from __future__ import annotations

import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libgpudrive_b200.so")
SOURCES = ["engine.cu", "scenario.cpp", "capi.cpp"]
HEADERS = ["engine.h", "gd_types.h", "scenario.h", "json_min.h", "glibc_trig.h", "device_math.cuh", "kernels_step.cuh",
           "kernels_obs.cuh", "kernels_obs_main.cuh", "kernels_reset.cuh", "kernels_bev.cuh", "kernels_policy.cuh",
           "kernels_env.cuh", "kernels_lidar_fix.cuh", os.path.join("..", "..", "include", "gpudrive_b200.h")]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    # bit-parity with the reference's CPU build: no FMA contraction on either side
    "--fmad=false", "-Xcompiler", "-fPIC,-ffp-contract=off,-Wall,-Wno-stringop-truncation,-pthread",
    "-shared",
]


def _nvcc():
    for c in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if c and os.path.exists(c):
            return c
    raise RuntimeError("nvcc not found")


def needs_build() -> bool:
    if not os.path.exists(OUT):
        return True
    t = os.path.getmtime(OUT)
    return any(os.path.getmtime(os.path.join(CSRC, f)) > t for f in SOURCES + HEADERS)


def build_engine(force: bool = False, verbose: bool = False) -> str:
    if not force and not needs_build():
        return OUT
    cmd = [_nvcc()] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", OUT] + SOURCES
    r = subprocess.run(cmd, cwd=CSRC, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + r.stdout + r.stderr)
    if verbose:
        print(r.stderr)
    return OUT


HEADLESS = os.path.join(HERE, "gd_headless")


def build_headless() -> str:
    """tools/gd_headless.cpp: the C++ counterpart of the reference's `headless` executable, linked against
    the C ABI only (plus cudart for two copies)."""
    src = os.path.join(HERE, "..", "tools", "gd_headless.cpp")
    cmd = [_nvcc(), "-O2", "-std=c++17", "-o", HEADLESS, src, "-L" + HERE, "-lgpudrive_b200",
           "-Xlinker", "-rpath," + HERE]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("gd_headless build failed:\n" + r.stdout + r.stderr)
    return HEADLESS
