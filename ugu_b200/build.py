"""Builds the native pieces in-tree (no JIT cache): the sm_100a CUDA library behind the C-ABI.

    python -m ugu_b200.build            # build if stale
    python -m ugu_b200.build --force

nvcc cross-compiles without a GPU.  `--fmad=false` is REQUIRED: the kernels reproduce the
reference's float results bit for bit and the reference object code contains no FMA
(SURVEY.md §8a B2).  Never add -use_fast_math.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent
REPO = ROOT.parent
CSRC = ROOT / "csrc"
LIBDIR = ROOT / "lib"
LIB = LIBDIR / "libugu_b200_raster.so"

SOURCES = [CSRC / "raster_api.cu", CSRC / "multiview_api.cu", CSRC / "static_scratch_api.cu",
           CSRC / "camera_host.cc"]
DEPS = SOURCES + sorted(CSRC.glob("*.cuh")) + [REPO / "include" / "ugu_b200_raster.h"]  # every kernel header

NVCC_FLAGS = [
    "-std=c++17", "-O3", "--fmad=false", "-lineinfo",
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-Xcompiler", "-fPIC,-O2,-Wall", "-shared", "-cudart", "static",
    "-Xptxas", "-v",
]


def find_nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found; the ugu_b200 CUDA library cannot be built")


def is_stale() -> bool:
    if not LIB.exists():
        return True
    t = LIB.stat().st_mtime
    return any(p.stat().st_mtime > t for p in DEPS)


def build_native(force: bool = False, verbose: bool = False) -> Path:
    if not force and not is_stale():
        return LIB
    LIBDIR.mkdir(parents=True, exist_ok=True)
    # Use the system host compiler: the image's default CXX wrapper is fine for nvcc's needs too.
    cmd = [find_nvcc(), *NVCC_FLAGS, f"-I{REPO / 'include'}", f"-I{CSRC}",
           "-o", str(LIB), *map(str, SOURCES)]
    proc = subprocess.run(cmd, capture_output=True, text=True)
    log = proc.stdout + proc.stderr
    (LIBDIR / "build.log").write_text(" ".join(cmd) + "\n" + log)
    if proc.returncode != 0:
        raise RuntimeError(f"nvcc failed ({proc.returncode}):\n{log}")
    if verbose:
        print(log)
    return LIB


if __name__ == "__main__":
    build_native(force="--force" in sys.argv, verbose=True)
    print(f"built {LIB}")
