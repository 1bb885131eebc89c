"""Builds the C++ adapter library — only possible where the reference headers exist (this
container; the GPU box receives the prebuilt, git-ignored binary).

    ugu_b200/cpp/build/librasterizer_cuda.so   ugu::RasterizerCuda, ugu::RaytracerCuda, ugu::VisibilityTesterCuda,
                                               ugu::MultiViewRasterizerCuda
                                               (links libugu_b200_raster.so)

Its differential harness against the reference Rasterizer is test infrastructure and lives in
tests/cpp/ (tests/cpp/build_check.py).
"""
from __future__ import annotations

import subprocess
from pathlib import Path

HERE = Path(__file__).resolve().parent
REPO = HERE.parents[1]
REF = Path("/root/reference")
OUT = HERE / "build"

DEFS = ["-DUGU_USE_OPENMP", "-DUGU_USE_STB", "-DUGU_USE_TINYOBJLOADER", "-DUGU_USE_LODEPNG",
        "-DUGU_USE_JSON", "-DUGU_USE_NANORT", "-DUGU_USE_TINYCOLORMAP", "-DUGU_USE_NANOFLANN", "-DNDEBUG"]


def compile_command() -> list:
    tp = REF / "third_party"
    incs = [f"-I{REF / 'include'}", f"-I{REF / 'src'}", f"-I{tp}", f"-I{tp / 'eigen'}", f"-I{tp / 'stb'}",
            f"-I{tp / 'tinyobjloader'}", f"-I{tp / 'lodepng'}", f"-I{tp / 'json' / 'single_include'}",
            f"-I{tp / 'nanort'}", f"-I{tp / 'tinycolormap' / 'include'}", f"-I{tp / 'nanoflann' / 'include'}",
            f"-I{REPO / 'include'}", f"-I{HERE}"]
    return ["g++", "-std=c++17", "-O2", "-fPIC", "-w", *DEFS, *incs]


def build() -> Path:
    if not (REF / "include" / "ugu").is_dir():
        raise ImportError("reference headers not present; the C++ adapter is prebuilt")
    OUT.mkdir(exist_ok=True)
    so = OUT / "librasterizer_cuda.so"
    subprocess.run([*compile_command(), "-shared", "-o", str(so), str(HERE / "rasterizer_cuda.cc"),
                    str(HERE / "visibility_tester_cuda.cc"), str(HERE / "multiview_cuda.cc"),
                    f"-L{REPO / 'ugu_b200' / 'lib'}", "-lugu_b200_raster", "-Wl,-rpath,$ORIGIN/../../lib"],
                   check=True, capture_output=True)
    return so


if __name__ == "__main__":
    print(build())
