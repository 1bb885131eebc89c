"""TEST INFRASTRUCTURE: builds tests/cpp/build/adapter_check, which links the product's C++ adapter
(ugu_b200/cpp/build/librasterizer_cuda.so) AND the compiled reference (oracle/_ref/libugu_ref.so)
to compare them behind the same ugu::RendererCpu interface.  Needs the reference headers, so it is
built in this container and the binary travels to the GPU box."""
from __future__ import annotations

import subprocess
import sys
from pathlib import Path

HERE = Path(__file__).resolve().parent
REPO = HERE.parents[1]
sys.path.insert(0, str(REPO))
OUT = HERE / "build"
EXE = OUT / "adapter_check"


# libgomp by path, like oracle/Makefile: the harness pins the reference's OpenMP thread count
GOMP = sorted(str(p) for p in list(Path("/usr/lib/gcc/x86_64-linux-gnu").glob("*/libgomp.so"))[:1]) or \
    ["/usr/lib/x86_64-linux-gnu/libgomp.so.1"]


def build() -> Path:
    from ugu_b200.cpp import build_adapter
    so = build_adapter.build()
    OUT.mkdir(exist_ok=True)
    subprocess.run([*build_adapter.compile_command(), "-o", str(EXE), str(HERE / "adapter_check.cc"),
                    f"-L{so.parent}", "-lrasterizer_cuda", f"-L{REPO / 'oracle' / '_ref'}", "-lugu_ref",
                    f"-L{REPO / 'ugu_b200' / 'lib'}", "-lugu_b200_raster", *GOMP,
                    "-Wl,-rpath,$ORIGIN/../../../ugu_b200/cpp/build:$ORIGIN/../../../ugu_b200/lib:"
                    "$ORIGIN/../../../oracle/_ref"],
                   check=True, capture_output=True)
    return EXE


if __name__ == "__main__":
    print(build())
