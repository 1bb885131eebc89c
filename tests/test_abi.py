"""CPU suite: the C-ABI library loads, exports every symbol include/ugu_b200_raster.h declares, and
refuses to compute without a GPU (no CPU fallback)."""
import ctypes as C
import re
from pathlib import Path

import numpy as np
import pytest

from ugu_b200 import _native as N

REPO = Path(__file__).resolve().parents[1]
HEADER = REPO / "include" / "ugu_b200_raster.h"


def declared_symbols():
    text = HEADER.read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(ugu_b200_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_all_exported_and_bound():
    lib = N.load()
    names = declared_symbols()
    assert len(names) >= 18
    for n in names:
        assert hasattr(lib, n), f"{n} declared in the header but not exported"
        assert n in N.SYMBOLS, f"{n} declared in the header but not bound in _native.py"
    assert sorted(N.SYMBOLS) == names


def test_abi_version():
    assert N.load().ugu_b200_abi_version() == 10


def test_struct_sizes_match_header_layout():
    assert C.sizeof(N.Option) == 28
    assert C.sizeof(N.Camera) == 8 + 16 + 36 + 12 + 36
    assert C.sizeof(N.Outputs) == 40
    assert C.sizeof(N.MeshDesc) == 32 + 12 * 8
    assert C.sizeof(N.Stats) == 8 * 4 + 8 * 8
    assert C.sizeof(N.VisOption) == 40
    assert C.sizeof(N.Keyframe) == 8 + 108 + 4 + 48 + 12 + 4 + 32


def test_library_carries_sm100a_code_only():
    import subprocess
    out = subprocess.run(["cuobjdump", "-lelf", str(N.LIB_PATH)], capture_output=True, text=True).stdout
    assert "sm_100a" in out
    assert not re.search(r"sm_(?!100a)\d+", out), out


def test_no_gpu_means_loud_failure_not_fallback():
    lib = N.load()
    if lib.ugu_b200_device_count() > 0:
        pytest.skip("a B200 is present")
    h = C.c_void_p()
    rc = lib.ugu_b200_rasterizer_create(0, C.byref(h))
    assert rc == N.E_CUDA and not h
    assert "no CPU fallback" in N.last_error()
    from ugu_b200 import Rasterizer
    with pytest.raises(RuntimeError):
        Rasterizer()


def test_depth_to_u16_truncates_like_convertto():
    lib = N.load()
    d = np.array([0.0, 0.99, 1.0, 1234.56, 65535.9, 65536.0, 70000.2, -1.0], np.float32)
    w = np.zeros(len(d), np.uint16)
    assert lib.ugu_b200_depth_to_u16(d.ctypes.data, w.ctypes.data, len(d)) == 0
    assert w.tolist() == [0, 0, 1, 1234, 65535, 0, 70000 - 65536, 65535]


def test_product_never_references_the_oracle():
    root = Path(__file__).resolve().parents[1] / "ugu_b200"
    for p in list(root.rglob("*.py")) + list(root.rglob("*.cu")) + list(root.rglob("*.cuh")) + \
            list(root.rglob("*.cc")) + list(root.rglob("*.h")):
        text = p.read_text()
        assert "liboracle" not in text and "libugu_ref" not in text, p
        assert not re.search(r"^\s*(from|import)\s+oracle\b", text, flags=re.M), p


def test_cpp_adapter_fails_loudly_without_gpu():
    """The C++ drop-in (ugu::RasterizerCuda) reports 'no CUDA rasterizer' instead of falling back."""
    import subprocess
    import torch
    exe = REPO / "tests" / "cpp" / "build" / "adapter_check"
    if not exe.exists() or torch.cuda.is_available():
        pytest.skip("adapter harness not built here, or a GPU is present")
    p = subprocess.run([str(exe)], capture_output=True, text=True, timeout=120)
    assert p.returncode == 2 and "no CUDA rasterizer" in p.stdout
