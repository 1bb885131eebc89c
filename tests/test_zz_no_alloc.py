"""ugu_b200_face_visibility_static — the allocation-free entry point (static __device__ scratch,
include/ugu_b200_raster.h) — against ugu_b200_face_visibility_device on the same device maps.
Integer work: every accumulator must be bit-identical.  Collected last on purpose (file name): the
entry point is an addition outside SURVEY.md §8 and must not hide other results.
"""
import ctypes as C

import numpy as np
import pytest


def _lib():
    from ugu_b200 import _native as N
    return N, N.load()


def test_static_entry_point_is_exported_and_declared():
    N, lib = _lib()
    assert hasattr(lib, "ugu_b200_face_visibility_static")
    from pathlib import Path
    header = (Path(__file__).resolve().parents[1] / "include" / "ugu_b200_raster.h").read_text()
    assert "ugu_b200_face_visibility_static(" in header
    assert "UGU_B200_STATIC_VIS_MAX_FACES (4 * 1024 * 1024)" in header


def _accumulators(torch, F):
    return (torch.zeros(F, dtype=torch.int32, device="cuda:0"), torch.zeros(F, dtype=torch.int32, device="cuda:0"),
            torch.zeros(F, dtype=torch.int64, device="cuda:0"))


def _run_static(lib, maps, F, first, acc):
    err = C.create_string_buffer(256)
    views, P = maps.shape[0], int(np.prod(maps.shape[1:]))
    rc = lib.ugu_b200_face_visibility_static(
        0, C.c_void_p(maps.data_ptr()), views, P, first, F, C.c_void_p(acc[0].data_ptr()),
        C.c_void_p(acc[1].data_ptr()), C.c_void_p(acc[2].data_ptr()), None, err, len(err))
    return rc, err.value.decode()


def _run_device(N, lib, handle, maps, F, first, acc):
    views, P = maps.shape[0], int(np.prod(maps.shape[1:]))
    rc = lib.ugu_b200_face_visibility_device(
        handle, C.c_void_p(maps.data_ptr()), views, P, first, F, C.c_void_p(acc[0].data_ptr()),
        C.c_void_p(acc[1].data_ptr()), C.c_void_p(acc[2].data_ptr()), None)
    assert rc == N.OK, N.last_error(handle)


@pytest.mark.gpu
def test_static_scratch_matches_device_form_on_consecutive_calls():
    import torch
    N, lib = _lib()
    handle = C.c_void_p()
    assert lib.ugu_b200_rasterizer_create(0, C.byref(handle)) == N.OK, N.last_error()
    try:
        rng = np.random.default_rng(7)
        # consecutive calls with different face counts / slot layouts / odd sizes (scalar loads),
        # the last one repeating the first: stale scratch from any of them would show
        cases = [((11, 37, 19), 1000, 0), ((20, 64, 64), 7, 5), ((3, 8, 16), 600_000, 0), ((9, 128, 128), 50, 100),
                 ((11, 37, 19), 1000, 0)]
        free_before = None
        for n, (shape, F, first) in enumerate(cases):
            maps = torch.from_numpy(rng.integers(-1, F + 2, shape).astype(np.int32)).to("cuda:0")
            want, got = _accumulators(torch, F), _accumulators(torch, F)
            torch.cuda.synchronize()  # the device form runs on the rasterizer's own stream
            _run_device(N, lib, handle, maps, F, first, want)
            torch.cuda.synchronize()
            if n >= 1:
                free_before = torch.cuda.mem_get_info(0)[0]
            for _ in range(2):  # two calls in a row accumulate into the same arrays, like the device form
                rc, msg = _run_static(lib, maps, F, first, got)
                assert rc == N.OK, msg
            torch.cuda.synchronize()
            if free_before is not None:  # the call allocated nothing on the device
                assert torch.cuda.mem_get_info(0)[0] == free_before
            _run_device(N, lib, handle, maps, F, first, want)
            torch.cuda.synchronize()
            for g, w, name in zip(got, want, ("pixel_count", "view_count", "best")):
                assert torch.equal(g, w), (name, shape, F)
            assert int(want[0].sum()) > 0
    finally:
        lib.ugu_b200_rasterizer_destroy(handle)


@pytest.mark.gpu
def test_static_scratch_rejects_sizes_beyond_the_array():
    import torch
    N, lib = _lib()
    limit = 4 * 1024 * 1024
    maps = torch.zeros((1, 4, 4), dtype=torch.int32, device="cuda:0")
    acc = _accumulators(torch, 16)  # never touched: the call must fail before any launch
    rc, msg = _run_static(lib, maps, limit + 1, 0, acc)
    assert rc == N.E_INVALID and "UGU_B200_STATIC_VIS_MAX_FACES" in msg
    # the limit itself is accepted (one view slot)
    F = limit
    maps = torch.tensor([[0, F - 1, F - 1, -1], [F - 1, 5, 5, 5]], dtype=torch.int32, device="cuda:0")
    acc = _accumulators(torch, F)
    rc, msg = _run_static(lib, maps, F, 0, acc)
    assert rc == N.OK, msg
    torch.cuda.synchronize()
    assert acc[0][F - 1].item() == 3 and acc[1][F - 1].item() == 2 and acc[0][5].item() == 3
    assert acc[2][F - 1].item() == (2 << 32) | 0xFFFFFFFF  # best: 2 pixels in view 0
