#!/usr/bin/env python
"""UV-space attribute rasterization on ONE GPU (raster_util.h:139-268).

Workloads (synthetic, the shapes of the reference's own callers):
  posmap   uv-sphere, ~1.0M faces, vertex positions -> 2048x2048 Image3f + mask  (app/posmap_gen.cc:74)
  vc2tex   spot, 5856 faces, vertex colours -> 4096x4096 Image3b + mask          (app/vc2tex.cc:138)
For each: device-resident time (inputs and outputs in HBM, L2 flushed between iterations, CUDA-event
pairs on the launching stream), the HBM roofline fraction on algorithmic bytes, the end-to-end host
form (H2D of the arrays + kernels + D2H of texture and mask), and the compiled reference on the
host (1 thread = the pinned deterministic order; all threads = its racy default) when oracle/_ref
travelled.  One JSON line per workload.

    python tools/bench_uvraster.py [posmap vc2tex] [--iters N] [--no-cpu]
"""
from __future__ import annotations

import ctypes as C
import json
import os
import sys
import time
from pathlib import Path

REPO = Path(__file__).resolve().parents[1]
for p in (str(REPO), str(REPO / "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np  # noqa: E402
import torch  # noqa: E402

import cases  # noqa: E402
from ugu_b200 import Image1b, Image3b, Image3f, Rasterizer, meshgen  # noqa: E402
from ugu_b200 import _native as N  # noqa: E402


def workload(name):
    if name == "posmap":
        m = meshgen.uv_sphere(708, 708)
        return dict(attrs=np.asarray(m.vertices, np.float32), attr_indices=np.asarray(m.vertex_indices, np.int32),
                    uv=np.asarray(m.uv, np.float32), uv_indices=np.asarray(m.uv_indices, np.int32),
                    width=2048, height=2048, texture_type=1,
                    desc="posmap: uv-sphere %d faces, positions -> 2048x2048 Image3f + mask" % len(m.vertex_indices))
    if name == "vc2tex":
        m = cases.load_spot()
        rng = np.random.default_rng(0)
        return dict(attrs=(rng.random((len(m.vertices), 3)) * 255).astype(np.float32),
                    attr_indices=np.asarray(m.vertex_indices, np.int32), uv=np.asarray(m.uv, np.float32),
                    uv_indices=np.asarray(m.uv_indices, np.int32), width=4096, height=4096, texture_type=0,
                    desc="vc2tex: spot 5856 faces, vertex colours -> 4096x4096 Image3b + mask")
    raise SystemExit(f"unknown workload {name}")


def run(name, iters, cpu):
    c = workload(name)
    w, h, tt = c["width"], c["height"], c["texture_type"]
    P, F = w * h, len(c["uv_indices"])
    texel = 12 if tt else 3
    algo = 12 * len(c["attrs"]) + 8 * len(c["uv"]) + 24 * F + P * (texel + 1)
    r = Rasterizer()
    dev = {k: torch.from_numpy(np.ascontiguousarray(c[k])).cuda() for k in ("attrs", "attr_indices", "uv", "uv_indices")}
    tex = torch.empty(P * texel, dtype=torch.uint8, device="cuda")
    mask = torch.empty(P, dtype=torch.uint8, device="cuda")
    d = N.UvRasterDesc(len(c["attrs"]), F, len(c["uv"]), 0, dev["attrs"].data_ptr(), dev["attr_indices"].data_ptr(),
                       dev["uv"].data_ptr(), dev["uv_indices"].data_ptr(), w, h, tt, 0, 0, 0)
    flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
    ts = torch.cuda.Stream()
    torch.cuda.set_stream(ts)
    for _ in range(3):
        flush.zero_()
        r.uv_raster_device(d, tex.data_ptr(), mask.data_ptr(), ts.cuda_stream)
    torch.cuda.synchronize()
    e0 = [torch.cuda.Event(enable_timing=True) for _ in range(iters)]
    e1 = [torch.cuda.Event(enable_timing=True) for _ in range(iters)]
    for i in range(iters):
        flush.zero_()
        e0[i].record()
        r.uv_raster_device(d, tex.data_ptr(), mask.data_ptr(), ts.cuda_stream)
        e1[i].record()
    torch.cuda.synchronize()
    ms = sum(a.elapsed_time(b) for a, b in zip(e0, e1)) / iters
    peak = json.loads((REPO / "MEASURED_PEAKS.json").read_text()).get("hbm_gbs", 6550.0) \
        if (REPO / "MEASURED_PEAKS.json").exists() else 6550.0
    line = {"workload": c["desc"], "iters": iters, "ms_per_call": ms, "mtexel_per_s": P / ms / 1e3,
            "mtri_per_s": F / ms / 1e3, "gpu_launches_per_call": 3,
            "roofline": {"bound": "hbm", "algorithmic_bytes": algo, "achieved": algo / ms / 1e6,
                         "peak": peak, "unit": "GB/s", "frac": algo / ms / 1e6 / peak},
            "l2": "flushed between timed iterations (512 MiB write)"}
    # end to end through the public host form (pinned images)
    timg = (Image3f if tt else Image3b)()
    mimg = Image1b()
    mimg._init(w, h)
    mimg.data[...] = 0
    assert r.RasterizeVertexAttributeToTexture(c["attrs"], c["attr_indices"], c["uv"], c["uv_indices"], timg, w, h, mimg)
    ne = max(3, min(iters, 10))
    t0 = time.perf_counter()
    for _ in range(ne):
        assert r.RasterizeVertexAttributeToTexture(c["attrs"], c["attr_indices"], c["uv"], c["uv_indices"], timg, w, h, mimg)
    line["e2e_ms_per_call"] = (time.perf_counter() - t0) / ne * 1e3
    line["e2e_h2d_bytes"] = 12 * len(c["attrs"]) + 8 * len(c["uv"]) + 24 * F + P
    line["e2e_d2h_bytes"] = P * (texel + 1)
    if cpu:
        import oracle
        if oracle.ref_available():
            lib = C.CDLL(str(oracle.REF_LIB))
            m0 = np.zeros((h, w), np.uint8)
            for threads, key in ((1, "cpu_reference_1thread_ms"), (-1, "cpu_reference_all_threads_ms")):
                lib.ugu_ref_uv_set_threads(threads)
                t0 = time.perf_counter()
                k = 0
                while k < 3 and time.perf_counter() - t0 < 15.0:
                    got = oracle.ref_uv_raster(c["attrs"], c["attr_indices"], c["uv"], c["uv_indices"], w, h, tt, None, m0)
                    k += 1
                line[key] = (time.perf_counter() - t0) / k * 1e3
            lib.ugu_ref_uv_set_threads(1)
            line["cpu_cores"] = os.cpu_count()
            line["cpu_note"] = "includes the shim's array -> std::vector conversion and image copies"
            # parity at the bench size: the 1-thread reference vs the device-resident result
            got = oracle.ref_uv_raster(c["attrs"], c["attr_indices"], c["uv"], c["uv_indices"], w, h, tt, None, m0)
            dt = np.float32 if tt else np.uint8
            mine = tex.cpu().numpy().view(dt).reshape(h, w, 3)
            line["matches_reference_bit_for_bit"] = bool(
                np.array_equal(mine.view(np.uint8), got[0].view(np.uint8)) and
                np.array_equal(mask.cpu().numpy().reshape(h, w), got[1]))
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    names = [a for a in sys.argv[1:] if not a.startswith("--") and not a.isdigit()] or ["posmap", "vc2tex"]
    iters = int(sys.argv[sys.argv.index("--iters") + 1]) if "--iters" in sys.argv else 30
    for nm in names:
        run(nm, iters, "--no-cpu" not in sys.argv)
