#!/usr/bin/env python
"""Raytracer parity mode on ONE GPU (ugu::Raytracer::Render, src/renderer/cpu/raytracer.cc:
117-225, nanort nearest hit).  Workload: config 2 (1.31M-triangle icosphere, vertex colours, 1920x1080, all
five outputs, culling on).  Device-resident time of one view (vertex, ray-table, raster, large, resolve
kernels; outputs left in HBM; L2 flushed between iterations; CUDA-event pairs on the launching stream), the
HBM roofline fraction on the view's algorithmic bytes, the end-to-end host form, and the compiled reference
Raytracer (BVH, OpenMP over rows) on the host with a bit-for-bit comparison of all five outputs.
One JSON line.

    python tools/bench_raytrace.py [--iters N] [--no-cpu] [--level L]
"""
from __future__ import annotations

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

from ugu_b200 import (DiffuseColor, Image1b, Image1f, Image1i, Image3b, Image3f, PinholeCamera, Raytracer,  # noqa: E402
                      RendererCpuOption, meshgen)


def main():
    iters = int(sys.argv[sys.argv.index("--iters") + 1]) if "--iters" in sys.argv else 30
    level = int(sys.argv[sys.argv.index("--level") + 1]) if "--level" in sys.argv else 8
    w, h = 1920, 1080
    opt = RendererCpuOption(diffuse_color=DiffuseColor.kVertex)
    c4 = "--c4" in sys.argv  # BASELINE config 4: 10 M-triangle textured uv-sphere at 4096x4096 (no CPU leg:
    if c4:                   # the reference's BVH needs tens of minutes on this mesh)
        w, h = 4096, 4096
        opt = RendererCpuOption(diffuse_color=DiffuseColor.kTexture)
        mesh = meshgen.uv_sphere(2237, 2236, tex_size=4096)
        sys.argv.append("--no-cpu")
    else:
        mesh = meshgen.icosphere(level)
    r = Raytracer(opt)
    r.set_mesh(mesh)
    assert r.PrepareMesh()
    cam = PinholeCamera(w, h, meshgen.orbit_poses((0, 0, 0), 2.7, 8)[3] if c4 else
                        meshgen.look_at_c2w((0, 0, -2.7), (0, 0, 0), (0, -1, 0)), fov_y_deg=45.0)
    P = w * h
    col = torch.empty(P * 3, dtype=torch.uint8, device="cuda")
    dep = torch.empty(P, dtype=torch.float32, device="cuda")
    nrm = torch.empty(P * 3, dtype=torch.float32, device="cuda")
    msk = torch.empty(P, dtype=torch.uint8, device="cuda")
    fid = torch.empty(P, dtype=torch.int32, device="cuda")
    flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
    ts = torch.cuda.Stream()
    torch.cuda.set_stream(ts)

    def run():
        r.render_device(cam, color=col.data_ptr(), depth=dep.data_ptr(), normal=nrm.data_ptr(), mask=msk.data_ptr(),
                        face_id=fid.data_ptr(), stream=ts.cuda_stream)

    for _ in range(3):
        flush.zero_()
        run()
    e0 = [torch.cuda.Event(enable_timing=True) for _ in range(iters)]
    e1 = [torch.cuda.Event(enable_timing=True) for _ in range(iters)]
    for i in range(iters):
        flush.zero_()
        e0[i].record()
        run()
        e1[i].record()
    torch.cuda.synchronize()
    ms = sum(a.elapsed_time(b) for a, b in zip(e0, e1)) / iters
    V, F = len(mesh.vertices), len(mesh.vertex_indices)
    algo = 12 * V + 12 * F + 12 * F + 12 * V + 12 * F + 12 * V + 24 * P  # as the rasterizer's view
    if c4:  # texture path: uv + uv indices + material ids + texels instead of vertex colours
        algo = 12 * V + 12 * F + 12 * F + 12 * V + 12 * F + 8 * len(mesh.uv) + 12 * F + 4 * F + 3 * 4096 * 4096 + 24 * P
    peak_file = REPO / "MEASURED_PEAKS.json"
    peak = json.loads(peak_file.read_text()).get("hbm_gbs", 6550.0) if peak_file.exists() else 6550.0
    covered = int((fid >= 0).sum().item())
    line = {"workload": f"Raytracer mode, {'uv-sphere, texture' if c4 else f'icosphere L{level}, vertex colours'} ({F} triangles), {w}x{h}, 5 outputs, "
                        f"{covered} covered pixels", "iters": iters, "ms_per_view": ms, "mpix_per_s": P / ms / 1e3,
            "mtri_per_s": F / ms / 1e3, "mrays_per_s": P / ms / 1e3, "gpu_launches_per_view": 5,
            "roofline": {"bound": "hbm", "scope": "whole view", "algorithmic_bytes": algo,
                         "achieved": algo / ms / 1e6, "peak": peak, "unit": "GB/s", "frac": algo / ms / 1e6 / peak},
            "l2": "flushed between timed iterations (512 MiB write)"}
    imgs = (Image3b(), Image1f(), Image3f(), Image1b(), Image1i())
    r.set_camera(cam)
    assert r.Render(*imgs)
    ne = max(3, min(iters, 10))
    t0 = time.perf_counter()
    for _ in range(ne):
        assert r.Render(*imgs)
    line["e2e_ms_per_view"] = (time.perf_counter() - t0) / ne * 1e3
    line["e2e_d2h_bytes"] = 24 * P
    if "--no-cpu" not in sys.argv:
        import oracle
        if oracle.ref_available():
            ref = oracle.OracleRef(mesh)
            want = ref.raytrace(cam, opt)  # builds the BVH, fills the ray tables
            first_ms = ref.last_raytrace_ms
            want = ref.raytrace(cam, opt)
            line["cpu_reference_ms"] = ref.last_raytrace_ms
            line["cpu_reference_first_call_ms"] = first_ms
            line["cpu_cores"] = os.cpu_count()
            line["cpu_note"] = "compiled ugu::Raytracer, nanort BVH, OpenMP over image rows (all cores)"
            keys = ("color", "depth", "normal", "mask", "face_id")
            diffs = {k: int((np.ascontiguousarray(im.data).view(np.uint8).reshape(-1) !=
                             np.ascontiguousarray(want[k]).view(np.uint8).reshape(-1)).sum())
                     for k, im in zip(keys, imgs)}
            line["differing_bytes_vs_reference"] = diffs
            line["matches_reference_bit_for_bit"] = not any(diffs.values())
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
