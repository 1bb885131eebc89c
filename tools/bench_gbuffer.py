#!/usr/bin/env python
"""G-buffer outputs on ONE GPU (ugu::GBuffer, include/ugu/renderer/base.h:34-77, as
RendererGl::ReadGbuf fills it, src/renderer/gl/renderer.cc:501-650).  Workload: config 2's view (1.31M-
triangle icosphere, vertex colours + vertex normals, 1920x1080), all eleven images.  Device-resident time
of the whole call (vertex, raster, binned, gbuffer, resolve kernels; outputs left in HBM; L2 flushed
between iterations; CUDA-event pairs on the launching stream), the same for a colour-only Render (the
difference is the gbuffer kernel + what the slimmer resolve saves), the HBM roofline fraction on
algorithmic bytes, the end-to-end host form, and the C restatement (oracle, "port") on the host.
One JSON line.

    python tools/bench_gbuffer.py [--iters N] [--no-cpu]
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

from ugu_b200 import DiffuseColor, GBuffer, PinholeCamera, Rasterizer, RendererCpuOption, meshgen  # noqa: E402

FIELDS = (("color", 3), ("normal_cam", 12), ("normal_wld", 12), ("pos_cam", 12), ("pos_wld", 12), ("depth_01", 4),
          ("stencil", 1), ("face_id", 4), ("uv", 12), ("bary", 12), ("geo_id", 4))


def main():
    iters = int(sys.argv[sys.argv.index("--iters") + 1]) if "--iters" in sys.argv else 30
    w, h, zn, zf = 1920, 1080, 0.5, 10.0
    opt = RendererCpuOption(diffuse_color=DiffuseColor.kVertex)
    mesh = meshgen.icosphere(8)
    r = Rasterizer(opt)
    r.set_mesh(mesh)
    assert r.PrepareMesh()
    cam = PinholeCamera(w, h, meshgen.look_at_c2w((0, 0, -2.7), (0, 0, 0), (0, -1, 0)), fov_y_deg=45.0)
    P = w * h
    bufs = {n: torch.empty(P * b, dtype=torch.uint8, device="cuda") for n, b in FIELDS}
    flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
    ts = torch.cuda.Stream()
    torch.cuda.set_stream(ts)
    ptrs = {n: b.data_ptr() for n, b in bufs.items()}

    def run_gbuf():
        r.render_gbuffer_device(cam, zn, zf, stream=ts.cuda_stream, **ptrs)

    def run_color():
        r.render_device(cam, color=ptrs["color"], stream=ts.cuda_stream)

    def timed(fn):
        for _ in range(3):
            flush.zero_()
            fn()
        e0 = [torch.cuda.Event(enable_timing=True) for _ in range(iters)]
        e1 = [torch.cuda.Event(enable_timing=True) for _ in range(iters)]
        for i in range(iters):
            flush.zero_()
            e0[i].record()
            fn()
            e1[i].record()
        torch.cuda.synchronize()
        return sum(a.elapsed_time(b) for a, b in zip(e0, e1)) / iters

    ms_color = timed(run_color)
    ms = timed(run_gbuf)
    V, F = len(mesh.vertices), len(mesh.vertex_indices)
    out_bytes = sum(b for _, b in FIELDS)
    # compulsory traffic of the gbuffer kernel: the key image once, every output once, positions /
    # normals / both index arrays once.  Of the whole call: + the view's own inputs (12V + 24F) + colours.
    algo_kernel = P * (8 + out_bytes - 3) + 12 * V + 12 * V + 24 * F
    algo_call = P * out_bytes + 12 * V + 24 * F + 12 * V + 12 * F + 12 * V
    peak_file = REPO / "MEASURED_PEAKS.json"
    peak = json.loads(peak_file.read_text()).get("hbm_gbs", 6550.0) if peak_file.exists() else 6550.0
    covered = int((bufs["stencil"] == 255).sum().item())
    line = {"workload": f"G-buffer (11 images, {out_bytes} B/pixel) of config 2's 1920x1080 view: "
                        f"{F} triangles, {covered} covered pixels", "iters": iters, "ms_per_call": ms,
            "ms_color_only_render": ms_color, "ms_gbuffer_increment": ms - ms_color,
            "mpix_per_s": P / ms / 1e3, "mtri_per_s": F / ms / 1e3, "gpu_launches_per_call": 5,
            "roofline": {"bound": "hbm", "scope": "whole call", "algorithmic_bytes": algo_call,
                         "achieved": algo_call / ms / 1e6, "peak": peak, "unit": "GB/s",
                         "frac": algo_call / ms / 1e6 / peak},
            "gbuffer_kernel_algorithmic_bytes": algo_kernel,
            "l2": "flushed between timed iterations (512 MiB write)"}
    # end to end: host images out through the public API
    r.set_camera(cam)
    gb = GBuffer()
    assert r.RenderGBuffer(gb, zn, zf)
    ne = max(3, min(iters, 10))
    t0 = time.perf_counter()
    for _ in range(ne):
        assert r.RenderGBuffer(gb, zn, zf)
    line["e2e_ms_per_call"] = (time.perf_counter() - t0) / ne * 1e3
    line["e2e_d2h_bytes"] = P * out_bytes
    if "--no-cpu" not in sys.argv:
        from oracle import GBUFFER_FIELDS, OraclePort, port_gbuffer
        port = OraclePort(mesh)
        t0 = time.perf_counter()
        k = 0
        while k < 3 and time.perf_counter() - t0 < 20.0:
            want = port_gbuffer(port, cam.to_native(), opt, zn, zf)
            k += 1
        line["cpu_port_ms"] = (time.perf_counter() - t0) / k * 1e3
        line["cpu_cores"] = 1
        line["cpu_note"] = "oracle_gbuffer (C restatement, single-threaded, two z-buffer passes); parity unpinned vs GL"
        same = all(np.array_equal(np.ascontiguousarray(getattr(gb, n).data).view(np.uint8),
                                  np.ascontiguousarray(want[n]).view(np.uint8)) for n, _, _ in GBUFFER_FIELDS)
        line["matches_restatement_bit_for_bit"] = bool(same)
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
