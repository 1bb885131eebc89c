#!/usr/bin/env python
"""Secondary measurements: BASELINE.json configs 1, 3, 4 and 5 on ONE GPU (bench.py measures the
headline config 2).  Device-resident loop (outputs left in HBM, L2 flushed between views, CUDA-event
pairs on the launching stream) with the per-stage breakdown, plus the compiled reference on the
host for a bounded sample when oracle/_ref travelled.  Prints one JSON line per config.

    python tools/bench_configs.py [c1 c3 c4 c5] [--views N] [--no-cpu]
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

import cases  # noqa: E402
from ugu_b200 import (DiffuseColor, PinholeCamera, Rasterizer, RendererCpuOption, meshgen)  # noqa: E402

ALL = ("color", "depth", "normal", "mask", "face_id")


def config(name):
    """-> mesh, option, cameras, outputs, description"""
    if name == "c1":
        mesh = cases.load_spot()
        poses = np.load(cases.GOLDEN_DIR / "spot_ex02_poses.npy")
        cams = [PinholeCamera(640, 480, p, principal_point=(318.6, 255.3), focal_length=(517.3, 516.5))
                for p in poses]
        return mesh, RendererCpuOption(diffuse_color=DiffuseColor.kTexture), cams, ALL, \
            "config1: spot 5856 tris, 640x480, ex02 poses, texture bilinear, 5 outputs"
    if name == "c3":
        mesh = cases.load_spot()
        c, ext = cases.bbox_center_and_extent(mesh)
        cams = [PinholeCamera(1024, 1024, p, fov_y_deg=45.0) for p in meshgen.orbit_poses(c, 1.5 * ext, 36)]
        return mesh, RendererCpuOption(diffuse_color=DiffuseColor.kTexture), cams, ALL, \
            "config3: spot 36-view turntable, 1024x1024, texture bilinear, 5 outputs"
    if name == "c4":
        mesh = meshgen.uv_sphere(2237, 2236, tex_size=4096)
        cams = [PinholeCamera(4096, 4096, p, fov_y_deg=45.0) for p in meshgen.orbit_poses((0, 0, 0), 2.7, 8)]
        return mesh, RendererCpuOption(diffuse_color=DiffuseColor.kTexture), cams, ALL, \
            f"config4: uv-sphere {len(mesh.vertex_indices)} tris, 4096x4096, 4096^2 texture bilinear, 5 outputs"
    if name == "c5":
        mesh = meshgen.icosphere(8)
        cams = [PinholeCamera(512, 512, p, fov_y_deg=45.0) for p in meshgen.random_sphere_poses(128, seed=0)]
        return mesh, RendererCpuOption(diffuse_color=DiffuseColor.kNone), cams, ("face_id",), \
            "config5: icosphere 1310720 tris, 128 of the 1024 random views, 512x512, face_id only"
    raise SystemExit(f"unknown config {name}")


def run(name, max_views, cpu):
    from bench import ClockSampler
    mesh, opt, cams, outs, desc = config(name)
    sampler = ClockSampler(0)
    sampler.start()
    W, H = cams[0].width(), cams[0].height()
    P, F = W * H, len(mesh.vertex_indices)
    r = Rasterizer(opt)
    r.set_mesh(mesh)
    assert r.PrepareMesh()
    spec = dict(color=(torch.uint8, 3), depth=(torch.float32, 1), normal=(torch.float32, 3),
                mask=(torch.uint8, 1), face_id=(torch.int32, 1))
    dev = {k: torch.empty(P * spec[k][1], dtype=spec[k][0], device="cuda") for k in outs}
    ptrs = {k: v.data_ptr() for k, v in dev.items()}
    flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
    ts = torch.cuda.Stream()
    torch.cuda.set_stream(ts)
    natives = [c.to_native() for c in cams]
    n = min(max_views, len(cams)) if max_views else len(cams)
    for i in range(3):
        flush.zero_()
        r.render_device(natives[i % len(natives)], stream=ts.cuda_stream, **ptrs)
    torch.cuda.synchronize()
    def timed_pass():
        e0 = [torch.cuda.Event(enable_timing=True) for _ in range(n)]
        e1 = [torch.cuda.Event(enable_timing=True) for _ in range(n)]
        for i in range(n):
            flush.zero_()
            e0[i].record()
            r.render_device(natives[i % len(natives)], stream=ts.cuda_stream, **ptrs)
            e1[i].record()
        torch.cuda.synchronize()
        return sum(a.elapsed_time(b) for a, b in zip(e0, e1)) / n

    ms = timed_pass()                      # the per-view time: only the view's kernels between the events
    r.reset_stats()
    r.set_profiling(True, (0, 1, 2, 3))    # second pass with per-stage event pairs for the breakdown
    timed_pass()
    st = r.stats()
    r.set_profiling(False, ())
    # the batched multi-view path: ONE render_batch call over all n views (launch sets of up to 16 views,
    # view = blockIdx.y), outputs device-resident, one event pair around the call; L2 flushed before the
    # batch, NOT between its views (a batch is one unit of work)
    bdev = {k: torch.empty(n * P * spec[k][1], dtype=spec[k][0], device="cuda") for k in outs}
    bptrs = {k: v.data_ptr() for k, v in bdev.items()}
    blist = [natives[i % len(natives)] for i in range(n)]
    r.render_batch(blist, on_device=True, stream=ts.cuda_stream, **bptrs)
    torch.cuda.synchronize()
    bt = []
    for rep in range(5):
        flush.zero_()
        b0, b1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        b0.record()
        r.render_batch(blist, on_device=True, stream=ts.cuda_stream, **bptrs)
        b1.record()
        torch.cuda.synchronize()
        bt.append(b0.elapsed_time(b1) / n)
    r.reset_stats()
    r.set_profiling(True, (0, 1, 2, 3))    # per-stage event pairs around each launch set of the batch
    r.render_batch(blist, on_device=True, stream=ts.cuda_stream, **bptrs)
    torch.cuda.synchronize()
    bst = r.stats()
    r.set_profiling(False, ())
    batch_stage_us = {k: round(v / max(1, bst["timed_views"]) * 1e3, 2) for k, v in bst["stage_ms"].items()}
    del bdev
    line = {"config": desc, "views": n, "ms_per_view": ms, "mpix_per_s": P / ms / 1e3,
            "batch_ms_per_view": {"min": min(bt), "median": sorted(bt)[len(bt) // 2], "max": max(bt),
                                  "stage_us_per_view": batch_stage_us,
                                  "note": "one render_batch call over all views, L2 flushed before the batch only"},
            "mtri_per_s": F / ms / 1e3,
            "stage_us": {k: round(v / max(1, st["timed_views"]) * 1e3, 2) for k, v in st["stage_ms"].items()},
            "large_triangles": st["large_triangles"], "tile_jobs": st["tile_jobs"]}
    # end to end through the public Render with pinned host images
    from ugu_b200 import Image1b, Image1f, Image1i, Image3b, Image3f
    imgs = dict(color=Image3b(), depth=Image1f(), normal=Image3f(), mask=Image1b(), face_id=Image1i())
    args = {k: (imgs[k] if k in outs else None) for k in ALL}
    r.set_camera(cams[0])
    assert r.Render(**args)
    t0 = time.perf_counter()
    ne = min(n, 16)
    for i in range(ne):
        r.set_camera(cams[i % len(cams)])
        assert r.Render(**args)
    line["e2e_ms_per_view"] = (time.perf_counter() - t0) / ne * 1e3
    line["clocks"] = sampler.stop()
    if cpu:
        import oracle
        if oracle.ref_available():
            os.environ["OMP_NUM_THREADS"] = str(os.cpu_count() or 1)
            ref = oracle.OracleRef(mesh)
            ref.render(cams[0], opt, want=outs)
            t0 = time.perf_counter()
            k = 0
            while k < min(n, 6) and time.perf_counter() - t0 < 25.0:
                ref.render(cams[(k + 1) % len(cams)], opt, want=outs)
                k += 1
            line["cpu_reference_ms_per_view"] = (time.perf_counter() - t0) / max(k, 1) * 1e3
            line["cpu_views"] = k
            line["cpu_cores"] = os.cpu_count()
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    names = [a for a in sys.argv[1:] if not a.startswith("--")] or ["c1", "c3", "c5", "c4"]
    views = 0
    if "--views" in sys.argv:
        views = int(sys.argv[sys.argv.index("--views") + 1])
        names = [x for x in names if x != str(views)]
    for nm in names:
        run(nm, views, "--no-cpu" not in sys.argv)
