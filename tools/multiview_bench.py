#!/usr/bin/env python
"""Multi-view dispatch over the GPUs of one box (BASELINE.json configs 3 and 5).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29511 tools/multiview_bench.py [c5|c3] [--views V] [--gather] [--check K] [--visibility] [--raytrace]

Every rank uploads the mesh to its own GPU (replicated, read-only), renders its contiguous slice of
the view list with ONE device-resident batch call (no data-path collective), optionally gathers the
per-view face_id maps to rank 0 over NCCL, and rank 0 prints one JSON line: whole-job views/s,
Mpix/s and Mtri/s (device time, max over ranks, barrier + synchronize on both sides).
`--check K` makes rank 0 re-render K of the gathered views on its own and compare bit for bit
(N-GPU result == 1-GPU result, SURVEY.md §4 iii).
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
import torch.distributed as dist  # noqa: E402


def main():
    args = sys.argv[1:]
    name = next((a for a in args if a in ("c3", "c5")), "c5")
    views = int(args[args.index("--views") + 1]) if "--views" in args else None
    check = int(args[args.index("--check") + 1]) if "--check" in args else 0
    gather = "--gather" in args or check > 0
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    import cases
    from ugu_b200 import DiffuseColor, PinholeCamera, Rasterizer, RendererCpuOption, meshgen
    from ugu_b200.multiview import MultiViewRenderer

    if name == "c5":
        mesh = meshgen.icosphere(8)
        n = views or 1024
        cams = [PinholeCamera(512, 512, p, fov_y_deg=45.0) for p in meshgen.random_sphere_poses(n, seed=0)]
        opt = RendererCpuOption(diffuse_color=DiffuseColor.kNone)
        outs = ("face_id",)
        desc = f"config5: icosphere 1310720 tris, {n} random views, 512x512, face_id only"
    else:
        mesh = cases.load_spot()
        c, ext = cases.bbox_center_and_extent(mesh)
        n = views or 36
        cams = [PinholeCamera(1024, 1024, p, fov_y_deg=45.0) for p in meshgen.orbit_poses(c, 1.5 * ext, n)]
        opt = RendererCpuOption(diffuse_color=DiffuseColor.kTexture)
        outs = ("color", "depth", "normal", "mask", "face_id")
        desc = f"config3: spot {n}-view turntable, 1024x1024, texture bilinear, 5 outputs"
    W, H = cams[0].width(), cams[0].height()
    F = len(mesh.vertex_indices)

    if "--raytrace" in args:  # the Raytracer parity mode (ugu::Raytracer outputs) through the same driver
        from ugu_b200 import Raytracer
        r = Raytracer(opt, device=local)
        desc += " [Raytracer mode]"
    else:
        r = Rasterizer(opt, device=local)
    r.set_mesh(mesh)
    assert r.PrepareMesh()
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    mv = MultiViewRenderer(r, mode="contiguous")

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    mv.render(cams, outs)  # warm-up (allocations, first launches)
    if gather:
        mv.render_and_gather(cams, "face_id", dst=0)
    barrier()
    e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
    t0 = time.perf_counter()
    e0.record()
    idx, dev = mv.render(cams, outs)
    e1.record()
    gathered = None
    if gather:
        from ugu_b200.multiview import gather_view_maps
        local_maps = dev["face_id"] if idx else torch.empty((0, H, W), dtype=torch.int32, device="cuda")
        gathered = gather_view_maps(local_maps, len(cams), "contiguous", 0)
    e2.record()
    barrier()
    wall = time.perf_counter() - t0
    t = torch.tensor([e0.elapsed_time(e1), e0.elapsed_time(e2)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    render_ms, total_ms = float(t[0]), float(t[1])

    vis_ms = vis_ok = None
    if "--visibility" in args:
        # SURVEY.md 8f consumer: per-face statistics over ALL views without moving a map — render the
        # local slice in chunks, fold into device accumulators, all-reduce 16 bytes per face
        mv.render_face_visibility(cams[: 2 * world], F)  # warm-up (scratch allocation)
        barrier()
        v0, v1 = (torch.cuda.Event(enable_timing=True) for _ in range(2))
        v0.record()
        vis = mv.render_face_visibility(cams, F, chunk_views=32)
        v1.record()
        barrier()
        tv = torch.tensor([v0.elapsed_time(v1)], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(tv, op=dist.ReduceOp.MAX)
        vis_ms = float(tv[0])
        if gathered is not None and rank == 0:  # against a histogram of the gathered maps
            ids = gathered.reshape(-1)
            want = torch.bincount(ids[ids >= 0].to(torch.int64), minlength=F)
            vis_ok = bool(torch.equal(want.to(torch.int32), vis["pixel_count"]))

    ok = None
    if check and rank == 0:
        sel = np.linspace(0, len(cams) - 1, check).astype(int)
        ok = True
        for i in sel:
            one = torch.empty((H, W), dtype=torch.int32, device="cuda")
            r.render_device(cams[int(i)] if "--raytrace" in args else cams[int(i)].to_native(),
                            stream=stream.cuda_stream, face_id=one.data_ptr())
            torch.cuda.synchronize()
            ok = ok and bool(torch.equal(one, gathered[int(i)]))
    if rank == 0:
        line = {"config": desc, "n_gpus": world, "views": len(cams), "render_ms": render_ms,
                "views_per_s": len(cams) / (render_ms * 1e-3),
                "mpix_per_s": len(cams) * W * H / (render_ms * 1e-3) / 1e6,
                "mtri_per_s": len(cams) * F / (render_ms * 1e-3) / 1e6,
                "ms_per_view_per_gpu": render_ms / max(1, -(-len(cams) // world)),
                "gather": gather, "render_plus_gather_ms": total_ms if gather else None,
                "gathered_bytes": int(gathered.numel() * 4) if gathered is not None else 0,
                "face_visibility_ms": vis_ms, "face_visibility_equals_histogram_of_gathered_maps": vis_ok,
                "n_gpu_equals_1_gpu": ok, "checked_views": check, "wall_s": wall,
                "timing": "CUDA events on the rendering stream, max over ranks, barrier on both sides; "
                          "views back to back (no L2 flush between views of a batch)"}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
