#!/usr/bin/env python
"""bench.py — the headline benchmark of the rasterizer hot path (BASELINE.json metric).

Workload (N = 1): BASELINE.json config 2 — synthetic 1 310 720-triangle subdivided icosphere with
vertex colours, one pinhole view at 1920x1080 per step, all five outputs (colour + depth + normal +
mask + face_id), back-face culling on.  A "step" is one Render of one view per GPU; the camera
orbits the mesh so every step has a different pose (callers mutate the camera between frames,
examples/ex09_renderer_realtime.cc:137-145).  With N > 1 the views shard across ranks with no
data-path collective (weak scaling: one view per rank per step).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

Prints ONE JSON line (see the keys at the end of main()).
"""
from __future__ import annotations

import argparse
import json
import os
import copy
import statistics
import sys
import threading
import time
from pathlib import Path

REPO = Path(__file__).resolve().parent
for p in (str(REPO), str(REPO / "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np  # noqa: E402

W, H = 1920, 1080
ICO_LEVEL = 8
N_POSES = 36
L2_FLUSH_BYTES = 512 << 20
WORKLOAD = "config2: icosphere L8 (V=655362, F=1310720) vertex colours, 1920x1080, 5 outputs"


def measured_peak_gbs():
    p = REPO / "MEASURED_PEAKS.json"
    if p.exists():
        try:
            return float(json.loads(p.read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def make_inputs(level=ICO_LEVEL):
    from ugu_b200 import DiffuseColor, PinholeCamera, RendererCpuOption, meshgen
    mesh = meshgen.icosphere(level)
    opt = RendererCpuOption(diffuse_color=DiffuseColor.kVertex)
    poses = meshgen.orbit_poses((0.0, 0.0, 0.0), 2.7, N_POSES)
    cams = [PinholeCamera(W, H, p, fov_y_deg=45.0) for p in poses]
    return mesh, opt, cams


def algorithmic_bytes(mesh, P):
    """Per-view and per-kernel ALGORITHMIC bytes (SURVEY.md §8d for the view)."""
    V, F = len(mesh.vertices), len(mesh.vertex_indices)
    Nn = len(mesh.normals)
    out = P * (3 + 4 + 12 + 1 + 4)
    view = 12 * V + 12 * F + 12 * F + (12 * Nn + 12 * F) + 12 * V + out
    blocks = (F + 255) // 256
    per_kernel = {
        # fused projection: the "vertex" stage is classify_kernel (block records in, block lists out) and the raster
        # kernel takes the positions in and writes the projected vertices out for the later kernels
        "vertex": 16 * blocks + 4 * blocks,
        "raster": 12 * V + 12 * F + 12 * F + 16 * V,   # positions + indices + face normals in, projected float4 out
        "binned": 0,                               # empty queue at config 2
        "resolve": 8 * P + 8 * P + out,            # visibility read + clear, outputs written once
    }
    return view, per_kernel


class ClockSampler:
    """Samples SM clock / throttle reasons with NVML while the timed regions run."""

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thr = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _run(self):
        nv = self.nv
        names = {"hw_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8),
                 "hw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40),
                 "sw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20),
                 "sw_power_cap": getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4)}
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            time.sleep(0.005)

    def start(self):
        if self.nv:
            self._thr = threading.Thread(target=self._run, daemon=True)
            self._thr.start()

    def stop(self):
        self._stop.set()
        if self._thr:
            self._thr.join()
        return {"sm_mhz": (statistics.median(self.samples) if self.samples else None),
                "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(self.samples)}


def time_reference(mesh, opt, cams, steps, warmup, budget_s=None, split=False):
    """The reference's own CPU implementation: oracle/_ref (compiled reference) when it travelled,
    else the plain-C oracle port.  One step = camera->set_c2w + Render, as callers do.  With `split`
    (compiled reference only) a few extra steps separate what a step is made of: the Render
    itself with the ray table warm, the per-pose ray-table rebuild (include/ugu/camera.h:482-513, the only
    OpenMP-parallel part), and the same with OMP_NUM_THREADS = 1."""
    # all the host threads the reference can use (torchrun exports OMP_NUM_THREADS=1 by default);
    # must be set before libgomp is loaded with the reference library
    os.environ["OMP_NUM_THREADS"] = str(os.cpu_count() or 1)
    import oracle
    extra = {}
    if oracle.ref_available():
        kind = "reference"
        runner = oracle.OracleRef(mesh)
        runner.set_num_threads(os.cpu_count() or 1)
        render = lambda cam: runner.render(cam, opt)
        cores = os.cpu_count() or 1
        note = ("unmodified ugu::Rasterizer; its Render loop is single-threaded, only the per-pose "
                "ray-table rebuild (camera.h:482-513) uses OpenMP on all cores")
    else:
        kind = "port"
        runner = oracle.OraclePort(mesh)
        render = lambda cam: runner.render(cam.to_native(), opt)
        cores = 1
        note = "plain-C scalar restatement (oracle/raster_oracle.c), one thread"
    for i in range(warmup):
        render(cams[i % len(cams)])
    t0 = time.perf_counter()
    done = 0
    for i in range(steps):
        render(cams[(warmup + i) % len(cams)])
        done += 1
        if budget_s is not None and time.perf_counter() - t0 > budget_s:
            break
    dt = time.perf_counter() - t0
    if split and kind == "reference":
        def cold_warm(n):
            cold, warm = [], []
            for i in range(n):
                runner.render(cams[(7 + i) % len(cams)], opt, repeat_warm=True)
                cold.append(runner.last_timing_ms[0])
                warm.append(runner.last_timing_ms[1])
            return statistics.median(cold), statistics.median(warm)
        c_all, w_all = cold_warm(3)
        runner.set_num_threads(1)
        c_one, w_one = cold_warm(3)
        runner.set_num_threads(os.cpu_count() or 1)
        extra = {"render_ms_ray_table_warm": w_all, "ray_table_rebuild_ms": max(0.0, c_all - w_all),
                 "omp_num_threads_1": {"ms_per_step": c_one, "render_ms_ray_table_warm": w_one,
                                       "ray_table_rebuild_ms": max(0.0, c_one - w_one)},
                 "split_sample": "median of 3 views each (set_c2w + Render, then Render again with the camera untouched)"}
    return dict(kind=kind, cores=cores, steps=done, seconds=dt, note=note, extra=extra)


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    mesh, opt, cams = make_inputs()
    steps = max(1, args.steps)
    r = time_reference(mesh, opt, cams, steps, max(0, min(args.warmup, 2)), budget_s=240.0, split=True)
    P = W * H
    F = len(mesh.vertex_indices)
    ms = r["seconds"] / r["steps"] * 1e3
    value = P * r["steps"] / r["seconds"] / 1e6
    line = {
        "impl": "reference", "metric": "Mpix/s (color+depth+normal+mask+face_id)", "value": value,
        "unit": "Mpix/s", "mtri_per_s": F * r["steps"] / r["seconds"] / 1e6, "n_gpus": args.gpus,
        "steps": r["steps"], "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "views_per_step": 1, "culling": True},
        "cpu_baseline": {"value": value, "unit": "Mpix/s", "cores": r["cores"], "kind": r["kind"],
                         "sample": f"{r['steps']} full views of the workload (set_c2w + Render each)",
                         "note": r["note"], **r["extra"]},
        "e2e": {"value": value, "unit": "Mpix/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


def multiview_records(world, rank, r_ico, ico_mesh, barrier):
    """BASELINE configs 5 and 3 — the configs the multi-GPU target is stated on — as STRONG scaling through the
    batched multi-view path (ugu_b200.multiview.MultiViewRenderer -> ugu_b200_rasterizer_render_batch): a fixed
    view list is split contiguously over the ranks; timed are (a) render only, outputs left on each GPU, (b)
    render + gather of every face_id map to rank 0 with the exchange overlapped (chunks sent as they finish,
    straight into the view-ordered destination), (c) config 5 only: render + per-GPU fold of the maps into per-face
    visibility statistics + an all-reduce of 16 bytes per face instead of the maps.  Device time between barriers,
    max over ranks (CUDA events on the current stream, which the library's work stream joins), median of 3 after
    one warm-up.  L2 is flushed before each measurement, not between the views of a batch (a batch is one unit of
    work; the views of a rank reuse the mesh in L2 by design)."""
    import torch
    import torch.distributed as dist
    import cases
    from ugu_b200 import DiffuseColor, PinholeCamera, Rasterizer, RendererCpuOption, meshgen
    from ugu_b200.multiview import MultiViewRenderer

    flush = torch.empty(L2_FLUSH_BYTES, dtype=torch.uint8, device="cuda")

    def timed(fn, reps=3):
        out = []
        for i in range(reps + 1):
            flush.zero_()
            barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            res = fn()
            e1.record()
            barrier()
            t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            if i:
                out.append(float(t[0]))
            del res
        return statistics.median(out)

    records = {}
    # ---- config 5: 1024 random views of the 1.31M-triangle icosphere at 512x512, face_id only -------------------
    n5 = 1024
    cams5 = [PinholeCamera(512, 512, p, fov_y_deg=45.0) for p in meshgen.random_sphere_poses(n5, seed=0)]
    wrap5 = [c.to_native() for c in cams5]   # camera blocks built once, as a caller looping over a fixed rig would
    saved = copy.copy(r_ico._option)
    r_ico.set_option(RendererCpuOption(diffuse_color=DiffuseColor.kNone))
    mv = MultiViewRenderer(r_ico)
    F = len(ico_mesh.vertex_indices)
    P5 = 512 * 512
    t_render = timed(lambda: mv.render(wrap5, ("face_id",)))
    t_gather = timed(lambda: mv.render_and_gather_overlapped(wrap5, "face_id", dst=0, chunk_views=16))
    t_fold = timed(lambda: mv.render_face_visibility(wrap5, F, chunk_views=32))
    gbytes = (n5 - len(mv.local_indices(n5)) if rank == 0 else 0) * P5 * 4
    records["config5"] = {
        "workload": "config5: icosphere L8 (F=1310720), 1024 random views, 512x512, face_id only",
        "scaling": "strong", "views": n5, "n_gpus": world,
        "render": {"ms": t_render, "views_per_s": n5 / (t_render * 1e-3), "us_per_view_per_gpu": t_render * 1e3 / (n5 / world),
                   "mpix_per_s": n5 * P5 / (t_render * 1e-3) / 1e6, "mtri_per_s": n5 * F / (t_render * 1e-3) / 1e6},
        "render_gather_to_rank0": {"ms": t_gather, "views_per_s": n5 / (t_gather * 1e-3),
                                   "bytes_into_root": gbytes,
                                   "exposed_gather_ms": max(0.0, t_gather - t_render),
                                   "note": "NCCL send/recv per 16-view chunk as it finishes, directly into the "
                                           "view-ordered buffer on rank 0"},
        "render_fold_allreduce": {"ms": t_fold, "views_per_s": n5 / (t_fold * 1e-3),
                                  "note": "per-GPU fold into per-face pixel / view counts and best view, then "
                                          "all-reduce of 16 B per face"},
        "algorithmic_bytes_per_view": 12 * len(ico_mesh.vertices) + 24 * F + 4 * P5,
    }
    r_ico.set_option(saved)
    del mv
    # ---- config 3: spot, 36-view turntable at 1024x1024, all five outputs ---------------------------------------
    spot = cases.load_spot()
    r3 = Rasterizer(RendererCpuOption(diffuse_color=DiffuseColor.kTexture), device=torch.cuda.current_device())
    r3.set_mesh(spot)
    assert r3.PrepareMesh()
    c, ext = cases.bbox_center_and_extent(spot)
    cams3 = [PinholeCamera(1024, 1024, p, fov_y_deg=45.0).to_native() for p in meshgen.orbit_poses(c, 1.5 * ext, 36)]
    mv3 = MultiViewRenderer(r3)
    outs = ("color", "depth", "normal", "mask", "face_id")
    t3 = timed(lambda: mv3.render(cams3, outs))
    t3g = timed(lambda: mv3.render_and_gather_overlapped(cams3, "face_id", dst=0, chunk_views=2))
    P3 = 1024 * 1024
    records["config3"] = {
        "workload": "config3: spot (F=5856) 36-view turntable, 1024x1024, texture bilinear, 5 outputs",
        "scaling": "strong", "views": 36, "n_gpus": world,
        "render": {"ms": t3, "views_per_s": 36 / (t3 * 1e-3), "us_per_view_per_gpu": t3 * 1e3 / -(-36 // world),
                   "mpix_per_s": 36 * P3 / (t3 * 1e-3) / 1e6},
        "render_gather_to_rank0": {"ms": t3g, "views_per_s": 36 / (t3g * 1e-3),
                                   "note": "face_id maps only are rendered and gathered (2-view chunks)"},
    }
    return records


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=300)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ugu_b200", choices=["ugu_b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-multiview", action="store_true", help="skip the config 3 / config 5 multi-view records")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference_arm(args)

    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a B200: ugu_b200 has no CPU fallback")
    torch.cuda.set_device(local)
    # host cores / pinned memory local to this rank's GPU (matters for e2e at N > 2: the D2H of N ranks)
    from ugu_b200.multiview import bind_rank_to_gpu_numa
    numa = bind_rank_to_gpu_numa(local) if world > 1 else {"bound": False, "note": "single rank: not bound"}
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    warmup = max(3, args.warmup)
    K = max(1, args.steps)

    from ugu_b200 import Image1b, Image1f, Image1i, Image3b, Image3f, Rasterizer
    from ugu_b200 import _native as N

    mesh, opt, cams = make_inputs()
    P = W * H
    V, F = len(mesh.vertices), len(mesh.vertex_indices)
    r = Rasterizer(opt, device=local)
    r.set_mesh(mesh)
    assert r.PrepareMesh()
    natives = [c.to_native() for c in cams]

    dev = dict(color=torch.empty(P * 3, dtype=torch.uint8, device="cuda"),
               depth=torch.empty(P, dtype=torch.float32, device="cuda"),
               normal=torch.empty(P * 3, dtype=torch.float32, device="cuda"),
               mask=torch.empty(P, dtype=torch.uint8, device="cuda"),
               face_id=torch.empty(P, dtype=torch.int32, device="cuda"))
    ptrs = {k: v.data_ptr() for k, v in dev.items()}
    flush = torch.empty(L2_FLUSH_BYTES, dtype=torch.uint8, device="cuda")
    # A real (non-default) stream: the kernels, the L2 flush and the timing events must all be on
    # ONE stream.  (Handle 0 would mean "the rasterizer's own stream" to the C-ABI, and torch events
    # recorded on the default stream would then not bracket the kernels at all.)
    tstream = torch.cuda.Stream()
    torch.cuda.set_stream(tstream)
    stream = tstream.cuda_stream
    assert stream != 0

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def device_step(i):
        # each rank takes its own view of the step (views shard across ranks, no collective)
        r.render_device(natives[(i * world + rank) % N_POSES], stream=stream, **ptrs)

    # ---- device-resident loop: inputs already in HBM, outputs left in HBM ------------------------
    for i in range(warmup):
        flush.zero_()
        device_step(i)
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    no_flush = os.environ.get("UGU_BENCH_NOFLUSH") == "1"   # diagnostic only: warm-L2 numbers are NOT bench values

    def timed_pass():
        ev0 = [torch.cuda.Event(enable_timing=True) for _ in range(K)]
        ev1 = [torch.cuda.Event(enable_timing=True) for _ in range(K)]
        barrier()
        w0 = time.perf_counter()
        for i in range(K):
            if not no_flush:
                flush.zero_()      # L2 flush between timed iterations (outside the per-step events)
            ev0[i].record()
            device_step(warmup + i)
            ev1[i].record()
        barrier()
        per = [a.elapsed_time(b) for a, b in zip(ev0, ev1)]
        return sum(per), time.perf_counter() - w0, per

    # pass 1 — the timed K steps behind `value`: nothing but the view's kernels between the events
    r.reset_stats()
    dev_ms, wall_dev, per_step = timed_pass()
    launches = r.stats()["kernel_launches"]
    # pass 2 — the same K steps again with the library's per-stage CUDA-event pairs switched on: the
    # per-kernel durations behind `roofline` / `stage_us` (the extra event records serialise the
    # kernels that pass 1 launches back to back, so this pass is slower and is not the bench value)
    r.reset_stats()
    r.set_profiling(True, (0, 1, 2, 3))
    prof_ms, _, _ = timed_pass()
    st = r.stats()
    r.set_profiling(False, ())

    # ---- end-to-end loop: the public Render call with HOST (pinned) images ------------------------
    imgs = (Image3b(), Image1f(), Image3f(), Image1b(), Image1i())
    r.set_camera(cams[0])
    for i in range(min(warmup, 5)):
        assert r.Render(*imgs)
    barrier()
    r.reset_stats()
    e2e_t = 0.0
    Ke = max(1, min(K, 200))
    for i in range(Ke):
        flush.zero_()
        torch.cuda.synchronize()
        cam = cams[((warmup + i) * world + rank) % N_POSES]
        t0 = time.perf_counter()
        r.set_camera(cam)         # the per-step input: the camera pose -> H2D as kernel constants
        ok = r.Render(*imgs)      # validates, renders, D2H of all five outputs, returns when landed
        e2e_t += time.perf_counter() - t0
        assert ok
    st_e = r.stats()
    barrier()
    # the same Render into PAGEABLE destinations — what ugu::Image really is (a std::vector, image.h:354): the
    # library stages the D2H through its pinned ring and memcpy's into the caller's images
    pimgs = (Image3b(pinned=False), Image1f(pinned=False), Image3f(pinned=False), Image1b(pinned=False),
             Image1i(pinned=False))
    for i in range(2):
        assert r.Render(*pimgs)
    e2e_pg = []
    for i in range(max(1, min(K, 50))):
        flush.zero_()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        r.set_camera(cams[((warmup + i) * world + rank) % N_POSES])
        ok = r.Render(*pimgs)
        e2e_pg.append(time.perf_counter() - t0)
        assert ok
    barrier()
    mv = multiview_records(world, rank, r, mesh, barrier) if not args.no_multiview else None
    clocks = sampler.stop()
    # context for e2e: the host link.  One pinned D2H copy of the same number of bytes a view returns.
    d2h_bytes = P * 24
    link_src = torch.empty(d2h_bytes, dtype=torch.uint8, device="cuda")
    link_dst = torch.empty(d2h_bytes, dtype=torch.uint8, pin_memory=True)
    link_dst.copy_(link_src, non_blocking=True)
    torch.cuda.synchronize()
    le0, le1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    le0.record()
    for _ in range(10):
        link_dst.copy_(link_src, non_blocking=True)
    le1.record()
    torch.cuda.synchronize()
    link_ms = le0.elapsed_time(le1) / 10

    # ---- max over ranks ---------------------------------------------------------------------------
    t = torch.tensor([dev_ms, e2e_t * 1e3, sum(e2e_pg) * 1e3], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms_max, e2e_ms_max, e2e_pg_ms_max = float(t[0]), float(t[1]), float(t[2])

    if rank == 0:
        peak, peak_src = measured_peak_gbs()
        view_bytes, kbytes = algorithmic_bytes(mesh, P)
        stage_ms = {k: v / max(1, st["timed_views"]) for k, v in st["stage_ms"].items()}
        dom = max(stage_ms, key=lambda k: stage_ms[k])
        achieved = kbytes[dom] / (stage_ms[dom] * 1e-3) / 1e9 if stage_ms[dom] > 0 else 0.0
        try:   # DRAM bytes per launch of each kernel from the committed ncu --set full capture
            traffic = json.loads((REPO / "profiles" / "traffic.json").read_text())
        except Exception:
            traffic = {}
        ms_per_step = dev_ms_max / K
        value = world * P * K / (dev_ms_max * 1e-3) / 1e6
        e2e_value = world * P * Ke / (e2e_ms_max * 1e-3) / 1e6
        line = {
            "metric": "Mpix/s (color+depth+normal+mask+face_id)", "value": value, "unit": "Mpix/s",
            "mtri_per_s": world * F * K / (dev_ms_max * 1e-3) / 1e6,
            "n_gpus": world, "steps": K, "warmup": warmup, "ms_per_step": ms_per_step,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": WORKLOAD, "views_per_step_per_gpu": 1, "culling": True,
                       "l2": ("NOT FLUSHED (diagnostic run, not a bench value)" if no_flush else
                              f"flushed between timed iterations ({L2_FLUSH_BYTES >> 20} MiB write)"),
                       "timing": "sum of per-step CUDA-event pairs on the launching stream, max over ranks",
                       "parallelism": f"views sharded over {world} GPU(s), no collective"},
            "e2e": {"value": e2e_value, "unit": "Mpix/s", "ms_per_step": e2e_ms_max / Ke, "steps": Ke,
                    "h2d_bytes_per_step": int(st_e["h2d_bytes"] // max(1, st_e["views"])),
                    "d2h_bytes_per_step": int(st_e["d2h_bytes"] // max(1, st_e["views"])),
                    "api": "ugu_b200.Rasterizer.Render -> ugu_b200_rasterizer_render, pinned host images",
                    "host_link": {"pinned_d2h_gbs": d2h_bytes / (link_ms * 1e-3) / 1e9,
                                  "ms_for_one_view_of_outputs": link_ms,
                                  "note": "a view returns 24 B/pixel; e2e is bounded by this copy; at N > 1 "
                                          "all ranks copy at the same time and share the host's memory system"},
                    "pageable": {"value": world * P * len(e2e_pg) / (e2e_pg_ms_max * 1e-3) / 1e6, "unit": "Mpix/s",
                                 "ms_per_step": e2e_pg_ms_max / len(e2e_pg), "steps": len(e2e_pg),
                                 "note": "same call, pageable destinations (the layout of ugu::Image, image.h:354): "
                                         "D2H through the library's pinned staging ring, drained by a few host threads "
                                         "(UGU_B200_COPY_THREADS, default min(6, 3/8 of the host threads))"},
                    "numa_rank0": numa},
            "gpu_launches": int(launches),
            "ms_per_step_spread": {"min": min(per_step), "median": statistics.median(per_step), "max": max(per_step),
                                   "note": "rank 0's per-step CUDA-event times of the timed pass"},
            "stage_us": {k: round(v * 1e3, 2) for k, v in stage_ms.items()},
            "ms_per_step_stage_timed_pass": prof_ms / K,
            "roofline": {"bound": "hbm", "kernel": dom + "_kernel", "achieved": achieved, "peak": peak,
                         "unit": "GB/s", "frac": achieved / peak, "traffic": traffic.get(dom + "_kernel"),
                         "traffic_source": traffic.get("source"),
                         "algorithmic_bytes_per_launch": kbytes[dom], "peak_source": peak_src,
                         "view": {"algorithmic_bytes": view_bytes,
                                  "achieved": view_bytes / (ms_per_step * 1e-3) / 1e9,
                                  "frac": view_bytes / (ms_per_step * 1e-3) / 1e9 / peak}},
            "clocks": clocks,
            "wall_s_device_loop": wall_dev,
        }
        if mv is not None:
            line["multiview"] = mv
        if world == 1 and not args.no_cpu_baseline:
            cb = time_reference(mesh, opt, cams, steps=24, warmup=1, budget_s=20.0, split=True)
            line["cpu_baseline"] = {
                "value": P * cb["steps"] / cb["seconds"] / 1e6, "unit": "Mpix/s", "cores": cb["cores"],
                "kind": cb["kind"], "ms_per_step": cb["seconds"] / cb["steps"] * 1e3,
                "sample": f"{cb['steps']} full views of the same workload (set_c2w + Render each)",
                "note": cb["note"], **cb["extra"]}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
