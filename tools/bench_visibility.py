#!/usr/bin/env python
"""The keyframe visibility pass (ugu::VisibilityTester; src/texturing/visibility_tester.cc:416-733)
on ONE GPU.  Workload: config 2's mesh (1.31M-triangle icosphere, vertex colours, vertex normals) seen by 16
keyframes at 1024x1024 on random sphere poses (config 5's pose generator); the keyframes' colour / depth / mask
images and the self depth maps are rendered on the device by the hot path and stay in HBM.  Timed: one
ugu_b200_visibility_test_device call over the 16 keyframes (fold into the accumulators, no dense records), CUDA-event
pairs on the launching stream, L2 flushed between iterations; a second figure lets the library render the self depth
maps itself.  Roofline: algorithmic bytes = per keyframe the vertex positions + normals (24 B) and the indices
(12 B per face) read once, plus one read-modify-write of the accumulators per call.  With oracle/_ref present, the
compiled reference tester on the host for a bounded sample of keyframes.  One JSON line.

    python tools/bench_visibility.py [--iters N] [--cpu] [--cpu-keyframes K]

(--cpu: the reference tester takes MINUTES per keyframe at this size — 408 s measured — so it is off by default.)
"""
from __future__ import annotations

import ctypes as C
import json
import sys
import time
from pathlib import Path

REPO = Path(__file__).resolve().parents[1]
for p in (str(REPO), str(REPO / "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np  # noqa: E402
import torch  # noqa: E402

from ugu_b200 import DiffuseColor, PinholeCamera, Rasterizer, RendererCpuOption, meshgen  # noqa: E402
from ugu_b200 import _native as N  # noqa: E402
from ugu_b200.visibility import Keyframe, VisibilityTester, native_keyframe  # noqa: E402


def arg(name, default):
    return type(default)(sys.argv[sys.argv.index(name) + 1]) if name in sys.argv else default


def main():
    iters, K, W = arg("--iters", 20), 16, 1024
    mesh = meshgen.icosphere(8)
    V, F, P = len(mesh.vertices), len(mesh.vertex_indices), W * W
    r = Rasterizer(RendererCpuOption(diffuse_color=DiffuseColor.kVertex, backface_culling=False))
    r.set_mesh(mesh)
    assert r.PrepareMesh()
    cams = [PinholeCamera(W, W, p, fov_y_deg=45.0) for p in meshgen.random_sphere_poses(K, seed=3)]
    col = torch.empty(K * P * 3, dtype=torch.uint8, device="cuda")
    dep = torch.empty(K * P, dtype=torch.float32, device="cuda")
    msk = torch.empty(K * P, dtype=torch.uint8, device="cuda")
    flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
    ts = torch.cuda.Stream()
    torch.cuda.set_stream(ts)
    torch.cuda.synchronize()
    # culling off: the depth image is also the self depth map (zbuf) of the occlusion test
    r.render_batch(cams, color=col.data_ptr(), depth=dep.data_ptr(), mask=msk.data_ptr(), on_device=True,
                   stream=ts.cuda_stream)
    t = VisibilityTester(rasterizer=r)
    t.set_data(mesh.vertices, mesh.normals, mesh.vertex_indices)
    assert t.PrepareData()
    lib = N.load()
    opt = t.native_option(True)

    def keyframes(with_zbuf):
        keep = []
        return (N.Keyframe * K)(*[native_keyframe(
            Keyframe(id=k, camera=cams[k], color=col.data_ptr() + k * P * 3, depth=dep.data_ptr() + k * P * 4,
                     mask=msk.data_ptr() + k * P, zbuf=(dep.data_ptr() + k * P * 4) if with_zbuf else None),
            keep, images_on_device=True) for k in range(K)])

    def timed(kfs):
        ms = []
        for i in range(iters + 3):
            lib.ugu_b200_visibility_reset(r._h)
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            rc = lib.ugu_b200_visibility_test_device(r._h, K, kfs, 0, C.byref(opt), None, C.c_void_p(ts.cuda_stream))
            assert rc == N.OK, N.last_error(r._h)
            e1.record()
            e1.synchronize()
            if i >= 3:
                ms.append(e0.elapsed_time(e1))
        return float(np.median(ms))

    ms_given = timed(keyframes(True))
    stats = t.stats(list(range(K)))
    ms_rendered = timed(keyframes(False))
    peak_file = REPO / "MEASURED_PEAKS.json"
    peak = json.loads(peak_file.read_text()).get("hbm_gbs", 6550.0) if peak_file.exists() else 6550.0
    acc_bytes = F * (4 + 8 + 8 + 8) + V * (4 + 16 + 8 + 8 + 8 + 8)
    alg = K * (24 * V + 12 * F) + 2 * acc_bytes
    out = {"workload": f"visibility pass: icosphere V={V} F={F}, {K} keyframes {W}x{W}, fold only, facewise",
           "iters": iters, "ms_per_call": ms_given, "us_per_keyframe": ms_given / K * 1e3,
           "mvertex_keyframes_per_s": V * K / ms_given / 1e3,
           "ms_per_call_self_depth_rendered_by_the_library": ms_rendered,
           "algorithmic_bytes_per_call": alg,
           "roofline": {"bound": "hbm", "achieved": alg / (ms_given * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                        "frac": alg / (ms_given * 1e-3) / 1e9 / peak},
           "faces_seen": int((stats["face_count"] > 0).sum()), "vertices_seen": int((stats["vertex_count"] > 0).sum()),
           "l2": "flushed between timed iterations (512 MiB write)"}
    if "--cpu" in sys.argv:
        try:
            import oracle
            if oracle.ref_available():
                kc = arg("--cpu-keyframes", 1)
                P1 = W * W
                host = [(k, cams[k], col[k * P1 * 3:(k + 1) * P1 * 3].cpu().numpy().reshape(W, W, 3),
                         dep[k * P1:(k + 1) * P1].cpu().numpy().reshape(W, W),
                         msk[k * P1:(k + 1) * P1].cpu().numpy().reshape(W, W)) for k in range(kc)]
                t0 = time.perf_counter()
                ref = oracle.ref_visibility(mesh.vertices, mesh.normals, mesh.vertex_indices, host)
                dt = time.perf_counter() - t0
                assert ref is not None
                out["cpu_baseline"] = {"kind": "reference", "ms_per_keyframe": dt / kc * 1e3, "keyframes": kc,
                                       "cores": 1, "sample": f"{kc} of the {K} keyframes (set_data + PrepareData's BVH "
                                       "build + Test + CalcStat, compiled ugu::VisibilityTester)",
                                       "speedup_per_keyframe": dt / kc * 1e3 / (ms_given / K)}
        except Exception as e:  # the baseline is optional
            out["cpu_baseline"] = {"error": repr(e)[:200]}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
