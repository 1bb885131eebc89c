#!/usr/bin/env python
"""ugu::Depth2Mesh on ONE GPU (src/util/rgbd_util.cc:94-282): the step ex02 runs on the
rendered depth and colour.  Workload: config 2's view (1.31M-triangle icosphere, vertex colours,
1920x1080) rendered on the device, then the textured overload with vertex colours, x/y step 1.
Device-resident time (depth and colour already in HBM, result left in HBM, L2 flushed between
iterations, CUDA-event pairs on the launching stream), the HBM roofline fraction on algorithmic bytes,
the end-to-end host form (H2D of depth + colour, kernels, D2H of every array), and the compiled
reference on the host when oracle/_ref travelled.  One JSON line.

    python tools/bench_depth2mesh.py [--iters N] [--no-cpu]
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

from ugu_b200 import (DiffuseColor, Image1f, Image3b, Image3f, Mesh, PinholeCamera, Rasterizer,  # noqa: E402
                      RendererCpuOption, meshgen)


def main():
    iters = int(sys.argv[sys.argv.index("--iters") + 1]) if "--iters" in sys.argv else 30
    w, h, th = 1920, 1080, 0.01
    r = Rasterizer(RendererCpuOption(diffuse_color=DiffuseColor.kVertex))
    r.set_mesh(meshgen.icosphere(8))
    assert r.PrepareMesh()
    cam = PinholeCamera(w, h, meshgen.look_at_c2w((0, 0, -2.7), (0, 0, 0), (0, -1, 0)), fov_y_deg=45.0)
    P = w * h
    depth = torch.empty(P, dtype=torch.float32, device="cuda")
    color = torch.empty(P * 3, dtype=torch.uint8, device="cuda")
    flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
    ts = torch.cuda.Stream()
    torch.cuda.set_stream(ts)
    r.render_device(cam, depth=depth.data_ptr(), color=color.data_ptr(), stream=ts.cuda_stream)

    def run():
        r.depth_to_mesh_device(cam, depth.data_ptr(), th, 1, 1, False, color.data_ptr(), w, h, True, ts.cuda_stream)

    for _ in range(3):
        flush.zero_()
        run()
    cnt = r.depth_to_mesh_counts()
    e0 = [torch.cuda.Event(enable_timing=True) for _ in range(iters)]
    e1 = [torch.cuda.Event(enable_timing=True) for _ in range(iters)]
    for i in range(iters):
        flush.zero_()
        e0[i].record()
        run()
        e1[i].record()
    torch.cuda.synchronize()
    ms = sum(a.elapsed_time(b) for a, b in zip(e0, e1)) / iters
    nv, nf = cnt["num_vertices"], cnt["num_faces"]
    # compulsory traffic: depth once, a colour texel per vertex, every output array once
    algo = 4 * P + 3 * nv + nv * (12 + 12 + 8 + 12) + nf * (12 + 12) + 2 * 12 * P
    peak_file = REPO / "MEASURED_PEAKS.json"
    peak = json.loads(peak_file.read_text()).get("hbm_gbs", 6550.0) if peak_file.exists() else 6550.0
    line = {"workload": f"Depth2Mesh (textured + vertex colours, step 1) of config 2's 1920x1080 view: "
                        f"{nv} vertices, {nf} faces", "iters": iters, "ms_per_call": ms,
            "mpix_per_s": P / ms / 1e3, "mtri_per_s": nf / ms / 1e3, "gpu_launches_per_call": 6 if os.environ.get("UGU_B200_D2M_FUSED", "1") == "0" else 4,
            "roofline": {"bound": "hbm", "algorithmic_bytes": algo, "achieved": algo / ms / 1e6, "peak": peak,
                         "unit": "GB/s", "frac": algo / ms / 1e6 / peak},
            "l2": "flushed between timed iterations (512 MiB write)"}
    # end to end: host images in, ugu::Mesh-style arrays + both organized images out
    dimg, cimg = Image1f(), Image3b()
    dimg._init(w, h)
    cimg._init(w, h)
    dimg.data[...] = depth.cpu().numpy().reshape(h, w)
    cimg.data[...] = color.cpu().numpy().reshape(h, w, 3)
    pc, nimg, mesh = Image3f(), Image3f(), Mesh()
    assert r.Depth2Mesh(dimg, cam, mesh, th, 1, 1, False, pc, nimg, cimg, True)
    ne = max(3, min(iters, 10))
    t0 = time.perf_counter()
    for _ in range(ne):
        assert r.Depth2Mesh(dimg, cam, mesh, th, 1, 1, False, pc, nimg, cimg, True)
    line["e2e_ms_per_call"] = (time.perf_counter() - t0) / ne * 1e3
    line["e2e_h2d_bytes"] = 7 * P
    line["e2e_d2h_bytes"] = nv * 44 + nf * 24 + 24 * P
    if "--no-cpu" not in sys.argv:
        import oracle
        if oracle.ref_available():
            t0 = time.perf_counter()
            k = 0
            while k < 3 and time.perf_counter() - t0 < 20.0:
                ref = oracle.ref_depth2mesh(cam, dimg.data, th, 1, 1, False, cimg.data, True)
                k += 1
            line["cpu_reference_ms"] = (time.perf_counter() - t0) / k * 1e3
            line["cpu_cores"] = os.cpu_count()
            line["cpu_note"] = "single-threaded in the reference; includes the shim's array copies"
            same = all(np.array_equal(np.ascontiguousarray(a).view(np.uint8), np.ascontiguousarray(b).view(np.uint8))
                       for a, b in ((mesh.vertices, ref["vertices"]), (mesh.vertex_indices, ref["faces"]),
                                    (mesh.normals, ref["normals"]), (mesh.face_normals, ref["face_normals"]),
                                    (mesh.uv, ref["uv"]), (mesh.vertex_colors, ref["colors"]),
                                    (pc.data, ref["point_cloud"]), (nimg.data, ref["normal_image"])))
            line["matches_reference_bit_for_bit"] = bool(same)
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
