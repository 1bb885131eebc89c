"""Generates tests/golden/d2m2cam_*.npz from the COMPILED, UNMODIFIED reference's two-camera ugu::Depth2Mesh
(include/ugu/util/rgbd_util.h:39-48; Depth2MeshImpl, src/util/rgbd_util.cc:283-466) through oracle/_ref.
Run only where /root/reference exists:

    make -C oracle ref && python tests/golden/make_golden_depth2mesh2.py

Depth = the reference-rendered depth of an existing golden case; the colour image is procedural
(tests/cases.py: procedural_color) and belongs to a second, wider pinhole camera of another size; each file
stores the parameters, depth2color and every output array.
"""
from __future__ import annotations

import sys
from pathlib import Path

import numpy as np

REPO = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(REPO))
sys.path.insert(0, str(REPO / "tests"))

import cases  # noqa: E402
import util  # noqa: E402
from oracle import ref_depth2mesh2  # noqa: E402

OUT = Path(__file__).resolve().parent

# (case, x_step, y_step, gl_coord, max_connect_z_diff, with_vertex_color, colour size, colour fov_y, variant of depth2color)
CASES = [
    ("spot_front_tex_bilinear", 1, 1, False, 0.05, True, (200, 160), 70.0, 0),
    ("ico4_vertex_cull", 2, 1, False, 0.05, False, (333, 251), 80.0, 1),
    ("kat_large_and_small_nocull", 1, 3, False, 1e9, True, (64, 64), 100.0, 2),
]

for i, (name, xs, ys, gl, th, wvc, (cw, ch), fov, var) in enumerate(CASES):
    _, cam, _, want, _ = util.load_golden(name)
    ccam, d2c, color = cases.two_camera_setup(cw, ch, fov, var)
    r = ref_depth2mesh2(cam, ccam, d2c, want["depth"], color, th, xs, ys, gl, wvc)
    d = {f"out_{k}": v for k, v in r.items()}
    d.update(case=np.str_(name), x_step=np.int32(xs), y_step=np.int32(ys), gl_coord=np.int32(gl),
             max_connect_z_diff=np.float32(th), with_vertex_color=np.int32(wvc), cw=np.int32(cw), ch=np.int32(ch),
             fov=np.float32(fov), variant=np.int32(var), depth2color=d2c)
    np.savez_compressed(OUT / f"d2m2cam_{i}_{name}.npz", **d)
    print(i, name, "nv", len(r["vertices"]), "nf", len(r["faces"]), "uv range", r["uv"].min(0), r["uv"].max(0))
