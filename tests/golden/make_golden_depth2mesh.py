"""Generates tests/golden/depth2mesh_*.npz from the COMPILED, UNMODIFIED reference's ugu::Depth2Mesh
(src/util/rgbd_util.cc:94-282,528-545; Mesh::CalcNormal, src/mesh.cc:297-349) through oracle/_ref.
Run only where /root/reference exists:

    make -C oracle ref && python tests/golden/make_golden_depth2mesh.py

Inputs are the reference-rendered depth / colour images of existing golden cases (golden_<case>.npz);
each file stores the parameters and every output array.
"""
from __future__ import annotations

import sys
from pathlib import Path

import numpy as np

REPO = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(REPO))
sys.path.insert(0, str(REPO / "tests"))

import util  # noqa: E402
from oracle import ref_depth2mesh  # noqa: E402

OUT = Path(__file__).resolve().parent

# (case, x_step, y_step, gl_coord, max_connect_z_diff, textured, with_vertex_color)
CASES = [
    ("spot_front_tex_bilinear", 1, 1, False, 0.05, False, False),
    ("spot_front_tex_bilinear", 2, 3, True, 0.5, True, True),
    ("ico4_vertex_cull", 1, 1, True, 0.05, True, False),
    ("ico4_vertex_cull", 3, 2, False, 0.002, False, False),
    ("kat_large_and_small_nocull", 1, 2, False, 1e9, True, True),
    ("kat_large_and_small_nocull", 1, 1, False, 0.0, False, False),   # nothing connects: no faces, no normals
]

for i, (name, xs, ys, gl, th, tex, wvc) in enumerate(CASES):
    _, cam, _, want, _ = util.load_golden(name)
    r = ref_depth2mesh(cam, want["depth"], th, xs, ys, gl, want["color"] if tex else None, wvc)
    d = {f"out_{k}": v for k, v in r.items()}
    d.update(case=np.str_(name), x_step=np.int32(xs), y_step=np.int32(ys), gl_coord=np.int32(gl),
             max_connect_z_diff=np.float32(th), textured=np.int32(tex), with_vertex_color=np.int32(wvc))
    np.savez_compressed(OUT / f"depth2mesh_{i}_{name}.npz", **d)
    print(i, name, "nv", len(r["vertices"]), "nf", len(r["faces"]), "normals", len(r["normals"]))
