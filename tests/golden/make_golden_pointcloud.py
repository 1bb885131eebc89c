"""Generates tests/golden/pointcloud_*.npz from the COMPILED, UNMODIFIED reference's
ugu::Depth2PointCloud (src/util/rgbd_util.cc:472-513) through oracle/_ref.  Run only where
/root/reference exists:

    make -C oracle ref && python tests/golden/make_golden_pointcloud.py

Inputs are the reference-rendered depth images of existing golden cases; each file holds the depth,
the camera and the reference outputs computed over a non-zero "previous" point-cloud image
(cases.pointcloud_prev; the reference leaves pixels without depth untouched) for gl_coord = false / true, plus the fresh-image output.
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
from oracle import ref_depth2pointcloud  # noqa: E402

OUT = Path(__file__).resolve().parent

for name in ("spot_front_tex_bilinear", "ico4_vertex_cull", "kat_large_and_small_nocull"):
    _, cam, _, want, _ = util.load_golden(name)
    depth = want["depth"]
    prev = cases.pointcloud_prev(depth.shape)
    d = dict(depth=depth)
    d.update(cases.camera_to_arrays(cam))
    d["pc_gl0"] = ref_depth2pointcloud(cam, depth, prev, False)
    d["pc_gl1"] = ref_depth2pointcloud(cam, depth, prev, True)
    d["pc_fresh_gl0"] = ref_depth2pointcloud(cam, depth, None, False)
    np.savez_compressed(OUT / f"pointcloud_{name}.npz", **d)
    print(name, depth.shape, "valid", int((depth > 0).sum()))
