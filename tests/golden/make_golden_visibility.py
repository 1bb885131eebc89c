#!/usr/bin/env python
"""Generates tests/golden/visibility_*.npz from the COMPILED REFERENCE (oracle/_ref): keyframes rendered by the
unmodified ugu::Rasterizer, then ugu::VisibilityTester::Test(keyframes, info, facewise) on them
(src/texturing/visibility_tester.cc:670-733; BVH rays), single-threaded (the face-wise mode's bookkeeping is
racy under OpenMP).  Stored: the mesh (or a reference to spot_mesh.npz), the keyframe cameras and images, the
self depth maps (reference Rasterizer, back-face culling off), the tester's camera constants, and the dense
per-keyframe records + CalcStat results for {facewise, not} x {NN, bilinear}.

    python tests/golden/make_golden_visibility.py     (needs /root/reference; run where oracle/_ref builds)
"""
import sys
from pathlib import Path

import numpy as np

REPO = Path(__file__).resolve().parents[2]
for p in (str(REPO), str(REPO / "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import cases  # noqa: E402
import oracle  # noqa: E402
from ugu_b200 import DiffuseColor, PinholeCamera, RendererCpuOption, meshgen  # noqa: E402


def build(name, mesh, mesh_ref, cams, render_opt, ids):
    ref = oracle.OracleRef(mesh)
    optz = RendererCpuOption(diffuse_color=DiffuseColor.kNone, backface_culling=False)
    kfs, zb = [], []
    for kid, cm in zip(ids, cams):
        o = ref.render(cm, render_opt)
        kfs.append((kid, cm, o["color"], o["depth"], o["mask"]))
        zb.append(ref.render(cm, optz, want=("depth",))["depth"])
    nrm = np.asarray(mesh.normals, np.float32)
    assert len(nrm) == len(mesh.vertices)
    d = {"mesh_ref": np.str_(mesh_ref), "ids": np.asarray(ids, np.int32),
         "cam_size": np.array([cams[0].width(), cams[0].height()], np.int32),
         "cam_c2w": np.stack([np.asarray(c.c2w(), np.float64) for c in cams]),
         "cam_pp": np.array(cams[0].principal_point(), np.float32),
         "cam_f": np.array(cams[0].focal_length(), np.float32),
         "color": np.stack([k[2] for k in kfs]), "depth": np.stack([k[3] for k in kfs]),
         "mask": np.stack([k[4] for k in kfs]), "zbuf": np.stack(zb)}
    if not mesh_ref:
        d.update(cases.mesh_to_arrays(mesh))
    consts = [oracle.ref_visibility_camera_constants(c) for c in cams]
    d["ref_w2c"] = np.stack([c[0] for c in consts])
    d["ref_c2w_R"] = np.stack([c[1] for c in consts])
    d["ref_c2w_t"] = np.stack([c[2] for c in consts])
    for facewise in (0, 1):
        for interp in (0, 1):
            r = oracle.ref_visibility(mesh.vertices, nrm, mesh.vertex_indices, kfs, facewise=bool(facewise),
                                      interp=interp, num_threads=1)
            assert r is not None
            for k, v in r.items():
                d[f"ref_fw{facewise}_i{interp}_{k}"] = v
            print(name, "facewise", facewise, "interp", interp, "visible vertices", r["v_visible"].sum(1),
                  "faces", r["f_visible"].sum(1))
    np.savez_compressed(cases.GOLDEN_DIR / f"visibility_{name}.npz", **d)


def main():
    spot = cases.load_spot()
    c, ext = cases.bbox_center_and_extent(spot)
    cams = [PinholeCamera(256, 192, p, fov_y_deg=45.0) for p in meshgen.orbit_poses(c, 1.5 * ext, 5)]
    build("spot_orbit5", spot, "spot_mesh.npz", cams, RendererCpuOption(diffuse_color=DiffuseColor.kTexture),
          [7, 3, 11, 4, 20])
    ico = meshgen.icosphere(4)
    cams = [PinholeCamera(200, 160, p, fov_y_deg=45.0) for p in meshgen.random_sphere_poses(4, seed=3)]
    build("ico4_random4", ico, "", cams, RendererCpuOption(diffuse_color=DiffuseColor.kVertex), [0, 1, 2, 3])


if __name__ == "__main__":
    main()
