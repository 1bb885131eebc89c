"""Generates tests/golden/*.npz from the COMPILED, UNMODIFIED reference (oracle/_ref/libugu_ref.so,
built from /root/reference by oracle/Makefile).  Run only where /root/reference exists:

    make -C oracle ref && python tests/golden/make_golden.py

Each golden_<name>.npz is self-contained: the mesh arrays, camera (size, c2w, intrinsics), option,
the reference's float camera constants (w2c_R, w2c_t, c2w_R as rasterizer.cc:118-119 /
camera.h:321 derive them with Eigen) and the five reference output images.  spot_mesh.npz is
data/spot as loaded by the reference's own Mesh::LoadObj (the model is Keenan Crane's CC0 "spot").
"""
from __future__ import annotations

import sys
from pathlib import Path

import numpy as np

REPO = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(REPO))
sys.path.insert(0, str(REPO / "tests"))

import cases  # noqa: E402
from oracle import OracleRef  # noqa: E402
from ugu_b200 import (ColorInterpolation, DiffuseColor, DiffuseShading, Material, Mesh,  # noqa: E402
                      PinholeCamera, RendererCpuOption, ShadingNormal)
from ugu_b200 import meshgen  # noqa: E402

OUT = Path(__file__).resolve().parent
SPOT_DIR = "/root/reference/data/spot/"


def save_case(name, mesh, camera, option, store_mesh=True, mesh_ref=""):
    ref = OracleRef(mesh)
    out = ref.render(camera, option)
    assert out is not None, name
    R, t, C = ref.camera_constants(camera.c2w())
    d = {}
    if store_mesh:
        d.update(cases.mesh_to_arrays(mesh))
    d["mesh_ref"] = np.array(mesh_ref)
    d.update(cases.camera_to_arrays(camera))
    d["option"] = cases.option_to_array(option)
    d["ref_w2c_R"], d["ref_w2c_t"], d["ref_c2w_R"] = R, t, C
    for k in cases.OUTPUT_KEYS:
        d[f"out_{k}"] = out[k]
    np.savez_compressed(OUT / f"golden_{name}.npz", **d)
    print(f"golden_{name}: covered={(out['face_id'] >= 0).sum()} mask={(out['mask'] > 0).sum()}")


def main():
    loader = OracleRef()
    a = loader.load_obj(SPOT_DIR + "spot_triangulated.obj", SPOT_DIR)
    spot = Mesh(vertices=a["vertices"], vertex_colors=a["vertex_colors"],
                vertex_indices=a["vertex_indices"], normals=a["normals"],
                face_normals=a["face_normals"], normal_indices=a["normal_indices"], uv=a["uv"],
                uv_indices=a["uv_indices"], material_ids=a["material_ids"],
                materials=[Material(diffuse_tex=a["tex0"])])
    print("spot:", {k: v.shape for k, v in a.items() if hasattr(v, "shape")})
    np.savez_compressed(OUT / "spot_mesh.npz", **cases.mesh_to_arrays(spot))

    # --- spot, ex02 intrinsics scaled to 160x120, poses front and top (ex02_renderer.cc:49-102)
    poses = cases.ex02_poses(spot, loader.lookat_c2w)
    np.save(OUT / "spot_ex02_poses.npy", np.stack(poses))
    r = 0.25
    for idx, tag in ((0, "front"), (4, "top")):
        cam = PinholeCamera(int(640 * r), int(480 * r), poses[idx],
                            principal_point=(np.float32(318.6) * np.float32(r), np.float32(255.3) * np.float32(r)),
                            focal_length=(np.float32(517.3) * np.float32(r), np.float32(516.5) * np.float32(r)))
        save_case(f"spot_{tag}_tex_bilinear", spot, cam,
                  RendererCpuOption(diffuse_color=DiffuseColor.kTexture), store_mesh=False,
                  mesh_ref="spot_mesh.npz")
    cam = PinholeCamera(160, 120, poses[2], fov_y_deg=30.0)
    save_case("spot_right_tex_nn_lambert_nocull", spot, cam,
              RendererCpuOption(diffuse_color=DiffuseColor.kTexture, interp=ColorInterpolation.kNn,
                                diffuse_shading=DiffuseShading.kLambertian, backface_culling=False),
              store_mesh=False, mesh_ref="spot_mesh.npz")

    # --- micro known-answer cases
    kcam = cases.kat_camera()
    for name, mesh in cases.kat_meshes().items():
        save_case(f"kat_{name}_cull", mesh, kcam, RendererCpuOption(diffuse_color=DiffuseColor.kVertex))
        save_case(f"kat_{name}_nocull", mesh, kcam,
                  RendererCpuOption(diffuse_color=DiffuseColor.kVertex, backface_culling=False,
                                    shading_normal=ShadingNormal.kFace))

    # --- icosphere level 4 (5120 faces), the config-2 camera scaled down, several options
    ico = meshgen.icosphere(4)
    c2w = meshgen.look_at_c2w((0, 0, -2.7), (0, 0, 0), (0, -1, 0))
    cam = PinholeCamera(192, 108, c2w, fov_y_deg=45.0)
    for tag, opt in cases.ALL_OPTIONS:
        if tag.startswith("tex"):
            continue
        save_case(f"ico4_{tag}", ico, cam, opt)

    # --- random soup (large + small triangles, flipped normals, two textures, all options)
    soup = cases.random_soup(400, seed=11)
    scam = PinholeCamera(96, 72, np.eye(4), fov_y_deg=60.0)
    for tag, opt in cases.ALL_OPTIONS:
        save_case(f"soup_{tag}", soup, scam, opt)


if __name__ == "__main__":
    main()
