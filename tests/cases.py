"""Shared case definitions for the parity tests and the golden-fixture generator.

A case = (mesh, camera, option).  Golden cases are stored self-contained in tests/golden/*.npz
(inputs + the compiled reference's outputs + the reference's float camera constants), written by
tests/golden/make_golden.py; the builders below are what that script runs.
"""
from __future__ import annotations

import math
from pathlib import Path
from typing import Dict, List, Tuple

import numpy as np

from ugu_b200 import (ColorInterpolation, DiffuseColor, DiffuseShading, Material, Mesh, PinholeCamera,
                      RendererCpuOption, ShadingNormal)
from ugu_b200 import meshgen

GOLDEN_DIR = Path(__file__).resolve().parent / "golden"
OUTPUT_KEYS = ("color", "depth", "normal", "mask", "face_id")

MESH_KEYS = ("vertices", "vertex_colors", "vertex_indices", "normals", "face_normals",
             "normal_indices", "uv", "uv_indices", "material_ids")


def mesh_to_arrays(mesh: Mesh) -> Dict[str, np.ndarray]:
    d = {f"mesh_{k}": np.asarray(getattr(mesh, k)) for k in MESH_KEYS}
    d["mesh_num_materials"] = np.int32(len(mesh.materials))
    for i, m in enumerate(mesh.materials):
        d[f"mesh_tex{i}"] = m.diffuse_tex if m.diffuse_tex is not None else np.zeros((0, 0, 3), np.uint8)
    return d


def mesh_from_arrays(d) -> Mesh:
    m = Mesh(**{k: np.asarray(d[f"mesh_{k}"]) for k in MESH_KEYS})
    m.materials = [Material(diffuse_tex=(np.asarray(d[f"mesh_tex{i}"]) if d[f"mesh_tex{i}"].size else None))
                   for i in range(int(d["mesh_num_materials"]))]
    return m


def option_to_array(o: RendererCpuOption) -> np.ndarray:
    return np.array([int(o.diffuse_color), int(o.interp), int(o.shading_normal),
                     int(o.diffuse_shading), 1 if o.backface_culling else 0], np.int32)


def option_from_array(a) -> RendererCpuOption:
    return RendererCpuOption(diffuse_color=DiffuseColor(int(a[0])), interp=ColorInterpolation(int(a[1])),
                             shading_normal=ShadingNormal(int(a[2])),
                             diffuse_shading=DiffuseShading(int(a[3])), backface_culling=bool(a[4]))


def camera_to_arrays(c: PinholeCamera) -> Dict[str, np.ndarray]:
    return {"cam_size": np.array([c.width(), c.height()], np.int32),
            "cam_c2w": np.asarray(c.c2w(), np.float64),
            "cam_pp": np.array(c.principal_point(), np.float32),
            "cam_f": np.array(c.focal_length(), np.float32)}


def camera_from_arrays(d) -> PinholeCamera:
    return PinholeCamera(int(d["cam_size"][0]), int(d["cam_size"][1]), np.asarray(d["cam_c2w"]),
                         principal_point=d["cam_pp"], focal_length=d["cam_f"])


def pointcloud_prev(shape) -> np.ndarray:
    """The deterministic non-zero "previous contents" of the in/out point-cloud image used by the
    Depth2PointCloud fixtures (tests/golden/make_golden_pointcloud.py)."""
    n = int(shape[0]) * int(shape[1]) * 3
    return ((np.arange(n) % 7) + 1).astype(np.float32).reshape(int(shape[0]), int(shape[1]), 3)


def load_spot() -> Mesh:
    """data/spot (2930 v / 5856 f / 1024^2 texture) as loaded by the reference's own
    Mesh::LoadObj, frozen into tests/golden/spot_mesh.npz by make_golden.py."""
    d = np.load(GOLDEN_DIR / "spot_mesh.npz")
    return mesh_from_arrays(d)


def bbox_center_and_extent(mesh: Mesh) -> Tuple[np.ndarray, float]:
    lo, hi = mesh.vertices.min(0), mesh.vertices.max(0)
    return ((lo + hi) * 0.5).astype(np.float32), float((hi - lo).max())


# ---------------------------------------------------------------------------------------------
# micro known-answer meshes
# ---------------------------------------------------------------------------------------------
def _tri_mesh(verts, faces, face_normals=None, colors=None) -> Mesh:
    v = np.asarray(verts, np.float32)
    f = np.asarray(faces, np.int32)
    fn = meshgen.calc_face_normals(v, f) if face_normals is None else np.asarray(face_normals, np.float32)
    m = Mesh(vertices=v, vertex_indices=f, face_normals=fn,
             normals=meshgen.calc_normals(v, f, fn), normal_indices=f.copy())
    if colors is None:
        rng = np.random.RandomState(7)
        colors = rng.uniform(0, 255, size=v.shape)
    m.vertex_colors = np.asarray(colors, np.float32)
    return m


def kat_meshes() -> Dict[str, Mesh]:
    """Camera at origin looking down +z, y down (identity c2w).  A face is FRONT when its normal
    points toward the camera (normal . ray <= 0)."""
    out = {}
    # (v1-v0)x(v2-v0) toward -z (against the +z viewing ray) => front-facing for an identity camera
    back = lambda z, s=1.0, ox=0.0, oy=0.0: [[-s + ox, -s + oy, z], [s + ox, -s + oy, z], [ox, s + oy, z]]
    front = lambda z, s=1.0, ox=0.0, oy=0.0: [[-s + ox, -s + oy, z], [ox, s + oy, z], [s + ox, -s + oy, z]]
    # 1. dummy face 0 far away, then two overlapping front triangles at different depths
    v = front(50.0, 0.01, 30, 30) + front(3.0) + front(2.0, 0.6, 0.3, 0.1)
    out["two_overlapping"] = _tri_mesh(v, [[0, 1, 2], [3, 4, 5], [6, 7, 8]])
    # 2. shared edge / identical depth: two coplanar triangles forming a quad + a duplicate of one
    q = [[-1, -1, 2.5], [-1, 1, 2.5], [1, 1, 2.5], [1, -1, 2.5]]
    out["shared_edge_tie"] = _tri_mesh(front(50.0, 0.01, 30, 30) + q,
                                       [[0, 1, 2], [3, 4, 5], [3, 5, 6], [3, 4, 5]])
    # 3. face-0 quirk: a lone front triangle as face 0 (depth/face_id written, mask/colour/normal not)
    out["face0_quirk"] = _tri_mesh(front(2.0), [[0, 1, 2]])
    # 4. back face occluding a front face (leaves a hole when culling is on)
    out["backface_occluder"] = _tri_mesh(front(50.0, 0.01, 30, 30) + front(3.0) + back(2.0, 0.5),
                                         [[0, 1, 2], [3, 4, 5], [6, 7, 8]])
    # 5. rejects: behind the camera, straddling z<0, off-screen, degenerate; plus one visible
    v = (front(50.0, 0.01, 30, 30) + front(-2.0) +
         [[-1, -1, 2.0], [1, -1, -0.5], [0, 1, 2.0]] + front(2.0, 0.2, 40.0, 0.0) +
         [[0, 0, 2.0], [0.5, 0.5, 2.0], [1.0, 1.0, 2.0]] + front(2.5, 0.7))
    out["rejects"] = _tri_mesh(v, [[0, 1, 2], [3, 4, 5], [6, 7, 8], [9, 10, 11], [12, 13, 14],
                                   [15, 16, 17]],
                               face_normals=[[0, 0, -1]] * 6)
    # 6. a large triangle covering the whole screen (binned path) behind a fan of small ones
    rng = np.random.RandomState(3)
    vs = front(50.0, 0.01, 30, 30) + front(6.0, 40.0)
    fs = [[0, 1, 2], [3, 4, 5]]
    for k in range(60):
        cx, cy = rng.uniform(-1.5, 1.5, 2)
        z = rng.uniform(2.0, 5.0)
        s = rng.uniform(0.02, 0.5)
        tri = front(z, s, cx, cy) if k % 3 else back(z, s, cx, cy)
        b = len(vs)
        vs += tri
        fs.append([b, b + 1, b + 2])
    out["large_and_small"] = _tri_mesh(vs, fs)
    return out


def kat_camera(w=64, h=48) -> PinholeCamera:
    return PinholeCamera(w, h, np.eye(4), principal_point=(w * 0.5 - 0.5, h * 0.5 - 0.5),
                         focal_length=(w * 0.55, w * 0.55))


def random_soup(n_tris: int, seed: int, big_fraction: float = 0.1) -> Mesh:
    """Random triangle soup in front of an identity camera: mostly small triangles, some large /
    screen-filling ones, random orientation, random (un-normalised, sometimes flipped) face normals."""
    rng = np.random.RandomState(seed)
    cen = np.stack([rng.uniform(-2, 2, n_tris), rng.uniform(-1.5, 1.5, n_tris),
                    rng.uniform(1.5, 8.0, n_tris)], 1)
    size = np.where(rng.uniform(size=n_tris) < big_fraction, rng.uniform(0.5, 6.0, n_tris),
                    rng.uniform(0.005, 0.15, n_tris))
    off = rng.normal(size=(n_tris, 3, 3)) * size[:, None, None]
    off[:, :, 2] *= 0.3
    v = (cen[:, None, :] + off).reshape(-1, 3).astype(np.float32)
    v[:, 2] = np.maximum(v[:, 2], 0.25)
    f = np.arange(n_tris * 3, dtype=np.int32).reshape(-1, 3)
    # face 0 is never shaded: make it a far dummy
    v[0:3] = np.array([[30, 30, 50], [30.01, 30, 50], [30, 30.01, 50]], np.float32)
    fn = meshgen.calc_face_normals(v, f)
    flip = rng.uniform(size=n_tris) < 0.15
    fn[flip] *= -1.0
    fn *= rng.uniform(0.5, 2.0, size=(n_tris, 1)).astype(np.float32)
    m = Mesh(vertices=v, vertex_indices=f, face_normals=fn.astype(np.float32),
             normals=rng.normal(size=(n_tris * 3, 3)).astype(np.float32), normal_indices=f.copy(),
             vertex_colors=rng.uniform(0, 255, size=(n_tris * 3, 3)).astype(np.float32),
             uv=rng.uniform(0.02, 0.98, size=(n_tris * 3, 2)).astype(np.float32), uv_indices=f.copy(),
             material_ids=rng.randint(0, 2, size=n_tris).astype(np.int32))
    m.materials = [Material(diffuse_tex=meshgen.xorshift32_texture(64, 48, seed=5)),
                   Material(diffuse_tex=meshgen.xorshift32_texture(33, 57, seed=9))]
    return m


ALL_OPTIONS: List[Tuple[str, RendererCpuOption]] = [
    ("vertex_cull", RendererCpuOption(diffuse_color=DiffuseColor.kVertex)),
    ("vertex_nocull_lambert", RendererCpuOption(diffuse_color=DiffuseColor.kVertex, backface_culling=False,
                                                diffuse_shading=DiffuseShading.kLambertian)),
    ("none_face_normal", RendererCpuOption(diffuse_color=DiffuseColor.kNone,
                                           shading_normal=ShadingNormal.kFace)),
    ("tex_bilinear", RendererCpuOption(diffuse_color=DiffuseColor.kTexture)),
    ("tex_nn_nocull", RendererCpuOption(diffuse_color=DiffuseColor.kTexture,
                                        interp=ColorInterpolation.kNn, backface_culling=False)),
    ("vertex_oren_nayar", RendererCpuOption(diffuse_color=DiffuseColor.kVertex,
                                            diffuse_shading=DiffuseShading.kOrenNayar)),
]


def ex02_poses(mesh: Mesh, lookat) -> List[np.ndarray]:
    """The six poses of examples/ex02_renderer.cc:49-102: eye = stats.center (the vertex mean,
    src/mesh.cc:187-205) +- 1.5 * the largest bbox edge along z, x, y.  `lookat(eye, center, up)`
    builds the c2w (the reference's ugu::c2w when the fixture generator runs,
    meshgen.look_at_c2w otherwise)."""
    c = mesh.vertices.astype(np.float64).mean(0).astype(np.float32)
    diff = mesh.vertices.max(0) - mesh.vertices.min(0)
    d = np.float32(np.float32(diff.max()) * np.float32(1.5))
    poses = []
    for axis, sign, up in ((2, 1, (0, 1, 0)), (2, -1, (0, 1, 0)), (0, 1, (0, 1, 0)),
                           (0, -1, (0, 1, 0)), (1, -1, (0, 0, 1)), (1, 1, (0, 0, -1))):
        eye = c.copy()
        eye[axis] += np.float32(sign) * d
        poses.append(lookat(eye, c, up))
    return poses


def procedural_color(cw: int, ch: int) -> np.ndarray:
    """A deterministic (ch, cw, 3) u8 image whose every texel differs from its neighbours (integer arithmetic only:
    identical wherever it is generated)."""
    y, x = np.mgrid[0:ch, 0:cw].astype(np.int64)
    return np.stack([(x * 7 + y * 13) % 256, (x * 3 + y * 29 + 11) % 256, (x * x + 5 * y) % 256], -1).astype(np.uint8)


def two_camera_setup(cw: int, ch: int, fov_y_deg: float, variant: int):
    """Colour camera (a wider pinhole of its own size), depth2color (4x4 float32: a small rotation + shift, built
    from exactly representable entries so that it is identical everywhere) and the colour image of the two-camera
    Depth2Mesh cases."""
    ccam = PinholeCamera(cw, ch, np.eye(4), fov_y_deg=fov_y_deg)
    m = np.eye(4, dtype=np.float32)
    if variant == 0:
        m[:3, :3] = np.array([[0.99951171875, 0.0, 0.03125], [0.0009765625, 1.0, 0.001953125],
                              [-0.03125, 0.0029296875, 0.99951171875]], np.float32)
        m[:3, 3] = np.array([0.015625, -0.0078125, 0.01171875], np.float32)
    elif variant == 1:
        m[:3, :3] = np.array([[1.0, -0.015625, 0.0], [0.015625, 1.0, 0.0078125], [0.0, -0.0078125, 1.0]], np.float32)
        m[:3, 3] = np.array([-0.03125, 0.0, 0.0625], np.float32)
    else:
        m[:3, 3] = np.array([0.0, 0.001953125, 0.25], np.float32)
    return ccam, m, procedural_color(cw, ch)
