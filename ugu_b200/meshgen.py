"""Deterministic synthetic meshes and camera poses for the configurations of BASELINE.json
(SURVEY.md §8d).  Host-side numpy only; these are INPUT generators — the same arrays are fed to the
CUDA path and to the oracle, so nothing here needs to match reference arithmetic.
"""
from __future__ import annotations

import math
from typing import List, Tuple

import numpy as np

from .renderer import Material, Mesh


def calc_face_normals(vertices: np.ndarray, faces: np.ndarray) -> np.ndarray:
    """Unit face normals, (v1 - v0) x (v2 - v0) normalised (as ugu::Mesh::CalcFaceNormal does)."""
    v0, v1, v2 = (vertices[faces[:, k]] for k in range(3))
    n = np.cross(v1 - v0, v2 - v0).astype(np.float32)
    ln = np.linalg.norm(n, axis=1, keepdims=True).astype(np.float32)
    ln[ln == 0] = 1.0
    return (n / ln).astype(np.float32)


def calc_normals(vertices: np.ndarray, faces: np.ndarray, face_normals: np.ndarray) -> np.ndarray:
    """Per-vertex normals = normalised sum of adjacent face normals (ugu::Mesh::CalcNormal)."""
    acc = np.zeros((len(vertices), 3), np.float64)
    for k in range(3):
        np.add.at(acc, faces[:, k], face_normals.astype(np.float64))
    ln = np.linalg.norm(acc, axis=1, keepdims=True)
    ln[ln == 0] = 1.0
    return (acc / ln).astype(np.float32)


def _icosahedron() -> Tuple[np.ndarray, np.ndarray]:
    t = (1.0 + math.sqrt(5.0)) / 2.0
    v = np.array([[-1, t, 0], [1, t, 0], [-1, -t, 0], [1, -t, 0], [0, -1, t], [0, 1, t],
                  [0, -1, -t], [0, 1, -t], [t, 0, -1], [t, 0, 1], [-t, 0, -1], [-t, 0, 1]],
                 np.float64)
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    f = np.array([[0, 11, 5], [0, 5, 1], [0, 1, 7], [0, 7, 10], [0, 10, 11], [1, 5, 9],
                  [5, 11, 4], [11, 10, 2], [10, 7, 6], [7, 1, 8], [3, 9, 4], [3, 4, 2],
                  [3, 2, 6], [3, 6, 8], [3, 8, 9], [4, 9, 5], [2, 4, 11], [6, 2, 10],
                  [8, 6, 7], [9, 8, 1]], np.int64)
    return v, f


def icosphere(level: int, radius: float = 1.0, vertex_color: bool = True) -> Mesh:
    """12-vertex icosahedron, `level` midpoint subdivisions re-normalised to `radius`
    (level 8: V = 655 362, F = 1 310 720 — config 2/5).  Outward CCW faces; vertex colour =
    255 * (0.5 * p / radius + 0.5)."""
    v, f = _icosahedron()
    for _ in range(level):
        nv = len(v)
        e = np.concatenate([f[:, [0, 1]], f[:, [1, 2]], f[:, [2, 0]]], axis=0)
        e.sort(axis=1)
        key = e[:, 0] * nv + e[:, 1]
        uniq, inv = np.unique(key, return_inverse=True)
        a, b = uniq // nv, uniq % nv
        mid = v[a] + v[b]
        mid /= np.linalg.norm(mid, axis=1, keepdims=True)
        v = np.concatenate([v, mid], axis=0)
        nf = len(f)
        m01, m12, m20 = nv + inv[:nf], nv + inv[nf:2 * nf], nv + inv[2 * nf:]
        f = np.concatenate([np.stack([f[:, 0], m01, m20], 1), np.stack([f[:, 1], m12, m01], 1),
                            np.stack([f[:, 2], m20, m12], 1), np.stack([m01, m12, m20], 1)], axis=0)
    verts = (v * radius).astype(np.float32)
    faces = f.astype(np.int32)
    fn = calc_face_normals(verts, faces)
    m = Mesh(vertices=verts, vertex_indices=faces, face_normals=fn,
             normals=calc_normals(verts, faces, fn), normal_indices=faces.copy())
    if vertex_color:
        m.vertex_colors = (255.0 * (0.5 * v + 0.5)).astype(np.float32)
    return m


def xorshift32_texture(width: int, height: int, seed: int = 1) -> np.ndarray:
    """height x width x 3 u8 procedural texture from a vectorised xorshift32 stream."""
    n = width * height * 3
    x = (np.arange(n, dtype=np.uint64) * np.uint64(2654435761) + np.uint64(seed)) & np.uint64(0xFFFFFFFF)
    x = x.astype(np.uint32)
    x[x == 0] = 1
    for _ in range(2):
        x ^= (x << np.uint32(13))
        x ^= (x >> np.uint32(17))
        x ^= (x << np.uint32(5))
    return (x >> np.uint32(24)).astype(np.uint8).reshape(height, width, 3)


def uv_sphere(n_stacks: int, n_slices: int, radius: float = 1.0, tex_size: int = 0,
              perturb: float = 0.03) -> Mesh:
    """UV-sphere topology with 2*n_slices*(n_stacks-1) faces (config 4: 2237 x 2236 -> 9 999 392),
    a seam-duplicated (n_stacks+1) x (n_slices+1) vertex grid, uv scaled into [0.001, 0.999] so the
    reference's unclamped bilinear taps stay inside the texture, radius perturbed by a fixed
    sinusoid, and an optional tex_size^2 procedural texture."""
    i = np.arange(n_stacks + 1, dtype=np.float64)[:, None]
    j = np.arange(n_slices + 1, dtype=np.float64)[None, :]
    theta = math.pi * i / n_stacks           # 0 (north) .. pi (south)
    phi = 2.0 * math.pi * j / n_slices
    r = radius * (1.0 + perturb * np.sin(5.0 * theta) * np.cos(7.0 * phi))
    x = r * np.sin(theta) * np.cos(phi)
    y = r * np.cos(theta) * np.ones_like(phi)
    z = r * np.sin(theta) * np.sin(phi)
    verts = np.stack([x, y, z], -1).reshape(-1, 3).astype(np.float32)
    u = 0.001 + 0.998 * (j / n_slices) * np.ones_like(i)
    v = 0.001 + 0.998 * (1.0 - i / n_stacks) * np.ones_like(j)
    uv = np.stack([u, v], -1).reshape(-1, 2).astype(np.float32)
    W = n_slices + 1
    ii, jj = np.meshgrid(np.arange(n_stacks), np.arange(n_slices), indexing="ij")
    a = (ii * W + jj).reshape(-1)
    b = a + 1
    c = a + W
    d = c + 1
    row = ii.reshape(-1)
    upper = np.stack([a, b, d], 1)[row > 0]             # skip degenerate triangles at the poles
    lower = np.stack([a, d, c], 1)[row < n_stacks - 1]
    faces = np.concatenate([upper, lower], 0).astype(np.int32)
    fn = calc_face_normals(verts, faces)
    # make every face normal point outward (the winding above is outward for this parametrisation
    # up to the sinusoidal perturbation; flip defensively so culling behaves like a closed surface)
    cen = verts[faces].mean(axis=1)
    flip = (fn * cen).sum(1) < 0
    faces[flip] = faces[flip][:, [0, 2, 1]]
    fn[flip] *= -1
    m = Mesh(vertices=verts, vertex_indices=faces, face_normals=fn,
             normals=calc_normals(verts, faces, fn), normal_indices=faces.copy(),
             uv=uv, uv_indices=faces.copy(),
             material_ids=np.zeros(len(faces), np.int32))
    if tex_size > 0:
        m.materials = [Material(diffuse_tex=xorshift32_texture(tex_size, tex_size, seed=1))]
    return m


def look_at_c2w(eye, center, up) -> np.ndarray:
    """Camera-to-world 4x4 (OpenCV convention: +z forward, +y down) like ugu::c2w
    (src/util/camera_util.cc:19-50): z = normalize(center - eye), x = normalize(z x up),
    y = normalize(z x x).  Computed in float32 then widened, as ex02/ex09 do."""
    eye = np.asarray(eye, np.float32)
    z = np.asarray(center, np.float32) - eye
    z = z / np.float32(np.linalg.norm(z))
    x = np.cross(z, np.asarray(up, np.float32)).astype(np.float32)
    x = x / np.float32(np.linalg.norm(x))
    y = np.cross(z, x).astype(np.float32)
    y = y / np.float32(np.linalg.norm(y))
    m = np.eye(4, dtype=np.float64)
    m[:3, 0], m[:3, 1], m[:3, 2], m[:3, 3] = x, y, z, eye
    return m


def orbit_poses(center, radius: float, n_views: int, up=(0.0, -1.0, 0.0)) -> List[np.ndarray]:
    """eye = center + radius * (sin a, 0, cos a), a = 2*pi*k/n (ex09_renderer_realtime.cc:137-142)."""
    c = np.asarray(center, np.float32)
    out = []
    for k in range(n_views):
        a = 2.0 * math.pi * k / n_views
        eye = c + np.float32(radius) * np.array([math.sin(a), 0.0, math.cos(a)], np.float32)
        out.append(look_at_c2w(eye, c, up))
    return out


def random_sphere_poses(n_views: int, radius: float = 2.7, seed: int = 0) -> List[np.ndarray]:
    """Eyes uniformly on a sphere of `radius` around the origin (config 5), looking at the origin."""
    rng = np.random.RandomState(seed)  # MT19937
    out = []
    while len(out) < n_views:
        d = rng.normal(size=3)
        n = np.linalg.norm(d)
        if n < 1e-6:
            continue
        d = d / n
        up = (0.0, 0.0, 1.0) if abs(d[2]) < 0.95 else (0.0, 1.0, 0.0)
        out.append(look_at_c2w(d * radius, (0.0, 0.0, 0.0), up))
    return out
