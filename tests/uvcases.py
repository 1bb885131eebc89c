"""Deterministic inputs of the UV-space rasterization cases (tests/golden/uvraster_*.npz).

A case is a dict: attrs (n, 3) f32, attr_indices (F, 3) i32 or None (per-face attributes), uv (m, 2)
f32, uv_indices (F, 3) i32, width, height, texture_type (0 = Image3b, 1 = Image3f), prev_texture /
prev_mask (arrays or None).  prev_texture not None = the reference's width/height <= 0 form
(rasterise over the existing image, raster_util.h:151-155); prev_mask not None = a mask is passed
(the reference never clears it)."""
from __future__ import annotations

from typing import Dict

import numpy as np

import cases


def _rng_floats(n, lo, hi, seed):
    """xorshift32-based floats, independent of numpy's generator versions."""
    out = np.empty(n, np.float64)
    s = np.uint32(seed * 2654435761 % (1 << 32) or 1)
    x = int(s)
    for i in range(n):
        x ^= (x << 13) & 0xFFFFFFFF
        x ^= x >> 17
        x ^= (x << 5) & 0xFFFFFFFF
        out[i] = x / 4294967296.0
    return (lo + (hi - lo) * out).astype(np.float32)


def kat_case() -> Dict:
    """Hand-made: two overlapping triangles (the later face owns the overlap), a shared edge, a
    zero-area triangle (area -> -FLT_MIN, raster_util.h:84-87), triangles partly / fully outside the
    texture and uv outside [0, 1], a sliver, opposite windings."""
    uv = np.array([
        [0.10, 0.10], [0.60, 0.15], [0.20, 0.70],    # 0-2  face 0
        [0.30, 0.20], [0.80, 0.30], [0.40, 0.90],    # 3-5  face 1 overlaps face 0
        [0.60, 0.15], [0.90, 0.60],                  # 6-7  face 2 shares the edge 1-? with face 0
        [0.50, 0.50], [0.70, 0.70], [0.90, 0.90],    # 8-10 face 3: collinear (zero area)
        [-0.30, 0.40], [0.20, 0.45], [-0.10, 0.95],  # 11-13 face 4: partly left of the texture
        [1.50, 1.50], [1.80, 1.50], [1.60, 1.90],    # 14-16 face 5: fully outside
        [0.05, 0.95], [0.95, 0.951], [0.05, 0.952],  # 17-19 face 6: sliver
        [0.55, 0.05], [0.95, 0.05], [0.75, 0.30],    # 20-22 face 7 / 8: same triangle, both windings
    ], np.float32)
    uv_idx = np.array([[0, 1, 2], [3, 4, 5], [1, 7, 2], [8, 9, 10], [11, 12, 13], [14, 15, 16],
                       [17, 18, 19], [20, 21, 22], [22, 21, 20]], np.int32)
    attrs = np.array([[255, 0, 0], [0, 255, 0], [0, 0, 255], [250.5, 10.5, 99.49], [-30, 300, 128],
                      [12.5, 13.5, 14.5], [1e6, -1e6, 77]], np.float32)
    attr_idx = np.array([[0, 1, 2], [3, 3, 3], [2, 1, 0], [0, 1, 2], [4, 5, 0], [1, 1, 1], [5, 5, 5],
                         [0, 2, 4], [6, 1, 3]], np.int32)
    prev_mask = np.full((48, 64), 7, np.uint8)
    return dict(attrs=attrs, attr_indices=attr_idx, uv=uv, uv_indices=uv_idx, width=64, height=48,
                texture_type=0, prev_texture=None, prev_mask=prev_mask)


def soup_case(n_tris=200, w=96, h=80, seed=3, texture_type=1, keep=True) -> Dict:
    uv = np.empty((n_tris * 3, 2), np.float32)
    c = _rng_floats(n_tris * 2, -0.1, 1.1, seed).reshape(-1, 2)
    r = _rng_floats(n_tris, 0.005, 0.2, seed + 1)
    d = _rng_floats(n_tris * 6, -1.0, 1.0, seed + 2).reshape(-1, 3, 2)
    for k in range(3):
        uv[k::3] = c + d[:, k] * r[:, None]
    uv_idx = np.arange(n_tris * 3, dtype=np.int32).reshape(-1, 3)
    attrs = _rng_floats(n_tris * 9, -5.0, 260.0, seed + 3).reshape(-1, 3)
    attr_idx = uv_idx.copy()
    prev_t = prev_m = None
    if keep:
        dt = np.float32 if texture_type else np.uint8
        prev_t = (_rng_floats(w * h * 3, 0, 255, seed + 4).reshape(h, w, 3)).astype(dt)
        prev_m = np.where(_rng_floats(w * h, 0, 1, seed + 5).reshape(h, w) > 0.5, 9, 0).astype(np.uint8)
    return dict(attrs=attrs, attr_indices=attr_idx, uv=uv, uv_indices=uv_idx, width=w, height=h,
                texture_type=texture_type, prev_texture=prev_t, prev_mask=prev_m)


def spot_case(kind: str) -> Dict:
    spot = cases.load_spot()
    uv = np.asarray(spot.uv, np.float32)
    ui = np.asarray(spot.uv_indices, np.int32)
    vi = np.asarray(spot.vertex_indices, np.int32)
    nv, nf = len(spot.vertices), len(ui)
    if kind == "vertex_u8_256":
        return dict(attrs=_rng_floats(nv * 3, 0, 255, 11).reshape(-1, 3), attr_indices=vi, uv=uv,
                    uv_indices=ui, width=256, height=256, texture_type=0, prev_texture=None,
                    prev_mask=np.zeros((256, 256), np.uint8))
    if kind == "vertex_f32_300x200":  # a position map, as app/posmap_gen.cc:74 makes
        return dict(attrs=np.asarray(spot.vertices, np.float32), attr_indices=vi, uv=uv, uv_indices=ui,
                    width=300, height=200, texture_type=1, prev_texture=None,
                    prev_mask=np.zeros((200, 300), np.uint8))
    if kind == "face_u8_128":  # out-of-range values exercise round + the wrapping u8 cast
        return dict(attrs=_rng_floats(nf * 3, -20, 280, 12).reshape(-1, 3), attr_indices=None, uv=uv,
                    uv_indices=ui, width=128, height=128, texture_type=0, prev_texture=None,
                    prev_mask=None)
    raise KeyError(kind)


def all_cases() -> Dict[str, Dict]:
    return {
        "kat": kat_case(),
        "soup_keep_f32": soup_case(),
        "soup_fresh_u8": soup_case(seed=8, texture_type=0, keep=False, w=50, h=33),
        "spot_vertex_u8_256": spot_case("vertex_u8_256"),
        "spot_vertex_f32_300x200": spot_case("vertex_f32_300x200"),
        "spot_face_u8_128": spot_case("face_u8_128"),
    }


def run_oracle(fn, case):
    """fn = oracle.port_uv_raster or oracle.ref_uv_raster -> (texture, mask or None) or None"""
    return fn(case["attrs"], case["attr_indices"], case["uv"], case["uv_indices"], case["width"],
              case["height"], case["texture_type"], case["prev_texture"], case["prev_mask"])
