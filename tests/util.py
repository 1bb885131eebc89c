"""Helpers shared by the parity tests."""
from __future__ import annotations

from typing import Dict, Iterable, Optional

import numpy as np

import cases
from ugu_b200 import (Image1b, Image1f, Image1i, Image3b, Image3f, Mesh, PinholeCamera, Rasterizer,
                      RendererCpuOption)

# BASELINE.json north_star tolerances
DEPTH_RTOL = 1e-5
NORMAL_ATOL = 1e-4
COLOR_LSB = 1


def golden_names():
    return sorted(p.stem[len("golden_"):] for p in cases.GOLDEN_DIR.glob("golden_*.npz"))


def load_golden(name: str):
    d = np.load(cases.GOLDEN_DIR / f"golden_{name}.npz")
    ref = str(d["mesh_ref"])
    mesh = cases.mesh_from_arrays(np.load(cases.GOLDEN_DIR / ref)) if ref else cases.mesh_from_arrays(d)
    cam = cases.camera_from_arrays(d)
    opt = cases.option_from_array(d["option"])
    out = {k: d[f"out_{k}"] for k in cases.OUTPUT_KEYS}
    consts = (d["ref_w2c_R"], d["ref_w2c_t"], d["ref_c2w_R"])
    return mesh, cam, opt, out, consts


class CudaRunner:
    """Drives the product exactly like a reference caller: set_mesh / PrepareMesh / set_camera /
    Render through the C-ABI."""

    def __init__(self, mesh: Mesh, option: Optional[RendererCpuOption] = None, pinned: bool = True):
        self.r = Rasterizer(option)
        self.r.set_mesh(mesh)
        assert self.r.PrepareMesh()
        self.imgs = dict(color=Image3b(pinned), depth=Image1f(pinned), normal=Image3f(pinned),
                         mask=Image1b(pinned), face_id=Image1i(pinned))

    def render(self, camera: PinholeCamera, option: RendererCpuOption,
               want: Iterable[str] = cases.OUTPUT_KEYS) -> Optional[Dict[str, np.ndarray]]:
        self.r.set_option(option)
        self.r.set_camera(camera)
        args = {k: (self.imgs[k] if k in want else None) for k in cases.OUTPUT_KEYS}
        if not self.r.Render(**args):
            return None
        return {k: self.imgs[k].data.copy() for k in want}


def bits(a: np.ndarray) -> np.ndarray:
    return np.ascontiguousarray(a).view(np.uint8)


def assert_parity(got: Dict[str, np.ndarray], want: Dict[str, np.ndarray], strict: bool = True,
                  color_lsb: int = 0, what: str = ""):
    """mask / face_id bit-exact; depth <= 1e-5 relative; normal <= 1e-4; colour +-1 LSB
    (BASELINE.json).  With strict=True additionally demands bit-exact depth / normal and colour
    within `color_lsb` (0 = bit-exact), which is what the kernels are built to deliver."""
    for k in ("face_id", "mask"):
        if k in want and k in got:
            bad = int((got[k] != want[k]).sum())
            assert bad == 0, f"{what}: {k} differs at {bad} pixels"
    if "depth" in want and "depth" in got:
        g, w = got["depth"].astype(np.float64), want["depth"].astype(np.float64)
        rel = np.abs(g - w) / np.maximum(np.abs(w), 1e-30)
        rel[(g == w)] = 0.0
        assert rel.max() <= DEPTH_RTOL, f"{what}: depth rel err {rel.max()}"
        if strict:
            assert np.array_equal(bits(got["depth"]), bits(want["depth"])), f"{what}: depth bits differ"
    if "normal" in want and "normal" in got:
        err = np.abs(got["normal"].astype(np.float64) - want["normal"].astype(np.float64)).max()
        assert err <= NORMAL_ATOL, f"{what}: normal abs err {err}"
        if strict:
            # -0.0 vs +0.0 is not a value difference
            assert np.array_equal(got["normal"], want["normal"]), f"{what}: normal values differ"
    if "color" in want and "color" in got:
        diff = np.abs(got["color"].astype(np.int32) - want["color"].astype(np.int32)).max()
        assert diff <= COLOR_LSB, f"{what}: colour differs by {diff} LSB"
        if strict:
            assert diff <= color_lsb, f"{what}: colour differs by {diff} LSB (strict {color_lsb})"
