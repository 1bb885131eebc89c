"""Generates tests/golden/raytrace_*.npz from the COMPILED, UNMODIFIED reference ugu::Raytracer
(src/renderer/cpu/raytracer.cc with the nanort BVH; oracle/_ref/libugu_ref.so, built from /root/reference
by oracle/Makefile).  Run only where /root/reference exists:

    make -C oracle ref && python tests/golden/make_golden_raytrace.py

Every fixture reuses the mesh / camera / option of the Rasterizer fixture golden_<name>.npz of the same
name (tests/golden/make_golden.py) and stores the five Raytracer output images; `depth_scale` holds the
RendererCpuOption::depth_scale used (the Raytracer applies it, raytracer.cc:171, the Rasterizer does not).
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
from oracle import OracleRef  # noqa: E402

OUT = Path(__file__).resolve().parent


def main():
    for name in util.golden_names():
        mesh, cam, opt, _, _ = util.load_golden(name)
        opt.depth_scale = 1000.0 if ("nocull" in name or "top" in name) else 1.0
        out = OracleRef(mesh).raytrace(cam, opt)
        assert out is not None, name
        d = {f"out_{k}": out[k] for k in cases.OUTPUT_KEYS}
        d["depth_scale"] = np.float32(opt.depth_scale)
        np.savez_compressed(OUT / f"raytrace_{name}.npz", **d)
        print(f"raytrace_{name}: covered={(out['face_id'] >= 0).sum()} depth_scale={opt.depth_scale}")


if __name__ == "__main__":
    main()
