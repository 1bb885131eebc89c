"""Generates tests/golden/uvraster_*.npz from the COMPILED, UNMODIFIED reference's
ugu::RasterizeVertexAttributeToTexture / RasterizeFaceAttributeToTexture / GenerateUvMask
(include/ugu/util/raster_util.h:139-268, src/util/raster_util.cc:9-27) through oracle/_ref, with
num_threads = 1 (the reference's only deterministic order).  Run only where /root/reference exists:

    make -C oracle ref && python tests/golden/make_golden_uvraster.py

Inputs come from tests/uvcases.py (deterministic); each file stores them next to the outputs so the
fixtures stay valid even if a generator changes.
"""
from __future__ import annotations

import sys
from pathlib import Path

import numpy as np

REPO = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(REPO))
sys.path.insert(0, str(REPO / "tests"))

import uvcases  # noqa: E402
from oracle import ref_uv_mask, ref_uv_raster  # noqa: E402

OUT = Path(__file__).resolve().parent

for name, c in uvcases.all_cases().items():
    tex, mask = uvcases.run_oracle(ref_uv_raster, c)
    d = {k: v for k, v in c.items() if v is not None}
    d["out_texture"] = tex
    if mask is not None:
        d["out_mask"] = mask
    if name.startswith("spot_"):  # the spot uv arrays live in spot_mesh.npz already
        d.pop("uv"), d.pop("uv_indices")
        if c["attr_indices"] is not None:
            d.pop("attr_indices")
        d["spot"] = np.int32(1)
    np.savez_compressed(OUT / f"uvraster_{name}.npz", **d)
    print(name, tex.shape, tex.dtype, "covered", None if mask is None else int((mask == 255).sum()))

c = uvcases.spot_case("face_u8_128")
m = ref_uv_mask(c["uv"], c["uv_indices"], 200, 100, 200)
np.savez_compressed(OUT / "uvmask_spot_200x100_c200.npz", out_mask=m, width=np.int32(200),
                    height=np.int32(100), color=np.int32(200))
print("uvmask", m.shape, int((m == 200).sum()))
