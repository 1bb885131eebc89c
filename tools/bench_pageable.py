#!/usr/bin/env python
"""Render into PAGEABLE host images (the layout of ugu::Image) at several host-copy thread counts:
    python tools/bench_pageable.py [threads ...]
One JSON line per setting: ms per Render of the bench workload (config 2), checked against the pinned result."""
import json
import os
import sys
import time
from pathlib import Path

REPO = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(REPO), str(REPO / "tests")]
import numpy as np  # noqa: E402
import torch  # noqa: E402

import bench  # noqa: E402
from ugu_b200 import Image1b, Image1f, Image1i, Image3b, Image3f, Rasterizer  # noqa: E402


def main():
    mesh, opt, cams = bench.make_inputs()
    flush = torch.empty(bench.L2_FLUSH_BYTES, dtype=torch.uint8, device="cuda")
    ref = None
    for threads in [int(a) for a in sys.argv[1:]] or [1, 2, 4, 8]:
        os.environ["UGU_B200_COPY_THREADS"] = str(threads)
        r = Rasterizer(opt)
        r.set_mesh(mesh)
        assert r.PrepareMesh()
        for pinned in (True, False):
            imgs = (Image3b(pinned=pinned), Image1f(pinned=pinned), Image3f(pinned=pinned), Image1b(pinned=pinned),
                    Image1i(pinned=pinned))
            ts = []
            for i in range(23):
                flush.zero_()
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                r.set_camera(cams[i % len(cams)])
                assert r.Render(*imgs)
                ts.append(time.perf_counter() - t0)
            ts = sorted(ts[3:])
            got = [np.array(im.data) for im in imgs]
            if ref is None:
                ref = got
            same = all(np.array_equal(a, b) for a, b in zip(ref, got))
            print(json.dumps({"copy_threads": threads, "pinned_destinations": pinned, "ms_min": ts[0] * 1e3,
                              "ms_median": ts[len(ts) // 2] * 1e3, "ms_max": ts[-1] * 1e3,
                              "equal_to_first_result": bool(same)}), flush=True)
        del r


if __name__ == "__main__":
    main()
