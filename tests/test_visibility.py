"""Face-visibility consumer of face_id maps (include/ugu_b200_raster.h
ugu_b200_face_visibility).  PARITY UNPINNED against the reference — it has no function that consumes
face_id maps (VisibilityTester shoots BVH rays, src/texturing/visibility_tester.cc:416-529) — so the
CUDA path is checked against the numpy restatement of the stated definition (oracle.face_visibility)
on hand-made maps, random maps and maps rendered by the (reference-pinned) rasterizer; the
multi-rank reduction is checked under gloo with world_size 2."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import face_visibility
from ugu_b200.multiview import MultiViewRenderer, decode_best


def test_definition_known_answers():
    maps = np.array([[[0, 0, 1, -1], [2, 2, 2, 7]],     # view 10: face 0 x2, 1 x1, 2 x3 (7 out of range)
                     [[1, 1, 1, 1], [0, 0, -1, -1]],    # view 11: face 1 x4, 0 x2
                     [[2, 2, 2, 3], [3, -1, -1, -1]]],  # view 12: face 2 x3, 3 x2
                    np.int32)
    got = face_visibility(maps, 5, first_view_id=10)
    assert got["pixel_count"].tolist() == [4, 5, 6, 2, 0]
    assert got["view_count"].tolist() == [2, 2, 2, 1, 0]
    assert got["best_view"].tolist() == [10, 11, 10, 12, -1]   # ties (faces 0 and 2) -> lowest view id
    assert got["best_count"].tolist() == [2, 4, 3, 2, 0]


def test_decode_best():
    best = torch.tensor([0, (3 << 32) | (0xFFFFFFFF - 10), (1 << 32) | 0xFFFFFFFF], dtype=torch.int64)
    v, c = decode_best(best)
    assert v.tolist() == [-1, 10, 0] and c.tolist() == [0, 3, 1]


# --- multi-rank reduction under gloo (fake renderer + numpy accumulate: no GPU here) ------------
class _Cam:
    def __init__(self, i):
        self.i = i

    def width(self):
        return 16

    def height(self):
        return 8


def _fake_map(i):
    rng = np.random.default_rng(100 + i)
    return rng.integers(-1, 12, (8, 16)).astype(np.int32)


def _fake_render(cameras, outputs):
    return {"face_id": torch.from_numpy(np.stack([_fake_map(c.i) for c in cameras]))}


def _numpy_accumulate(maps, first_view_id, num_faces, pix, views, best):
    r = face_visibility(maps.numpy(), num_faces, first_view_id)
    pix += torch.from_numpy(r["pixel_count"].astype(np.int32))
    views += torch.from_numpy(r["view_count"].astype(np.int32))
    packed = (r["best_count"].astype(np.int64) << 32) | (0xFFFFFFFF - r["best_view"].astype(np.int64))
    packed[r["best_view"] < 0] = 0
    torch.maximum(best, torch.from_numpy(packed), out=best)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n_views, mode, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        cams = [_Cam(i) for i in range(n_views)]
        mv = MultiViewRenderer(mode=mode, render_local=_fake_render, accumulate=_numpy_accumulate)
        got = mv.render_face_visibility(cams, 12, chunk_views=3)
        want = face_visibility(np.stack([_fake_map(i) for i in range(n_views)]), 12)
        q.put(all(np.array_equal(got[k].numpy().astype(np.int64), want[k].astype(np.int64)) for k in want))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n_views,mode", [(2, 9, "contiguous"), (2, 8, "round_robin"), (1, 5, "contiguous")])
def test_sharded_face_visibility_gloo(world, n_views, mode):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_views, mode, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert all(q.get(timeout=10) for _ in range(world))


# ---------------------------------------------------------------------------------------------- GPU
def _check(got, want):
    for k in want:
        assert np.array_equal(np.asarray(got[k]).astype(np.int64), want[k].astype(np.int64)), k


@pytest.mark.gpu
@pytest.mark.parametrize("shape,F", [((3, 8, 16), 12), ((1, 1, 1), 1), ((11, 37, 19), 50), ((20, 64, 64), 7),
                                     ((9, 5, 3), 1000)])
def test_cuda_face_visibility_random_maps(shape, F):
    from ugu_b200 import Rasterizer
    rng = np.random.default_rng(sum(shape) + F)
    maps = rng.integers(-1, F + 2, shape).astype(np.int32)   # includes -1 and ids >= F (ignored)
    r = Rasterizer()
    for first in (0, 1000):
        _check(r.face_visibility(maps, F, first), face_visibility(maps, F, first))
    _check(r.face_visibility(maps[:0], F), face_visibility(maps[:0], F))  # no views


@pytest.mark.gpu
def test_cuda_face_visibility_on_rendered_maps_device_form():
    """Rendered maps (icosphere, 24 views at 256^2 — config 5's shape, and spot at 512^2 with faces
    of hundreds of pixels) accumulated on the device in two calls; against numpy on the same maps."""
    import cases
    from ugu_b200 import DiffuseColor, PinholeCamera, Rasterizer, RendererCpuOption, meshgen
    for mesh, size, poses in ((meshgen.icosphere(6), 256, meshgen.random_sphere_poses(24, seed=1)),
                              (cases.load_spot(), 512, None)):
        if poses is None:
            c, ext = cases.bbox_center_and_extent(mesh)
            poses = meshgen.orbit_poses(c, 1.5 * ext, 10)
        F = len(mesh.vertex_indices)
        r = Rasterizer(RendererCpuOption(diffuse_color=DiffuseColor.kNone))
        r.set_mesh(mesh)
        assert r.PrepareMesh()
        cams = [PinholeCamera(size, size, p, fov_y_deg=45.0) for p in poses]
        mv = MultiViewRenderer(r)
        got = mv.render_face_visibility(cams, F, chunk_views=7)     # several accumulate calls
        _, outs = mv.render(cams, ("face_id",))
        maps = outs["face_id"].cpu().numpy()
        want = face_visibility(maps, F)
        _check({k: v.cpu().numpy() for k, v in got.items()}, want)
        assert (want["view_count"] > 0).sum() > F // 4
        # the host form agrees too
        _check(r.face_visibility(maps, F), want)
