"""Host logic of the multi-view dispatch (ugu_b200/multiview.py) on CPU: partitioning, and the
world_size-2/3 gather of ragged per-view maps under gloo.  The renderer is a fake (no GPU here);
the GPU suite checks the real one against single-view renders."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from ugu_b200.multiview import MultiViewRenderer, gather_view_maps, partition


@pytest.mark.parametrize("mode", ["contiguous", "round_robin"])
@pytest.mark.parametrize("n,world", [(36, 1), (36, 2), (36, 8), (5, 8), (0, 4), (1024, 8), (7, 3)])
def test_partition_covers_every_view_once(n, world, mode):
    parts = [partition(n, world, r, mode) for r in range(world)]
    flat = sorted(i for p in parts for i in p)
    assert flat == list(range(n))
    sizes = [len(p) for p in parts]
    assert max(sizes) - min(sizes) <= 1
    if mode == "contiguous":
        assert all(p == list(range(p[0], p[0] + len(p))) for p in parts if p)


class _Cam:
    def __init__(self, i):
        self.i = i

    def width(self):
        return 6

    def height(self):
        return 4


def _fake_render(cameras, outputs):
    # "face_id map" of view i = i * 100 + pixel index
    maps = [torch.arange(24, dtype=torch.int32).reshape(4, 6) + 100 * c.i for c in cameras]
    return {"face_id": torch.stack(maps) if maps else torch.empty((0, 4, 6), dtype=torch.int32)}


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
        mv = MultiViewRenderer(mode=mode, render_local=_fake_render)
        idx, outs = mv.render(cams)
        assert idx == partition(n_views, world, rank, mode)
        got = mv.render_and_gather(cams, "face_id", dst=0)
        # the chunked form: chunks sent as they finish, straight into the view-ordered destination
        got2 = mv.render_and_gather_overlapped(cams, "face_id", dst=0, chunk_views=2)
        if rank == 0:
            want = _fake_render(cams, ("face_id",))["face_id"]
            q.put(bool(torch.equal(got, want)) and bool(torch.equal(got2, want)))
        else:
            assert got is None and got2 is None
            q.put(True)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n_views,mode", [(2, 7, "contiguous"), (2, 6, "round_robin"),
                                                (3, 2, "contiguous"), (3, 11, "contiguous")])
def test_sharded_render_and_gather_gloo(world, n_views, mode):
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


def test_single_process_gather_is_identity():
    t = torch.arange(48, dtype=torch.int32).reshape(2, 4, 6)
    assert gather_view_maps(t, 2) is t


@pytest.mark.parametrize("mode", ["contiguous", "round_robin"])
@pytest.mark.parametrize("n,world", [(36, 1), (36, 2), (36, 8), (5, 8), (0, 4), (1024, 8), (7, 3)])
def test_c_abi_partition_matches_the_python_one(n, world, mode):
    """ugu_b200_partition (what C++ hosts and the library's own sharding use) == ugu_b200.multiview.partition."""
    import ctypes as C
    from ugu_b200 import _native as N
    lib = N.load()
    for r in range(world):
        cnt = C.c_int32(-1)
        idx = (C.c_int32 * max(n, 1))()
        assert lib.ugu_b200_partition(n, world, r, 1 if mode == "round_robin" else 0, idx, C.byref(cnt)) == N.OK
        assert list(idx[:cnt.value]) == partition(n, world, r, mode)
    cnt = C.c_int32()
    assert lib.ugu_b200_partition(4, 2, 2, 0, None, C.byref(cnt)) != N.OK      # rank out of range
    assert lib.ugu_b200_partition(4, 2, 0, 7, None, C.byref(cnt)) != N.OK      # unknown mode


def test_multiview_create_fails_loudly_without_a_gpu():
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from ugu_b200.multiview import NativeMultiView
    with pytest.raises(RuntimeError):
        NativeMultiView([0])


@pytest.mark.gpu
@pytest.mark.parametrize("transport", ["peer", "auto"])
def test_native_multiview_equals_single_renders(transport):
    """The single-process multi-GPU dispatch behind the C-ABI: every visible GPU (two handles on one GPU when there
    is only one) against per-view Render calls of one rasterizer — batch outputs, gathered face_id maps and the
    reduced per-face statistics, both partitions."""
    import numpy as np
    import cases
    import util
    from ugu_b200 import DiffuseColor, PinholeCamera, RendererCpuOption, meshgen
    from ugu_b200.multiview import NativeMultiView
    ndev = torch.cuda.device_count()
    devices = list(range(ndev)) if ndev > 1 else [0, 0]
    mesh = cases.load_spot()
    opt = RendererCpuOption(diffuse_color=DiffuseColor.kTexture)
    c, ext = cases.bbox_center_and_extent(mesh)
    cams = [PinholeCamera(200, 152, p, fov_y_deg=45.0) for p in meshgen.orbit_poses(c, 1.5 * ext, 13)]
    run = util.CudaRunner(mesh, opt)
    singles = [run.render(cm, opt) for cm in cams]
    mv = NativeMultiView(devices, opt, transport=transport)
    assert mv.size == len(devices)
    mv.prepare_mesh(mesh)
    F = len(mesh.vertex_indices)
    pix = np.zeros(F, np.int64); views = np.zeros(F, np.int64); bestc = np.zeros(F, np.int64); bestv = np.full(F, -1)
    for i, s in enumerate(singles):
        cnt = np.bincount(s["face_id"][s["face_id"] >= 0].ravel(), minlength=F)
        pix += cnt
        views += cnt > 0
        better = cnt > bestc
        bestv[better] = i
        bestc[better] = cnt[better]
    for mode in ("contiguous", "round_robin"):
        got = mv.render_batch(cams, mode=mode)
        for i, s in enumerate(singles):
            for k in cases.OUTPUT_KEYS:
                assert np.array_equal(util.bits(got[k][i]), util.bits(s[k])), (mode, i, k)
        for root in (0, len(devices) - 1):
            fid = mv.gather_face_ids(cams, root=root, chunk_views=3, mode=mode)
            for i, s in enumerate(singles):
                assert np.array_equal(fid[i], s["face_id"]), (mode, root, i)
        pc, vc, bv, bc = mv.face_visibility(cams, root=len(devices) - 1, chunk_views=4, mode=mode)
        assert np.array_equal(pc, pix) and np.array_equal(vc, views)
        assert np.array_equal(bc, bestc) and np.array_equal(bv, bestv)
