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
