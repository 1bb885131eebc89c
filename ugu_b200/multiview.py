"""Multi-view dispatch: a batch of camera views sharded over the GPUs of one box.

Views are independent given a replicated, read-only mesh, so rendering needs NO
data-path collective: rank r renders its own slice of the view list on its own GPU (one process per
GPU, `torch.distributed` only for the rendezvous).  The only exchange step is optional: gathering
the per-view face_id / visibility maps to one rank for a consumer (texturing / SfS-style visibility,
BASELINE.json configs 3 and 5) — NCCL over NVLink on GPUs, gloo in the CPU tests.

The model of the loop being replaced is the reference's sequential orbit loop
(examples/ex09_renderer_realtime.cc:125-182: set_c2w -> Render per view on one renderer).
"""
from __future__ import annotations

from typing import Callable, List, Optional, Sequence

import torch
import torch.distributed as dist


def partition(num_views: int, world: int, rank: int, mode: str = "contiguous") -> List[int]:
    """Indices of the views rank `rank` renders.  contiguous: ranks get ceil/floor(num/world)
    consecutive views (the first num % world ranks get one more); round_robin: i % world == rank."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("bad world/rank")
    if mode == "round_robin":
        return list(range(rank, num_views, world))
    if mode != "contiguous":
        raise ValueError(f"unknown partition mode {mode!r}")
    base, extra = divmod(num_views, world)
    start = rank * base + min(rank, extra)
    return list(range(start, start + base + (1 if rank < extra else 0)))


def _world_rank(group=None):
    if dist.is_available() and dist.is_initialized():
        return dist.get_world_size(group), dist.get_rank(group)
    return 1, 0


def gather_view_maps(local: torch.Tensor, num_views: int, mode: str = "contiguous", dst: int = 0,
                     group=None) -> Optional[torch.Tensor]:
    """Gather per-view maps (n_local, H, W) from every rank to `dst`, reassembled in GLOBAL view
    order -> (num_views, H, W) on dst, None elsewhere.  Ragged shards are padded to the largest
    shard for the collective and trimmed afterwards."""
    world, rank = _world_rank(group)
    counts = [len(partition(num_views, world, r, mode)) for r in range(world)]
    if local.shape[0] != counts[rank]:
        raise ValueError(f"rank {rank} holds {local.shape[0]} views, expected {counts[rank]}")
    if world == 1:
        return local
    cap = max(counts)
    padded = local
    if local.shape[0] < cap:
        pad = torch.zeros((cap - local.shape[0],) + tuple(local.shape[1:]), dtype=local.dtype,
                          device=local.device)
        padded = torch.cat([local, pad], 0)
    padded = padded.contiguous()
    bufs = [torch.empty_like(padded) for _ in range(world)] if rank == dst else None
    dist.gather(padded, bufs, dst=dst, group=group)
    if rank != dst:
        return None
    out = torch.empty((num_views,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    for r in range(world):
        idx = partition(num_views, world, r, mode)
        if idx:
            out[torch.as_tensor(idx, device=local.device)] = bufs[r][: len(idx)]
    return out


class MultiViewRenderer:
    """Shards `cameras` over the ranks of the default process group and renders the local slice
    with one device-resident batch call.

    `render_local(cameras, outputs) -> dict[str, Tensor]` is the per-rank renderer; the default one
    drives ugu_b200.Rasterizer.render_batch on this rank's GPU.  (The CPU tests inject a fake to
    exercise the sharding / gather logic under gloo.)"""

    def __init__(self, rasterizer=None, mode: str = "contiguous",
                 render_local: Optional[Callable] = None, group=None):
        self.rasterizer = rasterizer
        self.mode = mode
        self.group = group
        self._render_local = render_local or self._render_local_cuda

    def _render_local_cuda(self, cameras: Sequence, outputs: Sequence[str]):
        if self.rasterizer is None:
            raise RuntimeError("MultiViewRenderer needs a ugu_b200.Rasterizer (there is no CPU fallback)")
        n = len(cameras)
        if n == 0:
            return {}
        w, h = cameras[0].width(), cameras[0].height()
        P = w * h
        spec = dict(color=(torch.uint8, 3), depth=(torch.float32, 1), normal=(torch.float32, 3),
                    mask=(torch.uint8, 1), face_id=(torch.int32, 1))
        dev = {}
        for k in outputs:
            dt, ch = spec[k]
            shape = (n, h, w) if ch == 1 else (n, h, w, ch)
            dev[k] = torch.empty(shape, dtype=dt, device="cuda")
        stream = torch.cuda.current_stream().cuda_stream
        if (P % 16) == 0 or not ({"color", "mask"} & set(outputs)):
            self.rasterizer.render_batch(list(cameras), on_device=True, stream=stream,
                                         **{k: v.data_ptr() for k, v in dev.items()})
        else:  # odd sizes: per-view pointers would lose their 16-byte alignment -> one view at a time
            for i, cam in enumerate(cameras):
                tmp = {k: torch.empty_like(v[0]) for k, v in dev.items()}
                self.rasterizer.render_device(cam, stream=stream, **{k: v.data_ptr() for k, v in tmp.items()})
                for k in dev:
                    dev[k][i].copy_(tmp[k])
        if stream == 0:
            self.rasterizer.sync()  # handle 0 = the rasterizer's own stream
        return dev

    def local_indices(self, num_views: int) -> List[int]:
        world, rank = _world_rank(self.group)
        return partition(num_views, world, rank, self.mode)

    def render(self, cameras: Sequence, outputs: Sequence[str] = ("face_id",)):
        """Render this rank's slice; returns (global view indices, {output: (n_local, H, W[, C])})."""
        idx = self.local_indices(len(cameras))
        return idx, self._render_local([cameras[i] for i in idx], tuple(outputs))

    def render_and_gather(self, cameras: Sequence, output: str = "face_id", dst: int = 0):
        """Render, then gather one output of every view to rank `dst` in global view order."""
        idx, outs = self.render(cameras, (output,))
        if not idx:
            w, h = cameras[0].width(), cameras[0].height()
            local = torch.empty((0, h, w), dtype=torch.int32,
                                device="cuda" if torch.cuda.is_available() else "cpu")
        else:
            local = outs[output]
        return gather_view_maps(local, len(cameras), self.mode, dst, self.group)
