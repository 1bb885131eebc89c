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

import contextlib
from typing import Callable, List, Optional, Sequence

import torch
import torch.distributed as dist


def bind_rank_to_gpu_numa(device_index: int) -> dict:
    """Pin this process to the host cores that are local to GPU `device_index` (sysfs
    /sys/bus/pci/devices/<bdf>/local_cpulist), BEFORE it allocates pinned host memory: first-touch then
    places the D2H landing buffers on the GPU's own NUMA node, so that N ranks copying their views back
    at the same time do not all write through one socket's memory controllers.  No-op (reported as such)
    when the platform does not expose the topology (numa_node = -1, e.g. a flat VM)."""
    import os
    info = {"bound": False}
    try:
        props = torch.cuda.get_device_properties(device_index)
        bdf = f"{props.pci_domain_id:04x}:{props.pci_bus_id:02x}:{props.pci_device_id:02x}.0"
        base = f"/sys/bus/pci/devices/{bdf}"
        node = int(open(f"{base}/numa_node").read().strip())
        cpulist = open(f"{base}/local_cpulist").read().strip()
        info.update(pci=bdf, numa_node=node, local_cpulist=cpulist)
        cpus = set()
        for part in cpulist.split(","):
            if not part:
                continue
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        allowed = os.sched_getaffinity(0)
        cpus &= allowed
        if node >= 0 and cpus and cpus != allowed:
            os.sched_setaffinity(0, cpus)
            info["bound"] = True
        info["cpus"] = len(cpus) if cpus else len(allowed)
    except Exception as e:  # topology not exposed: leave the affinity alone
        info["error"] = f"{type(e).__name__}: {e}"[:120]
    return info


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


def decode_best(best: torch.Tensor):
    """(best_view, best_count) from the packed accumulator `count << 32 | (0xFFFFFFFF - view)`
    (int64; 0 = never visible -> view -1, count 0)."""
    count = best >> 32
    view = 0xFFFFFFFF - (best & 0xFFFFFFFF)
    view = torch.where(best == 0, torch.full_like(view, -1), view)
    return view.to(torch.int32), count.to(torch.int32)


def reduce_face_visibility(pixel_count: torch.Tensor, view_count: torch.Tensor, best: torch.Tensor,
                           group=None) -> None:
    """Combine per-rank face-visibility accumulators in place on every rank: SUM / SUM / MAX.
    Each rank accumulated over its own (disjoint) views, so this is exactly the statistic over all
    views.  Tensors are int32 / int32 / int64 views of the u32 / u32 / u64 device arrays (counts
    stay below 2^31).  16 bytes per face cross NVLink instead of every face_id map."""
    world, _ = _world_rank(group)
    if world == 1:
        return
    dist.all_reduce(pixel_count, op=dist.ReduceOp.SUM, group=group)
    dist.all_reduce(view_count, op=dist.ReduceOp.SUM, group=group)
    dist.all_reduce(best, op=dist.ReduceOp.MAX, group=group)


def _size(cam):
    """(width, height) of a PinholeCamera or of a prebuilt native camera block (ugu_b200._native.Camera) — callers
    that render the same view list repeatedly convert the cameras once (Camera.to_native) and pass the blocks."""
    w = cam.width
    return (w(), cam.height()) if callable(w) else (int(w), int(cam.height))


OUTPUT_SPEC = dict(color=(torch.uint8, 3), depth=(torch.float32, 1), normal=(torch.float32, 3),
                   mask=(torch.uint8, 1), face_id=(torch.int32, 1))


class MultiViewRenderer:
    """Shards `cameras` over the ranks of the default process group and renders the local slice
    with one device-resident batch call.

    `render_local(cameras, outputs) -> dict[str, Tensor]` is the per-rank renderer; the default one
    drives ugu_b200.Rasterizer.render_batch on this rank's GPU.  (The CPU tests inject a fake to
    exercise the sharding / gather logic under gloo.)"""

    def __init__(self, rasterizer=None, mode: str = "contiguous",
                 render_local: Optional[Callable] = None, group=None,
                 accumulate: Optional[Callable] = None):
        self.rasterizer = rasterizer
        self.mode = mode
        self.group = group
        self._render_local = render_local or self._render_local_cuda
        self._accumulate = accumulate or self._accumulate_cuda
        self._stream = None  # side stream of the CUDA path, see _on_work_stream

    def _render_local_cuda(self, cameras: Sequence, outputs: Sequence[str], into: Optional[dict] = None):
        """`into` = {output: contiguous (n, H, W[, C]) tensor} renders straight into caller storage (e.g. a slice
        of the gather destination)."""
        if self.rasterizer is None:
            raise RuntimeError("MultiViewRenderer needs a ugu_b200.Rasterizer (there is no CPU fallback)")
        n = len(cameras)
        if n == 0:
            return {}
        w, h = _size(cameras[0])
        dev = into if into is not None else {k: torch.empty(self.view_shape(n, h, w, k), dtype=OUTPUT_SPEC[k][0],
                                                            device="cuda") for k in outputs}
        with self._on_work_stream() as stream:
            # view slabs that are not 16-byte aligned (odd W*H) are handled inside the library
            self.rasterizer.render_batch(list(cameras), on_device=True, stream=stream,
                                         **{k: v.data_ptr() for k, v in dev.items()})
        return dev

    @staticmethod
    def view_shape(n: int, h: int, w: int, output: str):
        ch = OUTPUT_SPEC[output][1]
        return (n, h, w) if ch == 1 else (n, h, w, ch)

    @contextlib.contextmanager
    def _on_work_stream(self):
        """The library is never handed stream handle 0 from torch code (to the C-ABI, 0 means "the
        rasterizer's own stream", which torch knows nothing about): the work runs on a real side stream that
        first waits for torch's current stream (allocations, zero fills, earlier consumers) and that the
        current stream waits for afterwards, so whatever follows on torch's side — a collective, a decode,
        a copy to the host — is ordered after the library's kernels."""
        cur = torch.cuda.current_stream()
        if self._stream is None:
            self._stream = torch.cuda.Stream()
        self._stream.wait_stream(cur)
        try:
            yield self._stream.cuda_stream
        finally:
            cur.wait_stream(self._stream)

    def local_indices(self, num_views: int) -> List[int]:
        world, rank = _world_rank(self.group)
        return partition(num_views, world, rank, self.mode)

    def render(self, cameras: Sequence, outputs: Sequence[str] = ("face_id",)):
        """Render this rank's slice; returns (global view indices, {output: (n_local, H, W[, C])})."""
        idx = self.local_indices(len(cameras))
        return idx, self._render_local([cameras[i] for i in idx], tuple(outputs))

    def render_face_visibility(self, cameras: Sequence, num_faces: int, chunk_views: int = 32):
        """Per-face visibility statistics over ALL views: every rank renders the
        face_id maps of its slice `chunk_views` at a time, folds them into device accumulators
        (ugu_b200_face_visibility_device), and the accumulators are all-reduced — no map leaves its
        GPU.  Returns {pixel_count, view_count, best_view, best_count} (int32 tensors, every rank)."""
        idx = self.local_indices(len(cameras))
        dev = "cuda" if self.rasterizer is not None else "cpu"
        pix = torch.zeros(num_faces, dtype=torch.int32, device=dev)
        views = torch.zeros(num_faces, dtype=torch.int32, device=dev)
        best = torch.zeros(num_faces, dtype=torch.int64, device=dev)
        runs = []  # consecutive global view ids -> one accumulate call each
        for i in idx:
            if runs and runs[-1][-1] + 1 == i and len(runs[-1]) < chunk_views:
                runs[-1].append(i)
            else:
                runs.append([i])
        for run in runs:
            outs = self._render_local([cameras[i] for i in run], ("face_id",))
            maps = outs["face_id"]
            self._accumulate(maps, run[0], num_faces, pix, views, best)
        reduce_face_visibility(pix, views, best, self.group)
        bv, bc = decode_best(best)
        return dict(pixel_count=pix, view_count=views, best_view=bv, best_count=bc)

    def _accumulate_cuda(self, maps: torch.Tensor, first_view_id: int, num_faces: int, pix, views, best):
        if self.rasterizer is None:
            raise RuntimeError("face visibility needs a ugu_b200.Rasterizer (there is no CPU fallback)")
        n = maps.shape[0]
        with self._on_work_stream() as stream:
            self.rasterizer.face_visibility_device(
                maps.data_ptr(), n, maps[0].numel(), num_faces, pix.data_ptr(), views.data_ptr(),
                best.data_ptr(), first_view_id, stream)

    def render_and_gather(self, cameras: Sequence, output: str = "face_id", dst: int = 0):
        """Render, then gather one output of every view to rank `dst` in global view order."""
        idx, outs = self.render(cameras, (output,))
        if not idx:
            w, h = _size(cameras[0])
            local = torch.empty((0, h, w), dtype=torch.int32,
                                device="cuda" if torch.cuda.is_available() else "cpu")
        else:
            local = outs[output]
        return gather_view_maps(local, len(cameras), self.mode, dst, self.group)

    def render_and_gather_overlapped(self, cameras: Sequence, output: str = "face_id", dst: int = 0,
                                     chunk_views: int = 16):
        """Render + gather with the exchange hidden under the rendering: every rank renders its contiguous slice
        `chunk_views` at a time on the library's work stream and, as soon as a chunk is done, sends it (NCCL
        send / recv, which run on NCCL's own stream) straight into its final place in the view-ordered
        destination on `dst` — no padded staging buffers, no reordering pass on the root; the root renders its
        own views in place.  Returns the (num_views, H, W[, C]) tensor on `dst`, None elsewhere.  Round-robin
        partitions fall back to render_and_gather."""
        world, rank = _world_rank(self.group)
        if world == 1 or self.mode != "contiguous":
            return self.render_and_gather(cameras, output, dst)
        num = len(cameras)
        w, h = _size(cameras[0])
        dt = OUTPUT_SPEC[output][0]
        device = "cuda" if self.rasterizer is not None else "cpu"
        parts = [partition(num, world, r, self.mode) for r in range(world)]
        rounds = max((len(p) + chunk_views - 1) // chunk_views for p in parts)
        out = torch.empty(self.view_shape(num, h, w, output), dtype=dt, device=device) if rank == dst else None
        works, keep = [], []
        mine = parts[rank]
        for k in range(rounds):
            lo, hi = k * chunk_views, min((k + 1) * chunk_views, len(mine))
            ops = []
            if lo < hi:
                idx = mine[lo:hi]
                if rank == dst:  # in place
                    self._render_chunk([cameras[i] for i in idx], output, out[idx[0]: idx[-1] + 1])
                else:
                    buf = torch.empty(self.view_shape(len(idx), h, w, output), dtype=dt, device=device)
                    self._render_chunk([cameras[i] for i in idx], output, buf)
                    keep.append(buf)
                    ops.append(dist.P2POp(dist.isend, buf, dst, group=self.group))
            if rank == dst:
                for r in range(world):
                    if r == dst:
                        continue
                    plo, phi = k * chunk_views, min((k + 1) * chunk_views, len(parts[r]))
                    if plo < phi:
                        ops.append(dist.P2POp(dist.irecv, out[parts[r][plo]: parts[r][phi - 1] + 1], r,
                                              group=self.group))
            if ops:
                works.extend(dist.batch_isend_irecv(ops))
        for wk in works:
            wk.wait()
        return out

    def _render_chunk(self, cameras: Sequence, output: str, into: torch.Tensor) -> None:
        if self._render_local == self._render_local_cuda:
            self._render_local_cuda(cameras, (output,), into={output: into})
        else:  # injected renderer (CPU tests)
            into.copy_(self._render_local(cameras, (output,))[output])


class NativeMultiView:
    """Thin binding of the single-process multi-GPU dispatch behind the C-ABI (ugu_b200_multiview_*,
    include/ugu_b200_raster.h): one rasterizer handle per listed device, views sharded over them, the optional
    gather / reduction over NCCL or peer copies inside the library.  This is what a C++ host uses
    (ugu::MultiViewRasterizerCuda); the torch.distributed classes above are the one-process-per-GPU form."""

    PARTITION = {"contiguous": 0, "round_robin": 1}

    def __init__(self, devices: Sequence[int], option=None, transport: str = "auto"):
        import ctypes as C
        from . import _native as N
        from .renderer import RendererCpuOption
        self._C, self._N = C, N
        self._lib = N.load()
        devs = (C.c_int32 * len(devices))(*devices)
        h = C.c_void_p()
        rc = self._lib.ugu_b200_multiview_create(devs, len(devices), {"auto": 0, "peer": 1, "nccl": 2}[transport], C.byref(h))
        if rc != N.OK:
            msg = self._lib.ugu_b200_multiview_last_error(None)
            raise RuntimeError(msg.decode() if msg else "ugu_b200_multiview_create failed")
        self._h = h
        self._option = option or RendererCpuOption()
        self._num_faces = 0

    def __del__(self):
        h = getattr(self, "_h", None)
        if h:
            self._lib.ugu_b200_multiview_destroy(h)
            self._h = None

    def _check(self, rc):
        if rc != self._N.OK:
            msg = self._lib.ugu_b200_multiview_last_error(self._h)
            raise RuntimeError(msg.decode() if msg else f"status {rc}")

    @property
    def size(self) -> int:
        return int(self._lib.ugu_b200_multiview_size(self._h))

    @property
    def transport(self) -> str:
        return "nccl" if self._lib.ugu_b200_multiview_transport(self._h) == 2 else "peer"

    def set_option(self, option) -> None:
        self._option = option

    def prepare_mesh(self, mesh) -> None:
        from .renderer import mesh_desc
        d, keep = mesh_desc(mesh)
        self._check(self._lib.ugu_b200_multiview_upload_mesh(self._h, self._C.byref(d)))
        self._num_faces = int(d.num_faces)
        del keep

    def _begin(self, cameras):
        N, C = self._N, self._C
        o = self._option
        opt = N.Option(int(o.diffuse_color), int(o.interp), int(o.shading_normal), int(o.diffuse_shading),
                       1 if o.backface_culling else 0, float(o.depth_scale), float(o.oren_nayar_sigma))
        self._check(self._lib.ugu_b200_multiview_set_option(self._h, C.byref(opt)))
        natives = [c if isinstance(c, N.Camera) else c.to_native() for c in cameras]
        return (N.Camera * len(natives))(*natives), int(natives[0].width), int(natives[0].height)

    def render_batch(self, cameras, outputs=("color", "depth", "normal", "mask", "face_id"), mode="contiguous"):
        """-> dict of view-ordered host arrays (n, H, W[, 3])."""
        import numpy as np
        from .renderer import pinned_empty
        cams, w, h = self._begin(cameras)
        n = len(cameras)
        spec = dict(color=(np.uint8, (n, h, w, 3)), depth=(np.float32, (n, h, w)), normal=(np.float32, (n, h, w, 3)),
                    mask=(np.uint8, (n, h, w)), face_id=(np.int32, (n, h, w)))
        host = {k: pinned_empty(spec[k][1], spec[k][0]) for k in outputs}
        o = self._N.Outputs(*[host[k].ctypes.data if k in host else None
                              for k in ("color", "depth", "normal", "mask", "face_id")])
        self._check(self._lib.ugu_b200_multiview_render_batch(self._h, n, cams, self.PARTITION[mode],
                                                              self._C.byref(o)))
        return host

    def gather_face_ids(self, cameras, root=0, chunk_views=0, mode="contiguous", to_host=True):
        """Render + gather of the face_id maps on device slot `root`; -> (n, H, W) int32 host array, or None with
        to_host=False (the maps then stay in the library's buffer on the root device: timing runs)."""
        import numpy as np
        cams, w, h = self._begin(cameras)
        n = len(cameras)
        out = np.empty((n, h, w), np.int32) if to_host else None
        self._check(self._lib.ugu_b200_multiview_gather_face_ids(
            self._h, n, cams, self.PARTITION[mode], root, chunk_views, None, out.ctypes.data if to_host else None))
        return out

    def face_visibility(self, cameras, root=0, chunk_views=0, mode="contiguous"):
        """-> pixel_count, view_count, best_view, best_count per face over all views (reduced on `root`)."""
        import numpy as np
        cams, w, h = self._begin(cameras)
        F = self._num_faces
        pc, vc, bc = (np.zeros(F, np.uint32) for _ in range(3))
        bv = np.full(F, -1, np.int32)
        self._check(self._lib.ugu_b200_multiview_reduce_face_visibility(
            self._h, len(cameras), cams, self.PARTITION[mode], root, chunk_views, pc.ctypes.data, vc.ctypes.data,
            bv.ctypes.data, bc.ctypes.data))
        return pc, vc, bv, bc
