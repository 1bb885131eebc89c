"""Out-of-bounds write check for the device forms of the C-ABI (compute-sanitizer is closed on the GPU
pool, so this is the bounds check we own): every output lives in ONE device arena, each region sized to
the exact byte count the interface promises and surrounded by 0xA5 guard bytes.  After the call the
guards must be untouched and the regions must hold exactly what the host form (already pinned against
the reference / the restatement by the parity tests) returns.  Sizes are chosen to hit the staged
16-byte vector stores' tails: odd pixel counts, single rows, single pixels."""
import numpy as np
import pytest

import cases
from oracle import GBUFFER_FIELDS
from ugu_b200 import (DiffuseColor, GBuffer, Image1b, Image1f, Image1i, Image3b, Image3f, PinholeCamera,
                      RendererCpuOption, meshgen)

GUARD = 4096
FILL = 0xA5
OUTPUTS = [("color", np.uint8, 3), ("depth", np.float32, 1), ("normal", np.float32, 3),
           ("mask", np.uint8, 1), ("face_id", np.int32, 1)]
SIZES = [(160, 120), (37, 29), (33, 1), (1, 1), (1, 47), (255, 3)]


class Arena:
    """One uint8 device buffer: [guard][region 0][pad to 256 + guard][region 1]...[guard]."""

    def __init__(self, sizes):
        import torch
        self.off = {}
        pos = GUARD
        for name, nbytes in sizes.items():
            self.off[name] = (pos, nbytes)
            pos = (pos + nbytes + 255) // 256 * 256 + GUARD
        self.buf = torch.full((pos,), FILL, dtype=torch.uint8, device="cuda")
        self.inside = np.zeros(pos, bool)
        for o, n in self.off.values():
            self.inside[o:o + n] = True

    def ptr(self, name):
        return self.buf.data_ptr() + self.off[name][0]

    def ptrs(self):
        return {n: self.ptr(n) for n in self.off}

    def check(self):
        """-> {name: bytes}; asserts that every byte outside the regions still holds the fill."""
        host = self.buf.cpu().numpy()
        bad = np.flatnonzero((host != FILL) & ~self.inside)
        assert bad.size == 0, f"{bad.size} guard bytes overwritten, first at arena offset {bad[0]} " \
                              f"(regions: {self.off})"
        return {n: host[o:o + k] for n, (o, k) in self.off.items()}


def _case(w, h):
    mesh = meshgen.icosphere(3)
    cam = PinholeCamera(w, h, meshgen.look_at_c2w((0.2, 0.1, -2.2), (0, 0, 0), (0, -1, 0)), fov_y_deg=50.0)
    return mesh, cam, RendererCpuOption(diffuse_color=DiffuseColor.kVertex)


def _host_outputs(r, cam):
    ims = dict(color=Image3b(), depth=Image1f(), normal=Image3f(), mask=Image1b(), face_id=Image1i())
    r.set_camera(cam)
    assert r.Render(**ims)
    return {k: v.data for k, v in ims.items()}


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["raster", "raytrace"])
@pytest.mark.parametrize("size", SIZES, ids=lambda s: f"{s[0]}x{s[1]}")
def test_render_device_stays_inside_its_outputs(mode, size):
    import torch
    from ugu_b200 import Rasterizer, Raytracer
    w, h = size
    mesh, cam, opt = _case(w, h)
    r = (Rasterizer if mode == "raster" else Raytracer)(opt)
    r.set_mesh(mesh)
    assert r.PrepareMesh()
    want = _host_outputs(r, cam)
    P = w * h
    arena = Arena({n: P * ch * np.dtype(dt).itemsize for n, dt, ch in OUTPUTS})
    for _ in range(2):  # twice: the second call runs on the state the first one's resolve reset
        r.render_device(cam, stream=torch.cuda.current_stream().cuda_stream, **arena.ptrs())
    torch.cuda.synchronize()
    got = arena.check()
    for n, dt, ch in OUTPUTS:
        assert np.array_equal(got[n], want[n].reshape(-1).view(np.uint8)), n
    # subsets: an absent output must not be written anywhere either
    for keep in (("face_id",), ("color", "mask"), ("depth", "normal")):
        sub = Arena({n: P * ch * np.dtype(dt).itemsize for n, dt, ch in OUTPUTS if n in keep})
        r.render_device(cam, stream=torch.cuda.current_stream().cuda_stream, **sub.ptrs())
        torch.cuda.synchronize()
        got = sub.check()
        for n in keep:
            assert np.array_equal(got[n], want[n].reshape(-1).view(np.uint8)), (keep, n)


@pytest.mark.gpu
@pytest.mark.parametrize("size", SIZES, ids=lambda s: f"{s[0]}x{s[1]}")
def test_gbuffer_device_stays_inside_its_fields(size):
    import torch
    from ugu_b200 import Rasterizer
    w, h = size
    mesh, cam, opt = _case(w, h)
    r = Rasterizer(opt)
    r.set_mesh(mesh)
    assert r.PrepareMesh()
    r.set_camera(cam)
    gb = GBuffer()
    kw = dict(z_near=1.5, z_far=2.4, geo_id=2)  # the far plane cuts the sphere: erased winners too
    assert r.RenderGBuffer(gb, **kw)
    P = w * h
    arena = Arena({n: P * ch * np.dtype(dt).itemsize for n, dt, ch in GBUFFER_FIELDS})
    r.render_gbuffer_device(cam, kw["z_near"], kw["z_far"], geo_id_value=2,
                            stream=torch.cuda.current_stream().cuda_stream, **arena.ptrs())
    torch.cuda.synchronize()
    got = arena.check()
    for n, dt, ch in GBUFFER_FIELDS:
        assert np.array_equal(got[n], getattr(gb, n).data.reshape(-1).view(np.uint8)), n
    sub = Arena({n: P * ch * np.dtype(dt).itemsize for n, dt, ch in GBUFFER_FIELDS if n in ("uv", "stencil")})
    r.render_gbuffer_device(cam, kw["z_near"], kw["z_far"], geo_id_value=2,
                            stream=torch.cuda.current_stream().cuda_stream, **sub.ptrs())
    torch.cuda.synchronize()
    got = sub.check()
    for n in ("uv", "stencil"):
        assert np.array_equal(got[n], getattr(gb, n).data.reshape(-1).view(np.uint8)), n


@pytest.mark.gpu
def test_batch_device_stays_inside_its_outputs():
    import torch
    from ugu_b200 import Rasterizer, Raytracer
    w, h, n_views = 48, 31, 5  # 48*31 = 1488 = 16 * 93: per-view pointers keep their alignment
    mesh, _, opt = _case(w, h)
    cams = [PinholeCamera(w, h, c2w, fov_y_deg=45.0) for c2w in meshgen.orbit_poses((0, 0, 0), 2.5, n_views)]
    for cls in (Rasterizer, Raytracer):
        r = cls(opt)
        r.set_mesh(mesh)
        assert r.PrepareMesh()
        P = w * h
        arena = Arena({n: n_views * P * ch * np.dtype(dt).itemsize for n, dt, ch in OUTPUTS})
        r.render_batch(cams, on_device=True, stream=torch.cuda.current_stream().cuda_stream, **arena.ptrs())
        torch.cuda.synchronize()
        got = arena.check()
        for i, cam in enumerate(cams):
            want = _host_outputs(r, cam)
            for n, dt, ch in OUTPUTS:
                k = P * ch * np.dtype(dt).itemsize
                assert np.array_equal(got[n][i * k:(i + 1) * k], want[n].reshape(-1).view(np.uint8)), (cls, i, n)


@pytest.mark.gpu
@pytest.mark.parametrize("size", [(160, 120), (37, 29), (33, 1), (1, 1)], ids=lambda s: f"{s[0]}x{s[1]}")
def test_depth_consumers_stay_inside_their_outputs(size):
    """Depth2PointCloud's device form, and the face-visibility fold, on a rendered depth / face_id map."""
    import torch
    from ugu_b200 import Rasterizer
    w, h = size
    mesh, cam, opt = _case(w, h)
    r = Rasterizer(opt)
    r.set_mesh(mesh)
    assert r.PrepareMesh()
    want = _host_outputs(r, cam)
    P = w * h
    F = len(mesh.vertex_indices)
    arena = Arena(dict(depth=P * 4, face_id=P * 4, point_cloud=P * 12, pix=F * 4, views=F * 4, best=F * 8))
    st = torch.cuda.current_stream().cuda_stream
    r.render_device(cam, depth=arena.ptr("depth"), face_id=arena.ptr("face_id"), stream=st)
    r.depth_to_point_cloud_device(cam, arena.ptr("depth"), arena.ptr("point_cloud"), stream=st)
    for name, nbytes in (("pix", F * 4), ("views", F * 4), ("best", F * 8)):
        o = arena.off[name][0]
        arena.buf[o:o + nbytes].zero_()
    r.face_visibility_device(arena.ptr("face_id"), 1, P, F, arena.ptr("pix"), arena.ptr("views"),
                             arena.ptr("best"), 0, st)
    torch.cuda.synchronize()
    got = arena.check()
    pc = Image3f()
    dep = Image1f()
    dep._init(w, h)
    dep.data[...] = want["depth"]
    assert r.Depth2PointCloud(dep, cam, pc)
    assert np.array_equal(got["point_cloud"], pc.data.reshape(-1).view(np.uint8))
    fid = want["face_id"].reshape(-1)
    hist = np.bincount(fid[fid >= 0], minlength=F).astype(np.int32)
    assert np.array_equal(got["pix"].view(np.int32), hist)
    assert np.array_equal(got["views"].view(np.int32), (hist > 0).astype(np.int32))


@pytest.mark.gpu
@pytest.mark.parametrize("size", [(37, 29), (33, 1), (1, 1), (255, 3), (130, 67)], ids=lambda s: f"{s[0]}x{s[1]}")
def test_uv_raster_device_stays_inside_its_texture(size):
    """UV-space attribute raster, device form, at texture sizes that are not multiples of anything."""
    import torch
    import uvcases
    from oracle import port_uv_raster
    from ugu_b200 import Rasterizer, _native as N
    w, h = size
    c = uvcases.spot_case("vertex_u8_256")
    want = port_uv_raster(c["attrs"], c["attr_indices"], c["uv"], c["uv_indices"], w, h, 0, None,
                          np.zeros((h, w), np.uint8))
    dev = {k: torch.from_numpy(np.ascontiguousarray(c[k])).cuda() for k in ("attrs", "attr_indices", "uv", "uv_indices")}
    arena = Arena(dict(texture=w * h * 3, mask=w * h))
    r = Rasterizer()
    d = N.UvRasterDesc(len(c["attrs"]), len(c["uv_indices"]), len(c["uv"]), 0, dev["attrs"].data_ptr(),
                       dev["attr_indices"].data_ptr(), dev["uv"].data_ptr(), dev["uv_indices"].data_ptr(),
                       w, h, 0, 0, 0, 0)
    r.uv_raster_device(d, arena.ptr("texture"), arena.ptr("mask"), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    got = arena.check()
    assert np.array_equal(got["texture"], np.ascontiguousarray(want[0]).reshape(-1).view(np.uint8))
    assert np.array_equal(got["mask"], np.ascontiguousarray(want[1]).reshape(-1).view(np.uint8))
