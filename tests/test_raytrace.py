"""Raytracer parity mode: ugu::Raytracer::Render (src/renderer/cpu/raytracer.cc:117-225)
with the nanort nearest-hit query (include/ugu/accel/bvh_nanort.h:109-152, third_party/nanort/nanort.h).
Fixtures tests/golden/raytrace_*.npz come from the COMPILED reference Raytracer (make_golden_raytrace.py).

CPU: the brute-force C restatement (oracle_raytrace) against every fixture, bit for bit on all five
outputs.  GPU: the CUDA path through the C-ABI against the same fixtures and the restatement."""
import numpy as np
import pytest

import cases
import util
from oracle import OraclePort, port_raytrace, ref_available
from ugu_b200 import ColorInterpolation, DiffuseColor, PinholeCamera, RendererCpuOption, meshgen

NAMES = sorted(p.stem[len("raytrace_"):] for p in cases.GOLDEN_DIR.glob("raytrace_*.npz"))


def _load(name):
    mesh, cam, opt, _, _ = util.load_golden(name)
    d = np.load(cases.GOLDEN_DIR / f"raytrace_{name}.npz")
    opt.depth_scale = float(d["depth_scale"])
    return mesh, cam, opt, {k: d[f"out_{k}"] for k in cases.OUTPUT_KEYS}


def _org(cam):
    return np.asarray(cam.c2w(), np.float64)[:3, 3].astype(np.float32)  # camera.h:431-436 (c2w_t_f_)


def _assert_same(got, want, what):
    for k in cases.OUTPUT_KEYS:
        a, b = np.ascontiguousarray(got[k]).view(np.uint8), np.ascontiguousarray(want[k]).view(np.uint8)
        bad = int((a != b).sum())
        assert bad == 0, f"{what}: {k} differs in {bad} bytes"


def test_fixtures_exist():
    assert len(NAMES) >= 25


@pytest.mark.parametrize("name", NAMES)
def test_port_reproduces_reference_raytracer(name):
    mesh, cam, opt, want = _load(name)
    if len(mesh.vertex_indices) * cam.width() * cam.height() > 150e6:
        pytest.skip("brute force too slow for the CPU suite; covered by the GPU test against the fixture")
    got = port_raytrace(OraclePort(mesh), cam.to_native(), _org(cam), opt)
    assert got is not None
    _assert_same(got, want, name)


def test_raytracer_differs_from_rasterizer_where_it_should():
    """Semantics that distinguish the two CPU renderers: face 0 IS shaded, a back face in front of a
    front face hides it (then is culled), depth_scale is applied."""
    mesh, cam, opt, want = _load("kat_face0_quirk_cull")
    assert (want["face_id"] == 0).sum() > 0 and (want["mask"][want["face_id"] == 0] == 255).all()
    _, _, _, scaled = _load("kat_face0_quirk_nocull")
    vis = want["face_id"] >= 0
    assert np.allclose(scaled["depth"][vis], want["depth"][vis] * 1000.0, rtol=1e-6)
    _, _, _, ras, _ = util.load_golden("kat_face0_quirk_cull")
    assert (ras["mask"][ras["face_id"] == 0] == 0).all()  # the Rasterizer's `fid > 0` quirk


@pytest.mark.skipif(not ref_available(), reason="oracle/_ref not built here")
def test_port_vs_live_reference_orbit():
    from oracle import OracleRef
    spot = cases.load_spot()
    c, ext = cases.bbox_center_and_extent(spot)
    ref = OracleRef(spot)
    opt = RendererCpuOption(diffuse_color=DiffuseColor.kTexture, depth_scale=3.0)
    for k in (0, 2):
        cam = PinholeCamera(96, 72, meshgen.orbit_poses(c, 1.5 * ext, 4)[k], fov_y_deg=45.0)
        _assert_same(port_raytrace(OraclePort(spot), cam.to_native(), _org(cam), opt), ref.raytrace(cam, opt),
                     f"orbit {k}")


@pytest.mark.skipif(not ref_available(), reason="oracle/_ref not built here")
@pytest.mark.parametrize("seed", [31, 32, 33, 34])
def test_port_vs_live_reference_random_soups(seed):
    """Random triangle soups (large + sub-pixel triangles, flipped / unnormalised face normals, two
    textures) under every option set: the brute-force restatement against the reference's BVH query."""
    from oracle import OracleRef
    mesh = cases.random_soup(250, seed)
    ref, port = OracleRef(mesh), OraclePort(mesh)
    cam = cases.kat_camera(80, 60)
    for tag, opt in cases.ALL_OPTIONS:
        opt = RendererCpuOption(**{**opt.__dict__, "depth_scale": 1.0 + seed})
        _assert_same(port_raytrace(port, cam.to_native(), _org(cam), opt), ref.raytrace(cam, opt), f"soup {seed} {tag}")


# ---------------------------------------------------------------------------------------------- GPU
def _cuda_raytrace(mesh, cam, opt, want=cases.OUTPUT_KEYS, pinned=True):
    from ugu_b200 import Image1b, Image1f, Image1i, Image3b, Image3f, Raytracer
    r = Raytracer(opt)
    r.set_mesh(mesh)
    assert r.PrepareMesh()
    r.set_camera(cam)
    imgs = dict(color=Image3b(pinned), depth=Image1f(pinned), normal=Image3f(pinned), mask=Image1b(pinned),
                face_id=Image1i(pinned))
    assert r.Render(**{k: (imgs[k] if k in want else None) for k in cases.OUTPUT_KEYS})
    return r, {k: imgs[k].data.copy() for k in want}


@pytest.mark.gpu
@pytest.mark.parametrize("name", NAMES)
def test_cuda_raytracer_vs_reference_fixture(name):
    mesh, cam, opt, want = _load(name)
    _, got = _cuda_raytrace(mesh, cam, opt)
    _assert_same(got, want, name)


@pytest.mark.gpu
@pytest.mark.parametrize("seed", [41, 42, 43])
def test_cuda_raytracer_vs_port_random_soups(seed):
    """2000-triangle soups (10 % large: the queued path) under every option set, odd image sizes."""
    from ugu_b200 import Image1b, Image1f, Image1i, Image3b, Image3f, Raytracer
    mesh = cases.random_soup(2000, seed)
    port = OraclePort(mesh)
    cam = cases.kat_camera(131, 77)
    r = Raytracer()
    r.set_mesh(mesh)
    assert r.PrepareMesh()
    r.set_camera(cam)
    imgs = (Image3b(), Image1f(), Image3f(), Image1b(), Image1i())
    for tag, opt in cases.ALL_OPTIONS:
        opt = RendererCpuOption(**{**opt.__dict__, "depth_scale": 0.5 * seed})
        r.set_option(opt)
        assert r.Render(*imgs)
        got = {k: im.data for k, im in zip(cases.OUTPUT_KEYS, imgs)}
        _assert_same(got, port_raytrace(port, cam.to_native(), _org(cam), opt), f"soup {seed} {tag}")
    assert r.stats()["large_triangles"] > 0


@pytest.mark.gpu
def test_cuda_raytracer_subsets_reuse_and_rasterizer_interleave():
    """Single outputs, pageable destinations, repeated calls and a Rasterizer Render in between on the
    same handle's buffers (the nearest-hit buffer is left clean by every call)."""
    from ugu_b200 import Image1f, Image1i
    mesh, cam, opt, want = _load("soup_tex_nn_nocull")
    r, got = _cuda_raytrace(mesh, cam, opt, want=("depth", "face_id"), pinned=False)
    for k in ("depth", "face_id"):
        assert np.array_equal(got[k].view(np.uint8), want[k].view(np.uint8)), k
    fid = Image1i()
    assert r.RenderFaceId(fid) and np.array_equal(fid.data, want["face_id"])
    dep = Image1f()
    assert r.RenderDepth(dep) and np.array_equal(dep.data.view(np.uint32), want["depth"].view(np.uint32))
    # the plain Rasterizer on the same library afterwards is unaffected
    run = util.CudaRunner(mesh, opt)
    ras = run.render(cam, opt)
    util.assert_parity(ras, OraclePort(mesh).render(cam.to_native(), opt), strict=True, what="rasterizer after rt")


@pytest.mark.gpu
def test_cuda_raytracer_near_plane_and_large_triangles():
    """Triangles that straddle the camera plane (the Rasterizer drops them, a ray caster sees their visible
    part), a screen-filling triangle (queued path) and tiny ones, against the brute-force restatement."""
    v = np.array([[-1, -1, 3.0], [1, -1, -0.5], [0, 1, 3.0],          # straddles z = 0
                  [-40, -40, 9.0], [40, -40, 9.0], [0, 60, 9.0],       # fills the screen
                  [0.1, 0.1, 2.0], [0.12, 0.1, 2.0], [0.1, 0.12, 2.0],  # sub-pixel
                  [-0.5, 0.2, 1e-3], [0.5, 0.3, 4.0], [-0.2, -0.6, 4.0]], np.float32)  # a vertex almost on the plane
    f = np.arange(12, dtype=np.int32).reshape(-1, 3)
    mesh = cases._tri_mesh(v, f)
    opt = RendererCpuOption(diffuse_color=DiffuseColor.kVertex, backface_culling=False)
    for (w, h) in ((64, 48), (37, 23)):
        cam = cases.kat_camera(w, h)
        want = port_raytrace(OraclePort(mesh), cam.to_native(), _org(cam), opt)
        _, got = _cuda_raytrace(mesh, cam, opt)
        _assert_same(got, want, f"near-plane {w}x{h}")
        assert (want["face_id"] == 0).sum() > 0 and (want["face_id"] == 1).sum() > 0


@pytest.mark.gpu
@pytest.mark.skipif(not ref_available(), reason="oracle/_ref did not travel")
def test_cuda_raytracer_vs_live_reference_larger_views():
    """Sizes the brute-force port cannot do: the compiled reference Raytracer (BVH) on spot at 640x480 and
    an icosphere L6 (81920 faces) at 512x384; all five outputs bit for bit."""
    from oracle import OracleRef
    spot = cases.load_spot()
    c, ext = cases.bbox_center_and_extent(spot)
    cam = PinholeCamera(640, 480, meshgen.orbit_poses(c, 1.5 * ext, 6)[1], fov_y_deg=45.0)
    # nearest-texel colour: with the bilinear colourizer one pixel of this view taps the row past the end of
    # the texture (pixel_shader.h:318-343 does not clamp), where the reference reads whatever the heap holds
    # (this library and the port read 0); bilinear is covered by the fixtures, which stay inside
    opt = RendererCpuOption(diffuse_color=DiffuseColor.kTexture, interp=ColorInterpolation.kNn)
    _, got = _cuda_raytrace(spot, cam, opt)
    _assert_same(got, OracleRef(spot).raytrace(cam, opt), "spot 640x480")
    ico = meshgen.icosphere(6)
    cam = PinholeCamera(512, 384, meshgen.look_at_c2w((0.3, 0.2, -2.7), (0, 0, 0), (0, -1, 0)), fov_y_deg=45.0)
    opt = RendererCpuOption(diffuse_color=DiffuseColor.kVertex, depth_scale=1000.0)
    _, got = _cuda_raytrace(ico, cam, opt)
    _assert_same(got, OracleRef(ico).raytrace(cam, opt), "icosphere L6")


@pytest.mark.gpu
@pytest.mark.skipif(not ref_available(), reason="oracle/_ref did not travel")
def test_cuda_raytracer_full_size_configs_2_and_3_vs_live_reference():
    """BASELINE.json config 2 at full size (1.31 M triangles, vertex colours, 1920x1080) and three views of
    config 3 (spot, 1024x1024, texture, interpolated normals) against the compiled reference Raytracer —
    all five outputs bit for bit.  (Nearest-texel colour on spot: see the note in the 640x480 test.)"""
    from oracle import OracleRef
    ico = meshgen.icosphere(8)
    cam = PinholeCamera(1920, 1080, meshgen.look_at_c2w((0, 0, -2.7), (0, 0, 0), (0, -1, 0)), fov_y_deg=45.0)
    opt = RendererCpuOption(diffuse_color=DiffuseColor.kVertex)
    _, got = _cuda_raytrace(ico, cam, opt, pinned=False)
    want = OracleRef(ico).raytrace(cam, opt)
    assert (want["face_id"] >= 0).sum() > 800_000
    _assert_same(got, want, "config 2")
    del ico, got, want
    spot = cases.load_spot()  # config 3's content: spot, 1024x1024, texture + interpolated normals
    c, ext = cases.bbox_center_and_extent(spot)
    opt = RendererCpuOption(diffuse_color=DiffuseColor.kTexture, interp=ColorInterpolation.kNn)
    ref = OracleRef(spot)
    for k in (0, 7, 20):
        cam = PinholeCamera(1024, 1024, meshgen.orbit_poses(c, 1.5 * ext, 36)[k], fov_y_deg=45.0)
        _, got = _cuda_raytrace(spot, cam, opt, pinned=False)
        _assert_same(got, ref.raytrace(cam, opt), f"config 3 view {k}")


@pytest.mark.gpu
def test_cuda_raytracer_full_size_config4_agrees_with_the_rasterizer():
    """BASELINE.json config 4 (10 M triangles, 4096x4096, texture + interpolated normals): the reference
    Raytracer needs tens of minutes on a uv-sphere of this size (its BVH degenerates at the poles: 577 s
    for a 1 M-triangle uv-sphere at 2048x2048 against 6 s for config 2), so the check is size independent — ray casting
    and rasterization of the same closed mesh must pick the same face on (almost) every pixel, with the
    same depth up to rounding; they may differ only on silhouette / shared-edge pixels."""
    mesh = meshgen.uv_sphere(2237, 2236, tex_size=4096)
    assert mesh.vertex_indices.shape[0] == 9999392
    opt = RendererCpuOption(diffuse_color=DiffuseColor.kTexture)
    cam = PinholeCamera(4096, 4096, meshgen.orbit_poses((0, 0, 0), 2.7, 8)[3], fov_y_deg=45.0)
    _, rt = _cuda_raytrace(mesh, cam, opt, want=("depth", "face_id", "mask"), pinned=False)
    ras = util.CudaRunner(mesh, opt, pinned=False).render(cam, opt, want=("depth", "face_id", "mask"))
    both = (rt["face_id"] >= 0) & (ras["face_id"] >= 0)
    assert both.sum() > 5_000_000
    assert ((rt["face_id"] >= 0) != (ras["face_id"] >= 0)).sum() < 2e-3 * both.sum()
    same = rt["face_id"][both] == ras["face_id"][both]
    assert same.mean() > 0.97, same.mean()
    assert np.allclose(rt["depth"][both][same], ras["depth"][both][same], rtol=1e-4)


@pytest.mark.gpu
def test_cuda_raytracer_far_from_origin_and_tiny_scale():
    """The candidate bbox must stay a superset of nanort's hits when float cancellation makes the projected
    vertices coarse (mesh and camera thousands of units from the origin) and when everything is tiny
    (hits near t_min = 1e-4): compared with the brute-force restatement, which tests every triangle."""
    from ugu_b200 import Mesh
    base = meshgen.icosphere(4)
    opt = RendererCpuOption(diffuse_color=DiffuseColor.kVertex, backface_culling=False)
    for off, scale, dist in (((5000.0, -3000.0, 2000.0), 1.0, 2.7), ((0.0, 0.0, 0.0), 1e-3, 2.7e-3),
                             ((-40000.0, 100.0, 7.0), 30.0, 95.0)):
        off = np.asarray(off, np.float32)
        mesh = Mesh(vertices=(base.vertices * np.float32(scale) + off).astype(np.float32),
                    vertex_indices=base.vertex_indices, face_normals=base.face_normals, normals=base.normals,
                    normal_indices=base.normal_indices, vertex_colors=base.vertex_colors)
        eye = off.astype(np.float64) + np.array([0.3 * dist, 0.2 * dist, -dist])
        cam = PinholeCamera(96, 72, meshgen.look_at_c2w(eye, off.astype(np.float64), (0, -1, 0)), fov_y_deg=45.0)
        want = port_raytrace(OraclePort(mesh), cam.to_native(), _org(cam), opt)
        _, got = _cuda_raytrace(mesh, cam, opt)
        assert (want["face_id"] >= 0).sum() > 500, (off, scale)
        _assert_same(got, want, f"offset {off.tolist()} scale {scale}")


@pytest.mark.gpu
def test_cuda_raytracer_state_machine_and_failures():
    """false + log where the reference returns false: empty mesh at PrepareMesh (raytracer.cc:96-99), Render
    before PrepareMesh / without a camera, all-null outputs, texture colour without materials."""
    from ugu_b200 import Image1i, Image3b, Mesh, Raytracer
    mesh, cam, opt, _ = _load("ico4_vertex_cull")
    r = Raytracer(opt)
    r.set_mesh(Mesh(vertices=np.zeros((0, 3), np.float32), vertex_indices=np.zeros((0, 3), np.int32)))
    assert not r.PrepareMesh()
    r.set_mesh(mesh)
    fid = Image1i()
    r.set_camera(cam)
    assert not r.RenderFaceId(fid)            # set_mesh invalidates PrepareMesh (raytracer.cc:79)
    assert r.PrepareMesh()
    assert not r.Render(None, None, None, None, None)
    assert r.RenderFaceId(fid) and (fid.data >= 0).any()
    r.set_option(RendererCpuOption(diffuse_color=DiffuseColor.kTexture))
    assert not r.RenderColor(Image3b())       # no materials on the icosphere
    assert not r.RenderFaceId(fid)            # ... whichever output is asked for (util_private.cc:31-35)
    r.set_option(opt)
    assert r.RenderFaceId(fid)


@pytest.mark.gpu
def test_cuda_raytracer_device_form():
    import torch
    from ugu_b200 import Raytracer
    mesh, cam, opt, want = _load("ico4_vertex_cull")
    r = Raytracer(opt)
    r.set_mesh(mesh)
    assert r.PrepareMesh()
    P = cam.width() * cam.height()
    dep = torch.empty(P, dtype=torch.float32, device="cuda")
    fid = torch.empty(P, dtype=torch.int32, device="cuda")
    col = torch.empty(P * 3, dtype=torch.uint8, device="cuda")
    st = torch.cuda.Stream()
    r.render_device(cam, color=col.data_ptr(), depth=dep.data_ptr(), face_id=fid.data_ptr(), stream=st.cuda_stream)
    st.synchronize()
    assert np.array_equal(fid.cpu().numpy().reshape(want["face_id"].shape), want["face_id"])
    assert np.array_equal(dep.cpu().numpy().view(np.uint32).reshape(want["depth"].shape), want["depth"].view(np.uint32))
    assert np.array_equal(col.cpu().numpy().reshape(want["color"].shape), want["color"])


@pytest.fixture
def ordered(monkeypatch):
    """Raytracers created inside the test take the occlusion ordering of rt_raster_kernel whatever the mesh size."""
    monkeypatch.setenv("UGU_B200_TWOPASS", "1")


@pytest.mark.gpu
@pytest.mark.parametrize("name", NAMES)
def test_cuda_raytracer_ordered_vs_reference_fixture(ordered, name):
    mesh, cam, opt, want = _load(name)
    _, got = _cuda_raytrace(mesh, cam, opt)
    _assert_same(got, want, "ordered " + name)


@pytest.mark.gpu
@pytest.mark.parametrize("sliver", [False, True], ids=["shaped", "slivers"])
def test_cuda_raytracer_ordered_occluded_debris(ordered, sliver):
    """The wall-and-debris scene of the rasterizer's occlusion tests (hidden, partly visible and degenerate
    triangles behind a camera-facing wall) through the ray-cast mode with the ordering forced."""
    from test_parity_gpu import _wall_and_debris
    from ugu_b200 import ShadingNormal
    mesh = _wall_and_debris(17 + sliver, 12000, sliver)
    port = OraclePort(mesh)
    opt = RendererCpuOption(diffuse_color=DiffuseColor.kVertex, shading_normal=ShadingNormal.kFace,
                            backface_culling=False)
    for (w, h, fov) in ((320, 240, 60.0), (203, 117, 75.0)):
        cam = PinholeCamera(w, h, meshgen.look_at_c2w((0.02, -0.01, 0.0), (0, 0, 2.0), (0, -1, 0)), fov_y_deg=fov)
        _, got = _cuda_raytrace(mesh, cam, opt)
        _assert_same(got, port_raytrace(port, cam.to_native(), _org(cam), opt), f"rt debris {w}x{h}")


@pytest.mark.gpu
def test_cuda_raytracer_render_batch_host_form():
    """Raytracer.render_batch with HOST destinations (view i at slab i) equals per-view Render calls."""
    from ugu_b200 import Raytracer
    mesh, cam, opt, want = _load("ico4_vertex_cull")
    r = Raytracer(opt)
    r.set_mesh(mesh)
    assert r.PrepareMesh()
    cams = [PinholeCamera(cam.width(), cam.height(), p, fov_y_deg=45.0) for p in meshgen.orbit_poses((0, 0, 0), 2.7, 3)]
    h, w = cam.height(), cam.width()
    fid = np.empty((3, h, w), np.int32)
    dep = np.empty((3, h, w), np.float32)
    col = np.empty((3, h, w, 3), np.uint8)
    r.render_batch(cams, color=col.ctypes.data, depth=dep.ctypes.data, face_id=fid.ctypes.data, on_device=False)
    from ugu_b200 import Image1f, Image1i, Image3b
    for i, cm in enumerate(cams):
        r.set_camera(cm)
        c1, d1, f1 = Image3b(), Image1f(), Image1i()
        assert r.Render(color=c1, depth=d1, face_id=f1)
        assert np.array_equal(fid[i], f1.data) and np.array_equal(col[i], c1.data)
        assert np.array_equal(dep[i].view(np.uint32), d1.data.view(np.uint32))


@pytest.mark.gpu
def test_multiview_renderer_with_the_raytracer_and_visibility_fold():
    """The multi-view driver on ray-cast face_id maps (what the reference's own visibility pass uses: BVH
    rays, include/ugu/texturing/visibility_tester.h): batch == per-view renders, and the per-face
    visibility fold of those maps == a histogram of them."""
    import torch
    from oracle import face_visibility
    from ugu_b200 import Image1i, Raytracer
    from ugu_b200.multiview import MultiViewRenderer
    mesh = meshgen.icosphere(5)
    opt = RendererCpuOption(diffuse_color=DiffuseColor.kVertex)
    r = Raytracer(opt)
    r.set_mesh(mesh)
    assert r.PrepareMesh()
    cams = [PinholeCamera(128, 96, c2w, fov_y_deg=45.0) for c2w in meshgen.orbit_poses((0, 0, 0), 2.7, 6)]
    mv = MultiViewRenderer(r)
    idx, outs = mv.render(cams, ("face_id", "depth"))
    torch.cuda.synchronize()
    maps = outs["face_id"].cpu().numpy()
    fid = Image1i()
    for i, cam in zip(idx, cams):
        r.set_camera(cam)
        assert r.RenderFaceId(fid)
        assert np.array_equal(maps[i], fid.data), i
    F = len(mesh.vertex_indices)
    vis = mv.render_face_visibility(cams, F)
    torch.cuda.synchronize()
    want = face_visibility(maps, F)
    for k in ("pixel_count", "view_count", "best_view", "best_count"):
        assert np.array_equal(vis[k].cpu().numpy(), want[k]), k
