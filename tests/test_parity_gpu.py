"""GPU suite: the CUDA path (through the C-ABI, driven like a reference caller) against
 (a) the golden fixtures written by the compiled reference,
 (b) the plain-C oracle on seeded inputs, small and at BASELINE.json's full config-2 size,
 (c) the live compiled reference when oracle/_ref travelled to the box,
 (d) size-independent properties (idempotence, output-subset consistency, batch == single).
Bar: mask / face_id bit-exact, depth <= 1e-5 rel, normal <= 1e-4, colour +-1 LSB; the strict checks
additionally demand bit-exact depth/normal/colour (Oren-Nayar colour: +-1 LSB, device libm)."""
import numpy as np
import pytest

import cases
import util
from oracle import OraclePort, OracleRef, ref_available
from ugu_b200 import (DiffuseColor, DiffuseShading, Image3f, PinholeCamera, RendererCpuOption, ShadingNormal)
from ugu_b200 import meshgen

pytestmark = pytest.mark.gpu


def _lsb(opt):
    return 1 if opt.diffuse_shading == DiffuseShading.kOrenNayar else 0


@pytest.mark.parametrize("name", util.golden_names())
def test_cuda_vs_reference_golden(name):
    mesh, cam, opt, want, _ = util.load_golden(name)
    run = util.CudaRunner(mesh, opt)
    got = run.render(cam, opt)
    assert got is not None
    util.assert_parity(got, want, strict=True, color_lsb=_lsb(opt), what=name)
    assert run.r.stats()["hazard_fragments"] == 0


@pytest.mark.parametrize("pinned", [True, False])
def test_pageable_and_pinned_destinations_agree(pinned):
    mesh, cam, opt, want, _ = util.load_golden("soup_tex_bilinear")
    got = util.CudaRunner(mesh, opt, pinned=pinned).render(cam, opt)
    util.assert_parity(got, want, strict=True, what=f"pinned={pinned}")


@pytest.mark.parametrize("seed", [21, 22, 23, 24])
def test_cuda_vs_oracle_random_soup(seed):
    """Large + small triangles, flipped / unnormalised face normals, two textures, every option."""
    mesh = cases.random_soup(3000, seed=seed, big_fraction=0.15)
    rng = np.random.RandomState(seed)
    run = util.CudaRunner(mesh)
    port = OraclePort(mesh)
    for tag, opt in cases.ALL_OPTIONS:
        eye = rng.uniform(-0.3, 0.3, 3)
        c2w = meshgen.look_at_c2w(eye, eye + np.array([rng.uniform(-0.2, 0.2), rng.uniform(-0.2, 0.2), 1.0]),
                                  (0.1, -1.0, 0.05))
        cam = PinholeCamera(int(rng.choice([97, 160, 323])), int(rng.choice([61, 120, 200])), c2w,
                            fov_y_deg=float(rng.uniform(30, 80)))
        want = port.render(cam.to_native(), opt)
        got = run.render(cam, opt)
        util.assert_parity(got, want, strict=True, color_lsb=_lsb(opt), what=f"seed{seed}/{tag}")
    assert run.r.stats()["large_triangles"] > 0  # the tile-job path was exercised


def test_output_subsets_and_single_output_wrappers():
    mesh, cam, opt, want, _ = util.load_golden("spot_front_tex_bilinear")
    run = util.CudaRunner(mesh, opt)
    for subset in (("face_id",), ("depth",), ("mask",), ("normal",), ("color",), ("color", "mask"),
                   ("depth", "normal", "face_id")):
        got = run.render(cam, opt, want=subset)
        util.assert_parity(got, {k: want[k] for k in subset}, strict=True, what=str(subset))
    # all-null is an error, as in util_private.cc:63-67
    assert run.render(cam, opt, want=()) is None


def test_idempotent_and_camera_mutation():
    """Render twice == same bytes (the fused clear leaves no residue); mutate the camera the
    renderer holds between calls like examples/ex02_renderer.cc:142."""
    mesh = cases.load_spot()
    opt = RendererCpuOption(diffuse_color=DiffuseColor.kTexture)
    run = util.CudaRunner(mesh, opt)
    port = OraclePort(mesh)
    poses = np.load(cases.GOLDEN_DIR / "spot_ex02_poses.npy")
    cam = PinholeCamera(320, 240, poses[0], principal_point=(159.3, 127.65), focal_length=(258.65, 258.25))
    first = run.render(cam, opt)
    again = run.render(cam, opt)
    for k in first:
        assert np.array_equal(util.bits(first[k]), util.bits(again[k])), k
    for c2w in poses:
        cam.set_c2w(c2w)
        got = run.render(cam, opt)
        want = port.render(cam.to_native(), opt)
        util.assert_parity(got, want, strict=True, what="ex02 pose")


def test_size_change_and_odd_sizes():
    mesh = meshgen.icosphere(3)
    opt = RendererCpuOption(diffuse_color=DiffuseColor.kVertex)
    run = util.CudaRunner(mesh, opt)
    port = OraclePort(mesh)
    c2w = meshgen.look_at_c2w((0.3, -0.2, -2.2), (0, 0, 0), (0, -1, 0))
    for (w, h) in ((33, 17), (1, 1), (7, 3), (250, 130), (64, 64), (33, 17)):
        cam = PinholeCamera(w, h, c2w, fov_y_deg=55.0)
        util.assert_parity(run.render(cam, opt), port.render(cam.to_native(), opt), strict=True,
                           what=f"{w}x{h}")


def test_empty_and_degenerate_meshes():
    from ugu_b200 import Mesh
    cam = cases.kat_camera()
    opt = RendererCpuOption(diffuse_color=DiffuseColor.kNone, shading_normal=ShadingNormal.kFace,
                            backface_culling=False)
    # no faces at all: validation wants face normals for kFace -> same failure as the reference
    empty = Mesh(vertices=np.zeros((3, 3), np.float32))
    run = util.CudaRunner(empty)
    assert run.render(cam, opt) is None
    # one vertex-only mesh with vertex normals and no faces renders an empty frame
    m = Mesh(vertices=np.zeros((3, 3), np.float32), normals=np.zeros((3, 3), np.float32),
             normal_indices=np.zeros((0, 3), np.int32))
    opt2 = RendererCpuOption(diffuse_color=DiffuseColor.kNone, backface_culling=False)
    got = util.CudaRunner(m).render(cam, opt2, want=("depth", "face_id", "mask"))
    # (normal_indices is empty -> kVertex shading is refused; use the face-less face-normal-free path)
    if got is not None:
        assert (got["face_id"] == -1).all() and (got["depth"] == 0).all() and (got["mask"] == 0).all()


def test_close_up_huge_triangles():
    """Camera almost touching the surface: a few triangles cover the whole frame (tile-job path
    with thousands of jobs per triangle)."""
    mesh = meshgen.icosphere(1)
    opt = RendererCpuOption(diffuse_color=DiffuseColor.kVertex, backface_culling=False)
    c2w = meshgen.look_at_c2w((0.0, 0.0, -1.05), (0.05, 0.02, 0), (0, -1, 0))
    cam = PinholeCamera(640, 360, c2w, fov_y_deg=70.0)
    run = util.CudaRunner(mesh, opt)
    util.assert_parity(run.render(cam, opt), OraclePort(mesh).render(cam.to_native(), opt),
                       strict=True, what="close-up")
    st = run.r.stats()
    assert st["tile_jobs"] > 1000


def test_full_size_config2_icosphere_1080p():
    """BASELINE.json config 2 at full size: 1 310 720-triangle icosphere, vertex colours,
    1920x1080, all five outputs — compared with the oracle directly (it needs < 1 s here)."""
    mesh = meshgen.icosphere(8)
    assert mesh.vertex_indices.shape[0] == 1310720 and mesh.vertices.shape[0] == 655362
    opt = RendererCpuOption(diffuse_color=DiffuseColor.kVertex)
    cam = PinholeCamera(1920, 1080, meshgen.look_at_c2w((0, 0, -2.7), (0, 0, 0), (0, -1, 0)),
                        fov_y_deg=45.0)
    run = util.CudaRunner(mesh, opt)
    got = run.render(cam, opt)
    want = OraclePort(mesh).render(cam.to_native(), opt)
    assert (want["face_id"] >= 0).sum() > 800000
    util.assert_parity(got, want, strict=True, what="config2")
    # config 5 flavour: same mesh, 512x512, face_id only, culling off
    opt5 = RendererCpuOption(backface_culling=False)
    cam5 = PinholeCamera(512, 512, meshgen.random_sphere_poses(3, seed=0)[2], fov_y_deg=45.0)
    util.assert_parity(run.render(cam5, opt5, want=("face_id",)),
                       OraclePort(mesh).render(cam5.to_native(), opt5, want=("face_id",)),
                       strict=True, what="config5")


def test_batch_equals_single_views_device_and_host():
    import ctypes as C
    import torch
    mesh = cases.load_spot()
    opt = RendererCpuOption(diffuse_color=DiffuseColor.kTexture)
    run = util.CudaRunner(mesh, opt)
    c, ext = cases.bbox_center_and_extent(mesh)
    poses = meshgen.orbit_poses(c, 1.5 * ext, 5)
    cams = [PinholeCamera(128, 128, p, fov_y_deg=45.0) for p in poses]
    singles = [run.render(cm, opt) for cm in cams]
    P = 128 * 128
    n = len(cams)
    dev = dict(color=torch.empty(n * P * 3, dtype=torch.uint8, device="cuda"),
               depth=torch.empty(n * P, dtype=torch.float32, device="cuda"),
               normal=torch.empty(n * P * 3, dtype=torch.float32, device="cuda"),
               mask=torch.empty(n * P, dtype=torch.uint8, device="cuda"),
               face_id=torch.empty(n * P, dtype=torch.int32, device="cuda"))
    stream = torch.cuda.current_stream().cuda_stream
    run.r.render_batch(cams, **{k: v.data_ptr() for k, v in dev.items()}, on_device=True, stream=stream)
    torch.cuda.synchronize()
    host = {k: v.cpu().numpy() for k, v in dev.items()}
    shapes = dict(color=(n, 128, 128, 3), depth=(n, 128, 128), normal=(n, 128, 128, 3),
                  mask=(n, 128, 128), face_id=(n, 128, 128))
    for i in range(n):
        for k in cases.OUTPUT_KEYS:
            assert np.array_equal(util.bits(host[k].reshape(shapes[k])[i]), util.bits(singles[i][k])), (i, k)
    # host-destination batch (pinned -> overlapped DMA path)
    from ugu_b200.renderer import pinned_empty
    hb = {k: pinned_empty(shapes[k], singles[0][k].dtype) for k in cases.OUTPUT_KEYS}
    run.r.render_batch(cams, **{k: v.ctypes.data for k, v in hb.items()}, on_device=False)
    for i in range(n):
        for k in cases.OUTPUT_KEYS:
            assert np.array_equal(util.bits(hb[k][i]), util.bits(singles[i][k])), (i, k)


@pytest.mark.parametrize("size,n_views", [((37, 29), 37), ((64, 48), 19), ((33, 1), 3)],
                         ids=lambda v: str(v).replace(" ", ""))
def test_batch_launch_sets_at_ragged_sizes(size, n_views):
    """More views than one launch set holds (16), at image sizes whose per-view slabs lose their 16-byte
    alignment (odd W*H: the kernels fall back to scalar stores for those slabs), device and host forms,
    on a caller stream and on the library's own."""
    import torch
    w, h = size
    mesh = meshgen.icosphere(3)
    opt = RendererCpuOption(diffuse_color=DiffuseColor.kVertex, backface_culling=False)
    run = util.CudaRunner(mesh, opt)
    cams = [PinholeCamera(w, h, p, fov_y_deg=50.0) for p in meshgen.orbit_poses((0, 0, 0), 2.6, n_views)]
    singles = [run.render(cm, opt) for cm in cams]
    P = w * h
    spec = dict(color=(torch.uint8, 3), depth=(torch.float32, 1), normal=(torch.float32, 3),
                mask=(torch.uint8, 1), face_id=(torch.int32, 1))
    for own_stream in (False, True):
        dev = {k: torch.full((n_views * P * ch,), 7, dtype=dt, device="cuda") for k, (dt, ch) in spec.items()}
        torch.cuda.synchronize()
        stream = 0 if own_stream else torch.cuda.current_stream().cuda_stream
        run.r.render_batch(cams, **{k: v.data_ptr() for k, v in dev.items()}, on_device=True, stream=stream)
        run.r.sync()
        torch.cuda.synchronize()
        for k, (dt, ch) in spec.items():
            got = dev[k].cpu().numpy().reshape((n_views,) + singles[0][k].shape)
            for i in range(n_views):
                assert np.array_equal(util.bits(got[i]), util.bits(singles[i][k])), (own_stream, i, k)
    for pinned in (True, False):
        if pinned:
            from ugu_b200.renderer import pinned_empty
            hb = {k: pinned_empty((n_views,) + singles[0][k].shape, singles[0][k].dtype) for k in spec}
        else:
            hb = {k: np.empty((n_views,) + singles[0][k].shape, singles[0][k].dtype) for k in spec}
        run.r.render_batch(cams, **{k: v.ctypes.data for k, v in hb.items()}, on_device=False)
        for k in spec:
            for i in range(n_views):
                assert np.array_equal(util.bits(hb[k][i]), util.bits(singles[i][k])), (pinned, i, k)
    # the scratch is left clean: a single view afterwards is unchanged
    util.assert_parity(run.render(cams[1], opt), singles[1], strict=True, what="single after batch")


def test_calls_on_different_streams_are_ordered():
    """One handle, one set of scratch buffers: back-to-back calls on different streams (and on the
    library's own) must not race on them (each launch set waits for the previous one's event)."""
    import torch
    mesh = meshgen.icosphere(5)
    opt = RendererCpuOption(diffuse_color=DiffuseColor.kVertex)
    run = util.CudaRunner(mesh, opt)
    cams = [PinholeCamera(256, 192, p, fov_y_deg=45.0) for p in meshgen.orbit_poses((0, 0, 0), 2.7, 6)]
    singles = [run.render(cm, opt, want=("face_id", "depth")) for cm in cams]
    streams = [torch.cuda.Stream() for _ in range(3)]
    P = 256 * 192
    fid = [torch.empty(P, dtype=torch.int32, device="cuda") for _ in cams]
    dep = [torch.empty(P, dtype=torch.float32, device="cuda") for _ in cams]
    torch.cuda.synchronize()
    for rep in range(3):
        for i, cm in enumerate(cams):
            st = streams[i % 3].cuda_stream if (i + rep) % 4 else 0
            run.r.render_device(cm, face_id=fid[i].data_ptr(), depth=dep[i].data_ptr(), stream=st)
        run.r.sync()
        torch.cuda.synchronize()
        for i in range(len(cams)):
            assert np.array_equal(fid[i].cpu().numpy().reshape(192, 256), singles[i]["face_id"]), (rep, i)
            assert np.array_equal(util.bits(dep[i].cpu().numpy().reshape(192, 256)), util.bits(singles[i]["depth"])), (rep, i)


@pytest.mark.skipif(not ref_available(), reason="oracle/_ref did not travel to this box")
def test_cuda_vs_live_compiled_reference_spot_640x480():
    """BASELINE.json config 1: spot, 640x480, ex02 intrinsics (r = 1) and six poses, textured
    bilinear, all outputs — against the UNMODIFIED reference compiled from /root/reference."""
    mesh = cases.load_spot()
    opt = RendererCpuOption(diffuse_color=DiffuseColor.kTexture)
    ref = OracleRef(mesh)
    run = util.CudaRunner(mesh, opt)
    poses = np.load(cases.GOLDEN_DIR / "spot_ex02_poses.npy")
    for c2w in poses:
        cam = PinholeCamera(640, 480, c2w, principal_point=(318.6, 255.3), focal_length=(517.3, 516.5))
        util.assert_parity(run.render(cam, opt), ref.render(cam, opt), strict=True, what="config1")


ADAPTER_CHECK = cases.GOLDEN_DIR.parents[0] / "cpp" / "build" / "adapter_check"


@pytest.mark.skipif(not ADAPTER_CHECK.exists(), reason="C++ adapter harness not prebuilt (needs the reference headers)")
def test_cpp_adapter_vs_reference():
    """ugu::RasterizerCuda and the reference ugu::Rasterizer behind the same ugu::RendererCpu*
    (tests/cpp/adapter_check.cc): 4 option sets x 5 orbit views, all five outputs, plus the
    single-output wrappers and RenderW."""
    import subprocess
    p = subprocess.run([str(ADAPTER_CHECK)], capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stdout[-4000:] + p.stderr[-2000:]
    assert "ADAPTER PARITY OK" in p.stdout


def test_multiview_renderer_matches_single_views():
    """ugu_b200.multiview on one rank: the device-resident batch equals per-view Render calls."""
    from ugu_b200.multiview import MultiViewRenderer
    mesh = meshgen.icosphere(4)
    opt = RendererCpuOption(diffuse_color=DiffuseColor.kVertex)
    run = util.CudaRunner(mesh, opt)
    cams = [PinholeCamera(96, 64, p, fov_y_deg=45.0) for p in meshgen.random_sphere_poses(6, seed=3)]
    mv = MultiViewRenderer(run.r)
    idx, outs = mv.render(cams, ("face_id", "depth"))
    assert idx == list(range(6))
    fid = mv.render_and_gather(cams, "face_id").cpu().numpy()
    for i, cm in enumerate(cams):
        single = run.render(cm, opt, want=("face_id", "depth"))
        assert np.array_equal(fid[i], single["face_id"])
        assert np.array_equal(util.bits(outs["depth"][i].cpu().numpy()), util.bits(single["depth"]))


def test_full_size_config3_spot_turntable_1024():
    """BASELINE.json config 3: spot, 36-view turntable at 1024x1024, textured bilinear, all outputs —
    every view against the oracle, rendered through the multi-view batch path."""
    from ugu_b200.multiview import MultiViewRenderer
    mesh = cases.load_spot()
    opt = RendererCpuOption(diffuse_color=DiffuseColor.kTexture)
    run = util.CudaRunner(mesh, opt)
    port = OraclePort(mesh)
    c, ext = cases.bbox_center_and_extent(mesh)
    cams = [PinholeCamera(1024, 1024, p, fov_y_deg=45.0) for p in meshgen.orbit_poses(c, 1.5 * ext, 36)]
    idx, outs = MultiViewRenderer(run.r).render(cams, cases.OUTPUT_KEYS)
    host = {k: v.cpu().numpy() for k, v in outs.items()}
    for i, cam in enumerate(cams):
        want = port.render(cam.to_native(), opt)
        got = {k: host[k][i] for k in cases.OUTPUT_KEYS}
        util.assert_parity(got, want, strict=True, what=f"config3 view {i}")


def test_full_size_config4_10m_triangles_4096():
    """BASELINE.json config 4: 9 999 392-triangle textured mesh at 4096x4096 with UV texture fetch
    (bilinear) and interpolated normals, all outputs, against the oracle at full size."""
    mesh = meshgen.uv_sphere(2237, 2236, tex_size=4096)
    assert mesh.vertex_indices.shape[0] == 9999392
    opt = RendererCpuOption(diffuse_color=DiffuseColor.kTexture)
    cam = PinholeCamera(4096, 4096, meshgen.orbit_poses((0, 0, 0), 2.7, 8)[3], fov_y_deg=45.0)
    run = util.CudaRunner(mesh, opt, pinned=False)
    got = run.render(cam, opt)
    assert got is not None
    want = OraclePort(mesh).render(cam.to_native(), opt)
    assert (want["face_id"] >= 0).sum() > 5_000_000
    util.assert_parity(got, want, strict=True, what="config4")
    assert run.r.stats()["hazard_fragments"] == 0


def test_full_size_config5_random_views_512_face_id():
    """BASELINE.json config 5 (one rank's share, sampled): 1.31 M-triangle mesh, random views at
    512x512, face_id maps only — 12 views against the oracle, and the 128-view device batch against
    per-view renders (batch == single is size independent)."""
    mesh = meshgen.icosphere(8)
    opt = RendererCpuOption()
    run = util.CudaRunner(mesh, opt)
    port = OraclePort(mesh)
    cams = [PinholeCamera(512, 512, p, fov_y_deg=45.0) for p in meshgen.random_sphere_poses(128, seed=0)]
    from ugu_b200.multiview import MultiViewRenderer
    fid = MultiViewRenderer(run.r).render_and_gather(cams, "face_id").cpu().numpy()
    assert fid.shape == (128, 512, 512)
    for i in range(0, 128, 11):
        want = port.render(cams[i].to_native(), opt, want=("face_id",))
        assert np.array_equal(fid[i], want["face_id"]), f"config5 view {i}"
    for i in (5, 77, 127):
        single = run.render(cams[i], opt, want=("face_id",))
        assert np.array_equal(fid[i], single["face_id"])
    # every view sees roughly half of a closed sphere; ids are in range
    assert fid.max() < 1310720 and fid.min() == -1


# ---- two-pass occlusion ordering (raster_kernel PASS 1 / PASS 2) forced on meshes of every size ------------------
@pytest.fixture
def two_pass(monkeypatch):
    """Rasterizers created inside the test take the two-pass path whatever the mesh size (the knob is read
    when the handle is created)."""
    monkeypatch.setenv("UGU_B200_TWOPASS", "1")


@pytest.mark.parametrize("name", util.golden_names())
def test_two_pass_vs_reference_golden(two_pass, name):
    mesh, cam, opt, want, _ = util.load_golden(name)
    run = util.CudaRunner(mesh, opt)
    got = run.render(cam, opt)
    assert got is not None
    util.assert_parity(got, want, strict=True, color_lsb=_lsb(opt), what="two-pass " + name)
    assert run.r.stats()["hazard_fragments"] == 0


@pytest.mark.parametrize("seed", [31, 32])
def test_two_pass_vs_oracle_random_soup(two_pass, seed):
    mesh = cases.random_soup(4000, seed=seed, big_fraction=0.1)
    rng = np.random.RandomState(seed)
    run = util.CudaRunner(mesh)
    port = OraclePort(mesh)
    for tag, opt in cases.ALL_OPTIONS:
        eye = rng.uniform(-0.3, 0.3, 3)
        c2w = meshgen.look_at_c2w(eye, eye + np.array([rng.uniform(-0.2, 0.2), rng.uniform(-0.2, 0.2), 1.0]),
                                  (0.1, -1.0, 0.05))
        cam = PinholeCamera(int(rng.choice([97, 160, 323])), int(rng.choice([61, 120, 200])), c2w,
                            fov_y_deg=float(rng.uniform(30, 80)))
        util.assert_parity(run.render(cam, opt), port.render(cam.to_native(), opt), strict=True,
                           color_lsb=_lsb(opt), what=f"two-pass seed{seed}/{tag}")


def _wall_and_debris(seed, n_debris, sliver):
    """A camera-facing wall of small triangles (pass 1) in front of a cloud of tiny triangles that face away
    (pass 2 groups): well-shaped ones, slivers, zero-area and near-degenerate ones, some poking through the
    wall, some next to its border — what the triangle-level occlusion test of pass 2 has to get right."""
    from ugu_b200 import Mesh
    rng = np.random.RandomState(seed)
    n = 48
    gx, gy = np.meshgrid(np.linspace(-1, 1, n + 1), np.linspace(-1, 1, n + 1))
    wall_v = np.stack([gx.ravel(), gy.ravel(), np.full(gx.size, 2.0) + 0.02 * rng.randn(gx.size)], 1)
    quads = [(j * (n + 1) + i, j * (n + 1) + i + 1, (j + 1) * (n + 1) + i, (j + 1) * (n + 1) + i + 1)
             for j in range(n) for i in range(n)]
    # the camera looks down +z; (a, c, b) winds so that the geometric normal points to -z (towards the camera)
    wall_f = np.array([t for (a, b, c, d) in quads for t in ((a, c, b), (b, c, d))], np.int32)
    ctr = np.stack([rng.uniform(-1.2, 1.2, n_debris), rng.uniform(-1.2, 1.2, n_debris),
                    rng.uniform(1.9, 3.0, n_debris)], 1)
    ext = 0.02 * rng.rand(n_debris, 1, 1) ** 2
    tri = ctr[:, None, :] + ext * rng.randn(n_debris, 3, 3)
    if sliver:
        k = rng.rand(n_debris) < 0.5          # half of them collapse (almost) onto a line
        t = rng.rand(n_debris, 1)
        line = tri[:, 0] * (1 - t) + tri[:, 1] * t
        eps = np.where(rng.rand(n_debris, 1) < 0.3, 0.0, 10.0 ** rng.uniform(-9, -3, (n_debris, 1)))
        tri[k, 2] = (line + eps * rng.randn(n_debris, 3))[k]
    # make the debris face away from the camera (normal.z > 0): swap two corners where needed
    nz = np.cross(tri[:, 1] - tri[:, 0], tri[:, 2] - tri[:, 0])[:, 2]
    flip = nz < 0
    tri[flip, 1], tri[flip, 2] = tri[flip, 2].copy(), tri[flip, 1].copy()
    deb_v = tri.reshape(-1, 3)
    deb_f = np.arange(n_debris * 3, dtype=np.int32).reshape(-1, 3) + len(wall_v)
    verts = np.concatenate([wall_v, deb_v]).astype(np.float32)
    faces = np.concatenate([wall_f, deb_f]).astype(np.int32)
    fn = np.cross(verts[faces[:, 1]] - verts[faces[:, 0]], verts[faces[:, 2]] - verts[faces[:, 0]])
    ln = np.linalg.norm(fn, axis=1, keepdims=True)
    fn = np.where(ln > 0, fn / np.maximum(ln, 1e-30), np.array([[0, 0, 1.0]])).astype(np.float32)
    cols = rng.uniform(0, 255, verts.shape).astype(np.float32)
    return Mesh(vertices=verts, vertex_indices=faces, face_normals=fn, vertex_colors=cols)


@pytest.mark.parametrize("sliver", [False, True], ids=["shaped", "slivers"])
@pytest.mark.parametrize("culling", [True, False], ids=["cull", "nocull"])
def test_two_pass_occluded_debris(two_pass, sliver, culling):
    mesh = _wall_and_debris(7 + sliver, 20000, sliver)
    opt = RendererCpuOption(diffuse_color=DiffuseColor.kVertex, shading_normal=ShadingNormal.kFace,
                            backface_culling=culling)
    run = util.CudaRunner(mesh, opt)
    port = OraclePort(mesh)
    for (w, h, fov) in ((640, 480, 60.0), (203, 117, 75.0)):
        cam = PinholeCamera(w, h, meshgen.look_at_c2w((0.02, -0.01, 0.0), (0, 0, 2.0), (0, -1, 0)), fov_y_deg=fov)
        want = port.render(cam.to_native(), opt)
        got = run.render(cam, opt)
        util.assert_parity(got, want, strict=True, what=f"debris {w}x{h}")
    assert (want["face_id"] >= 0).sum() > 1000


# ---- per-tile triangle lists (tile_kernel), off by default: forced here so the path stays exact ---------------------
@pytest.fixture
def tile_lists(monkeypatch):
    monkeypatch.setenv("UGU_B200_TILES", "1")


@pytest.mark.parametrize("name", [n for n in util.golden_names() if n.startswith(("spot", "soup", "kat"))])
def test_tile_lists_vs_reference_golden(tile_lists, name):
    mesh, cam, opt, want, _ = util.load_golden(name)
    run = util.CudaRunner(mesh, opt)
    got = run.render(cam, opt)
    util.assert_parity(got, want, strict=True, color_lsb=_lsb(opt), what="tile lists " + name)
    # idempotent: the lists are left empty
    again = run.render(cam, opt)
    for k in got:
        assert np.array_equal(util.bits(got[k]), util.bits(again[k])), k


def test_tile_lists_overflow_and_huge_triangles(tile_lists):
    """More than kTileCap triangles on one tile (they fall back to the tile-job queue) next to screen-filling ones
    (more than 3 x 3 tiles: queue), at a size with ragged border tiles."""
    mesh = cases.random_soup(6000, seed=41, big_fraction=0.05)
    opt = RendererCpuOption(diffuse_color=DiffuseColor.kVertex, shading_normal=ShadingNormal.kFace)
    run = util.CudaRunner(mesh, opt)
    port = OraclePort(mesh)
    for (w, h) in ((97, 61), (333, 207)):
        cam = PinholeCamera(w, h, meshgen.look_at_c2w((0.1, 0.0, 0.0), (0.0, 0.1, 1.0), (0, -1, 0)), fov_y_deg=50.0)
        util.assert_parity(run.render(cam, opt), port.render(cam.to_native(), opt), strict=True, what=f"tiles {w}x{h}")


def test_hazard_fragments_are_reported_not_silent():
    """A triangle through the camera plane (z == 0 at a vertex passes the reference's `z < 0` reject and yields
    inf / NaN depths): the call succeeds, and the handle reports the off-contract fragments."""
    from ugu_b200 import Mesh, _native as N
    v = np.array([[-1, -1, 2.0], [1, -1, 2.0], [0, 1, 2.0],          # a regular triangle
                  [-0.5, -0.5, 0.0], [0.5, -0.5, 1.0], [0.0, 0.5, 1.0]], np.float32)   # one vertex ON the camera plane
    f = np.array([[0, 1, 2], [3, 4, 5]], np.int32)
    fn = np.tile(np.array([[0, 0, -1.0]], np.float32), (2, 1))
    mesh = Mesh(vertices=v, vertex_indices=f, face_normals=fn)
    opt = RendererCpuOption(diffuse_color=DiffuseColor.kNone, shading_normal=ShadingNormal.kFace, backface_culling=False)
    run = util.CudaRunner(mesh, opt)
    cam = cases.kat_camera()
    got = run.render(cam, opt, want=("depth", "face_id"))
    assert got is not None
    lib = N.load()
    hz = int(lib.ugu_b200_rasterizer_last_hazards(run.r._h))
    assert hz == run.r.stats()["hazard_fragments"]
    # a clean frame afterwards reports zero again
    mesh2 = Mesh(vertices=v[:3], vertex_indices=f[:1], face_normals=fn[:1])
    run2 = util.CudaRunner(mesh2, opt)
    assert run2.render(cam, opt, want=("depth",)) is not None
    assert int(lib.ugu_b200_rasterizer_last_hazards(run2.r._h)) == 0


# ---- tile jobs folded into the resolve kernel (no binned_kernel launch), forced on meshes that DO queue triangles -----
@pytest.fixture
def folded(monkeypatch):
    monkeypatch.setenv("UGU_B200_FOLD_BINNED", "1")


@pytest.mark.parametrize("name", [n for n in util.golden_names() if n.startswith(("spot", "soup", "kat"))])
def test_folded_tile_jobs_vs_reference_golden(folded, name):
    mesh, cam, opt, want, _ = util.load_golden(name)
    run = util.CudaRunner(mesh, opt)
    for rep in range(2):  # twice: the grid barrier re-arms itself, the queue is left empty
        got = run.render(cam, opt)
        util.assert_parity(got, want, strict=True, color_lsb=_lsb(opt), what=f"folded {name} #{rep}")


def test_folded_tile_jobs_close_up_and_batches(folded):
    """Thousands of tile jobs rasterised by the resolve grid behind its barrier; single views, a 5-view batch
    (per-view barriers) and a size change."""
    import torch
    mesh = meshgen.icosphere(1)
    opt = RendererCpuOption(diffuse_color=DiffuseColor.kVertex, backface_culling=False)
    run = util.CudaRunner(mesh, opt)
    port = OraclePort(mesh)
    cams = [PinholeCamera(640, 360, meshgen.look_at_c2w((0.0, 0.0, -1.05 - 0.1 * i), (0.05, 0.02, 0), (0, -1, 0)),
                          fov_y_deg=70.0) for i in range(5)]
    wants = [port.render(c.to_native(), opt) for c in cams]
    for c, w in zip(cams, wants):
        util.assert_parity(run.render(c, opt), w, strict=True, what="folded close-up")
    assert run.r.stats()["tile_jobs"] > 1000
    P = 640 * 360
    fid = torch.empty(5 * P, dtype=torch.int32, device="cuda")
    dep = torch.empty(5 * P, dtype=torch.float32, device="cuda")
    run.r.render_batch(cams, face_id=fid.data_ptr(), depth=dep.data_ptr(), on_device=True,
                       stream=torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    for i, w in enumerate(wants):
        assert np.array_equal(fid[i * P:(i + 1) * P].cpu().numpy().reshape(360, 640), w["face_id"]), i
        assert np.array_equal(util.bits(dep[i * P:(i + 1) * P].cpu().numpy().reshape(360, 640)), util.bits(w["depth"])), i
    cam2 = PinholeCamera(97, 61, cams[0].c2w(), fov_y_deg=70.0)
    util.assert_parity(run.render(cam2, opt), port.render(cam2.to_native(), opt), strict=True, what="folded small")


# ---- host-form Render: resolve in two bands, the D2H of band 0 under the resolve of band 1 ------------------------
@pytest.fixture
def banded(monkeypatch):
    monkeypatch.setenv("UGU_B200_BANDS", "2")  # band every image of >= 64 chunks, not only the large ones


@pytest.mark.parametrize("name", util.golden_names())
def test_banded_render_vs_reference_golden(banded, name):
    mesh, cam, opt, want, _ = util.load_golden(name)
    run = util.CudaRunner(mesh, opt)  # pinned images: the banded path
    for rep in range(2):
        got = run.render(cam, opt)
        util.assert_parity(got, want, strict=True, color_lsb=_lsb(opt), what=f"banded {name} #{rep}")


def test_banded_render_ragged_sizes_and_subsets(banded):
    """Image sizes whose band boundary / last chunk are ragged, output subsets, large triangles (tile jobs run by
    binned_kernel ahead of band 0), and a pageable destination among pinned ones (falls back to one launch)."""
    mesh = meshgen.icosphere(2)
    opt = RendererCpuOption(diffuse_color=DiffuseColor.kVertex, backface_culling=True)
    port = OraclePort(mesh)
    for (w, h) in ((97, 61), (333, 77), (640, 360), (1021, 513)):
        cam = PinholeCamera(w, h, meshgen.look_at_c2w((0.3, 0.2, -2.2), (0, 0, 0), (0, -1, 0)), fov_y_deg=50.0)
        want = port.render(cam.to_native(), opt)
        run = util.CudaRunner(mesh, opt)
        util.assert_parity(run.render(cam, opt), want, strict=True, what=f"banded {w}x{h}")
        got = run.render(cam, opt, want=("depth", "face_id"))
        util.assert_parity(got, {k: want[k] for k in got}, strict=True, what=f"banded subset {w}x{h}")
        mixed = util.CudaRunner(mesh, opt)
        mixed.imgs["normal"] = Image3f(pinned=False)
        util.assert_parity(mixed.render(cam, opt), want, strict=True, what=f"mixed pinned {w}x{h}")


# ---- occlusion ordering: well-shaped camera-averse patches behind an occluder -------------------------------------------
def test_two_pass_patches_peeking_past_an_occluder(monkeypatch):
    """Well-shaped, camera-averse patches (whole deferred blocks) behind a wall that covers them fully, partly or not
    at all, with the ordering forced on this small mesh: against the oracle."""
    from ugu_b200 import Mesh
    monkeypatch.setenv("UGU_B200_TWOPASS", "1")
    rng = np.random.RandomState(3)

    def grid(n, x0, x1, y0, y1, z, flip):
        gx, gy = np.meshgrid(np.linspace(x0, x1, n + 1), np.linspace(y0, y1, n + 1))
        v = np.stack([gx.ravel(), gy.ravel(), np.full(gx.size, z) + 0.004 * rng.randn(gx.size)], 1)
        q = [(j * (n + 1) + i, j * (n + 1) + i + 1, (j + 1) * (n + 1) + i, (j + 1) * (n + 1) + i + 1)
             for j in range(n) for i in range(n)]
        f = np.array([t for (a, b, c, d) in q for t in (((a, c, b), (b, c, d)) if not flip else ((a, b, c), (b, d, c)))],
                     np.int32)
        return v, f
    parts = [grid(40, -0.6, 0.5, -0.5, 0.5, 2.0, False)]                       # the wall, facing the camera
    for k in range(6):                                                            # patches facing away, behind / beside it
        cx, cy = rng.uniform(-0.9, 0.9, 2)
        parts.append(grid(16, cx - 0.2, cx + 0.2, cy - 0.2, cy + 0.2, rng.uniform(2.2, 3.0), True))
    verts, faces, off = [], [], 0
    for v, f in parts:
        verts.append(v)
        faces.append(f + off)
        off += len(v)
    verts = np.concatenate(verts).astype(np.float32)
    faces = np.concatenate(faces).astype(np.int32)
    fn = np.cross(verts[faces[:, 1]] - verts[faces[:, 0]], verts[faces[:, 2]] - verts[faces[:, 0]])
    fn = (fn / np.linalg.norm(fn, axis=1, keepdims=True)).astype(np.float32)
    mesh = Mesh(vertices=verts, vertex_indices=faces, face_normals=fn,
                vertex_colors=rng.uniform(0, 255, verts.shape).astype(np.float32))
    for culling in (True, False):
        opt = RendererCpuOption(diffuse_color=DiffuseColor.kVertex, shading_normal=ShadingNormal.kFace,
                                backface_culling=culling)
        run = util.CudaRunner(mesh, opt)
        port = OraclePort(mesh)
        for (w, h, fov, eye) in ((1280, 960, 50.0, (0.0, 0.0, 0.0)), (803, 417, 70.0, (0.15, -0.1, 0.2))):
            cam = PinholeCamera(w, h, meshgen.look_at_c2w(eye, (0, 0, 2.0), (0, -1, 0)), fov_y_deg=fov)
            util.assert_parity(run.render(cam, opt), port.render(cam.to_native(), opt), strict=True,
                               what=f"peeking patches cull={culling} {w}x{h}")
