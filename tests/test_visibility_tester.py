"""Keyframe visibility pass: ugu::VisibilityTester with the BVH query replaced by the
rasterizer's own depth map.

Pinned against the COMPILED REFERENCE tester (fixtures tests/golden/visibility_*.npz, made by
tests/golden/make_golden_visibility.py from oracle/_ref): on the (vertex, keyframe) / (face, keyframe) pairs
BOTH methods call visible the records must agree — projected position, colour, intensity, viewing angle and
projected area bit for bit (CPU restatement; the CUDA path's viewing angle within a few ulp: device acosf),
distance within the tester's own eps and, for all but the thin-geometry cases, to float tolerance — and the
arg-best keyframe ids of FaceInfo::CalcStat exactly wherever the visible sets agree.  The DISAGREEMENT SET
(sub-pixel BVH rays vs the nearest-pixel depth map at silhouettes and thin parts) is measured and bounded.
"""
import ctypes as C

import numpy as np
import pytest

import cases
import oracle
from oracle import ref_available
from ugu_b200 import _native as N

FIXTURES = ("spot_orbit5", "ico4_random4")
MAX_DISAGREE = 0.02     # fraction of the (element, keyframe) decisions that may differ from the BVH tester
ULP_ANGLE = 4           # device acosf vs glibc acosf


def load(name):
    d = np.load(cases.GOLDEN_DIR / f"visibility_{name}.npz")
    ref = str(d["mesh_ref"])
    mesh = cases.load_spot() if ref else cases.mesh_from_arrays(d)
    return d, mesh


def cameras(d):
    from ugu_b200 import PinholeCamera
    w, h = int(d["cam_size"][0]), int(d["cam_size"][1])
    return [PinholeCamera(w, h, c2w, principal_point=d["cam_pp"], focal_length=d["cam_f"]) for c2w in d["cam_c2w"]]


def port_run(d, mesh, facewise, interp):
    """The C restatement on every keyframe of a fixture -> dense arrays shaped like the reference's."""
    K = len(d["ids"])
    V, F = len(mesh.vertices), len(mesh.vertex_indices)
    nrm = np.asarray(mesh.normals, np.float32)
    vo = oracle.vis_option(facewise=facewise, interp=interp)
    w, h = int(d["cam_size"][0]), int(d["cam_size"][1])
    out = dict(v_visible=np.zeros((K, V), np.uint8), v_rec=np.zeros((K, V), oracle.VIS_VERTEX_DTYPE),
               f_visible=np.zeros((K, F), np.uint8), f_area=np.zeros((K, F), np.float32),
               f_viewing_angle=np.zeros((K, F), np.float32), f_distance=np.zeros((K, F), np.float32))
    for k in range(K):
        vc = oracle.vis_camera((d["ref_w2c"][k], d["ref_c2w_R"][k], d["ref_c2w_t"][k]), w, h,
                               float(d["cam_f"][0]), float(d["cam_f"][1]), float(d["cam_pp"][0]), float(d["cam_pp"][1]))
        p = oracle.port_visibility_keyframe(mesh.vertices, nrm, mesh.vertex_indices, vc, vo, d["color"][k],
                                            d["depth"][k], d["mask"][k], d["zbuf"][k])
        assert p is not None
        for key in out:
            out[key][k] = p[key]
    return out


def bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


def ulp_diff(a, b):
    return np.abs(bits(a).astype(np.int64) - bits(b).astype(np.int64))


def check_against_reference(d, got, facewise, interp, angle_ulp, what):
    """`got`: dict like port_run's.  Returns the measured disagreement numbers."""
    p = f"ref_fw{int(facewise)}_i{interp}_"
    rv, gv = d[p + "v_visible"].astype(bool), got["v_visible"].astype(bool)
    rf, gf = d[p + "f_visible"].astype(bool), got["f_visible"].astype(bool)
    both_v, both_f = rv & gv, rf & gf
    dis_v = (rv ^ gv).sum() / max(1, (rv | gv).sum())
    dis_f = (rf ^ gf).sum() / max(1, (rf | gf).sum())
    assert dis_v <= MAX_DISAGREE and dis_f <= 2 * MAX_DISAGREE, (what, dis_v, dis_f)
    rec = got["v_rec"]
    proj_key = "projected" if "projected" in rec.dtype.names else "projected_pos"
    # vertex records: in the face-wise mode a vertex keeps the record of the first passing (face, corner); one
    # upstream disagreement can hand a vertex another corner's offset, so a few records may differ there
    same_proj = (bits(d[p + "v_projected"])[both_v] == bits(rec[proj_key])[both_v]).all(-1)
    if not facewise:
        assert same_proj.all(), what
    else:
        assert (~same_proj).mean() <= 0.01, (what, (~same_proj).sum())
    sel = np.zeros_like(both_v)
    sel[both_v] = same_proj
    assert np.array_equal(d[p + "v_color"][sel], rec["color"][sel]), what
    assert np.array_equal(d[p + "v_intensity"][sel], rec["intensity"][sel]), what
    assert ulp_diff(d[p + "v_viewing_angle"][sel], rec["viewing_angle"][sel]).max() <= angle_ulp, what
    # distance: the BVH reports the nearest hit, which is nearer than the vertex when a thin part lies within eps
    dr, dg = d[p + "v_distance"][sel], rec["distance"][sel]
    finite = dr < 1e30  # the reference leaves FLT_MAX when its ray hits nothing (visibility_tester.cc:459-477)
    err = np.abs(dr[finite] - dg[finite])
    assert err.max() <= 0.1 + 1e-5, (what, err.max())
    assert (err > 1e-4 * np.maximum(dg[finite], 1.0)).mean() <= 0.03, what
    # face records on faces both see and whose three corner records are the same ones
    tri = mesh_faces = None
    f_ok = both_f.copy()
    assert ulp_diff(d[p + "f_viewing_angle"][f_ok], got["f_viewing_angle"][f_ok]).max() <= max(angle_ulp, 1) * 2 \
        or facewise, what
    if not facewise:
        assert np.array_equal(bits(d[p + "f_area"][f_ok]), bits(got["f_area"][f_ok])), what
    else:
        assert (bits(d[p + "f_area"][f_ok]) != bits(got["f_area"][f_ok])).mean() <= 0.01, what
    return dict(vertex_disagreement=float(dis_v), face_disagreement=float(dis_f),
                vertex_pairs=int((rv | gv).sum()), face_pairs=int((rf | gf).sum()),
                thin_geometry_distance_pairs=int((err > 1e-4 * np.maximum(dg[finite], 1.0)).sum()))


def fold_faces(f_visible, area, angle, dist, ids):
    """FaceInfo::CalcStat (visibility_tester.cc:185-218): first max / first min over the keyframes in order."""
    K, F = f_visible.shape
    out = np.full((F, 3), -1, np.int32)
    vis = f_visible.astype(bool)
    a = np.where(vis, area, -np.inf)
    g = np.where(vis, angle, np.inf)
    dd = np.where(vis, dist, np.inf)
    any_ = vis.any(0)
    ids = np.asarray(ids, np.int32)
    out[any_, 0] = ids[np.argmax(a, 0)][any_]
    out[any_, 1] = ids[np.argmin(g, 0)][any_]
    out[any_, 2] = ids[np.argmin(dd, 0)][any_]
    return out


def compare_face_ids(d, p, got, ours, min_same):
    """max_area_id / min_viewing_angle_id must equal the reference's on every face whose visible set, areas and
    angles are the reference's bit for bit; min_distance_id on those whose best distance is clear of the runner-up
    by more than the float noise between the BVH's t and |camera_p|."""
    vis = d[p + "f_visible"].astype(bool)
    same = (d[p + "f_visible"] == got["f_visible"]).all(0) & vis.any(0)
    same &= ((bits(d[p + "f_area"]) == bits(got["f_area"])) | ~vis).all(0)
    same &= ((bits(d[p + "f_viewing_angle"]) == bits(got["f_viewing_angle"])) | ~vis).all(0)
    assert same.sum() >= min_same * vis.any(0).sum(), (same.sum(), vis.any(0).sum())
    assert np.array_equal(ours[same][:, :2], d[p + "face_ids"][same][:, :2])
    dist = np.sort(np.where(vis, got["f_distance"], np.inf), 0)
    with np.errstate(invalid="ignore"):
        clear = same & ((dist[1] - dist[0] > 1e-3) if len(dist) > 1 else True)
    clear &= (np.abs(np.where(vis, d[p + "f_distance"] - got["f_distance"], 0.0)) < 1e-4).all(0)
    assert clear.sum() > 0.5 * same.sum()
    assert np.array_equal(ours[clear][:, 2], d[p + "face_ids"][clear][:, 2])


@pytest.mark.parametrize("name", FIXTURES)
@pytest.mark.parametrize("facewise", [False, True], ids=["vertexwise", "facewise"])
@pytest.mark.parametrize("interp", [0, 1], ids=["nn", "bilinear"])
def test_port_vs_reference_tester(name, facewise, interp):
    d, mesh = load(name)
    got = port_run(d, mesh, facewise, interp)
    m = check_against_reference(d, got, facewise, interp, angle_ulp=0, what=(name, facewise, interp))
    print(name, "facewise" if facewise else "vertexwise", m)
    # FaceInfo::CalcStat on faces whose visible set agrees in every keyframe and whose records match bit for bit
    p = f"ref_fw{int(facewise)}_i{interp}_"
    ours = fold_faces(got["f_visible"], got["f_area"], got["f_viewing_angle"], got["f_distance"], d["ids"])
    compare_face_ids(d, p, got, ours, min_same=0.9)


def same_constants(got, want):
    """Bit-equal, except entries that are pure cancellation noise in double (true value 0, |x| < 1e-6): there the
    double rounding of a different association order may show (< 4e-15 absolute, ~1e-7 of a float ulp of any
    coordinate it multiplies).  Returns (ok, exact)."""
    got, want = np.asarray(got, np.float32).reshape(-1), np.asarray(want, np.float32).reshape(-1)
    neq = bits(got) != bits(want)
    ok = bool(np.all((np.abs(want[neq]) < 1e-6) & (np.abs(got[neq].astype(np.float64) - want[neq]) < 4e-15)))
    return ok, not neq.any()


def test_keyframe_camera_constants_match_the_reference():
    """ugu_b200_keyframe_camera_from_c2w restates Eigen's Affine3d inverse + cast: equal to the constants the
    compiled reference derived for every fixture camera (bit for bit above the cancellation noise)."""
    lib = N.load()
    for name in FIXTURES:
        d, _ = load(name)
        w, h = int(d["cam_size"][0]), int(d["cam_size"][1])
        for k, c2w in enumerate(d["cam_c2w"]):
            kf = N.Keyframe()
            rc = lib.ugu_b200_keyframe_camera_from_c2w((C.c_double * 16)(*c2w.reshape(-1)), w, h, float(d["cam_f"][0]),
                                                       float(d["cam_f"][1]), float(d["cam_pp"][0]),
                                                       float(d["cam_pp"][1]), C.byref(kf))
            assert rc == N.OK
            assert same_constants(kf.w2c[:], d["ref_w2c"][k])[0], (name, k)
            assert same_constants(kf.c2w_t[:], d["ref_c2w_t"][k])[1], (name, k)
            assert same_constants(kf.cam.c2w_R[:], d["ref_c2w_R"][k])[1], (name, k)


@pytest.mark.skipif(not ref_available(), reason="oracle/_ref is not built here")
def test_camera_constants_sweep_against_the_live_reference():
    """500 random and orbit poses: the host-side constants (rasterizer and tester flavours) against Eigen's
    (ADVICE r1).  c2w_R / c2w_t are plain casts and must be bit-equal; w2c_R (own polar decomposition vs Eigen's
    JacobiSVD rotation()), w2c_t and the affine w2c must be bit-equal wherever an entry is above the double
    cancellation noise.  The rate of poses with a noise-level difference is printed: the parity bar for hosts
    WITHOUT Eigen is therefore "bit-exact given identical camera constants"; C++ hosts take the constants from
    ugu::Camera itself (ugu_b200/cpp/rasterizer_cuda.cc) and are not affected."""
    from ugu_b200 import PinholeCamera, meshgen
    lib = N.load()
    poses = list(meshgen.random_sphere_poses(300, seed=11)) + list(meshgen.orbit_poses((0.1, -0.2, 0.3), 2.2, 200))
    rng = np.random.default_rng(5)
    inexact = 0
    for i, c2w in enumerate(poses):
        c2w = np.array(c2w, np.float64)
        if i % 3 == 0:  # off-centre, non-unit distances
            c2w[:3, 3] = c2w[:3, 3] * rng.uniform(0.3, 40.0) + rng.normal(0, 1, 3)
        cam = PinholeCamera(320, 240, c2w, fov_y_deg=50.0)
        w2c, R, t = oracle.ref_visibility_camera_constants(cam)
        kf = N.Keyframe()
        assert lib.ugu_b200_keyframe_camera_from_c2w((C.c_double * 16)(*c2w.reshape(-1)), 320, 240, 1.0, 1.0, 0.0, 0.0,
                                                     C.byref(kf)) == N.OK
        rw2c_R, rw2c_t, _ = oracle.ref_camera_constants(c2w)
        assert same_constants(kf.c2w_t[:], t)[1] and same_constants(kf.cam.c2w_R[:], R)[1], i
        exact = True
        for got, want in ((kf.w2c[:], w2c), (kf.cam.w2c_R[:], rw2c_R), (kf.cam.w2c_t[:], rw2c_t)):
            ok, ex = same_constants(got, want)
            assert ok, (i, np.array(got[:]), want)
            exact &= ex
        inexact += 0 if exact else 1
    print(f"camera constants: {inexact} of {len(poses)} poses differ from Eigen's in a cancellation-noise entry")


# ---- GPU ------------------------------------------------------------------------------------------------------
def cuda_run(d, mesh, facewise, interp, give_zbuf=True, records=True):
    from ugu_b200 import ColorInterpolation, DiffuseColor, Rasterizer, RendererCpuOption
    from ugu_b200.visibility import Keyframe, VisibilityInfo, VisibilityTester, VisibilityTesterOption
    r = Rasterizer(RendererCpuOption(diffuse_color=DiffuseColor.kNone))
    r.set_mesh(mesh)
    assert r.PrepareMesh()
    t = VisibilityTester(VisibilityTesterOption(interp=ColorInterpolation(interp)), rasterizer=r)
    t.set_data(mesh.vertices, mesh.normals, mesh.vertex_indices)
    assert t.PrepareData()
    cams = cameras(d)
    kfs = [Keyframe(id=int(d["ids"][k]), color=d["color"][k], depth=d["depth"][k], mask=d["mask"][k], camera=cams[k],
                    zbuf=(d["zbuf"][k] if give_zbuf else None)) for k in range(len(cams))]
    info = VisibilityInfo()
    assert t.Test(kfs, info, facewise, records=records)
    return info, t


def as_port_layout(info):
    return dict(v_visible=info.vertex_visible, v_rec=info.vertex, f_visible=info.face_visible,
                f_area=info.face["area"], f_viewing_angle=info.face["viewing_angle"], f_distance=info.face["distance"])


@pytest.mark.gpu
@pytest.mark.parametrize("name", FIXTURES)
@pytest.mark.parametrize("facewise", [False, True], ids=["vertexwise", "facewise"])
@pytest.mark.parametrize("interp", [0, 1], ids=["nn", "bilinear"])
def test_cuda_vs_port_and_reference_tester(name, facewise, interp):
    d, mesh = load(name)
    info, tester = cuda_run(d, mesh, facewise, interp)
    got = as_port_layout(info)
    want = port_run(d, mesh, facewise, interp)
    # CUDA vs the C restatement: identical decisions; every field bit-exact except what depends on acosf
    assert np.array_equal(got["v_visible"], want["v_visible"])
    assert np.array_equal(got["f_visible"], want["f_visible"])
    vis = want["v_visible"].astype(bool)
    for k_c, k_p in (("color", "color"), ("intensity", "intensity"), ("projected_pos", "projected"),
                     ("distance", "distance")):
        assert np.array_equal(bits(got["v_rec"][k_c][vis]), bits(want["v_rec"][k_p][vis])), k_c
    assert ulp_diff(got["v_rec"]["viewing_angle"][vis], want["v_rec"]["viewing_angle"][vis]).max() <= ULP_ANGLE
    fv = want["f_visible"].astype(bool)
    assert np.array_equal(bits(got["f_area"][fv]), bits(want["f_area"][fv]))
    assert np.array_equal(bits(got["f_distance"][fv]), bits(want["f_distance"][fv]))
    assert ulp_diff(got["f_viewing_angle"][fv], want["f_viewing_angle"][fv]).max() <= 2 * ULP_ANGLE
    assert not got["v_rec"]["distance"][~vis].any() and not got["f_area"][~fv].any()  # invisible rows are zero
    # CUDA vs the compiled reference tester
    m = check_against_reference(d, got, facewise, interp, angle_ulp=ULP_ANGLE, what=(name, facewise, interp))
    print(name, "facewise" if facewise else "vertexwise", m)
    # folds: FaceInfo::CalcStat ids from the device accumulators == a host fold of the device records
    st = info.stats
    ours = fold_faces(got["f_visible"], got["f_area"], got["f_viewing_angle"], got["f_distance"], d["ids"])
    assert np.array_equal(st["face_max_area_id"], ours[:, 0])
    assert np.array_equal(st["face_min_viewing_angle_id"], ours[:, 1])
    assert np.array_equal(st["face_min_distance_id"], ours[:, 2])
    assert np.array_equal(st["face_count"], got["f_visible"].sum(0))
    # ... and == the reference's wherever its visible sets and records are the same
    p = f"ref_fw{int(facewise)}_i{interp}_"
    compare_face_ids(d, p, got, ours, min_same=0.5)
    # vertex folds (VertexInfo::CalcStat, min / mean subset) against a host fold of the device records
    K, V = got["v_visible"].shape
    visb = got["v_visible"].astype(bool)
    cnt = visb.sum(0)
    assert np.array_equal(st["vertex_count"], cnt)
    ids = np.asarray(d["ids"], np.int32)
    ang = np.where(visb, got["v_rec"]["viewing_angle"], np.inf)
    dist = np.where(visb, got["v_rec"]["distance"], np.inf)
    anyv = cnt > 0
    assert np.array_equal(st["vertex_min_viewing_angle_id"][anyv], ids[np.argmin(ang, 0)][anyv])
    assert np.array_equal(st["vertex_min_distance_id"][anyv], ids[np.argmin(dist, 0)][anyv])
    assert (st["vertex_min_viewing_angle_id"][~anyv] == -1).all()
    mean_c = np.zeros((V, 3), np.float32)
    mean_a = np.zeros(V, np.float32)
    for k in range(K):  # float sums in keyframe order, like the reference's +=
        mean_c[visb[k]] += got["v_rec"]["color"][k][visb[k]]
        mean_a[visb[k]] += got["v_rec"]["viewing_angle"][k][visb[k]]
    mean_c[anyv] /= cnt[anyv, None].astype(np.float32)
    mean_a[anyv] /= cnt[anyv].astype(np.float32)
    assert np.array_equal(bits(st["vertex_mean_color"]), bits(mean_c))
    assert np.array_equal(bits(st["vertex_mean_viewing_angle"]), bits(mean_a))
    best = np.argmin(ang, 0)
    assert np.array_equal(st["vertex_min_viewing_angle_color"][anyv],
                          got["v_rec"]["color"][best, np.arange(V)][anyv])
    # reference VertexInfo::CalcStat where the vertex's visible set and records are identical
    same_v = (d[p + "v_visible"] == got["v_visible"]).all(0) & anyv
    same_v &= ((bits(d[p + "v_viewing_angle"]) == bits(got["v_rec"]["viewing_angle"])) | ~visb).all(0)
    same_v &= ((d[p + "v_color"] == got["v_rec"]["color"]).all(-1) | ~visb).all(0)
    assert same_v.sum() > 0.3 * anyv.sum()
    assert np.array_equal(st["vertex_min_viewing_angle_id"][same_v], d[p + "vertex_ids"][same_v, 0])
    assert np.array_equal(bits(st["vertex_mean_color"][same_v]), bits(d[p + "vertex_colors"][same_v, 0]))
    assert np.array_equal(bits(st["vertex_mean_viewing_angle"][same_v]), bits(d[p + "vertex_scalars"][same_v, 0]))


@pytest.mark.gpu
def test_library_rendered_zbuf_equals_the_given_one_and_chunks_accumulate():
    """zbuf = None makes the library render the self depth map from the uploaded rasterizer mesh (depth only,
    back-face culling off) — bit-identical to the reference Rasterizer's, so nothing changes; and two calls of
    three + two keyframes fold to the same statistics as one call of five."""
    d, mesh = load("spot_orbit5")
    a, _ = cuda_run(d, mesh, True, 1, give_zbuf=True)
    b, _ = cuda_run(d, mesh, True, 1, give_zbuf=False)
    for k in ("vertex_visible", "vertex", "face_visible", "face"):
        assert np.array_equal(getattr(a, k), getattr(b, k)), k
    for k, v in a.stats.items():
        assert np.array_equal(v, b.stats[k]), k
    from ugu_b200 import DiffuseColor, Rasterizer, RendererCpuOption
    from ugu_b200.visibility import Keyframe, VisibilityInfo, VisibilityTester
    r = Rasterizer(RendererCpuOption(diffuse_color=DiffuseColor.kNone))
    r.set_mesh(mesh)
    assert r.PrepareMesh()
    t = VisibilityTester(rasterizer=r)
    t.set_data(mesh.vertices, mesh.normals, mesh.vertex_indices)
    assert t.PrepareData()
    cams = cameras(d)
    kfs = [Keyframe(id=int(d["ids"][k]), color=d["color"][k], depth=d["depth"][k], mask=d["mask"][k], camera=cams[k])
           for k in range(5)]
    info = VisibilityInfo()
    assert t.Test(kfs[:3], info, True, records=False) and t.Test(kfs[3:], info, True, records=False)
    for k, v in a.stats.items():
        assert np.array_equal(v, info.stats[k]), k
    # error behaviour mirrors ValidateAndInitBeforeTest (visibility_tester.cc:274-300)
    bad = VisibilityInfo()
    assert not t.Test([Keyframe(id=1, color=None, camera=cams[0])], bad, True)
    assert not t.Test([Keyframe(id=1, color=d["color"][0], depth=None, mask=d["mask"][0], camera=cams[0])], bad, True)
    t2 = VisibilityTester(rasterizer=r)
    assert not t2.Test(kfs[:1], VisibilityInfo(), True)  # PrepareData not called


@pytest.mark.gpu
def test_config3_turntable_on_the_device():
    """BASELINE config 3: spot, 36 keyframes at 1024x1024 rendered by the hot path (colour / depth / mask stay in
    HBM), then the visibility pass on the device images with library-rendered self depth maps; face counts are
    checked for plausibility against the face_id-map consumer (every face the tester calls visible from a
    keyframe owns pixels or is sub-pixel there) and the per-keyframe time is printed."""
    import time
    import torch
    from ugu_b200 import DiffuseColor, PinholeCamera, Rasterizer, RendererCpuOption, meshgen
    from ugu_b200.visibility import Keyframe, VisibilityTester, native_keyframe
    mesh = cases.load_spot()
    r = Rasterizer(RendererCpuOption(diffuse_color=DiffuseColor.kTexture))
    r.set_mesh(mesh)
    assert r.PrepareMesh()
    c, ext = cases.bbox_center_and_extent(mesh)
    K, W = 36, 1024
    cams = [PinholeCamera(W, W, p, fov_y_deg=45.0) for p in meshgen.orbit_poses(c, 1.5 * ext, K)]
    P = W * W
    col = torch.empty(K * P * 3, dtype=torch.uint8, device="cuda")
    dep = torch.empty(K * P, dtype=torch.float32, device="cuda")
    msk = torch.empty(K * P, dtype=torch.uint8, device="cuda")
    ts = torch.cuda.Stream()
    torch.cuda.synchronize()
    r.render_batch(cams, color=col.data_ptr(), depth=dep.data_ptr(), mask=msk.data_ptr(), on_device=True,
                   stream=ts.cuda_stream)
    t = VisibilityTester(rasterizer=r)
    t.set_data(mesh.vertices, mesh.normals, mesh.vertex_indices)
    assert t.PrepareData()
    keep = []
    kfs = (N.Keyframe * K)(*[native_keyframe(
        Keyframe(id=k, camera=cams[k], color=col.data_ptr() + k * P * 3, depth=dep.data_ptr() + k * P * 4,
                 mask=msk.data_ptr() + k * P), keep, images_on_device=True) for k in range(K)])
    opt = t.native_option(True)
    lib = N.load()
    lib.ugu_b200_visibility_reset(r._h)
    for rep in range(2):
        if rep:
            lib.ugu_b200_visibility_reset(r._h)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        rc = lib.ugu_b200_visibility_test_device(r._h, K, kfs, 0, C.byref(opt), None, C.c_void_p(ts.cuda_stream))
        assert rc == N.OK, N.last_error(r._h)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
    st = t.stats(list(range(K)))
    seen = st["face_count"] > 0
    print(f"config 3 visibility pass: {dt / K * 1e6:.1f} us per keyframe (incl. the self depth render), "
          f"{int(seen.sum())} of {len(seen)} faces seen by some keyframe")
    assert seen.mean() > 0.8   # a horizontal turntable sees all but the faces turned up or down
    assert (st["face_max_area_id"][seen] >= 0).all() and (st["face_max_area_id"][~seen] == -1).all()
    assert st["face_count"].max() <= K


@pytest.mark.gpu
@pytest.mark.parametrize("name", FIXTURES)
@pytest.mark.parametrize("facewise", [False, True], ids=["vertexwise", "facewise"])
def test_face_fold_inside_the_face_kernels_equals_the_fold_kernel(monkeypatch, name, facewise):
    """Without dense face records the face kernels fold into the accumulators themselves (packed-key atomics);
    the statistics must equal, bit for bit, those of the record rows + vt_fold_face_kernel path
    (UGU_B200_VT_FOLD=0) and those of a call that also returns the records."""
    d, mesh = load(name)
    fused, _ = cuda_run(d, mesh, facewise, 1, records=False)
    monkeypatch.setenv("UGU_B200_VT_FOLD", "0")
    rows, _ = cuda_run(d, mesh, facewise, 1, records=False)
    monkeypatch.delenv("UGU_B200_VT_FOLD")
    full, _ = cuda_run(d, mesh, facewise, 1, records=True)
    assert int(fused.stats["face_count"].sum()) > 0
    for k in fused.stats:
        for other, what in ((rows, "fold kernel"), (full, "with records")):
            a, b = np.ascontiguousarray(fused.stats[k]), np.ascontiguousarray(other.stats[k])
            assert np.array_equal(a.view(np.uint8), b.view(np.uint8)), (k, what)
