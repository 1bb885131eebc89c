"""G-buffer outputs: ugu::GBuffer (include/ugu/renderer/base.h:34-77) as
RendererGl::ReadGbuf fills it (src/renderer/gl/renderer.cc:501-650) from the gbuf shaders, evaluated on
the Rasterizer's winners.  PARITY UNPINNED against the reference's OpenGL output (no graphics stack on
this box; GL samples differently anyway): the checker is oracle_gbuffer, a restatement of the shader /
read-back arithmetic on the reference-pinned winners, weights and depths of the port.

CPU: the restatement's consistency with the (pinned) Rasterizer outputs and its own invariants.
GPU: the CUDA path through the C-ABI against the restatement, bit for bit on every field."""
import numpy as np
import pytest

import cases
from oracle import GBUFFER_FIELDS, OraclePort, port_gbuffer
from ugu_b200 import (ColorInterpolation, DiffuseColor, DiffuseShading, Mesh, PinholeCamera,
                      RendererCpuOption, meshgen)

FIELDS = [n for n, _, _ in GBUFFER_FIELDS]


def _spot_case(w=200, h=150, k=1):
    spot = cases.load_spot()
    c, ext = cases.bbox_center_and_extent(spot)
    cam = PinholeCamera(w, h, meshgen.orbit_poses(c, 1.5 * ext, 4)[k], fov_y_deg=45.0)
    return spot, cam, RendererCpuOption(diffuse_color=DiffuseColor.kTexture)


def _ico_case(w=160, h=120):
    mesh = meshgen.icosphere(4)
    cam = PinholeCamera(w, h, meshgen.look_at_c2w((0, 0, -2.7), (0, 0, 0), (0, -1, 0)), fov_y_deg=45.0)
    return mesh, cam, RendererCpuOption(diffuse_color=DiffuseColor.kVertex)


def _soup_case(w=97, h=61, cull=False):
    mesh = cases.random_soup(400, 11)
    return mesh, cases.kat_camera(w, h), RendererCpuOption(diffuse_color=DiffuseColor.kTexture,
                                                           interp=ColorInterpolation.kNn,
                                                           backface_culling=cull)


CASES = {
    "spot": (_spot_case, dict(z_near=0.01, z_far=100.0)),
    "spot_flat": (_spot_case, dict(z_near=0.01, z_far=100.0, flat_normal=True, geo_id=3)),
    "ico_clip": (_ico_case, dict(z_near=1.9, z_far=2.4, geo_id=1)),  # near AND far cut into the sphere
    "soup_nocull": (_soup_case, dict(z_near=0.3, z_far=6.0)),
    "soup_cull": (lambda: _soup_case(cull=True), dict(z_near=0.3, z_far=6.0, flat_normal=True)),
}


def _oracle(mesh, cam, opt, kw, fields=None):
    return port_gbuffer(OraclePort(mesh), cam.to_native(), opt, fields=fields, **kw)


# ---------------------------------------------------------------------------------------------- CPU
@pytest.mark.parametrize("name", sorted(CASES))
def test_restatement_is_consistent_with_the_rasterizer(name):
    make, kw = CASES[name]
    mesh, cam, opt = make()
    g = _oracle(mesh, cam, opt, kw)
    assert g is not None
    r = OraclePort(mesh).render(cam.to_native(), opt)
    vis = g["stencil"] == 255
    assert vis.any() and set(np.unique(g["stencil"])) <= {0, 255}
    # winners: face_id + 1 (gbuf.frag:36) wherever the Rasterizer kept one and its depth is inside (near, far)
    inside = (r["face_id"] >= 0) & (r["depth"] > kw["z_near"]) & (r["depth"] < kw["z_far"])
    border = np.isclose(r["depth"], kw["z_near"], rtol=1e-4) | np.isclose(r["depth"], kw["z_far"], rtol=1e-4)
    assert np.array_equal(vis[~border], inside[~border])
    assert np.array_equal(g["face_id"][vis] - 1, r["face_id"][vis]) and not g["face_id"][~vis].any()
    assert (g["geo_id"][vis] == kw.get("geo_id", 0)).all() and not g["geo_id"][~vis].any()
    # pos_cam.z is the Rasterizer depth up to the rounding of two evaluation routes; pos_cam = R pos_wld + t
    n = cam.to_native()
    R, t = np.array(n.w2c_R, np.float64).reshape(3, 3), np.array(n.w2c_t, np.float64)
    assert np.allclose(g["pos_cam"][vis][:, 2], r["depth"][vis], rtol=2e-5)
    assert np.allclose(g["pos_cam"][vis], g["pos_wld"][vis].astype(np.float64) @ R.T + t, atol=1e-5)
    assert np.allclose(g["normal_cam"][vis], g["normal_wld"][vis].astype(np.float64) @ R.T, atol=1e-6)
    nn = np.linalg.norm(g["normal_wld"][vis].astype(np.float64), axis=-1)
    assert np.allclose(nn[nn > 0], 1.0, atol=1e-6) and (nn > 0).mean() > 0.99
    assert np.allclose(g["bary"][vis].sum(-1), 1.0, atol=1e-6)
    if mesh.uv is not None and len(mesh.uv):
        assert np.allclose(g["uv"][vis].sum(-1), 1.0, atol=1e-5)
    else:
        assert not g["uv"].any()
    # window depth: inside (0, 1) on winners, 1 on background, monotone in z
    d = g["depth_01"]
    assert ((d[vis] > 0) & (d[vis] < 1)).all() and (d[r["face_id"] < 0] == 1.0).all()
    zs, ds = r["depth"][vis], d[vis]
    o = np.argsort(zs)
    assert (np.diff(ds[o]) >= -1e-6).all()
    for k in ("normal_cam", "normal_wld", "pos_cam", "pos_wld", "uv", "bary"):
        assert not g[k][~vis].any()
    # colour: the Rasterizer's, with the clipped winners erased
    assert np.array_equal(g["color"][vis], r["color"][vis]) and not g["color"][~vis].any()


def test_projected_world_position_lands_on_its_pixel():
    mesh, cam, opt = _spot_case()
    g = _oracle(mesh, cam, opt, dict(z_near=0.01, z_far=100.0))
    vis = g["stencil"] == 255
    n = cam.to_native()
    pc = g["pos_cam"][vis].astype(np.float64)
    ys, xs = np.nonzero(vis)
    assert np.allclose(n.fx * pc[:, 0] / pc[:, 2] + n.cx, xs, atol=2e-3)
    assert np.allclose(n.fy * pc[:, 1] / pc[:, 2] + n.cy, ys, atol=2e-3)
    # uv interpolated like the texture colourizer: texel lookup at that uv reproduces the NN colour
    tex = mesh.materials[0].diffuse_tex
    g2 = _oracle(mesh, cam, RendererCpuOption(diffuse_color=DiffuseColor.kTexture, interp=ColorInterpolation.kNn),
                 dict(z_near=0.01, z_far=100.0))
    shaded = vis & (g2["face_id"] > 1)  # face 0 is never coloured by the Rasterizer
    uv = g2["uv"][shaded]
    tx = np.round(uv[:, 0] * np.float32(tex.shape[1] - 1)).astype(int)
    ty = np.round((np.float32(1.0) - uv[:, 1]) * np.float32(tex.shape[0] - 1)).astype(int)
    assert np.array_equal(tex[ty, tx], g2["color"][shaded])


def test_restatement_rejects_bad_near_far_and_handles_missing_attributes():
    mesh, cam, opt = _ico_case()
    assert _oracle(mesh, cam, opt, dict(z_near=0.0, z_far=10.0)) is None
    assert _oracle(mesh, cam, opt, dict(z_near=2.0, z_far=1.0)) is None
    g = _oracle(mesh, cam, opt, dict(z_near=0.1, z_far=10.0))  # the icosphere has no uv
    assert not g["uv"].any() and (g["stencil"] == 255).any()
    bare = Mesh(vertices=mesh.vertices, vertex_indices=mesh.vertex_indices, face_normals=mesh.face_normals)
    g = _oracle(bare, cam, RendererCpuOption(diffuse_color=DiffuseColor.kNone), dict(z_near=0.1, z_far=10.0),
                fields=[f for f in FIELDS if f != "color"])
    assert not g["normal_wld"].any() and g["pos_wld"].any()   # no vertex normals -> zero normals


# ---------------------------------------------------------------------------------------------- GPU
def _same(a, b):
    return np.array_equal(np.ascontiguousarray(a).view(np.uint8), np.ascontiguousarray(b).view(np.uint8))


def _cuda(mesh, cam, opt, kw, fields=None, pinned=True):
    from ugu_b200 import GBuffer, Rasterizer
    r = Rasterizer(opt)
    r.set_mesh(mesh)
    assert r.PrepareMesh()
    r.set_camera(cam)
    gb = GBuffer(pinned)
    assert r.RenderGBuffer(gb, fields=fields, **kw)
    return r, gb


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(CASES))
def test_cuda_gbuffer_equals_restatement_bit_for_bit(name):
    make, kw = CASES[name]
    mesh, cam, opt = make()
    want = _oracle(mesh, cam, opt, kw)
    r, gb = _cuda(mesh, cam, opt, kw)
    for f in FIELDS:
        assert _same(getattr(gb, f).data, want[f]), f
    # the visibility buffer is left clean: an ordinary Render afterwards is unaffected
    from ugu_b200 import Image1i
    fid = Image1i()
    assert r.RenderFaceId(fid)
    assert np.array_equal(fid.data, OraclePort(mesh).render(cam.to_native(), opt, want=("face_id",))["face_id"])
    # and a second G-buffer call gives the same answer (idempotence)
    assert r.RenderGBuffer(gb, **kw)
    for f in FIELDS:
        assert _same(getattr(gb, f).data, want[f]), f


@pytest.mark.gpu
def test_cuda_gbuffer_subset_pageable_and_ragged_size():
    mesh, cam, opt = _soup_case(w=33, h=7)  # 231 pixels: ragged last chunk
    kw = dict(z_near=0.3, z_far=6.0)
    want = _oracle(mesh, cam, opt, kw)
    sub = ["pos_wld", "stencil", "face_id", "bary"]
    r, gb = _cuda(mesh, cam, opt, kw, fields=sub, pinned=False)
    for f in sub:
        assert _same(getattr(gb, f).data, want[f]), f
    assert gb.color.empty() and gb.uv.empty()
    r2, gb2 = _cuda(mesh, cam, opt, kw, fields=["color"])
    assert _same(gb2.color.data, want["color"])


@pytest.mark.gpu
def test_cuda_gbuffer_full_size_config2_properties_and_errors():
    from ugu_b200 import GBuffer, Image1f, Image1i, Rasterizer
    mesh = meshgen.icosphere(8)
    opt = RendererCpuOption(diffuse_color=DiffuseColor.kVertex)
    cam = PinholeCamera(1920, 1080, meshgen.look_at_c2w((0, 0, -2.7), (0, 0, 0), (0, -1, 0)), fov_y_deg=45.0)
    r, gb = _cuda(mesh, cam, opt, dict(z_near=0.5, z_far=10.0))
    fid, dep = Image1i(), Image1f()
    assert r.Render(None, dep, None, None, fid)
    vis = gb.stencil.data == 255
    assert np.array_equal(vis, fid.data >= 0) and vis.sum() > 800000
    assert np.array_equal(gb.face_id.data[vis] - 1, fid.data[vis])
    assert np.allclose(gb.pos_cam.data[vis][:, 2], dep.data[vis], rtol=2e-5)
    assert np.allclose(np.linalg.norm(gb.pos_wld.data[vis].astype(np.float64), axis=-1), 1.0, atol=2e-4)
    assert np.allclose(np.linalg.norm(gb.normal_wld.data[vis].astype(np.float64), axis=-1), 1.0, atol=1e-6)
    # failures: false + log, like Render's preconditions
    assert not r.RenderGBuffer(gb, z_near=0.0, z_far=1.0)
    assert not r.RenderGBuffer(gb, z_near=2.0, z_far=1.0)
    r3 = Rasterizer(RendererCpuOption(diffuse_color=DiffuseColor.kTexture))
    r3.set_mesh(mesh)
    assert r3.PrepareMesh()
    r3.set_camera(cam)
    assert not r3.RenderGBuffer(GBuffer())                      # texture colour without materials
    assert r3.RenderGBuffer(GBuffer(), fields=["face_id"])      # ... matters only when colour is requested


@pytest.mark.gpu
def test_cuda_gbuffer_device_form_chains_on_a_stream():
    import torch
    mesh, cam, opt = _spot_case(w=256, h=192)
    kw = dict(z_near=0.01, z_far=100.0, geo_id=3)
    want = _oracle(mesh, cam, opt, kw)
    from ugu_b200 import Rasterizer
    r = Rasterizer(opt)
    r.set_mesh(mesh)
    assert r.PrepareMesh()
    P = cam.width() * cam.height()
    # pre-filled, so that a field the call skipped cannot pass by landing on fresh zeroed memory
    bufs = {n: torch.full((P * ch * np.dtype(dt).itemsize,), 0xA5, dtype=torch.uint8, device="cuda")
            for n, dt, ch in GBUFFER_FIELDS}
    st = torch.cuda.Stream()
    st.wait_stream(torch.cuda.current_stream())
    r.render_gbuffer_device(cam, 0.01, 100.0, geo_id_value=3, stream=st.cuda_stream,
                            **{n: b.data_ptr() for n, b in bufs.items()})
    with pytest.raises(TypeError):
        r.render_gbuffer_device(cam, 0.01, 100.0, shaded=bufs["color"].data_ptr())
    st.synchronize()
    for n, dt, ch in GBUFFER_FIELDS:
        got = bufs[n].cpu().numpy().view(dt).reshape(want[n].shape)
        assert _same(got, want[n]), n
