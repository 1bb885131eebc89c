"""CPU suite: the plain-C oracle (oracle/raster_oracle.c) against the golden fixtures produced by
the compiled reference, and — where oracle/_ref exists — against the live reference."""
import numpy as np
import pytest

import cases
import util
from oracle import OraclePort, OracleRef, ref_available
from ugu_b200 import DiffuseShading, PinholeCamera
from ugu_b200 import _native as N
from ugu_b200 import meshgen


class _Pod:
    """Camera constants as the REFERENCE derived them (from the golden file)."""

    def __init__(self, cam, consts):
        self.width, self.height = cam.width(), cam.height()
        self.fx, self.fy = float(cam.focal_length()[0]), float(cam.focal_length()[1])
        self.cx, self.cy = float(cam.principal_point()[0]), float(cam.principal_point()[1])
        self.w2c_R = (N.C.c_float * 9)(*consts[0])
        self.w2c_t = (N.C.c_float * 3)(*consts[1])
        self.c2w_R = (N.C.c_float * 9)(*consts[2])


@pytest.mark.parametrize("name", util.golden_names())
def test_port_reproduces_reference_golden(name):
    mesh, cam, opt, want, consts = util.load_golden(name)
    got = OraclePort(mesh).render(_Pod(cam, consts), opt)
    assert got is not None
    # Oren-Nayar goes through libm acosf/sinf/tanf in both: same libm here, so still bit-exact.
    util.assert_parity(got, want, strict=True, color_lsb=0, what=name)
    for k in cases.OUTPUT_KEYS:
        assert np.array_equal(util.bits(got[k]), util.bits(want[k])), f"{name}: {k} bytes differ"


@pytest.mark.parametrize("name", util.golden_names())
def test_host_camera_constants_match_reference(name):
    """ugu_b200_camera_from_c2w (pure host code) against Eigen's values stored in the fixture.
    Compared as float VALUES: the sign of a zero is not pinned (see camera_host.cc)."""
    _, cam, _, _, consts = util.load_golden(name)
    pod = cam.to_native()
    assert np.array_equal(np.array(pod.w2c_R, np.float32), consts[0])
    assert np.array_equal(np.array(pod.w2c_t, np.float32), consts[1])
    assert np.array_equal(np.array(pod.c2w_R, np.float32), consts[2])


def test_port_with_host_constants_equals_golden_faceid():
    """End to end on the CPU side: host camera math + port == reference, for every golden case."""
    for name in util.golden_names():
        mesh, cam, opt, want, _ = util.load_golden(name)
        got = OraclePort(mesh).render(cam.to_native(), opt)
        util.assert_parity(got, want, strict=True, what=name)


def test_validation_failures_mirror_reference():
    mesh = cases.kat_meshes()["two_overlapping"]
    cam = cases.kat_camera().to_native()
    port = OraclePort(mesh)
    from ugu_b200 import DiffuseColor, RendererCpuOption
    assert port.render(cam, RendererCpuOption(diffuse_color=DiffuseColor.kTexture)) is None  # no materials
    assert port.render(cam, RendererCpuOption(diffuse_color=DiffuseColor.kVertex), want=()) is None  # all null


@pytest.mark.skipif(not ref_available(), reason="oracle/_ref not built (needs /root/reference)")
@pytest.mark.parametrize("seed", [1, 2, 3])
def test_port_vs_live_reference_random(seed):
    mesh = cases.random_soup(600, seed=seed, big_fraction=0.2)
    rng = np.random.RandomState(seed)
    ref = OracleRef(mesh)
    port = OraclePort(mesh)
    for tag, opt in cases.ALL_OPTIONS:
        eye = rng.uniform(-0.3, 0.3, 3)
        c2w = meshgen.look_at_c2w(eye, eye + np.array([rng.uniform(-0.2, 0.2), rng.uniform(-0.2, 0.2), 1.0]),
                                  (0.1, -1.0, 0.05))
        cam = PinholeCamera(80 + 8 * seed, 60, c2w, fov_y_deg=50.0 + seed)
        want = ref.render(cam, opt)
        R, t, Cm = ref.camera_constants(cam.c2w())
        got = port.render(_Pod(cam, (R, t, Cm)), opt)
        util.assert_parity(got, want, strict=True, what=f"seed{seed}/{tag}")


@pytest.mark.skipif(not ref_available(), reason="oracle/_ref not built (needs /root/reference)")
def test_host_camera_constants_vs_live_reference_many():
    ref = OracleRef()
    rng = np.random.RandomState(0)
    bad = 0
    for i in range(2000):
        eye = rng.normal(size=3) * 3
        c2w = meshgen.look_at_c2w(eye, rng.normal(size=3) * 0.3, rng.normal(size=3))
        if i % 3 == 0:  # also poses that are float-rounded rotations with double noise
            c2w[:3, :3] += rng.normal(size=(3, 3)) * 1e-9
        R, t, Cm = ref.camera_constants(c2w)
        pod = PinholeCamera(8, 8, c2w, fov_y_deg=45).to_native()
        ok = (np.array_equal(np.array(pod.w2c_R, np.float32), R) and
              np.array_equal(np.array(pod.w2c_t, np.float32), t) and
              np.array_equal(np.array(pod.c2w_R, np.float32), Cm))
        bad += 0 if ok else 1
    assert bad == 0, f"{bad} / 2000 poses differ from Eigen after the float cast"
