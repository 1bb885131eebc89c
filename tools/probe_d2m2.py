"""Probe: float op order of depth2color * p and Project in the two-camera Depth2Mesh (compiled reference)."""
import sys
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import numpy as np
import oracle
from ugu_b200 import PinholeCamera, meshgen
f32 = np.float32
W, H, CW, CH = 64, 48, 96, 80
rng = np.random.RandomState(0)
depth = (1.0 + rng.rand(H, W)).astype(f32)
color = rng.randint(0, 255, (CH, CW, 3)).astype(np.uint8)
dcam = PinholeCamera(W, H, np.eye(4), fov_y_deg=50.0)
ccam = PinholeCamera(CW, CH, np.eye(4), fov_y_deg=65.0)
a = 0.05
R = np.array([[np.cos(a), 0, np.sin(a)], [0.01, 1, 0.02], [-np.sin(a), 0.03, np.cos(a)]], f32)
t = np.array([0.02, -0.013, 0.007], f32)
M = np.eye(4, dtype=f32); M[:3, :3] = R; M[:3, 3] = t
for gl in (False,):
    ref = oracle.ref_depth2mesh2(dcam, ccam, M, depth, color, 10.0, 1, 1, gl, True)
    v = ref["vertices"]
    x, y, z = v[:, 0], v[:, 1], v[:, 2]
    def order(kind):
        out = []
        for i in range(3):
            a0, a1, a2 = R[i, 0] * x, R[i, 1] * y, R[i, 2] * z
            if kind == "t+(a0+(a1+a2))": out.append(t[i] + (a0 + (a1 + a2)))
            if kind == "((t+a0)+a1)+a2": out.append(((t[i] + a0) + a1) + a2)
            if kind == "((a0+a1)+a2)+t": out.append(((a0 + a1) + a2) + t[i])
            if kind == "(a0+(a1+a2))+t": out.append((a0 + (a1 + a2)) + t[i])
            if kind == "t+((a0+a1)+a2)": out.append(t[i] + ((a0 + a1) + a2))
        return out
    fx, fy = [f32(q) for q in ccam.focal_length()]
    cx, cy = [f32(q) for q in ccam.principal_point()]
    for kind in ("t+(a0+(a1+a2))", "((t+a0)+a1)+a2", "((a0+a1)+a2)+t", "(a0+(a1+a2))+t", "t+((a0+a1)+a2)"):
        px, py, pz = order(kind)
        ix = fx / pz * px + cx
        iy = fy / pz * py + cy
        u = (ix + f32(0.5)) / f32(CW)
        vv = f32(1.0) - (iy + f32(0.5)) / f32(CH)
        bad = int((u.view(np.uint32) != ref["uv"][:, 0].view(np.uint32)).sum() + (vv.view(np.uint32) != ref["uv"][:, 1].view(np.uint32)).sum())
        print(kind, "mismatching uv words:", bad, "of", 2 * len(u))

# the C port against the compiled reference, bit for bit
for gl in (False, True):
    for xs, ys in ((1, 1), (2, 3)):
        ref = oracle.ref_depth2mesh2(dcam, ccam, M, depth, color, 0.4, xs, ys, gl, True)
        port = oracle.port_depth2mesh(dcam.to_native(), depth, 0.4, xs, ys, gl, color, True,
                                      two_cam=oracle.two_camera_constants(M, ccam))
        for k in ref:
            a, b = np.asarray(ref[k]), np.asarray(port[k])
            assert a.shape == b.shape and np.array_equal(a.view(np.uint8), b.view(np.uint8)), (k, gl, xs, ys)
        print("port == reference", gl, xs, ys, len(ref["vertices"]), len(ref["faces"]))
