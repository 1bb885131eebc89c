"""TEST INFRASTRUCTURE — python loaders for the two checkers.

  OraclePort     oracle/liboracle.so        plain-C restatement (oracle/raster_oracle.c)
  OracleRef      oracle/_ref/libugu_ref.so  the unmodified reference Rasterizer + oracle/ref_shim.cc

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this
package.  Nothing under ugu_b200/ does.
"""
from __future__ import annotations

import ctypes as C
import subprocess
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
PORT_LIB = HERE / "liboracle.so"
REF_LIB = HERE / "_ref" / "libugu_ref.so"


def build_port(force: bool = False) -> Path:
    src = HERE / "raster_oracle.c"
    if force or not PORT_LIB.exists() or PORT_LIB.stat().st_mtime < src.stat().st_mtime:
        subprocess.run(["make", "-C", str(HERE), "oracle"], check=True, capture_output=True)
    return PORT_LIB


def build_ref() -> Path:
    """Compile the reference from /root/reference (only possible where that tree exists)."""
    subprocess.run(["make", "-C", str(HERE), "-j8", "ref"], check=True, capture_output=True)
    return REF_LIB


def ref_available() -> bool:
    return REF_LIB.exists()


class _MeshDesc(C.Structure):
    _fields_ = [("num_vertices", C.c_int32), ("num_normals", C.c_int32), ("num_colors", C.c_int32),
                ("num_faces", C.c_int32), ("num_face_normals", C.c_int32), ("num_uv", C.c_int32),
                ("num_materials", C.c_int32), ("pad_", C.c_int32),
                ("vertices", C.c_void_p), ("normals", C.c_void_p), ("vertex_colors", C.c_void_p),
                ("vertex_indices", C.c_void_p), ("normal_indices", C.c_void_p),
                ("face_normals", C.c_void_p), ("uv", C.c_void_p), ("uv_indices", C.c_void_p),
                ("material_ids", C.c_void_p), ("tex_data", C.POINTER(C.c_void_p)),
                ("tex_w", C.POINTER(C.c_int32)), ("tex_h", C.POINTER(C.c_int32))]


class _PortCamera(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("fx", C.c_float), ("fy", C.c_float),
                ("cx", C.c_float), ("cy", C.c_float), ("w2c_R", C.c_float * 9),
                ("w2c_t", C.c_float * 3), ("c2w_R", C.c_float * 9)]


class _PortOption(C.Structure):
    _fields_ = [("diffuse_color", C.c_int32), ("interp", C.c_int32), ("shading_normal", C.c_int32),
                ("diffuse_shading", C.c_int32), ("backface_culling", C.c_int32),
                ("depth_scale", C.c_float), ("oren_nayar_sigma", C.c_float)]


class _RefCamera(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("fx", C.c_float), ("fy", C.c_float),
                ("cx", C.c_float), ("cy", C.c_float), ("c2w", C.c_double * 16)]


class _RefOption(C.Structure):
    _fields_ = [("diffuse_color", C.c_int32), ("interp", C.c_int32), ("shading_normal", C.c_int32),
                ("diffuse_shading", C.c_int32), ("backface_culling", C.c_int32),
                ("depth_scale", C.c_float)]


class _MeshHolder:
    """Keeps the numpy arrays alive and exposes the flat descriptor both checkers share."""

    def __init__(self, mesh):
        f32 = lambda a, w: np.ascontiguousarray(np.asarray(a, np.float32).reshape(-1, w))
        i32 = lambda a, w: np.ascontiguousarray(np.asarray(a, np.int32).reshape(-1, w))
        self.v, self.n, self.c = f32(mesh.vertices, 3), f32(mesh.normals, 3), f32(mesh.vertex_colors, 3)
        self.vi, self.ni, self.ui = (i32(mesh.vertex_indices, 3), i32(mesh.normal_indices, 3),
                                     i32(mesh.uv_indices, 3))
        self.fn, self.uv = f32(mesh.face_normals, 3), f32(mesh.uv, 2)
        self.mid = np.ascontiguousarray(np.asarray(mesh.material_ids, np.int32).reshape(-1))
        self.texs = [None if (t.diffuse_tex is None or t.diffuse_tex.size == 0)
                     else np.ascontiguousarray(t.diffuse_tex, np.uint8) for t in mesh.materials]
        F, M = len(self.vi), len(self.texs)
        p = lambda a: a.ctypes.data_as(C.c_void_p)
        d = _MeshDesc()
        d.num_vertices, d.num_normals, d.num_colors = len(self.v), len(self.n), len(self.c)
        d.num_faces, d.num_face_normals, d.num_uv, d.num_materials = F, len(self.fn), len(self.uv), M
        d.vertices, d.normals, d.vertex_colors = p(self.v), p(self.n), p(self.c)
        d.vertex_indices = p(self.vi)
        d.normal_indices = p(self.ni) if len(self.ni) == F and F > 0 else None
        d.face_normals, d.uv = p(self.fn), p(self.uv)
        d.uv_indices = p(self.ui) if len(self.ui) == F and F > 0 else None
        d.material_ids = p(self.mid) if len(self.mid) == F and F > 0 else None
        self._tp = (C.c_void_p * max(M, 1))(*[None if t is None else t.ctypes.data for t in self.texs])
        self._tw = (C.c_int32 * max(M, 1))(*[0 if t is None else t.shape[1] for t in self.texs])
        self._th = (C.c_int32 * max(M, 1))(*[0 if t is None else t.shape[0] for t in self.texs])
        d.tex_data = C.cast(self._tp, C.POINTER(C.c_void_p))
        d.tex_w = C.cast(self._tw, C.POINTER(C.c_int32))
        d.tex_h = C.cast(self._th, C.POINTER(C.c_int32))
        self.desc = d


def _alloc_outputs(W, H, want):
    out = {}
    if "color" in want:
        out["color"] = np.empty((H, W, 3), np.uint8)
    if "depth" in want:
        out["depth"] = np.empty((H, W), np.float32)
    if "normal" in want:
        out["normal"] = np.empty((H, W, 3), np.float32)
    if "mask" in want:
        out["mask"] = np.empty((H, W), np.uint8)
    if "face_id" in want:
        out["face_id"] = np.empty((H, W), np.int32)
    return out


ALL_OUTPUTS = ("color", "depth", "normal", "mask", "face_id")


def _optr(out, k):
    return out[k].ctypes.data_as(C.c_void_p) if k in out else None


class OraclePort:
    """The plain-C restatement.  Takes the float camera constants (a ugu_b200 `Camera` POD or any
    object with the same fields) so that it and the CUDA path see identical inputs."""

    def __init__(self, mesh):
        build_port()
        self.lib = C.CDLL(str(PORT_LIB))
        self.lib.oracle_render.restype = C.c_int
        self.lib.oracle_render.argtypes = [C.POINTER(_MeshDesc), C.POINTER(_PortCamera),
                                           C.POINTER(_PortOption)] + [C.c_void_p] * 5
        self.holder = _MeshHolder(mesh)

    def render(self, cam_pod, option, want=ALL_OUTPUTS):
        cam = _PortCamera()
        for name, _ in _PortCamera._fields_:
            setattr(cam, name, getattr(cam_pod, name))
        opt = _PortOption(int(option.diffuse_color), int(option.interp), int(option.shading_normal),
                          int(option.diffuse_shading), 1 if option.backface_culling else 0,
                          float(option.depth_scale), 0.3)  # CopyTo never copies sigma
        out = _alloc_outputs(cam.width, cam.height, want)
        rc = self.lib.oracle_render(C.byref(self.holder.desc), C.byref(cam), C.byref(opt),
                                    *[_optr(out, k) for k in ALL_OUTPUTS])
        if rc != 0:
            return None
        return out


def port_raytrace(port: "OraclePort", cam_pod, org_w, option, want=ALL_OUTPUTS):
    """oracle_raytrace (oracle/raster_oracle.c): ugu::Raytracer::Render restated with a brute-force nearest
    hit over nanort's intersector.  `org_w` = c2w translation as float32.  None on failure."""
    lib = port.lib
    lib.oracle_raytrace.restype = C.c_int
    lib.oracle_raytrace.argtypes = [C.POINTER(_MeshDesc), C.POINTER(_PortCamera), C.POINTER(C.c_float),
                                    C.POINTER(_PortOption)] + [C.c_void_p] * 5
    cam = _PortCamera()
    for name, _ in _PortCamera._fields_:
        setattr(cam, name, getattr(cam_pod, name))
    opt = _PortOption(int(option.diffuse_color), int(option.interp), int(option.shading_normal),
                      int(option.diffuse_shading), 1 if option.backface_culling else 0,
                      float(option.depth_scale), 0.3)
    org = (C.c_float * 3)(*[float(np.float32(x)) for x in org_w])
    out = _alloc_outputs(cam.width, cam.height, want)
    rc = lib.oracle_raytrace(C.byref(port.holder.desc), C.byref(cam), org, C.byref(opt),
                             *[_optr(out, k) for k in ALL_OUTPUTS])
    return out if rc == 0 else None


GBUFFER_FIELDS = (("color", np.uint8, 3), ("normal_cam", np.float32, 3), ("normal_wld", np.float32, 3),
                  ("pos_cam", np.float32, 3), ("pos_wld", np.float32, 3), ("depth_01", np.float32, 1),
                  ("stencil", np.uint8, 1), ("face_id", np.int32, 1), ("uv", np.float32, 3),
                  ("bary", np.float32, 3), ("geo_id", np.int32, 1))


class _PortGBuffer(C.Structure):
    _fields_ = [(n, C.c_void_p) for n, _, _ in GBUFFER_FIELDS]


def port_gbuffer(port: "OraclePort", cam_pod, option, z_near, z_far, flat_normal=False, geo_id=0, fields=None):
    """oracle_gbuffer (oracle/raster_oracle.c): the G-buffer restatement on the port's winners.  PARITY
    UNPINNED against the reference's OpenGL output (see the C header comment).  None on failure."""
    lib = port.lib
    lib.oracle_gbuffer.restype = C.c_int
    lib.oracle_gbuffer.argtypes = [C.POINTER(_MeshDesc), C.POINTER(_PortCamera), C.POINTER(_PortOption),
                                   C.c_float, C.c_float, C.c_int32, C.c_int32, C.POINTER(_PortGBuffer)]
    cam = _PortCamera()
    for name, _ in _PortCamera._fields_:
        setattr(cam, name, getattr(cam_pod, name))
    opt = _PortOption(int(option.diffuse_color), int(option.interp), int(option.shading_normal),
                      int(option.diffuse_shading), 1 if option.backface_culling else 0,
                      float(option.depth_scale), 0.3)
    out, g = {}, _PortGBuffer()
    for n, dt, ch in GBUFFER_FIELDS:
        if fields is not None and n not in fields:
            continue
        out[n] = np.empty((cam.height, cam.width) if ch == 1 else (cam.height, cam.width, ch), dt)
        setattr(g, n, out[n].ctypes.data)
    rc = lib.oracle_gbuffer(C.byref(port.holder.desc), C.byref(cam), C.byref(opt), float(z_near), float(z_far),
                            1 if flat_normal else 0, int(geo_id), C.byref(g))
    return out if rc == 0 else None


def port_depth2pointcloud(cam_pod, depth, point_cloud=None, gl_coord=False):
    """Plain-C restatement of ugu::Depth2PointCloud (organized Image3f form, rgbd_util.cc:472-513).
    `point_cloud` (H, W, 3) is the in/out image; None = a zero image."""
    build_port()
    lib = C.CDLL(str(PORT_LIB))
    lib.oracle_depth2pointcloud.restype = C.c_int
    lib.oracle_depth2pointcloud.argtypes = [C.POINTER(_PortCamera), C.c_void_p, C.c_void_p, C.c_int]
    cam = _PortCamera()
    for name, _ in _PortCamera._fields_:
        setattr(cam, name, getattr(cam_pod, name))
    d = np.ascontiguousarray(depth, np.float32)
    pc = np.zeros((cam.height, cam.width, 3), np.float32) if point_cloud is None else \
        np.array(point_cloud, np.float32, copy=True, order="C")
    rc = lib.oracle_depth2pointcloud(C.byref(cam), d.ctypes.data, pc.ctypes.data, 1 if gl_coord else 0)
    return pc if rc == 0 else None


def ref_depth2pointcloud(camera, depth, point_cloud=None, gl_coord=False):
    """The compiled reference's ugu::Depth2PointCloud through oracle/_ref."""
    lib = C.CDLL(str(REF_LIB))
    lib.ugu_ref_depth2pointcloud.restype = C.c_int
    lib.ugu_ref_depth2pointcloud.argtypes = [C.POINTER(_RefCamera), C.c_void_p, C.c_void_p, C.c_int]
    cam = _RefCamera()
    cam.width, cam.height = camera.width(), camera.height()
    cam.fx, cam.fy = float(camera.focal_length()[0]), float(camera.focal_length()[1])
    cam.cx, cam.cy = float(camera.principal_point()[0]), float(camera.principal_point()[1])
    cam.c2w = (C.c_double * 16)(*np.asarray(camera.c2w(), np.float64).reshape(-1))
    d = np.ascontiguousarray(depth, np.float32)
    pc = np.zeros((cam.height, cam.width, 3), np.float32) if point_cloud is None else \
        np.array(point_cloud, np.float32, copy=True, order="C")
    rc = lib.ugu_ref_depth2pointcloud(C.byref(cam), d.ctypes.data, pc.ctypes.data, 1 if gl_coord else 0)
    return pc if rc == 0 else None


def _uv_raster_call(lib, fn_name, attrs, attr_indices, uv, uv_indices, width, height, texture_type,
                    texture, mask):
    fn = getattr(lib, fn_name)
    fn.restype = C.c_int
    fn.argtypes = [C.c_int32] * 4 + [C.c_void_p] * 4 + [C.c_int32] * 4 + [C.c_void_p] * 2
    a = np.ascontiguousarray(attrs, np.float32).reshape(-1, 3)
    ui = np.ascontiguousarray(uv_indices, np.int32).reshape(-1, 3)
    u = np.ascontiguousarray(uv, np.float32).reshape(-1, 2)
    per_face = attr_indices is None
    ai = None if per_face else np.ascontiguousarray(attr_indices, np.int32).reshape(-1, 3)
    if ai is not None and len(ai) != len(ui):
        return None  # raster_util.h:148-151
    dt = np.uint8 if texture_type == 0 else np.float32
    keep = texture is not None
    tex = np.array(texture, dt, copy=True, order="C").reshape(height, width, 3) if keep else \
        np.zeros((height, width, 3), dt)
    m = None if mask is None else np.array(mask, np.uint8, copy=True, order="C").reshape(height, width)
    rc = fn(len(a), len(ui), len(u), 1 if per_face else 0, a.ctypes.data,
            None if ai is None else ai.ctypes.data, u.ctypes.data, ui.ctypes.data, width, height,
            texture_type, 1 if keep else 0, tex.ctypes.data, None if m is None else m.ctypes.data)
    return (tex, m) if rc == 0 else None


def port_uv_raster(attrs, attr_indices, uv, uv_indices, width, height, texture_type=0, texture=None,
                   mask=None):
    """Plain-C restatement of ugu::RasterizeVertexAttributeToTexture (attr_indices given) /
    RasterizeFaceAttributeToTexture (attr_indices None) with num_threads = 1
    (raster_util.h:139-268).  texture_type 0 = Image3b, 1 = Image3f; `texture` not None = rasterise
    over its contents (the reference's width/height <= 0 form); `mask` = in/out Image1b or None.
    Returns (texture, mask) or None when the reference returns false."""
    build_port()
    return _uv_raster_call(C.CDLL(str(PORT_LIB)), "oracle_uv_raster", attrs, attr_indices, uv,
                           uv_indices, width, height, texture_type, texture, mask)


def ref_uv_raster(attrs, attr_indices, uv, uv_indices, width, height, texture_type=0, texture=None,
                  mask=None):
    """The compiled reference's templates through oracle/_ref (num_threads = 1)."""
    return _uv_raster_call(C.CDLL(str(REF_LIB)), "ugu_ref_uv_raster", attrs, attr_indices, uv,
                           uv_indices, width, height, texture_type, texture, mask)


def _uv_mask_call(lib, fn_name, uv, uv_indices, width, height, color):
    fn = getattr(lib, fn_name)
    fn.restype = C.c_int
    fn.argtypes = [C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_void_p]
    ui = np.ascontiguousarray(uv_indices, np.int32).reshape(-1, 3)
    u = np.ascontiguousarray(uv, np.float32).reshape(-1, 2)
    m = np.zeros((height, width), np.uint8)
    rc = fn(len(ui), len(u), u.ctypes.data, ui.ctypes.data, width, height, int(color), m.ctypes.data)
    return m if rc == 0 else None


def port_uv_mask(uv, uv_indices, width, height, color=255):
    """Plain-C restatement of ugu::GenerateUvMask (Image1b form, src/util/raster_util.cc:9-27)."""
    build_port()
    return _uv_mask_call(C.CDLL(str(PORT_LIB)), "oracle_uv_mask", uv, uv_indices, width, height, color)


def ref_uv_mask(uv, uv_indices, width, height, color=255):
    return _uv_mask_call(C.CDLL(str(REF_LIB)), "ugu_ref_uv_mask", uv, uv_indices, width, height, color)


def face_visibility(face_id_maps, num_faces, first_view_id=0):
    """numpy restatement of the face-visibility consumer (include/ugu_b200_raster.h,
    ugu_b200_face_visibility).  PARITY UNPINNED: the reference has no function that consumes face_id
    maps (its VisibilityTester shoots BVH rays), so this checks the CUDA path against the stated
    definition only: pixel_count / view_count / best_view (most pixels, ties -> lowest id, -1 =
    never) / best_count."""
    maps = np.asarray(face_id_maps, np.int32)
    if maps.ndim == 2:
        maps = maps[None]
    pix = np.zeros(num_faces, np.int64)
    views = np.zeros(num_faces, np.int64)
    best_view = np.full(num_faces, -1, np.int64)
    best_count = np.zeros(num_faces, np.int64)
    for v, m in enumerate(maps):
        ids = m[(m >= 0) & (m < num_faces)]
        c = np.bincount(ids, minlength=num_faces)
        pix += c
        views += c > 0
        better = c > best_count  # strict: an equal count keeps the earlier (lower) view id
        best_view[better] = first_view_id + v
        best_count[better] = c[better]
    return dict(pixel_count=pix.astype(np.uint32), view_count=views.astype(np.uint32),
                best_view=best_view.astype(np.int32), best_count=best_count.astype(np.uint32))


def _d2m_pack(nv, nf, has_uv, has_col, have_normals, bufs):
    v, f, n, fn, uv, col, pc, ni = bufs
    return dict(vertices=v[:nv].copy(), faces=f[:nf].copy(),
                normals=n[:nv].copy() if have_normals else np.zeros((0, 3), np.float32),
                face_normals=fn[:nf].copy(), uv=uv[:nv].copy() if has_uv else np.zeros((0, 2), np.float32),
                colors=col[:nv].copy() if has_col else np.zeros((0, 3), np.float32),
                point_cloud=pc, normal_image=ni)


def _d2m_bufs(W, H, x_step, y_step):
    G = max(1, ((W - 1) // max(x_step, 1)) * ((H - 1) // max(y_step, 1)))
    return (np.zeros((G, 3), np.float32), np.zeros((2 * G, 3), np.int32), np.zeros((G, 3), np.float32),
            np.zeros((2 * G, 3), np.float32), np.zeros((G, 2), np.float32), np.zeros((G, 3), np.float32),
            np.zeros((H, W, 3), np.float32), np.zeros((H, W, 3), np.float32))


def two_camera_constants(depth2color, color_camera):
    """The 16 floats the two-camera Depth2Mesh needs beside the depth camera: depth2color's rotation (row-major 9)
    and translation (3), then fx, fy, cx, cy of the colour camera."""
    m = np.asarray(depth2color, np.float32).reshape(4, 4)
    f, c = color_camera.focal_length(), color_camera.principal_point()
    return np.concatenate([m[:3, :3].reshape(-1), m[:3, 3], np.asarray([f[0], f[1], c[0], c[1]], np.float32)]
                          ).astype(np.float32)


def port_depth2mesh(cam_pod, depth, max_connect_z_diff, x_step=1, y_step=1, gl_coord=False, color=None,
                    with_vertex_color=False, two_cam=None):
    """Plain-C restatement of ugu::Depth2Mesh (rgbd_util.cc:94-282 + Mesh::CalcNormal).  color None =
    the depth-only overload, else the textured one; two_cam (two_camera_constants) = the two-camera overload
    (rgbd_util.cc:283-466).  Returns a dict of arrays or None (= false)."""
    build_port()
    lib = C.CDLL(str(PORT_LIB))
    fn = lib.oracle_depth2mesh if two_cam is None else lib.oracle_depth2mesh2
    fn.restype = C.c_int
    fn.argtypes = ([C.POINTER(_PortCamera)] + ([] if two_cam is None else [C.c_void_p]) +
                   [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_float,
                    C.c_int32, C.c_int32, C.c_int32] + [C.c_void_p] * 9)
    cam = _PortCamera()
    for name, _ in _PortCamera._fields_:
        setattr(cam, name, getattr(cam_pod, name))
    d = np.ascontiguousarray(depth, np.float32)
    col = None if color is None else np.ascontiguousarray(color, np.uint8)
    bufs = _d2m_bufs(cam.width, cam.height, x_step, y_step)
    counts = np.zeros(3, np.int32)
    tc = None if two_cam is None else np.ascontiguousarray(two_cam, np.float32)
    rc = fn(C.byref(cam), *([] if tc is None else [tc.ctypes.data]), d.ctypes.data,
            None if col is None else col.ctypes.data,
            0 if col is None else col.shape[1], 0 if col is None else col.shape[0],
            1 if with_vertex_color else 0, float(max_connect_z_diff), x_step, y_step, 1 if gl_coord else 0,
            *[b.ctypes.data for b in bufs], counts.ctypes.data)
    if rc != 0:
        return None
    return _d2m_pack(int(counts[0]), int(counts[1]), col is not None, col is not None and with_vertex_color,
                     bool(counts[2]), bufs)


def ref_depth2mesh(camera, depth, max_connect_z_diff, x_step=1, y_step=1, gl_coord=False, color=None,
                   with_vertex_color=False):
    """The compiled reference's ugu::Depth2Mesh through oracle/_ref."""
    lib = C.CDLL(str(REF_LIB))
    lib.ugu_ref_depth2mesh.restype = C.c_int
    lib.ugu_ref_depth2mesh.argtypes = [C.POINTER(_RefCamera), C.c_void_p, C.c_void_p, C.c_int32, C.c_int32,
                                       C.c_int32, C.c_float, C.c_int32, C.c_int32, C.c_int32, C.c_void_p]
    lib.ugu_ref_depth2mesh_copy.restype = C.c_int
    lib.ugu_ref_depth2mesh_copy.argtypes = [C.c_void_p] * 8
    cam = _RefCamera()
    cam.width, cam.height = camera.width(), camera.height()
    cam.fx, cam.fy = float(camera.focal_length()[0]), float(camera.focal_length()[1])
    cam.cx, cam.cy = float(camera.principal_point()[0]), float(camera.principal_point()[1])
    cam.c2w = (C.c_double * 16)(*np.asarray(camera.c2w(), np.float64).reshape(-1))
    d = np.ascontiguousarray(depth, np.float32)
    col = None if color is None else np.ascontiguousarray(color, np.uint8)
    out = np.zeros(6, np.int32)
    rc = lib.ugu_ref_depth2mesh(C.byref(cam), d.ctypes.data, None if col is None else col.ctypes.data,
                                0 if col is None else col.shape[1], 0 if col is None else col.shape[0],
                                1 if with_vertex_color else 0, float(max_connect_z_diff), x_step, y_step,
                                1 if gl_coord else 0, out.ctypes.data)
    if rc != 0:
        return None
    nv, nf = int(out[0]), int(out[1])
    bufs = _d2m_bufs(cam.width, cam.height, x_step, y_step)
    have_normals = nf > 0
    ptrs = [b.ctypes.data for b in bufs]
    if not have_normals:
        ptrs[2] = None
        ptrs[7] = None  # the reference reads normals()[i] of an empty vector here: undefined
    bad = lib.ugu_ref_depth2mesh_copy(*ptrs)
    assert bad == 0, "normal_indices / uv_indices differ from vertex_indices"
    return _d2m_pack(nv, nf, bool(out[2]), bool(out[3]), have_normals, bufs)


def _ref_camera(camera):
    cam = _RefCamera()
    cam.width, cam.height = camera.width(), camera.height()
    cam.fx, cam.fy = float(camera.focal_length()[0]), float(camera.focal_length()[1])
    cam.cx, cam.cy = float(camera.principal_point()[0]), float(camera.principal_point()[1])
    cam.c2w = (C.c_double * 16)(*np.asarray(camera.c2w(), np.float64).reshape(-1))
    return cam


def ref_depth2mesh2(depth_camera, color_camera, depth2color, depth, color, max_connect_z_diff, x_step=1,
                    y_step=1, gl_coord=False, with_vertex_color=False):
    """The compiled reference's two-camera ugu::Depth2Mesh (rgbd_util.h:39-48) through oracle/_ref.
    depth2color: 4x4 float32 (Eigen::Affine3f); color: (color_camera.height, color_camera.width, 3) u8."""
    lib = C.CDLL(str(REF_LIB))
    lib.ugu_ref_depth2mesh2.restype = C.c_int
    lib.ugu_ref_depth2mesh2.argtypes = [C.POINTER(_RefCamera), C.POINTER(_RefCamera), C.c_void_p, C.c_void_p,
                                        C.c_void_p, C.c_int32, C.c_float, C.c_int32, C.c_int32, C.c_int32,
                                        C.c_void_p]
    lib.ugu_ref_depth2mesh_copy.restype = C.c_int
    lib.ugu_ref_depth2mesh_copy.argtypes = [C.c_void_p] * 8
    dcam, ccam = _ref_camera(depth_camera), _ref_camera(color_camera)
    d = np.ascontiguousarray(depth, np.float32)
    col = np.ascontiguousarray(color, np.uint8)
    assert col.shape[:2] == (ccam.height, ccam.width)
    m = np.ascontiguousarray(depth2color, np.float32).reshape(4, 4)
    out = np.zeros(6, np.int32)
    rc = lib.ugu_ref_depth2mesh2(C.byref(dcam), C.byref(ccam), m.ctypes.data, d.ctypes.data, col.ctypes.data,
                                 1 if with_vertex_color else 0, float(max_connect_z_diff), x_step, y_step,
                                 1 if gl_coord else 0, out.ctypes.data)
    if rc != 0:
        return None
    nv, nf = int(out[0]), int(out[1])
    bufs = _d2m_bufs(dcam.width, dcam.height, x_step, y_step)
    have_normals = nf > 0
    ptrs = [b.ctypes.data for b in bufs]
    if not have_normals:
        ptrs[2] = None
        ptrs[7] = None
    bad = lib.ugu_ref_depth2mesh_copy(*ptrs)
    assert bad == 0, "normal_indices / uv_indices differ from vertex_indices"
    return _d2m_pack(nv, nf, bool(out[2]), bool(out[3]), have_normals, bufs)


class OracleRef:
    """The compiled, unmodified reference ugu::Rasterizer (present only where oracle/_ref was built)."""

    def __init__(self, mesh=None):
        if not REF_LIB.exists():
            raise FileNotFoundError(f"{REF_LIB} not built (needs /root/reference: make -C oracle ref)")
        lib = C.CDLL(str(REF_LIB))
        lib.ugu_ref_create.restype = C.c_void_p
        lib.ugu_ref_create.argtypes = [C.POINTER(_MeshDesc)]
        lib.ugu_ref_destroy.argtypes = [C.c_void_p]
        lib.ugu_ref_render.restype = C.c_int
        lib.ugu_ref_render.argtypes = [C.c_void_p, C.POINTER(_RefCamera), C.POINTER(_RefOption)] + \
            [C.c_void_p] * 5 + [C.c_int, C.POINTER(C.c_double)]
        lib.ugu_ref_camera_constants.argtypes = [C.POINTER(C.c_double)] + [C.POINTER(C.c_float)] * 3
        lib.ugu_ref_lookat_c2w.argtypes = [C.POINTER(C.c_float)] * 3 + [C.POINTER(C.c_double)]
        lib.ugu_ref_mesh_load_obj.restype = C.c_void_p
        lib.ugu_ref_mesh_load_obj.argtypes = [C.c_char_p, C.c_char_p]
        lib.ugu_ref_mesh_free.argtypes = [C.c_void_p]
        lib.ugu_ref_mesh_sizes.argtypes = [C.c_void_p, C.POINTER(C.c_int32)]
        lib.ugu_ref_mesh_copy.argtypes = [C.c_void_p] * 11
        self.lib = lib
        self.h = None
        self.holder = None
        self.last_timing_ms = (0.0, 0.0)
        if mesh is not None:
            self.holder = _MeshHolder(mesh)
            self.h = lib.ugu_ref_create(C.byref(self.holder.desc))
            if not self.h:
                raise RuntimeError("ugu_ref_create failed")

    def __del__(self):
        if getattr(self, "h", None):
            self.lib.ugu_ref_destroy(self.h)
            self.h = None

    def render(self, camera, option, want=ALL_OUTPUTS, repeat_warm=False):
        """`camera` is a ugu_b200.PinholeCamera-like object (width(), height(), c2w(),
        principal_point(), focal_length())."""
        cam = _RefCamera()
        cam.width, cam.height = camera.width(), camera.height()
        cam.fx, cam.fy = float(camera.focal_length()[0]), float(camera.focal_length()[1])
        cam.cx, cam.cy = float(camera.principal_point()[0]), float(camera.principal_point()[1])
        cam.c2w = (C.c_double * 16)(*np.asarray(camera.c2w(), np.float64).reshape(-1))
        opt = _RefOption(int(option.diffuse_color), int(option.interp), int(option.shading_normal),
                         int(option.diffuse_shading), 1 if option.backface_culling else 0,
                         float(option.depth_scale))
        out = _alloc_outputs(cam.width, cam.height, want)
        timing = (C.c_double * 2)()
        rc = self.lib.ugu_ref_render(self.h, C.byref(cam), C.byref(opt),
                                     *[_optr(out, k) for k in ALL_OUTPUTS],
                                     1 if repeat_warm else 0, timing)
        self.last_timing_ms = (timing[0], timing[1])
        if rc != 0:
            return None
        return out

    def set_num_threads(self, n: int) -> None:
        """OpenMP threads of the compiled reference from now on (the ray-table rebuild is its only parallel part)."""
        self.lib.ugu_ref_set_num_threads.argtypes = [C.c_int]
        self.lib.ugu_ref_set_num_threads(int(n))

    def raytrace(self, camera, option, want=ALL_OUTPUTS):
        """The compiled reference ugu::Raytracer (src/renderer/cpu/raytracer.cc, nanort BVH) on this mesh."""
        self.lib.ugu_ref_raytrace.restype = C.c_int
        self.lib.ugu_ref_raytrace.argtypes = [C.c_void_p, C.POINTER(_RefCamera), C.POINTER(_RefOption)] + \
            [C.c_void_p] * 5 + [C.POINTER(C.c_double)]
        cam = _RefCamera()
        cam.width, cam.height = camera.width(), camera.height()
        cam.fx, cam.fy = float(camera.focal_length()[0]), float(camera.focal_length()[1])
        cam.cx, cam.cy = float(camera.principal_point()[0]), float(camera.principal_point()[1])
        cam.c2w = (C.c_double * 16)(*np.asarray(camera.c2w(), np.float64).reshape(-1))
        opt = _RefOption(int(option.diffuse_color), int(option.interp), int(option.shading_normal),
                         int(option.diffuse_shading), 1 if option.backface_culling else 0,
                         float(option.depth_scale))
        out = _alloc_outputs(cam.width, cam.height, want)
        timing = (C.c_double * 1)()
        rc = self.lib.ugu_ref_raytrace(self.h, C.byref(cam), C.byref(opt),
                                       *[_optr(out, k) for k in ALL_OUTPUTS], timing)
        self.last_raytrace_ms = timing[0]
        return out if rc == 0 else None

    def camera_constants(self, c2w):
        m = (C.c_double * 16)(*np.asarray(c2w, np.float64).reshape(-1))
        R, t, Cm = (C.c_float * 9)(), (C.c_float * 3)(), (C.c_float * 9)()
        self.lib.ugu_ref_camera_constants(m, R, t, Cm)
        return (np.array(R, np.float32), np.array(t, np.float32), np.array(Cm, np.float32))

    def lookat_c2w(self, eye, center, up):
        f3 = lambda a: (C.c_float * 3)(*[float(x) for x in a])
        m = (C.c_double * 16)()
        self.lib.ugu_ref_lookat_c2w(f3(eye), f3(center), f3(up), m)
        return np.array(m, np.float64).reshape(4, 4)

    def load_obj(self, obj_path: str, mtl_dir: str):
        """Load an OBJ through the reference's own loader; returns a dict of numpy arrays."""
        h = self.lib.ugu_ref_mesh_load_obj(obj_path.encode(), mtl_dir.encode())
        if not h:
            raise RuntimeError(f"reference LoadObj failed for {obj_path}")
        sz = (C.c_int32 * 9)()
        self.lib.ugu_ref_mesh_sizes(h, sz)
        V, Nn, Nc, F, Nfn, Nuv, M, tw, th = list(sz)
        a = dict(vertices=np.zeros((V, 3), np.float32), normals=np.zeros((Nn, 3), np.float32),
                 vertex_colors=np.zeros((Nc, 3), np.float32),
                 vertex_indices=np.zeros((F, 3), np.int32), normal_indices=np.zeros((F, 3), np.int32),
                 face_normals=np.zeros((Nfn, 3), np.float32), uv=np.zeros((Nuv, 2), np.float32),
                 uv_indices=np.zeros((F, 3), np.int32), material_ids=np.zeros((F,), np.int32),
                 tex0=np.zeros((th, tw, 3), np.uint8))
        p = lambda k: a[k].ctypes.data_as(C.c_void_p) if a[k].size else None
        self.lib.ugu_ref_mesh_copy(h, p("vertices"), p("normals"), p("vertex_colors"),
                                   p("vertex_indices"), p("normal_indices"), p("face_normals"),
                                   p("uv"), p("uv_indices"), p("material_ids"), p("tex0"))
        self.lib.ugu_ref_mesh_free(h)
        a["num_materials"] = M
        return a


# ---- ugu::VisibilityTester through oracle/_ref (reference-pinned side of the visibility consumer) -------------
class _RefKeyframe(C.Structure):
    _fields_ = [("id", C.c_int32), ("pad_", C.c_int32), ("cam", _RefCamera), ("color", C.c_void_p),
                ("depth", C.c_void_p), ("mask", C.c_void_p)]


class _RefVisOption(C.Structure):
    _fields_ = [("backface_culling_angle_rad_th", C.c_float), ("use_mask", C.c_int32), ("use_depth", C.c_int32),
                ("interp", C.c_int32), ("facewise", C.c_int32), ("num_threads", C.c_int32)]


def _ref_camera(camera) -> "_RefCamera":
    cam = _RefCamera()
    cam.width, cam.height = camera.width(), camera.height()
    cam.fx, cam.fy = float(camera.focal_length()[0]), float(camera.focal_length()[1])
    cam.cx, cam.cy = float(camera.principal_point()[0]), float(camera.principal_point()[1])
    cam.c2w = (C.c_double * 16)(*np.asarray(camera.c2w(), np.float64).reshape(-1))
    return cam


VERTEX_STAT_COLORS = ("mean", "median", "min_viewing_angle", "mean_viewing_angle", "median_viewing_angle",
                      "min_distance", "mean_distance", "median_distance", "median_intensity", "min_intensity")
VERTEX_STAT_SCALARS = ("mean_viewing_angle", "median_viewing_angle", "mean_distance", "median_distance",
                       "min_intensity", "median_intensity", "num_visible")


def ref_visibility(vertices, normals, indices, keyframes, backface_culling_angle_rad_th=np.pi / 2, use_mask=True,
                   use_depth=True, interp=1, facewise=True, num_threads=1):
    """The compiled reference's ugu::VisibilityTester::Test(keyframes, info, facewise)
    (src/texturing/visibility_tester.cc:657-733) on `keyframes` = [(id, camera, color, depth, mask)] (host
    images; depth / mask may be None).  Returns dense per-(keyframe, vertex) and per-(keyframe, face) records
    plus the CalcStat results, or None when the reference returns false."""
    lib = C.CDLL(str(REF_LIB))
    lib.ugu_ref_visibility_run.restype = C.c_void_p
    lib.ugu_ref_visibility_run.argtypes = [C.c_int32, C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_int32,
                                           C.POINTER(_RefKeyframe), C.POINTER(_RefVisOption)]
    v = np.ascontiguousarray(vertices, np.float32).reshape(-1, 3)
    n = np.ascontiguousarray(normals, np.float32).reshape(-1, 3)
    f = np.ascontiguousarray(indices, np.int32).reshape(-1, 3)
    V, F, K = len(v), len(f), len(keyframes)
    keep = []
    kfs = (_RefKeyframe * K)()
    for k, (kid, cam, color, depth, mask) in enumerate(keyframes):
        kfs[k].id = int(kid)
        kfs[k].cam = _ref_camera(cam)
        c = np.ascontiguousarray(color, np.uint8)
        d = None if depth is None else np.ascontiguousarray(depth, np.float32)
        m = None if mask is None else np.ascontiguousarray(mask, np.uint8)
        keep += [c, d, m]
        kfs[k].color = c.ctypes.data
        kfs[k].depth = None if d is None else d.ctypes.data
        kfs[k].mask = None if m is None else m.ctypes.data
    opt = _RefVisOption(float(backface_culling_angle_rad_th), int(bool(use_mask)), int(bool(use_depth)), int(interp),
                        int(bool(facewise)), int(num_threads))
    h = lib.ugu_ref_visibility_run(V, v.ctypes.data, n.ctypes.data, F, f.ctypes.data, K, kfs, C.byref(opt))
    if not h:
        return None
    h = C.c_void_p(h)
    out = dict(
        v_visible=np.zeros((K, V), np.uint8), v_color=np.zeros((K, V, 3), np.float32),
        v_intensity=np.zeros((K, V), np.float32), v_projected=np.zeros((K, V, 2), np.float32),
        v_viewing_angle=np.zeros((K, V), np.float32), v_distance=np.zeros((K, V), np.float32),
        f_visible=np.zeros((K, F), np.uint8), f_area=np.zeros((K, F), np.float32),
        f_viewing_angle=np.zeros((K, F), np.float32), f_distance=np.zeros((K, F), np.float32),
        f_projected_tri=np.zeros((K, F, 6), np.float32),
        face_ids=np.zeros((F, 3), np.int32), vertex_ids=np.zeros((V, 2), np.int32),
        vertex_scalars=np.zeros((V, 7), np.float32), vertex_colors=np.zeros((V, 10, 3), np.float32))
    p = lambda name: out[name].ctypes.data_as(C.c_void_p)
    lib.ugu_ref_visibility_vertex_records.argtypes = [C.c_void_p] * 7
    lib.ugu_ref_visibility_vertex_records(h, p("v_visible"), p("v_color"), p("v_intensity"), p("v_projected"),
                                          p("v_viewing_angle"), p("v_distance"))
    lib.ugu_ref_visibility_face_records.argtypes = [C.c_void_p] * 6
    lib.ugu_ref_visibility_face_records(h, p("f_visible"), p("f_area"), p("f_viewing_angle"), p("f_distance"),
                                        p("f_projected_tri"))
    lib.ugu_ref_visibility_stats.argtypes = [C.c_void_p] * 5
    lib.ugu_ref_visibility_stats(h, p("face_ids"), p("vertex_ids"), p("vertex_scalars"), p("vertex_colors"))
    lib.ugu_ref_visibility_free.argtypes = [C.c_void_p]
    lib.ugu_ref_visibility_free(h)
    return out


def ref_visibility_camera_constants(camera):
    """(w2c 3x4 float, c2w_R 3x3 float, c2w_t 3 float) exactly as the tester reads them from ugu::PinholeCamera."""
    lib = C.CDLL(str(REF_LIB))
    lib.ugu_ref_visibility_camera_constants.argtypes = [C.POINTER(_RefCamera), C.c_void_p, C.c_void_p, C.c_void_p]
    w2c, R, t = np.zeros((3, 4), np.float32), np.zeros((3, 3), np.float32), np.zeros(3, np.float32)
    cam = _ref_camera(camera)
    lib.ugu_ref_visibility_camera_constants(C.byref(cam), w2c.ctypes.data, R.ctypes.data, t.ctypes.data)
    return w2c, R, t


class _VisCamera(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("fx", C.c_float), ("fy", C.c_float),
                ("cx", C.c_float), ("cy", C.c_float), ("w2c", C.c_float * 12), ("c2w_R", C.c_float * 9),
                ("c2w_t", C.c_float * 3)]


class _VisOption(C.Structure):
    _fields_ = [("backface_culling_angle_rad_th", C.c_float), ("use_mask", C.c_int32), ("use_depth", C.c_int32),
                ("interp", C.c_int32), ("max_viewing_angle", C.c_float),
                ("max_distance_from_camera_to_vertex", C.c_float), ("max_depth_difference", C.c_float),
                ("eps", C.c_float), ("facewise", C.c_int32), ("reserved", C.c_int32)]


VIS_VERTEX_DTYPE = np.dtype([("color", np.float32, 3), ("intensity", np.float32), ("projected", np.float32, 2),
                             ("viewing_angle", np.float32), ("distance", np.float32)])


def vis_option(backface_culling_angle_rad_th=None, use_mask=True, use_depth=True, interp=1, facewise=True,
               max_viewing_angle=None, max_distance=None, max_depth_difference=None, eps=0.1):
    """Defaults of ugu::VisibilityTesterOption (visibility_tester.h:39-58)."""
    half_pi = float(np.float32(90.0) * np.float32(np.pi / 180.0))  # ugu::radians(90.0f)
    fmax = float(np.finfo(np.float32).max)
    return _VisOption(half_pi if backface_culling_angle_rad_th is None else float(backface_culling_angle_rad_th),
                      int(bool(use_mask)), int(bool(use_depth)), int(interp),
                      half_pi if max_viewing_angle is None else float(max_viewing_angle),
                      fmax if max_distance is None else float(max_distance),
                      fmax if max_depth_difference is None else float(max_depth_difference), float(eps),
                      int(bool(facewise)), 0)


def vis_camera(cam_consts, width, height, fx, fy, cx, cy):
    """cam_consts = (w2c 3x4, c2w_R 3x3, c2w_t 3) float32."""
    w2c, R, t = cam_consts
    return _VisCamera(int(width), int(height), float(fx), float(fy), float(cx), float(cy),
                      (C.c_float * 12)(*np.asarray(w2c, np.float32).reshape(-1)),
                      (C.c_float * 9)(*np.asarray(R, np.float32).reshape(-1)),
                      (C.c_float * 3)(*np.asarray(t, np.float32).reshape(-1)))


def port_visibility_keyframe(vertices, normals, indices, cam: "_VisCamera", opt: "_VisOption", color, depth, mask,
                             zbuf):
    """oracle_visibility_keyframe (oracle/raster_oracle.c): one keyframe of the visibility pass with the
    z-buffer in place of the BVH query.  Returns dict(v_visible, v_rec (structured), f_visible, f_area,
    f_viewing_angle, f_distance, f_projected_tri)."""
    build_port()
    lib = C.CDLL(str(PORT_LIB))
    fn = lib.oracle_visibility_keyframe
    fn.restype = C.c_int
    fn.argtypes = [C.c_int32, C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.POINTER(_VisCamera),
                   C.POINTER(_VisOption)] + [C.c_void_p] * 11
    v = np.ascontiguousarray(vertices, np.float32).reshape(-1, 3)
    n = np.ascontiguousarray(normals, np.float32).reshape(-1, 3)
    f = np.ascontiguousarray(indices, np.int32).reshape(-1, 3)
    V, F = len(v), len(f)
    c = np.ascontiguousarray(color, np.uint8)
    d = None if depth is None else np.ascontiguousarray(depth, np.float32)
    m = None if mask is None else np.ascontiguousarray(mask, np.uint8)
    z = None if zbuf is None else np.ascontiguousarray(zbuf, np.float32)
    out = dict(v_visible=np.zeros(V, np.uint8), v_rec=np.zeros(V, VIS_VERTEX_DTYPE),
               f_visible=np.zeros(F, np.uint8), f_area=np.zeros(F, np.float32),
               f_viewing_angle=np.zeros(F, np.float32), f_distance=np.zeros(F, np.float32),
               f_projected_tri=np.zeros((F, 6), np.float32))
    q = lambda a: None if a is None else a.ctypes.data
    rc = fn(V, v.ctypes.data, n.ctypes.data, F, f.ctypes.data, C.byref(cam), C.byref(opt), q(c), q(d), q(m), q(z),
            q(out["v_visible"]), q(out["v_rec"]), q(out["f_visible"]), q(out["f_area"]),
            q(out["f_viewing_angle"]), q(out["f_distance"]), q(out["f_projected_tri"]))
    return out if rc == 0 else None


def ref_camera_constants(c2w):
    """(w2c_R 9, w2c_t 3, c2w_R 9) float32 as the reference Rasterizer derives them (rasterizer.cc:118-119,
    camera.h:321) — Eigen's inverse() / rotation() through oracle/_ref, no renderer needed."""
    lib = C.CDLL(str(REF_LIB))
    lib.ugu_ref_camera_constants.argtypes = [C.POINTER(C.c_double)] + [C.POINTER(C.c_float)] * 3
    m = (C.c_double * 16)(*np.asarray(c2w, np.float64).reshape(-1))
    R, t, Cm = (C.c_float * 9)(), (C.c_float * 3)(), (C.c_float * 9)()
    lib.ugu_ref_camera_constants(m, R, t, Cm)
    return np.array(R, np.float32), np.array(t, np.float32), np.array(Cm, np.float32)
