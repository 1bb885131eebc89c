"""Host-side mirror of the reference renderer interface, on top of the C-ABI.

Same names, argument meaning, call order and error behaviour as the reference's
`ugu::RendererCpu` / `ugu::Rasterizer` (include/ugu/renderer/cpu/renderer.h:17-76,
src/renderer/cpu/rasterizer.cc:73-103,309-348), `ugu::PinholeCamera`
(include/ugu/camera.h:257-331,341-387) and the read side of `ugu::Mesh` (include/ugu/mesh.h:137-148):

    renderer = Rasterizer(option)      # ctor(option)
    renderer.set_mesh(mesh)            # set_mesh
    renderer.PrepareMesh()             # PrepareMesh -> the one-time H2D upload
    renderer.set_camera(camera)        # set_camera
    camera.set_c2w(c2w)                # callers mutate the camera the renderer holds
    ok = renderer.Render(color, depth, normal, mask, face_id)   # bool + logged message

The C++ flavour of the same thing, compiled against the reference headers, is
ugu_b200/cpp/rasterizer_cuda.{h,cc}.  Everything computed here runs in the CUDA library; without
it (or without a B200) construction fails loudly — there is no CPU path.
"""
from __future__ import annotations

import ctypes as C
import enum
import math
import sys
from dataclasses import dataclass, field
from typing import List, Optional, Sequence

import numpy as np

from . import _native as N


class DiffuseColor(enum.IntEnum):      # renderer/base.h:11-15
    kNone = 0
    kTexture = 1
    kVertex = 2


class ColorInterpolation(enum.IntEnum):  # common.h:50-53
    kNn = 0
    kBilinear = 1


class ShadingNormal(enum.IntEnum):     # renderer/base.h:19-22
    kFace = 0
    kVertex = 1


class DiffuseShading(enum.IntEnum):    # renderer/base.h:26-32
    kNone = 0
    kLambertian = 1
    kOrenNayar = 2


@dataclass
class RendererCpuOption:
    """renderer/cpu/renderer.h:17-38 (same defaults)."""
    diffuse_color: DiffuseColor = DiffuseColor.kNone
    interp: ColorInterpolation = ColorInterpolation.kBilinear
    shading_normal: ShadingNormal = ShadingNormal.kVertex
    diffuse_shading: DiffuseShading = DiffuseShading.kNone
    depth_scale: float = 1.0
    backface_culling: bool = True
    oren_nayar_sigma: float = 0.3

    def CopyTo(self, dst: "RendererCpuOption") -> None:
        # renderer.h:30-37 — oren_nayar_sigma is NOT copied (stays at the destination's 0.3).
        dst.diffuse_color = self.diffuse_color
        dst.depth_scale = self.depth_scale
        dst.interp = self.interp
        dst.shading_normal = self.shading_normal
        dst.diffuse_shading = self.diffuse_shading
        dst.backface_culling = self.backface_culling


def LOGE(msg: str) -> None:
    """The reference reports failures with LOGE(...) + `return false` (util_private.cc:19-67)."""
    sys.stderr.write("[ugu_b200 E] " + msg.rstrip("\n") + "\n")


def LOGW(msg: str) -> None:
    sys.stderr.write("[ugu_b200 W] " + msg.rstrip("\n") + "\n")


@dataclass
class Material:
    """The part of ugu::ObjMaterial the renderer reads: diffuse_tex (Image3b, H x W x 3 u8)."""
    diffuse_tex: Optional[np.ndarray] = None


@dataclass
class Mesh:
    """Read side of ugu::Mesh (mesh.h:83-98,137-148): the same AoS arrays, as numpy."""
    vertices: np.ndarray = field(default_factory=lambda: np.zeros((0, 3), np.float32))
    vertex_colors: np.ndarray = field(default_factory=lambda: np.zeros((0, 3), np.float32))
    vertex_indices: np.ndarray = field(default_factory=lambda: np.zeros((0, 3), np.int32))
    normals: np.ndarray = field(default_factory=lambda: np.zeros((0, 3), np.float32))
    face_normals: np.ndarray = field(default_factory=lambda: np.zeros((0, 3), np.float32))
    normal_indices: np.ndarray = field(default_factory=lambda: np.zeros((0, 3), np.int32))
    uv: np.ndarray = field(default_factory=lambda: np.zeros((0, 2), np.float32))
    uv_indices: np.ndarray = field(default_factory=lambda: np.zeros((0, 3), np.int32))
    material_ids: np.ndarray = field(default_factory=lambda: np.zeros((0,), np.int32))
    materials: List[Material] = field(default_factory=list)


class PinholeCamera:
    """ugu::PinholeCamera (camera.h:79-183): size, c2w pose (4x4 double), pixel-scale principal
    point and focal length.  The float constants Render needs are derived by the C-ABI
    (ugu_b200_camera_from_c2w) exactly when the renderer reads the camera."""

    def __init__(self, width: int, height: int, c2w: Optional[np.ndarray] = None,
                 principal_point: Optional[Sequence[float]] = None,
                 focal_length: Optional[Sequence[float]] = None,
                 fov_y_deg: Optional[float] = None):
        self._width, self._height = int(width), int(height)
        self.set_c2w(np.eye(4) if c2w is None else c2w)
        if principal_point is None:  # camera.h:276-277
            principal_point = (np.float32(self._width) * np.float32(0.5) - np.float32(0.5),
                               np.float32(self._height) * np.float32(0.5) - np.float32(0.5))
        self._pp = (np.float32(principal_point[0]), np.float32(principal_point[1]))
        if focal_length is not None:
            self._f = (np.float32(focal_length[0]), np.float32(focal_length[1]))
        else:
            self.set_fov_y(45.0 if fov_y_deg is None else fov_y_deg)

    def width(self) -> int:
        return self._width

    def height(self) -> int:
        return self._height

    def c2w(self) -> np.ndarray:
        return self._c2w

    def set_size(self, width: int, height: int) -> None:
        self._width, self._height = int(width), int(height)

    def set_c2w(self, c2w: np.ndarray) -> None:
        c2w = np.asarray(c2w, dtype=np.float64)
        if c2w.shape != (4, 4):
            raise ValueError("c2w must be 4x4")
        self._c2w = np.ascontiguousarray(c2w)

    def principal_point(self):
        return self._pp

    def focal_length(self):
        return self._f

    def set_principal_point(self, pp) -> None:
        self._pp = (np.float32(pp[0]), np.float32(pp[1]))

    def set_focal_length(self, f) -> None:
        self._f = (np.float32(f[0]), np.float32(f[1]))

    def set_fov_y(self, fov_y_deg: float) -> None:
        # camera.h:372-378: height * 0.5f / static_cast<float>(std::tan(radians(fov) * 0.5))
        rad = np.float32(np.float32(fov_y_deg) * np.float32(math.pi / 180.0))  # radians(float)
        t = np.float32(math.tan(float(rad) * 0.5))
        fy = np.float32(np.float32(self._height) * np.float32(0.5)) / t
        self._f = (np.float32(fy), np.float32(fy))

    def to_native(self) -> N.Camera:
        cam = N.Camera()
        c2w = (C.c_double * 16)(*self._c2w.reshape(-1))
        rc = N.load().ugu_b200_camera_from_c2w(c2w, self._width, self._height, float(self._f[0]),
                                               float(self._f[1]), float(self._pp[0]),
                                               float(self._pp[1]), C.byref(cam))
        if rc != N.OK:
            raise ValueError("c2w is not invertible")
        return cam


class _PinnedBlock:
    """Owner of one cudaHostAlloc block, exposed to numpy through __array_interface__: every array / view
    made from it keeps it alive through `.base`, and the block is freed when the last one is gone."""

    def __init__(self, lib, nbytes: int):
        self._lib = lib
        self.ptr = lib.ugu_b200_host_alloc(max(nbytes, 1))
        if not self.ptr:
            raise MemoryError(N.last_error())
        self.__array_interface__ = {"shape": (max(nbytes, 1),), "typestr": "|u1", "data": (self.ptr, False),
                                    "version": 3}

    def __del__(self):
        ptr, self.ptr = getattr(self, "ptr", None), None
        if ptr:
            try:
                self._lib.ugu_b200_host_free(ptr)
            except Exception:
                pass


def color_camera_native(color_camera: "PinholeCamera", depth2color) -> "N.ColorCamera":
    """ugu_b200_color_camera of the two-camera Depth2Mesh: depth2color (4x4, as Eigen::Affine3f) and the colour
    camera's intrinsics, all rounded to float like the reference's arguments."""
    m = np.asarray(depth2color, np.float32).reshape(4, 4)
    cc = N.ColorCamera()
    cc.depth2color_R = (C.c_float * 9)(*m[:3, :3].reshape(-1))
    cc.depth2color_t = (C.c_float * 3)(*m[:3, 3])
    f, c = color_camera.focal_length(), color_camera.principal_point()
    cc.fx, cc.fy, cc.cx, cc.cy = float(f[0]), float(f[1]), float(c[0]), float(c[1])
    return cc


def pinned_empty(shape, dtype) -> np.ndarray:
    """A numpy array over pinned host memory from the C-ABI (so Render's D2H is one DMA).  The memory
    lives exactly as long as the array and the views taken from it."""
    dtype = np.dtype(dtype)
    count = int(np.prod(shape))
    raw = np.asarray(_PinnedBlock(N.load(), count * dtype.itemsize))
    return raw[: count * dtype.itemsize].view(dtype).reshape(shape)


class Image:
    """Minimal ugu::Image<T> stand-in: `.data` is a contiguous numpy array (rows, cols[, ch]) that
    Render (re)allocates only when the size differs, like ugu::Init (image.h:963-985)."""

    dtype = np.uint8
    channels = 1

    def __init__(self, pinned: bool = True):
        self.rows, self.cols = 0, 0
        self.pinned = pinned
        self.data = np.zeros((0, 0) if self.channels == 1 else (0, 0, self.channels), self.dtype)

    def empty(self) -> bool:
        return self.rows < 1 or self.cols < 1

    def _init(self, width: int, height: int) -> None:
        if self.cols != width or self.rows != height:
            shape = (height, width) if self.channels == 1 else (height, width, self.channels)
            self.data = pinned_empty(shape, self.dtype) if self.pinned else np.empty(shape, self.dtype)
            self.rows, self.cols = height, width


class Image3b(Image):
    dtype, channels = np.uint8, 3


class Image1f(Image):
    dtype, channels = np.float32, 1


class Image3f(Image):
    dtype, channels = np.float32, 3


class Image1b(Image):
    dtype, channels = np.uint8, 1


class Image1i(Image):
    dtype, channels = np.int32, 1


class Image1w(Image):
    dtype, channels = np.uint16, 1


class GBuffer:
    """ugu::GBuffer (include/ugu/renderer/base.h:34-77) without `shaded` (the GL viewer's deferred
    lighting pass): the images RendererGl::ReadGbuf fills (src/renderer/gl/renderer.cc:501-650)."""

    TYPES = {"color": Image3b, "normal_cam": Image3f, "normal_wld": Image3f, "pos_cam": Image3f,
             "pos_wld": Image3f, "depth_01": Image1f, "stencil": Image1b, "face_id": Image1i, "uv": Image3f,
             "bary": Image3f, "geo_id": Image1i}

    def __init__(self, pinned: bool = True):
        for name, cls in self.TYPES.items():
            setattr(self, name, cls(pinned))

    def Init(self, w: int, h: int) -> None:  # base.h:48-61 (contents are rewritten by every render)
        for name in self.TYPES:
            getattr(self, name)._init(w, h)


def _ptr(a: Optional[np.ndarray]):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def mesh_desc(m: "Mesh"):
    """The C-ABI mesh descriptor of a Mesh (AoS host arrays as ugu::Mesh holds them, mesh.h:83-98) and the objects
    that must stay alive while the descriptor is in use."""
    f32 = lambda a, w: np.ascontiguousarray(np.asarray(a, np.float32).reshape(-1, w))
    i32 = lambda a, w: np.ascontiguousarray(np.asarray(a, np.int32).reshape(-1, w))
    v, n, c = f32(m.vertices, 3), f32(m.normals, 3), f32(m.vertex_colors, 3)
    vi, ni, ui = i32(m.vertex_indices, 3), i32(m.normal_indices, 3), i32(m.uv_indices, 3)
    fn, uv = f32(m.face_normals, 3), f32(m.uv, 2)
    mid = np.ascontiguousarray(np.asarray(m.material_ids, np.int32).reshape(-1))
    F = len(vi)
    texs = [None if (t.diffuse_tex is None or t.diffuse_tex.size == 0)
            else np.ascontiguousarray(t.diffuse_tex, np.uint8) for t in m.materials]
    M = len(texs)
    d = N.MeshDesc()
    d.num_vertices, d.num_normals, d.num_colors = len(v), len(n), len(c)
    d.num_faces, d.num_face_normals, d.num_uv, d.num_materials = F, len(fn), len(uv), M
    d.vertices, d.normals, d.vertex_colors = _ptr(v), _ptr(n), _ptr(c)
    d.vertex_indices = _ptr(vi)
    d.normal_indices = _ptr(ni) if len(ni) == F and F > 0 else None
    d.face_normals, d.uv = _ptr(fn), _ptr(uv)
    d.uv_indices = _ptr(ui) if len(ui) == F and F > 0 else None
    d.material_ids = _ptr(mid) if len(mid) == F and F > 0 else None
    tex_ptrs = (C.c_void_p * max(M, 1))(*[None if t is None else t.ctypes.data for t in texs])
    tex_w = (C.c_int32 * max(M, 1))(*[0 if t is None else t.shape[1] for t in texs])
    tex_h = (C.c_int32 * max(M, 1))(*[0 if t is None else t.shape[0] for t in texs])
    d.tex_data = C.cast(tex_ptrs, C.POINTER(C.c_void_p))
    d.tex_w = C.cast(tex_w, C.POINTER(C.c_int32))
    d.tex_h = C.cast(tex_h, C.POINTER(C.c_int32))
    return d, (v, n, c, vi, ni, ui, fn, uv, mid, texs, tex_ptrs, tex_w, tex_h)


class Rasterizer:
    """The CUDA-backed drop-in for ugu::Rasterizer (rasterizer.h:14-53)."""

    def __init__(self, option: Optional[RendererCpuOption] = None, device: int = 0):
        self._lib = N.load()
        self._h = C.c_void_p()
        rc = self._lib.ugu_b200_rasterizer_create(int(device), C.byref(self._h))
        if rc != N.OK:
            raise RuntimeError(f"ugu_b200: cannot create the CUDA rasterizer: {N.last_error()}")
        self._option = RendererCpuOption()
        self._mesh: Optional[Mesh] = None
        self._mesh_initialized = False
        self._camera: Optional[PinholeCamera] = None
        self._keep = None
        if option is not None:
            self.set_option(option)

    def __del__(self):
        h = getattr(self, "_h", None)
        if h:
            self._lib.ugu_b200_rasterizer_destroy(h)
            self._h = None

    # -- RendererCpu interface -------------------------------------------------------------------
    def set_option(self, option: RendererCpuOption) -> None:
        option.CopyTo(self._option)  # rasterizer.cc:73-75

    def set_mesh(self, mesh: Mesh) -> None:
        self._mesh_initialized = False  # rasterizer.cc:78
        self._mesh = mesh
        if len(mesh.face_normals) == 0:
            LOGW("face normal is empty. culling and shading may not work")
        if len(mesh.normals) == 0:
            LOGW("vertex normal is empty. shading may not work")

    def PrepareMesh(self) -> bool:
        if self._mesh is None:
            LOGE("mesh has not been set")
            return False
        d, keep = mesh_desc(self._mesh)
        rc = self._lib.ugu_b200_rasterizer_upload_mesh(self._h, C.byref(d))
        if rc != N.OK:
            LOGE(N.last_error(self._h))
            return False
        self._mesh_initialized = True
        return True

    def set_camera(self, camera: PinholeCamera) -> None:
        self._camera = camera

    def _native_option(self) -> N.Option:
        o = self._option
        return N.Option(int(o.diffuse_color), int(o.interp), int(o.shading_normal),
                        int(o.diffuse_shading), 1 if o.backface_culling else 0,
                        float(o.depth_scale), float(o.oren_nayar_sigma))

    def Render(self, color: Optional[Image3b] = None, depth: Optional[Image1f] = None,
               normal: Optional[Image3f] = None, mask: Optional[Image1b] = None,
               face_id: Optional[Image1i] = None) -> bool:
        if self._camera is None:
            LOGE("camera has not been set")
            return False
        if not self._mesh_initialized:
            LOGE("mesh has not been initialized")
            return False
        if not isinstance(self._camera, PinholeCamera):
            LOGE("only PinholeCamera is supported by the CUDA rasterizer")
            return False
        opt = self._native_option()
        self._lib.ugu_b200_rasterizer_set_option(self._h, C.byref(opt))
        cam = self._camera.to_native()
        imgs = (color, depth, normal, mask, face_id)
        if all(i is None for i in imgs):
            LOGE("all arguments are nullptr. nothing to do")
            return False
        for im in imgs:  # ugu::Init: (re)allocate only when the size differs (image.h:970-985)
            if im is not None:
                im._init(cam.width, cam.height)
        out = N.Outputs(*[None if im is None else im.data.ctypes.data for im in imgs])
        rc = self._render_call(cam, out)
        if rc != N.OK:
            LOGE(N.last_error(self._h))
            return False
        hz = int(self._lib.ugu_b200_rasterizer_last_hazards(self._h))
        if hz:
            LOGW(f"{hz} fragment(s) with depth < FLT_MIN or NaN: the reference's result is order dependent there "
                 "(rasterizer.cc:229) and this frame may differ from it at those pixels")
        return True

    def _render_call(self, cam: "N.Camera", out: "N.Outputs") -> int:
        return self._lib.ugu_b200_rasterizer_render(self._h, C.byref(cam), C.byref(out))

    def RenderColor(self, color: Image3b) -> bool:
        return self.Render(color, None, None, None, None)

    def RenderDepth(self, depth: Image1f) -> bool:
        return self.Render(None, depth, None, None, None)

    def RenderNormal(self, normal: Image3f) -> bool:
        return self.Render(None, None, normal, None, None)

    def RenderMask(self, mask: Image1b) -> bool:
        return self.Render(None, None, None, mask, None)

    def RenderFaceId(self, face_id: Image1i) -> bool:
        return self.Render(None, None, None, None, face_id)

    def RenderW(self, color: Optional[Image3b], depth: Optional[Image1w], normal: Optional[Image3f],
                mask: Optional[Image1b], face_id: Optional[Image1i]) -> bool:
        if depth is None:  # rasterizer.cc:331-334
            LOGE("depth is nullptr")
            return False
        f_depth = getattr(self, "_w_depth", None)  # one scratch image per renderer, not one per call
        if f_depth is None:
            f_depth = self._w_depth = Image1f()
        ok = self.Render(color, f_depth, normal, mask, face_id)
        if ok:
            depth._init(f_depth.cols, f_depth.rows)
            self._lib.ugu_b200_depth_to_u16(f_depth.data.ctypes.data, depth.data.ctypes.data,
                                            f_depth.data.size)
        return ok

    def RenderDepthW(self, depth: Image1w) -> bool:
        return self.RenderW(None, depth, None, None, None)

    # -- the step after Render in examples/ex02_renderer.cc:156-158 --------------------------------
    def Depth2PointCloud(self, depth: Image1f, camera: PinholeCamera, point_cloud: Image3f,
                         gl_coord: bool = False) -> bool:
        """ugu::Depth2PointCloud(const Image1f&, const Camera&, Image3f*, bool gl_coord)
        (src/util/rgbd_util.cc:472-513): organized camera-space point cloud of a depth image.  Like
        the reference, a point_cloud of another size is re-initialised to 0 and pixels with depth <
        FLT_MIN are left as they were."""
        if depth.cols != camera.width() or depth.rows != camera.height():
            LOGE(f"Depth2PointCloud depth size ({depth.cols}, {depth.rows}) and camera size "
                 f"({camera.width()}, {camera.height()}) are different")
            return False
        fresh = point_cloud.cols != depth.cols or point_cloud.rows != depth.rows
        point_cloud._init(depth.cols, depth.rows)
        cam = camera.to_native()
        rc = self._lib.ugu_b200_depth_to_point_cloud(self._h, C.byref(cam), depth.data.ctypes.data,
                                                     point_cloud.data.ctypes.data,
                                                     1 if gl_coord else 0, 1 if fresh else 0)
        if rc != N.OK:
            LOGE(N.last_error(self._h))
            return False
        return True

    def depth_to_point_cloud_device(self, camera: PinholeCamera, depth: int, point_cloud: int,
                                    gl_coord: bool = False, zero_invalid: bool = True,
                                    stream: int = 0) -> None:
        """Device-resident form: `depth` / `point_cloud` are raw 16-byte aligned device pointers."""
        cam = camera.to_native() if isinstance(camera, PinholeCamera) else camera
        rc = self._lib.ugu_b200_depth_to_point_cloud_device(
            self._h, C.byref(cam), C.c_void_p(depth), C.c_void_p(point_cloud), 1 if gl_coord else 0,
            1 if zero_invalid else 0, C.c_void_p(stream))
        if rc != N.OK:
            raise RuntimeError(N.last_error(self._h))

    # -- UV-space attribute rasterization (include/ugu/util/raster_util.h:139-268) -----------------
    def _uv_raster(self, attrs, attr_indices, uvs, uv_indices, texture, width, height, mask) -> bool:
        a = np.ascontiguousarray(attrs, np.float32).reshape(-1, 3)
        ui = np.ascontiguousarray(uv_indices, np.int32).reshape(-1, 3)
        u = np.ascontiguousarray(uvs, np.float32).reshape(-1, 2)
        per_face = attr_indices is None
        ai = None if per_face else np.ascontiguousarray(attr_indices, np.int32).reshape(-1, 3)
        # raster_util.h:148-151 / :238-240 — return false, before the texture is touched
        if per_face:
            if len(a) == 0 or len(a) != len(ui):
                return False
        elif len(ai) == 0 or len(ai) != len(ui):
            return False
        keep = not (width > 0 and height > 0)
        if keep:
            if texture.empty():
                return False  # raster_util.h:153-155
            width, height = texture.cols, texture.rows
        else:
            texture._init(width, height)
        if mask is not None and (mask.cols != width or mask.rows != height):
            LOGE("mask size differs from the texture size")  # out-of-bounds writes in the reference
            return False
        d = N.UvRasterDesc(len(a), len(ui), len(u), 1 if per_face else 0, a.ctypes.data,
                           None if ai is None else ai.ctypes.data, u.ctypes.data, ui.ctypes.data,
                           width, height, 1 if texture.dtype == np.float32 else 0, 1 if keep else 0,
                           1, 0)  # the reference never clears the mask: in/out
        rc = self._lib.ugu_b200_uv_raster(self._h, C.byref(d), texture.data.ctypes.data,
                                          None if mask is None else mask.data.ctypes.data)
        if rc != N.OK:
            LOGE(N.last_error(self._h))
            return False
        return True

    def RasterizeVertexAttributeToTexture(self, vertex_attrs, vertex_attr_indices, uvs, uv_indices,
                                          texture, width: int = 512, height: int = 512,
                                          mask: Optional["Image1b"] = None) -> bool:
        """ugu::RasterizeVertexAttributeToTexture<T> (raster_util.h:139-226), T = Vec3b (Image3b) or
        Vec3f (Image3f), in the reference's sequential order (num_threads = 1): the highest covering
        face index owns a texel.  width/height <= 0 rasterises over the existing texture."""
        return self._uv_raster(vertex_attrs, vertex_attr_indices, uvs, uv_indices, texture, width,
                               height, mask)

    def RasterizeFaceAttributeToTexture(self, face_attrs, uvs, uv_indices, texture, width: int = 512,
                                        height: int = 512, mask: Optional["Image1b"] = None) -> bool:
        """ugu::RasterizeFaceAttributeToTexture<T> (raster_util.h:228-268)."""
        return self._uv_raster(face_attrs, None, uvs, uv_indices, texture, width, height, mask)

    def GenerateUvMask(self, uvs, uv_indices, mask: "Image1b", color: int = 255, width: int = 512,
                       height: int = 512) -> bool:
        """ugu::GenerateUvMask(uvs, uv_indices, Image1b&, uint8_t color, w, h)
        (src/util/raster_util.cc:9-27)."""
        ui = np.ascontiguousarray(uv_indices, np.int32).reshape(-1, 3)
        u = np.ascontiguousarray(uvs, np.float32).reshape(-1, 2)
        if width <= 0 or height <= 0:
            return False
        mask._init(width, height)
        if len(ui) == 0:  # the reference returns false but still hands back a zero mask
            mask.data[...] = 0
            return False
        rc = self._lib.ugu_b200_uv_mask(self._h, len(ui), len(u), u.ctypes.data, ui.ctypes.data, width,
                                        height, int(color) & 0xFF, mask.data.ctypes.data)
        if rc != N.OK:
            LOGE(N.last_error(self._h))
            return False
        return True

    def uv_raster_device(self, desc: "N.UvRasterDesc", texture: int, mask: int = 0, stream: int = 0) -> None:
        """Device-resident form: every pointer in `desc`, `texture` and `mask` are raw 16-byte
        aligned device pointers; asynchronous on `stream`."""
        rc = self._lib.ugu_b200_uv_raster_device(self._h, C.byref(desc), C.c_void_p(texture),
                                                 C.c_void_p(mask) if mask else None, C.c_void_p(stream))
        if rc != N.OK:
            raise RuntimeError(N.last_error(self._h))

    # -- face-visibility consumer of face_id maps (no reference counterpart) --------
    def face_visibility(self, face_id_maps: np.ndarray, num_faces: int, first_view_id: int = 0) -> dict:
        """Per-face statistics over a stack of face_id maps (n_views, H, W) int32 in host memory:
        pixel_count, view_count, best_view (-1 = never visible), best_count — see
        include/ugu_b200_raster.h (ugu_b200_face_visibility)."""
        maps = np.ascontiguousarray(face_id_maps, np.int32)
        if maps.ndim == 2:
            maps = maps[None]
        n = maps.shape[0]
        P = int(np.prod(maps.shape[1:])) if n else 1
        out = dict(pixel_count=np.zeros(num_faces, np.uint32), view_count=np.zeros(num_faces, np.uint32),
                   best_view=np.full(num_faces, -1, np.int32), best_count=np.zeros(num_faces, np.uint32))
        rc = self._lib.ugu_b200_face_visibility(
            self._h, maps.ctypes.data if n else None, n, P, first_view_id, num_faces,
            out["pixel_count"].ctypes.data, out["view_count"].ctypes.data, out["best_view"].ctypes.data,
            out["best_count"].ctypes.data)
        if rc != N.OK:
            raise RuntimeError(N.last_error(self._h))
        return out

    def face_visibility_device(self, face_id: int, num_views: int, pixels_per_view: int, num_faces: int,
                               pixel_count: int, view_count: int, best: int, first_view_id: int = 0,
                               stream: int = 0) -> None:
        """Device-resident form: raw device pointers; accumulates into caller-zeroed arrays of
        num_faces u32 / u32 / u64 elements, asynchronously on `stream`."""
        rc = self._lib.ugu_b200_face_visibility_device(
            self._h, C.c_void_p(face_id), num_views, pixels_per_view, first_view_id, num_faces,
            C.c_void_p(pixel_count), C.c_void_p(view_count), C.c_void_p(best), C.c_void_p(stream))
        if rc != N.OK:
            raise RuntimeError(N.last_error(self._h))

    # -- ugu::Depth2Mesh (include/ugu/util/rgbd_util.h:27-37, src/util/rgbd_util.cc:94-282) -------
    def Depth2Mesh(self, depth: "Image1f", camera: "PinholeCamera", mesh: "Mesh", max_connect_z_diff: float,
                   x_step: int = 1, y_step: int = 1, gl_coord: bool = False,
                   point_cloud: Optional["Image3f"] = None, normal: Optional["Image3f"] = None,
                   color: Optional["Image3b"] = None, with_vertex_color: bool = False,
                   color_camera: Optional["PinholeCamera"] = None, depth2color=None) -> bool:
        """The overloads of ugu::Depth2Mesh: `color` None = depth only, else the textured one (uv,
        material = the colour image, vertex colours when asked); with `color_camera` and `depth2color`
        (4x4, the reference's Eigen::Affine3f) the two-camera overload (rgbd_util.h:39-48): uv and colours
        from the projection of depth2color * p into the colour camera.  Fills `mesh` like the reference:
        vertices, vertex_indices, normals + normal_indices + face_normals (Mesh::CalcNormal), uv +
        uv_indices + materials + material_ids, vertex_colors; `point_cloud` / `normal` are the
        organized images.  false + log on the reference's error conditions."""
        if depth.cols != camera.width() or depth.rows != camera.height():  # rgbd_util.cc:113-119
            LOGE(f"Depth2Mesh depth size ({depth.cols}, {depth.rows}) and camera size "
                 f"({camera.width()}, {camera.height()}) are different")
            return False
        cam = camera.to_native()
        cnt = N.Depth2MeshCounts()
        col = None if color is None else np.ascontiguousarray(color.data, np.uint8)
        if color_camera is not None:
            if col is None or depth2color is None:
                LOGE("Depth2Mesh with a colour camera needs the colour image and depth2color")
                return False
            rc = self._lib.ugu_b200_depth_to_mesh_two_camera(
                self._h, C.byref(cam), C.byref(color_camera_native(color_camera, depth2color)),
                np.ascontiguousarray(depth.data, np.float32).ctypes.data, col.ctypes.data, color.cols, color.rows,
                1 if with_vertex_color else 0, float(max_connect_z_diff), int(x_step), int(y_step),
                1 if gl_coord else 0, C.byref(cnt))
        else:
            rc = self._lib.ugu_b200_depth_to_mesh(
                self._h, C.byref(cam), np.ascontiguousarray(depth.data, np.float32).ctypes.data,
                None if col is None else col.ctypes.data, 0 if col is None else color.cols,
                0 if col is None else color.rows, 1 if with_vertex_color else 0, float(max_connect_z_diff),
                int(x_step), int(y_step), 1 if gl_coord else 0, C.byref(cnt))
        if rc != N.OK:
            LOGE(N.last_error(self._h))
            return False
        nv, nf = cnt.num_vertices, cnt.num_faces
        v = np.zeros((nv, 3), np.float32)
        f = np.zeros((nf, 3), np.int32)
        n = np.zeros((nv if cnt.has_normals else 0, 3), np.float32)
        fn = np.zeros((nf, 3), np.float32)
        uv = np.zeros((nv if cnt.has_uv else 0, 2), np.float32)
        vc = np.zeros((nv if cnt.has_colors else 0, 3), np.float32)
        if point_cloud is not None:
            point_cloud._init(depth.cols, depth.rows)
        if normal is not None:
            normal._init(depth.cols, depth.rows)
        rc = self._lib.ugu_b200_depth_to_mesh_fetch(
            self._h, None, v.ctypes.data, f.ctypes.data, n.ctypes.data if n.size else None,
            fn.ctypes.data if nf else None, uv.ctypes.data if uv.size else None,
            vc.ctypes.data if vc.size else None,
            None if point_cloud is None else point_cloud.data.ctypes.data,
            None if normal is None else normal.data.ctypes.data)
        if rc != N.OK:
            LOGE(N.last_error(self._h))
            return False
        mesh.vertices, mesh.vertex_indices = v, f
        mesh.normals, mesh.face_normals = n, fn
        mesh.normal_indices = f.copy() if cnt.has_normals else np.zeros((0, 3), np.int32)
        mesh.vertex_colors = vc
        mesh.uv = uv
        mesh.uv_indices = f.copy() if cnt.has_uv else np.zeros((0, 3), np.int32)
        mesh.material_ids = np.zeros(nf if cnt.has_uv else 0, np.int32)
        mesh.materials = [Material(diffuse_tex=color.data.copy())] if color is not None else []
        return True

    def depth_to_mesh_device(self, camera, depth: int, max_connect_z_diff: float, x_step: int = 1,
                             y_step: int = 1, gl_coord: bool = False, color: int = 0, color_width: int = 0,
                             color_height: int = 0, with_vertex_color: bool = False, stream: int = 0) -> None:
        """Device-resident form: `depth` / `color` are raw device pointers (e.g. outputs of
        render_device); asynchronous on `stream`; the result stays on the device (depth_to_mesh_counts)."""
        cam = camera.to_native() if isinstance(camera, PinholeCamera) else camera
        rc = self._lib.ugu_b200_depth_to_mesh_device(
            self._h, C.byref(cam), C.c_void_p(depth), C.c_void_p(color) if color else None, color_width,
            color_height, 1 if with_vertex_color else 0, float(max_connect_z_diff), int(x_step), int(y_step),
            1 if gl_coord else 0, C.c_void_p(stream))
        if rc != N.OK:
            raise RuntimeError(N.last_error(self._h))

    def depth_to_mesh_counts(self) -> dict:
        """Waits for the last depth_to_mesh call; {num_vertices, num_faces, has_normals, ...}."""
        cnt = N.Depth2MeshCounts()
        rc = self._lib.ugu_b200_depth_to_mesh_fetch(self._h, C.byref(cnt), *([None] * 8))
        if rc != N.OK:
            raise RuntimeError(N.last_error(self._h))
        return {k: getattr(cnt, k) for k, _ in N.Depth2MeshCounts._fields_ if k != "reserved"}

    def RenderGBuffer(self, gbuf: GBuffer, z_near: float = 0.01, z_far: float = 1000.0,
                      flat_normal: bool = False, geo_id: int = 0, fields: Optional[Sequence[str]] = None) -> bool:
        """The G-buffer of RendererGl::ReadGbuf (src/renderer/gl/renderer.cc:501-650; near / far as
        RendererGl::SetNearFar, :751-754) from this rasterizer's winners — definition in
        include/ugu_b200_raster.h (ugu_b200_rasterizer_render_gbuffer).  `fields` restricts the images
        written (default: all of GBuffer.TYPES); False + log on the Render precondition failures."""
        if self._camera is None:
            LOGE("camera has not been set")
            return False
        if not self._mesh_initialized:
            LOGE("mesh has not been initialized")
            return False
        opt = self._native_option()
        self._lib.ugu_b200_rasterizer_set_option(self._h, C.byref(opt))
        cam = self._camera.to_native()
        names = list(GBuffer.TYPES) if fields is None else list(fields)
        out = N.GBuffer()
        for n in names:
            im = getattr(gbuf, n)
            im._init(cam.width, cam.height)
            setattr(out, n, im.data.ctypes.data)
        go = N.GBufferOption(float(z_near), float(z_far), 1 if flat_normal else 0, int(geo_id))
        rc = self._lib.ugu_b200_rasterizer_render_gbuffer(self._h, C.byref(cam), C.byref(go), C.byref(out))
        if rc != N.OK:
            LOGE(N.last_error(self._h))
            return False
        return True

    def render_gbuffer_device(self, camera: PinholeCamera, z_near: float, z_far: float, flat_normal: bool = False,
                              geo_id_value: int = 0, stream: int = 0, **fields: int) -> None:
        """Device form: `fields` are raw 16-byte aligned DEVICE pointers by GBuffer field name (`geo_id`
        is the geo_id IMAGE; the id written into it is `geo_id_value`)."""
        unknown = set(fields) - set(GBuffer.TYPES)
        if unknown:
            raise TypeError(f"unknown G-buffer field(s): {sorted(unknown)}")
        opt = self._native_option()
        self._lib.ugu_b200_rasterizer_set_option(self._h, C.byref(opt))
        cam = camera.to_native() if isinstance(camera, PinholeCamera) else camera
        out = N.GBuffer(**fields)
        go = N.GBufferOption(float(z_near), float(z_far), 1 if flat_normal else 0, int(geo_id_value))
        rc = self._lib.ugu_b200_rasterizer_render_gbuffer_device(self._h, C.byref(cam), C.byref(go), C.byref(out),
                                                                 C.c_void_p(stream))
        if rc != N.OK:
            raise RuntimeError(N.last_error(self._h))

    # -- device-resident / batch extensions (no reference counterpart) --------
    def render_device(self, camera: PinholeCamera, color=None, depth=None, normal=None, mask=None,
                      face_id=None, stream: int = 0) -> None:
        """Render into DEVICE buffers given as raw device pointers (ints), asynchronously on the
        cudaStream_t `stream` (0 = the rasterizer's own stream)."""
        opt = self._native_option()
        self._lib.ugu_b200_rasterizer_set_option(self._h, C.byref(opt))
        cam = camera.to_native() if isinstance(camera, PinholeCamera) else camera
        out = N.Outputs(color, depth, normal, mask, face_id)
        rc = self._lib.ugu_b200_rasterizer_render_device(self._h, C.byref(cam), C.byref(out),
                                                         C.c_void_p(stream))
        if rc != N.OK:
            raise RuntimeError(N.last_error(self._h))

    def render_batch(self, cameras: Sequence, color=None, depth=None, normal=None, mask=None,
                     face_id=None, on_device: bool = False, stream: int = 0) -> None:
        """Render len(cameras) views; view i lands at element offset i*W*H of each output.  Outputs
        are raw pointers (ints): device pointers when on_device, else host pointers."""
        opt = self._native_option()
        self._lib.ugu_b200_rasterizer_set_option(self._h, C.byref(opt))
        n = len(cameras)
        cams = (N.Camera * n)(*[c.to_native() if isinstance(c, PinholeCamera) else c for c in cameras])
        out = N.Outputs(color, depth, normal, mask, face_id)
        rc = self._lib.ugu_b200_rasterizer_render_batch(self._h, n, cams, C.byref(out),
                                                        1 if on_device else 0, C.c_void_p(stream))
        if rc != N.OK:
            raise RuntimeError(N.last_error(self._h))

    def sync(self) -> None:
        if self._lib.ugu_b200_rasterizer_sync(self._h) != N.OK:
            raise RuntimeError(N.last_error(self._h))

    def set_profiling(self, enable: bool, stages: Sequence[int] = (0, 1, 2, 3)) -> None:
        mask = 0
        for s in stages:
            mask |= 1 << s
        self._lib.ugu_b200_rasterizer_set_profiling(self._h, 1 if enable else 0, mask)

    def stats(self) -> dict:
        st = N.Stats()
        if self._lib.ugu_b200_rasterizer_get_stats(self._h, C.byref(st)) != N.OK:
            raise RuntimeError(N.last_error(self._h))
        d = {k: int(getattr(st, k)) for k in ("timed_views", "views", "kernel_launches", "h2d_bytes",
                                              "d2h_bytes", "large_triangles", "tile_jobs",
                                              "hazard_fragments")}
        d["stage_ms"] = {N.STAGE_NAMES[i]: float(st.stage_ms[i]) for i in range(N.NUM_STAGES)}
        return d

    def reset_stats(self) -> None:
        self._lib.ugu_b200_rasterizer_reset_stats(self._h)


class Raytracer(Rasterizer):
    """The CUDA-backed drop-in for ugu::Raytracer (include/ugu/renderer/cpu/raytracer.h,
    src/renderer/cpu/raytracer.cc:117-225): same interface and call order as the Rasterizer, the outputs
    are those of the reference's nanort ray caster (nearest hit of the primary ray, back-face test on
    the hit, depth_scale applied, face 0 shaded) — see ugu_b200_rasterizer_raytrace in the C header."""

    def PrepareMesh(self) -> bool:
        m = self._mesh
        if m is not None and (m.vertices is None or len(m.vertices) == 0 or m.vertex_indices is None
                              or len(m.vertex_indices) == 0):
            LOGE("mesh is empty")  # Raytracer::Impl::PrepareMesh, raytracer.cc:96-99
            return False
        return super().PrepareMesh()

    def _org(self, camera: PinholeCamera):
        o = np.asarray(camera.c2w(), np.float64)[:3, 3].astype(np.float32)  # c2w_t_f_, camera.h:431-436
        return (C.c_float * 3)(float(o[0]), float(o[1]), float(o[2]))

    def _render_call(self, cam: "N.Camera", out: "N.Outputs") -> int:
        return self._lib.ugu_b200_rasterizer_raytrace(self._h, C.byref(cam), self._org(self._camera), C.byref(out))

    def render_device(self, camera: PinholeCamera, color=None, depth=None, normal=None, mask=None,
                      face_id=None, stream: int = 0) -> None:
        opt = self._native_option()
        self._lib.ugu_b200_rasterizer_set_option(self._h, C.byref(opt))
        out = N.Outputs(color, depth, normal, mask, face_id)
        rc = self._lib.ugu_b200_rasterizer_raytrace_device(self._h, C.byref(camera.to_native()), self._org(camera),
                                                           C.byref(out), C.c_void_p(stream))
        if rc != N.OK:
            raise RuntimeError(N.last_error(self._h))

    def render_batch(self, cameras: Sequence, color=None, depth=None, normal=None, mask=None,
                     face_id=None, on_device: bool = False, stream: int = 0) -> None:
        """len(cameras) views back to back on one stream; view i lands at element offset i*W*H of each
        output (DEVICE pointers; the multi-view driver's path, ugu_b200.multiview.MultiViewRenderer — the
        reference's own visibility pass casts BVH rays, include/ugu/texturing/visibility_tester.h)."""
        if not cameras:
            return
        P = cameras[0].width() * cameras[0].height()
        if not on_device:  # host destinations: one synchronous ray-cast call per view into the view's slab
            opt = self._native_option()
            self._lib.ugu_b200_rasterizer_set_option(self._h, C.byref(opt))
            for i, cam in enumerate(cameras):
                out = N.Outputs(None if color is None else color + i * P * 3,
                                None if depth is None else depth + i * P * 4,
                                None if normal is None else normal + i * P * 12,
                                None if mask is None else mask + i * P,
                                None if face_id is None else face_id + i * P * 4)
                rc = self._lib.ugu_b200_rasterizer_raytrace(self._h, C.byref(cam.to_native()), self._org(cam),
                                                            C.byref(out))
                if rc != N.OK:
                    raise RuntimeError(N.last_error(self._h))
            return
        for i, cam in enumerate(cameras):
            self.render_device(cam,
                               color=None if color is None else color + i * P * 3,
                               depth=None if depth is None else depth + i * P * 4,
                               normal=None if normal is None else normal + i * P * 12,
                               mask=None if mask is None else mask + i * P,
                               face_id=None if face_id is None else face_id + i * P * 4, stream=stream)
