"""Host-side mirror of ugu::VisibilityTester for Python hosts (include/ugu/texturing/visibility_tester.h:216-265):
same class / method names and call order (set_data -> PrepareData -> Test(keyframes, info, facewise)), the work
done by the CUDA kernels behind the C-ABI (ugu_b200_visibility_*, include/ugu_b200_raster.h).  There is no CPU
fallback.  What differs from the reference — the BVH ray is replaced by the rasterizer's own depth map — is stated
in the C header."""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from typing import List, Optional, Sequence

import numpy as np

from . import _native as N
from .renderer import ColorInterpolation, LOGE, PinholeCamera, Rasterizer

VERTEX_REC_DTYPE = np.dtype([("color", np.float32, 3), ("intensity", np.float32), ("projected_pos", np.float32, 2),
                             ("viewing_angle", np.float32), ("distance", np.float32)])
FACE_REC_DTYPE = np.dtype([("area", np.float32), ("viewing_angle", np.float32), ("distance", np.float32),
                           ("projected_tri", np.float32, 6)])

_HALF_PI = float(np.float32(90.0) * np.float32(np.pi / 180.0))  # ugu::radians(90.0f)
_FLT_MAX = float(np.finfo(np.float32).max)


@dataclass
class VisibilityTesterOption:  # visibility_tester.h:39-58
    backface_culling_angle_rad_th: float = _HALF_PI
    use_mask: bool = True
    use_depth: bool = True
    interp: ColorInterpolation = ColorInterpolation.kBilinear
    calc_stat_vertex_info: bool = True
    collect_face_info: bool = True
    calc_stat_face_info: bool = True
    max_viewing_angle: float = _HALF_PI
    max_distance_from_camera_to_vertex: float = _FLT_MAX
    max_depth_difference: float = _FLT_MAX
    eps: float = 0.1

    def CopyTo(self, dst: "VisibilityTesterOption") -> None:
        # the reference copies only these seven (visibility_tester.cc:43-53, "todo")
        dst.backface_culling_angle_rad_th = self.backface_culling_angle_rad_th
        dst.use_mask = self.use_mask
        dst.use_depth = self.use_depth
        dst.interp = self.interp
        dst.calc_stat_vertex_info = self.calc_stat_vertex_info
        dst.collect_face_info = self.collect_face_info
        dst.calc_stat_face_info = self.calc_stat_face_info


@dataclass
class Keyframe:  # visibility_tester.h:201-214
    id: int = -1
    color: Optional[np.ndarray] = None   # (H, W, 3) u8
    depth: Optional[np.ndarray] = None   # (H, W) f32
    mask: Optional[np.ndarray] = None    # (H, W) u8
    camera: Optional[PinholeCamera] = None
    zbuf: Optional[np.ndarray] = None    # (H, W) f32: the mesh's own depth map (None = rendered by the library)

    @staticmethod
    def Create() -> "Keyframe":
        return Keyframe()


@dataclass
class VisibilityInfo:
    """ugu::VisibilityInfo (visibility_tester.h:183-199) as dense arrays: row k of the per-keyframe records
    belongs to keyframe_ids[k]; `stats` holds what CalcStatFaceInfo / CalcStatVertexInfo fold (ids, not orders)."""
    keyframe_ids: List[int] = field(default_factory=list)
    vertex_visible: Optional[np.ndarray] = None  # (K, V) u8
    vertex: Optional[np.ndarray] = None          # (K, V) VERTEX_REC_DTYPE
    face_visible: Optional[np.ndarray] = None    # (K, F) u8
    face: Optional[np.ndarray] = None            # (K, F) FACE_REC_DTYPE
    stats: dict = field(default_factory=dict)
    has_vertex_stat: bool = False
    has_face_stat: bool = False


def native_keyframe(kf: Keyframe, keep: list, images_on_device: bool = False) -> "N.Keyframe":
    cam = kf.camera
    out = N.Keyframe()
    c2w = (C.c_double * 16)(*np.asarray(cam.c2w(), np.float64).reshape(-1))
    f, pp = cam.focal_length(), cam.principal_point()
    rc = N.load().ugu_b200_keyframe_camera_from_c2w(c2w, cam.width(), cam.height(), float(f[0]), float(f[1]),
                                                    float(pp[0]), float(pp[1]), C.byref(out))
    if rc != N.OK:
        raise ValueError("c2w is not invertible")
    out.id = int(kf.id)
    if images_on_device:  # ints = device pointers
        out.color, out.depth, out.mask, out.zbuf = kf.color, kf.depth, kf.mask, kf.zbuf
        return out
    for name, dt in (("color", np.uint8), ("depth", np.float32), ("mask", np.uint8), ("zbuf", np.float32)):
        a = getattr(kf, name)
        if a is None or np.size(a) == 0:
            setattr(out, name, None)
            continue
        a = np.ascontiguousarray(a, dt)
        keep.append(a)
        setattr(out, name, a.ctypes.data)
    return out


class VisibilityTester:
    """The CUDA-backed drop-in for ugu::VisibilityTester.  `rasterizer` is the ugu_b200.Rasterizer whose
    uploaded mesh supplies the self depth maps (keyframes without a zbuf) — the same geometry as set_data's."""

    def __init__(self, option: Optional[VisibilityTesterOption] = None, rasterizer: Optional[Rasterizer] = None,
                 self_occlusion: bool = True):
        self._lib = N.load()
        self._r = rasterizer if rasterizer is not None else Rasterizer()
        self._option = VisibilityTesterOption()
        self._data = None
        self._initialized = False
        self._keyframe: Optional[Keyframe] = None
        self.self_occlusion = self_occlusion
        if option is not None:
            self.set_option(option)

    def set_option(self, option: VisibilityTesterOption) -> None:
        option.CopyTo(self._option)

    def set_data(self, vertices, normals, indices) -> None:
        self._initialized = False
        self._data = (np.ascontiguousarray(vertices, np.float32).reshape(-1, 3),
                      np.ascontiguousarray(normals, np.float32).reshape(-1, 3),
                      np.ascontiguousarray(indices, np.int32).reshape(-1, 3))

    def PrepareData(self) -> bool:
        if self._data is None or any(len(a) == 0 for a in self._data) or len(self._data[0]) != len(self._data[1]):
            LOGE("mesh is empty")
            return False
        v, n, f = self._data
        rc = self._lib.ugu_b200_visibility_set_data(self._r._h, len(v), v.ctypes.data, n.ctypes.data, len(f),
                                                    f.ctypes.data)
        if rc != N.OK:
            LOGE(N.last_error(self._r._h))
            return False
        self._initialized = True
        return True

    def set_keyframe(self, keyframe: Keyframe) -> None:
        self._keyframe = keyframe

    def native_option(self, facewise: bool) -> "N.VisOption":
        o = self._option
        return N.VisOption(float(o.backface_culling_angle_rad_th), int(bool(o.use_mask)), int(bool(o.use_depth)),
                           int(o.interp), float(o.max_viewing_angle), float(o.max_distance_from_camera_to_vertex),
                           float(o.max_depth_difference), float(o.eps), int(bool(facewise)),
                           int(bool(self.self_occlusion)))

    def Test(self, keyframes_or_info, info: Optional[VisibilityInfo] = None, facewise: bool = True,
             records: bool = True) -> bool:
        """Test(info, facewise) on the keyframe of set_keyframe, or Test(keyframes, info, facewise)
        (visibility_tester.cc:670-733).  Fills `info` (dense per-keyframe records when `records`, and the
        folded statistics)."""
        if isinstance(keyframes_or_info, VisibilityInfo):
            info, keyframes = keyframes_or_info, [self._keyframe]
        else:
            keyframes = list(keyframes_or_info)
        if info is None:
            LOGE("info is nullptr")
            return False
        if not keyframes or keyframes[0] is None:
            LOGE("keyframe has not been set")
            return False
        if not self._initialized:
            LOGE("mesh has not been initialized")
            return False
        V, F, K = len(self._data[0]), len(self._data[2]), len(keyframes)
        keep: list = []
        for kf in keyframes:
            if kf.color is None or np.size(kf.color) == 0:
                LOGE("keyframe color is empty")
                return False
        kfs = (N.Keyframe * K)(*[native_keyframe(kf, keep) for kf in keyframes])
        opt = self.native_option(facewise)
        rec = N.VisRecords()
        if records:
            info.vertex_visible = np.zeros((K, V), np.uint8)
            info.vertex = np.zeros((K, V), VERTEX_REC_DTYPE)
            info.face_visible = np.zeros((K, F), np.uint8)
            info.face = np.zeros((K, F), FACE_REC_DTYPE)
            rec = N.VisRecords(info.vertex_visible.ctypes.data, info.vertex.ctypes.data,
                               info.face_visible.ctypes.data, info.face.ctypes.data)
        first = len(info.keyframe_ids)
        if first == 0:
            self._lib.ugu_b200_visibility_reset(self._r._h)
        rc = self._lib.ugu_b200_visibility_test(self._r._h, K, kfs, first, C.byref(opt), C.byref(rec))
        if rc != N.OK:
            LOGE(N.last_error(self._r._h))
            return False
        info.keyframe_ids.extend(int(kf.id) for kf in keyframes)
        info.stats = self.stats(info.keyframe_ids)
        info.has_face_stat = info.has_vertex_stat = True
        return True

    def stats(self, keyframe_ids: Sequence[int]) -> dict:
        """The folded statistics with orders translated to keyframe ids (-1 = never visible)."""
        V, F = len(self._data[0]), len(self._data[2])
        shapes = dict(face_count=((F,), np.uint32), face_max_area_order=((F,), np.int32),
                      face_min_viewing_angle_order=((F,), np.int32), face_min_distance_order=((F,), np.int32),
                      face_max_area=((F,), np.float32), face_min_viewing_angle=((F,), np.float32),
                      face_min_distance=((F,), np.float32), vertex_count=((V,), np.uint32),
                      vertex_min_viewing_angle_order=((V,), np.int32), vertex_min_distance_order=((V,), np.int32),
                      vertex_mean_color=((V, 3), np.float32), vertex_mean_viewing_angle=((V,), np.float32),
                      vertex_mean_distance=((V,), np.float32), vertex_min_viewing_angle_color=((V, 3), np.float32),
                      vertex_min_distance_color=((V, 3), np.float32))
        out = {k: np.zeros(s, d) for k, (s, d) in shapes.items()}
        st = N.VisStats(*[out[n].ctypes.data for n in N.VisStats.FIELDS])
        rc = self._lib.ugu_b200_visibility_stats(self._r._h, C.byref(st))
        if rc != N.OK:
            raise RuntimeError(N.last_error(self._r._h))
        ids = np.asarray(list(keyframe_ids) + [-1], np.int32)  # order -1 -> id -1
        for name in ("face_max_area", "face_min_viewing_angle", "face_min_distance", "vertex_min_viewing_angle",
                     "vertex_min_distance"):
            out[name + "_id"] = ids[out[name + "_order"]]
        return out
