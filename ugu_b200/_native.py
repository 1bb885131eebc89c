"""ctypes binding of the C-ABI in include/ugu_b200_raster.h (the product's only native entry).

The library is built in-tree by ugu_b200.build (nvcc, sm_100a).  Loading never falls back to a CPU
implementation: if the shared object is missing or unloadable, importing raises.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path

LIB_PATH = Path(__file__).resolve().parent / "lib" / "libugu_b200_raster.so"

OK, E_INVALID, E_CUDA, E_NOMEM, E_STATE = 0, 1, 2, 3, 4
NUM_STAGES = 4
STAGE_NAMES = ("vertex", "raster", "binned", "resolve")


class Option(C.Structure):
    _fields_ = [("diffuse_color", C.c_int32), ("interp", C.c_int32), ("shading_normal", C.c_int32),
                ("diffuse_shading", C.c_int32), ("backface_culling", C.c_int32),
                ("depth_scale", C.c_float), ("oren_nayar_sigma", C.c_float)]


class MeshDesc(C.Structure):
    _fields_ = [("num_vertices", C.c_int32), ("num_normals", C.c_int32), ("num_colors", C.c_int32),
                ("num_faces", C.c_int32), ("num_face_normals", C.c_int32), ("num_uv", C.c_int32),
                ("num_materials", C.c_int32), ("reserved", C.c_int32),
                ("vertices", C.c_void_p), ("normals", C.c_void_p), ("vertex_colors", C.c_void_p),
                ("vertex_indices", C.c_void_p), ("normal_indices", C.c_void_p),
                ("face_normals", C.c_void_p), ("uv", C.c_void_p), ("uv_indices", C.c_void_p),
                ("material_ids", C.c_void_p), ("tex_data", C.POINTER(C.c_void_p)),
                ("tex_w", C.POINTER(C.c_int32)), ("tex_h", C.POINTER(C.c_int32))]


class Camera(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("fx", C.c_float), ("fy", C.c_float),
                ("cx", C.c_float), ("cy", C.c_float), ("w2c_R", C.c_float * 9),
                ("w2c_t", C.c_float * 3), ("c2w_R", C.c_float * 9)]


class Outputs(C.Structure):
    _fields_ = [("color", C.c_void_p), ("depth", C.c_void_p), ("normal", C.c_void_p),
                ("mask", C.c_void_p), ("face_id", C.c_void_p)]


class UvRasterDesc(C.Structure):
    _fields_ = [("num_attrs", C.c_int32), ("num_faces", C.c_int32), ("num_uv", C.c_int32),
                ("per_face", C.c_int32), ("attrs", C.c_void_p), ("attr_indices", C.c_void_p),
                ("uv", C.c_void_p), ("uv_indices", C.c_void_p), ("width", C.c_int32),
                ("height", C.c_int32), ("texture_type", C.c_int32), ("keep_texture", C.c_int32),
                ("keep_mask", C.c_int32), ("reserved", C.c_int32)]


class ColorCamera(C.Structure):
    _fields_ = [("depth2color_R", C.c_float * 9), ("depth2color_t", C.c_float * 3), ("fx", C.c_float),
                ("fy", C.c_float), ("cx", C.c_float), ("cy", C.c_float)]


class Depth2MeshCounts(C.Structure):
    _fields_ = [("num_vertices", C.c_int32), ("num_faces", C.c_int32), ("has_normals", C.c_int32),
                ("has_uv", C.c_int32), ("has_colors", C.c_int32), ("reserved", C.c_int32)]


class GBuffer(C.Structure):
    FIELDS = ("color", "normal_cam", "normal_wld", "pos_cam", "pos_wld", "depth_01", "stencil", "face_id",
              "uv", "bary", "geo_id")
    _fields_ = [(n, C.c_void_p) for n in FIELDS]


class GBufferOption(C.Structure):
    _fields_ = [("z_near", C.c_float), ("z_far", C.c_float), ("flat_normal", C.c_int32), ("geo_id", C.c_int32)]


class Stats(C.Structure):
    _fields_ = [("stage_ms", C.c_double * NUM_STAGES), ("timed_views", C.c_uint64),
                ("views", C.c_uint64), ("kernel_launches", C.c_uint64), ("h2d_bytes", C.c_uint64),
                ("d2h_bytes", C.c_uint64), ("large_triangles", C.c_uint64),
                ("tile_jobs", C.c_uint64), ("hazard_fragments", C.c_uint64)]


class VisOption(C.Structure):
    _fields_ = [("backface_culling_angle_rad_th", C.c_float), ("use_mask", C.c_int32), ("use_depth", C.c_int32),
                ("interp", C.c_int32), ("max_viewing_angle", C.c_float),
                ("max_distance_from_camera_to_vertex", C.c_float), ("max_depth_difference", C.c_float),
                ("eps", C.c_float), ("facewise", C.c_int32), ("self_occlusion", C.c_int32)]


class Keyframe(C.Structure):
    _fields_ = [("id", C.c_int32), ("reserved", C.c_int32), ("cam", Camera), ("w2c", C.c_float * 12),
                ("c2w_t", C.c_float * 3), ("reserved2", C.c_float), ("color", C.c_void_p),
                ("depth", C.c_void_p), ("mask", C.c_void_p), ("zbuf", C.c_void_p)]


class VisRecords(C.Structure):
    _fields_ = [("vertex_visible", C.c_void_p), ("vertex", C.c_void_p), ("face_visible", C.c_void_p),
                ("face", C.c_void_p)]


class VisStats(C.Structure):
    FIELDS = ("face_count", "face_max_area_order", "face_min_viewing_angle_order", "face_min_distance_order",
              "face_max_area", "face_min_viewing_angle", "face_min_distance", "vertex_count",
              "vertex_min_viewing_angle_order", "vertex_min_distance_order", "vertex_mean_color",
              "vertex_mean_viewing_angle", "vertex_mean_distance", "vertex_min_viewing_angle_color",
              "vertex_min_distance_color")
    _fields_ = [(n, C.c_void_p) for n in FIELDS]


class VisAccumulators(C.Structure):
    FIELDS = ("face_count", "face_max_area", "face_min_angle", "face_min_dist", "vertex_count",
              "vertex_sum_color", "vertex_sum_angle_dist", "vertex_min_angle", "vertex_min_dist",
              "vertex_min_colors")
    _fields_ = [(n, C.c_void_p) for n in FIELDS]


# every symbol include/ugu_b200_raster.h declares: name -> (restype, argtypes)
SYMBOLS = {
    "ugu_b200_abi_version": (C.c_int, []),
    "ugu_b200_device_count": (C.c_int, []),
    "ugu_b200_rasterizer_create": (C.c_int, [C.c_int, C.POINTER(C.c_void_p)]),
    "ugu_b200_rasterizer_destroy": (None, [C.c_void_p]),
    "ugu_b200_rasterizer_set_option": (C.c_int, [C.c_void_p, C.POINTER(Option)]),
    "ugu_b200_rasterizer_upload_mesh": (C.c_int, [C.c_void_p, C.POINTER(MeshDesc)]),
    "ugu_b200_rasterizer_render": (C.c_int, [C.c_void_p, C.POINTER(Camera), C.POINTER(Outputs)]),
    "ugu_b200_rasterizer_render_device": (C.c_int, [C.c_void_p, C.POINTER(Camera),
                                                    C.POINTER(Outputs), C.c_void_p]),
    "ugu_b200_rasterizer_render_batch": (C.c_int, [C.c_void_p, C.c_int32, C.POINTER(Camera),
                                                   C.POINTER(Outputs), C.c_int32, C.c_void_p]),
    "ugu_b200_rasterizer_sync": (C.c_int, [C.c_void_p]),
    "ugu_b200_rasterizer_last_hazards": (C.c_uint64, [C.c_void_p]),
    "ugu_b200_depth_to_u16": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t]),
    "ugu_b200_depth_to_point_cloud_device": (C.c_int, [C.c_void_p, C.POINTER(Camera), C.c_void_p,
                                                       C.c_void_p, C.c_int32, C.c_int32, C.c_void_p]),
    "ugu_b200_depth_to_point_cloud": (C.c_int, [C.c_void_p, C.POINTER(Camera), C.c_void_p, C.c_void_p,
                                                C.c_int32, C.c_int32]),
    "ugu_b200_uv_raster": (C.c_int, [C.c_void_p, C.POINTER(UvRasterDesc), C.c_void_p, C.c_void_p]),
    "ugu_b200_uv_raster_device": (C.c_int, [C.c_void_p, C.POINTER(UvRasterDesc), C.c_void_p, C.c_void_p,
                                            C.c_void_p]),
    "ugu_b200_uv_mask": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_int32,
                                   C.c_int32, C.c_int32, C.c_void_p]),
    "ugu_b200_face_visibility_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int64, C.c_int32,
                                                  C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "ugu_b200_face_visibility": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int64, C.c_int32, C.c_int32,
                                           C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "ugu_b200_face_visibility_static": (C.c_int, [C.c_int32, C.c_void_p, C.c_int32, C.c_int64, C.c_int32,
                                                  C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                                  C.c_char_p, C.c_size_t]),
    "ugu_b200_visibility_set_data": (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p]),
    "ugu_b200_visibility_reset": (C.c_int, [C.c_void_p]),
    "ugu_b200_visibility_test": (C.c_int, [C.c_void_p, C.c_int32, C.POINTER(Keyframe), C.c_int32,
                                           C.POINTER(VisOption), C.POINTER(VisRecords)]),
    "ugu_b200_visibility_test_device": (C.c_int, [C.c_void_p, C.c_int32, C.POINTER(Keyframe), C.c_int32,
                                                  C.POINTER(VisOption), C.POINTER(VisRecords), C.c_void_p]),
    "ugu_b200_visibility_stats": (C.c_int, [C.c_void_p, C.POINTER(VisStats)]),
    "ugu_b200_visibility_accumulators": (C.c_int, [C.c_void_p, C.POINTER(VisAccumulators)]),
    "ugu_b200_keyframe_camera_from_c2w": (C.c_int, [C.POINTER(C.c_double), C.c_int32, C.c_int32, C.c_float,
                                                    C.c_float, C.c_float, C.c_float, C.POINTER(Keyframe)]),
    "ugu_b200_depth_to_mesh": (C.c_int, [C.c_void_p, C.POINTER(Camera), C.c_void_p, C.c_void_p, C.c_int32,
                                         C.c_int32, C.c_int32, C.c_float, C.c_int32, C.c_int32, C.c_int32,
                                         C.POINTER(Depth2MeshCounts)]),
    "ugu_b200_depth_to_mesh_device": (C.c_int, [C.c_void_p, C.POINTER(Camera), C.c_void_p, C.c_void_p,
                                                C.c_int32, C.c_int32, C.c_int32, C.c_float, C.c_int32,
                                                C.c_int32, C.c_int32, C.c_void_p]),
    "ugu_b200_depth_to_mesh_two_camera": (C.c_int, [C.c_void_p, C.POINTER(Camera), C.POINTER(ColorCamera), C.c_void_p,
                                                    C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_float,
                                                    C.c_int32, C.c_int32, C.c_int32, C.POINTER(Depth2MeshCounts)]),
    "ugu_b200_depth_to_mesh_two_camera_device": (C.c_int, [C.c_void_p, C.POINTER(Camera), C.POINTER(ColorCamera),
                                                           C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32,
                                                           C.c_float, C.c_int32, C.c_int32, C.c_int32, C.c_void_p]),
    "ugu_b200_depth_to_mesh_fetch": (C.c_int, [C.c_void_p, C.POINTER(Depth2MeshCounts)] + [C.c_void_p] * 8),
    "ugu_b200_rasterizer_render_gbuffer": (C.c_int, [C.c_void_p, C.POINTER(Camera), C.POINTER(GBufferOption),
                                                     C.POINTER(GBuffer)]),
    "ugu_b200_rasterizer_render_gbuffer_device": (C.c_int, [C.c_void_p, C.POINTER(Camera),
                                                            C.POINTER(GBufferOption), C.POINTER(GBuffer),
                                                            C.c_void_p]),
    "ugu_b200_rasterizer_raytrace": (C.c_int, [C.c_void_p, C.POINTER(Camera), C.POINTER(C.c_float),
                                               C.POINTER(Outputs)]),
    "ugu_b200_rasterizer_raytrace_device": (C.c_int, [C.c_void_p, C.POINTER(Camera), C.POINTER(C.c_float),
                                                      C.POINTER(Outputs), C.c_void_p]),
    "ugu_b200_rasterizer_set_profiling": (C.c_int, [C.c_void_p, C.c_int32, C.c_uint32]),
    "ugu_b200_rasterizer_get_stats": (C.c_int, [C.c_void_p, C.POINTER(Stats)]),
    "ugu_b200_rasterizer_reset_stats": (C.c_int, [C.c_void_p]),
    "ugu_b200_camera_from_c2w": (C.c_int, [C.POINTER(C.c_double), C.c_int32, C.c_int32, C.c_float,
                                           C.c_float, C.c_float, C.c_float, C.POINTER(Camera)]),
    "ugu_b200_host_alloc": (C.c_void_p, [C.c_size_t]),
    "ugu_b200_host_free": (None, [C.c_void_p]),
    "ugu_b200_last_error": (C.c_char_p, [C.c_void_p]),
    # multi-view dispatch over several GPUs from one process
    "ugu_b200_partition": (C.c_int, [C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.POINTER(C.c_int32),
                                     C.POINTER(C.c_int32)]),
    "ugu_b200_multiview_create": (C.c_int, [C.POINTER(C.c_int32), C.c_int32, C.c_int32, C.POINTER(C.c_void_p)]),
    "ugu_b200_multiview_destroy": (None, [C.c_void_p]),
    "ugu_b200_multiview_size": (C.c_int32, [C.c_void_p]),
    "ugu_b200_multiview_transport": (C.c_int32, [C.c_void_p]),
    "ugu_b200_multiview_rasterizer": (C.c_void_p, [C.c_void_p, C.c_int32]),
    "ugu_b200_multiview_last_error": (C.c_char_p, [C.c_void_p]),
    "ugu_b200_multiview_set_option": (C.c_int, [C.c_void_p, C.POINTER(Option)]),
    "ugu_b200_multiview_upload_mesh": (C.c_int, [C.c_void_p, C.POINTER(MeshDesc)]),
    "ugu_b200_multiview_render_batch": (C.c_int, [C.c_void_p, C.c_int32, C.POINTER(Camera), C.c_int32,
                                                  C.POINTER(Outputs)]),
    "ugu_b200_multiview_gather_face_ids": (C.c_int, [C.c_void_p, C.c_int32, C.POINTER(Camera), C.c_int32, C.c_int32,
                                                     C.c_int32, C.c_void_p, C.c_void_p]),
    "ugu_b200_multiview_gathered": (C.c_void_p, [C.c_void_p]),
    "ugu_b200_multiview_reduce_face_visibility": (C.c_int, [C.c_void_p, C.c_int32, C.POINTER(Camera), C.c_int32,
                                                            C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p,
                                                            C.c_void_p]),
}

_lib = None


def load() -> C.CDLL:
    """Load libugu_b200_raster.so (once).  Raises if it is absent — there is no fallback."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -m ugu_b200.build` "
            "(ugu_b200 has no CPU fallback)")
    lib = C.CDLL(str(LIB_PATH))
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)  # AttributeError if the library does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    if lib.ugu_b200_abi_version() != 10:
        raise ImportError("libugu_b200_raster.so ABI version mismatch")
    _lib = lib
    return lib


def last_error(handle=None) -> str:
    msg = load().ugu_b200_last_error(handle)
    return msg.decode() if msg else ""
