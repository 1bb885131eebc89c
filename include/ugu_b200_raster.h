/*
 * ugu_b200_raster.h — C-ABI of the B200-native (sm_100a) software mesh rasterizer.
 *
 * This is the drop-in boundary for ONE path of unclearness/ugu: ugu::Rasterizer behind the
 * ugu::RendererCpu interface (reference include/ugu/renderer/cpu/renderer.h:42-76,
 * include/ugu/renderer/cpu/rasterizer.h:14-53, src/renderer/cpu/rasterizer.cc:105-307).
 * The reference has no FFI; the host-side C++ class that a maintainer adds to the reference tree
 * (ugu::RasterizerCuda : ugu::RendererCpu, ugu_b200/cpp/rasterizer_cuda.{h,cc}) binds exactly
 * these entry points.  See INTEGRATION.md.
 *
 * Conventions: plain C, POD descriptors, opaque handle, int status (0 = ok, see UGU_B200_E_*),
 * no exceptions, no C++/torch types.  All matrices are ROW-MAJOR.  All images are contiguous,
 * row-major, channels interleaved — the layout of ugu::ImageBase (reference
 * include/ugu/image.h:374-382,410-413).  There is NO CPU fallback: every entry point that computes
 * fails with UGU_B200_E_CUDA when no sm_100 device / driver is available.
 */
#ifndef UGU_B200_RASTER_H_
#define UGU_B200_RASTER_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define UGU_B200_ABI_VERSION 10

enum {
  UGU_B200_OK = 0,
  UGU_B200_E_INVALID = 1,   /* failed validation (mirrors util_private.cc:19-67 returning false) */
  UGU_B200_E_CUDA = 2,      /* CUDA runtime/driver error, or no usable device */
  UGU_B200_E_NOMEM = 3,
  UGU_B200_E_STATE = 4      /* call order violated (mesh not prepared, ...) */
};

/* Enumerators mirror the reference enums value for value. */
enum { UGU_B200_COLOR_NONE = 0, UGU_B200_COLOR_TEXTURE = 1, UGU_B200_COLOR_VERTEX = 2 }; /* renderer/base.h:11-15 */
enum { UGU_B200_INTERP_NN = 0, UGU_B200_INTERP_BILINEAR = 1 };                           /* common.h:50-53 */
enum { UGU_B200_NORMAL_FACE = 0, UGU_B200_NORMAL_VERTEX = 1 };                           /* renderer/base.h:19-22 */
enum { UGU_B200_SHADING_NONE = 0, UGU_B200_SHADING_LAMBERTIAN = 1, UGU_B200_SHADING_OREN_NAYAR = 2 }; /* base.h:26-32 */

/* Replaces ugu::RendererCpuOption (renderer/cpu/renderer.h:17-38). */
typedef struct ugu_b200_option {
  int32_t diffuse_color;
  int32_t interp;
  int32_t shading_normal;
  int32_t diffuse_shading;
  int32_t backface_culling;
  float depth_scale;       /* accepted and ignored, as the reference Rasterizer does */
  float oren_nayar_sigma;  /* the reference's CopyTo never copies it; hosts pass 0.3f (renderer.h:30-37) */
} ugu_b200_option;

/* Replaces the read side of ugu::Mesh (mesh.h:137-148): AoS host arrays exactly as
 * std::vector<Eigen::Vector3f/3i/2f>::data() lays them out.  Repacked to device SoA on upload. */
typedef struct ugu_b200_mesh_desc {
  int32_t num_vertices;
  int32_t num_normals;
  int32_t num_colors;        /* 0 or num_vertices */
  int32_t num_faces;
  int32_t num_face_normals;  /* 0 or num_faces */
  int32_t num_uv;
  int32_t num_materials;
  int32_t reserved;
  const float* vertices;          /* num_vertices * 3 */
  const float* normals;           /* num_normals * 3 */
  const float* vertex_colors;     /* num_colors * 3, RGB as float 0..255 */
  const int32_t* vertex_indices;  /* num_faces * 3 */
  const int32_t* normal_indices;  /* num_faces * 3, or NULL */
  const float* face_normals;      /* num_face_normals * 3 */
  const float* uv;                /* num_uv * 2 */
  const int32_t* uv_indices;      /* num_faces * 3, or NULL */
  const int32_t* material_ids;    /* num_faces, or NULL */
  const uint8_t* const* tex_data; /* num_materials pointers to RGB8 (Image3b) texels, NULL = empty */
  const int32_t* tex_w;
  const int32_t* tex_h;
} ugu_b200_mesh_desc;

/* Replaces what Render reads from ugu::PinholeCamera each call: size, intrinsics
 * (camera.h:341-347) and the float pose constants derived as rasterizer.cc:118-119
 * (w2c().rotation().cast<float>(), w2c().translation().cast<float>()) and camera.h:321
 * (c2w().matrix().block<3,3>(0,0).cast<float>()). */
typedef struct ugu_b200_camera {
  int32_t width, height;
  float fx, fy, cx, cy;
  float w2c_R[9];
  float w2c_t[3];
  float c2w_R[9];
} ugu_b200_camera;

/* The five outputs of RendererCpu::Render (renderer.h:60-61).  Any pointer may be NULL; all NULL is
 * an error (util_private.cc:63-67).  Sizes: color W*H*3 u8 (RGB), depth W*H f32, normal W*H*3 f32
 * (camera space), mask W*H u8 (0/255), face_id W*H i32 (-1 = none). */
typedef struct ugu_b200_outputs {
  uint8_t* color;
  float* depth;
  float* normal;
  uint8_t* mask;
  int32_t* face_id;
} ugu_b200_outputs;

/* Per-stage device timings (CUDA events on the rendering stream) and counters, accumulated since
 * the last ugu_b200_rasterizer_reset_stats(). */
enum {
  UGU_B200_STAGE_VERTEX = 0,   /* vertex transform + projection              (rasterizer.cc:125-136) */
  UGU_B200_STAGE_RASTER = 1,   /* triangle setup, small-triangle raster, large-triangle enqueue (:162-241) */
  UGU_B200_STAGE_BINNED = 2,   /* large-triangle tile-job raster              (:194-241) */
  UGU_B200_STAGE_RESOLVE = 3,  /* resolve + attributes + shading + packing    (:244-301) */
  UGU_B200_NUM_STAGES = 4
};
typedef struct ugu_b200_stats {
  double stage_ms[UGU_B200_NUM_STAGES]; /* summed event time of each stage over timed views */
  uint64_t timed_views;                 /* views that contributed to stage_ms */
  uint64_t views;                       /* views rendered */
  uint64_t kernel_launches;             /* kernels launched by this library */
  uint64_t h2d_bytes;                   /* per-view host->device bytes (camera constants) */
  uint64_t d2h_bytes;                   /* device->host bytes of outputs */
  uint64_t large_triangles;             /* last synchronous view: triangles sent to the tile-job path */
  uint64_t tile_jobs;                   /* last synchronous view: 8x4-pixel tile jobs */
  uint64_t hazard_fragments;            /* last synchronous view: fragments with z < FLT_MIN or NaN */
} ugu_b200_stats;

typedef struct ugu_b200_rasterizer ugu_b200_rasterizer;

int ugu_b200_abi_version(void);

/* Number of usable sm_100 devices (0 if none / no driver).  Never fails. */
int ugu_b200_device_count(void);

/* ctor of ugu::Rasterizer (rasterizer.cc:351-357).  `device` is a CUDA ordinal. */
int ugu_b200_rasterizer_create(int device, ugu_b200_rasterizer** out);
void ugu_b200_rasterizer_destroy(ugu_b200_rasterizer* r);

/* Rasterizer::set_option (rasterizer.cc:73-75). */
int ugu_b200_rasterizer_set_option(ugu_b200_rasterizer* r, const ugu_b200_option* opt);

/* Rasterizer::set_mesh + PrepareMesh (rasterizer.cc:77-99): the one-time H2D upload + SoA repack.
 * The host arrays are not referenced after the call returns. */
int ugu_b200_rasterizer_upload_mesh(ugu_b200_rasterizer* r, const ugu_b200_mesh_desc* mesh);

/* Rasterizer::Render (rasterizer.cc:105-307) into HOST buffers: validates, renders, copies the
 * requested outputs device->host and returns when they are complete.  Pinned buffers (from
 * ugu_b200_host_alloc or cudaHostRegister) are written by DMA directly; pageable ones through an
 * internal pinned staging ring. */
int ugu_b200_rasterizer_render(ugu_b200_rasterizer* r, const ugu_b200_camera* cam,
                               const ugu_b200_outputs* host_out);

/* Same, outputs left in DEVICE memory (16-byte aligned device pointers), asynchronous on `stream`
 * (a cudaStream_t passed as void*; NULL = the rasterizer's own stream). */
int ugu_b200_rasterizer_render_device(ugu_b200_rasterizer* r, const ugu_b200_camera* cam,
                                      const ugu_b200_outputs* dev_out, void* stream);

/* Multi-view batch (the ex09 orbit loop, examples/ex09_renderer_realtime.cc:125-182): view i writes
 * at element offset i*W*H of every non-NULL output.  All cameras must share width/height.
 * `outputs_on_device` selects device-resident (async on `stream`) or host (synchronous) outputs. */
int ugu_b200_rasterizer_render_batch(ugu_b200_rasterizer* r, int32_t num_views,
                                     const ugu_b200_camera* cams, const ugu_b200_outputs* out,
                                     int32_t outputs_on_device, void* stream);

/* Off-contract fragments (depth < FLT_MIN or NaN, where the reference's sequential `d < FLT_MIN || z < d` rule
 * becomes order dependent, rasterizer.cc:229) met by the last HOST-form render call on this handle (render,
 * render_batch with host outputs, render_gbuffer, raytrace): 0 = the frame is the reference's frame; > 0 = it may
 * differ at those fragments.  The call itself still returns UGU_B200_OK; ugu::RasterizerCuda logs a warning. */
uint64_t ugu_b200_rasterizer_last_hazards(const ugu_b200_rasterizer* r);

/* Block until everything queued by this rasterizer has finished; returns the first deferred error. */
int ugu_b200_rasterizer_sync(ugu_b200_rasterizer* r);

/* Rasterizer::RenderW's depth conversion (rasterizer.cc:329-344 + ConvertTo, image_proc.h:453-460):
 * truncating f32 -> u16 cast of a host depth image. */
int ugu_b200_depth_to_u16(const float* depth, uint16_t* depth_w, size_t count);

/* ugu::Depth2PointCloud(const Image1f& depth, const Camera&, Image3f* point_cloud, bool gl_coord)
 * (src/util/rgbd_util.cc:472-513; PinholeCamera::Unproject, include/ugu/camera.h:406-413) — the step
 * examples/ex02_renderer.cc:156-158 runs on the rendered depth.  point_cloud is W*H*3 f32 in
 * ugu::Image3f layout, camera space.  The reference writes only pixels with depth >= FLT_MIN and
 * leaves the others as they were: `zero_invalid` = 0 keeps that behaviour (the buffer is in/out),
 * 1 writes zeros there (what a freshly Init()-ed image holds; saves the upload of the old image).
 * Only size and intrinsics of `cam` are used.
 * _device: depth / point_cloud are 16-byte aligned DEVICE pointers (e.g. the depth output of
 * render_device), asynchronous on `stream`.  The host form copies in, runs, copies out, returns
 * when the image is complete. */
int ugu_b200_depth_to_point_cloud_device(ugu_b200_rasterizer* r, const ugu_b200_camera* cam,
                                         const float* depth_dev, float* point_cloud_dev,
                                         int32_t gl_coord, int32_t zero_invalid, void* stream);
int ugu_b200_depth_to_point_cloud(ugu_b200_rasterizer* r, const ugu_b200_camera* cam,
                                  const float* depth_host, float* point_cloud_host, int32_t gl_coord,
                                  int32_t zero_invalid);

/* UV-space attribute rasterization: replaces
 *   ugu::RasterizeVertexAttributeToTexture<T>  (include/ugu/util/raster_util.h:139-226)  per_face = 0
 *   ugu::RasterizeFaceAttributeToTexture<T>    (raster_util.h:228-268)                   per_face = 1
 * with T = Vec3b (texture_type 0, W*H*3 u8) or Vec3f (texture_type 1, W*H*3 f32), in the reference's
 * SEQUENTIAL semantics (num_threads = 1: faces in ascending index, the last writer of a texel wins;
 * with more threads the reference itself is racy).  Arrays are the AoS data() of the reference's
 * std::vector<Eigen::Vector3f/3i/2f> arguments.  The texture is zero-initialised like
 * Image<T>::zeros(height, width) unless keep_texture = 1, which is the reference's width/height <= 0
 * form (rasterise over the existing contents, raster_util.h:151-155).  `mask` (W*H u8, may be NULL)
 * gets 255 on covered texels; the reference never initialises it, so keep_mask = 1 leaves the other
 * texels as they were (in/out) and keep_mask = 0 writes 0 there (a fresh zeros image).
 * Returns UGU_B200_E_INVALID where the reference returns false (no faces / mismatched sizes,
 * raster_util.h:148-151,238-240) and for out-of-range indices (undefined behaviour in the reference). */
typedef struct ugu_b200_uv_raster_desc {
  int32_t num_attrs;            /* per_face: must equal num_faces */
  int32_t num_faces;
  int32_t num_uv;
  int32_t per_face;
  const float* attrs;           /* num_attrs * 3 */
  const int32_t* attr_indices;  /* num_faces * 3; ignored when per_face */
  const float* uv;              /* num_uv * 2 */
  const int32_t* uv_indices;    /* num_faces * 3 */
  int32_t width, height;
  int32_t texture_type;         /* 0 = Image3b, 1 = Image3f */
  int32_t keep_texture;
  int32_t keep_mask;
  int32_t reserved;
} ugu_b200_uv_raster_desc;

/* Host form: arrays and images in host memory; uploads, runs, copies back, returns when complete. */
int ugu_b200_uv_raster(ugu_b200_rasterizer* r, const ugu_b200_uv_raster_desc* d, void* texture_host,
                       uint8_t* mask_host);
/* Device form: every pointer of `d`, `texture_dev` and `mask_dev` are 16-byte aligned DEVICE
 * pointers (index ranges are NOT validated); asynchronous on `stream`. */
int ugu_b200_uv_raster_device(ugu_b200_rasterizer* r, const ugu_b200_uv_raster_desc* d,
                              void* texture_dev, uint8_t* mask_dev, void* stream);
/* ugu::GenerateUvMask(uvs, uv_indices, Image1b& mask, uint8_t color, width, height)
 * (src/util/raster_util.cc:9-27): `color` where any face covers the texel, else 0.  Host arrays. */
int ugu_b200_uv_mask(ugu_b200_rasterizer* r, int32_t num_faces, int32_t num_uv, const float* uv,
                     const int32_t* uv_indices, int32_t width, int32_t height, int32_t color,
                     uint8_t* mask_host);

/* Face-visibility consumer of rendered face_id maps: the rasterizer-side counterpart
 * of the per-keyframe face statistics the reference's texturing pipeline collects with BVH rays
 * (ugu::FaceInfo / FaceInfoPerKeyframe, include/ugu/texturing/visibility_tester.h:135-181;
 * FaceInfo::CalcStat's max_area_id, src/texturing/visibility_tester.cc:185-218).  There is no
 * reference function that consumes face_id maps, so the definition below is this library's own:
 *   pixel_count[f]  pixels owned by face f, summed over the views
 *   view_count[f]   number of views in which f owns at least one pixel
 *   best[f]         max over views of ((pixels of f in the view) << 32 | (0xFFFFFFFF - view id)),
 *                   0 = never visible: the view that sees most of f, ties to the lowest view id
 * Ids outside [0, num_faces) (-1 = background) are ignored.  The three vectors combine across ranks
 * by SUM / SUM / MAX, so a multi-GPU job all-reduces 16 bytes per face instead of gathering maps.
 * _device: `face_id_dev` holds num_views maps of pixels_per_view int32 each (the face_id output of
 * render_batch); the accumulators are DEVICE arrays of num_faces elements that the caller zeroes
 * once and may accumulate into over several calls (view ids first_view_id .. + num_views - 1);
 * asynchronous on `stream`. */
int ugu_b200_face_visibility_device(ugu_b200_rasterizer* r, const int32_t* face_id_dev,
                                    int32_t num_views, int64_t pixels_per_view, int32_t first_view_id,
                                    int32_t num_faces, uint32_t* pixel_count_dev,
                                    uint32_t* view_count_dev, uint64_t* best_dev, void* stream);
/* Host form: maps and results in host memory; results are computed from zero for these maps.
 * best_view[f] = -1 and best_count[f] = 0 for faces never visible.  Any output may be NULL. */
int ugu_b200_face_visibility(ugu_b200_rasterizer* r, const int32_t* face_id_host, int32_t num_views,
                             int64_t pixels_per_view, int32_t first_view_id, int32_t num_faces,
                             uint32_t* pixel_count_host, uint32_t* view_count_host,
                             int32_t* best_view_host, uint32_t* best_count_host);

/* Allocation-free variant of ugu_b200_face_visibility_device.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture,
 * tight serving loops, harnesses that forbid it).  Same inputs, accumulators and results as the _device
 * form, but it takes no rasterizer handle and its histogram scratch is a file-scope __device__ array of
 * UGU_B200_STATIC_VIS_MAX_FACES 32-bit words (16 MiB) that every GPU loading the library holds until the
 * process exits.  The call performs no cudaMalloc / cudaFree / allocating library routine: it makes `device`
 * current, clears the part of the array it uses on `stream`, and queues the kernels there (asynchronous;
 * `stream` NULL = the default stream).  num_faces > UGU_B200_STATIC_VIS_MAX_FACES returns
 * UGU_B200_E_INVALID.  Views are processed min(8, MAX_FACES / num_faces) at a time.  ALL calls on a device
 * share the array: calls that overlap on the GPU (different streams, concurrent threads) race — serialise
 * them.  On failure a message is written to `err` (err_len bytes, may be NULL). */
#define UGU_B200_STATIC_VIS_MAX_FACES (4 * 1024 * 1024)
int ugu_b200_face_visibility_static(int32_t device, const int32_t* face_id_dev, int32_t num_views,
                                    int64_t pixels_per_view, int32_t first_view_id, int32_t num_faces,
                                    uint32_t* pixel_count_dev, uint32_t* view_count_dev, uint64_t* best_dev,
                                    void* stream, char* err, size_t err_len);

/* Keyframe visibility pass: replaces ugu::VisibilityTester
 * (include/ugu/texturing/visibility_tester.h:216-265; src/texturing/visibility_tester.cc:416-529 TestVertex,
 * :531-549 TestVertices, :551-604 TestFaces, :606-668 TestFacewiseVertices, :670-733 Test) and the folds of
 * ugu::VisibilityInfo (FaceInfo::CalcStat :185-218; the min / mean part of VertexInfo::CalcStat :66-152).
 * Per keyframe (a pinhole camera + colour / depth / mask images) every mesh vertex is projected, bounds-,
 * mask- and depth-checked at its nearest pixel, back-face tested with its vertex normal against the ray through
 * its sub-pixel projection, occlusion-tested, and sampled for colour (NN or the reference's bilinear) and
 * intensity; a face is visible when its three vertices are, with the mean viewing angle / distance and the
 * projected area of FaceInfoPerKeyframe.  `facewise` selects TestFacewiseVertices (the reference's default:
 * each corner pulled 1e-3 of the way to the face centre; a vertex keeps the record of the first passing (face,
 * corner) in face order — the reference's sequential order, its OpenMP form is racy).
 * ONE substitution: the reference asks a BVH for the nearest hit of the vertex ray (:457-483); this library reads
 * `zbuf`, the depth map of the mesh itself from the keyframe camera as ITS rasterizer renders it with back-face
 * culling off, at the nearest pixel: with t_v = |camera_p| and t_s = zbuf(nn) * (t_v / camera_p.z), the vertex
 * is occluded iff zbuf(nn) >= FLT_MIN and t_v - t_s > eps; `distance` = t_v (reference: the hit's t, which is
 * smaller when a nearer surface lies within eps); the depth check uses camera_p.z.  On the (vertex, keyframe)
 * pairs both methods call visible, projected_pos, colour, intensity and area are bit-identical to the compiled
 * reference, viewing_angle to a few ulp (device acosf), distance to float tolerance; the decisions differ on
 * < 1 % of the pairs (sub-pixel rays vs nearest-pixel depth at silhouettes; tests/test_visibility_tester.py
 * states the set).  Field names and defaults: ugu::VisibilityTesterOption (visibility_tester.h:39-58) — note
 * that the reference's CopyTo (:43-53) copies only the first seven fields, so through its public API
 * max_viewing_angle / max_distance / max_depth_difference / eps always keep their defaults. */
typedef struct ugu_b200_vis_option {
  float backface_culling_angle_rad_th;        /* default radians(90) */
  int32_t use_mask, use_depth;                /* default 1, 1 */
  int32_t interp;                             /* UGU_B200_INTERP_*, default bilinear */
  float max_viewing_angle;                    /* default radians(90) */
  float max_distance_from_camera_to_vertex;   /* default FLT_MAX */
  float max_depth_difference;                 /* default FLT_MAX */
  float eps;                                  /* default 0.1 */
  int32_t facewise;                           /* Test(..., facewise): default 1 */
  int32_t self_occlusion;                     /* 1: use the keyframe's zbuf, or render it from the uploaded
                                                 rasterizer mesh when the pointer is NULL; 0: no occlusion test */
} ugu_b200_vis_option;
typedef struct ugu_b200_vis_vertex {          /* ugu::VertexInfoPerKeyframe (visibility_tester.h:60-72) */
  float color[3];
  float intensity;
  float projected_pos[2];
  float viewing_angle;
  float distance;
} ugu_b200_vis_vertex;
typedef struct ugu_b200_vis_face {            /* ugu::FaceInfoPerKeyframe (visibility_tester.h:130-144) */
  float area, viewing_angle, distance;
  float projected_tri[6];
} ugu_b200_vis_face;
/* ugu::Keyframe (visibility_tester.h:201-214) + the float constants the tester reads from its camera:
 * cam.{width,height,fx,fy,cx,cy,c2w_R}, w2c = camera->w2c().cast<float>() (row-major 3x4), c2w_t = c2w_t_f_.
 * cam.w2c_R / cam.w2c_t are used only when the library renders zbuf itself.  Images: colour W*H*3 u8, depth W*H
 * f32 (may be NULL when !use_depth), mask W*H u8 (NULL when !use_mask), zbuf W*H f32 (see self_occlusion). */
typedef struct ugu_b200_keyframe {
  int32_t id;
  int32_t reserved;
  ugu_b200_camera cam;
  float w2c[12];
  float c2w_t[3];
  float reserved2;
  const uint8_t* color;
  const float* depth;
  const uint8_t* mask;
  const float* zbuf;
} ugu_b200_keyframe;
/* Dense per-keyframe records: row k (keyframe k of the call) holds num_vertices / num_faces entries; entries of
 * invisible elements are zero.  Any pointer may be NULL. */
typedef struct ugu_b200_vis_records {
  uint8_t* vertex_visible;           /* K * V */
  ugu_b200_vis_vertex* vertex;       /* K * V, 16-byte aligned */
  uint8_t* face_visible;             /* K * F */
  ugu_b200_vis_face* face;           /* K * F */
} ugu_b200_vis_records;
/* VisibilityInfo after CalcStatFaceInfo / CalcStatVertexInfo, restricted to what the device folds: per face the
 * number of keyframes that see it and the FIRST keyframe (in call order, `order` = first_order + position) with
 * the largest area / smallest viewing angle / smallest distance — FaceInfo::max_area_id, min_viewing_angle_id,
 * min_distance_id are keyframes[order].id; per vertex the count, VertexInfo::mean_color / mean_viewing_angle /
 * mean_distance, min_viewing_angle_id / _color, min_distance_id / _color.  -1 / 0 where never visible.  Medians,
 * weighted medians and modes need the per-keyframe lists: take them from the dense records.  HOST arrays, any NULL. */
typedef struct ugu_b200_vis_stats {
  uint32_t* face_count;
  int32_t* face_max_area_order;
  int32_t* face_min_viewing_angle_order;
  int32_t* face_min_distance_order;
  float* face_max_area;
  float* face_min_viewing_angle;
  float* face_min_distance;
  uint32_t* vertex_count;
  int32_t* vertex_min_viewing_angle_order;
  int32_t* vertex_min_distance_order;
  float* vertex_mean_color;               /* V * 3 */
  float* vertex_mean_viewing_angle;
  float* vertex_mean_distance;
  float* vertex_min_viewing_angle_color;  /* V * 3 */
  float* vertex_min_distance_color;       /* V * 3 */
} ugu_b200_vis_stats;
/* Raw DEVICE accumulators (for multi-GPU jobs: ranks fold disjoint keyframes and combine with
 * SUM / MAX / MIN / MIN, SUM / SUM / SUM / MIN / MIN; see ugu_b200/csrc/visibility_tester_kernels.cuh). */
typedef struct ugu_b200_vis_accumulators {
  uint32_t* face_count;        /* F, SUM */
  uint64_t* face_max_area;     /* F, MAX: area bits << 32 | (0xFFFFFFFF - order); 0 = never */
  uint64_t* face_min_angle;    /* F, MIN: angle bits << 32 | order; ~0 = never */
  uint64_t* face_min_dist;     /* F, MIN */
  uint32_t* vertex_count;      /* V, SUM */
  float* vertex_sum_color;     /* V * 4 (r g b -), SUM */
  float* vertex_sum_angle_dist;/* V * 2, SUM */
  uint64_t* vertex_min_angle;  /* V, MIN */
  uint64_t* vertex_min_dist;   /* V, MIN */
  uint32_t* vertex_min_colors; /* V * 2: packed r | g << 8 | b << 16 of the min_angle / min_dist winner */
} ugu_b200_vis_accumulators;

/* VisibilityTester::set_data + PrepareData (visibility_tester.cc:236-270): uploads vertices (V*3), per-vertex
 * normals (V*3) and indices (F*3) and clears the accumulators.  E_INVALID for an empty mesh (PrepareData's false). */
int ugu_b200_visibility_set_data(ugu_b200_rasterizer* r, int32_t num_vertices, const float* vertices,
                                 const float* normals, int32_t num_faces, const int32_t* indices);
/* A fresh VisibilityInfo: clears the accumulators. */
int ugu_b200_visibility_reset(ugu_b200_rasterizer* r);
/* VisibilityTester::Test over `num_keyframes` keyframes with HOST images (uploads them, runs, copies the
 * requested dense records to HOST arrays, returns when complete) and folds them into the accumulators with
 * orders first_order .. first_order + num_keyframes - 1.  E_INVALID where the reference returns false (colour
 * empty, use_mask / use_depth without the image, no data set). */
int ugu_b200_visibility_test(ugu_b200_rasterizer* r, int32_t num_keyframes, const ugu_b200_keyframe* keyframes,
                             int32_t first_order, const ugu_b200_vis_option* opt,
                             const ugu_b200_vis_records* host_records);
/* Same with DEVICE images and DEVICE record arrays (e.g. the colour / depth / mask outputs of render_batch),
 * asynchronous on `stream`. */
int ugu_b200_visibility_test_device(ugu_b200_rasterizer* r, int32_t num_keyframes,
                                    const ugu_b200_keyframe* keyframes, int32_t first_order,
                                    const ugu_b200_vis_option* opt, const ugu_b200_vis_records* dev_records,
                                    void* stream);
/* Finalises the accumulators (means = sums / count in float) into HOST arrays; waits for pending work. */
int ugu_b200_visibility_stats(ugu_b200_rasterizer* r, const ugu_b200_vis_stats* host_out);
int ugu_b200_visibility_accumulators(ugu_b200_rasterizer* r, ugu_b200_vis_accumulators* out);
/* The float constants of a keyframe camera from a row-major 4x4 double c2w, as PinholeCamera::set_c2w derives
 * them (camera.h:317-331): fills kf->cam (like ugu_b200_camera_from_c2w), kf->w2c and kf->c2w_t.  Host only. */
int ugu_b200_keyframe_camera_from_c2w(const double* c2w16, int32_t width, int32_t height, float fx, float fy,
                                      float cx, float cy, ugu_b200_keyframe* kf);

/* ugu::Depth2Mesh (include/ugu/util/rgbd_util.h:27-37; Depth2MeshImpl, src/util/rgbd_util.cc:94-282,
 * 528-545; Mesh::CalcNormal, src/mesh.cc:297-349) — the step examples/ex02_renderer.cc:156-158 runs on
 * the rendered depth and colour: grid points (x = x_step, 2 x_step, .. < W; y likewise) with depth >=
 * FLT_MIN become vertices in raster order (camera space; y, z negated for gl_coord), each connects to
 * its upper-left / upper / left neighbours with up to two triangles when the depth differences are
 * below max_connect_z_diff (and — sic — the neighbours are not vertex 0), then face and vertex
 * normals.  `color` NULL = the depth-only overload; else the textured overload: uv per vertex, and
 * nearest-texel vertex colours (float 0..255) when with_vertex_color.  Only size and intrinsics of
 * `cam` are used; cam->width/height are the depth image's.
 * Returns UGU_B200_E_INVALID where the reference returns false (negative max_connect_z_diff, steps <
 * 1, aspect ratios of depth and colour differing by more than 0.01).
 * The result stays in device memory until the next call; `counts` tells how much to fetch. */
typedef struct ugu_b200_depth2mesh_counts {
  int32_t num_vertices;
  int32_t num_faces;
  int32_t has_normals;   /* 0 when there are no faces: Mesh::CalcNormal returns early (mesh.cc:298-300) */
  int32_t has_uv;
  int32_t has_colors;
  int32_t reserved;
} ugu_b200_depth2mesh_counts;

/* Host form: uploads depth (and colour), runs, returns the counts. */
int ugu_b200_depth_to_mesh(ugu_b200_rasterizer* r, const ugu_b200_camera* cam, const float* depth_host,
                           const uint8_t* color_host, int32_t color_width, int32_t color_height,
                           int32_t with_vertex_color, float max_connect_z_diff, int32_t x_step,
                           int32_t y_step, int32_t gl_coord, ugu_b200_depth2mesh_counts* counts);
/* Device form: depth_dev / color_dev are DEVICE pointers (e.g. outputs of render_device);
 * asynchronous on `stream`; the counts come with the fetch. */
int ugu_b200_depth_to_mesh_device(ugu_b200_rasterizer* r, const ugu_b200_camera* cam,
                                  const float* depth_dev, const uint8_t* color_dev, int32_t color_width,
                                  int32_t color_height, int32_t with_vertex_color,
                                  float max_connect_z_diff, int32_t x_step, int32_t y_step,
                                  int32_t gl_coord, void* stream);
/* The two-camera overload (rgbd_util.h:39-48; Depth2MeshImpl, src/util/rgbd_util.cc:283-466): the colour image
 * belongs to a second pinhole camera; uv and nearest-texel vertex colours come from the projection of
 * depth2color * p (p = the vertex in the depth camera's frame, AFTER the gl_coord flip, as the reference does) into
 * it (PinholeCamera::Project, camera.h:380-387), then X2U / Y2V over the colour image's size.  depth2color_R is the
 * row-major 3x3 linear part and depth2color_t the translation of the reference's Eigen::Affine3f.  No aspect-ratio
 * check in this overload; the colour image is always required; a vertex that projects outside the colour image
 * reads out of bounds in the reference — here its texel coordinates are clamped.  The reference Init()s the two
 * organized images only when their size differs from the depth image's; this library always returns what fresh
 * images would hold.  Results are fetched with ugu_b200_depth_to_mesh_fetch. */
typedef struct ugu_b200_color_camera {
  float depth2color_R[9];
  float depth2color_t[3];
  float fx, fy, cx, cy;   /* colour camera intrinsics; its size is color_width x color_height */
} ugu_b200_color_camera;
int ugu_b200_depth_to_mesh_two_camera(ugu_b200_rasterizer* r, const ugu_b200_camera* depth_cam,
                                      const ugu_b200_color_camera* color_cam, const float* depth_host,
                                      const uint8_t* color_host, int32_t color_width, int32_t color_height,
                                      int32_t with_vertex_color, float max_connect_z_diff, int32_t x_step,
                                      int32_t y_step, int32_t gl_coord, ugu_b200_depth2mesh_counts* counts);
int ugu_b200_depth_to_mesh_two_camera_device(ugu_b200_rasterizer* r, const ugu_b200_camera* depth_cam,
                                             const ugu_b200_color_camera* color_cam, const float* depth_dev,
                                             const uint8_t* color_dev, int32_t color_width, int32_t color_height,
                                             int32_t with_vertex_color, float max_connect_z_diff, int32_t x_step,
                                             int32_t y_step, int32_t gl_coord, void* stream);
/* Waits for the last depth_to_mesh call and copies its result to HOST arrays (any may be NULL):
 * vertices / normals / colors num_vertices*3 f32, uv num_vertices*2, faces num_faces*3 i32 (they
 * are the mesh's vertex_indices, normal_indices and uv_indices alike), face_normals num_faces*3,
 * point_cloud / normal_image W*H*3 f32 (the organized images of rgbd_util.cc:257-277, zero where
 * there is no vertex).  `counts` (may be NULL) receives the counts. */
int ugu_b200_depth_to_mesh_fetch(ugu_b200_rasterizer* r, ugu_b200_depth2mesh_counts* counts,
                                 float* vertices, int32_t* faces, float* normals, float* face_normals,
                                 float* uv, float* colors, float* point_cloud, float* normal_image);

/* G-buffer outputs: the fields of ugu::GBuffer (include/ugu/renderer/base.h:34-77) as
 * RendererGl::ReadGbuf fills them (src/renderer/gl/renderer.cc:501-650) from the gbuf shaders
 * (src/shader/glsl/vert/gbuf.vert:42-56, geom/dummy.geom:26-80, frag/gbuf.frag:21-44), computed from the
 * winners of THIS rasterizer (the CPU Rasterizer's sampling rule, rasterizer.cc:194-237) — the GL
 * pipeline itself is not available headless.  Exact per-field definition and float op order: the header
 * comment of ugu_b200/csrc/gbuffer_kernels.cuh.  Summary, per pixel with a visible winner (face f, depth
 * z, weights (1-u-v, u, v)) whose window depth d lies in (0, 1):
 *   stencil 255, depth_01 = d, face_id = f + 1, geo_id = opt.geo_id, bary = (u, v, 1-u-v),
 *   pos_wld, pos_cam = w2c_t + w2c_R pos_wld, normal_wld (normalised interpolated vertex normal, or the
 *   flat face normal), normal_cam = w2c_R normal_wld, uv = (s, t, 1-s-t); color = the Render colour.
 * Elsewhere: depth_01 = clamp(d, 0, 1) (1 on background), everything else 0 (face_id 0, not -1).
 * Sizes: 3-float images W*H*3 f32, depth_01 W*H f32, stencil W*H u8, face_id / geo_id W*H i32, color
 * W*H*3 u8.  Any pointer may be NULL; all NULL is UGU_B200_E_INVALID.  GBuffer::shaded (the viewer's
 * deferred lighting pass) is not produced. */
typedef struct ugu_b200_gbuffer {
  uint8_t* color;
  float* normal_cam;
  float* normal_wld;
  float* pos_cam;
  float* pos_wld;
  float* depth_01;
  uint8_t* stencil;
  int32_t* face_id;
  float* uv;
  float* bary;
  int32_t* geo_id;
} ugu_b200_gbuffer;
typedef struct ugu_b200_gbuffer_option {
  float z_near, z_far;   /* RendererGl::SetNearFar (renderer.cc:751-754); 0 < z_near < z_far */
  int32_t flat_normal;   /* RendererGl::SetFlatNormal: face normal from the world positions */
  int32_t geo_id;        /* written to geo_id where stencil == 255 (one geometry per rasterizer) */
} ugu_b200_gbuffer_option;
/* Host form (synchronous, HOST pointers) and device form (16-byte aligned DEVICE pointers,
 * asynchronous on `stream`).  Uses the options of set_option for colour and back-face culling. */
int ugu_b200_rasterizer_render_gbuffer(ugu_b200_rasterizer* r, const ugu_b200_camera* cam,
                                       const ugu_b200_gbuffer_option* opt, const ugu_b200_gbuffer* host_out);
int ugu_b200_rasterizer_render_gbuffer_device(ugu_b200_rasterizer* r, const ugu_b200_camera* cam,
                                              const ugu_b200_gbuffer_option* opt,
                                              const ugu_b200_gbuffer* dev_out, void* stream);

/* Raytracer parity mode: the outputs of ugu::Raytracer::Render
 * (src/renderer/cpu/raytracer.cc:117-225) — per pixel the nearest hit of the primary ray as
 * BvhNanort::Intersect(ray, test_all = false) finds it (include/ugu/accel/bvh_nanort.h:109-152; nanort's
 * watertight TriangleIntersector, third_party/nanort/nanort.h:1050-1199, t in [1e-4, 1e30)), back-face
 * test on the hit (face_normal . ray_w > 0 leaves the pixel empty), face_id = the hit face (face 0
 * included), mask 255, depth = (w2c_R (org + ray_w t) + w2c_t).z * option.depth_scale, normal and colour
 * from the intersector's (u, v) through the same pixel shaders.  No BVH on the device: triangles are
 * scattered to the pixels their projection can cover and reduced with a 64-bit atomicMin on (t, ~face)
 * (ugu_b200/csrc/raytrace_kernels.cuh).  Two triangles with bit-identical t at a pixel: the highest face
 * index wins (nanort: the last one its traversal visits).
 * `org_w` = 3 floats, PinholeCamera::org_ray_w = the c2w translation cast to float (camera.h:431-436).
 * Same options, validation, outputs and error behaviour as ugu_b200_rasterizer_render. */
int ugu_b200_rasterizer_raytrace(ugu_b200_rasterizer* r, const ugu_b200_camera* cam, const float* org_w,
                                 const ugu_b200_outputs* host_out);
int ugu_b200_rasterizer_raytrace_device(ugu_b200_rasterizer* r, const ugu_b200_camera* cam, const float* org_w,
                                        const ugu_b200_outputs* dev_out, void* stream);

/* Enable (1) / disable (0) per-stage CUDA-event timing; `stage_mask` bit s times stage s. */
int ugu_b200_rasterizer_set_profiling(ugu_b200_rasterizer* r, int32_t enable, uint32_t stage_mask);
int ugu_b200_rasterizer_get_stats(ugu_b200_rasterizer* r, ugu_b200_stats* out);
int ugu_b200_rasterizer_reset_stats(ugu_b200_rasterizer* r);

/* Host-side camera constants from a row-major 4x4 double c2w, as PinholeCamera::set_c2w
 * (camera.h:317-331) and rasterizer.cc:118-119 derive them.  Pure host code, needs no GPU. */
int ugu_b200_camera_from_c2w(const double* c2w16, int32_t width, int32_t height, float fx, float fy,
                             float cx, float cy, ugu_b200_camera* out);

/* Pinned host memory for outputs (so Render's D2H is a single DMA). */
void* ugu_b200_host_alloc(size_t bytes);
void ugu_b200_host_free(void* p);

/* Message of the last error on this rasterizer (or of the last handle-less call when r == NULL). */
const char* ugu_b200_last_error(const ugu_b200_rasterizer* r);

/* ---- Multi-view dispatch over several GPUs of one box from ONE host process (SURVEY.md 8b last row, 8e) --------
 * The C++ replacement of the reference's sequential orbit loop (examples/ex09_renderer_realtime.cc:125-182:
 * set_c2w -> Render per view on one renderer): one rasterizer handle per listed device, the view list sharded
 * over them (no data-path collective: views are independent given the replicated mesh), each device driven by
 * its own host thread through the batched launch sets of ugu_b200_rasterizer_render_batch.  The optional exchange
 * step — gathering the per-view face_id maps on one device for a consumer — sends every chunk of views as soon as it
 * is rendered, straight to its final place in the view-ordered destination, over NCCL send/recv (libnccl.so.2 is
 * dlopen-ed at create) or cudaMemcpyPeerAsync; the per-face visibility fold reduces 16 bytes per face instead.
 * A device may be listed more than once (several handles on one GPU: the sharding logic on a one-GPU box);
 * NCCL needs distinct devices, the peer transport does not. */
typedef struct ugu_b200_multiview ugu_b200_multiview;
enum { UGU_B200_PARTITION_CONTIGUOUS = 0, UGU_B200_PARTITION_ROUND_ROBIN = 1 };
enum { UGU_B200_TRANSPORT_AUTO = 0, UGU_B200_TRANSPORT_PEER = 1, UGU_B200_TRANSPORT_NCCL = 2 };

/* Global indices of the views slot `rank` of `world` renders.  contiguous: ceil/floor(num_views / world)
 * consecutive views (the first num_views % world slots get one more); round robin: i % world == rank.
 * `indices` may be NULL (count only).  Pure host code. */
int ugu_b200_partition(int32_t num_views, int32_t world, int32_t rank, int32_t mode, int32_t* indices,
                       int32_t* count);

/* transport AUTO: NCCL when libnccl.so.2 loads and the devices are distinct, else peer copies (the
 * environment variable UGU_B200_MV_TRANSPORT=peer|nccl overrides the argument). */
int ugu_b200_multiview_create(const int32_t* devices, int32_t num_devices, int32_t transport,
                              ugu_b200_multiview** out);
void ugu_b200_multiview_destroy(ugu_b200_multiview* mv);
int32_t ugu_b200_multiview_size(const ugu_b200_multiview* mv);
int32_t ugu_b200_multiview_transport(const ugu_b200_multiview* mv); /* UGU_B200_TRANSPORT_PEER or _NCCL */
ugu_b200_rasterizer* ugu_b200_multiview_rasterizer(ugu_b200_multiview* mv, int32_t slot);
const char* ugu_b200_multiview_last_error(const ugu_b200_multiview* mv);

/* set_option / PrepareMesh on every device (the mesh is replicated). */
int ugu_b200_multiview_set_option(ugu_b200_multiview* mv, const ugu_b200_option* opt);
int ugu_b200_multiview_upload_mesh(ugu_b200_multiview* mv, const ugu_b200_mesh_desc* mesh);

/* The orbit loop, sharded: view i lands at element offset i * W*H of every non-NULL HOST output, whatever device
 * rendered it.  All cameras share one image size.  Synchronous. */
int ugu_b200_multiview_render_batch(ugu_b200_multiview* mv, int32_t num_views, const ugu_b200_camera* cams,
                                    int32_t partition, const ugu_b200_outputs* host_out);

/* Render the face_id map of every view (sharded, `chunk_views` views per launch set and message; 0 = 16) and
 * gather the maps in view order on device slot `root`: into `face_id_root_dev` (a 16-byte aligned device pointer
 * on that device, num_views * W*H int32) or, when it is NULL, into an internal buffer
 * (ugu_b200_multiview_gathered).  `face_id_host` (optional) additionally receives the gathered maps.  Synchronous. */
int ugu_b200_multiview_gather_face_ids(ugu_b200_multiview* mv, int32_t num_views, const ugu_b200_camera* cams,
                                       int32_t partition, int32_t root, int32_t chunk_views,
                                       int32_t* face_id_root_dev, int32_t* face_id_host);
const int32_t* ugu_b200_multiview_gathered(const ugu_b200_multiview* mv);

/* Render the face_id maps (sharded), fold them per device into the per-face statistics of
 * ugu_b200_face_visibility (global view ids) and reduce the accumulators to slot `root` — SUM / SUM / MAX of 16
 * bytes per face (ncclReduce, or peer copies and a combine kernel) instead of the maps.  Results as in the host
 * form of ugu_b200_face_visibility; any output may be NULL.  Synchronous. */
int ugu_b200_multiview_reduce_face_visibility(ugu_b200_multiview* mv, int32_t num_views,
                                              const ugu_b200_camera* cams, int32_t partition, int32_t root,
                                              int32_t chunk_views, uint32_t* pixel_count_host,
                                              uint32_t* view_count_host, int32_t* best_view_host,
                                              uint32_t* best_count_host);

#ifdef __cplusplus
}
#endif
#endif /* UGU_B200_RASTER_H_ */
