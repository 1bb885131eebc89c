// TEST INFRASTRUCTURE — NOT PART OF THE PRODUCT.
//
// A thin extern "C" shim around the UNMODIFIED reference ugu::Rasterizer, compiled from the
// sources where they lie under /root/reference by oracle/Makefile into oracle/_ref/libugu_ref.so.
// It exists so that tests/ and bench.py's cpu_baseline / `--impl reference` legs can drive the
// real reference implementation (include/ugu/renderer/cpu/rasterizer.h:14-53,
// src/renderer/cpu/rasterizer.cc:105-307) from plain arrays.  Nothing under ugu_b200/ may load it.
//
// The shim only converts flat arrays <-> ugu::Mesh / ugu::PinholeCamera / ugu::Image and calls the
// reference's public API in the call order of examples/ex02_renderer.cc:245-278,142-143.

#include <chrono>
#ifdef _OPENMP
#include <omp.h>
#endif
#include <cstdint>
#include <cstring>
#include <memory>
#include <vector>

#include "ugu/camera.h"
#include "ugu/log.h"
#include "ugu/mesh.h"
#include "ugu/renderer/cpu/rasterizer.h"
#include "ugu/renderer/cpu/raytracer.h"
#include "ugu/texturing/visibility_tester.h"
#include "ugu/util/raster_util.h"
#include "ugu/util/rgbd_util.h"

extern "C" {

struct UguRefMeshDesc {
  int32_t num_vertices;
  int32_t num_normals;
  int32_t num_colors;
  int32_t num_faces;
  int32_t num_face_normals;
  int32_t num_uv;
  int32_t num_materials;
  int32_t pad_;
  const float* vertices;          // num_vertices * 3
  const float* normals;           // num_normals * 3
  const float* vertex_colors;     // num_colors * 3 (0..255 as float)
  const int32_t* vertex_indices;  // num_faces * 3
  const int32_t* normal_indices;  // num_faces * 3 or null
  const float* face_normals;      // num_face_normals * 3
  const float* uv;                // num_uv * 2
  const int32_t* uv_indices;      // num_faces * 3 or null
  const int32_t* material_ids;    // num_faces or null
  const uint8_t* const* tex_data; // num_materials pointers, RGB8 row-major
  const int32_t* tex_w;
  const int32_t* tex_h;
};

struct UguRefCamera {
  int32_t width, height;
  float fx, fy, cx, cy;
  double c2w[16];  // row-major 4x4
};

struct UguRefOption {
  int32_t diffuse_color;    // ugu::DiffuseColor
  int32_t interp;           // ugu::ColorInterpolation
  int32_t shading_normal;   // ugu::ShadingNormal
  int32_t diffuse_shading;  // ugu::DiffuseShading
  int32_t backface_culling;
  float depth_scale;
};

}  // extern "C"

namespace {

struct RefHandle {
  std::shared_ptr<ugu::Mesh> mesh;
  std::shared_ptr<ugu::PinholeCamera> camera;
  std::unique_ptr<ugu::Rasterizer> renderer;
  std::unique_ptr<ugu::Raytracer> raytracer;  // created by the first ugu_ref_raytrace (builds the nanort BVH)
  ugu::Image3b color;
  ugu::Image1f depth;
  ugu::Image3f normal;
  ugu::Image1b mask;
  ugu::Image1i face_id;
};

template <typename V>
std::vector<V> ToVec(const typename V::Scalar* p, int n) {
  std::vector<V> out(static_cast<size_t>(n));
  constexpr int D = V::RowsAtCompileTime;
  for (int i = 0; i < n; i++) {
    for (int k = 0; k < D; k++) out[i][k] = p[static_cast<size_t>(i) * D + k];
  }
  return out;
}

Eigen::Affine3d ToAffine(const double* m) {
  Eigen::Affine3d a = Eigen::Affine3d::Identity();
  for (int r = 0; r < 4; r++)
    for (int c = 0; c < 4; c++) a.matrix()(r, c) = m[r * 4 + c];
  return a;
}

double NowMs() {
  return std::chrono::duration<double, std::milli>(
             std::chrono::steady_clock::now().time_since_epoch())
      .count();
}

}  // namespace

extern "C" {

void* ugu_ref_create(const UguRefMeshDesc* d) {
  ugu::set_log_level(ugu::LogLevel::kError);
  auto h = new RefHandle;
  h->mesh = std::make_shared<ugu::Mesh>();
  ugu::Mesh& m = *h->mesh;
  m.set_vertices(ToVec<Eigen::Vector3f>(d->vertices, d->num_vertices));
  if (d->num_normals > 0) m.set_normals(ToVec<Eigen::Vector3f>(d->normals, d->num_normals));
  if (d->num_colors > 0)
    m.set_vertex_colors(ToVec<Eigen::Vector3f>(d->vertex_colors, d->num_colors));
  m.set_vertex_indices(ToVec<Eigen::Vector3i>(d->vertex_indices, d->num_faces));
  if (d->normal_indices)
    m.set_normal_indices(ToVec<Eigen::Vector3i>(d->normal_indices, d->num_faces));
  if (d->num_face_normals > 0)
    m.set_face_normals(ToVec<Eigen::Vector3f>(d->face_normals, d->num_face_normals));
  if (d->num_uv > 0) m.set_uv(ToVec<Eigen::Vector2f>(d->uv, d->num_uv));
  if (d->uv_indices) m.set_uv_indices(ToVec<Eigen::Vector3i>(d->uv_indices, d->num_faces));
  if (d->num_materials > 0) {
    std::vector<ugu::ObjMaterial> mats(d->num_materials);
    for (int i = 0; i < d->num_materials; i++) {
      mats[i].name = "mat" + std::to_string(i);
      if (d->tex_data[i] != nullptr && d->tex_w[i] > 0 && d->tex_h[i] > 0) {
        mats[i].diffuse_tex = ugu::Image3b::zeros(d->tex_h[i], d->tex_w[i]);
        std::memcpy(mats[i].diffuse_tex.data, d->tex_data[i],
                    static_cast<size_t>(d->tex_w[i]) * d->tex_h[i] * 3);
      }
    }
    m.set_materials(mats);
  }
  if (d->material_ids) {
    m.set_material_ids(std::vector<int>(d->material_ids, d->material_ids + d->num_faces));
  }
  h->renderer = std::make_unique<ugu::Rasterizer>();
  h->renderer->set_mesh(h->mesh);
  if (!h->renderer->PrepareMesh()) {
    delete h;
    return nullptr;
  }
  return h;
}

void ugu_ref_destroy(void* handle) { delete static_cast<RefHandle*>(handle); }

// ugu::Raytracer (src/renderer/cpu/raytracer.cc:117-225, BVH = BvhNanort, include/ugu/accel/bvh_nanort.h)
// on the same mesh / camera conventions as ugu_ref_render.  `timing_ms` (optional, 1 double): wall ms of
// set_c2w + Render.
int ugu_ref_raytrace(void* handle, const UguRefCamera* cam, const UguRefOption* opt, uint8_t* color,
                     float* depth, float* normal, uint8_t* mask, int32_t* face_id, double* timing_ms) {
  RefHandle* h = static_cast<RefHandle*>(handle);
  if (h == nullptr) return -1;
  if (!h->raytracer) {
    h->raytracer = std::make_unique<ugu::Raytracer>();
    h->raytracer->set_mesh(h->mesh);
    if (!h->raytracer->PrepareMesh()) {
      h->raytracer.reset();
      return 2;
    }
  }
  ugu::RendererCpuOption o;
  o.diffuse_color = static_cast<ugu::DiffuseColor>(opt->diffuse_color);
  o.interp = static_cast<ugu::ColorInterpolation>(opt->interp);
  o.shading_normal = static_cast<ugu::ShadingNormal>(opt->shading_normal);
  o.diffuse_shading = static_cast<ugu::DiffuseShading>(opt->diffuse_shading);
  o.backface_culling = opt->backface_culling != 0;
  o.depth_scale = opt->depth_scale;
  h->raytracer->set_option(o);
  auto camera = std::make_shared<ugu::PinholeCamera>(cam->width, cam->height, ToAffine(cam->c2w),
                                                     Eigen::Vector2f(cam->cx, cam->cy),
                                                     Eigen::Vector2f(cam->fx, cam->fy));
  h->raytracer->set_camera(camera);
  ugu::Image3b c;
  ugu::Image1f d;
  ugu::Image3f n;
  ugu::Image1b m;
  ugu::Image1i f;
  const double t0 = NowMs();
  camera->set_c2w(ToAffine(cam->c2w));
  const bool ok = h->raytracer->Render(color ? &c : nullptr, depth ? &d : nullptr, normal ? &n : nullptr,
                                       mask ? &m : nullptr, face_id ? &f : nullptr);
  if (timing_ms) timing_ms[0] = NowMs() - t0;
  if (!ok) return 1;
  const size_t P = static_cast<size_t>(cam->width) * cam->height;
  if (color) std::memcpy(color, c.data, P * 3);
  if (depth) std::memcpy(depth, d.data, P * 4);
  if (normal) std::memcpy(normal, n.data, P * 12);
  if (mask) std::memcpy(mask, m.data, P);
  if (face_id) std::memcpy(face_id, f.data, P * 4);
  return 0;
}

// OpenMP threads of the reference from now on (only PinholeCamera::InitRayTable is parallel on the Rasterizer path,
// include/ugu/camera.h:482-513): lets the bench report the OMP_NUM_THREADS=1 figure in the same process.
void ugu_ref_set_num_threads(int n) {
#ifdef _OPENMP
  if (n > 0) omp_set_num_threads(n);
#endif
}

// Renders one view.  Any output pointer may be null.  `timing_ms` (optional, 2 doubles) receives
// [0] wall ms of the whole caller-visible step (set_c2w + Render, i.e. including the lazy ray-table
// rebuild of include/ugu/camera.h:474-513) and [1] the same again for an immediately repeated Render
// with the camera untouched (steady state, ray table warm) when `repeat_warm` != 0.
int ugu_ref_render(void* handle, const UguRefCamera* cam, const UguRefOption* opt, uint8_t* color,
                   float* depth, float* normal, uint8_t* mask, int32_t* face_id, int repeat_warm,
                   double* timing_ms) {
  RefHandle* h = static_cast<RefHandle*>(handle);
  if (h == nullptr) return -1;
  ugu::RendererCpuOption o;
  o.diffuse_color = static_cast<ugu::DiffuseColor>(opt->diffuse_color);
  o.interp = static_cast<ugu::ColorInterpolation>(opt->interp);
  o.shading_normal = static_cast<ugu::ShadingNormal>(opt->shading_normal);
  o.diffuse_shading = static_cast<ugu::DiffuseShading>(opt->diffuse_shading);
  o.backface_culling = opt->backface_culling != 0;
  o.depth_scale = opt->depth_scale;
  h->renderer->set_option(o);

  if (!h->camera || h->camera->width() != cam->width || h->camera->height() != cam->height) {
    h->camera = std::make_shared<ugu::PinholeCamera>(
        cam->width, cam->height, Eigen::Affine3d::Identity(), Eigen::Vector2f(cam->cx, cam->cy),
        Eigen::Vector2f(cam->fx, cam->fy));
    h->renderer->set_camera(h->camera);
  }
  h->camera->set_principal_point(Eigen::Vector2f(cam->cx, cam->cy));
  h->camera->set_focal_length(Eigen::Vector2f(cam->fx, cam->fy));

  ugu::Image3b* pc = color ? &h->color : nullptr;
  ugu::Image1f* pd = depth ? &h->depth : nullptr;
  ugu::Image3f* pn = normal ? &h->normal : nullptr;
  ugu::Image1b* pm = mask ? &h->mask : nullptr;
  ugu::Image1i* pf = face_id ? &h->face_id : nullptr;

  double t0 = NowMs();
  h->camera->set_c2w(ToAffine(cam->c2w));
  bool ok = h->renderer->Render(pc, pd, pn, pm, pf);
  double t1 = NowMs();
  if (timing_ms) timing_ms[0] = t1 - t0;
  if (ok && repeat_warm) {
    double t2 = NowMs();
    ok = h->renderer->Render(pc, pd, pn, pm, pf);
    double t3 = NowMs();
    if (timing_ms) timing_ms[1] = t3 - t2;
  } else if (timing_ms) {
    timing_ms[1] = 0.0;
  }
  if (!ok) return 1;
  const size_t P = static_cast<size_t>(cam->width) * cam->height;
  if (color) std::memcpy(color, h->color.data, P * 3);
  if (depth) std::memcpy(depth, h->depth.data, P * 4);
  if (normal) std::memcpy(normal, h->normal.data, P * 12);
  if (mask) std::memcpy(mask, h->mask.data, P);
  if (face_id) std::memcpy(face_id, h->face_id.data, P * 4);
  return 0;
}

// The float camera constants exactly as the reference derives them
// (src/renderer/cpu/rasterizer.cc:118-119, include/ugu/camera.h:317-331).  Row-major 3x3.
int ugu_ref_camera_constants(const double* c2w16, float* w2c_R, float* w2c_t, float* c2w_R) {
  ugu::PinholeCamera cam(4, 4, ToAffine(c2w16), 45.0f);
  const Eigen::Matrix3f R = cam.w2c().rotation().cast<float>();
  const Eigen::Vector3f t = cam.w2c().translation().cast<float>();
  const Eigen::Matrix3f C = cam.c2w().matrix().block<3, 3>(0, 0).cast<float>();
  for (int r = 0; r < 3; r++) {
    for (int c = 0; c < 3; c++) {
      w2c_R[r * 3 + c] = R(r, c);
      c2w_R[r * 3 + c] = C(r, c);
    }
    w2c_t[r] = t[r];
  }
  return 0;
}

// ugu::c2w look-at (include/ugu/util/camera_util.h:26-34) so fixtures can use the very poses of
// examples/ex02_renderer.cc:49-102 and ex09_renderer_realtime.cc:137-142.  Row-major 4x4 out.
int ugu_ref_lookat_c2w(const float* eye, const float* center, const float* up, double* c2w16) {
  Eigen::Matrix4f m;
  ugu::c2w(Eigen::Vector3f(eye[0], eye[1], eye[2]), Eigen::Vector3f(center[0], center[1], center[2]),
           Eigen::Vector3f(up[0], up[1], up[2]), &m);
  for (int r = 0; r < 4; r++)
    for (int c = 0; c < 4; c++) c2w16[r * 4 + c] = static_cast<double>(m(r, c));
  return 0;
}

// ugu::Depth2PointCloud(const Image1f&, const Camera&, Image3f*, bool gl_coord)
// (src/util/rgbd_util.cc:472-513), the step ex02 runs on the rendered depth
// (examples/ex02_renderer.cc:156-158).  `pc` (W*H*3 floats) is IN/OUT: the reference only writes
// pixels whose depth is >= FLT_MIN and leaves the others as they were.
int ugu_ref_depth2pointcloud(const UguRefCamera* cam, const float* depth, float* pc, int gl_coord) {
  ugu::PinholeCamera camera(cam->width, cam->height, ToAffine(cam->c2w),
                            Eigen::Vector2f(cam->cx, cam->cy), Eigen::Vector2f(cam->fx, cam->fy));
  const size_t P = static_cast<size_t>(cam->width) * cam->height;
  ugu::Image1f d;
  ugu::Image3f out;
  ugu::Init(&d, cam->width, cam->height, 0.0f);
  ugu::Init(&out, cam->width, cam->height, 0.0f);
  std::memcpy(d.data, depth, P * 4);
  std::memcpy(out.data, pc, P * 12);
  if (!ugu::Depth2PointCloud(d, camera, &out, gl_coord != 0)) return 1;
  std::memcpy(pc, out.data, P * 12);
  return 0;
}

// ugu::RasterizeVertexAttributeToTexture / RasterizeFaceAttributeToTexture
// (include/ugu/util/raster_util.h:139-268) with num_threads = 1 — the reference's only deterministic
// order (faces in ascending index, last writer wins; with more threads it is racy).
//   per_face     0: attrs indexed through attr_indices (vertex attributes); 1: attrs[f] per face
//   texture_type 0: Image3b, 1: Image3f
//   keep_texture 1: the reference is called with width = height = -1 and rasterises over the
//                existing contents of `texture` (raster_util.h:151-155); 0: zero-initialised
//   mask         W*H u8 in/out (the reference never initialises it), or null
static int g_uv_threads = 1;
// threads the reference's parallel_for uses in the UV-raster calls below (1 = the deterministic
// order the parity tests pin; -1 = the reference's default, all hardware threads, racy on overlaps)
void ugu_ref_uv_set_threads(int n) { g_uv_threads = n; }

int ugu_ref_uv_raster(int32_t num_attrs, int32_t num_faces, int32_t num_uv, int32_t per_face,
                      const float* attrs, const int32_t* attr_indices, const float* uv,
                      const int32_t* uv_indices, int32_t width, int32_t height, int32_t texture_type,
                      int32_t keep_texture, void* texture, uint8_t* mask) {
  std::vector<Eigen::Vector3f> a(num_attrs);
  for (int i = 0; i < num_attrs; i++) a[i] = Eigen::Vector3f(attrs[i * 3], attrs[i * 3 + 1], attrs[i * 3 + 2]);
  std::vector<Eigen::Vector3i> ai(per_face ? 0 : num_faces), ui(num_faces);
  for (int i = 0; i < num_faces; i++) {
    if (!per_face) ai[i] = Eigen::Vector3i(attr_indices[i * 3], attr_indices[i * 3 + 1], attr_indices[i * 3 + 2]);
    ui[i] = Eigen::Vector3i(uv_indices[i * 3], uv_indices[i * 3 + 1], uv_indices[i * 3 + 2]);
  }
  std::vector<Eigen::Vector2f> u(num_uv);
  for (int i = 0; i < num_uv; i++) u[i] = Eigen::Vector2f(uv[i * 2], uv[i * 2 + 1]);
  const size_t P = static_cast<size_t>(width) * height;
  ugu::Image1b m;
  if (mask) {
    ugu::Init(&m, width, height, static_cast<unsigned char>(0));
    std::memcpy(m.data, mask, P);
  }
  const int w = keep_texture ? -1 : width, h = keep_texture ? -1 : height;
  bool ok;
  if (texture_type == 0) {
    ugu::Image3b t = ugu::Image3b::zeros(height, width);
    if (keep_texture) std::memcpy(t.data, texture, P * 3);
    ok = per_face ? ugu::RasterizeFaceAttributeToTexture(a, u, ui, t, w, h, mask ? &m : nullptr, g_uv_threads)
                  : ugu::RasterizeVertexAttributeToTexture(a, ai, u, ui, t, w, h, mask ? &m : nullptr, g_uv_threads);
    if (ok) std::memcpy(texture, t.data, P * 3);
  } else {
    ugu::Image3f t = ugu::Image3f::zeros(height, width);
    if (keep_texture) std::memcpy(t.data, texture, P * 12);
    ok = per_face ? ugu::RasterizeFaceAttributeToTexture(a, u, ui, t, w, h, mask ? &m : nullptr, g_uv_threads)
                  : ugu::RasterizeVertexAttributeToTexture(a, ai, u, ui, t, w, h, mask ? &m : nullptr, g_uv_threads);
    if (ok) std::memcpy(texture, t.data, P * 12);
  }
  if (ok && mask) std::memcpy(mask, m.data, P);
  return ok ? 0 : 1;
}

// ugu::GenerateUvMask(uvs, uv_indices, Image1b&, uint8_t color, w, h, num_threads)
// (src/util/raster_util.cc:9-27)
int ugu_ref_uv_mask(int32_t num_faces, int32_t num_uv, const float* uv, const int32_t* uv_indices,
                    int32_t width, int32_t height, int32_t color, uint8_t* mask) {
  std::vector<Eigen::Vector3i> ui(num_faces);
  for (int i = 0; i < num_faces; i++)
    ui[i] = Eigen::Vector3i(uv_indices[i * 3], uv_indices[i * 3 + 1], uv_indices[i * 3 + 2]);
  std::vector<Eigen::Vector2f> u(num_uv);
  for (int i = 0; i < num_uv; i++) u[i] = Eigen::Vector2f(uv[i * 2], uv[i * 2 + 1]);
  ugu::Image1b m;
  if (!ugu::GenerateUvMask(u, ui, m, static_cast<uint8_t>(color), width, height, 1)) return 1;
  std::memcpy(mask, m.data, static_cast<size_t>(width) * height);
  return 0;
}

// ugu::Depth2Mesh (include/ugu/util/rgbd_util.h:27-37, src/util/rgbd_util.cc:94-282,528-545): the step
// examples/ex02_renderer.cc:156-158 runs on the rendered depth (and colour).  `color` null -> the
// depth-only overload; else the textured overload (uv + material; vertex colours when asked).
// The result is held until the next call; out[0..5] = Nv, Nf, has_uv, has_color, W, H.
static ugu::Mesh g_d2m_mesh;
static ugu::Image3f g_d2m_pc, g_d2m_nrm;
int ugu_ref_depth2mesh(const UguRefCamera* cam, const float* depth, const uint8_t* color, int32_t cw,
                       int32_t ch, int32_t with_vertex_color, float max_connect_z_diff, int32_t x_step,
                       int32_t y_step, int32_t gl_coord, int32_t* out) {
  ugu::set_log_level(ugu::LogLevel::kError);
  ugu::PinholeCamera camera(cam->width, cam->height, ToAffine(cam->c2w),
                            Eigen::Vector2f(cam->cx, cam->cy), Eigen::Vector2f(cam->fx, cam->fy));
  const size_t P = static_cast<size_t>(cam->width) * cam->height;
  ugu::Image1f d;
  ugu::Init(&d, cam->width, cam->height, 0.0f);
  std::memcpy(d.data, depth, P * 4);
  g_d2m_mesh.Clear();
  bool ok;
  if (color == nullptr) {
    ok = ugu::Depth2Mesh(d, camera, &g_d2m_mesh, max_connect_z_diff, x_step, y_step, gl_coord != 0,
                         &g_d2m_pc, &g_d2m_nrm);
  } else {
    ugu::Image3b c = ugu::Image3b::zeros(ch, cw);
    std::memcpy(c.data, color, static_cast<size_t>(cw) * ch * 3);
    ok = ugu::Depth2Mesh(d, c, camera, &g_d2m_mesh, max_connect_z_diff, x_step, y_step, gl_coord != 0,
                         "Depth2Mesh_mat", with_vertex_color != 0, &g_d2m_pc, &g_d2m_nrm);
  }
  if (!ok) return 1;
  out[0] = static_cast<int32_t>(g_d2m_mesh.vertices().size());
  out[1] = static_cast<int32_t>(g_d2m_mesh.vertex_indices().size());
  out[2] = g_d2m_mesh.uv().empty() ? 0 : 1;
  out[3] = g_d2m_mesh.vertex_colors().empty() ? 0 : 1;
  out[4] = g_d2m_pc.cols;
  out[5] = g_d2m_pc.rows;
  return 0;
}

// The two-camera overload (rgbd_util.h:39-48; Depth2MeshImpl, rgbd_util.cc:283-466): uv / vertex colours from
// the projection of depth2color * p into the colour camera.  d2c16 = row-major 4x4 float (an Eigen::Affine3f).
// The organized images are fresh (the overload only Init()s them when the size differs).
int ugu_ref_depth2mesh2(const UguRefCamera* dcam, const UguRefCamera* ccam, const float* d2c16, const float* depth,
                        const uint8_t* color, int32_t with_vertex_color, float max_connect_z_diff, int32_t x_step,
                        int32_t y_step, int32_t gl_coord, int32_t* out) {
  ugu::set_log_level(ugu::LogLevel::kError);
  ugu::PinholeCamera depth_camera(dcam->width, dcam->height, ToAffine(dcam->c2w),
                                  Eigen::Vector2f(dcam->cx, dcam->cy), Eigen::Vector2f(dcam->fx, dcam->fy));
  ugu::PinholeCamera color_camera(ccam->width, ccam->height, ToAffine(ccam->c2w),
                                  Eigen::Vector2f(ccam->cx, ccam->cy), Eigen::Vector2f(ccam->fx, ccam->fy));
  Eigen::Affine3f d2c = Eigen::Affine3f::Identity();
  for (int i = 0; i < 4; i++)
    for (int j = 0; j < 4; j++) d2c.matrix()(i, j) = d2c16[i * 4 + j];
  const size_t P = static_cast<size_t>(dcam->width) * dcam->height;
  ugu::Image1f d;
  ugu::Init(&d, dcam->width, dcam->height, 0.0f);
  std::memcpy(d.data, depth, P * 4);
  ugu::Image3b c = ugu::Image3b::zeros(ccam->height, ccam->width);
  std::memcpy(c.data, color, static_cast<size_t>(ccam->width) * ccam->height * 3);
  g_d2m_mesh.Clear();
  g_d2m_pc = ugu::Image3f();
  g_d2m_nrm = ugu::Image3f();
  const bool ok = ugu::Depth2Mesh(d, c, depth_camera, color_camera, d2c, &g_d2m_mesh, max_connect_z_diff, x_step,
                                  y_step, gl_coord != 0, "Depth2Mesh_mat", with_vertex_color != 0, &g_d2m_pc,
                                  &g_d2m_nrm);
  if (!ok) return 1;
  out[0] = static_cast<int32_t>(g_d2m_mesh.vertices().size());
  out[1] = static_cast<int32_t>(g_d2m_mesh.vertex_indices().size());
  out[2] = g_d2m_mesh.uv().empty() ? 0 : 1;
  out[3] = g_d2m_mesh.vertex_colors().empty() ? 0 : 1;
  out[4] = g_d2m_pc.cols;
  out[5] = g_d2m_pc.rows;
  return 0;
}

// Copies the held result (any pointer may be null): vertices / normals Nv*3, faces Nf*3 (== the
// mesh's normal_indices and uv_indices), face normals Nf*3, uv Nv*2, colours Nv*3, the organized
// point-cloud and normal images W*H*3.  Returns 1 if an index array differs from vertex_indices.
int ugu_ref_depth2mesh_copy(float* vertices, int32_t* faces, float* normals, float* face_normals,
                            float* uv, float* colors, float* point_cloud, float* normal_image) {
  const ugu::Mesh& m = g_d2m_mesh;
  const size_t nv = m.vertices().size(), nf = m.vertex_indices().size();
  for (size_t i = 0; i < nv; i++)
    for (int k = 0; k < 3; k++) {
      if (vertices) vertices[i * 3 + k] = m.vertices()[i][k];
      if (normals && i < m.normals().size()) normals[i * 3 + k] = m.normals()[i][k];
      if (colors && i < m.vertex_colors().size()) colors[i * 3 + k] = m.vertex_colors()[i][k];
    }
  if (uv)
    for (size_t i = 0; i < m.uv().size(); i++) {
      uv[i * 2] = m.uv()[i][0];
      uv[i * 2 + 1] = m.uv()[i][1];
    }
  int bad = 0;
  for (size_t i = 0; i < nf; i++)
    for (int k = 0; k < 3; k++) {
      if (faces) faces[i * 3 + k] = m.vertex_indices()[i][k];
      if (face_normals && i < m.face_normals().size()) face_normals[i * 3 + k] = m.face_normals()[i][k];
      if (i < m.normal_indices().size() && m.normal_indices()[i][k] != m.vertex_indices()[i][k]) bad = 1;
      if (!m.uv().empty() && (i >= m.uv_indices().size() || m.uv_indices()[i][k] != m.vertex_indices()[i][k])) bad = 1;
    }
  if (nf > 0 && m.normal_indices().size() != nf) bad = 1;
  if (!m.uv().empty() && m.material_ids().size() != nf) bad = 1;
  const size_t P = static_cast<size_t>(g_d2m_pc.cols) * g_d2m_pc.rows;
  if (point_cloud) std::memcpy(point_cloud, g_d2m_pc.data, P * 12);
  if (normal_image) std::memcpy(normal_image, g_d2m_nrm.data, P * 12);
  return bad;
}

// --- OBJ loading through the reference (fixture generation only) -------------------------------
void* ugu_ref_mesh_load_obj(const char* obj_path, const char* mtl_dir) {
  ugu::set_log_level(ugu::LogLevel::kError);
  auto m = new ugu::Mesh;
  if (!m->LoadObj(obj_path, mtl_dir)) {
    delete m;
    return nullptr;
  }
  return m;
}

void ugu_ref_mesh_free(void* mesh) { delete static_cast<ugu::Mesh*>(mesh); }

// out[0..8] = V, Nn, Nc, F, Nfn, Nuv, M, tex_w(0), tex_h(0)
void ugu_ref_mesh_sizes(void* mesh, int32_t* out) {
  const ugu::Mesh& m = *static_cast<ugu::Mesh*>(mesh);
  out[0] = static_cast<int32_t>(m.vertices().size());
  out[1] = static_cast<int32_t>(m.normals().size());
  out[2] = static_cast<int32_t>(m.vertex_colors().size());
  out[3] = static_cast<int32_t>(m.vertex_indices().size());
  out[4] = static_cast<int32_t>(m.face_normals().size());
  out[5] = static_cast<int32_t>(m.uv().size());
  out[6] = static_cast<int32_t>(m.materials().size());
  out[7] = m.materials().empty() ? 0 : m.materials()[0].diffuse_tex.cols;
  out[8] = m.materials().empty() ? 0 : m.materials()[0].diffuse_tex.rows;
}

void ugu_ref_mesh_copy(void* mesh, float* vertices, float* normals, float* colors,
                       int32_t* vertex_indices, int32_t* normal_indices, float* face_normals,
                       float* uv, int32_t* uv_indices, int32_t* material_ids, uint8_t* tex0) {
  const ugu::Mesh& m = *static_cast<ugu::Mesh*>(mesh);
  auto cp3f = [](const std::vector<Eigen::Vector3f>& v, float* o) {
    if (!o) return;
    for (size_t i = 0; i < v.size(); i++)
      for (int k = 0; k < 3; k++) o[i * 3 + k] = v[i][k];
  };
  auto cp3i = [](const std::vector<Eigen::Vector3i>& v, int32_t* o) {
    if (!o) return;
    for (size_t i = 0; i < v.size(); i++)
      for (int k = 0; k < 3; k++) o[i * 3 + k] = v[i][k];
  };
  cp3f(m.vertices(), vertices);
  cp3f(m.normals(), normals);
  cp3f(m.vertex_colors(), colors);
  cp3i(m.vertex_indices(), vertex_indices);
  cp3i(m.normal_indices(), normal_indices);
  cp3f(m.face_normals(), face_normals);
  if (uv)
    for (size_t i = 0; i < m.uv().size(); i++) {
      uv[i * 2] = m.uv()[i][0];
      uv[i * 2 + 1] = m.uv()[i][1];
    }
  cp3i(m.uv_indices(), uv_indices);
  if (material_ids)
    for (size_t i = 0; i < m.material_ids().size(); i++) material_ids[i] = m.material_ids()[i];
  if (tex0 && !m.materials().empty()) {
    const ugu::Image3b& t = m.materials()[0].diffuse_tex;
    std::memcpy(tex0, t.data, static_cast<size_t>(t.rows) * t.cols * 3);
  }
}


// ---- ugu::VisibilityTester (include/ugu/texturing/visibility_tester.h:216-265,
// src/texturing/visibility_tester.cc:416-733): the reference's per-keyframe vertex / face visibility pass
// (BVH rays) and the statistics of VisibilityInfo.  The keyframes carry HOST images (colour, depth, mask)
// and a pinhole camera; the tester is driven through its public API only (so, like every reference caller,
// only the seven option fields VisibilityTesterOption::CopyTo copies take effect).  `num_threads` = 1 pins
// the order of the face-wise mode, whose `v_succeeded` bookkeeping is racy under OpenMP
// (visibility_tester.cc:608-629).
struct UguRefKeyframe {
  int32_t id;
  int32_t pad_;
  UguRefCamera cam;
  const uint8_t* color;  // H*W*3
  const float* depth;    // H*W or null
  const uint8_t* mask;   // H*W or null
};
struct UguRefVisOption {
  float backface_culling_angle_rad_th;
  int32_t use_mask, use_depth, interp;
  int32_t facewise;
  int32_t num_threads;  // <= 0: leave OpenMP alone
};

struct RefVisHandle {
  ugu::VisibilityInfo info;
  std::vector<int> kf_ids;
  int V = 0, F = 0;
};

void* ugu_ref_visibility_run(int32_t V, const float* vertices, const float* normals, int32_t F,
                             const int32_t* indices, int32_t K, const UguRefKeyframe* kfs,
                             const UguRefVisOption* o) {
  ugu::set_log_level(ugu::LogLevel::kError);
#ifdef _OPENMP
  if (o->num_threads > 0) omp_set_num_threads(o->num_threads);
#endif
  ugu::VisibilityTesterOption opt;
  opt.backface_culling_angle_rad_th = o->backface_culling_angle_rad_th;
  opt.use_mask = o->use_mask != 0;
  opt.use_depth = o->use_depth != 0;
  opt.interp = static_cast<ugu::ColorInterpolation>(o->interp);
  ugu::VisibilityTester tester(opt);
  tester.set_data(ToVec<Eigen::Vector3f>(vertices, V), ToVec<Eigen::Vector3f>(normals, V),
                  ToVec<Eigen::Vector3i>(indices, F));
  if (!tester.PrepareData()) return nullptr;
  std::vector<std::shared_ptr<ugu::Keyframe>> keyframes;
  auto h = new RefVisHandle;
  h->V = V;
  h->F = F;
  for (int k = 0; k < K; k++) {
    auto kf = ugu::Keyframe::Create();
    const UguRefKeyframe& s = kfs[k];
    kf->id = s.id;
    h->kf_ids.push_back(s.id);
    const int W = s.cam.width, H = s.cam.height;
    kf->color = ugu::Image3b::zeros(H, W);
    std::memcpy(kf->color.data, s.color, static_cast<size_t>(W) * H * 3);
    if (s.depth) {
      kf->depth = ugu::Image1f::zeros(H, W);
      std::memcpy(kf->depth.data, s.depth, static_cast<size_t>(W) * H * 4);
    }
    if (s.mask) {
      kf->mask = ugu::Image1b::zeros(H, W);
      std::memcpy(kf->mask.data, s.mask, static_cast<size_t>(W) * H);
    }
    kf->camera = std::make_shared<ugu::PinholeCamera>(W, H, ToAffine(s.cam.c2w), Eigen::Vector2f(s.cam.cx, s.cam.cy),
                                                      Eigen::Vector2f(s.cam.fx, s.cam.fy));
    keyframes.push_back(kf);
  }
  h->info.vertex_info_list.resize(V);
  h->info.face_info_list.resize(F);
  if (!tester.Test(keyframes, &h->info, o->facewise != 0)) {
    delete h;
    return nullptr;
  }
  return h;
}

void ugu_ref_visibility_free(void* handle) { delete static_cast<RefVisHandle*>(handle); }

// Dense per-(keyframe, vertex) records, K*V each: visible (u8), colour (3 f32), intensity, projected_pos (2),
// viewing_angle, distance.  Rows of keyframes in which the vertex is not visible are left untouched.
void ugu_ref_visibility_vertex_records(void* handle, uint8_t* visible, float* color, float* intensity,
                                       float* projected, float* viewing_angle, float* distance) {
  auto* h = static_cast<RefVisHandle*>(handle);
  const size_t V = h->V;
  for (size_t v = 0; v < V; v++) {
    for (const auto& r : h->info.vertex_info_list[v].visible_keyframes) {
      size_t k = 0;
      while (k < h->kf_ids.size() && h->kf_ids[k] != r.kf_id) k++;
      const size_t i = k * V + v;
      visible[i] = 1;
      for (int c = 0; c < 3; c++) color[i * 3 + c] = r.color[c];
      intensity[i] = r.intensity;
      projected[i * 2 + 0] = r.projected_pos[0];
      projected[i * 2 + 1] = r.projected_pos[1];
      viewing_angle[i] = r.viewing_angle;
      distance[i] = r.distance;
    }
  }
}

// Dense per-(keyframe, face) records, K*F each: visible (u8), area, viewing_angle, distance, projected_tri (6).
void ugu_ref_visibility_face_records(void* handle, uint8_t* visible, float* area, float* viewing_angle,
                                     float* distance, float* projected_tri) {
  auto* h = static_cast<RefVisHandle*>(handle);
  const size_t F = h->F;
  for (size_t f = 0; f < F; f++) {
    for (const auto& r : h->info.face_info_list[f].visible_keyframes) {
      size_t k = 0;
      while (k < h->kf_ids.size() && h->kf_ids[k] != r.kf_id) k++;
      const size_t i = k * F + f;
      visible[i] = 1;
      area[i] = r.area;
      viewing_angle[i] = r.viewing_angle;
      distance[i] = r.distance;
      for (int j = 0; j < 3; j++) {
        projected_tri[i * 6 + 2 * j] = r.projected_tri[j][0];
        projected_tri[i * 6 + 2 * j + 1] = r.projected_tri[j][1];
      }
    }
  }
}

// FaceInfo::CalcStat (visibility_tester.cc:185-218): F x {max_area_id, min_viewing_angle_id, min_distance_id}
// (-1 where the face was never visible), and VertexInfo::CalcStat (visibility_tester.cc:66-152): per vertex
// ids[2] = {min_viewing_angle_id, min_distance_id}, scalars[7] = {mean_viewing_angle, median_viewing_angle,
// mean_distance, median_distance, min_intensity, median_intensity, number of visible keyframes}, colors[10][3] =
// {mean, median, min_viewing_angle, mean_viewing_angle, median_viewing_angle, min_distance, mean_distance,
//  median_distance, median_intensity, min_intensity}_color.
void ugu_ref_visibility_stats(void* handle, int32_t* face_ids, int32_t* vertex_ids, float* vertex_scalars,
                              float* vertex_colors) {
  auto* h = static_cast<RefVisHandle*>(handle);
  for (int f = 0; f < h->F; f++) {
    const auto& fi = h->info.face_info_list[f];
    face_ids[f * 3 + 0] = fi.max_area_id;
    face_ids[f * 3 + 1] = fi.min_viewing_angle_id;
    face_ids[f * 3 + 2] = fi.min_distance_id;
  }
  for (int v = 0; v < h->V; v++) {
    const auto& vi = h->info.vertex_info_list[v];
    vertex_ids[v * 2 + 0] = vi.min_viewing_angle_id;
    vertex_ids[v * 2 + 1] = vi.min_distance_id;
    float* s = vertex_scalars + static_cast<size_t>(v) * 7;
    s[0] = vi.mean_viewing_angle;
    s[1] = vi.median_viewing_angle;
    s[2] = vi.mean_distance;
    s[3] = vi.median_distance;
    s[4] = vi.min_intensity;
    s[5] = vi.median_intensity;
    s[6] = static_cast<float>(vi.visible_keyframes.size());
    const Eigen::Vector3f* cs[10] = {&vi.mean_color, &vi.median_color, &vi.min_viewing_angle_color,
                                     &vi.mean_viewing_angle_color, &vi.median_viewing_angle_color,
                                     &vi.min_distance_color, &vi.mean_distance_color, &vi.median_distance_color,
                                     &vi.median_intensity_color, &vi.min_intensity_color};
    const bool any = !vi.visible_keyframes.empty();
    for (int c = 0; c < 10; c++)
      for (int k = 0; k < 3; k++)
        // the last two are never initialised by the reference for vertices without a visible keyframe
        vertex_colors[(static_cast<size_t>(v) * 10 + c) * 3 + k] = (any || c < 8) ? (*cs[c])[k] : 0.0f;
  }
}

// The float camera constants the tester reads: w2c().cast<float>() as a row-major 3x4, c2w_R_f_ (3x3) and
// c2w_t_f_ (camera.h:317-331).
void ugu_ref_visibility_camera_constants(const UguRefCamera* cam, float* w2c34, float* c2w_R, float* c2w_t) {
  ugu::PinholeCamera c(cam->width, cam->height, ToAffine(cam->c2w), Eigen::Vector2f(cam->cx, cam->cy),
                       Eigen::Vector2f(cam->fx, cam->fy));
  const Eigen::Affine3f w2c = c.w2c().cast<float>();
  for (int r = 0; r < 3; r++)
    for (int k = 0; k < 4; k++) w2c34[r * 4 + k] = w2c.matrix()(r, k);
  const Eigen::Affine3f c2w = c.c2w().cast<float>();
  for (int r = 0; r < 3; r++) {
    for (int k = 0; k < 3; k++) c2w_R[r * 3 + k] = c2w.matrix()(r, k);
    c2w_t[r] = c2w.matrix()(r, 3);
  }
}

}  // extern "C"
