// ugu::RasterizerCuda: host C++ over the C-ABI (include/ugu_b200_raster.h).  Compiled against the
// reference headers; mirrors the state machine and error behaviour of Rasterizer::Impl
// (src/renderer/cpu/rasterizer.cc:35-103,309-404) and leaves every computation to the CUDA library.
#include "rasterizer_cuda.h"

#include <cmath>
#include <cstring>
#include <limits>
#include <vector>

#include "ugu/log.h"
#include "ugu/util/raster_util.h"
#include "ugu_b200_raster.h"

namespace ugu {

class RasterizerCuda::Impl {
 public:
  ugu_b200_rasterizer* handle_ = nullptr;
  bool mesh_initialized_ = false;
  std::shared_ptr<const Camera> camera_;
  std::shared_ptr<const Mesh> mesh_;
  RendererCpuOption option_;
  bool raytrace_ = false;  // RaytracerCuda: Render answers like ugu::Raytracer

  explicit Impl(int device) {
    if (ugu_b200_rasterizer_create(device, &handle_) != UGU_B200_OK) {
      LOGE("RasterizerCuda: %s\n", ugu_b200_last_error(nullptr));
      handle_ = nullptr;
    }
  }
  ~Impl() { ugu_b200_rasterizer_destroy(handle_); }

  bool PrepareMesh() {
    if (mesh_ == nullptr) {
      LOGE("mesh has not been set\n");
      return false;
    }
    if (raytrace_ && (mesh_->vertices().empty() || mesh_->vertex_indices().empty())) {
      LOGE("mesh is empty\n");  // Raytracer::Impl::PrepareMesh, raytracer.cc:96-99
      return false;
    }
    if (handle_ == nullptr) {
      LOGE("RasterizerCuda: no CUDA rasterizer (no CPU fallback)\n");
      return false;
    }
    const Mesh& m = *mesh_;
    const size_t F = m.vertex_indices().size();
    ugu_b200_mesh_desc d;
    std::memset(&d, 0, sizeof(d));
    d.num_vertices = static_cast<int32_t>(m.vertices().size());
    d.num_normals = static_cast<int32_t>(m.normals().size());
    d.num_colors = static_cast<int32_t>(m.vertex_colors().size());
    d.num_faces = static_cast<int32_t>(F);
    d.num_face_normals = static_cast<int32_t>(m.face_normals().size());
    d.num_uv = static_cast<int32_t>(m.uv().size());
    d.num_materials = static_cast<int32_t>(m.materials().size());
    // std::vector<Eigen::Vector3f/3i/2f> is a dense AoS array (no padding: 12 / 12 / 8 bytes)
    static_assert(sizeof(Eigen::Vector3f) == 12 && sizeof(Eigen::Vector3i) == 12 &&
                      sizeof(Eigen::Vector2f) == 8,
                  "unexpected Eigen fixed-vector layout");
    d.vertices = d.num_vertices ? m.vertices()[0].data() : nullptr;
    d.normals = d.num_normals ? m.normals()[0].data() : nullptr;
    d.vertex_colors = d.num_colors ? m.vertex_colors()[0].data() : nullptr;
    d.vertex_indices = F ? m.vertex_indices()[0].data() : nullptr;
    d.normal_indices = (F && m.normal_indices().size() == F) ? m.normal_indices()[0].data() : nullptr;
    d.face_normals = d.num_face_normals ? m.face_normals()[0].data() : nullptr;
    d.uv = d.num_uv ? m.uv()[0].data() : nullptr;
    d.uv_indices = (F && m.uv_indices().size() == F) ? m.uv_indices()[0].data() : nullptr;
    d.material_ids = (F && m.material_ids().size() == F) ? m.material_ids().data() : nullptr;
    std::vector<const uint8_t*> tex(m.materials().size(), nullptr);
    std::vector<int32_t> tw(m.materials().size(), 0), th(m.materials().size(), 0);
    for (size_t i = 0; i < m.materials().size(); i++) {
      const Image3b& t = m.materials()[i].diffuse_tex;
      if (!t.empty()) {
        tex[i] = t.data;
        tw[i] = t.cols;
        th[i] = t.rows;
      }
    }
    d.tex_data = tex.data();
    d.tex_w = tw.data();
    d.tex_h = th.data();
    if (ugu_b200_rasterizer_upload_mesh(handle_, &d) != UGU_B200_OK) {
      LOGE("%s\n", ugu_b200_last_error(handle_));
      return false;
    }
    mesh_initialized_ = true;
    return true;
  }

  // Preconditions shared by Render and RenderGBuffer, in the order of ValidateAndInitBeforeRender
  // (util_private.cc:19-67; the mesh/option dependent ones are checked by the library against the
  // uploaded mesh), then the option and camera PODs of the C-ABI.
  bool Begin(ugu_b200_camera* cam) const {
    if (camera_ == nullptr) {
      LOGE("camera has not been set\n");
      return false;
    }
    if (!mesh_initialized_) {
      LOGE("mesh has not been initialized\n");
      return false;
    }
    const PinholeCamera* pin = dynamic_cast<const PinholeCamera*>(camera_.get());
    if (pin == nullptr || dynamic_cast<const OpenCvCamera*>(camera_.get()) != nullptr) {
      LOGE("RasterizerCuda supports PinholeCamera only (no CPU fallback)\n");
      return false;
    }
    ugu_b200_option o;
    o.diffuse_color = static_cast<int32_t>(option_.diffuse_color);
    o.interp = static_cast<int32_t>(option_.interp);
    o.shading_normal = static_cast<int32_t>(option_.shading_normal);
    o.diffuse_shading = static_cast<int32_t>(option_.diffuse_shading);
    o.backface_culling = option_.backface_culling ? 1 : 0;
    o.depth_scale = option_.depth_scale;
    o.oren_nayar_sigma = option_.oren_nayar_sigma;
    ugu_b200_rasterizer_set_option(handle_, &o);

    // Camera constants exactly as rasterizer.cc:118-119 and camera.h:321 derive them.
    ugu_b200_camera& c = *cam;
    c.width = camera_->width();
    c.height = camera_->height();
    c.fx = pin->focal_length().x();
    c.fy = pin->focal_length().y();
    c.cx = pin->principal_point().x();
    c.cy = pin->principal_point().y();
    const Eigen::Matrix3f w2c_R = camera_->w2c().rotation().cast<float>();
    const Eigen::Vector3f w2c_t = camera_->w2c().translation().cast<float>();
    const Eigen::Matrix3f c2w_R = camera_->c2w().matrix().block<3, 3>(0, 0).cast<float>();
    for (int i = 0; i < 3; i++) {
      for (int j = 0; j < 3; j++) {
        c.w2c_R[i * 3 + j] = w2c_R(i, j);
        c.c2w_R[i * 3 + j] = c2w_R(i, j);
      }
      c.w2c_t[i] = w2c_t[i];
    }
    return true;
  }

  // ugu::GBuffer as RendererGl::ReadGbuf fills it (src/renderer/gl/renderer.cc:501-650), from this
  // rasterizer's winners.  Like ReadGbuf (:515-518) the images are (re)allocated when the size differs.
  bool RenderGBuffer(GBuffer* g, float z_near, float z_far, bool flat_normal, int geo_id) const {
    if (g == nullptr) {
      LOGE("gbuffer is nullptr\n");
      return false;
    }
    ugu_b200_camera c;
    if (!Begin(&c)) return false;
    if (g->color.cols != c.width || g->color.rows != c.height) g->Init(c.width, c.height);
    ugu_b200_gbuffer out;
    out.color = g->color.data;
    out.normal_cam = reinterpret_cast<float*>(g->normal_cam.data);
    out.normal_wld = reinterpret_cast<float*>(g->normal_wld.data);
    out.pos_cam = reinterpret_cast<float*>(g->pos_cam.data);
    out.pos_wld = reinterpret_cast<float*>(g->pos_wld.data);
    out.depth_01 = reinterpret_cast<float*>(g->depth_01.data);
    out.stencil = g->stencil.data;
    out.face_id = reinterpret_cast<int32_t*>(g->face_id.data);
    out.uv = reinterpret_cast<float*>(g->uv.data);
    out.bary = reinterpret_cast<float*>(g->bary.data);
    out.geo_id = reinterpret_cast<int32_t*>(g->geo_id.data);
    ugu_b200_gbuffer_option go;
    go.z_near = z_near;
    go.z_far = z_far;
    go.flat_normal = flat_normal ? 1 : 0;
    go.geo_id = geo_id;
    if (ugu_b200_rasterizer_render_gbuffer(handle_, &c, &go, &out) != UGU_B200_OK) {
      LOGE("%s\n", ugu_b200_last_error(handle_));
      return false;
    }
    g->shaded.setTo(0.0);  // the viewer's deferred lighting pass is not produced
    return true;
  }

  bool Render(Image3b* color, Image1f* depth, Image3f* normal, Image1b* mask,
              Image1i* face_id) const {
    ugu_b200_camera c;
    if (!Begin(&c)) return false;
    if (color == nullptr && depth == nullptr && normal == nullptr && mask == nullptr &&
        face_id == nullptr) {
      LOGE("all arguments are nullptr. nothing to do\n");
      return false;
    }
    // ugu::Init: (re)allocate only when the size differs (image.h:963-985).  The library writes
    // every pixel of every requested image, so the clear is part of the kernel.
    const int w = c.width, h = c.height;
    if (color != nullptr && (color->cols != w || color->rows != h)) *color = Image3b::zeros(h, w);
    if (depth != nullptr && (depth->cols != w || depth->rows != h)) *depth = Image1f::zeros(h, w);
    if (normal != nullptr && (normal->cols != w || normal->rows != h)) *normal = Image3f::zeros(h, w);
    if (mask != nullptr && (mask->cols != w || mask->rows != h)) *mask = Image1b::zeros(h, w);
    if (face_id != nullptr && (face_id->cols != w || face_id->rows != h)) *face_id = Image1i::zeros(h, w);
    ugu_b200_outputs out;
    out.color = color ? color->data : nullptr;
    out.depth = depth ? reinterpret_cast<float*>(depth->data) : nullptr;
    out.normal = normal ? reinterpret_cast<float*>(normal->data) : nullptr;
    out.mask = mask ? mask->data : nullptr;
    out.face_id = face_id ? reinterpret_cast<int32_t*>(face_id->data) : nullptr;
    int rc;
    if (raytrace_) {  // ugu::Raytracer semantics; ray origin = PinholeCamera::org_ray_w (camera.h:431-436)
      const Eigen::Vector3f org = camera_->c2w().translation().cast<float>();
      rc = ugu_b200_rasterizer_raytrace(handle_, &c, org.data(), &out);
    } else {
      rc = ugu_b200_rasterizer_render(handle_, &c, &out);
    }
    if (rc != UGU_B200_OK) {
      LOGE("%s\n", ugu_b200_last_error(handle_));
      return false;
    }
    if (const uint64_t hz = ugu_b200_rasterizer_last_hazards(handle_))
      LOGW("%llu fragment(s) with depth < FLT_MIN or NaN: the reference's result is order dependent there "
           "(rasterizer.cc:229) and this frame may differ from it at those pixels\n", (unsigned long long)hz);
    return true;
  }
};

RasterizerCuda::RasterizerCuda() : pimpl_(new Impl(0)) {}
RasterizerCuda::~RasterizerCuda() {}
RasterizerCuda::RasterizerCuda(const RendererCpuOption& option, int device)
    : pimpl_(new Impl(device)) {
  set_option(option);
}
bool RasterizerCuda::valid() const { return pimpl_->handle_ != nullptr; }

bool RasterizerCuda::Depth2PointCloud(const Image1f& depth, const Camera& camera, Image3f* point_cloud,
                                      bool gl_coord) const {
  if (pimpl_->handle_ == nullptr) {
    LOGE("RasterizerCuda: no CUDA rasterizer (no CPU fallback)\n");
    return false;
  }
  if (depth.cols != camera.width() || depth.rows != camera.height()) {  // rgbd_util.cc:474-480
    LOGE("Depth2PointCloud depth size (%d, %d) and camera size (%d, %d) are different\n", depth.cols,
         depth.rows, camera.width(), camera.height());
    return false;
  }
  const auto* pin = dynamic_cast<const PinholeCamera*>(&camera);
  if (pin == nullptr || dynamic_cast<const OpenCvCamera*>(&camera) != nullptr || point_cloud == nullptr) {
    LOGE("RasterizerCuda::Depth2PointCloud supports PinholeCamera only\n");
    return false;
  }
  const bool fresh = point_cloud->cols != depth.cols || point_cloud->rows != depth.rows;
  if (fresh) Init(point_cloud, depth.cols, depth.rows, 0.0f);  // rgbd_util.cc:482-484
  ugu_b200_camera c{};
  c.width = camera.width();
  c.height = camera.height();
  c.fx = pin->focal_length().x();
  c.fy = pin->focal_length().y();
  c.cx = pin->principal_point().x();
  c.cy = pin->principal_point().y();
  if (ugu_b200_depth_to_point_cloud(pimpl_->handle_, &c, reinterpret_cast<const float*>(depth.data),
                                    reinterpret_cast<float*>(point_cloud->data), gl_coord ? 1 : 0,
                                    fresh ? 1 : 0) != UGU_B200_OK) {
    LOGE("%s\n", ugu_b200_last_error(pimpl_->handle_));
    return false;
  }
  return true;
}

namespace {
// Depth2PointCloudImpl (rgbd_util.cc:12-93) on top of the organized device form.
bool PointCloudMesh(const RasterizerCuda& self, const Image1f& depth, const Image3b* color, const Camera& camera,
                    Mesh* point_cloud, bool gl_coord) {
  if (depth.cols != camera.width() || depth.rows != camera.height()) {
    LOGE("Depth2PointCloud depth size (%d, %d) and camera size (%d, %d) are different\n", depth.cols, depth.rows,
         camera.width(), camera.height());
    return false;
  }
  if (color != nullptr) {  // rgbd_util.cc:23-39
    const float depth_aspect_ratio = static_cast<float>(depth.cols) / static_cast<float>(depth.rows);
    const float color_aspect_ratio = static_cast<float>(color->cols) / static_cast<float>(color->rows);
    if (std::abs(depth_aspect_ratio - color_aspect_ratio) > 0.01f) {
      LOGE("Depth2PointCloud depth aspect ratio %f and color aspect ratio %f are very different\n",
           depth_aspect_ratio, color_aspect_ratio);
      return false;
    }
  }
  if (point_cloud == nullptr) return false;
  Image3f organized;
  if (!self.Depth2PointCloud(depth, camera, &organized, gl_coord)) return false;
  point_cloud->Clear();
  std::vector<Eigen::Vector3f> vertices, vertex_colors;
  for (int y = 0; y < camera.height(); y++) {
    for (int x = 0; x < camera.width(); x++) {
      if (depth.at<float>(y, x) < std::numeric_limits<float>::min()) continue;
      const Vec3f& p = organized.at<Vec3f>(y, x);
      vertices.emplace_back(p[0], p[1], p[2]);
      if (color != nullptr) {  // nearest neighbour, rgbd_util.cc:62-78
        const Eigen::Vector2f uv(X2U(static_cast<float>(x), depth.cols), Y2V(static_cast<float>(y), depth.rows, false));
        const Eigen::Vector2i pos(static_cast<int>(std::round(U2X(uv.x(), color->cols))),
                                  static_cast<int>(std::round(V2Y(uv.y(), color->rows, false))));
        const Vec3b& c = color->at<Vec3b>(pos.y(), pos.x());
        vertex_colors.emplace_back(static_cast<float>(c[0]), static_cast<float>(c[1]), static_cast<float>(c[2]));
      }
    }
  }
  point_cloud->set_vertices(vertices);
  if (color != nullptr) point_cloud->set_vertex_colors(vertex_colors);
  return true;
}
}  // namespace

bool RasterizerCuda::Depth2PointCloud(const Image1f& depth, const Camera& camera, Mesh* point_cloud,
                                      bool gl_coord) const {
  return PointCloudMesh(*this, depth, nullptr, camera, point_cloud, gl_coord);
}
bool RasterizerCuda::Depth2PointCloud(const Image1f& depth, const Image3b& color, const Camera& camera,
                                      Mesh* point_cloud, bool gl_coord) const {
  return PointCloudMesh(*this, depth, &color, camera, point_cloud, gl_coord);
}

namespace {
// Shared body of the four Rasterize*AttributeToTexture overloads.  The Eigen vectors are plain
// packed float/int triples (and pairs), so data() is exactly the AoS layout the C-ABI takes.
template <typename ImageT>
bool UvRaster(ugu_b200_rasterizer* h, const std::vector<Eigen::Vector3f>& attrs,
              const std::vector<Eigen::Vector3i>* attr_indices, const std::vector<Eigen::Vector2f>& uvs,
              const std::vector<Eigen::Vector3i>& uv_indices, ImageT& texture, int width, int height,
              Image1b* mask, int texture_type) {
  if (h == nullptr) {
    LOGE("RasterizerCuda: no CUDA rasterizer (no CPU fallback)\n");
    return false;
  }
  if (attr_indices != nullptr) {  // raster_util.h:148-151
    if (attr_indices->empty() || attr_indices->size() != uv_indices.size()) return false;
  } else if (attrs.empty() || attrs.size() != uv_indices.size()) {  // raster_util.h:238-240
    return false;
  }
  const bool keep = !(width > 0 && height > 0);
  if (!keep) {
    texture = ImageT::zeros(height, width);  // raster_util.h:152
  } else if (texture.empty()) {
    return false;  // raster_util.h:153-155
  }
  if (mask != nullptr && (mask->cols != texture.cols || mask->rows != texture.rows)) {
    LOGE("RasterizerCuda: mask size differs from the texture size\n");
    return false;
  }
  ugu_b200_uv_raster_desc d{};
  d.num_attrs = static_cast<int32_t>(attrs.size());
  d.num_faces = static_cast<int32_t>(uv_indices.size());
  d.num_uv = static_cast<int32_t>(uvs.size());
  d.per_face = attr_indices == nullptr ? 1 : 0;
  d.attrs = attrs.empty() ? nullptr : attrs[0].data();
  d.attr_indices = attr_indices == nullptr ? nullptr : (*attr_indices)[0].data();
  d.uv = uvs.empty() ? nullptr : uvs[0].data();
  d.uv_indices = uv_indices[0].data();
  d.width = texture.cols;
  d.height = texture.rows;
  d.texture_type = texture_type;
  d.keep_texture = keep ? 1 : 0;
  d.keep_mask = 1;  // the reference never clears the mask
  if (ugu_b200_uv_raster(h, &d, texture.data, mask != nullptr ? mask->data : nullptr) != UGU_B200_OK) {
    LOGE("%s\n", ugu_b200_last_error(h));
    return false;
  }
  return true;
}
}  // namespace

bool RasterizerCuda::RasterizeVertexAttributeToTexture(
    const std::vector<Eigen::Vector3f>& vertex_attrs, const std::vector<Eigen::Vector3i>& vertex_attr_indices,
    const std::vector<Eigen::Vector2f>& uvs, const std::vector<Eigen::Vector3i>& uv_indices, Image3b& texture,
    int width, int height, Image1b* mask) const {
  return UvRaster(pimpl_->handle_, vertex_attrs, &vertex_attr_indices, uvs, uv_indices, texture, width, height, mask, 0);
}
bool RasterizerCuda::RasterizeVertexAttributeToTexture(
    const std::vector<Eigen::Vector3f>& vertex_attrs, const std::vector<Eigen::Vector3i>& vertex_attr_indices,
    const std::vector<Eigen::Vector2f>& uvs, const std::vector<Eigen::Vector3i>& uv_indices, Image3f& texture,
    int width, int height, Image1b* mask) const {
  return UvRaster(pimpl_->handle_, vertex_attrs, &vertex_attr_indices, uvs, uv_indices, texture, width, height, mask, 1);
}
bool RasterizerCuda::RasterizeFaceAttributeToTexture(
    const std::vector<Eigen::Vector3f>& face_attrs, const std::vector<Eigen::Vector2f>& uvs,
    const std::vector<Eigen::Vector3i>& uv_indices, Image3b& texture, int width, int height, Image1b* mask) const {
  return UvRaster(pimpl_->handle_, face_attrs, nullptr, uvs, uv_indices, texture, width, height, mask, 0);
}
bool RasterizerCuda::RasterizeFaceAttributeToTexture(
    const std::vector<Eigen::Vector3f>& face_attrs, const std::vector<Eigen::Vector2f>& uvs,
    const std::vector<Eigen::Vector3i>& uv_indices, Image3f& texture, int width, int height, Image1b* mask) const {
  return UvRaster(pimpl_->handle_, face_attrs, nullptr, uvs, uv_indices, texture, width, height, mask, 1);
}

bool RasterizerCuda::GenerateUvMask(const std::vector<Eigen::Vector2f>& uvs,
                                    const std::vector<Eigen::Vector3i>& uv_indices, Image1b& mask, uint8_t color,
                                    int width, int height) const {
  if (pimpl_->handle_ == nullptr) {
    LOGE("RasterizerCuda: no CUDA rasterizer (no CPU fallback)\n");
    return false;
  }
  if (width <= 0 || height <= 0) return false;
  mask = Image1b::zeros(height, width);  // raster_util.cc:19,26: a fresh image even on failure
  if (uv_indices.empty()) return false;
  if (ugu_b200_uv_mask(pimpl_->handle_, static_cast<int32_t>(uv_indices.size()), static_cast<int32_t>(uvs.size()),
                       uvs.empty() ? nullptr : uvs[0].data(), uv_indices[0].data(), width, height, color,
                       mask.data) != UGU_B200_OK) {
    LOGE("%s\n", ugu_b200_last_error(pimpl_->handle_));
    return false;
  }
  return true;
}

namespace {
// `color_camera` / `depth2color` != nullptr: the two-camera overload (rgbd_util.cc:283-466)
bool Depth2MeshCuda(ugu_b200_rasterizer* h, const Image1f& depth, const Image3b* color, const Camera& camera,
                    Mesh* mesh, float max_connect_z_diff, int x_step, int y_step, bool gl_coord,
                    const std::string& material_name, bool with_vertex_color, Image3f* point_cloud,
                    Image3f* normal, const Camera* color_camera = nullptr,
                    const Eigen::Affine3f* depth2color = nullptr) {
  if (h == nullptr) {
    LOGE("RasterizerCuda: no CUDA rasterizer (no CPU fallback)\n");
    return false;
  }
  const auto* pin = dynamic_cast<const PinholeCamera*>(&camera);
  if (pin == nullptr || dynamic_cast<const OpenCvCamera*>(&camera) != nullptr || mesh == nullptr) {
    LOGE("RasterizerCuda::Depth2Mesh supports PinholeCamera only\n");
    return false;
  }
  // the reference checks its scalar arguments first (rgbd_util.cc:100-112), then the sizes (:113-119)
  if (max_connect_z_diff >= 0 && x_step >= 1 && y_step >= 1 &&
      (depth.cols != camera.width() || depth.rows != camera.height())) {
    LOGE("Depth2Mesh depth size (%d, %d) and camera size (%d, %d) are different\n", depth.cols, depth.rows,
         camera.width(), camera.height());
    return false;
  }
  ugu_b200_camera c{};
  c.width = depth.cols;
  c.height = depth.rows;
  c.fx = pin->focal_length().x();
  c.fy = pin->focal_length().y();
  c.cx = pin->principal_point().x();
  c.cy = pin->principal_point().y();
  ugu_b200_depth2mesh_counts cnt{};
  int rc;
  if (color_camera != nullptr) {
    const auto* cpin = dynamic_cast<const PinholeCamera*>(color_camera);
    if (cpin == nullptr || dynamic_cast<const OpenCvCamera*>(color_camera) != nullptr || color == nullptr ||
        depth2color == nullptr) {
      LOGE("RasterizerCuda::Depth2Mesh supports PinholeCamera only\n");
      return false;
    }
    ugu_b200_color_camera cc{};
    for (int i = 0; i < 3; i++) {
      for (int j = 0; j < 3; j++) cc.depth2color_R[i * 3 + j] = depth2color->matrix()(i, j);
      cc.depth2color_t[i] = depth2color->matrix()(i, 3);
    }
    cc.fx = cpin->focal_length().x();
    cc.fy = cpin->focal_length().y();
    cc.cx = cpin->principal_point().x();
    cc.cy = cpin->principal_point().y();
    rc = ugu_b200_depth_to_mesh_two_camera(h, &c, &cc, reinterpret_cast<const float*>(depth.data), color->data,
                                           color->cols, color->rows, with_vertex_color ? 1 : 0, max_connect_z_diff,
                                           x_step, y_step, gl_coord ? 1 : 0, &cnt);
  } else {
    rc = ugu_b200_depth_to_mesh(h, &c, reinterpret_cast<const float*>(depth.data), color ? color->data : nullptr,
                                color ? color->cols : 0, color ? color->rows : 0, with_vertex_color ? 1 : 0,
                                max_connect_z_diff, x_step, y_step, gl_coord ? 1 : 0, &cnt);
  }
  if (rc != UGU_B200_OK) {
    LOGE("%s\n", ugu_b200_last_error(h));
    return false;
  }
  mesh->Clear();  // rgbd_util.cc:138
  const size_t nv = static_cast<size_t>(cnt.num_vertices), nf = static_cast<size_t>(cnt.num_faces);
  std::vector<Eigen::Vector3f> vertices(nv), normals(cnt.has_normals ? nv : 0), face_normals(nf),
      colors(cnt.has_colors ? nv : 0);
  std::vector<Eigen::Vector3i> faces(nf);
  std::vector<Eigen::Vector2f> uvs(cnt.has_uv ? nv : 0);
  // The one-camera overloads Init() the organized images (rgbd_util.cc:258,269); the two-camera one only when
  // their size differs (:441-443,454-456) and otherwise overwrites the vertices' pixels of what it was given.
  Image3f pc_fresh, nrm_fresh;
  Image3f* pc_dst = point_cloud;
  Image3f* nrm_dst = normal;
  auto keeps = [&](const Image3f* im) {
    return color_camera != nullptr && im != nullptr && im->cols == depth.cols && im->rows == depth.rows;
  };
  if (keeps(point_cloud)) pc_dst = &pc_fresh;
  if (keeps(normal)) nrm_dst = &nrm_fresh;
  if (pc_dst != nullptr) Init(pc_dst, depth.cols, depth.rows, 0.0f);
  if (nrm_dst != nullptr) Init(nrm_dst, depth.cols, depth.rows, 0.0f);
  if (ugu_b200_depth_to_mesh_fetch(
          h, nullptr, nv ? vertices[0].data() : nullptr, nf ? faces[0].data() : nullptr,
          normals.empty() ? nullptr : normals[0].data(), nf ? face_normals[0].data() : nullptr,
          uvs.empty() ? nullptr : uvs[0].data(), colors.empty() ? nullptr : colors[0].data(),
          pc_dst ? reinterpret_cast<float*>(pc_dst->data) : nullptr,
          nrm_dst ? reinterpret_cast<float*>(nrm_dst->data) : nullptr) != UGU_B200_OK) {
    LOGE("%s\n", ugu_b200_last_error(h));
    return false;
  }
  if (pc_dst == &pc_fresh || nrm_dst == &nrm_fresh) {
    for (int y = y_step; y < depth.rows; y += y_step)
      for (int x = x_step; x < depth.cols; x += x_step) {
        if (depth.at<float>(y, x) < std::numeric_limits<float>::min()) continue;
        if (pc_dst == &pc_fresh) point_cloud->at<Vec3f>(y, x) = pc_fresh.at<Vec3f>(y, x);
        if (nrm_dst == &nrm_fresh && cnt.has_normals) normal->at<Vec3f>(y, x) = nrm_fresh.at<Vec3f>(y, x);
      }
  }
  mesh->set_vertices(vertices);
  mesh->set_vertex_indices(faces);
  if (cnt.has_normals) {  // what Mesh::CalcNormal leaves behind (mesh.cc:297-337)
    mesh->set_face_normals(face_normals);
    mesh->set_normals(normals);
    mesh->set_normal_indices(faces);
  }
  if (color != nullptr) {  // rgbd_util.cc:241-253
    mesh->set_uv(uvs);
    mesh->set_uv_indices(faces);
    ObjMaterial material;
    color->copyTo(material.diffuse_tex);
    material.name = material_name;
    mesh->set_materials({material});
    mesh->set_material_ids(std::vector<int>(nf, 0));
  }
  if (cnt.has_colors) mesh->set_vertex_colors(colors);
  return true;
}
}  // namespace

bool RasterizerCuda::Depth2Mesh(const Image1f& depth, const Camera& camera, Mesh* mesh, float max_connect_z_diff,
                                int x_step, int y_step, bool gl_coord, Image3f* point_cloud,
                                Image3f* normal) const {
  return Depth2MeshCuda(pimpl_->handle_, depth, nullptr, camera, mesh, max_connect_z_diff, x_step, y_step,
                        gl_coord, "illegal_material", false, point_cloud, normal);
}
bool RasterizerCuda::Depth2Mesh(const Image1f& depth, const Image3b& color, const Camera& camera, Mesh* mesh,
                                float max_connect_z_diff, int x_step, int y_step, bool gl_coord,
                                const std::string& material_name, bool with_vertex_color, Image3f* point_cloud,
                                Image3f* normal) const {
  return Depth2MeshCuda(pimpl_->handle_, depth, &color, camera, mesh, max_connect_z_diff, x_step, y_step,
                        gl_coord, material_name, with_vertex_color, point_cloud, normal);
}

bool RasterizerCuda::Depth2Mesh(const Image1f& depth, const Image3b& color, const Camera& depth_camera,
                                const Camera& color_camera, const Eigen::Affine3f depth2color, Mesh* mesh,
                                float max_connect_z_diff, int x_step, int y_step, bool gl_coord,
                                const std::string& material_name, bool with_vertex_color, Image3f* point_cloud,
                                Image3f* normal) const {
  return Depth2MeshCuda(pimpl_->handle_, depth, &color, depth_camera, mesh, max_connect_z_diff, x_step, y_step,
                        gl_coord, material_name, with_vertex_color, point_cloud, normal, &color_camera,
                        &depth2color);
}

void RasterizerCuda::set_option(const RendererCpuOption& option) {
  option.CopyTo(&pimpl_->option_);  // rasterizer.cc:73-75 (oren_nayar_sigma is not copied)
}

void RasterizerCuda::set_mesh(std::shared_ptr<const Mesh> mesh) {
  pimpl_->mesh_initialized_ = false;  // rasterizer.cc:78
  pimpl_->mesh_ = mesh;
  if (mesh->face_normals().empty()) {
    LOGW("face normal is empty. culling and shading may not work\n");
  }
  if (mesh->normals().empty()) {
    LOGW("vertex normal is empty. shading may not work\n");
  }
}

bool RasterizerCuda::PrepareMesh() { return pimpl_->PrepareMesh(); }
void RasterizerCuda::set_camera(std::shared_ptr<const Camera> camera) { pimpl_->camera_ = camera; }

bool RasterizerCuda::Render(Image3b* color, Image1f* depth, Image3f* normal, Image1b* mask,
                            Image1i* face_id) const {
  return pimpl_->Render(color, depth, normal, mask, face_id);
}
bool RasterizerCuda::RenderColor(Image3b* color) const {
  return Render(color, nullptr, nullptr, nullptr, nullptr);
}
bool RasterizerCuda::RenderDepth(Image1f* depth) const {
  return Render(nullptr, depth, nullptr, nullptr, nullptr);
}
bool RasterizerCuda::RenderNormal(Image3f* normal) const {
  return Render(nullptr, nullptr, normal, nullptr, nullptr);
}
bool RasterizerCuda::RenderMask(Image1b* mask) const {
  return Render(nullptr, nullptr, nullptr, mask, nullptr);
}
bool RasterizerCuda::RenderFaceId(Image1i* face_id) const {
  return Render(nullptr, nullptr, nullptr, nullptr, face_id);
}

void RasterizerCuda::set_raytrace_mode(bool on) { pimpl_->raytrace_ = on; }

RaytracerCuda::RaytracerCuda() : RasterizerCuda() { set_raytrace_mode(true); }
RaytracerCuda::RaytracerCuda(const RendererCpuOption& option, int device) : RasterizerCuda(option, device) {
  set_raytrace_mode(true);
}
RaytracerCuda::~RaytracerCuda() = default;

bool RasterizerCuda::RenderGBuffer(GBuffer* gbuf, float z_near, float z_far, bool flat_normal, int geo_id) const {
  return pimpl_->RenderGBuffer(gbuf, z_near, z_far, flat_normal, geo_id);
}

bool RasterizerCuda::RenderW(Image3b* color, Image1w* depth, Image3f* normal, Image1b* mask,
                             Image1i* face_id) const {
  if (depth == nullptr) {  // rasterizer.cc:331-334
    LOGE("depth is nullptr");
    return false;
  }
  Image1f f_depth;
  const bool ok = Render(color, &f_depth, normal, mask, face_id);
  if (ok) {
    if (depth->cols != f_depth.cols || depth->rows != f_depth.rows)
      *depth = Image1w::zeros(f_depth.rows, f_depth.cols);
    ugu_b200_depth_to_u16(reinterpret_cast<const float*>(f_depth.data),
                          reinterpret_cast<uint16_t*>(depth->data),
                          static_cast<size_t>(f_depth.rows) * f_depth.cols);
  }
  return ok;
}

bool RasterizerCuda::RenderDepthW(Image1w* depth) const {
  return RenderW(nullptr, depth, nullptr, nullptr, nullptr);
}

}  // namespace ugu
