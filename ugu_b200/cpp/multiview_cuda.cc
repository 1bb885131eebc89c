// ugu::MultiViewRasterizerCuda: host C++ over the multi-view C-ABI (include/ugu_b200_raster.h,
// ugu_b200_multiview_*).  Compiled against the reference headers; every computation happens in the CUDA library.
#include "multiview_cuda.h"

#include <cstring>
#include <thread>

#include "ugu/log.h"
#include "ugu_b200_raster.h"

namespace ugu {

namespace {

// View-ordered landing buffer for the DMA of all devices.  Kept by the renderer and only ever grown: pinning
// hundreds of megabytes costs more than rendering them.
struct PinnedBlock {
  void* p = nullptr;
  size_t bytes = 0;
  PinnedBlock() = default;
  ~PinnedBlock() { ugu_b200_host_free(p); }
  PinnedBlock(const PinnedBlock&) = delete;
  PinnedBlock& operator=(const PinnedBlock&) = delete;
  void* Ensure(size_t need) {
    if (need > bytes) {
      ugu_b200_host_free(p);
      p = ugu_b200_host_alloc(need);
      bytes = p ? need : 0;
    }
    return need ? p : nullptr;
  }
};

ugu_b200_option ToOption(const RendererCpuOption& option) {
  ugu_b200_option o;
  o.diffuse_color = static_cast<int32_t>(option.diffuse_color);
  o.interp = static_cast<int32_t>(option.interp);
  o.shading_normal = static_cast<int32_t>(option.shading_normal);
  o.diffuse_shading = static_cast<int32_t>(option.diffuse_shading);
  o.backface_culling = option.backface_culling ? 1 : 0;
  o.depth_scale = option.depth_scale;
  o.oren_nayar_sigma = option.oren_nayar_sigma;
  return o;
}

// Camera constants exactly as rasterizer.cc:118-119 and camera.h:321 derive them.
bool ToCamera(const Camera& camera, ugu_b200_camera* c) {
  const PinholeCamera* pin = dynamic_cast<const PinholeCamera*>(&camera);
  if (pin == nullptr || dynamic_cast<const OpenCvCamera*>(&camera) != nullptr) return false;
  c->width = camera.width();
  c->height = camera.height();
  c->fx = pin->focal_length().x();
  c->fy = pin->focal_length().y();
  c->cx = pin->principal_point().x();
  c->cy = pin->principal_point().y();
  const Eigen::Matrix3f w2c_R = camera.w2c().rotation().cast<float>();
  const Eigen::Vector3f w2c_t = camera.w2c().translation().cast<float>();
  const Eigen::Matrix3f c2w_R = camera.c2w().matrix().block<3, 3>(0, 0).cast<float>();
  for (int i = 0; i < 3; i++) {
    for (int j = 0; j < 3; j++) {
      c->w2c_R[i * 3 + j] = w2c_R(i, j);
      c->c2w_R[i * 3 + j] = c2w_R(i, j);
    }
    c->w2c_t[i] = w2c_t[i];
  }
  return true;
}

template <typename ImageT>
void Scatter(const void* src, size_t view_bytes, int w, int h, std::vector<ImageT>* dst) {
  for (size_t i = 0; i < dst->size(); i++) {
    ImageT& im = (*dst)[i];
    if (im.cols != w || im.rows != h) im = ImageT::zeros(h, w);
    std::memcpy(im.data, static_cast<const char*>(src) + i * view_bytes, view_bytes);
  }
}

}  // namespace

class MultiViewRasterizerCuda::Impl {
 public:
  PinnedBlock land_[5];  // colour, depth, normal, mask, face_id
  ugu_b200_multiview* mv_ = nullptr;
  std::shared_ptr<const Mesh> mesh_;
  RendererCpuOption option_;
  bool mesh_initialized_ = false;

  explicit Impl(const std::vector<int>& devices) {
    std::vector<int32_t> devs(devices.begin(), devices.end());
    if (devs.empty())
      for (int i = 0; i < ugu_b200_device_count(); i++) devs.push_back(i);
    if (devs.empty() ||
        ugu_b200_multiview_create(devs.data(), static_cast<int32_t>(devs.size()), UGU_B200_TRANSPORT_AUTO, &mv_) !=
            UGU_B200_OK) {
      LOGE("MultiViewRasterizerCuda: %s\n", devs.empty() ? "no usable CUDA device" : ugu_b200_multiview_last_error(nullptr));
      mv_ = nullptr;
    }
  }
  ~Impl() { ugu_b200_multiview_destroy(mv_); }

  bool Cameras(const std::vector<std::shared_ptr<const Camera>>& cameras, std::vector<ugu_b200_camera>* out) const {
    if (mv_ == nullptr) {
      LOGE("MultiViewRasterizerCuda: no CUDA rasterizer (no CPU fallback)\n");
      return false;
    }
    if (!mesh_initialized_) {
      LOGE("mesh has not been initialized\n");
      return false;
    }
    if (cameras.empty()) {
      LOGE("camera has not been set\n");
      return false;
    }
    out->resize(cameras.size());
    for (size_t i = 0; i < cameras.size(); i++) {
      if (cameras[i] == nullptr || !ToCamera(*cameras[i], &(*out)[i])) {
        LOGE("MultiViewRasterizerCuda supports PinholeCamera only (no CPU fallback)\n");
        return false;
      }
    }
    const ugu_b200_option o = ToOption(option_);
    ugu_b200_multiview_set_option(mv_, &o);
    return true;
  }
};

MultiViewRasterizerCuda::MultiViewRasterizerCuda(const RendererCpuOption& option, const std::vector<int>& devices)
    : pimpl_(new Impl(devices)) {
  set_option(option);
}
MultiViewRasterizerCuda::~MultiViewRasterizerCuda() = default;

void MultiViewRasterizerCuda::set_option(const RendererCpuOption& option) { option.CopyTo(&pimpl_->option_); }

void MultiViewRasterizerCuda::set_mesh(std::shared_ptr<const Mesh> mesh) {
  pimpl_->mesh_initialized_ = false;
  pimpl_->mesh_ = mesh;
  if (mesh != nullptr && mesh->face_normals().empty()) LOGW("face normal is empty. culling and shading may not work\n");
  if (mesh != nullptr && mesh->normals().empty()) LOGW("vertex normal is empty. shading may not work\n");
}

bool MultiViewRasterizerCuda::PrepareMesh() {
  Impl& p = *pimpl_;
  if (p.mesh_ == nullptr) {
    LOGE("mesh has not been set\n");
    return false;
  }
  if (p.mv_ == nullptr) {
    LOGE("MultiViewRasterizerCuda: no CUDA rasterizer (no CPU fallback)\n");
    return false;
  }
  const Mesh& m = *p.mesh_;
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
  static_assert(sizeof(Eigen::Vector3f) == 12 && sizeof(Eigen::Vector3i) == 12 && sizeof(Eigen::Vector2f) == 8,
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
  if (ugu_b200_multiview_upload_mesh(p.mv_, &d) != UGU_B200_OK) {
    LOGE("%s\n", ugu_b200_multiview_last_error(p.mv_));
    return false;
  }
  p.mesh_initialized_ = true;
  return true;
}

bool MultiViewRasterizerCuda::RenderBatch(const std::vector<std::shared_ptr<const Camera>>& cameras,
                                          std::vector<Image3b>* color, std::vector<Image1f>* depth,
                                          std::vector<Image3f>* normal, std::vector<Image1b>* mask,
                                          std::vector<Image1i>* face_id, bool round_robin) const {
  std::vector<ugu_b200_camera> cams;
  if (!pimpl_->Cameras(cameras, &cams)) return false;
  if (!color && !depth && !normal && !mask && !face_id) {
    LOGE("all arguments are nullptr. nothing to do\n");
    return false;
  }
  const size_t n = cams.size();
  const int w = cams[0].width, h = cams[0].height;
  const size_t P = static_cast<size_t>(w) * h;
  PinnedBlock* L = pimpl_->land_;
  ugu_b200_outputs o;
  std::memset(&o, 0, sizeof(o));
  o.color = static_cast<uint8_t*>(L[0].Ensure(color ? n * P * 3 : 0));
  o.depth = static_cast<float*>(L[1].Ensure(depth ? n * P * 4 : 0));
  o.normal = static_cast<float*>(L[2].Ensure(normal ? n * P * 12 : 0));
  o.mask = static_cast<uint8_t*>(L[3].Ensure(mask ? n * P : 0));
  o.face_id = static_cast<int32_t*>(L[4].Ensure(face_id ? n * P * 4 : 0));
  if ((color && !o.color) || (depth && !o.depth) || (normal && !o.normal) || (mask && !o.mask) ||
      (face_id && !o.face_id)) {
    LOGE("MultiViewRasterizerCuda: pinned host allocation failed\n");
    return false;
  }
  if (ugu_b200_multiview_render_batch(pimpl_->mv_, static_cast<int32_t>(n), cams.data(),
                                      round_robin ? UGU_B200_PARTITION_ROUND_ROBIN : UGU_B200_PARTITION_CONTIGUOUS,
                                      &o) != UGU_B200_OK) {
    LOGE("%s\n", ugu_b200_multiview_last_error(pimpl_->mv_));
    return false;
  }
  // into the caller's images: one host thread per output
  std::vector<std::thread> th;
  if (color) { color->resize(n); th.emplace_back([&] { Scatter(o.color, P * 3, w, h, color); }); }
  if (depth) { depth->resize(n); th.emplace_back([&] { Scatter(o.depth, P * 4, w, h, depth); }); }
  if (normal) { normal->resize(n); th.emplace_back([&] { Scatter(o.normal, P * 12, w, h, normal); }); }
  if (mask) { mask->resize(n); th.emplace_back([&] { Scatter(o.mask, P, w, h, mask); }); }
  if (face_id) { face_id->resize(n); th.emplace_back([&] { Scatter(o.face_id, P * 4, w, h, face_id); }); }
  for (auto& t : th) t.join();
  return true;
}

bool MultiViewRasterizerCuda::GatherFaceIds(const std::vector<std::shared_ptr<const Camera>>& cameras,
                                            std::vector<Image1i>* face_id, int root, int chunk_views,
                                            bool round_robin) const {
  std::vector<ugu_b200_camera> cams;
  if (!pimpl_->Cameras(cameras, &cams)) return false;
  if (face_id == nullptr) {
    LOGE("all arguments are nullptr. nothing to do\n");
    return false;
  }
  const size_t n = cams.size();
  const int w = cams[0].width, h = cams[0].height;
  const size_t P = static_cast<size_t>(w) * h;
  int32_t* land = static_cast<int32_t*>(pimpl_->land_[4].Ensure(n * P * 4));
  if (land == nullptr) {
    LOGE("MultiViewRasterizerCuda: pinned host allocation failed\n");
    return false;
  }
  if (ugu_b200_multiview_gather_face_ids(pimpl_->mv_, static_cast<int32_t>(n), cams.data(),
                                         round_robin ? UGU_B200_PARTITION_ROUND_ROBIN : UGU_B200_PARTITION_CONTIGUOUS,
                                         root, chunk_views, nullptr, land) != UGU_B200_OK) {
    LOGE("%s\n", ugu_b200_multiview_last_error(pimpl_->mv_));
    return false;
  }
  face_id->resize(n);
  Scatter(land, P * 4, w, h, face_id);
  return true;
}

bool MultiViewRasterizerCuda::FaceVisibility(const std::vector<std::shared_ptr<const Camera>>& cameras,
                                             std::vector<uint32_t>* pixel_count, std::vector<uint32_t>* view_count,
                                             std::vector<int32_t>* best_view, std::vector<uint32_t>* best_count,
                                             int root, int chunk_views, bool round_robin) const {
  std::vector<ugu_b200_camera> cams;
  if (!pimpl_->Cameras(cameras, &cams)) return false;
  const size_t F = pimpl_->mesh_->vertex_indices().size();
  if (pixel_count) pixel_count->assign(F, 0u);
  if (view_count) view_count->assign(F, 0u);
  if (best_view) best_view->assign(F, -1);
  if (best_count) best_count->assign(F, 0u);
  if (ugu_b200_multiview_reduce_face_visibility(
          pimpl_->mv_, static_cast<int32_t>(cams.size()), cams.data(),
          round_robin ? UGU_B200_PARTITION_ROUND_ROBIN : UGU_B200_PARTITION_CONTIGUOUS, root, chunk_views,
          pixel_count ? pixel_count->data() : nullptr, view_count ? view_count->data() : nullptr,
          best_view ? best_view->data() : nullptr, best_count ? best_count->data() : nullptr) != UGU_B200_OK) {
    LOGE("%s\n", ugu_b200_multiview_last_error(pimpl_->mv_));
    return false;
  }
  return true;
}

int MultiViewRasterizerCuda::size() const { return pimpl_->mv_ ? ugu_b200_multiview_size(pimpl_->mv_) : 0; }
bool MultiViewRasterizerCuda::valid() const { return pimpl_->mv_ != nullptr; }
const char* MultiViewRasterizerCuda::transport() const {
  return pimpl_->mv_ && ugu_b200_multiview_transport(pimpl_->mv_) == UGU_B200_TRANSPORT_NCCL ? "nccl" : "peer";
}

}  // namespace ugu
