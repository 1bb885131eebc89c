// ugu::VisibilityTesterCuda: host C++ over the C-ABI (ugu_b200_visibility_*).  Mirrors the state machine,
// validation and option handling of ugu::VisibilityTester (src/texturing/visibility_tester.cc:220-300,670-733);
// the per-keyframe vertex / face tests run on the device, the list building and CalcStat are the reference's.
#include "visibility_tester_cuda.h"

#include <cstring>

#include "ugu/log.h"
#include "ugu_b200_raster.h"

namespace ugu {

class VisibilityTesterCuda::Impl {
 public:
  ugu_b200_rasterizer* handle_ = nullptr;
  bool mesh_initialized_ = false;
  std::shared_ptr<const Keyframe> keyframe_;
  VisibilityTesterOption option_;
  std::vector<Eigen::Vector3f> vertices_, normals_;
  std::vector<Eigen::Vector3i> indices_;

  explicit Impl(int device) {
    if (ugu_b200_rasterizer_create(device, &handle_) != UGU_B200_OK) {
      LOGE("VisibilityTesterCuda: %s\n", ugu_b200_last_error(nullptr));
      handle_ = nullptr;
    }
  }
  ~Impl() { ugu_b200_rasterizer_destroy(handle_); }

  bool PrepareData() {
    if (vertices_.empty() || normals_.empty() || indices_.empty()) {
      LOGE("mesh is empty\n");
      return false;
    }
    if (handle_ == nullptr) {
      LOGE("VisibilityTesterCuda: no CUDA device (no CPU fallback)\n");
      return false;
    }
    if (vertices_.size() != normals_.size()) {
      LOGE("VisibilityTesterCuda: %d normals for %d vertices\n", static_cast<int>(normals_.size()),
           static_cast<int>(vertices_.size()));
      return false;
    }
    // the geometry also goes to the rasterizer of this handle: it renders the self depth maps (the stand-in
    // for the reference's BVH, built at this point by bvh_->Build())
    ugu_b200_mesh_desc d;
    std::memset(&d, 0, sizeof(d));
    d.num_vertices = static_cast<int32_t>(vertices_.size());
    d.num_faces = static_cast<int32_t>(indices_.size());
    d.vertices = vertices_[0].data();
    d.vertex_indices = indices_[0].data();
    if (ugu_b200_rasterizer_upload_mesh(handle_, &d) != UGU_B200_OK ||
        ugu_b200_visibility_set_data(handle_, d.num_vertices, vertices_[0].data(), normals_[0].data(), d.num_faces,
                                     indices_[0].data()) != UGU_B200_OK) {
      LOGE("%s\n", ugu_b200_last_error(handle_));
      return false;
    }
    mesh_initialized_ = true;
    return true;
  }

  // ValidateAndInitBeforeTest (visibility_tester.cc:274-340)
  bool Validate(const Keyframe* kf, VisibilityInfo* info) const {
    if (info == nullptr) {
      LOGE("info is nullptr\n");
      return false;
    }
    if (kf == nullptr) {
      LOGE("keyframe has not been set\n");
      return false;
    }
    if (!mesh_initialized_) {
      LOGE("mesh has not been initialized\n");
      return false;
    }
    if (kf->color.empty()) {
      LOGE("keyframe color is empty\n");
      return false;
    }
    if (option_.use_mask && kf->mask.empty()) {
      LOGE("use_mask is true but keyframe mask is empty\n");
      return false;
    }
    if (option_.use_depth && kf->depth.empty()) {
      LOGE("use_depth is true but keyframe depth is empty\n");
      return false;
    }
    if (info->vertex_info_list.size() != vertices_.size()) {
      LOGE("info->vertex_info_list size %d is different from mesh %d\n",
           static_cast<int>(info->vertex_info_list.size()), static_cast<int>(vertices_.size()));
      return false;
    }
    if (info->face_info_list.size() != indices_.size()) {
      LOGE("info->face_info_list size %d is different from mesh %d\n",
           static_cast<int>(info->face_info_list.size()), static_cast<int>(indices_.size()));
      return false;
    }
    if (dynamic_cast<const PinholeCamera*>(kf->camera.get()) == nullptr ||
        dynamic_cast<const OpenCvCamera*>(kf->camera.get()) != nullptr) {
      LOGE("VisibilityTesterCuda supports PinholeCamera only (no CPU fallback)\n");
      return false;
    }
    for (auto& ii : info->vertex_info_list) {
      ii.min_viewing_angle_color = option_.invalid_color;
      ii.min_distance_color = option_.invalid_color;
      ii.mean_color = option_.invalid_color;
      ii.median_color = option_.invalid_color;
      ii.mean_viewing_angle_color = option_.invalid_color;
      ii.median_viewing_angle_color = option_.invalid_color;
      ii.mean_distance_color = option_.invalid_color;
      ii.median_distance_color = option_.invalid_color;
    }
    return true;
  }

  static void FillKeyframe(const Keyframe& kf, ugu_b200_keyframe* out) {
    std::memset(out, 0, sizeof(*out));
    const PinholeCamera& pin = *dynamic_cast<const PinholeCamera*>(kf.camera.get());
    out->id = kf.id;
    ugu_b200_camera& c = out->cam;
    c.width = pin.width();
    c.height = pin.height();
    c.fx = pin.focal_length().x();
    c.fy = pin.focal_length().y();
    c.cx = pin.principal_point().x();
    c.cy = pin.principal_point().y();
    const Eigen::Matrix3f w2c_R = pin.w2c().rotation().cast<float>();  // for the self depth render (rasterizer.cc:118)
    const Eigen::Vector3f w2c_t = pin.w2c().translation().cast<float>();
    const Eigen::Affine3f w2c = pin.w2c().cast<float>();                // what TestVertex multiplies by (:422)
    const Eigen::Affine3f c2w = pin.c2w().cast<float>();
    for (int i = 0; i < 3; i++) {
      for (int j = 0; j < 3; j++) {
        c.w2c_R[i * 3 + j] = w2c_R(i, j);
        c.c2w_R[i * 3 + j] = c2w.matrix()(i, j);
        out->w2c[i * 4 + j] = w2c.matrix()(i, j);
      }
      c.w2c_t[i] = w2c_t[i];
      out->w2c[i * 4 + 3] = w2c.matrix()(i, 3);
      out->c2w_t[i] = c2w.matrix()(i, 3);
    }
    out->color = kf.color.data;
    out->depth = kf.depth.empty() ? nullptr : reinterpret_cast<const float*>(kf.depth.data);
    out->mask = kf.mask.empty() ? nullptr : kf.mask.data;
    out->zbuf = nullptr;  // rendered by the library from the uploaded mesh
  }

  // One keyframe: device tests -> dense records -> the reference's list updates, in index order.
  bool TestOne(const Keyframe& kf, VisibilityInfo* info, bool facewise) const {
    ugu_b200_keyframe k;
    FillKeyframe(kf, &k);
    ugu_b200_vis_option o;
    o.backface_culling_angle_rad_th = option_.backface_culling_angle_rad_th;
    o.use_mask = option_.use_mask ? 1 : 0;
    o.use_depth = option_.use_depth ? 1 : 0;
    o.interp = static_cast<int32_t>(option_.interp);
    o.max_viewing_angle = option_.max_viewing_angle;
    o.max_distance_from_camera_to_vertex = option_.max_distance_from_camera_to_vertex;
    o.max_depth_difference = option_.max_depth_difference;
    o.eps = option_.eps;
    o.facewise = facewise ? 1 : 0;
    o.self_occlusion = 1;
    const size_t V = vertices_.size(), F = indices_.size();
    std::vector<uint8_t> vv(V), fv(F);
    std::vector<ugu_b200_vis_vertex> vr(V + 1);  // + 1: room to align to 16 bytes
    std::vector<ugu_b200_vis_face> fr(F);
    ugu_b200_vis_records rec;
    rec.vertex_visible = vv.data();
    rec.vertex = vr.data();
    rec.face_visible = fv.data();
    rec.face = fr.data();
    if (ugu_b200_visibility_test(handle_, 1, &k, 0, &o, &rec) != UGU_B200_OK) {
      LOGE("%s\n", ugu_b200_last_error(handle_));
      return false;
    }
    for (size_t v = 0; v < V; v++) {
      if (!vv[v]) continue;
      VertexInfoPerKeyframe vi;
      vi.kf_id = kf.id;
      vi.color = Eigen::Vector3f(vr[v].color[0], vr[v].color[1], vr[v].color[2]);
      vi.intensity = vr[v].intensity;
      vi.projected_pos = Eigen::Vector2f(vr[v].projected_pos[0], vr[v].projected_pos[1]);
      vi.viewing_angle = vr[v].viewing_angle;
      vi.distance = vr[v].distance;
      info->Update(static_cast<int>(v), vi);
    }
    if (facewise || option_.collect_face_info) {
      for (size_t f = 0; f < F; f++) {
        if (!fv[f]) continue;
        FaceInfoPerKeyframe fi;
        fi.kf_id = kf.id;
        fi.face_id = static_cast<int>(f);
        fi.area = fr[f].area;
        fi.viewing_angle = fr[f].viewing_angle;
        fi.distance = fr[f].distance;
        for (int j = 0; j < 3; j++)
          fi.projected_tri[j] = Eigen::Vector2f(fr[f].projected_tri[2 * j], fr[f].projected_tri[2 * j + 1]);
        info->Update(static_cast<int>(f), fi);
      }
    }
    return true;
  }

  // Test(info, facewise) (visibility_tester.cc:670-700)
  bool Test(VisibilityInfo* info, bool facewise, std::function<void(VertexInfo&)> func) const {
    if (!Validate(keyframe_.get(), info)) return false;
    if (!TestOne(*keyframe_, info, facewise)) return false;
    if (option_.calc_stat_vertex_info) info->CalcStatVertexInfo(func);
    if ((facewise || option_.collect_face_info) && option_.calc_stat_face_info) info->CalcStatFaceInfo();
    return true;
  }
};

VisibilityTesterCuda::VisibilityTesterCuda() : pimpl_(new Impl(0)) {}
VisibilityTesterCuda::~VisibilityTesterCuda() {}
VisibilityTesterCuda::VisibilityTesterCuda(const VisibilityTesterOption& option, int device) : pimpl_(new Impl(device)) {
  set_option(option);
}
void VisibilityTesterCuda::set_option(const VisibilityTesterOption& option) { option.CopyTo(&pimpl_->option_); }
void VisibilityTesterCuda::set_data(const std::vector<Eigen::Vector3f>& vertices,
                                    const std::vector<Eigen::Vector3f>& normals,
                                    const std::vector<Eigen::Vector3i>& indices) {
  pimpl_->mesh_initialized_ = false;
  pimpl_->vertices_ = vertices;
  pimpl_->normals_ = normals;
  pimpl_->indices_ = indices;
}
bool VisibilityTesterCuda::PrepareData() { return pimpl_->PrepareData(); }
void VisibilityTesterCuda::set_keyframe(std::shared_ptr<Keyframe> keyframe) { pimpl_->keyframe_ = keyframe; }
bool VisibilityTesterCuda::Test(VisibilityInfo* info, bool facewise, std::function<void(VertexInfo&)> func) const {
  return pimpl_->Test(info, facewise, func);
}
// Test(keyframes, info, facewise) (visibility_tester.cc:702-733): statistics only after the last keyframe
bool VisibilityTesterCuda::Test(std::vector<std::shared_ptr<Keyframe>> keyframes, VisibilityInfo* info, bool facewise,
                                std::function<void(VertexInfo&)> func) {
  if (keyframes.empty()) {
    LOGE("keyframe has not been set\n");
    return false;
  }
  bool ret = true;
  VisibilityTesterOption org_option;
  pimpl_->option_.CopyTo(&org_option);
  pimpl_->option_.calc_stat_face_info = false;
  pimpl_->option_.calc_stat_vertex_info = false;
  for (size_t i = 0; i + 1 < keyframes.size(); i++) {
    set_keyframe(keyframes[i]);
    if (!Test(info, facewise, func)) {
      ret = false;
      break;
    }
  }
  set_option(org_option);
  set_keyframe(keyframes.back());
  if (!Test(info, facewise, func)) ret = false;
  return ret;
}
bool VisibilityTesterCuda::valid() const { return pimpl_->handle_ != nullptr; }

}  // namespace ugu
