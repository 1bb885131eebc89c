// ugu::VisibilityTesterCuda — the drop-in a UGU maintainer adds next to ugu::VisibilityTester
// (include/ugu/texturing/visibility_tester.h:216-265): same public interface and call order
//
//     VisibilityTesterCuda tester(option);
//     tester.set_data(vertices, normals, indices);  tester.PrepareData();
//     tester.Test(keyframes, &info, facewise);        // or set_keyframe + Test(&info, facewise)
//
// and it fills the same ugu::VisibilityInfo: per-vertex / per-face `visible_keyframes` lists in keyframe order,
// then the reference's OWN VertexInfo::CalcStat / FaceInfo::CalcStat (host code of the reference tree) under the
// same option flags.  The per-keyframe tests run in libugu_b200_raster.so through the C-ABI
// (ugu_b200_visibility_*, include/ugu_b200_raster.h); the one substitution — the BVH ray replaced by this
// library's own depth map of the mesh — is stated there.  No CPU fallback: without an sm_100 device
// PrepareData() / Test() return false and log.  Pinhole cameras only.
#pragma once

#include <functional>
#include <memory>
#include <vector>

#include "ugu/texturing/visibility_tester.h"

namespace ugu {

class VisibilityTesterCuda {
  class Impl;
  std::unique_ptr<Impl> pimpl_;

 public:
  VisibilityTesterCuda();
  ~VisibilityTesterCuda();
  explicit VisibilityTesterCuda(const VisibilityTesterOption& option, int device = 0);
  void set_option(const VisibilityTesterOption& option);

  void set_data(const std::vector<Eigen::Vector3f>& vertices, const std::vector<Eigen::Vector3f>& normals,
                const std::vector<Eigen::Vector3i>& indices);
  bool PrepareData();
  void set_keyframe(std::shared_ptr<Keyframe> keyframe);

  bool Test(VisibilityInfo* info, bool facewise = true,
            std::function<void(VertexInfo&)> vert_custom_func = nullptr) const;
  bool Test(std::vector<std::shared_ptr<Keyframe>> keyframes, VisibilityInfo* info, bool facewise = true,
            std::function<void(VertexInfo&)> vert_custom_func = nullptr);

  bool valid() const;
};

}  // namespace ugu
