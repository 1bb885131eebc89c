// ugu::RasterizerCuda — the drop-in a UGU maintainer adds next to ugu::Rasterizer.
//
// Same interface as the reference's CPU rasterizer (include/ugu/renderer/cpu/rasterizer.h:14-53):
// it derives from ugu::RendererCpu (include/ugu/renderer/cpu/renderer.h:42-76), so the swap at a
// construction site (examples/ex02_renderer.cc:249-254) is one line:
//
//     std::unique_ptr<ugu::RendererCpu> renderer = std::make_unique<ugu::RasterizerCuda>(option);
//
// All work happens in libugu_b200_raster.so through the C-ABI of include/ugu_b200_raster.h.  There
// is no CPU fallback: if no sm_100 device is usable, PrepareMesh()/Render() return false and log.
#pragma once

#include <memory>
#include <string>
#include <vector>

#include "ugu/image.h"  // renderer/base.h (GBuffer) uses the image types without including them
#include "ugu/renderer/base.h"
#include "ugu/renderer/cpu/renderer.h"

namespace ugu {

class RasterizerCuda : public RendererCpu {
  class Impl;
  std::unique_ptr<Impl> pimpl_;

 public:
  RasterizerCuda();
  ~RasterizerCuda() override;
  explicit RasterizerCuda(const RendererCpuOption& option, int device = 0);

  void set_option(const RendererCpuOption& option) override;
  void set_mesh(std::shared_ptr<const Mesh> mesh) override;
  bool PrepareMesh() override;
  void set_camera(std::shared_ptr<const Camera> camera) override;

  bool Render(Image3b* color, Image1f* depth, Image3f* normal, Image1b* mask,
              Image1i* face_id) const override;
  bool RenderColor(Image3b* color) const override;
  bool RenderDepth(Image1f* depth) const override;
  bool RenderNormal(Image3f* normal) const override;
  bool RenderMask(Image1b* mask) const override;
  bool RenderFaceId(Image1i* face_id) const override;
  bool RenderW(Image3b* color, Image1w* depth, Image3f* normal, Image1b* mask,
               Image1i* face_id) const override;
  bool RenderDepthW(Image1w* depth) const override;

  // ugu::Depth2PointCloud(depth, camera, &point_cloud, gl_coord) (include/ugu/util/rgbd_util.h:17-18,
  // src/util/rgbd_util.cc:472-513) on this renderer's device — the step examples/ex02_renderer.cc
  // runs on the rendered depth.  Same contract: false on a size mismatch, point_cloud re-initialised
  // to 0 when its size differs, pixels with depth < FLT_MIN left as they were.  Pinhole cameras only.
  bool Depth2PointCloud(const Image1f& depth, const Camera& camera, Image3f* point_cloud,
                        bool gl_coord = false) const;
  // The Mesh-output overloads (rgbd_util.h:20-25, Depth2PointCloudImpl rgbd_util.cc:12-93): the valid pixels in
  // raster order as `vertices` (and, with `color`, their nearest colour texels as `vertex_colors`).  The
  // unprojection runs on the device (the organized form above); compaction and the colour lookup are host loops
  // over the reference's own X2U / Y2V / U2X / V2Y expressions.  Same false returns.
  bool Depth2PointCloud(const Image1f& depth, const Camera& camera, Mesh* point_cloud, bool gl_coord = false) const;
  bool Depth2PointCloud(const Image1f& depth, const Image3b& color, const Camera& camera, Mesh* point_cloud,
                        bool gl_coord = false) const;

  // ugu::RasterizeVertexAttributeToTexture<T> / RasterizeFaceAttributeToTexture<T> / GenerateUvMask
  // (include/ugu/util/raster_util.h:139-278, src/util/raster_util.cc:9-27) on this renderer's device,
  // T = Vec3b or Vec3f, in the reference's sequential semantics (num_threads = 1: the highest
  // covering face index owns a texel; the reference is racy with more threads).  Same contract:
  // false for empty / mismatched index arrays, width/height <= 0 rasterises over the existing
  // texture, the mask (same size as the texture) is in/out.  Out-of-range indices return false.
  bool RasterizeVertexAttributeToTexture(const std::vector<Eigen::Vector3f>& vertex_attrs,
                                         const std::vector<Eigen::Vector3i>& vertex_attr_indices,
                                         const std::vector<Eigen::Vector2f>& uvs,
                                         const std::vector<Eigen::Vector3i>& uv_indices, Image3b& texture,
                                         int width = 512, int height = 512, Image1b* mask = nullptr) const;
  bool RasterizeVertexAttributeToTexture(const std::vector<Eigen::Vector3f>& vertex_attrs,
                                         const std::vector<Eigen::Vector3i>& vertex_attr_indices,
                                         const std::vector<Eigen::Vector2f>& uvs,
                                         const std::vector<Eigen::Vector3i>& uv_indices, Image3f& texture,
                                         int width = 512, int height = 512, Image1b* mask = nullptr) const;
  bool RasterizeFaceAttributeToTexture(const std::vector<Eigen::Vector3f>& face_attrs,
                                       const std::vector<Eigen::Vector2f>& uvs,
                                       const std::vector<Eigen::Vector3i>& uv_indices, Image3b& texture,
                                       int width = 512, int height = 512, Image1b* mask = nullptr) const;
  bool RasterizeFaceAttributeToTexture(const std::vector<Eigen::Vector3f>& face_attrs,
                                       const std::vector<Eigen::Vector2f>& uvs,
                                       const std::vector<Eigen::Vector3i>& uv_indices, Image3f& texture,
                                       int width = 512, int height = 512, Image1b* mask = nullptr) const;
  bool GenerateUvMask(const std::vector<Eigen::Vector2f>& uvs, const std::vector<Eigen::Vector3i>& uv_indices,
                      Image1b& mask, uint8_t color, int width = 512, int height = 512) const;

  // ugu::Depth2Mesh, both overloads (include/ugu/util/rgbd_util.h:27-37, src/util/rgbd_util.cc:94-282,
  // 528-545) on this renderer's device: raster-order vertex / face compaction as a prefix sum, face
  // and vertex normals as Mesh::CalcNormal computes them.  Fills `mesh` exactly like the reference
  // (vertices, vertex_indices, normals / normal_indices / face_normals, and for the textured
  // overload uv / uv_indices / materials / material_ids and optional vertex colours) and the
  // organized point-cloud / normal images.  Same false returns.  Pinhole cameras only.
  bool Depth2Mesh(const Image1f& depth, const Camera& camera, Mesh* mesh, float max_connect_z_diff,
                  int x_step = 1, int y_step = 1, bool gl_coord = false, Image3f* point_cloud = nullptr,
                  Image3f* normal = nullptr) const;
  bool Depth2Mesh(const Image1f& depth, const Image3b& color, const Camera& camera, Mesh* mesh,
                  float max_connect_z_diff, int x_step = 1, int y_step = 1, bool gl_coord = false,
                  const std::string& material_name = "Depth2Mesh_mat", bool with_vertex_color = false,
                  Image3f* point_cloud = nullptr, Image3f* normal = nullptr) const;
  // the two-camera overload (rgbd_util.h:39-48, rgbd_util.cc:283-466): uv / vertex colours from the projection of
  // depth2color * p into `color_camera`; like the reference it keeps the pixels without a vertex of organized
  // images that already have the depth image's size
  bool Depth2Mesh(const Image1f& depth, const Image3b& color, const Camera& depth_camera,
                  const Camera& color_camera, const Eigen::Affine3f depth2color, Mesh* mesh,
                  float max_connect_z_diff, int x_step = 1, int y_step = 1, bool gl_coord = false,
                  const std::string& material_name = "Depth2Mesh_mat", bool with_vertex_color = false,
                  Image3f* point_cloud = nullptr, Image3f* normal = nullptr) const;

  // ugu::GBuffer (include/ugu/renderer/base.h:34-77) as RendererGl::ReadGbuf fills it
  // (src/renderer/gl/renderer.cc:501-650) from the gbuf shaders (src/shader/glsl/{vert/gbuf.vert,
  // geom/dummy.geom,frag/gbuf.frag}), computed headless from THIS rasterizer's winners: stencil,
  // depth_01 (near / far as RendererGl::SetNearFar), face_id (+1, 0 = background), geo_id, bary, uv,
  // pos_wld / pos_cam, normal_wld / normal_cam, color (= Render's); `shaded` is zeroed.  The sampling
  // rule is the CPU Rasterizer's (integer pixel centres, inclusive edges), not OpenGL's; definition in
  // include/ugu_b200_raster.h.  false + log on Render's precondition failures or bad near / far.
  bool RenderGBuffer(GBuffer* gbuf, float z_near = 0.01f, float z_far = 1000.0f, bool flat_normal = false,
                     int geo_id = 0) const;

  // false when the CUDA library could not create a rasterizer (no B200 / no driver).
  bool valid() const;

 protected:
  void set_raytrace_mode(bool on);  // Render / RenderW answer like ugu::Raytracer (see RaytracerCuda)
};

// Drop-in for ugu::Raytracer (include/ugu/renderer/cpu/raytracer.h:14-53, src/renderer/cpu/raytracer.cc:
// 117-225): the same RendererCpu interface and call order; Render produces the outputs of the reference's
// nanort ray caster — nearest hit of each pixel's primary ray (watertight intersector,
// third_party/nanort/nanort.h:1050-1199, t in [1e-4, 1e30)), back-face test on the hit, face 0 shaded,
// depth = camera-space z of the hit point * depth_scale, normal / colour from the intersector's (u, v).
// No BVH is built: PrepareMesh uploads the mesh as for the rasterizer (ugu_b200_rasterizer_raytrace).
class RaytracerCuda : public RasterizerCuda {
 public:
  RaytracerCuda();
  ~RaytracerCuda() override;
  explicit RaytracerCuda(const RendererCpuOption& option, int device = 0);
};

}  // namespace ugu
