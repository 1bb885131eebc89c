// ugu::MultiViewRasterizerCuda — the orbit loop of examples/ex09_renderer_realtime.cc:125-182 (one renderer,
// set_c2w -> Render per view) as one call over all the B200s of the box, from one C++ host process.
//
// Same set-up calls as ugu::RendererCpu (set_option / set_mesh / PrepareMesh); the cameras of a batch are
// passed as a list because every view is rendered concurrently (the reference's shared, mutated camera cannot
// describe that).  Views shard over the devices with no collective (the mesh is replicated at PrepareMesh); the
// only exchange is the optional gather of the face_id maps on one device, overlapped with rendering, over NCCL
// or peer copies (include/ugu_b200_raster.h, ugu_b200_multiview_*).  No CPU fallback: false + log when no B200
// is usable.
#pragma once

#include <memory>
#include <vector>

#include "ugu/camera.h"
#include "ugu/image.h"
#include "ugu/mesh.h"
#include "ugu/renderer/cpu/renderer.h"

namespace ugu {

class MultiViewRasterizerCuda {
  class Impl;
  std::unique_ptr<Impl> pimpl_;

 public:
  // devices: CUDA device ordinals, one rasterizer each (a device may repeat); empty = every usable B200
  explicit MultiViewRasterizerCuda(const RendererCpuOption& option = RendererCpuOption(),
                                   const std::vector<int>& devices = {});
  ~MultiViewRasterizerCuda();

  void set_option(const RendererCpuOption& option);
  void set_mesh(std::shared_ptr<const Mesh> mesh);
  bool PrepareMesh();

  // view i -> element i of every non-null output vector (resized to cameras.size(); images (re)allocated to the
  // camera size).  All cameras must be PinholeCameras of one image size.  All-null is an error, as in Render.
  bool RenderBatch(const std::vector<std::shared_ptr<const Camera>>& cameras, std::vector<Image3b>* color,
                   std::vector<Image1f>* depth, std::vector<Image3f>* normal, std::vector<Image1b>* mask,
                   std::vector<Image1i>* face_id, bool round_robin = false) const;

  // the face_id maps of all views gathered on device slot `root` (chunks sent while the next ones render),
  // then copied into `face_id`
  bool GatherFaceIds(const std::vector<std::shared_ptr<const Camera>>& cameras, std::vector<Image1i>* face_id,
                     int root = 0, int chunk_views = 0, bool round_robin = false) const;

  // per-face statistics over all views (pixels owned, views seen in, view that sees most of the face; -1 = never)
  // folded per device and reduced to `root`: 16 bytes per face cross the links instead of the maps
  bool FaceVisibility(const std::vector<std::shared_ptr<const Camera>>& cameras,
                      std::vector<uint32_t>* pixel_count, std::vector<uint32_t>* view_count,
                      std::vector<int32_t>* best_view, std::vector<uint32_t>* best_count, int root = 0,
                      int chunk_views = 0, bool round_robin = false) const;

  int size() const;                // rasterizers in use (0 = none could be created)
  bool valid() const;
  const char* transport() const;   // "nccl" or "peer"
};

}  // namespace ugu
