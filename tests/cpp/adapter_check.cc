// TEST HARNESS (not product): drives the reference ugu::Rasterizer and ugu::RasterizerCuda through
// the SAME ugu::RendererCpu* interface, in the call order of examples/ex02_renderer.cc:245-278 and
// the orbit loop of examples/ex09_renderer_realtime.cc:125-182, and compares all five outputs.
// Built only where the reference headers exist (tests/cpp/build_check.py); the binary travels
// to the GPU box and is run by tests/test_parity_gpu.py::test_cpp_adapter_vs_reference.
// Exit code 0 = parity (face_id / mask / depth / normal bit-exact, colour bit-exact or ±1 LSB for
// Oren-Nayar), 1 = mismatch, 2 = no CUDA rasterizer.
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <memory>
#include <vector>

#include "rasterizer_cuda.h"
#include "multiview_cuda.h"
#include "ugu_b200_raster.h"
#include "visibility_tester_cuda.h"
#include "ugu/camera.h"
#include "ugu/log.h"
#include "ugu/mesh.h"
#include "ugu/renderer/cpu/rasterizer.h"
#include "ugu/renderer/cpu/raytracer.h"
#include "ugu/util/raster_util.h"
#include "ugu/util/rgbd_util.h"
#include "ugu/util/camera_util.h"
#include "ugu/util/geom_util.h"

extern "C" void omp_set_num_threads(int);  // libgomp, linked through the compiled reference

namespace {

std::shared_ptr<ugu::Mesh> MakeTestMesh(int n_stacks, int n_slices) {
  ugu::MeshPtr mesh = ugu::MakeUvSphere(n_stacks, n_slices);
  const int V = static_cast<int>(mesh->vertices().size());
  // deterministic bumps so that depth ties are rare and silhouettes are irregular
  std::vector<Eigen::Vector3f> v = mesh->vertices();
  for (int i = 0; i < V; i++) {
    const float s = 1.0f + 0.08f * std::sin(7.0f * v[i].x()) * std::cos(5.0f * v[i].y() + 3.0f * v[i].z());
    v[i] *= s;
  }
  mesh->set_vertices(v);
  std::vector<Eigen::Vector2f> uv(V);
  std::vector<Eigen::Vector3f> col(V);
  for (int i = 0; i < V; i++) {
    uv[i] = Eigen::Vector2f(0.05f + 0.9f * (0.5f + 0.5f * v[i].x() / 1.1f),
                            0.05f + 0.9f * (0.5f + 0.5f * v[i].y() / 1.1f));
    uv[i] = uv[i].cwiseMax(0.02f).cwiseMin(0.98f);
    col[i] = Eigen::Vector3f(127.5f + 127.0f * std::sin(3.0f * v[i].x()),
                             127.5f + 127.0f * std::sin(4.0f * v[i].y()),
                             127.5f + 127.0f * std::sin(5.0f * v[i].z()));
  }
  mesh->set_uv(uv);
  mesh->set_uv_indices(mesh->vertex_indices());
  mesh->set_vertex_colors(col);
  std::vector<ugu::ObjMaterial> mats(1);
  mats[0].diffuse_tex = ugu::Image3b::zeros(256, 192);
  uint32_t s = 1u;
  for (int y = 0; y < 256; y++)
    for (int x = 0; x < 192; x++) {
      s ^= s << 13; s ^= s >> 17; s ^= s << 5;
      mats[0].diffuse_tex.at<ugu::Vec3b>(y, x) =
          ugu::Vec3b({uint8_t(s & 255u), uint8_t((s >> 8) & 255u), uint8_t((s >> 16) & 255u)});
    }
  mesh->set_materials(mats);
  mesh->set_material_ids(std::vector<int>(mesh->vertex_indices().size(), 0));
  mesh->CalcFaceNormal();
  mesh->CalcNormal();
  return mesh;
}

template <typename T>
long CountDiff(const ugu::Image<T>& a, const ugu::Image<T>& b, int tol) {
  if (a.rows != b.rows || a.cols != b.cols) return -1;
  const size_t n = static_cast<size_t>(a.rows) * a.cols * sizeof(T);
  long bad = 0;
  if (tol == 0) {
    // value compare on the raw element type would treat -0.0f == 0.0f; bytes are stricter, and the
    // normals of uncovered pixels are +0 in both, so compare floats by value, integers by bytes
    for (size_t i = 0; i < n; i++) bad += a.data[i] != b.data[i];
  } else {
    for (size_t i = 0; i < n; i++) bad += std::abs(int(a.data[i]) - int(b.data[i])) > tol;
  }
  return bad;
}

long CountDiffFloat(const float* a, const float* b, size_t n) {
  long bad = 0;
  for (size_t i = 0; i < n; i++) bad += !(a[i] == b[i]);
  return bad;
}

struct Outs {
  ugu::Image3b color;
  ugu::Image1f depth;
  ugu::Image3f normal;
  ugu::Image1b mask;
  ugu::Image1i face_id;
};

}  // namespace

int main(int argc, char** argv) {
  const bool timing = argc > 1 && std::strcmp(argv[1], "--time") == 0;
  ugu::set_log_level(ugu::LogLevel::kWarning);
  auto mesh = MakeTestMesh(timing ? 700 : 90, timing ? 900 : 120);
  const int W = timing ? 1920 : 320, H = timing ? 1080 : 240;
  auto camera = std::make_shared<ugu::PinholeCamera>(W, H, 50.0f);

  struct Case {
    const char* name;
    ugu::DiffuseColor color;
    ugu::ColorInterpolation interp;
    ugu::ShadingNormal normal;
    ugu::DiffuseShading shading;
    bool cull;
    int color_tol;
  };
  const Case cases[] = {
      {"vertex/vnormal/none/cull", ugu::DiffuseColor::kVertex, ugu::ColorInterpolation::kBilinear,
       ugu::ShadingNormal::kVertex, ugu::DiffuseShading::kNone, true, 0},
      {"texture-bilinear/vnormal/lambert/nocull", ugu::DiffuseColor::kTexture,
       ugu::ColorInterpolation::kBilinear, ugu::ShadingNormal::kVertex,
       ugu::DiffuseShading::kLambertian, false, 0},
      {"texture-nn/fnormal/none/cull", ugu::DiffuseColor::kTexture, ugu::ColorInterpolation::kNn,
       ugu::ShadingNormal::kFace, ugu::DiffuseShading::kNone, true, 0},
      {"white/vnormal/oren-nayar/cull", ugu::DiffuseColor::kNone, ugu::ColorInterpolation::kBilinear,
       ugu::ShadingNormal::kVertex, ugu::DiffuseShading::kOrenNayar, true, 1},
  };
  int failures = 0;
  for (const Case& cs : cases) {
    ugu::RendererCpuOption option;
    option.diffuse_color = cs.color;
    option.interp = cs.interp;
    option.shading_normal = cs.normal;
    option.diffuse_shading = cs.shading;
    option.backface_culling = cs.cull;
    // the two renderers behind the same interface pointer type
    std::unique_ptr<ugu::RendererCpu> ref = std::make_unique<ugu::Rasterizer>(option);
    auto cuda_owner = std::make_unique<ugu::RasterizerCuda>(option);
    if (!cuda_owner->valid()) {
      std::printf("no CUDA rasterizer available\n");
      return 2;
    }
    std::unique_ptr<ugu::RendererCpu> cuda = std::move(cuda_owner);
    for (ugu::RendererCpu* r : {ref.get(), cuda.get()}) {
      r->set_mesh(mesh);
      if (!r->PrepareMesh()) return 1;
      r->set_camera(camera);
    }
    const int n_views = timing ? 6 : 5;
    Outs o[2];  // reused across the views like a caller's frame loop (Render re-initialises them, util_private.cc:72-87)
    for (int k = 0; k < n_views; k++) {
      const double a = 2.0 * 3.14159265358979323846 * k / n_views + 0.3;
      Eigen::Matrix4d c2w;
      ugu::c2w(Eigen::Vector3d(2.9 * std::sin(a), 0.4 * std::cos(2 * a), 2.9 * std::cos(a)),
               Eigen::Vector3d(0.02, -0.03, 0.01), Eigen::Vector3d(0, -1, 0), &c2w);
      camera->set_c2w(Eigen::Affine3d(c2w));  // callers mutate the camera the renderers hold
      double ms[2];
      int idx = 0;
      for (ugu::RendererCpu* r : {ref.get(), cuda.get()}) {
        const auto t0 = std::chrono::steady_clock::now();
        if (!r->Render(&o[idx].color, &o[idx].depth, &o[idx].normal, &o[idx].mask, &o[idx].face_id)) return 1;
        ms[idx] = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        idx++;
      }
      const size_t P = static_cast<size_t>(W) * H;
      const long d_fid = CountDiff(o[0].face_id, o[1].face_id, 0);
      const long d_mask = CountDiff(o[0].mask, o[1].mask, 0);
      const long d_depth = CountDiffFloat(reinterpret_cast<float*>(o[0].depth.data),
                                          reinterpret_cast<float*>(o[1].depth.data), P);
      const long d_normal = CountDiffFloat(reinterpret_cast<float*>(o[0].normal.data),
                                           reinterpret_cast<float*>(o[1].normal.data), P * 3);
      const long d_color = CountDiff(o[0].color, o[1].color, cs.color_tol);
      long covered = 0;
      for (size_t i = 0; i < P; i++) covered += reinterpret_cast<int32_t*>(o[0].face_id.data)[i] >= 0;
      std::printf("%-42s view %d: covered %ld  diff face_id %ld mask %ld depth %ld normal %ld color %ld"
                  "   ref %.2f ms  cuda %.2f ms\n",
                  cs.name, k, covered, d_fid, d_mask, d_depth, d_normal, d_color, ms[0], ms[1]);
      if (d_fid || d_mask || d_depth || d_normal || d_color || covered == 0) failures++;
    }
    // single-output wrappers and RenderW
    ugu::Image1i f0, f1;
    ugu::Image1w w0, w1;
    if (!ref->RenderFaceId(&f0) || !cuda->RenderFaceId(&f1) || CountDiff(f0, f1, 0)) failures++;
    if (!ref->RenderDepthW(&w0) || !cuda->RenderDepthW(&w1) || CountDiff(w0, w1, 0)) failures++;
    if (cuda->Render(nullptr, nullptr, nullptr, nullptr, nullptr)) failures++;  // all-null is an error
    {  // the step after Render in ex02: Depth2PointCloud, reference vs CUDA, fresh and in/out images
      ugu::Image1f dref;
      if (!ref->RenderDepth(&dref)) failures++;
      for (int gl = 0; gl < 2; gl++) {
        ugu::Image3f p0, p1;
        if (!ugu::Depth2PointCloud(dref, *camera, &p0, gl != 0)) failures++;
        if (!static_cast<ugu::RasterizerCuda*>(cuda.get())->Depth2PointCloud(dref, *camera, &p1, gl != 0)) failures++;
        if (CountDiff(p0, p1, 0)) failures++;
        {  // the Mesh-output overloads (raster-order compaction, nearest colour)
          ugu::Mesh m0, m1, m2, m3;
          ugu::Image3b cref;
          if (!ref->RenderColor(&cref)) failures++;
          if (!ugu::Depth2PointCloud(dref, *camera, &m0, gl != 0)) failures++;
          if (!static_cast<ugu::RasterizerCuda*>(cuda.get())->Depth2PointCloud(dref, *camera, &m1, gl != 0)) failures++;
          if (!ugu::Depth2PointCloud(dref, cref, *camera, &m2, gl != 0)) failures++;
          if (!static_cast<ugu::RasterizerCuda*>(cuda.get())->Depth2PointCloud(dref, cref, *camera, &m3, gl != 0)) failures++;
          auto same = [](const std::vector<Eigen::Vector3f>& a, const std::vector<Eigen::Vector3f>& b) {
            return a.size() == b.size() && (a.empty() || std::memcmp(a.data(), b.data(), a.size() * sizeof(Eigen::Vector3f)) == 0);
          };
          if (m0.vertices().empty() || !same(m0.vertices(), m1.vertices()) || !same(m2.vertices(), m3.vertices()) ||
              !same(m2.vertex_colors(), m3.vertex_colors()) || m3.vertex_colors().size() != m3.vertices().size())
            failures++;
        }
        p0.setTo(ugu::Vec3f({1.f, 2.f, 3.f}));
        p1.setTo(ugu::Vec3f({1.f, 2.f, 3.f}));
        if (!ugu::Depth2PointCloud(dref, *camera, &p0, gl != 0)) failures++;
        if (!static_cast<ugu::RasterizerCuda*>(cuda.get())->Depth2PointCloud(dref, *camera, &p1, gl != 0)) failures++;
        if (CountDiff(p0, p1, 0)) failures++;
      }
    }
  }
  {  // UV-space rasterization: the reference templates (num_threads = 1) vs the CUDA adapter
    ugu::RasterizerCuda uv_owner{ugu::RendererCpuOption()};
    if (!uv_owner.valid()) return 2;
    auto* rc = &uv_owner;
    const auto& uv = mesh->uv();
    const auto& uvi = mesh->uv_indices();
    const auto& vi = mesh->vertex_indices();
    const auto& col = mesh->vertex_colors();
    {
      ugu::Image3b t0, t1;
      ugu::Image1b m0 = ugu::Image1b::zeros(300, 200), m1 = ugu::Image1b::zeros(300, 200);
      if (!ugu::RasterizeVertexAttributeToTexture(col, vi, uv, uvi, t0, 200, 300, &m0, 1)) failures++;
      if (!rc->RasterizeVertexAttributeToTexture(col, vi, uv, uvi, t1, 200, 300, &m1)) failures++;
      if (CountDiff(t0, t1, 0) || CountDiff(m0, m1, 0)) failures++;
      // over the existing texture (width/height <= 0), per-face attributes
      std::vector<Eigen::Vector3f> fa(uvi.size());
      for (size_t i = 0; i < fa.size(); i++) fa[i] = Eigen::Vector3f(float(i % 251), float(i % 13) * 20.5f, 300.0f - float(i % 7));
      std::vector<Eigen::Vector3i> half_i(uvi.begin(), uvi.begin() + uvi.size() / 2);
      std::vector<Eigen::Vector3f> half_a(fa.begin(), fa.begin() + fa.size() / 2);
      if (!ugu::RasterizeFaceAttributeToTexture(half_a, uv, half_i, t0, -1, -1, nullptr, 1)) failures++;
      if (!rc->RasterizeFaceAttributeToTexture(half_a, uv, half_i, t1, -1, -1, nullptr)) failures++;
      if (CountDiff(t0, t1, 0)) failures++;
    }
    {
      ugu::Image3f p0, p1;  // a position map, as app/posmap_gen.cc:74
      if (!ugu::RasterizeVertexAttributeToTexture(mesh->vertices(), vi, uv, uvi, p0, 257, 129, nullptr, 1)) failures++;
      if (!rc->RasterizeVertexAttributeToTexture(mesh->vertices(), vi, uv, uvi, p1, 257, 129, nullptr)) failures++;
      if (CountDiff(p0, p1, 0)) failures++;
    }
    {
      ugu::Image1b g0, g1;
      if (!ugu::GenerateUvMask(uv, uvi, g0, 200, 128, 96, 1)) failures++;
      if (!rc->GenerateUvMask(uv, uvi, g1, 200, 128, 96)) failures++;
      if (CountDiff(g0, g1, 0)) failures++;
      ugu::Image3b e;
      std::vector<Eigen::Vector3i> none;
      if (rc->RasterizeVertexAttributeToTexture(col, none, uv, none, e, 8, 8, nullptr)) failures++;  // false like the reference
      if (!e.empty()) failures++;
    }
  }
  {  // Depth2Mesh on a rendered view: the reference (rgbd_util.cc:94-282) vs the CUDA adapter, both overloads
    ugu::RendererCpuOption o;
    o.diffuse_color = ugu::DiffuseColor::kVertex;
    ugu::Rasterizer rr(o);
    rr.set_mesh(mesh);
    if (!rr.PrepareMesh()) return 1;
    rr.set_camera(camera);
    ugu::Image3b col;
    ugu::Image1f dep;
    if (!rr.Render(&col, &dep, nullptr, nullptr, nullptr)) return 1;
    ugu::RasterizerCuda rc{ugu::RendererCpuOption()};
    if (!rc.valid()) return 2;
    auto same3f = [](const std::vector<Eigen::Vector3f>& a, const std::vector<Eigen::Vector3f>& b) {
      return a.size() == b.size() && (a.empty() || std::memcmp(a[0].data(), b[0].data(), a.size() * 12) == 0);
    };
    auto same3i = [](const std::vector<Eigen::Vector3i>& a, const std::vector<Eigen::Vector3i>& b) {
      return a.size() == b.size() && (a.empty() || std::memcmp(a[0].data(), b[0].data(), a.size() * 12) == 0);
    };
    for (int variant = 0; variant < 4; variant++) {
      const int xs = variant == 1 ? 2 : 1, ys = variant == 2 ? 3 : 1;
      const bool gl = variant & 1, tex = variant >= 2, wvc = variant == 3;
      const float th = variant == 0 ? 0.01f : 0.2f;
      ugu::Mesh m0, m1;
      ugu::Image3f p0, p1, n0, n1;
      const bool ok0 = tex ? ugu::Depth2Mesh(dep, col, *camera, &m0, th, xs, ys, gl, "mat", wvc, &p0, &n0)
                           : ugu::Depth2Mesh(dep, *camera, &m0, th, xs, ys, gl, &p0, &n0);
      const bool ok1 = tex ? rc.Depth2Mesh(dep, col, *camera, &m1, th, xs, ys, gl, "mat", wvc, &p1, &n1)
                           : rc.Depth2Mesh(dep, *camera, &m1, th, xs, ys, gl, &p1, &n1);
      int bad = 0;
      if (!ok0 || !ok1) bad++;
      if (!same3f(m0.vertices(), m1.vertices()) || !same3i(m0.vertex_indices(), m1.vertex_indices())) bad++;
      if (!same3f(m0.normals(), m1.normals()) || !same3f(m0.face_normals(), m1.face_normals())) bad++;
      if (!same3i(m0.normal_indices(), m1.normal_indices()) || !same3i(m0.uv_indices(), m1.uv_indices())) bad++;
      if (!same3f(m0.vertex_colors(), m1.vertex_colors())) bad++;
      if (m0.uv().size() != m1.uv().size() ||
          (!m0.uv().empty() && std::memcmp(m0.uv()[0].data(), m1.uv()[0].data(), m0.uv().size() * 8) != 0)) bad++;
      if (m0.material_ids() != m1.material_ids() || m0.materials().size() != m1.materials().size()) bad++;
      if (CountDiff(p0, p1, 0) || CountDiff(n0, n1, 0)) bad++;
      std::printf("Depth2Mesh variant %d: %zu vertices %zu faces  %s\n", variant, m0.vertices().size(),
                  m0.vertex_indices().size(), bad ? "MISMATCH" : "identical");
      if (bad || m0.vertex_indices().empty()) failures++;
    }
    {  // the two-camera overload (rgbd_util.h:39-48): a wider colour camera, slightly rotated and shifted; the
       // organized images arrive with the depth image's size and stale contents, which the reference keeps
      auto color_camera = std::make_shared<ugu::PinholeCamera>(col.cols, col.rows, 60.0f);
      Eigen::Affine3f d2c = Eigen::Affine3f::Identity();
      d2c.linear() = (Eigen::AngleAxisf(0.03f, Eigen::Vector3f(0.2f, 1.0f, 0.1f).normalized())).toRotationMatrix();
      d2c.translation() = Eigen::Vector3f(0.02f, -0.01f, 0.015f);
      for (int wvc = 0; wvc < 2; wvc++) {
        ugu::Mesh m0, m1;
        ugu::Image3f p0, p1, n0, n1;
        ugu::Init(&p0, dep.cols, dep.rows, 7.0f);
        ugu::Init(&p1, dep.cols, dep.rows, 7.0f);
        const int xs = wvc ? 2 : 1;
        const bool ok0 = ugu::Depth2Mesh(dep, col, *camera, *color_camera, d2c, &m0, 0.05f, xs, 1, false, "mat",
                                         wvc != 0, &p0, &n0);
        const bool ok1 = rc.Depth2Mesh(dep, col, *camera, *color_camera, d2c, &m1, 0.05f, xs, 1, false, "mat",
                                       wvc != 0, &p1, &n1);
        int bad = 0;
        if (!ok0 || !ok1) bad++;
        if (!same3f(m0.vertices(), m1.vertices()) || !same3i(m0.vertex_indices(), m1.vertex_indices())) bad++;
        if (!same3f(m0.normals(), m1.normals()) || !same3f(m0.face_normals(), m1.face_normals())) bad++;
        if (!same3i(m0.uv_indices(), m1.uv_indices()) || !same3f(m0.vertex_colors(), m1.vertex_colors())) bad++;
        if (m0.uv().size() != m1.uv().size() ||
            (!m0.uv().empty() && std::memcmp(m0.uv()[0].data(), m1.uv()[0].data(), m0.uv().size() * 8) != 0)) bad++;
        if (CountDiff(p0, p1, 0) || CountDiff(n0, n1, 0)) bad++;
        std::printf("Depth2Mesh two cameras (vertex colours %d): %zu vertices %zu faces  %s\n", wvc,
                    m0.vertices().size(), m0.vertex_indices().size(), bad ? "MISMATCH" : "identical");
        if (bad || m0.vertex_indices().empty() || m0.uv().empty()) failures++;
      }
    }
    ugu::Mesh mm;
    if (rc.Depth2Mesh(dep, *camera, &mm, -1.0f)) failures++;  // false like the reference
  }
  {  // Raytracer parity mode: ugu::Raytracer (nanort BVH) vs ugu::RaytracerCuda behind the
     // same RendererCpu interface; all five outputs compared bit for bit (texture colour: nearest texel,
     // the unclamped bilinear taps of pixel_shader.h:318-343 read past the texture on this mesh)
    for (int variant = 0; variant < 2; variant++) {
      ugu::RendererCpuOption o;
      o.diffuse_color = variant == 0 ? ugu::DiffuseColor::kVertex : ugu::DiffuseColor::kTexture;
      o.interp = ugu::ColorInterpolation::kNn;
      o.shading_normal = variant == 0 ? ugu::ShadingNormal::kVertex : ugu::ShadingNormal::kFace;
      o.diffuse_shading = variant == 0 ? ugu::DiffuseShading::kLambertian : ugu::DiffuseShading::kNone;
      o.backface_culling = variant == 0;
      o.depth_scale = variant == 0 ? 1.0f : 1000.0f;
      std::unique_ptr<ugu::RendererCpu> ref = std::make_unique<ugu::Raytracer>(o);
      std::unique_ptr<ugu::RendererCpu> cuda = std::make_unique<ugu::RaytracerCuda>(o);
      for (ugu::RendererCpu* r : {ref.get(), cuda.get()}) {
        r->set_mesh(mesh);
        if (!r->PrepareMesh()) return 1;
        r->set_camera(camera);
      }
      for (int k = 0; k < 3; k++) {
        const double a = 2.0 * 3.14159265358979323846 * k / 3 + 0.7;
        Eigen::Matrix4d c2w;
        ugu::c2w(Eigen::Vector3d(2.9 * std::sin(a), 0.5 * std::cos(2 * a), 2.9 * std::cos(a)),
                 Eigen::Vector3d(0.02, -0.03, 0.01), Eigen::Vector3d(0, -1, 0), &c2w);
        camera->set_c2w(Eigen::Affine3d(c2w));
        Outs o0, o1;
        if (!ref->Render(&o0.color, &o0.depth, &o0.normal, &o0.mask, &o0.face_id)) return 1;
        if (!cuda->Render(&o1.color, &o1.depth, &o1.normal, &o1.mask, &o1.face_id)) return 1;
        const size_t P = static_cast<size_t>(W) * H;
        const long d_fid = CountDiff(o0.face_id, o1.face_id, 0), d_mask = CountDiff(o0.mask, o1.mask, 0);
        const long d_depth = CountDiffFloat(reinterpret_cast<float*>(o0.depth.data),
                                            reinterpret_cast<float*>(o1.depth.data), P);
        const long d_normal = CountDiffFloat(reinterpret_cast<float*>(o0.normal.data),
                                             reinterpret_cast<float*>(o1.normal.data), P * 3);
        const long d_color = CountDiff(o0.color, o1.color, 0);
        long covered = 0;
        for (size_t i = 0; i < P; i++) covered += reinterpret_cast<int32_t*>(o0.face_id.data)[i] >= 0;
        std::printf("Raytracer variant %d view %d: covered %ld  diff face_id %ld mask %ld depth %ld normal %ld "
                    "color %ld\n", variant, k, covered, d_fid, d_mask, d_depth, d_normal, d_color);
        if (d_fid || d_mask || d_depth || d_normal || d_color || covered == 0) failures++;
      }
      ugu::Image1w w0, w1;
      if (!ref->RenderDepthW(&w0) || !cuda->RenderDepthW(&w1) || CountDiff(w0, w1, 0)) failures++;
    }
  }
  {  // G-buffer: no GL reference can run here, so RasterizerCuda::RenderGBuffer is checked
     // for consistency with the reference Rasterizer on the same mesh / camera (winners, depth, colour)
    ugu::RendererCpuOption o;
    o.diffuse_color = ugu::DiffuseColor::kTexture;
    ugu::Rasterizer rr(o);
    rr.set_mesh(mesh);
    if (!rr.PrepareMesh()) return 1;
    rr.set_camera(camera);
    ugu::Image3b col;
    ugu::Image1f dep;
    ugu::Image1i fid;
    if (!rr.Render(&col, &dep, nullptr, nullptr, &fid)) return 1;
    ugu::RasterizerCuda rc(o);
    rc.set_mesh(mesh);
    if (!rc.PrepareMesh()) return 2;
    rc.set_camera(camera);
    ugu::GBuffer g;
    int bad = 0;
    if (!rc.RenderGBuffer(&g)) bad++;
    if (rc.RenderGBuffer(&g, 1.0f, 0.5f)) bad++;  // far < near -> false
    size_t covered = 0;
    const Eigen::Matrix3f R = camera->w2c().rotation().cast<float>();
    for (int y = 0; y < dep.rows && !bad; y++)
      for (int x = 0; x < dep.cols; x++) {
        const int f = fid.at<int>(y, x);
        const bool vis = g.stencil.at<uint8_t>(y, x) == 255;
        if (vis != (f >= 0) || g.face_id.at<int>(y, x) != (vis ? f + 1 : 0)) bad++;
        if (std::memcmp(&g.color.at<ugu::Vec3b>(y, x), &col.at<ugu::Vec3b>(y, x), 3) != 0) bad++;
        if (!vis) continue;
        covered++;
        const ugu::Vec3f pc = g.pos_cam.at<ugu::Vec3f>(y, x), nw = g.normal_wld.at<ugu::Vec3f>(y, x);
        const ugu::Vec3f nc = g.normal_cam.at<ugu::Vec3f>(y, x), b = g.bary.at<ugu::Vec3f>(y, x);
        const float d = dep.at<float>(y, x);
        if (std::abs(pc[2] - d) > 2e-5f * d) bad++;
        const Eigen::Vector3f ncr = R * Eigen::Vector3f(nw[0], nw[1], nw[2]);
        if (ncr[0] != nc[0] || ncr[1] != nc[1] || ncr[2] != nc[2]) bad++;  // renderer.cc:583, same Eigen expression
        if (std::abs(b[0] + b[1] + b[2] - 1.0f) > 1e-6f) bad++;
        const float d01 = g.depth_01.at<float>(y, x);
        if (!(d01 > 0.0f && d01 < 1.0f)) bad++;
      }
    std::printf("GBuffer vs Rasterizer: %zu covered pixels  %s\n", covered, bad ? "MISMATCH" : "consistent");
    if (bad || covered == 0) failures++;
  }
  if (!timing) {
    // ugu::VisibilityTesterCuda vs the reference ugu::VisibilityTester (BVH rays) on keyframes rendered by the
    // reference Rasterizer: both fill a ugu::VisibilityInfo through the same interface; compared are the visible
    // sets (the disagreement of sub-pixel rays vs nearest-pixel depth must stay below 2 %), the per-keyframe
    // records on the pairs both call visible, and the CalcStat ids on faces whose lists are identical.
    const std::vector<Eigen::Vector3f> eyes = {{0.f, 0.f, 2.7f}, {2.2f, 0.4f, 1.2f}, {-1.5f, -1.1f, -1.9f}, {0.3f, 2.6f, 0.5f}};
    std::vector<std::shared_ptr<ugu::Keyframe>> kfs;
    ugu::RendererCpuOption o;
    o.diffuse_color = ugu::DiffuseColor::kVertex;
    ugu::Rasterizer rr(o);
    rr.set_mesh(mesh);
    rr.PrepareMesh();
    for (size_t i = 0; i < eyes.size(); i++) {
      Eigen::Matrix4f c2w_f;
      ugu::c2w(eyes[i], Eigen::Vector3f(0.f, 0.f, 0.f), Eigen::Vector3f(0.f, 1.f, 0.1f).normalized(), &c2w_f);
      Eigen::Affine3d c2w = Eigen::Affine3d(c2w_f.cast<double>());
      auto cam = std::make_shared<ugu::PinholeCamera>(W, H, 50.0f);
      cam->set_c2w(c2w);
      rr.set_camera(cam);
      auto kf = ugu::Keyframe::Create();
      kf->id = 100 + 7 * static_cast<int>(i);
      kf->camera = cam;
      if (!rr.Render(&kf->color, &kf->depth, nullptr, &kf->mask, nullptr)) return 1;
      kfs.push_back(kf);
    }
    for (int facewise = 0; facewise < 2; facewise++) {
      omp_set_num_threads(1);  // the reference's face-wise bookkeeping is racy under OpenMP
      ugu::VisibilityTesterOption vo;
      ugu::VisibilityTester ref_t(vo);
      ref_t.set_data(mesh->vertices(), mesh->normals(), mesh->vertex_indices());
      ugu::VisibilityTesterCuda cu_t(vo);
      cu_t.set_data(mesh->vertices(), mesh->normals(), mesh->vertex_indices());
      ugu::VisibilityInfo ri(*mesh), ci(*mesh);
      if (!ref_t.PrepareData() || !cu_t.PrepareData()) return 2;
      if (!ref_t.Test(kfs, &ri, facewise != 0) || !cu_t.Test(kfs, &ci, facewise != 0)) return 1;
      long pairs = 0, differ = 0, rec_bad = 0, stat_faces = 0, stat_bad = 0;
      for (size_t v = 0; v < ri.vertex_info_list.size(); v++) {
        const auto& a = ri.vertex_info_list[v].visible_keyframes;
        const auto& b = ci.vertex_info_list[v].visible_keyframes;
        for (const auto& kf : kfs) {
          const int ia = ri.vertex_info_list[v].VisibleFrom(kf->id), ib = ci.vertex_info_list[v].VisibleFrom(kf->id);
          if (ia < 0 && ib < 0) continue;
          pairs++;
          if ((ia < 0) != (ib < 0)) {
            differ++;
            continue;
          }
          if (facewise) continue;  // a vertex may keep another corner's record after an upstream disagreement
          if (a[ia].projected_pos != b[ib].projected_pos || a[ia].color != b[ib].color ||
              a[ia].intensity != b[ib].intensity || std::abs(a[ia].viewing_angle - b[ib].viewing_angle) > 2e-6f)
            rec_bad++;
        }
      }
      for (size_t f = 0; f < ri.face_info_list.size(); f++) {
        const auto& a = ri.face_info_list[f].visible_keyframes;
        const auto& b = ci.face_info_list[f].visible_keyframes;
        bool same = a.size() == b.size() && !a.empty();
        for (size_t i = 0; same && i < a.size(); i++)
          same = a[i].kf_id == b[i].kf_id && a[i].area == b[i].area && a[i].viewing_angle == b[i].viewing_angle;
        if (!same) continue;
        stat_faces++;
        if (ri.face_info_list[f].max_area_id != ci.face_info_list[f].max_area_id ||
            ri.face_info_list[f].min_viewing_angle_id != ci.face_info_list[f].min_viewing_angle_id)
          stat_bad++;
      }
      const bool ok = pairs > 0 && differ * 50 <= pairs && rec_bad == 0 && stat_bad == 0 && stat_faces > 0 &&
                      ci.has_vertex_stat && ci.has_face_stat;
      std::printf("VisibilityTester %-10s: %ld (vertex, keyframe) pairs, %ld decided differently (%.2f %%), %ld record "
                  "mismatches, CalcStat ids on %ld identical face lists: %ld wrong  %s\n",
                  facewise ? "face-wise" : "vertex-wise", pairs, differ, 100.0 * differ / (pairs ? pairs : 1), rec_bad,
                  stat_faces, stat_bad, ok ? "ok" : "MISMATCH");
      if (!ok) failures++;
    }
  }

  {  // ---- the ex09 orbit loop (examples/ex09_renderer_realtime.cc:125-182) sharded over every visible GPU --------
    // config-3 shape: 36-view turntable at 1024 x 1024, textured, all five outputs.  N-device result == the
    // per-view Render of ONE RasterizerCuda, bit for bit, for both partitions; gathered face_id maps == the
    // rendered ones; the reduced per-face statistics == a histogram of those maps.
    const int MW = timing ? 1024 : 1024, MH = MW, NVIEWS = 36;
    ugu::RendererCpuOption option;
    option.diffuse_color = ugu::DiffuseColor::kTexture;
    std::vector<std::shared_ptr<const ugu::Camera>> cams;
    for (int k = 0; k < NVIEWS; k++) {
      const double a = 2.0 * 3.14159265358979323846 * k / NVIEWS;
      Eigen::Matrix4d c2w;
      ugu::c2w(Eigen::Vector3d(2.9 * std::sin(a), 0.0, 2.9 * std::cos(a)), Eigen::Vector3d(0, 0, 0),
               Eigen::Vector3d(0, -1, 0), &c2w);
      auto cm = std::make_shared<ugu::PinholeCamera>(MW, MH, 45.0f);
      cm->set_c2w(Eigen::Affine3d(c2w));
      cams.push_back(cm);
    }
    std::vector<int> devices;
    for (int i = 0; i < ugu_b200_device_count(); i++) devices.push_back(i);
    if (devices.size() == 1) devices.push_back(devices[0]);  // one GPU: two handles on it, same sharding logic
    ugu::MultiViewRasterizerCuda mv(option, devices);
    auto single = std::make_unique<ugu::RasterizerCuda>(option);
    bool ok = mv.valid() && single->valid();
    if (ok) {
      mv.set_mesh(mesh);
      single->set_mesh(mesh);
      ok = mv.PrepareMesh() && single->PrepareMesh();
    }
    std::vector<Outs> want(NVIEWS);
    for (int k = 0; ok && k < NVIEWS; k++) {
      single->set_camera(cams[k]);
      ok = single->Render(&want[k].color, &want[k].depth, &want[k].normal, &want[k].mask, &want[k].face_id);
    }
    long bad = 0;
    double ms_batch[2] = {0, 0};
    for (int rr = 0; ok && rr < 2; rr++) {
      std::vector<ugu::Image3b> c;
      std::vector<ugu::Image1f> d;
      std::vector<ugu::Image3f> n;
      std::vector<ugu::Image1b> m;
      std::vector<ugu::Image1i> f;
      const auto t0 = std::chrono::steady_clock::now();
      ok = mv.RenderBatch(cams, &c, &d, &n, &m, &f, rr != 0);
      ms_batch[rr] = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
      for (int k = 0; ok && k < NVIEWS; k++)
        bad += CountDiff(want[k].color, c[k], 0) + CountDiff(want[k].depth, d[k], 0) +
               CountDiff(want[k].normal, n[k], 0) + CountDiff(want[k].mask, m[k], 0) +
               CountDiff(want[k].face_id, f[k], 0);
    }
    long gbad = 0;
    double ms_gather = 0;
    for (int rr = 0; ok && rr < 2; rr++) {
      std::vector<ugu::Image1i> f;
      const int root = rr == 0 ? 0 : mv.size() - 1;
      const auto t0 = std::chrono::steady_clock::now();
      ok = mv.GatherFaceIds(cams, &f, root, rr == 0 ? 4 : 3, rr != 0);
      if (rr == 0) ms_gather = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
      for (int k = 0; ok && k < NVIEWS; k++) gbad += CountDiff(want[k].face_id, f[k], 0);
    }
    long vbad = 0;
    if (ok) {
      const size_t F = mesh->vertex_indices().size();
      std::vector<uint32_t> pix(F, 0), views(F, 0), bestc(F, 0);
      std::vector<int32_t> bestv(F, -1);
      std::vector<uint32_t> cnt(F);
      for (int k = 0; k < NVIEWS; k++) {
        std::fill(cnt.begin(), cnt.end(), 0u);
        const int32_t* fid = reinterpret_cast<const int32_t*>(want[k].face_id.data);
        for (size_t i = 0; i < static_cast<size_t>(MW) * MH; i++)
          if (fid[i] >= 0) cnt[fid[i]]++;
        for (size_t f = 0; f < F; f++) {
          if (!cnt[f]) continue;
          pix[f] += cnt[f];
          views[f]++;
          if (cnt[f] > bestc[f]) {  // ties keep the lowest view id
            bestc[f] = cnt[f];
            bestv[f] = k;
          }
        }
      }
      for (int rr = 0; ok && rr < 2; rr++) {
        std::vector<uint32_t> p2, v2, c2;
        std::vector<int32_t> b2;
        ok = mv.FaceVisibility(cams, &p2, &v2, &b2, &c2, rr == 0 ? 0 : mv.size() - 1, 5, rr != 0);
        for (size_t f = 0; ok && f < F; f++)
          vbad += (p2[f] != pix[f]) + (v2[f] != views[f]) + (b2[f] != bestv[f]) + (c2[f] != bestc[f]);
      }
    }
    std::printf("MultiViewRasterizerCuda: %d handle(s) on %d device(s), transport %s, %d views %dx%d: batch diff %ld "
                "(contiguous %.1f ms, round robin %.1f ms incl. host copies), gather diff %ld (%.1f ms), visibility "
                "diff %ld  %s\n",
                mv.size(), ugu_b200_device_count(), mv.transport(), NVIEWS, MW, MH, bad, ms_batch[0], ms_batch[1], gbad,
                ms_gather, vbad, (ok && !bad && !gbad && !vbad) ? "ok" : "MISMATCH");
    if (!ok || bad || gbad || vbad) failures++;
  }
  std::printf(failures ? "ADAPTER PARITY FAILED (%d)\n" : "ADAPTER PARITY OK\n", failures);
  return failures ? 1 : 0;
}
