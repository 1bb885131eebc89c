// sm_100a kernels of the keyframe visibility pass: the per-keyframe vertex / face tests of
// ugu::VisibilityTester (src/texturing/visibility_tester.cc:416-529 TestVertex, :531-549 TestVertices, :551-604
// TestFaces, :606-668 TestFacewiseVertices) and the folds of FaceInfo::CalcStat (:185-218) / the min / mean
// part of VertexInfo::CalcStat (:66-152), with ONE substitution: where the reference asks a BVH for the nearest
// hit of the ray through the vertex's sub-pixel projection (:457-483), these kernels read `zbuf` — the mesh's own
// depth map rendered from the keyframe camera by THIS rasterizer (back-face culling off) — at the nearest pixel:
//     t_v = |camera_p|                      distance of the (offset) vertex along its own ray
//     t_s = zbuf(nn) * (t_v / camera_p.z)   distance of the surface seen at that pixel along that ray
//     occluded  iff  zbuf(nn) >= FLT_MIN  and  t_v - t_s > eps        (reference: |hit - vertex| > eps)
//     distance  =  t_v                                               (reference: the BVH hit's t)
//     the depth check uses camera_p.z                                 (reference: the hit's camera z)
// Every other expression is the reference's, single round-to-nearest float operations in its order
// (--fmad=false): Affine3f * Vector3f = ((m0 x + m1 y) + m2 z) + t,
// PinholeCamera::Project (camera.h:380-387), ray_w(float, float) (camera.h:438-448), std::round / std::floor,
// BilinearInterpolation (util/image_util.h:156-193), Color2Gray in double (image_util.h:228-244).  acosf is the
// device libm's (<= 2 ulp from glibc's), so viewing_angle is held to a few ulp, everything else bit-exact against
// the C restatement (oracle_visibility_keyframe) and, on the pairs both call visible, against the compiled
// reference tester.
//
// Work items: a thread per vertex or per face, blockIdx.y = keyframe of the launch set (parameter blocks travel
// as a kernel-parameter array, like the view batches of raster_kernels.cuh).  Streaming, HBM-bound: per
// (vertex, keyframe) 24 bytes in, 33 bytes out and four image gathers.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace ugu_b200 {

constexpr int kVtMaxKeyframes = 16;  // keyframes per launch set
constexpr unsigned kVtNoOwner = 0xFFFFFFFFu;

struct VtCamera {
  int W, H;
  float fx, fy, cx, cy;
  float w2c[12];   // row-major 3x4: camera->w2c().cast<float>()
  float c2w_R[9];  // c2w_R_f_
};

struct VtKeyframe {
  VtCamera cam;
  const uint8_t* color;  // W*H*3
  const float* depth;    // W*H or null
  const uint8_t* mask;   // W*H or null
  const float* zbuf;     // W*H or null: the mesh's own depth map from this camera (occlusion test)
  unsigned order;        // position of the keyframe in the caller's sequence (tie-breaks of the folds)
  int pad;
};

struct VtBatch {
  VtKeyframe k[kVtMaxKeyframes];
};

struct VtOption {
  float backface_culling_angle_rad_th;
  int use_mask, use_depth, interp;
  float max_viewing_angle, max_distance, max_depth_difference, eps;
};

struct __align__(16) VtVertexRec {  // ugu::VertexInfoPerKeyframe without the id (visibility_tester.h:60-72)
  float color[3];
  float intensity;
  float projected[2];
  float viewing_angle;
  float distance;
};

struct VtFaceRec {  // ugu::FaceInfoPerKeyframe (visibility_tester.h:130-144)
  float area, viewing_angle, distance;
  float tri[6];
};

struct VtMesh {
  int V, F;
  const float* pos;  // V*3
  const float* nrm;  // V*3 (vertex normals, indexed like vertices)
  const int* idx;    // F*3
};

struct VtOut {  // dense per-keyframe records, keyframe kf of the launch set at row (row0 + kf)
  uint8_t* v_visible;   // rows * V
  VtVertexRec* v_rec;   // rows * V
  uint8_t* f_visible;   // rows * F
  VtFaceRec* f_rec;     // rows * F
  unsigned* owner;      // kVtMaxKeyframes * V scratch of the face-wise mode, kVtNoOwner between calls
  unsigned row0;
  // No dense face records wanted: the face kernels fold straight into the accumulators (VtAcc::f_*) with the packed
  // keys of vt_fold_face_kernel — no rows * F records written and read back, no fold launch.  NULL = write rows.
  unsigned* fold_count;
  unsigned long long* fold_max_area;
  unsigned long long* fold_min_angle;
  unsigned long long* fold_min_dist;
};

__device__ __forceinline__ unsigned vt_f2u8(float f) {  // static_cast<unsigned char>(float), x86-64 style
  if (!(f > -2147483904.0f && f < 2147483648.0f)) return 0u;
  return (unsigned)__float2int_rz(f) & 0xFFu;
}

// visibility_tester.cc:416-529 with the z-buffer substitution.  world_p = vertex + offset.
__device__ __forceinline__ bool vt_test_vertex(const VtKeyframe& kf, const VtOption& opt, float vx, float vy, float vz,
                                               float nx, float ny, float nz, float ox, float oy, float oz,
                                               VtVertexRec& out) {
  const VtCamera& c = kf.cam;
  const float wx = vx + ox, wy = vy + oy, wz = vz + oz;
  // Affine3f * Vector3f as Eigen evaluates it: ((m0 x + m1 y) + m2 z) + t (column-wise accumulation, then the
  // translation; probed against the compiled reference)
  const float px = ((c.w2c[0] * wx + c.w2c[1] * wy) + c.w2c[2] * wz) + c.w2c[3];
  const float py = ((c.w2c[4] * wx + c.w2c[5] * wy) + c.w2c[6] * wz) + c.w2c[7];
  const float pz = ((c.w2c[8] * wx + c.w2c[9] * wy) + c.w2c[10] * wz) + c.w2c[11];
  const float ix = c.fx / pz * px + c.cx;
  const float iy = c.fy / pz * py + c.cy;
  if (pz < 0.0f || ix < 0.0f || (float)(c.W - 1) < ix || iy < 0.0f || (float)(c.H - 1) < iy) return false;
  if (!(ix == ix) || !(iy == iy)) return false;  // NaN projections index nothing (undefined in the reference)
  const int nn_x = (int)roundf(ix), nn_y = (int)roundf(iy);
  const size_t nn = (size_t)nn_y * c.W + nn_x;
  if (opt.use_mask && __ldg(kf.mask + nn) != 255) return false;
  float kd = 0.0f;
  if (opt.use_depth) {
    kd = __ldg(kf.depth + nn);
    if (kd < 1.1920929e-7f) return false;  // std::numeric_limits<float>::epsilon()
  }
  float d0 = (ix - c.cx) / c.fx, d1 = (iy - c.cy) / c.fy, d2 = 1.0f;
  const float n2 = d0 * d0 + (d1 * d1 + d2 * d2);
  if (n2 > 0.0f) {
    const float s = sqrtf(n2);
    d0 = d0 / s;
    d1 = d1 / s;
    d2 = d2 / s;
  }
  const float r0 = c.c2w_R[0] * d0 + (c.c2w_R[1] * d1 + c.c2w_R[2] * d2);
  const float r1 = c.c2w_R[3] * d0 + (c.c2w_R[4] * d1 + c.c2w_R[5] * d2);
  const float r2 = c.c2w_R[6] * d0 + (c.c2w_R[7] * d1 + c.c2w_R[8] * d2);
  const float dot = -(nx * r0 + (ny * r1 + nz * r2));
  const float viewing_angle = acosf(dot);
  if (opt.backface_culling_angle_rad_th < viewing_angle) return false;
  const float t_v = sqrtf(px * px + (py * py + pz * pz));
  if (kf.zbuf) {
    const float z_s = __ldg(kf.zbuf + nn);
    if (z_s >= 1.17549435e-38f) {
      const float t_s = z_s * (t_v / pz);
      if (t_v - t_s > opt.eps) return false;
    }
  }
  if (opt.use_depth) {
    const float diff = fabsf(pz - kd);
    if (opt.max_depth_difference < diff) return false;
  }
  if (t_v > opt.max_distance || viewing_angle > opt.max_viewing_angle) return false;
  out.projected[0] = ix;
  out.projected[1] = iy;
  out.viewing_angle = viewing_angle;
  out.distance = t_v;
  if (opt.interp == 0) {
    const uint8_t* q = kf.color + nn * 3;
    out.color[0] = (float)__ldg(q);
    out.color[1] = (float)__ldg(q + 1);
    out.color[2] = (float)__ldg(q + 2);
  } else {  // BilinearInterpolation (image_util.h:156-193)
    int x0 = (int)floorf(ix), y0 = (int)floorf(iy);
    int x1 = x0 + 1, y1 = y0 + 1;
    if (x0 < 0) x0 = 0;
    if (y0 < 0) y0 = 0;
    if (c.W <= x1) x1 = c.W - 1;
    if (c.H <= y1) y1 = c.H - 1;
    const float lu = ix - (float)x0, lv = iy - (float)y0;
    const uint8_t* q00 = kf.color + ((size_t)y0 * c.W + x0) * 3;
    const uint8_t* q01 = kf.color + ((size_t)y0 * c.W + x1) * 3;
    const uint8_t* q10 = kf.color + ((size_t)y1 * c.W + x0) * 3;
    const uint8_t* q11 = kf.color + ((size_t)y1 * c.W + x1) * 3;
#pragma unroll
    for (int i = 0; i < 3; i++) {
      const float v = (1.0f - lu) * (1.0f - lv) * (float)__ldg(q00 + i) + lu * (1.0f - lv) * (float)__ldg(q01 + i) +
                      (1.0f - lu) * lv * (float)__ldg(q10 + i) + lu * lv * (float)__ldg(q11 + i);
      out.color[i] = (float)vt_f2u8(v);
    }
  }
  out.intensity = (float)(0.2989 * (double)out.color[0] + 0.5870 * (double)out.color[1] + 0.1140 * (double)out.color[2]);
  return true;
}

__device__ __forceinline__ void vt_store_vertex(VtVertexRec* dst, const VtVertexRec& r) {
  float4* q = reinterpret_cast<float4*>(dst);
  q[0] = make_float4(r.color[0], r.color[1], r.color[2], r.intensity);
  q[1] = make_float4(r.projected[0], r.projected[1], r.viewing_angle, r.distance);
}
__device__ __forceinline__ VtVertexRec vt_load_vertex(const VtVertexRec* src) {
  const float4* q = reinterpret_cast<const float4*>(src);
  const float4 a = q[0], b = q[1];
  VtVertexRec r;
  r.color[0] = a.x; r.color[1] = a.y; r.color[2] = a.z; r.intensity = a.w;
  r.projected[0] = b.x; r.projected[1] = b.y; r.viewing_angle = b.z; r.distance = b.w;
  return r;
}

// TestVertices (:531-549): a thread per vertex, blockIdx.y = keyframe.  Invisible rows are written as zeros.
__global__ void __launch_bounds__(256) vt_vertex_kernel(const VtBatch kb, const VtOption opt, const VtMesh m,
                                                        const VtOut out) {
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= m.V) return;
  const VtKeyframe& kf = kb.k[blockIdx.y];
  const size_t row = (size_t)(out.row0 + blockIdx.y) * m.V + v;
  VtVertexRec r;
  const bool ok = vt_test_vertex(kf, opt, __ldg(m.pos + 3 * (size_t)v), __ldg(m.pos + 3 * (size_t)v + 1),
                                 __ldg(m.pos + 3 * (size_t)v + 2), __ldg(m.nrm + 3 * (size_t)v),
                                 __ldg(m.nrm + 3 * (size_t)v + 1), __ldg(m.nrm + 3 * (size_t)v + 2), 0.f, 0.f, 0.f, r);
  if (!ok) r = VtVertexRec{};
  out.v_visible[row] = ok ? 1 : 0;
  vt_store_vertex(out.v_rec + row, r);
}

__device__ __forceinline__ void vt_store_face(const VtOut& out, size_t row, int f, unsigned order, bool ok,
                                              float angle, float dist, const float* tri) {
  if (out.fold_count != nullptr) {  // same values, same keys as vt_fold_face_kernel; max / min are order independent
    if (!ok) return;
    const float a = 0.5f * fabsf((tri[4] - tri[0]) * (tri[3] - tri[1]) - (tri[5] - tri[1]) * (tri[2] - tri[0]));
    atomicAdd(out.fold_count + f, 1u);
    atomicMax(out.fold_max_area + f, ((unsigned long long)__float_as_uint(a) << 32) | (0xFFFFFFFFu - order));
    atomicMin(out.fold_min_angle + f, ((unsigned long long)__float_as_uint(angle / 3.0f) << 32) | order);
    atomicMin(out.fold_min_dist + f, ((unsigned long long)__float_as_uint(dist / 3.0f) << 32) | order);
    return;
  }
  out.f_visible[row] = ok ? 1 : 0;
  VtFaceRec fr{};
  if (ok) {
    fr.viewing_angle = angle / 3.0f;
    fr.distance = dist / 3.0f;
    // 0.5f * |EdgeFunction(p0, p1, p2)| (raster_util.h:31-33)
    fr.area = 0.5f * fabsf((tri[4] - tri[0]) * (tri[3] - tri[1]) - (tri[5] - tri[1]) * (tri[2] - tri[0]));
#pragma unroll
    for (int k = 0; k < 6; k++) fr.tri[k] = tri[k];
  }
  out.f_rec[row] = fr;
}

// TestFaces (:551-604): a thread per face reads the records of its three vertices.
__global__ void __launch_bounds__(256) vt_face_kernel(const VtBatch kb, const VtMesh m, const VtOut out) {
  const int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= m.F) return;
  const size_t vrow = (size_t)(out.row0 + blockIdx.y) * m.V;
  float angle = 0.0f, dist = 0.0f, tri[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  bool ok = true;
#pragma unroll
  for (int j = 0; j < 3; j++) {
    const int vid = __ldg(m.idx + 3 * (size_t)f + j);
    if (ok && out.v_visible[vrow + vid]) {
      const VtVertexRec r = vt_load_vertex(out.v_rec + vrow + vid);
      angle += r.viewing_angle;
      dist += r.distance;
      tri[2 * j] = r.projected[0];
      tri[2 * j + 1] = r.projected[1];
    } else {
      ok = false;
    }
  }
  vt_store_face(out, (size_t)(out.row0 + blockIdx.y) * m.F + f, f, kb.k[blockIdx.y].order, ok, angle, dist, tri);
}

// TestFacewiseVertices (:606-668): a thread per face tests its corners, each pulled 1e-3 of the way toward the
// face centre, in order, stopping at the first failure; a passing (face, corner) bids for the vertex with
// atomicMin(face * 4 + corner) — the sequential "update only once per vertex" rule (the first passing test in
// face order) made order independent.  vt_owner_kernel then recomputes the winner's record.
__device__ __forceinline__ void vt_face_center(const VtMesh& m, int f, int* vid, float* vx, float* c) {
  c[0] = c[1] = c[2] = 0.0f;
#pragma unroll
  for (int j = 0; j < 3; j++) {
    vid[j] = __ldg(m.idx + 3 * (size_t)f + j);
#pragma unroll
    for (int k = 0; k < 3; k++) {
      vx[3 * j + k] = __ldg(m.pos + 3 * (size_t)vid[j] + k);
      c[k] += vx[3 * j + k];
    }
  }
#pragma unroll
  for (int k = 0; k < 3; k++) c[k] /= 3.0f;
}

__global__ void __launch_bounds__(256) vt_facewise_kernel(const VtBatch kb, const VtOption opt, const VtMesh m,
                                                          const VtOut out) {
  const int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= m.F) return;
  const VtKeyframe& kf = kb.k[blockIdx.y];
  int vid[3];
  float vx[9], c[3];
  vt_face_center(m, f, vid, vx, c);
  float angle = 0.0f, dist = 0.0f, tri[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  bool ok = true;
  for (int j = 0; j < 3 && ok; j++) {
    VtVertexRec r;
    const float* p = vx + 3 * j;
    ok = vt_test_vertex(kf, opt, p[0], p[1], p[2], __ldg(m.nrm + 3 * (size_t)vid[j]),
                        __ldg(m.nrm + 3 * (size_t)vid[j] + 1), __ldg(m.nrm + 3 * (size_t)vid[j] + 2),
                        (c[0] - p[0]) * 1e-3f, (c[1] - p[1]) * 1e-3f, (c[2] - p[2]) * 1e-3f, r);
    if (ok) {
      atomicMin(out.owner + (size_t)blockIdx.y * m.V + vid[j], (unsigned)f * 4u + (unsigned)j);
      angle += r.viewing_angle;
      dist += r.distance;
      tri[2 * j] = r.projected[0];
      tri[2 * j + 1] = r.projected[1];
    }
  }
  vt_store_face(out, (size_t)(out.row0 + blockIdx.y) * m.F + f, f, kb.k[blockIdx.y].order, ok, angle, dist, tri);
}

__global__ void __launch_bounds__(256) vt_owner_kernel(const VtBatch kb, const VtOption opt, const VtMesh m,
                                                       const VtOut out) {
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= m.V) return;
  const VtKeyframe& kf = kb.k[blockIdx.y];
  unsigned* own = out.owner + (size_t)blockIdx.y * m.V + v;
  const unsigned key = *own;
  const size_t row = (size_t)(out.row0 + blockIdx.y) * m.V + v;
  VtVertexRec r{};
  bool ok = false;
  if (key != kVtNoOwner) {
    *own = kVtNoOwner;  // leave the scratch clean
    int vid[3];
    float vx[9], c[3];
    vt_face_center(m, (int)(key >> 2), vid, vx, c);
    const float* p = vx + 3 * (key & 3u);
    ok = vt_test_vertex(kf, opt, p[0], p[1], p[2], __ldg(m.nrm + 3 * (size_t)v), __ldg(m.nrm + 3 * (size_t)v + 1),
                        __ldg(m.nrm + 3 * (size_t)v + 2), (c[0] - p[0]) * 1e-3f, (c[1] - p[1]) * 1e-3f,
                        (c[2] - p[2]) * 1e-3f, r);
    if (!ok) r = VtVertexRec{};  // cannot happen: the same test passed in vt_facewise_kernel
  }
  out.v_visible[row] = ok ? 1 : 0;
  vt_store_vertex(out.v_rec + row, r);
}

// ---- folds over the keyframes of a launch set, in keyframe order ---------------------------------------------
// Per face (FaceInfo::CalcStat, :185-218; std::max_element / min_element keep the FIRST extreme):
//   count, max_area key = area_bits << 32 | (0xFFFFFFFF - order)   (max; 0 = never visible)
//   min_angle key = angle_bits << 32 | order, min_dist key likewise (min; ~0 = never visible)
// Per vertex (VertexInfo::CalcStat, :66-152, the min / mean subset): count, float sums of colour / viewing angle /
// distance in keyframe order, min_angle / min_dist keys and the colour of each winner packed r | g << 8 | b << 16
// (keyframe colours are bytes).  One thread owns an element, so plain read-modify-write; partial accumulators of
// different ranks combine with SUM / MAX / MIN (+ the winner's colour from the rank that holds the key).
struct VtAcc {
  unsigned* f_count;
  unsigned long long* f_max_area;
  unsigned long long* f_min_angle;
  unsigned long long* f_min_dist;
  unsigned* v_count;
  float4* v_sum_color;                 // r g b, w unused
  float2* v_sum_ad;                    // viewing angle, distance
  unsigned long long* v_min_angle;
  unsigned long long* v_min_dist;
  uint2* v_min_colors;                 // packed colour of the min_angle / min_dist winner
};

__device__ __forceinline__ unsigned long long vt_key(float value, unsigned low) {
  return ((unsigned long long)__float_as_uint(value) << 32) | (unsigned long long)low;
}

__global__ void __launch_bounds__(256) vt_fold_face_kernel(const VtBatch kb, const int nk, const VtMesh m,
                                                           const VtOut out, const VtAcc acc) {
  const int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= m.F) return;
  unsigned count = acc.f_count[f];
  unsigned long long a = acc.f_max_area[f], g = acc.f_min_angle[f], d = acc.f_min_dist[f];
  for (int k = 0; k < nk; k++) {
    const size_t row = (size_t)(out.row0 + k) * m.F + f;
    if (!out.f_visible[row]) continue;
    const VtFaceRec r = out.f_rec[row];
    const unsigned order = kb.k[k].order;
    count++;
    const unsigned long long ka = vt_key(r.area, 0xFFFFFFFFu - order);
    const unsigned long long kg = vt_key(r.viewing_angle, order), kd = vt_key(r.distance, order);
    if (ka > a) a = ka;
    if (kg < g) g = kg;
    if (kd < d) d = kd;
  }
  acc.f_count[f] = count;
  acc.f_max_area[f] = a;
  acc.f_min_angle[f] = g;
  acc.f_min_dist[f] = d;
}

__global__ void __launch_bounds__(256) vt_fold_vertex_kernel(const VtBatch kb, const int nk, const VtMesh m,
                                                             const VtOut out, const VtAcc acc) {
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= m.V) return;
  unsigned count = acc.v_count[v];
  float4 sc = acc.v_sum_color[v];
  float2 sad = acc.v_sum_ad[v];
  unsigned long long g = acc.v_min_angle[v], d = acc.v_min_dist[v];
  uint2 mc = acc.v_min_colors[v];
  for (int k = 0; k < nk; k++) {
    const size_t row = (size_t)(out.row0 + k) * m.V + v;
    if (!out.v_visible[row]) continue;
    const VtVertexRec r = vt_load_vertex(out.v_rec + row);
    const unsigned order = kb.k[k].order;
    count++;
    sc.x += r.color[0];
    sc.y += r.color[1];
    sc.z += r.color[2];
    sad.x += r.viewing_angle;
    sad.y += r.distance;
    const unsigned packed = (unsigned)r.color[0] | ((unsigned)r.color[1] << 8) | ((unsigned)r.color[2] << 16);
    const unsigned long long kg = vt_key(r.viewing_angle, order), kd = vt_key(r.distance, order);
    if (kg < g) {
      g = kg;
      mc.x = packed;
    }
    if (kd < d) {
      d = kd;
      mc.y = packed;
    }
  }
  acc.v_count[v] = count;
  acc.v_sum_color[v] = sc;
  acc.v_sum_ad[v] = sad;
  acc.v_min_angle[v] = g;
  acc.v_min_dist[v] = d;
  acc.v_min_colors[v] = mc;
}

__global__ void vt_fill_u32(unsigned* p, size_t n, unsigned v) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}
__global__ void vt_fill_u64(unsigned long long* p, size_t n, unsigned long long v) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}

}  // namespace ugu_b200
