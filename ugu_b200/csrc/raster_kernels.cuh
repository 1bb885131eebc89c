// Hand-written sm_100a kernels of the B200-native mesh rasterizer.
//
// Every float expression is evaluated with single round-to-nearest IEEE operations in the
// association order of the compiled reference (SURVEY.md §8a "verified float op-order"), so the file
// MUST be compiled with --fmad=false and without -use_fast_math (see ugu_b200/build.py).  The
// visibility test is a 64-bit atomicMin on (float_bits(z) << 32 | face_id), which is
// order-isomorphic to the reference's sequential `z < d` rule with its lowest-face-index tie-break
// (src/renderer/cpu/rasterizer.cc:228-237) for every fragment with z >= FLT_MIN.
//
// Data layout in HBM (built once by upload, ugu_b200/csrc/raster_api.cu):
//   pos_x/pos_y/pos_z   V floats each (SoA planes, float4-vectorised coalesced loads)
//   idx0/idx1/idx2      F int32 each (SoA planes)
//   fn_x/fn_y/fn_z      F floats each (face normals, SoA planes)
//   nrm4 / col4         float4 per normal / per vertex colour (one LDG.128 per gather)
//   frec                2 x int4 per face = one 32-byte sector (one LDG.256) per winning fragment:
//                       {vidx0, vidx1, vidx2, material_id}, {nidx0, nidx1, nidx2, 0}
//   vattr               when normal_indices == vertex_indices: {nx ny nz 0, r g b 0} per vertex,
//                       32 bytes = one LDG.256 instead of two gathers
//   uvidx               int4 per face (3 indices + pad)
//   uv2                 float2 per uv
//   tex (per material)  uchar4 RGBA8 texels, linear index = y * w + x like Image3b::at
//   proj                float4 per vertex (image x, image y, camera z, 0), rewritten every view
//   vis                 uint64 per pixel, the packed visibility buffer
//                       float_bits(z) << 32 | face_id << 1 | backface; EMPTY = ~0
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace ugu_b200 {

constexpr unsigned long long kVisEmpty = 0xFFFFFFFFFFFFFFFFull;
constexpr int kCoopMaxSamplesLimit = 256;  // hard limit of the warp-cooperative path (8-bit dx/dy)
constexpr int kJobW = 8;              // large-triangle tile job: 8 x 4 pixels = one warp
constexpr int kJobH = 4;
constexpr int kQueueJobBits = 38;     // packed queue counter: [63:38] entries, [37:0] tile jobs
constexpr unsigned long long kQueueJobMask = (1ull << kQueueJobBits) - 1ull;
constexpr unsigned long long kQueueOverflowFlag = 1ull << 48;  // added to the hazard counter when the job field carries
constexpr int kJobsPerChunk = 16;     // tile jobs a warp claims at a time in the binned kernel
constexpr int kTile = 16;             // screen tile of the per-tile triangle lists: 16 x 16 pixels = one CTA of tile_kernel
constexpr int kTileShift = 4;
constexpr int kTileCap = 64;          // entries per tile list; a triangle that finds a list full takes the tile-job queue
constexpr int kTileStage = 64;        // triangles staged in shared memory at a time

struct ViewParams {
  int W, H;
  float fx, fy, cx, cy;
  float wm1, hm1;  // (float)(W - 1), (float)(H - 1)
  unsigned w_magic, w_shift;  // p / W == __umulhi(p, w_magic) >> w_shift for p < 2^31 (W > 1)
  float facing_eps;           // 64 * 2^-24 * max_i sum_j |c2w_R[i][j]|  (see facing_backface)
  float inv_fx, inv_fy;       // 1 / fx, 1 / fy for the approximate ray numerators of facing_backface
  float w2c_R[9];
  float w2c_t[3];
  float c2w_R[9];
  float eye[3];  // camera centre in world space (-c2w_R * w2c_t); only the group ordering heuristic reads it
};

// A batch of views rendered by ONE launch set (blockIdx.y = view of the batch; the vertex kernel loops
// over the views itself).  The parameter blocks travel as a kernel-parameter array: indexing it with
// the CTA-uniform view index compiles to LDCU c[0x0][UR + offset] — constant-bank speed, no local copy.
constexpr int kMaxBatch = 16;
template <int NV>
struct ViewBatch {
  ViewParams v[NV];
};

struct MeshDev {
  int V, F;
  int num_normals, num_colors, num_face_normals, num_uv, num_materials;
  const float* pos_x;
  const float* pos_y;
  const float* pos_z;
  const float4* pos4;   // the same positions as one 16-byte record per vertex (gathers by vertex index)
  const int* idx0;
  const int* idx1;
  const int* idx2;
  const float* fn_x;
  const float* fn_y;
  const float* fn_z;
  const float4* nrm4;
  const float4* col4;
  const float4* vattr;  // optional merged per-vertex record {nx ny nz 0, r g b 0} (32 B), see upload
  // raster order (slot -> face, see BuildFaceOrder in raster_api.cu): the streams raster_kernel reads with
  // coalesced loads, and the face id of each slot (NULL: slot == face)
  const int* r_idx0;
  const int* r_idx1;
  const int* r_idx2;
  const float* r_fn_x;
  const float* r_fn_y;
  const float* r_fn_z;
  const int* r_face;
  const float4* grp;    // per block of 256 consecutive slots: {sum of face normals, their dot with the mean centroid}
  // fused projection (raster_kernel): per block its <= 256 distinct vertices {x, y, z, global id | -1} and per slot
  // three 8-bit indices into them (0xFFFFFFFF: the block has more, its lanes project their own corners); NULL = off
  const float4* lv;
  const unsigned* lidx;
  const int4* frec;
  const int4* uvidx;
  const float2* uv2;
  const uchar4* const* tex;  // device array of per-material texel pointers
  const int2* tex_size;      // (w, h) per material
};

struct FrameDev {
  unsigned long long* host_counters;  // mapped pinned host memory (3 words) or null, see resolve_kernel
  float4* proj;                  // V
  float* ray_dx;                 // W: ((float)x - cx) / fx
  float* ray_dy;                 // H: ((float)y - cy) / fy
  unsigned long long* vis;       // W*H
  unsigned long long* counters;  // [0] packed large queue counter, [1] hazard fragments, [2] chunk cursor
  int* q_face;                   // F
  unsigned long long* q_job;     // F : first tile job of each queued triangle
  int coop_max;                  // bbox sample count up to which a triangle takes the cooperative path
  int* tile_cnt;                 // per-tile triangle lists (tile_kernel): entries per tile, NULL = off; left 0 by tile_kernel
  int* tile_bin;                 // kTileCap face ids per tile
  int tiles_x;                   // tiles per row
  int* order;                    // occlusion ordering of raster_kernel: block (256 slots) processed by CTA i; NULL = off
  // element strides between the per-view slabs of a batch (view v uses base + v * stride); the
  // counters slab of a view is 8 words: [0] packed large queue counter, [1] hazard fragments, [2] chunk
  // cursor, [3] / [7] camera-facing / deferred blocks of the occlusion ordering (vertex kernel), [4] / [5] the
  // view's (queue, hazards) as saved by resolve_kernel for the host.  Word 6 of slab 0: hazard fragments
  // summed over every view since the handle was created (never reset; the host reports deltas).
  unsigned long long s_proj, s_ray_dx, s_ray_dy, s_vis, s_q, s_order, s_tiles;
};
constexpr int kCounterSlab = 16;  // words [8] / [9]: arrive / leave counters of resolve_kernel's grid barrier
constexpr int kTotalsOffset = 6;

__device__ __forceinline__ FrameDev frame_view(const FrameDev& b, int view) {
  FrameDev f = b;
  const unsigned long long v = (unsigned long long)view;
  f.proj = b.proj + v * b.s_proj;
  f.ray_dx = b.ray_dx + v * b.s_ray_dx;
  f.ray_dy = b.ray_dy + v * b.s_ray_dy;
  f.vis = b.vis + v * b.s_vis;
  f.counters = b.counters + v * (unsigned long long)kCounterSlab;
  f.q_face = b.q_face + v * b.s_q;
  f.q_job = b.q_job + v * b.s_q;
  if (b.order) f.order = b.order + v * b.s_order;
  if (b.tile_cnt) {
    f.tile_cnt = b.tile_cnt + v * b.s_tiles;
    f.tile_bin = b.tile_bin + v * b.s_tiles * (unsigned long long)kTileCap;
  }
  return f;
}

struct ShadeParams {
  int diffuse_color, interp, shading_normal, diffuse_shading, backface_culling;
  float on_A, on_B;
  uint8_t* color;
  float* depth;
  float* normal;
  uint8_t* mask;
  int* face_id;
  // batch: view v of the launch set writes at element offset (first_view + v) * view_pixels of every output
  unsigned long long view_pixels;
  unsigned first_view;
  int fold_binned;  // no binned_kernel was launched for this launch set: the resolve kernel owns the tile-job queue
  // single view resolved in bands (host-form Render: the D2H of band k runs under the resolve of band k + 1): this
  // launch covers the 32-pixel chunks [chunk_begin, chunk_end); chunk_end == 0 = to the end of the image
  unsigned chunk_begin, chunk_end;
};

__device__ __forceinline__ ShadeParams shade_view(const ShadeParams& b, int view) {
  ShadeParams s = b;
  const unsigned long long off = (unsigned long long)(b.first_view + (unsigned)view) * b.view_pixels;
  if (b.color) s.color = b.color + off * 3ull;
  if (b.depth) s.depth = b.depth + off;
  if (b.normal) s.normal = b.normal + off * 3ull;
  if (b.mask) s.mask = b.mask + off;
  if (b.face_id) s.face_id = b.face_id + off;
  return s;
}

// ---------------------------------------------------------------------------------------------
// Exact-order float helpers (Eigen fixed-size reduction order: a0*b0 + (a1*b1 + a2*b2)).
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float dot3(float a0, float a1, float a2, float b0, float b1, float b2) {
  return a0 * b0 + (a1 * b1 + a2 * b2);
}

// 256-bit read-only global load (sm_100: LDG.E.256): one request for a whole 32-byte sector
__device__ __forceinline__ void ldg256(const void* p, float4& lo, float4& hi) {
  asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=f"(lo.x), "=f"(lo.y), "=f"(lo.z), "=f"(lo.w), "=f"(hi.x), "=f"(hi.y), "=f"(hi.z),
                 "=f"(hi.w)
               : "l"(p));
}

// Programmatic dependent launch (sm_90+): every per-view kernel is launched with
// cudaLaunchAttributeProgrammaticStreamSerialization, signals its dependents as soon as it starts
// (they are scheduled once the LAST CTA of this grid has started, i.e. into this grid's tail) and
// waits here for the complete, flushed results of its predecessors before touching per-view data.
// Both are no-ops for a plain launch.
__device__ __forceinline__ void pdl_launch_dependents() {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

struct Tri {
  float v0x, v0y, v0z, v1x, v1y, v1z, v2x, v2y, v2z;
  float e0x, e0y;  // v2 - v1
  float e1x, e1y;  // v0 - v2
  float e2x, e2y;  // v1 - v0
  float area;
};

__device__ __forceinline__ void tri_setup(Tri& t, const float4& a, const float4& b, const float4& c) {
  t.v0x = a.x; t.v0y = a.y; t.v0z = a.z;
  t.v1x = b.x; t.v1y = b.y; t.v1z = b.z;
  t.v2x = c.x; t.v2y = c.y; t.v2z = c.z;
  t.e0x = c.x - b.x; t.e0y = c.y - b.y;
  t.e1x = a.x - c.x; t.e1y = a.y - c.y;
  t.e2x = b.x - a.x; t.e2y = b.y - a.y;
  // EdgeFunction(v0, v1, v2) = (v2.x - v0.x)*(v1.y - v0.y) - (v2.y - v0.y)*(v1.x - v0.x)
  t.area = (c.x - a.x) * t.e2y - (c.y - a.y) * t.e2x;
}

// EdgeFunction(a, b, p) with (b - a) precomputed: (p.x - a.x)*(b.y - a.y) - (p.y - a.y)*(b.x - a.x)
__device__ __forceinline__ void tri_edges(const Tri& t, float px, float py, float& w0, float& w1,
                                          float& w2) {
  w0 = (px - t.v1x) * t.e0y - (py - t.v1y) * t.e0x;
  w1 = (px - t.v2x) * t.e1y - (py - t.v2y) * t.e1x;
  w2 = (px - t.v0x) * t.e2y - (py - t.v0y) * t.e2x;
}

// PinholeCamera::ray_w (camera.h:438-448): normalised camera ray rotated to world.
__device__ __forceinline__ void ray_world(const ViewParams& vp, float d0, float d1, float& r0,
                                          float& r1, float& r2) {
  float c0 = d0, c1 = d1, c2 = 1.0f;
  const float n2 = d0 * d0 + (d1 * d1 + 1.0f);  // d2*d2 == 1.0f exactly
  if (n2 > 0.0f) {
    const float s = sqrtf(n2);
    c0 = d0 / s;
    c1 = d1 / s;
    c2 = 1.0f / s;
  }
  r0 = dot3(vp.c2w_R[0], vp.c2w_R[1], vp.c2w_R[2], c0, c1, c2);
  r1 = dot3(vp.c2w_R[3], vp.c2w_R[4], vp.c2w_R[5], c0, c1, c2);
  r2 = dot3(vp.c2w_R[6], vp.c2w_R[7], vp.c2w_R[8], c0, c1, c2);
}

// rasterizer.cc:208-224 — perspective-correct depth and weights of a covered sample.
__device__ __forceinline__ float tri_depth(const Tri& t, float& w0, float& w1, float& w2) {
  w0 = w0 / t.area;
  w1 = w1 / t.area;
  w2 = w2 / t.area;
  w0 = w0 / t.v0z;
  w1 = w1 / t.v1z;
  w2 = w2 / t.v2z;
  const float z = 1.0f / ((w0 + w1) + w2);
  w0 = w0 * z;
  w1 = w1 * z;
  w2 = w2 * z;
  return z;
}

__device__ __forceinline__ void vis_update(unsigned long long* __restrict__ vis, size_t pix,
                                           unsigned long long seen, float z, int face,
                                           bool backface, unsigned long long* hazard) {
  // Fragments with z < FLT_MIN (zero, negative, denormal) or NaN make the reference's sequential
  // rule order dependent (rasterizer.cc:229, `d < FLT_MIN || z < d`); count them so the host can
  // report the view as off-contract instead of silently diverging.
  if (!(z >= 1.17549435e-38f)) {
    atomicAdd(hazard, 1ull);
    return;
  }
  // (z, face) orders exactly like the reference's sequential rule; the fragment's backface flag
  // (rasterizer.cc:200,235) rides in bit 0 — it is a function of (face, pixel), so it never
  // decides a comparison.
  const unsigned long long key =
      ((unsigned long long)__float_as_uint(z) << 32) |
      (unsigned long long)(((unsigned int)face << 1) | (backface ? 1u : 0u));
  // `seen` is an earlier read of vis[pix] (values only ever decrease): most occluded fragments
  // never issue the atomic.
  if (key < seen) atomicMin(vis + pix, key);
}

// Occlusion ordering (see raster_kernel): does the block face away from the camera?
__device__ __forceinline__ bool group_deferred(const ViewParams& vp, const float4& g) {
  // sum_f n_f . (centroid - eye) > 0: the group faces away from the camera (NaN -> not deferred)
  return g.w - ((g.x * vp.eye[0] + g.y * vp.eye[1]) + g.z * vp.eye[2]) > 0.0f;
}

// ---------------------------------------------------------------------------------------------
// K1: vertex transform + projection (rasterizer.cc:125-136, camera.h:380-387) and the per-view ray
// numerator tables.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float4 project_vertex(const ViewParams& vp, float x, float y, float z) {
  float px = dot3(vp.w2c_R[0], vp.w2c_R[1], vp.w2c_R[2], x, y, z);
  float py = dot3(vp.w2c_R[3], vp.w2c_R[4], vp.w2c_R[5], x, y, z);
  float pz = dot3(vp.w2c_R[6], vp.w2c_R[7], vp.w2c_R[8], x, y, z);
  px = px + vp.w2c_t[0];
  py = py + vp.w2c_t[1];
  pz = pz + vp.w2c_t[2];
  float4 o;
  o.x = vp.fx / pz * px + vp.cx;
  o.y = vp.fy / pz * py + vp.cy;
  o.z = pz;
  o.w = 0.0f;
  return o;
}

// Occlusion ordering of raster_kernel: camera-facing blocks fill the view's list from the front, the others
// from the back (one atomicAdd per warp and list; a warp's 32 consecutive blocks stay neighbours).  Thread t
// classifies block t for each of the n views; whole warps call it.
template <int NV>
__device__ __forceinline__ void classify_blocks(const ViewBatch<NV>& vb, const int n, const MeshDev& m,
                                                const FrameDev& fr, const int t) {
  const int nblocks = (m.F + 255) >> 8;
  if (fr.order == nullptr || (t & ~31) >= nblocks) return;
  const bool blk_valid = t < nblocks;
  const float4 rec = blk_valid ? __ldg(m.grp + t) : make_float4(0.f, 0.f, 0.f, 0.f);
  const unsigned lane = threadIdx.x & 31u;
  const unsigned lt = (1u << lane) - 1u;
  for (int v = 0; v < n; v++) {
    const bool def = blk_valid && group_deferred(vb.v[v], rec);
    const unsigned bf = __ballot_sync(0xffffffffu, blk_valid && !def), bd = __ballot_sync(0xffffffffu, def);
    unsigned long long* cnt = fr.counters + (unsigned long long)v * kCounterSlab;
    unsigned basef = 0u, based = 0u;
    if (lane == 0u) {
      if (bf) basef = (unsigned)atomicAdd(cnt + 3, (unsigned long long)__popc(bf));
      if (bd) based = (unsigned)atomicAdd(cnt + 7, (unsigned long long)__popc(bd));
    }
    basef = __shfl_sync(0xffffffffu, basef, 0);
    based = __shfl_sync(0xffffffffu, based, 0);
    int* order = fr.order + (unsigned long long)v * fr.s_order;
    if (blk_valid) {
      if (def) order[nblocks - 1 - (int)(based + __popc(bd & lt))] = t;
      else order[basef + __popc(bf & lt)] = t;
    }
  }
}

// One vertex per thread (fully coalesced 4-byte loads, 16-byte stores; the kernel is a few
// microseconds of pure latency at mesh sizes that fit one wave).  A batch loads every position once
// and projects it through each view of the batch: 12 bytes in, 16 bytes out per (vertex, view).
template <int NV>
__global__ void __launch_bounds__(256) vertex_kernel_b(const ViewBatch<NV> vb, const int nviews,
                                                       const MeshDev m, const FrameDev fr) {
  pdl_launch_dependents();
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  float x = 0.f, y = 0.f, z = 0.f;
  if (t < m.V) {  // the mesh is immutable: its loads may precede the wait on the previous batch's kernels
    x = __ldg(m.pos_x + t);
    y = __ldg(m.pos_y + t);
    z = __ldg(m.pos_z + t);
  }
  pdl_wait();
  const int n = (NV == 1) ? 1 : nviews;
  for (int v = 0; v < n; v++) {
    const ViewParams& vp = vb.v[v];
    if (t < m.V) fr.proj[(unsigned long long)v * fr.s_proj + t] = project_vertex(vp, x, y, z);
    // ray tables: the first W + H threads
    if (t < vp.W) fr.ray_dx[(unsigned long long)v * fr.s_ray_dx + t] = ((float)t - vp.cx) / vp.fx;
    else if (t < vp.W + vp.H)
      fr.ray_dy[(unsigned long long)v * fr.s_ray_dy + (t - vp.W)] = ((float)(t - vp.W) - vp.cy) / vp.fy;
  }
  classify_blocks<NV>(vb, n, m, fr, t);
}

// The block lists alone, for the fused projection (no vertex kernel).
template <int NV>
__global__ void __launch_bounds__(256) classify_kernel(const ViewBatch<NV> vb, const int nviews, const MeshDev m,
                                                       const FrameDev fr) {
  pdl_launch_dependents();
  pdl_wait();  // counters [3] / [7] are reset by the previous view's resolve
  classify_blocks<NV>(vb, (NV == 1) ? 1 : nviews, m, fr, (int)(blockIdx.x * blockDim.x + threadIdx.x));
}

// ---------------------------------------------------------------------------------------------
// Triangle setup shared by K2 and K3: rejects + clamped bbox (rasterizer.cc:168-193).
// Returns false when the triangle is rejected or has no pixel sample inside its bbox.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ int sat_int(float f) {
  // cvt.rzi.s32.f32 saturates and maps NaN to 0; the clamps below make the result identical to the
  // reference for every on-contract (finite, in-int-range) coordinate.
  return __float2int_rz(f);
}

__device__ __forceinline__ bool tri_bounds(const Tri& t, const ViewParams& vp, int& x0, int& y0,
                                           int& x1, int& y1) {
  if (t.v0z < 0.0f || t.v1z < 0.0f || t.v2z < 0.0f) return false;
  const float xmin = fminf(fminf(t.v0x, t.v1x), t.v2x);
  const float ymin = fminf(fminf(t.v0y, t.v1y), t.v2y);
  const float xmax = fmaxf(fmaxf(t.v0x, t.v1x), t.v2x);
  const float ymax = fmaxf(fmaxf(t.v0y, t.v1y), t.v2y);
  if (xmin > vp.wm1 || xmax < 0.0f || ymin > vp.hm1 || ymax < 0.0f) return false;
  x0 = max(0, sat_int(ceilf(xmin)));
  x1 = min(vp.W - 1, sat_int(floorf(xmax)));
  y0 = max(0, sat_int(ceilf(ymin)));
  y1 = min(vp.H - 1, sat_int(floorf(ymax)));
  if (fabsf(t.area) < 1.17549435e-38f) return false;
  return x1 >= x0 && y1 >= y0;
}

// ---------------------------------------------------------------------------------------------
// Occlusion ordering (PASS 1 of raster_kernel).
//
// Back faces take part in the reference's depth test (rasterizer.cc:198-200,228-237), so nothing may be
// culled by orientation.  But ORDER is free — the 64-bit atomicMin is order independent — and a fragment
// that provably loses against what is already stored need not be produced at all.  The vertex kernel sorts
// the blocks of 256 faces (= one CTA) into two lists per view: blocks whose summed face normal points
// towards the camera (on a closed mesh: the visible side) are dispatched first, the others ("deferred")
// after them, and a deferred block drops a triangle in phase A — before its shared-memory record, its bbox
// samples and its fragments exist — when
//   (i)  |area| >= 0.004 * ex * ey   (ex, ey = extents of its image-space bbox), and
//   (ii) every pixel of its (at most 4 x 4) bbox already stores a depth zs < 0.998 * min(z0, z1, z2).
// Whatever camera-facing block is still in flight when a deferred block looks only makes (ii) fail more
// often; which blocks are deferred, and when, affects speed only.
// Proof that every fragment of such a triangle loses.  Let E = ex * ey, u = 2^-24.  For an integer sample p
// inside the bbox all coordinate differences are bounded by ex / ey, so each float edge function w_k and
// `area` carry an absolute error <= 8 u E, and (the three exact edge functions sum to the exact area)
// the float sum s3 = (w0 + w1) + w2 differs from `area` by <= 44 u E <= 0.001 |area| under (i) (44 u / 0.004
// = 6.6e-4).  So a covered sample (all w_k >= 0 or all <= 0) cannot have all w_k == 0, `area` has the sign of
// the w_k, and s3 <= 1.001 |area|: the premises of raster_fragments' early-out, whose bound gives
// z >= 0.998999 * min z_k > 0.998f * min z_k > zs for the reference's z = 1 / sum_k (w_k / area / z_k).  Stored
// depths only decrease, so the fragment is strictly behind the final winner whatever its facing (and it is
// not a hazard fragment: zs >= FLT_MIN).
// ---------------------------------------------------------------------------------------------

// 0.998 * min z of a triangle that satisfies (i), as float bits (positive floats order like unsigned
// integers; EMPTY pixels read 0xFFFFFFFF and never pass); 0 when the triangle may not be dropped.
__device__ __forceinline__ unsigned tri_cull_bits(const Tri& t) {
  const float ex = fmaxf(fmaxf(t.v0x, t.v1x), t.v2x) - fminf(fminf(t.v0x, t.v1x), t.v2x);
  const float ey = fmaxf(fmaxf(t.v0y, t.v1y), t.v2y) - fminf(fminf(t.v0y, t.v1y), t.v2y);
  const float E = ex * ey;
  const float zmin = fminf(fminf(t.v0z, t.v1z), t.v2z);
  if (!(E > 1e-25f && E < 1e25f && fabsf(t.area) >= 0.004f * E && zmin > 1e-30f && zmin < 1e30f)) return 0u;
  return __float_as_uint(0.998f * zmin);
}

// (ii): is every pixel of [x0, x0 + nx) x [y0, y0 + ny), nx, ny <= 4, nearer than `zbits`?
__device__ __forceinline__ bool bbox_occluded(const FrameDev& fr, int W, int x0, int y0, int nx, int ny,
                                              unsigned zbits) {
  const int xa = x0 + min(1, nx - 1), xb = x0 + min(2, nx - 1), xc = x0 + nx - 1;
  for (int yy = y0; yy < y0 + ny; yy++) {
    // high word of the packed key = float bits of the stored depth (plain loads: per-view data)
    const unsigned* row = reinterpret_cast<const unsigned*>(fr.vis + (size_t)yy * W) + 1;
    const unsigned m = max(max(row[2 * x0], row[2 * xa]), max(row[2 * xb], row[2 * xc]));
    if (m >= zbits) return false;
  }
  return true;
}

// ---------------------------------------------------------------------------------------------
// K2: triangle setup + warp-cooperative fine raster.
//
// A warp takes a group of 64 consecutive faces.  Phase A (lane = face, 2 faces per lane): gather
// the projected vertices, apply the reference's rejects, and either (a) compact the triangle into
// the warp's shared-memory record table when its clamped bbox holds <= coop_max samples, or (b)
// append it to the large-triangle queue (warp prefix sum over a packed (entries, tile jobs)
// counter, one atomicAdd per warp).  Phase B (lane = bbox sample): the samples of all recorded
// triangles form one index space [0, T) that the warp walks 32 at a time, so lanes stay busy
// whatever the mix of triangle sizes; the owner of a sample is found from a REDUX.OR bitmask of the
// triangle start offsets.  Samples that pass the (cheap) inclusive edge test are compacted into a
// ring queue, and the expensive part — per-pixel world ray, facing, the seven IEEE divisions of the
// perspective-correct depth, the 64-bit atomicMin — runs on full warps of queued fragments.
// ---------------------------------------------------------------------------------------------
// HALVES = faces per lane in a group (group = 32 * HALVES faces per warp)
template <int HALVES>
struct __align__(16) WarpTris {
  // per recorded triangle (slot), SoA of 16-byte fields -> one LDS.128 each
  float4 a[32 * HALVES];  // v0x v0y v1x v1y
  float4 b[32 * HALVES];  // v2x v2y (int)(x0 | y0 << 16) (int)face
  float4 c[32 * HALVES];  // v0z v1z v2z area
  float4 d[32 * HALVES];  // g = c2w_R^T * face_normal (3), facing threshold
  int2 meta[32 * HALVES];  // first sample of the triangle in the group's sample space; nx | ceil(65536 / nx) << 9
  uint4 q[64];  // fragment ring: {slot | dx << 6 | dy << 14 | pos << 22 | neg << 23, w0, w1, w2}
};

// exact int -> float for 0 <= v < 2^23 without the (quarter-rate) I2F
__device__ __forceinline__ float small_int_to_float(int v) {
  return __uint_as_float(0x4B000000u | (unsigned)v) - 8388608.0f;
}

// Per-pixel facing (rasterizer.cc:197-200): backface = face_normal . ray_w(x, y) > 0 with
// ray_w = c2w_R * normalize(d), d = ((x - cx) / fx, (y - cy) / fy, 1).  Only the SIGN of that float
// expression is needed, and it equals the sign of val = g . da with g = c2w_R^T * face_normal and
// da = ((x - cx) * (1 / fx), (y - cy) * (1 / fy), 1) whenever |val| is clear of the rounding error
// of both evaluations (u = 2^-24, rho = max row 1-norm of c2w_R, D = |d|, dm = max(|d0|, |d1|, 1)):
//   reference    |t   - T|   <= 10.2 u rho |f|_1       t its float result, T the real value
//   cheap path   |val - D T| <= (6.1 + 2.5) u rho |f|_1 dm   (2.5 u: da versus d), D <= sqrt(3) dm
// so |val| > 26.3 u rho |f|_1 dm implies sign(t) == sign(val) and t != 0.  The stored threshold
// uses 64 instead of 26.3 (and an absolute floor for the denormal range); anything closer to
// edge-on, or NaN, evaluates the reference's expression verbatim (from the exact ray tables).
__device__ __forceinline__ bool facing_backface(const ViewParams& vp, const MeshDev& m,
                                                const FrameDev& fr, const float4& g, int face,
                                                int x, int y) {
  const float d0 = (small_int_to_float(x) - vp.cx) * vp.inv_fx;
  const float d1 = (small_int_to_float(y) - vp.cy) * vp.inv_fy;
  const float val = (g.x * d0 + g.y * d1) + g.z;
  const float dm = fmaxf(fmaxf(fabsf(d0), fabsf(d1)), 1.0f);
  if (fabsf(val) > g.w * dm) return val > 0.0f;
  float r0, r1, r2;
  // the ray-table expressions of vertex_kernel_b, evaluated here: the fused projection runs no vertex kernel
  ray_world(vp, (small_int_to_float(x) - vp.cx) / vp.fx, (small_int_to_float(y) - vp.cy) / vp.fy, r0, r1, r2);
  return dot3(__ldg(m.fn_x + face), __ldg(m.fn_y + face), __ldg(m.fn_z + face), r0, r1, r2) > 0.0f;
}

template <int HALVES>
__device__ __forceinline__ void raster_fragments(const ViewParams& vp, const MeshDev& m,
                                                 const FrameDev& fr, WarpTris<HALVES>& S, bool has_fn,
                                                 int qhead, int n, unsigned lane) {
  if ((int)lane < n) {
    const uint4 e = S.q[(qhead + (int)lane) & 63];
    const unsigned packed = e.x;
    float w0 = __uint_as_float(e.y), w1 = __uint_as_float(e.z), w2 = __uint_as_float(e.w);
    const int slot = (int)(packed & 63u);
    const float4 rb = S.b[slot];
    const unsigned xy = __float_as_uint(rb.z);
    const int x = (int)(xy & 0xFFFFu) + (int)((packed >> 6) & 255u);
    const int y = (int)(xy >> 16) + (int)((packed >> 14) & 255u);
    const int face = __float_as_int(rb.w);
    const size_t pix = (size_t)y * vp.W + x;
    // issued now, consumed after the divisions: a stale value only makes the early-outs less sharp
    const unsigned long long seen = fr.vis[pix];
    const float4 rc = S.c[slot];
    const bool pos = (packed >> 22) & 1u, neg = (packed >> 23) & 1u;

    // Conservative early-out.  For a one-orientation sample with |w0+w1+w2| <= 1.001 |area| (same
    // sign), the reference's z = 1 / (sum_k (w_k / area) / z_k) is >= 0.998999 * min z_k (each
    // correctly rounded op moves a bound by at most 2^-24), so a fragment whose 0.998 * zmin is
    // already behind the stored depth cannot win — whatever its facing — and the stored depth only
    // ever decreases.
    {
      const float zs = __uint_as_float((unsigned)(seen >> 32));  // NaN while the pixel is EMPTY
      const float zmin = fminf(fminf(rc.x, rc.y), rc.z);
      const float s3 = (w0 + w1) + w2;
      const float sa = pos ? s3 : -s3, aa = pos ? rc.w : -rc.w;
      if (pos != neg && aa > 0.0f && sa <= 1.001f * aa && zmin < 1e30f && 0.998f * zmin > zs) return;
    }
    bool backface = false;
    if (has_fn)
      backface = facing_backface(vp, m, fr, S.d[slot], face, x, y);
    if (backface ? neg : pos) {
      // rasterizer.cc:208-224
      w0 = w0 / rc.w;
      w1 = w1 / rc.w;
      w2 = w2 / rc.w;
      w0 = w0 / rc.x;
      w1 = w1 / rc.y;
      w2 = w2 / rc.z;
      const float z = 1.0f / ((w0 + w1) + w2);
      vis_update(fr.vis, pix, seen, z, face, backface, fr.counters + 1);
    }
  }
}

// Everything a warp does for its 32 * HALVES faces in ONE view once their indices, (optional) normals and projected
// vertices are in registers: rejects, occlusion test of deferred blocks, queueing of large triangles, the
// shared-memory table, phase B and the fragment stage.
template <int HALVES, int PASS, bool TILES>
__device__ __forceinline__ void raster_warp_view(const ViewParams& vp, const MeshDev& m, const FrameDev& fr,
                                                 WarpTris<HALVES>& S, const unsigned lane, const bool has_fn,
                                                 const bool deferred, const int (&fidx)[HALVES], int (&face)[HALVES],
                                                 float (&fnx)[HALVES], float (&fny)[HALVES], float (&fnz)[HALVES],
                                                 const float4 (&pa)[HALVES], const float4 (&pb)[HALVES],
                                                 const float4 (&pc)[HALVES]) {
  const unsigned lt_mask = (1u << lane) - 1u;
  int nslots = 0, T = 0;
  unsigned long long my[HALVES];  // packed (1 entry, njobs) when the face is queued
#pragma unroll
  for (int h = 0; h < HALVES; h++) {
    my[h] = 0ull;
    const int f = fidx[h];
    int cnt = 0;
    int x0 = 0, y0 = 0, nx = 1;
    Tri t;
    if (f < m.F) {
      tri_setup(t, pa[h], pb[h], pc[h]);
      int x1, y1;
      if (tri_bounds(t, vp, x0, y0, x1, y1)) {
        nx = x1 - x0 + 1;
        const int ny = y1 - y0 + 1;
        const long long ns = (long long)nx * ny;
        if (ns <= (long long)fr.coop_max) {
          cnt = (int)ns;
          if (PASS != 0 && deferred && nx <= 4 && ny <= 4) {
            const unsigned zbits = tri_cull_bits(t);
            if (zbits != 0u && bbox_occluded(fr, vp.W, x0, y0, nx, ny, zbits)) cnt = 0;
          }
        } else {
          bool listed = false;
          if (TILES) {
            const int tx0 = x0 >> kTileShift, ty0 = y0 >> kTileShift;
            const int ntx = (x1 >> kTileShift) - tx0 + 1, nty = (y1 >> kTileShift) - ty0 + 1;
            if (ntx <= 3 && nty <= 3) {
              const int fid = (PASS != 0 && deferred && m.r_face) ? __ldg(m.r_face + f) : face[h];
              int pos[9];  // all the appends in flight together: one round trip, not nine
#pragma unroll
              for (int k = 0; k < 9; k++) {
                const int dx = k % 3, dy = k / 3;
                pos[k] = (dx < ntx && dy < nty) ? atomicAdd(fr.tile_cnt + (ty0 + dy) * fr.tiles_x + tx0 + dx, 1) : -1;
              }
              listed = true;
#pragma unroll
              for (int k = 0; k < 9; k++) {
                const int dx = k % 3, dy = k / 3;
                if (dx < ntx && dy < nty) {
                  if (pos[k] < kTileCap)
                    fr.tile_bin[(size_t)((ty0 + dy) * fr.tiles_x + tx0 + dx) * kTileCap + pos[k]] = fid;
                  else
                    listed = false;  // (tiles it did enter may rasterise it twice: the atomicMin is idempotent)
                }
              }
            }
          }
          if (!listed) {
            const unsigned long long njobs = (unsigned long long)((nx + kJobW - 1) / kJobW) *
                                             (unsigned long long)((ny + kJobH - 1) / kJobH);
            my[h] = (1ull << kQueueJobBits) | njobs;
          }
        }
      }
    }
    if (PASS != 0 && deferred && (cnt > 0 || my[h] != 0ull)) {
      if (m.r_face) face[h] = __ldg(m.r_face + f);
      if (has_fn) {
        fnx[h] = __ldg(m.r_fn_x + f);
        fny[h] = __ldg(m.r_fn_y + f);
        fnz[h] = __ldg(m.r_fn_z + f);
      }
    }
    // slot = rank among the recorded triangles; start = exclusive prefix of the sample counts
    const unsigned rec = __ballot_sync(0xffffffffu, cnt > 0);
    if (rec == 0u) continue;  // (warp-uniform) nothing to record: sub-pixel, rejected or occluded triangles only
    int incl = cnt;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int o = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= (unsigned)d) incl += o;
    }
    if (cnt > 0) {
      const int slot = nslots + __popc(rec & lt_mask);
      S.a[slot] = make_float4(t.v0x, t.v0y, t.v1x, t.v1y);
      S.b[slot] = make_float4(t.v2x, t.v2y, __int_as_float(x0 | (y0 << 16)), __int_as_float(face[h]));
      S.c[slot] = make_float4(t.v0z, t.v1z, t.v2z, t.area);
      // g = c2w_R^T * fn and the facing threshold (see facing_backface)
      S.d[slot] = make_float4(
          (vp.c2w_R[0] * fnx[h] + vp.c2w_R[3] * fny[h]) + vp.c2w_R[6] * fnz[h],
          (vp.c2w_R[1] * fnx[h] + vp.c2w_R[4] * fny[h]) + vp.c2w_R[7] * fnz[h],
          (vp.c2w_R[2] * fnx[h] + vp.c2w_R[5] * fny[h]) + vp.c2w_R[8] * fnz[h],
          fmaxf(vp.facing_eps * ((fabsf(fnx[h]) + fabsf(fny[h])) + fabsf(fnz[h])), 1e-30f));
      // ceil(65536 / nx) through an exact float division (nx <= 256): no integer-divide sequence
      const int recip = (int)(65536.0f / small_int_to_float(nx));
      S.meta[slot] = make_int2(T + incl - cnt, nx | ((recip + ((recip * nx) < 65536 ? 1 : 0)) << 9));
    }
    nslots += __popc(rec);
    T += __shfl_sync(0xffffffffu, incl, 31);
  }
  {  // large-triangle queue: warp-level prefix sum of the packed counter
    unsigned long long mine = 0ull;
#pragma unroll
    for (int h = 0; h < HALVES; h++) mine += my[h];
    if (__any_sync(0xffffffffu, mine != 0ull)) {
      unsigned long long incl = mine;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const unsigned long long o = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= (unsigned)d) incl += o;
      }
      const unsigned long long total = __shfl_sync(0xffffffffu, incl, 31);
      unsigned long long qbase = 0ull;
      if (lane == 31u) {
        qbase = atomicAdd(fr.counters, total);
        // 2^38 tile jobs in one view (hundreds of thousands of screen-filling triangles) would carry into the
        // entry count: flag the view (the host reports kQueueOverflowFlag in the hazard counter as an error);
        // the guards below and in binned_kernel keep every access in bounds meanwhile
        if (((qbase & kQueueJobMask) + (total & kQueueJobMask)) >> kQueueJobBits) atomicAdd(fr.counters + 1, kQueueOverflowFlag);
      }
      qbase = __shfl_sync(0xffffffffu, qbase, 31);
      unsigned long long excl = qbase + incl - mine;
#pragma unroll
      for (int h = 0; h < HALVES; h++) {
        if (my[h] != 0ull) {
          const unsigned qslot = (unsigned)(excl >> kQueueJobBits);
          if (qslot < (unsigned)m.F) {
            fr.q_face[qslot] = face[h];
            fr.q_job[qslot] = excl & kQueueJobMask;
          }
          excl += my[h];
        }
      }
    }
  }
  if (T == 0) return;
  __syncwarp();

  // ---- Phase B ---------------------------------------------------------------------------------
  int st[HALVES];
#pragma unroll
  for (int h = 0; h < HALVES; h++)
    st[h] = ((int)lane + 32 * h < nslots) ? S.meta[lane + 32 * h].x : 0x7fffffff;
  const unsigned le_mask = lt_mask | (1u << lane);
  int started = 0, qhead = 0, qcount = 0;
  for (int sb = 0; sb < T; sb += 32) {
    unsigned bits = 0u;
#pragma unroll
    for (int h = 0; h < HALVES; h++) {
      const unsigned rr = (unsigned)(st[h] - sb);
      if (rr < 32u) bits |= 1u << rr;
    }
    const unsigned starts = __reduce_or_sync(0xffffffffu, bits);
    const int s = sb + (int)lane;
    bool cand = false;
    uint4 e = make_uint4(0u, 0u, 0u, 0u);
    if (s < T) {
      const int slot = started + __popc(starts & le_mask) - 1;
      const int2 mt = S.meta[slot];
      const float4 ra = S.a[slot];
      const float4 rb = S.b[slot];
      const int k = s - mt.x;
      const int nx = mt.y & 511;
      const int dy = (k * (mt.y >> 9)) >> 16;  // k / nx, exact for k, nx <= 256
      const int dx = k - dy * nx;
      const unsigned xy = __float_as_uint(rb.z);
      const float px = small_int_to_float((int)(xy & 0xFFFFu) + dx);
      const float py = small_int_to_float((int)(xy >> 16) + dy);
      // EdgeFunction(v1,v2,p), (v2,v0,p), (v0,v1,p) — same expressions as tri_setup/tri_edges
      const float w0 = (px - ra.z) * (rb.y - ra.w) - (py - ra.w) * (rb.x - ra.z);
      const float w1 = (px - rb.x) * (ra.y - rb.y) - (py - rb.y) * (ra.x - rb.x);
      const float w2 = (px - ra.x) * (ra.w - ra.y) - (py - ra.y) * (ra.z - ra.x);
      const bool pos = (w0 >= 0.0f) && (w1 >= 0.0f) && (w2 >= 0.0f);
      const bool neg = (w0 <= 0.0f) && (w1 <= 0.0f) && (w2 <= 0.0f);
      cand = pos || neg;
      e = make_uint4((unsigned)slot | ((unsigned)dx << 6) | ((unsigned)dy << 14) |
                         (pos ? (1u << 22) : 0u) | (neg ? (1u << 23) : 0u),
                     __float_as_uint(w0), __float_as_uint(w1), __float_as_uint(w2));
    }
    started += __popc(starts);
    const unsigned bal = __ballot_sync(0xffffffffu, cand);
    if (cand) S.q[(qhead + qcount + __popc(bal & lt_mask)) & 63] = e;
    qcount += __popc(bal);
    if (qcount >= 32) {
      __syncwarp();
      raster_fragments(vp, m, fr, S, has_fn, qhead, 32, lane);
      __syncwarp();
      qhead = (qhead + 32) & 63;
      qcount -= 32;
    }
  }
  if (qcount > 0) {
    __syncwarp();
    raster_fragments(vp, m, fr, S, has_fn, qhead, qcount, lane);
  }
}

// PASS 0: CTA i takes block i.  PASS 1: the occlusion ordering above — CTA i takes block order[i] and applies
// the occlusion test when i is past the camera-facing blocks (HALVES == 1).
// TILES: triangles too large for the cooperative path that overlap at most 3 x 3 screen tiles are appended to the
// per-tile lists of tile_kernel; the rest (and whatever finds a list full) goes to the tile-job queue.
template <int HALVES, int MINB, int NV, int PASS, bool TILES = false>
__global__ void __launch_bounds__(256, MINB) raster_kernel(const ViewBatch<NV> vb, const MeshDev m,
                                                           const FrameDev fr0) {
  static_assert(PASS == 0 || HALVES == 1, "block records are per 8 warps x 32 faces");
  const int view = (NV == 1) ? 0 : (int)blockIdx.y;
  const ViewParams& vp = vb.v[view];
  const FrameDev fr = (NV == 1) ? fr0 : frame_view(fr0, view);
  __shared__ WarpTris<HALVES> sh[8];
  WarpTris<HALVES>& S = sh[threadIdx.x >> 5];
  // Fused projection: the CTA projects the <= 256 distinct vertices of its block (thread = vertex) into shared
  // memory and into fr.proj for the later kernels, then its faces pick their corners by 8-bit local index.
  const bool fused = HALVES == 1 && m.lv != nullptr;
  __shared__ float4 s_proj[HALVES == 1 ? 256 : 1];
  const unsigned lane = threadIdx.x & 31u;
  pdl_launch_dependents();
  unsigned blk = blockIdx.x;
  bool deferred = false;
  if (PASS == 1) {
    pdl_wait();  // the block order of this view comes from the vertex kernel
    blk = (unsigned)fr.order[blockIdx.x];
    deferred = (unsigned long long)blockIdx.x >= fr.counters[3];
  }
  if (PASS >= 2) {  // two launches over all blocks: PASS 2 takes the camera-facing ones, PASS 3 the others
    deferred = group_deferred(vp, __ldg(m.grp + blk));
    if (deferred != (PASS == 3)) {
      // not this pass's block.  CTA 0 waits all the same: a grid none of whose threads waited could complete
      // before its predecessor and break the chain for its successor.
      if (blockIdx.x == 0) pdl_wait();
      return;
    }
  }
  const int base = (int)((blk * 8u + (threadIdx.x >> 5)) * (unsigned)(32 * HALVES));
  if (!fused && base >= m.F) return;  // whole warp; the fused path has one block-level barrier below
  const bool has_fn = m.num_face_normals > 0;
  float4 lvx = make_float4(0.f, 0.f, 0.f, __int_as_float(-1));
  if (fused) lvx = __ldg(m.lv + (size_t)blk * 256u + threadIdx.x);  // immutable: before the wait

  // ---- Phase A ---------------------------------------------------------------------------------
  int fidx[HALVES], face[HALVES];
  int i0[HALVES], i1[HALVES], i2[HALVES];
  unsigned li[HALVES];
  float4 pa[HALVES], pb[HALVES], pc[HALVES];
  float fnx[HALVES], fny[HALVES], fnz[HALVES];
#pragma unroll
  for (int h = 0; h < HALVES; h++) {
    fnx[h] = fny[h] = fnz[h] = 0.f;
    i0[h] = i1[h] = i2[h] = 0;
    fidx[h] = base + h * 32 + (int)lane;  // slot in raster order
    face[h] = fidx[h];
    li[h] = 0xFFFFFFFFu;
    if (fidx[h] < m.F) {  // immutable mesh data: may be read before the vertex kernel has finished
      if (fused) li[h] = __ldg(m.lidx + fidx[h]);
      if (li[h] == 0xFFFFFFFFu) {
        i0[h] = __ldg(m.r_idx0 + fidx[h]);
        i1[h] = __ldg(m.r_idx1 + fidx[h]);
        i2[h] = __ldg(m.r_idx2 + fidx[h]);
      }
      if (PASS != 0 && deferred) continue;  // face id and normal only for the triangles that survive
      if (m.r_face) face[h] = __ldg(m.r_face + fidx[h]);
      if (has_fn) {  // same round trip as the indices, not a third dependent one
        fnx[h] = __ldg(m.r_fn_x + fidx[h]);
        fny[h] = __ldg(m.r_fn_y + fidx[h]);
        fnz[h] = __ldg(m.r_fn_z + fidx[h]);
      }
    }
  }
  if (PASS != 1) pdl_wait();  // projected vertices of this view; visibility buffer / counters left clean by resolve
  if (fused) {
    const int vid = __float_as_int(lvx.w);
    if (vid >= 0) {
      const float4 P = project_vertex(vp, lvx.x, lvx.y, lvx.z);
      s_proj[threadIdx.x] = P;
      fr.proj[vid] = P;  // binned / resolve gather by global id; blocks sharing a vertex store the same value
    }
    __syncthreads();
    if (base >= m.F) return;
  }
#pragma unroll
  for (int h = 0; h < HALVES; h++) {
    if (fidx[h] < m.F) {
      if (li[h] != 0xFFFFFFFFu) {
        pa[h] = s_proj[li[h] & 255u];
        pb[h] = s_proj[(li[h] >> 8) & 255u];
        pc[h] = s_proj[(li[h] >> 16) & 255u];
      } else if (fused) {  // block with more than 256 distinct vertices: every face projects its own corners
        const float4 A = __ldg(m.pos4 + i0[h]), B = __ldg(m.pos4 + i1[h]), Cc = __ldg(m.pos4 + i2[h]);
        pa[h] = project_vertex(vp, A.x, A.y, A.z);
        pb[h] = project_vertex(vp, B.x, B.y, B.z);
        pc[h] = project_vertex(vp, Cc.x, Cc.y, Cc.z);
        fr.proj[i0[h]] = pa[h];
        fr.proj[i1[h]] = pb[h];
        fr.proj[i2[h]] = pc[h];
      } else {
        // plain (coherent) loads: proj is rewritten every view by the preceding kernel
        pa[h] = fr.proj[i0[h]];
        pb[h] = fr.proj[i1[h]];
        pc[h] = fr.proj[i2[h]];
      }
    }
  }
  raster_warp_view<HALVES, PASS, TILES>(vp, m, fr, S, lane, has_fn, deferred, fidx, face, fnx, fny, fnz, pa, pb, pc);
}

// ---------------------------------------------------------------------------------------------
// K3: large triangles.  The queue orders triangles by their first tile job, so the global job
// space [0, T) is load-balanced regardless of triangle size: each warp claims chunks of jobs (chunk
// size adapted to T so that every resident warp gets work: small meshes of mid-size triangles would
// otherwise leave most of the machine idle), finds the owning queue entry with a warp-cooperative
// 32-ary search, then walks entries; a job is an 8x4 pixel block of the triangle's bbox, one lane
// per pixel.  Persistent grid (multiple of the SM count); exits at once when the queue is empty.
// ---------------------------------------------------------------------------------------------
// The tile jobs of one view, shared by `nwarps` warps of which this is warp `gw` (binned_kernel's grid, or the
// resolve kernel's when the separate launch is folded away).
// L2ONLY: the early-z read of the visibility buffer bypasses L1 — inside resolve_kernel the same SM reads those
// lines again after atomics of other SMs have changed them in L2, and must not find a stale copy.
template <bool L2ONLY>
__device__ __forceinline__ void tile_jobs(const ViewParams& vp, const MeshDev& m, const FrameDev& fr,
                                          unsigned long long packed, unsigned long long gw,
                                          unsigned long long nwarps) {
  const unsigned long long T = packed & kQueueJobMask;
  const unsigned n = min((unsigned)(packed >> kQueueJobBits), (unsigned)m.F);  // (clamped: see the overflow flag)
  if (T == 0ull || n == 0u) return;
  const unsigned lane = threadIdx.x & 31u;
  // Jobs per chunk: every warp makes `rounds` passes over interleaved chunks of at most
  // kJobsPerChunk jobs, sized so that the chunk count is (just under) a multiple of the warp count
  // — with a fixed chunk size a job count slightly above nwarps * kJobsPerChunk would hand a few
  // warps twice the work of the others.
  const unsigned long long rounds =
      (T + nwarps * (unsigned long long)kJobsPerChunk - 1) / (nwarps * (unsigned long long)kJobsPerChunk);
  unsigned long long per = (T + nwarps * rounds - 1) / (nwarps * rounds);
  per = per < 2ull ? 2ull : per;
  const unsigned long long nchunks = (T + per - 1) / per;
  const bool has_fn = m.num_face_normals > 0;
  const int lx = (int)(lane & (kJobW - 1)), ly = (int)(lane >> 3);

  for (unsigned long long chunk = gw; chunk < nchunks; chunk += nwarps) {
    const unsigned long long jb = chunk * per;
    const unsigned long long je = (jb + per < T) ? jb + per : T;
    // 32-ary search: largest e with q_job[e] <= jb (q_job[0] == 0).
    unsigned lo = 0, hi = n;
    while (hi - lo > 1u) {
      const unsigned span = hi - lo;
      const unsigned step = (span + 31u) / 32u;
      const unsigned idx = lo + lane * step;
      const bool pred = (idx < hi) && (fr.q_job[idx] <= jb);
      const unsigned bal = __ballot_sync(0xffffffffu, pred);
      const unsigned cnt = (unsigned)__popc(bal);  // >= 1: lane 0 probes lo
      const unsigned nlo = lo + (cnt - 1u) * step;
      const unsigned nhi = lo + cnt * step;
      lo = nlo;
      hi = (nhi < hi) ? nhi : hi;
    }
    unsigned e = lo;
    unsigned long long j = jb;
    while (j < je) {
      const int f = fr.q_face[e];
      const unsigned long long js = fr.q_job[e];
      const unsigned long long jn = (e + 1u < n) ? fr.q_job[e + 1u] : T;
      const unsigned long long jend = (jn < je) ? jn : je;
      const int i0 = __ldg(m.idx0 + f), i1 = __ldg(m.idx1 + f), i2 = __ldg(m.idx2 + f);
      float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
      if (has_fn) {  // g = c2w_R^T * fn and the facing threshold (see facing_backface)
        const float fnx = __ldg(m.fn_x + f), fny = __ldg(m.fn_y + f), fnz = __ldg(m.fn_z + f);
        g = make_float4((vp.c2w_R[0] * fnx + vp.c2w_R[3] * fny) + vp.c2w_R[6] * fnz,
                        (vp.c2w_R[1] * fnx + vp.c2w_R[4] * fny) + vp.c2w_R[7] * fnz,
                        (vp.c2w_R[2] * fnx + vp.c2w_R[5] * fny) + vp.c2w_R[8] * fnz,
                        fmaxf(vp.facing_eps * ((fabsf(fnx) + fabsf(fny)) + fabsf(fnz)), 1e-30f));
      }
      Tri t;
      tri_setup(t, fr.proj[i0], fr.proj[i1], fr.proj[i2]);
      int x0, y0, x1, y1;
      tri_bounds(t, vp, x0, y0, x1, y1);  // queued triangles always pass
      const unsigned ntx = (unsigned)((x1 - x0 + kJobW) / kJobW);
      // tile coordinates of the first job, then stepped (one integer division per triangle, not per job)
      unsigned ty = (unsigned)(j - js) / ntx, tx = (unsigned)(j - js) - ty * ntx;
      for (unsigned long long jj = j; jj < jend; ++jj, ++tx) {
        if (tx == ntx) {
          tx = 0u;
          ++ty;
        }
        const int x = x0 + (int)tx * kJobW + lx;
        const int y = y0 + (int)ty * kJobH + ly;
        if (x <= x1 && y <= y1) {
          float w0, w1, w2;
          tri_edges(t, small_int_to_float(x), small_int_to_float(y), w0, w1, w2);
          const bool pos = (w0 >= 0.0f) && (w1 >= 0.0f) && (w2 >= 0.0f);
          const bool neg = (w0 <= 0.0f) && (w1 <= 0.0f) && (w2 <= 0.0f);
          if (pos || neg) {
            const size_t pix = (size_t)y * vp.W + x;
            const unsigned long long seen = L2ONLY ? __ldcg(fr.vis + pix) : fr.vis[pix];  // consumed after the divisions
            const bool bf = has_fn ? facing_backface(vp, m, fr, g, f, x, y) : false;
            if (bf ? neg : pos) {
              const float z = tri_depth(t, w0, w1, w2);
              vis_update(fr.vis, pix, seen, z, f, bf, fr.counters + 1);
            }
          }
        }
      }
      j = jend;
      ++e;
    }
  }
}

// (8 CTAs of 4 warps per SM = at most 64 registers: the batch instantiation would otherwise take 71 and lose a CTA
//  per SM — config 3 batched 38.6 -> 35.9 us per view, profiles/r03g_binned_launch_bounds.jsonl)
template <int NV>
__global__ void __launch_bounds__(128, 8) binned_kernel(const ViewBatch<NV> vb, const MeshDev m,
                                                     const FrameDev fr0) {
  const int view = (NV == 1) ? 0 : (int)blockIdx.y;
  const ViewParams& vp = vb.v[view];
  const FrameDev fr = (NV == 1) ? fr0 : frame_view(fr0, view);
  pdl_launch_dependents();
  pdl_wait();
  tile_jobs<false>(vp, m, fr, fr.counters[0], (unsigned long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5),
            (unsigned long long)gridDim.x * (blockDim.x >> 5));
}

// ---------------------------------------------------------------------------------------------
// K3t: per-tile triangle lists (rasterizer.cc:194-241 for the triangles raster_kernel listed).  One CTA per
// 16 x 16 pixel screen tile, one thread per pixel.  The tile's list is staged through shared memory 64
// triangles at a time — thread k < 64 does the dependent loads (face record -> projected vertices) and the
// setup of one triangle exactly once per (triangle, tile) — then every pixel walks the staged records
// (broadcast LDS; a warp owns two 16-pixel rows and skips records whose bbox misses them) with the depth test
// in a register: the same (z, face, backface) key as everywhere else, minimum kept per pixel, and a single
// atomicMin per covered pixel merges it with what raster_kernel / binned_kernel put into the visibility
// buffer.  Empty tiles exit on their count; every tile leaves its count 0 for the next view.
// ---------------------------------------------------------------------------------------------
struct __align__(16) TileTris {
  float4 a[kTileStage];  // v0x v0y v1x v1y
  float4 b[kTileStage];  // v2x v2y (int)(x0 | x1 << 16) (int)(y0 | y1 << 16)   (clamped bbox)
  float4 c[kTileStage];  // v0z v1z v2z area
  float4 d[kTileStage];  // g = c2w_R^T * face_normal (3), facing threshold
  int face[kTileStage];
};

template <int NV>
__global__ void __launch_bounds__(256) tile_kernel(const ViewBatch<NV> vb, const MeshDev m, const FrameDev fr0) {
  const int view = (NV == 1) ? 0 : (int)blockIdx.y;
  const ViewParams& vp = vb.v[view];
  const FrameDev fr = (NV == 1) ? fr0 : frame_view(fr0, view);
  __shared__ TileTris S;
  pdl_launch_dependents();
  pdl_wait();
  const int tile = (int)blockIdx.x;
  const int total = fr.tile_cnt[tile];
  if (total == 0) return;  // (CTA-uniform)
  const int n = min(total, kTileCap);
  const int ty = tile / fr.tiles_x, tx = tile - ty * fr.tiles_x;
  const int x = tx * kTile + (int)(threadIdx.x & 15u), y = ty * kTile + (int)(threadIdx.x >> 4);
  const int wy0 = ty * kTile + (int)(threadIdx.x >> 5) * 2;  // the warp's two rows
  const bool inside = x < vp.W && y < vp.H;
  const float px = small_int_to_float(x), py = small_int_to_float(y);
  const bool has_fn = m.num_face_normals > 0;
  unsigned long long best = kVisEmpty;
  unsigned long long hazards = 0ull;
  for (int c0 = 0; c0 < n; c0 += kTileStage) {
    const int cn = min(kTileStage, n - c0);
    if ((int)threadIdx.x < cn) {
      const int k = (int)threadIdx.x;
      const int f = fr.tile_bin[(size_t)tile * kTileCap + c0 + k];
      const int4 vr = __ldg(m.frec + 2 * (size_t)f);
      float fnx = 0.f, fny = 0.f, fnz = 0.f;
      if (has_fn) {
        fnx = __ldg(m.fn_x + f);
        fny = __ldg(m.fn_y + f);
        fnz = __ldg(m.fn_z + f);
      }
      Tri t;
      tri_setup(t, fr.proj[vr.x], fr.proj[vr.y], fr.proj[vr.z]);
      int bx0, by0, bx1, by1;
      tri_bounds(t, vp, bx0, by0, bx1, by1);  // listed triangles always pass
      S.a[k] = make_float4(t.v0x, t.v0y, t.v1x, t.v1y);
      S.b[k] = make_float4(t.v2x, t.v2y, __int_as_float(bx0 | (bx1 << 16)), __int_as_float(by0 | (by1 << 16)));
      S.c[k] = make_float4(t.v0z, t.v1z, t.v2z, t.area);
      S.d[k] = make_float4((vp.c2w_R[0] * fnx + vp.c2w_R[3] * fny) + vp.c2w_R[6] * fnz,
                           (vp.c2w_R[1] * fnx + vp.c2w_R[4] * fny) + vp.c2w_R[7] * fnz,
                           (vp.c2w_R[2] * fnx + vp.c2w_R[5] * fny) + vp.c2w_R[8] * fnz,
                           fmaxf(vp.facing_eps * ((fabsf(fnx) + fabsf(fny)) + fabsf(fnz)), 1e-30f));
      S.face[k] = f;
    }
    __syncthreads();
    for (int k = 0; k < cn; k++) {
      const float4 rb = S.b[k];
      const unsigned bx = __float_as_uint(rb.z), by = __float_as_uint(rb.w);
      const int bx0 = (int)(bx & 0xFFFFu), bx1 = (int)(bx >> 16), by0 = (int)(by & 0xFFFFu), by1 = (int)(by >> 16);
      if (by1 < wy0 || by0 > wy0 + 1 || bx1 < tx * kTile || bx0 > tx * kTile + (kTile - 1)) continue;  // warp-uniform
      if (!inside || x < bx0 || x > bx1 || y < by0 || y > by1) continue;
      const float4 ra = S.a[k];
      // EdgeFunction(v1,v2,p), (v2,v0,p), (v0,v1,p) — the expressions of tri_setup / tri_edges
      float w0 = (px - ra.z) * (rb.y - ra.w) - (py - ra.w) * (rb.x - ra.z);
      float w1 = (px - rb.x) * (ra.y - rb.y) - (py - rb.y) * (ra.x - rb.x);
      float w2 = (px - ra.x) * (ra.w - ra.y) - (py - ra.y) * (ra.z - ra.x);
      const bool pos = (w0 >= 0.0f) && (w1 >= 0.0f) && (w2 >= 0.0f);
      const bool neg = (w0 <= 0.0f) && (w1 <= 0.0f) && (w2 <= 0.0f);
      if (!(pos || neg)) continue;
      const float4 rc = S.c[k];
      {  // conservative early-out against the register winner (the bound of raster_fragments)
        const float zs = __uint_as_float((unsigned)(best >> 32));  // NaN while there is no winner
        const float zmin = fminf(fminf(rc.x, rc.y), rc.z);
        const float s3 = (w0 + w1) + w2;
        const float sa = pos ? s3 : -s3, aa = pos ? rc.w : -rc.w;
        if (pos != neg && aa > 0.0f && sa <= 1.001f * aa && zmin < 1e30f && 0.998f * zmin > zs) continue;
      }
      const int face = S.face[k];
      const bool bf = has_fn ? facing_backface(vp, m, fr, S.d[k], face, x, y) : false;
      if (!(bf ? neg : pos)) continue;
      w0 = w0 / rc.w;
      w1 = w1 / rc.w;
      w2 = w2 / rc.w;
      w0 = w0 / rc.x;
      w1 = w1 / rc.y;
      w2 = w2 / rc.z;
      const float z = 1.0f / ((w0 + w1) + w2);
      if (!(z >= 1.17549435e-38f)) {  // see vis_update
        hazards++;
        continue;
      }
      const unsigned long long key = ((unsigned long long)__float_as_uint(z) << 32) |
                                     (unsigned long long)(((unsigned)face << 1) | (bf ? 1u : 0u));
      best = key < best ? key : best;
    }
    __syncthreads();  // the staging area is rewritten by the next chunk
  }
  if (best != kVisEmpty) atomicMin(fr.vis + (size_t)y * vp.W + x, best);
  if (hazards) atomicAdd(fr.counters + 1, hazards);
  if (threadIdx.x == 0) fr.tile_cnt[tile] = 0;
}

// ---------------------------------------------------------------------------------------------
// K4: resolve + attributes + shading + packing (rasterizer.cc:244-301, pixel_shader.h:222-422),
// fused with the clear of every requested output (util_private.cc:72-87) and with the reset of the
// visibility buffer and queue counters for the next view.  One pixel per thread (maximum number of
// independent gather chains in flight); a warp owns 32 consecutive pixels and stages its colour /
// normal / mask values through shared memory so that every global store is a full-width vector
// store.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned f2u8(float f) {
  // static_cast<unsigned char>(float) as the x86-64 reference compiles it: cvttss2si, low byte;
  // out-of-int-range and NaN give 0x80000000 -> 0.
  if (!(f > -2147483904.0f && f < 2147483648.0f)) return 0u;
  return (unsigned)__float2int_rz(f) & 0xFFu;
}

__device__ __forceinline__ float3 fetch_texel(const uchar4* __restrict__ tex, int w, int h, int x,
                                              int y) {
  const long long idx = (long long)y * w + x;  // linear, like Image3b::at without bounds checks
  if (idx < 0 || idx >= (long long)w * h) return make_float3(0.f, 0.f, 0.f);
  const uchar4 t = __ldg(tex + idx);
  return make_float3((float)t.x, (float)t.y, (float)t.z);
}

struct PixelOut {
  unsigned color;  // r | g << 8 | b << 16
  float depth;
  float nx, ny, nz;
  bool mask;
  int face_id;
};

// Does the winner stored in `key` get shaded (mask = 255, rasterizer.cc:254-266)?  False for EMPTY
// pixels, culled back-face winners (rasterizer.cc:248-252) and face 0 (sic).
__device__ __forceinline__ bool key_shaded(unsigned long long key, bool cull) {
  if (key == kVisEmpty) return false;
  const unsigned low = (unsigned)(key & 0xFFFFFFFFull);
  return !((cull && (low & 1u)) || (low >> 1) == 0u);
}

// The resolve kernel is compiled per (COLOR, NMODE) so that each instantiation carries only the
// gathers and registers of its own attribute path:
//   COLOR 0  no interpolated colour (white, or colour not requested)   pixel_shader.h:222-230
//         1  texture (nearest / bilinear chosen at run time)           pixel_shader.h:260-359
//         2  vertex colour                                              pixel_shader.h:234-256
//   NMODE 0  no normal needed (no normal output, no shading)
//         1  face normals                                               rasterizer.cc:268-270
//         2  vertex normals through normal_indices                      rasterizer.cc:271-279
//         3  vertex normals from the merged per-vertex records (normal_indices == vertex_indices)
template <int COLOR, int NMODE>
struct ResolveTraits {
  static constexpr bool kWeights = (NMODE >= 2) || (COLOR != 0);  // needs the winner's barycentrics
  static constexpr bool kNi = NMODE == 2;
  static constexpr bool kUv = COLOR == 1;
};

// Per-lane state that travels down the software pipeline with a pixel chunk.
template <int COLOR, int NMODE>
struct ResolveStage {
  unsigned long long key;
  int4 vr;            // vertex indices + material id          (stage 2, from frec)
  int4 ni;            // normal indices                         (stage 2, NMODE 2)
  int4 ui;            // uv indices                             (stage 2, COLOR 1)
  float4 p0, p1, p2;  // projected vertices of the winner       (stage 3)
  float2 t0, t1, t2;  // uv of the winner's corners             (stage 3, COLOR 1)
};

// Stage 1: the packed key of pixel chunk*32 + lane (EMPTY outside the image).  Loads are only
// ISSUED by the stage functions; nothing consumes the values before the next loop iteration.
__device__ __forceinline__ unsigned long long resolve_load_key(const FrameDev& fr, unsigned chunk,
                                                               unsigned nchunks, unsigned lane,
                                                               unsigned P) {
  unsigned long long key = kVisEmpty;
  if (chunk < nchunks) {
    const unsigned p = chunk * 32u + lane;
    if (p < P) key = fr.vis[p];
  }
  return key;
}

// Stage 2: the winner's face record (and uv indices), indexed by the face id in the key.
template <int COLOR, int NMODE>
__device__ __forceinline__ void resolve_stage2(const MeshDev& m, bool cull, ResolveStage<COLOR, NMODE>& s) {
  using T = ResolveTraits<COLOR, NMODE>;
  if (!T::kWeights) return;
  s.vr = make_int4(0, 0, 0, 0);
  if (T::kNi) s.ni = make_int4(0, 0, 0, 0);
  if (T::kUv) s.ui = make_int4(0, 0, 0, 0);
  if (!key_shaded(s.key, cull)) return;
  const unsigned fid = (unsigned)(s.key & 0xFFFFFFFFull) >> 1;
  const int4* rec = m.frec + 2 * (size_t)fid;
  if (T::kNi) {
    float4 lo, hi;
    ldg256(rec, lo, hi);
    s.vr = make_int4(__float_as_int(lo.x), __float_as_int(lo.y), __float_as_int(lo.z), __float_as_int(lo.w));
    s.ni = make_int4(__float_as_int(hi.x), __float_as_int(hi.y), __float_as_int(hi.z), __float_as_int(hi.w));
  } else {
    s.vr = __ldg(rec);
  }
  if (T::kUv) s.ui = __ldg(m.uvidx + fid);
}

// Stage 3: the winner's projected vertices (per-view data: plain loads) and corner uvs.
template <int COLOR, int NMODE>
__device__ __forceinline__ void resolve_stage3(const MeshDev& m, const FrameDev& fr, bool cull,
                                               ResolveStage<COLOR, NMODE>& s) {
  using T = ResolveTraits<COLOR, NMODE>;
  if (!T::kWeights) return;
  s.p0 = s.p1 = s.p2 = make_float4(0.f, 0.f, 0.f, 0.f);
  if (T::kUv) s.t0 = s.t1 = s.t2 = make_float2(0.f, 0.f);
  if (!key_shaded(s.key, cull)) return;
  s.p0 = fr.proj[s.vr.x];
  s.p1 = fr.proj[s.vr.y];
  s.p2 = fr.proj[s.vr.z];
  if (T::kUv) {
    s.t0 = __ldg(m.uv2 + s.ui.x);
    s.t1 = __ldg(m.uv2 + s.ui.y);
    s.t2 = __ldg(m.uv2 + s.ui.z);
  }
}

// Stage 4: everything else of rasterizer.cc:244-301 for one pixel.
template <int COLOR, int NMODE>
__device__ __forceinline__ void resolve_pixel(const ViewParams& vp, const MeshDev& m,
                                              const FrameDev& fr, const ShadeParams& sp,
                                              const ResolveStage<COLOR, NMODE>& s, int x, int y,
                                              PixelOut& o) {
  using T = ResolveTraits<COLOR, NMODE>;
  o.color = 0u;
  o.depth = 0.0f;
  o.nx = o.ny = o.nz = 0.0f;
  o.mask = false;
  o.face_id = -1;
  const unsigned long long key = s.key;
  if (key == kVisEmpty) return;
  const unsigned low = (unsigned)(key & 0xFFFFFFFFull);
  const int fid = (int)(low >> 1);
  // the winner's backface flag (rasterizer.cc:200,235), carried by the key
  if (sp.backface_culling && (low & 1u)) return;  // depth = 0, face_id = -1 (rasterizer.cc:248-252)
  o.depth = __uint_as_float((unsigned)(key >> 32));
  o.face_id = fid;
  if (fid <= 0) return;  // sic: face 0 is never shaded (rasterizer.cc:254)
  o.mask = true;

  const bool need_color = sp.color != nullptr;
  const bool need_normal = sp.normal != nullptr;
  const bool shade = need_color && sp.diffuse_shading != 0;
  if (NMODE == 0 && !need_color) return;

  // attribute gathers first: they are in flight while the weights are recomputed
  float4 n0 = make_float4(0.f, 0.f, 0.f, 0.f), n1 = n0, n2 = n0;
  float4 c0 = n0, c1 = n0, c2 = n0;
  float nwx = 0.f, nwy = 0.f, nwz = 0.f;
  if (NMODE == 1) {
    nwx = __ldg(m.fn_x + fid);
    nwy = __ldg(m.fn_y + fid);
    nwz = __ldg(m.fn_z + fid);
  } else if (NMODE == 2) {
    n0 = __ldg(m.nrm4 + s.ni.x);
    n1 = __ldg(m.nrm4 + s.ni.y);
    n2 = __ldg(m.nrm4 + s.ni.z);
  } else if (NMODE == 3) {  // merged vertex records deliver the colours together with the normals
    ldg256(m.vattr + 2 * (size_t)s.vr.x, n0, c0);
    ldg256(m.vattr + 2 * (size_t)s.vr.y, n1, c1);
    ldg256(m.vattr + 2 * (size_t)s.vr.z, n2, c2);
  }
  if (COLOR == 2 && NMODE != 3 && need_color) {
    c0 = __ldg(m.col4 + s.vr.x);
    c1 = __ldg(m.col4 + s.vr.y);
    c2 = __ldg(m.col4 + s.vr.z);
  }

  float w0 = 0.f, w1 = 0.f, w2 = 0.f;
  if (T::kWeights) {  // recomputed bit-identically to the raster kernels (no weight image)
    Tri t;
    tri_setup(t, s.p0, s.p1, s.p2);
    tri_edges(t, (float)x, (float)y, w0, w1, w2);
    // rasterizer.cc:208-224 without its seventh division: z = 1 / ((w0 + w1) + w2) of exactly these operands is
    // what the raster kernel stored in the key (o.depth)
    w0 = w0 / t.area;
    w1 = w1 / t.area;
    w2 = w2 / t.area;
    w0 = w0 / t.v0z;
    w1 = w1 / t.v1z;
    w2 = w2 / t.v2z;
    w0 = w0 * o.depth;
    w1 = w1 * o.depth;
    w2 = w2 * o.depth;
  }
  if (NMODE >= 2) {  // rasterizer.cc:271-279
    nwx = (w0 * n0.x + w1 * n1.x) + w2 * n2.x;
    nwy = (w0 * n0.y + w1 * n1.y) + w2 * n2.y;
    nwz = (w0 * n0.z + w1 * n1.z) + w2 * n2.z;
  }
  if (NMODE != 0 && need_normal) {  // rasterizer.cc:281-289
    o.nx = dot3(vp.w2c_R[0], vp.w2c_R[1], vp.w2c_R[2], nwx, nwy, nwz);
    o.ny = dot3(vp.w2c_R[3], vp.w2c_R[4], vp.w2c_R[5], nwx, nwy, nwz);
    o.nz = dot3(vp.w2c_R[6], vp.w2c_R[7], vp.w2c_R[8], nwx, nwy, nwz);
  }
  if (!need_color) return;

  unsigned cr = 255u, cg = 255u, cb = 255u;  // DiffuseDefaultColorizer
  const float u = w1, v = w2;
  if (COLOR == 2) {  // DiffuseVertexColorColorizer (pixel_shader.h:234-256)
    const float k0 = 1.0f - u - v;
    cr = f2u8((k0 * c0.x + u * c1.x) + v * c2.x);
    cg = f2u8((k0 * c0.y + u * c1.y) + v * c2.y);
    cb = f2u8((k0 * c0.z + u * c1.z) + v * c2.z);
  } else if (COLOR == 1) {  // texture colourizers (pixel_shader.h:260-359)
    const int mat = s.vr.w;
    const uchar4* tex = m.tex[mat];
    const int2 ts = m.tex_size[mat];
    const float k0 = 1.0f - u - v;
    const float iu = (k0 * s.t0.x + u * s.t1.x) + v * s.t2.x;
    const float iv = (k0 * s.t0.y + u * s.t1.y) + v * s.t2.y;
    const float tx = iu * (float)(ts.x - 1);
    const float ty = (1.0f - iv) * (float)(ts.y - 1);
    if (sp.interp == 0) {
      const float3 c = fetch_texel(tex, ts.x, ts.y, sat_int(roundf(tx)), sat_int(roundf(ty)));
      cr = f2u8(c.x); cg = f2u8(c.y); cb = f2u8(c.z);
    } else {
      const int xm = sat_int(floorf(tx)), ym = sat_int(floorf(ty));
      const float lu = tx - (float)xm, lv = ty - (float)ym;
      const float3 a = fetch_texel(tex, ts.x, ts.y, xm, ym);
      const float3 b = fetch_texel(tex, ts.x, ts.y, xm + 1, ym);
      const float3 c = fetch_texel(tex, ts.x, ts.y, xm, ym + 1);
      const float3 d = fetch_texel(tex, ts.x, ts.y, xm + 1, ym + 1);
      const float k00 = (1.0f - lu) * (1.0f - lv), k10 = lu * (1.0f - lv);
      const float k01 = (1.0f - lu) * lv, k11 = lu * lv;
      cr = f2u8(((k00 * a.x + k10 * b.x) + k01 * c.x) + k11 * d.x);
      cg = f2u8(((k00 * a.y + k10 * b.y) + k01 * c.y) + k11 * d.y);
      cb = f2u8(((k00 * a.z + k10 * b.z) + k01 * c.z) + k11 * d.z);
    }
  }
  if (NMODE != 0 && shade) {
    float r0, r1, r2;  // light_dir == ray_w (rasterizer.cc:291-298)
    ray_world(vp, ((float)x - vp.cx) / vp.fx, ((float)y - vp.cy) / vp.fy, r0, r1, r2);  // the ray tables' expressions
    float coeff;
    if (sp.diffuse_shading == 1) {  // DiffuseLambertianShader (pixel_shader.h:367-386)
      coeff = -dot3(r0, r1, r2, nwx, nwy, nwz);
      if (coeff < 0.0f) coeff = 0.0f;
    } else {  // DiffuseOrenNayarShader (pixel_shader.h:390-422)
      const float dot_light = -dot3(r0, r1, r2, nwx, nwy, nwz);
      const float theta_i = acosf(dot_light);
      const float theta_r = theta_i;  // dot_ray is the same expression: light_dir == ray_w
      const float blx = (-nwx) * dot_light - r0;
      const float bly = (-nwy) * dot_light - r1;
      const float blz = (-nwz) * dot_light - r2;
      const float phi = fmaxf(0.0f, dot3(blx, bly, blz, blx, bly, blz));
      const float alpha = theta_i, beta = theta_r;
      coeff = fmaxf(0.0f, dot_light) * (sp.on_A + (sp.on_B * phi * sinf(alpha) * tanf(beta)));
    }
    cr = f2u8(coeff * (float)cr);
    cg = f2u8(coeff * (float)cg);
    cb = f2u8(coeff * (float)cb);
  }
  o.color = cr | (cg << 8) | (cb << 16);
}

// Persistent grid: warp w resolves the 32-pixel chunks w, w + nwarps, ... through a four-stage
// software pipeline.  While chunk i gathers its vertex attributes, shades and stores, the projected
// vertices (and uvs) of chunk i+1, the face record of chunk i+2 and the key of chunk i+3 are in
// flight, so of the dependent chain key -> face record -> projected vertices / attributes -> texels
// only the last links are exposed, and those overlap the weight arithmetic of the same chunk.
// FOLD: no binned_kernel ran for this launch set; this kernel owns the tile-job queue (see below).
template <int COLOR, int NMODE, int MINB, bool DEEP, int NV, bool FOLD = false>
__global__ void __launch_bounds__(256, MINB) resolve_kernel(const ViewBatch<NV> vb, const MeshDev m,
                                                            const FrameDev fr0, const ShadeParams sp0) {
  const int view = (NV == 1) ? 0 : (int)blockIdx.y;
  const ViewParams& vp = vb.v[view];
  const FrameDev fr = (NV == 1) ? fr0 : frame_view(fr0, view);
  // (the single-view instantiation is only launched with first_view == 0: everything stays in the constant bank)
  const ShadeParams sp = (NV == 1) ? sp0 : shade_view(sp0, view);
  // full-width vector stores need the view's slab of every output 16-byte aligned: the bases are (validated
  // by the host), so every slab is when W*H is a multiple of 16; other slabs take the scalar stores
  const bool vec_ok =
      (NV == 1) || ((((unsigned long long)(sp0.first_view + (unsigned)view) * sp0.view_pixels) & 15ull) == 0ull);
  __shared__ __align__(16) float s_n[8][96];
  __shared__ __align__(16) unsigned char s_c[8][96];
  const unsigned P = (unsigned)vp.W * (unsigned)vp.H;  // W*H < 2^31 is validated by the host
  const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
  const unsigned nchunks = (NV == 1 && sp.chunk_end != 0u) ? sp.chunk_end : ((P + 31u) >> 5);
  const unsigned nwarps = gridDim.x * 8u;
  pdl_launch_dependents();
  pdl_wait();
  const bool cull = sp.backface_culling != 0;
  using Stage = ResolveStage<COLOR, NMODE>;
  unsigned chunk = ((NV == 1) ? sp.chunk_begin : 0u) + blockIdx.x * 8u + warp;
  Stage a{}, b{}, c{};  // chunk, chunk + nwarps, chunk + 2 nwarps
  // FOLD: the queue counter travels with the first keys (one round trip, not two); the keys are speculative
  const unsigned long long packed = FOLD ? fr.counters[0] : 0ull;
  a.key = resolve_load_key(fr, chunk, nchunks, lane, P);
  b.key = resolve_load_key(fr, chunk + nwarps, nchunks, lane, P);
  c.key = resolve_load_key(fr, chunk + 2u * nwarps, nchunks, lane, P);
  if (FOLD && (packed & kQueueJobMask) != 0ull) {
    // No binned_kernel ran (meshes of small triangles: the launch would read one counter and exit, ~3 us of every
    // view), but this view queued triangles after all: this grid rasterises them first — it is persistent and
    // co-resident (<= 4 CTAs per SM, and its PDL successor cannot start before all of its CTAs have), so a grid
    // barrier is safe — and only then resolves.  Every thread reads the same counter: the branch is grid-uniform;
    // the counter is reset after the barrier.
    tile_jobs<true>(vp, m, fr, packed, (unsigned long long)blockIdx.x * 8u + warp, (unsigned long long)nwarps);
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence();
      atomicAdd(fr.counters + 8, 1ull);
      while (*(volatile unsigned long long*)(fr.counters + 8) < (unsigned long long)gridDim.x) {
      }
      __threadfence();
      // the last CTA to leave re-arms the barrier for the next view
      if (atomicAdd(fr.counters + 9, 1ull) == (unsigned long long)gridDim.x - 1ull) {
        fr.counters[8] = 0ull;
        fr.counters[9] = 0ull;
      }
    }
    __syncthreads();
    // the speculative keys are stale, and so are their lines in this SM's L1: reload from L2
    a.key = b.key = c.key = kVisEmpty;
    const unsigned pa = chunk * 32u + lane, pb = (chunk + nwarps) * 32u + lane, pc = (chunk + 2u * nwarps) * 32u + lane;
    if (chunk < nchunks && pa < P) a.key = __ldcg(fr.vis + pa);
    if (chunk + nwarps < nchunks && pb < P) b.key = __ldcg(fr.vis + pb);
    if (chunk + 2u * nwarps < nchunks && pc < P) c.key = __ldcg(fr.vis + pc);
  }
  if (blockIdx.x == 0 && threadIdx.x == 0 && (NV != 1 || sp.chunk_begin == 0u)) {
    // the queue / hazard counters are consumed: snapshot them for the host (counters[4..5]) and
    // reset them for the next view (by the first band's launch when the view is resolved in bands)
    fr.counters[4] = fr.counters[0];
    fr.counters[5] = fr.counters[1];
    if (fr.counters[1] != 0ull) atomicAdd(fr0.counters + kTotalsOffset, fr.counters[1]);
    if (NV == 1 && fr0.host_counters != nullptr) {
      // the same three words straight into the host's (mapped, pinned) snapshot: the host form of a single view
      // needs no 24-byte DMA between the kernels and the image copies
      volatile unsigned long long* h = fr0.host_counters;
      h[0] = fr.counters[0];
      h[1] = fr.counters[1];
      h[2] = *(volatile unsigned long long*)(fr0.counters + kTotalsOffset);
    }
    fr.counters[0] = 0ull;
    fr.counters[1] = 0ull;
    fr.counters[2] = 0ull;
    fr.counters[3] = 0ull;
    fr.counters[7] = 0ull;
  }
  resolve_stage2(m, cull, a);
  resolve_stage2(m, cull, b);
  if (DEEP) resolve_stage3(m, fr, cull, a);
  for (; chunk < nchunks; chunk += nwarps) {
    const unsigned long long key_d = resolve_load_key(fr, chunk + 3u * nwarps, nchunks, lane, P);
    resolve_stage2(m, cull, c);
    if (DEEP) resolve_stage3(m, fr, cull, b);
    else resolve_stage3(m, fr, cull, a);

    const unsigned wbase = chunk * 32u;  // first pixel of this chunk
    const unsigned p = wbase + lane;
    const unsigned y = (vp.W == 1) ? p : (__umulhi(p, vp.w_magic) >> vp.w_shift);
    const unsigned x = p - y * (unsigned)vp.W;
    if (vec_ok && wbase + 32u <= P && __all_sync(0xffffffffu, a.key == kVisEmpty)) {
      // Nothing landed on these 32 pixels (most of a typical frame): write the cleared values
      // (util_private.cc:72-87) with full-width stores and move on.
      if (sp.depth) sp.depth[p] = 0.0f;
      if (sp.face_id) sp.face_id[p] = -1;
      if (sp.mask && lane < 8u) *reinterpret_cast<unsigned*>(sp.mask + wbase + lane * 4u) = 0u;
      if (sp.normal && lane < 24u)
        reinterpret_cast<float4*>(sp.normal + (size_t)wbase * 3u)[lane] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (sp.color && lane < 6u)
        reinterpret_cast<uint4*>(sp.color + (size_t)wbase * 3u)[lane] = make_uint4(0u, 0u, 0u, 0u);
      a = b;
      b = c;
      c.key = key_d;
      continue;
    }
    if (a.key != kVisEmpty) fr.vis[p] = kVisEmpty;  // leave the buffer EMPTY for the next view
    PixelOut o;
    resolve_pixel<COLOR, NMODE>(vp, m, fr, sp, a, (int)x, (int)y, o);

    if (vec_ok && wbase + 32u <= P) {
      if (sp.depth) sp.depth[p] = o.depth;
      if (sp.face_id) sp.face_id[p] = o.face_id;
      if (sp.mask) {
        const unsigned bal = __ballot_sync(0xffffffffu, o.mask);
        if (lane < 8u) {
          // nibble -> 4 bytes of 0x00 / 0xFF
          const unsigned nb = (bal >> (lane * 4u)) & 15u;
          const unsigned word = ((nb * 0x00204081u) & 0x01010101u) * 0xFFu;
          *reinterpret_cast<unsigned*>(sp.mask + wbase + lane * 4u) = word;
        }
      }
      if (sp.normal) {
        s_n[warp][lane * 3u + 0u] = o.nx;
        s_n[warp][lane * 3u + 1u] = o.ny;
        s_n[warp][lane * 3u + 2u] = o.nz;
      }
      if (sp.color) {
        s_c[warp][lane * 3u + 0u] = (unsigned char)(o.color & 0xFFu);
        s_c[warp][lane * 3u + 1u] = (unsigned char)((o.color >> 8) & 0xFFu);
        s_c[warp][lane * 3u + 2u] = (unsigned char)((o.color >> 16) & 0xFFu);
      }
      __syncwarp();
      if (sp.normal && lane < 24u)
        reinterpret_cast<float4*>(sp.normal + (size_t)wbase * 3u)[lane] =
            reinterpret_cast<const float4*>(s_n[warp])[lane];
      if (sp.color && lane < 6u)
        reinterpret_cast<uint4*>(sp.color + (size_t)wbase * 3u)[lane] =
            reinterpret_cast<const uint4*>(s_c[warp])[lane];
      __syncwarp();  // the staging rows are rewritten by the next chunk
    } else if (p < P) {  // ragged last chunk of the image
      if (sp.color) {
        sp.color[(size_t)p * 3 + 0] = (uint8_t)(o.color & 0xFFu);
        sp.color[(size_t)p * 3 + 1] = (uint8_t)((o.color >> 8) & 0xFFu);
        sp.color[(size_t)p * 3 + 2] = (uint8_t)((o.color >> 16) & 0xFFu);
      }
      if (sp.depth) sp.depth[p] = o.depth;
      if (sp.normal) {
        sp.normal[(size_t)p * 3 + 0] = o.nx;
        sp.normal[(size_t)p * 3 + 1] = o.ny;
        sp.normal[(size_t)p * 3 + 2] = o.nz;
      }
      if (sp.mask) sp.mask[p] = o.mask ? 255 : 0;
      if (sp.face_id) sp.face_id[p] = o.face_id;
    }
    a = b;
    b = c;
    c.key = key_d;
  }
}

// ---------------------------------------------------------------------------------------------
// K5: depth image -> organized point cloud, ugu::Depth2PointCloud(const Image1f&, const Camera&,
// Image3f*, bool gl_coord) (src/util/rgbd_util.cc:472-513) with PinholeCamera::Unproject
// (include/ugu/camera.h:406-413): camera_p = ((x - cx) * d / fx, (y - cy) * d / fy, d), y and z
// negated for gl_coord.  Pixels with d < FLT_MIN are skipped by the reference (the output keeps what
// it held); `zero_invalid` writes zeros there instead, which is what a freshly Init()-ed image
// holds.  Streaming: 4 pixels per thread, one 16-byte load and three 16-byte stores.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float3 unproject_pixel(const ViewParams& vp, unsigned p, float d, bool gl) {
  const unsigned y = (vp.W == 1) ? p : (__umulhi(p, vp.w_magic) >> vp.w_shift);
  const unsigned x = p - y * (unsigned)vp.W;
  float3 c;
  c.x = ((float)x - vp.cx) * d / vp.fx;
  c.y = ((float)y - vp.cy) * d / vp.fy;
  c.z = d;
  if (gl) {
    c.y = -c.y;
    c.z = -c.z;
  }
  return c;
}

__global__ void __launch_bounds__(256) depth_to_point_cloud_kernel(const ViewParams vp,
                                                                   const float* __restrict__ depth,
                                                                   float* __restrict__ pc, int gl_coord,
                                                                   int zero_invalid) {
  const unsigned P = (unsigned)vp.W * (unsigned)vp.H;
  const unsigned p0 = (blockIdx.x * blockDim.x + threadIdx.x) * 4u;
  if (p0 >= P) return;
  const bool gl = gl_coord != 0;
  if (p0 + 4u <= P) {
    const float4 d4 = *reinterpret_cast<const float4*>(depth + p0);
    const float d[4] = {d4.x, d4.y, d4.z, d4.w};
    bool valid[4];
    bool all = true;
#pragma unroll
    for (int k = 0; k < 4; k++) {
      valid[k] = !(d[k] < 1.17549435e-38f);
      all = all && valid[k];
    }
    if (all || zero_invalid) {
      float o[12];
#pragma unroll
      for (int k = 0; k < 4; k++) {
        float3 c = make_float3(0.f, 0.f, 0.f);
        if (valid[k]) c = unproject_pixel(vp, p0 + k, d[k], gl);
        o[3 * k + 0] = c.x;
        o[3 * k + 1] = c.y;
        o[3 * k + 2] = c.z;
      }
      float4* dst = reinterpret_cast<float4*>(pc + (size_t)p0 * 3);
      dst[0] = make_float4(o[0], o[1], o[2], o[3]);
      dst[1] = make_float4(o[4], o[5], o[6], o[7]);
      dst[2] = make_float4(o[8], o[9], o[10], o[11]);
      return;
    }
#pragma unroll
    for (int k = 0; k < 4; k++) {
      if (!valid[k]) continue;
      const float3 c = unproject_pixel(vp, p0 + k, d[k], gl);
      float* q = pc + (size_t)(p0 + k) * 3;
      q[0] = c.x;
      q[1] = c.y;
      q[2] = c.z;
    }
    return;
  }
  for (unsigned p = p0; p < P; p++) {  // ragged tail
    const float d = depth[p];
    float* q = pc + (size_t)p * 3;
    if (d < 1.17549435e-38f) {
      if (zero_invalid) q[0] = q[1] = q[2] = 0.f;
      continue;
    }
    const float3 c = unproject_pixel(vp, p, d, gl);
    q[0] = c.x;
    q[1] = c.y;
    q[2] = c.z;
  }
}

// ---------------------------------------------------------------------------------------------
// Upload-time repack kernels (AoS host layout -> device SoA / padded vectors).  Run once per mesh.
// ---------------------------------------------------------------------------------------------
__global__ void repack_vec3_to_planes(const float* __restrict__ aos, int n, float* __restrict__ px,
                                      float* __restrict__ py, float* __restrict__ pz) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    px[i] = aos[(size_t)i * 3 + 0];
    py[i] = aos[(size_t)i * 3 + 1];
    pz[i] = aos[(size_t)i * 3 + 2];
  }
}

__global__ void repack_ivec3_to_planes(const int* __restrict__ aos, int n, int* __restrict__ p0,
                                       int* __restrict__ p1, int* __restrict__ p2) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    p0[i] = aos[(size_t)i * 3 + 0];
    p1[i] = aos[(size_t)i * 3 + 1];
    p2[i] = aos[(size_t)i * 3 + 2];
  }
}

__global__ void repack_vec3_to_float4(const float* __restrict__ aos, int n, float4* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n)
    out[i] = make_float4(aos[(size_t)i * 3 + 0], aos[(size_t)i * 3 + 1], aos[(size_t)i * 3 + 2], 0.f);
}

__global__ void repack_ivec3_to_int4(const int* __restrict__ aos, int n, int4* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n)
    out[i] = make_int4(aos[(size_t)i * 3 + 0], aos[(size_t)i * 3 + 1], aos[(size_t)i * 3 + 2], 0);
}

// {vidx0, vidx1, vidx2, material_id}, {nidx0, nidx1, nidx2, 0}: one 32-byte sector per face
__global__ void build_face_records(const int* __restrict__ vidx, const int* __restrict__ nidx,
                                   const int* __restrict__ mat, int n, int4* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    out[2 * (size_t)i] = make_int4(vidx[(size_t)i * 3 + 0], vidx[(size_t)i * 3 + 1],
                                   vidx[(size_t)i * 3 + 2], mat ? mat[i] : 0);
    out[2 * (size_t)i + 1] = nidx ? make_int4(nidx[(size_t)i * 3 + 0], nidx[(size_t)i * 3 + 1],
                                              nidx[(size_t)i * 3 + 2], 0)
                                  : make_int4(0, 0, 0, 0);
  }
}

// {nx ny nz 0, r g b 0} per vertex (normals indexed like vertices); colours may be absent
__global__ void build_vertex_attr(const float* __restrict__ nrm, const float* __restrict__ col, int n,
                                  float4* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    out[2 * (size_t)i] = make_float4(nrm[(size_t)i * 3 + 0], nrm[(size_t)i * 3 + 1], nrm[(size_t)i * 3 + 2], 0.f);
    out[2 * (size_t)i + 1] = col ? make_float4(col[(size_t)i * 3 + 0], col[(size_t)i * 3 + 1],
                                               col[(size_t)i * 3 + 2], 0.f)
                                 : make_float4(0.f, 0.f, 0.f, 0.f);
  }
}

__global__ void repack_rgb_to_rgba(const uint8_t* __restrict__ rgb, size_t n, uchar4* __restrict__ out) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = make_uchar4(rgb[i * 3 + 0], rgb[i * 3 + 1], rgb[i * 3 + 2], 255);
}

// dst plane k [slot] = src plane k [slot_face[slot]]
template <typename T>
__global__ void gather_planes(const int* __restrict__ slot_face, int n, const T* __restrict__ a,
                              const T* __restrict__ b, const T* __restrict__ c, T* __restrict__ oa,
                              T* __restrict__ ob, T* __restrict__ oc) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    const int f = slot_face[i];
    oa[i] = a[f];
    ob[i] = b[f];
    oc[i] = c[f];
  }
}

// Block records of the occlusion ordering: a CTA per 256 consecutive slots (= one raster CTA); {sum of the
// geometric (area-weighted) face normals, its dot product with the mean face centroid}.
__global__ void __launch_bounds__(256) block_records_kernel(const float4* __restrict__ pos4,
                                                            const int* __restrict__ i0, const int* __restrict__ i1,
                                                            const int* __restrict__ i2, int F,
                                                            float4* __restrict__ grp) {
  __shared__ float red[8][7];
  const int f = (int)(blockIdx.x * 256u + threadIdx.x);
  const int lane = (int)(threadIdx.x & 31u), warp = (int)(threadIdx.x >> 5);
  float v[7] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};  // normal, centroid sum, count
  if (f < F) {
    const float4 a = pos4[i0[f]], b = pos4[i1[f]], c = pos4[i2[f]];
    const float ux = b.x - a.x, uy = b.y - a.y, uz = b.z - a.z;
    const float wx = c.x - a.x, wy = c.y - a.y, wz = c.z - a.z;
    v[0] = uy * wz - uz * wy;
    v[1] = uz * wx - ux * wz;
    v[2] = ux * wy - uy * wx;
    v[3] = (a.x + b.x + c.x) * (1.0f / 3.0f);
    v[4] = (a.y + b.y + c.y) * (1.0f / 3.0f);
    v[5] = (a.z + b.z + c.z) * (1.0f / 3.0f);
    v[6] = 1.f;
  }
#pragma unroll
  for (int k = 0; k < 7; k++) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v[k] += __shfl_xor_sync(0xffffffffu, v[k], d);
    if (lane == 0) red[warp][k] = v[k];
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    float t[7];
    for (int k = 0; k < 7; k++) {
      t[k] = 0.f;
      for (int w = 0; w < 8; w++) t[k] += red[w][k];
    }
    const float inv = 1.0f / t[6];
    grp[blockIdx.x] = make_float4(t[0], t[1], t[2], (t[0] * (t[3] * inv) + t[1] * (t[4] * inv)) + t[2] * (t[5] * inv));
  }
}

__global__ void fill_u64(unsigned long long* __restrict__ p, size_t n, unsigned long long v) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}

}  // namespace ugu_b200
