// sm_100a kernels for ugu::Depth2Mesh: Depth2MeshImpl, reference
// src/util/rgbd_util.cc:94-282, with PinholeCamera::Unproject (include/ugu/camera.h:406-413), the uv /
// nearest-colour rules of rgbd_util.cc:171-196 and Mesh::CalcNormal / CalcFaceNormal
// (src/mesh.cc:297-349) — the step examples/ex02_renderer.cc:156-158 runs on the rendered depth.
//
// The reference is one sequential raster scan over the grid points (x = x_step, 2 x_step, .. < W;
// y likewise): a valid depth appends a vertex, and up to two triangles to its upper-left / upper /
// left neighbours, so vertex and face numbering are a raster-order compaction.  Here that is a
// prefix sum: every grid point contributes (1 vertex, 0..2 faces) packed in one 64-bit value;
// block sums -> single-block scan -> per-block scan give each point its vertex id and first face id.
// The `index > 0` quirk of rgbd_util.cc:222-234 (vertex 0 never joins a face) only needs to know
// WHICH grid point becomes vertex 0 — the first valid one, found with an atomicMin beforehand.
// Vertex normals: Mesh::CalcNormal adds face normals to their three corners in face order; a grid
// vertex touches at most six faces (its own two, the right neighbour's second, the lower
// neighbour's first, the lower-right neighbour's two) and raster order is exactly that order, so
// each vertex sums its (up to) six face normals itself — no atomics, bit-identical sums.
//
// Float discipline as in raster_kernels.cuh (--fmad=false, reference association order).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace ugu_b200 {

constexpr int kD2MItems = 4;                       // grid points per thread
constexpr int kD2MBlock = 256 * kD2MItems;         // grid points per CTA

struct D2MParams {
  int W, H, xs, ys, GX, GY;  // image size, steps, grid points per row / column
  float fx, fy, cx, cy, th;
  int gl, with_uv, with_color, cw, ch;
  const float* depth;
  const uint8_t* color;      // cw * ch * 3 or null
  int* first;                // [0] = INT_MAX - (first valid grid point), 0 if none: see d2m_first()
  unsigned long long* block_sums;  // per CTA packed (faces << 32 | vertices), scanned in place
  unsigned long long* totals;      // [0] packed totals
  int* vid;                  // per grid point: vertex id or -1
  unsigned* fref;            // per grid point: first face id | has_t1 << 30 | has_t2 << 31
  float* vertices;           // Nv * 3
  int* faces;                // Nf * 3
  float* normals;            // Nv * 3
  float* face_normals;       // Nf * 3
  float* uv;                 // Nv * 2
  float* colors;             // Nv * 3
  float* point_cloud;        // W*H*3 (pre-zeroed unless zero_fill) or null
  float* normal_image;       // W*H*3 (pre-zeroed unless zero_fill) or null
  // single-pass path (d2m_scan_emit_kernel)
  unsigned long long* state; // per chunk: flag << 62 | faces << 31 | vertices (decoupled look-back), zeroed per call
  int* ticket;               // chunk ticket counter, zeroed per call
  int zero_fill;             // steps (1, 1): the kernels write the zeros of the two images themselves (no memsets)
  // two-camera overload (rgbd_util.cc:283-466): uv / colour from the projection of depth2color * p into the colour
  // camera.  d2c = rotation (row-major 9) + translation (3) of the Eigen::Affine3f
  unsigned gx_magic, gx_shift;  // g / GX == __umulhi(g, gx_magic) >> gx_shift for g < 2^31 (GX > 1), as ViewParams::w_magic
  int two_cam;
  float d2c[12];
  float cfx, cfy, ccx, ccy;
};

__device__ __forceinline__ void d2m_pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void d2m_pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// The first valid grid point (INT_MAX if none).  Stored inverted, so that "none" is 0 and one memset of zeros
// initialises every small word of a call.
__device__ __forceinline__ int d2m_first(const D2MParams& p) { return 0x7fffffff - *(volatile int*)p.first; }

// row of grid point g (g < 2^31) without an integer division
__device__ __forceinline__ int d2m_row(const D2MParams& p, int g) {
  return (p.GX == 1) ? g : (int)(__umulhi((unsigned)g, p.gx_magic) >> p.gx_shift);
}

__device__ __forceinline__ float d2m_depth(const D2MParams& p, int gx, int gy) {
  return __ldg(p.depth + (size_t)((gy + 1) * p.ys) * p.W + (size_t)((gx + 1) * p.xs));
}
// the pixel (gx, gy) - one step; gx / gy may be -1 (pixel column / row 0, never a vertex)
__device__ __forceinline__ bool d2m_valid(float d) { return !(d < 1.17549435e-38f); }  // rgbd_util.cc:158

__device__ __forceinline__ void d2m_unproject(const D2MParams& p, int gx, int gy, float d, float (&o)[3]) {
  const float x = (float)((gx + 1) * p.xs), y = (float)((gy + 1) * p.ys);
  o[0] = (x - p.cx) * d / p.fx;  // camera.h:406-413
  o[1] = (y - p.cy) * d / p.fy;
  o[2] = d;
  if (p.gl) {
    o[1] = -o[1];
    o[2] = -o[2];
  }
}

// Which of the two triangles of grid point (gx, gy) exist (rgbd_util.cc:222-234).  A neighbour's
// "index > 0" == it is a vertex and not vertex 0 (= grid point `first`).  Neighbours in pixel row /
// column 0 are never vertices, but their depth still enters the difference test.
__device__ __forceinline__ unsigned d2m_tris(const D2MParams& p, int gx, int gy, float d, int first) {
  const size_t px = (size_t)((gx + 1) * p.xs), py = (size_t)((gy + 1) * p.ys);
  const float d_ul = __ldg(p.depth + (py - p.ys) * p.W + (px - p.xs));
  const float d_up = __ldg(p.depth + (py - p.ys) * p.W + px);
  const float d_lf = __ldg(p.depth + py * p.W + (px - p.xs));
  const int g = gy * p.GX + gx;
  const bool ul = gx > 0 && gy > 0 && d2m_valid(d_ul) && (g - p.GX - 1) != first;
  const bool up = gy > 0 && d2m_valid(d_up) && (g - p.GX) != first;
  const bool lf = gx > 0 && d2m_valid(d_lf) && (g - 1) != first;
  const bool c_ul = fabsf(d_ul - d) < p.th, c_up = fabsf(d_up - d) < p.th, c_lf = fabsf(d_lf - d) < p.th;
  return ((ul && up && c_ul && c_up) ? 1u : 0u) | ((ul && lf && c_ul && c_lf) ? 2u : 0u);
}

// packed contribution of grid point g: vertices in the low word, faces in the high word
__device__ __forceinline__ unsigned long long d2m_contrib(const D2MParams& p, int g, int G, int first,
                                                          unsigned& tris) {
  tris = 0u;
  if (g >= G) return 0ull;
  const int gy = d2m_row(p, g), gx = g - gy * p.GX;
  const float d = d2m_depth(p, gx, gy);
  if (!d2m_valid(d)) return 0ull;
  tris = d2m_tris(p, gx, gy, d, first);
  return 1ull | ((unsigned long long)__popc(tris) << 32);
}

// exclusive scan of one value per thread over a 256-thread CTA; returns the CTA total in `total`
__device__ __forceinline__ unsigned long long d2m_block_scan(unsigned long long v, unsigned long long& total) {
  __shared__ unsigned long long s_w[8];
  const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
  unsigned long long incl = v;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const unsigned long long o = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= (unsigned)d) incl += o;
  }
  __syncthreads();  // s_w may still be read by a previous call
  if (lane == 31u) s_w[warp] = incl;
  __syncthreads();
  unsigned long long base = 0ull, tot = 0ull;
#pragma unroll
  for (int w = 0; w < 8; w++) {
    if ((unsigned)w < warp) base += s_w[w];
    tot += s_w[w];
  }
  total = tot;
  return base + incl - v;
}

// D2M-K1: index of the first valid grid point (it becomes vertex 0).  A resident grid walks the
// kD2MBlock-point chunks in ascending order; a warp stops as soon as a grid point below its next chunk is
// known to be valid, so the image is read only up to (one round past) the first valid depth.
__global__ void __launch_bounds__(256) d2m_first_kernel(const D2MParams p) {
  __shared__ int s_known;
  __shared__ int s_min[8];
  const int G = p.GX * p.GY;
  const int nchunks = (G + kD2MBlock - 1) / kD2MBlock;
  for (int chunk = blockIdx.x; chunk < nchunks; chunk += gridDim.x) {
    const int base = chunk * kD2MBlock;
    if (threadIdx.x == 0) s_known = d2m_first(p);
    __syncthreads();
    if (base > s_known) break;  // CTA-uniform
    int mine = 0x7fffffff;
#pragma unroll
    for (int k = kD2MItems - 1; k >= 0; k--) {
      const int g = base + k * 256 + (int)threadIdx.x;
      if (g < G) {
        const int gy = d2m_row(p, g), gx = g - gy * p.GX;
        if (d2m_valid(d2m_depth(p, gx, gy))) mine = g;  // descending k: the lowest valid g survives
      }
    }
    mine = __reduce_min_sync(0xffffffffu, mine);
    if ((threadIdx.x & 31u) == 0u) s_min[threadIdx.x >> 5] = mine;
    __syncthreads();
    if (threadIdx.x == 0) {
      int m = s_min[0];
#pragma unroll
      for (int w = 1; w < 8; w++) m = min(m, s_min[w]);
      // one same-address atomic per CTA at most, and only if it can still lower the (re-read) value
      if (m < d2m_first(p)) atomicMax(p.first, 0x7fffffff - m);
    }
  }
}

// D2M-K2: per-CTA sums of the packed contributions
__global__ void __launch_bounds__(256) d2m_count_kernel(const D2MParams p) {
  const int G = p.GX * p.GY;
  const int first = d2m_first(p);
  const int g0 = blockIdx.x * kD2MBlock + threadIdx.x * kD2MItems;
  unsigned long long sum = 0ull;
#pragma unroll
  for (int k = 0; k < kD2MItems; k++) {
    unsigned tris;
    sum += d2m_contrib(p, g0 + k, G, first, tris);
  }
  unsigned long long total;
  (void)d2m_block_scan(sum, total);
  if (threadIdx.x == 0) p.block_sums[blockIdx.x] = total;
}

// D2M-K3: exclusive scan of the CTA sums (one CTA, running carry) + the totals
__global__ void __launch_bounds__(256) d2m_scan_kernel(const D2MParams p, int nblocks) {
  unsigned long long carry = 0ull;
  for (int base = 0; base < nblocks; base += 256) {
    const int i = base + (int)threadIdx.x;
    const unsigned long long v = i < nblocks ? p.block_sums[i] : 0ull;
    unsigned long long total;
    const unsigned long long excl = d2m_block_scan(v, total);
    if (i < nblocks) p.block_sums[i] = carry + excl;
    carry += total;
  }
  if (threadIdx.x == 0) p.totals[0] = carry;
}

// Everything that belongs to grid point g = (gx, gy) alone, in two parts.
// d2m_emit_image: what needs no prefix — the point-cloud texel (and, with zero_fill, the zeros of the image),
// and the "no vertex" marks of an invalid point.
__device__ __forceinline__ void d2m_zero3(float* q) { q[0] = 0.f; q[1] = 0.f; q[2] = 0.f; }
__device__ __forceinline__ void d2m_emit_image(const D2MParams& p, int g, int gx, int gy, bool valid, float d) {
  const int x = (gx + 1) * p.xs, y = (gy + 1) * p.ys;
  if (p.zero_fill && p.point_cloud) {  // steps (1, 1): pixel row / column 0 belong to the grid points beside them
    if (gx == 0) d2m_zero3(p.point_cloud + 3 * ((size_t)y * p.W));
    if (gy == 0) d2m_zero3(p.point_cloud + 3 * (size_t)x);
    if (g == 0) d2m_zero3(p.point_cloud);
  }
  if (!valid) {
    p.vid[g] = -1;
    p.fref[g] = 0u;
    if (p.zero_fill && p.point_cloud) d2m_zero3(p.point_cloud + 3 * ((size_t)y * p.W + x));
    return;
  }
  if (p.point_cloud) {  // rgbd_util.cc:257-266
    float pos[3];
    d2m_unproject(p, gx, gy, d, pos);
    float* q = p.point_cloud + 3 * ((size_t)y * p.W + x);
    q[0] = pos[0]; q[1] = pos[1]; q[2] = pos[2];
  }
}
// d2m_emit_vertex: vertex id / first face id images, position, uv, nearest colour of a VALID point.
// `run` = packed (faces << 32 | vertices) exclusive prefix of g.
__device__ __forceinline__ void d2m_emit_vertex(const D2MParams& p, int g, int gx, int gy, float d,
                                                unsigned long long run, unsigned tris) {
  const int x = (gx + 1) * p.xs, y = (gy + 1) * p.ys;
  const int v = (int)(run & 0xFFFFFFFFull);
  const unsigned f0 = (unsigned)(run >> 32);
  p.vid[g] = v;
  p.fref[g] = f0 | (tris << 30);
  float pos[3];
  d2m_unproject(p, gx, gy, d, pos);
  p.vertices[3 * (size_t)v + 0] = pos[0];
  p.vertices[3 * (size_t)v + 1] = pos[1];
  p.vertices[3 * (size_t)v + 2] = pos[2];
  if (p.with_uv || p.with_color) {
    float u = ((float)x + 0.5f) / (float)p.W;   // X2U, raster_util.h:22-24
    float w = ((float)y + 0.5f) / (float)p.H;   // Y2V, flip = false
    if (p.two_cam) {
      // rgbd_util.cc:345-350: Affine3f * Vector3f evaluates ((R_i0 x + R_i1 y) + R_i2 z) + t_i (pinned against the
      // compiled reference, tools/probe_d2m2.py), then PinholeCamera::Project (camera.h:380-387), X2U / Y2V
      const float qx = ((p.d2c[0] * pos[0] + p.d2c[1] * pos[1]) + p.d2c[2] * pos[2]) + p.d2c[9];
      const float qy = ((p.d2c[3] * pos[0] + p.d2c[4] * pos[1]) + p.d2c[5] * pos[2]) + p.d2c[10];
      const float qz = ((p.d2c[6] * pos[0] + p.d2c[7] * pos[1]) + p.d2c[8] * pos[2]) + p.d2c[11];
      const float ix = p.cfx / qz * qx + p.ccx;
      const float iy = p.cfy / qz * qy + p.ccy;
      u = (ix + 0.5f) / (float)p.cw;
      w = (iy + 0.5f) / (float)p.ch;
    }
    if (p.with_color) {  // nearest colour texel, rgbd_util.cc:178-190
      const int tx = __float2int_rz(roundf(u * (float)p.cw - 0.5f));
      const int ty = __float2int_rz(roundf(w * (float)p.ch - 0.5f));
      const int txc = min(max(tx, 0), p.cw - 1), tyc = min(max(ty, 0), p.ch - 1);  // in range on contract
      const uint8_t* t = p.color + 3 * ((size_t)tyc * p.cw + txc);
      p.colors[3 * (size_t)v + 0] = (float)t[0];
      p.colors[3 * (size_t)v + 1] = (float)t[1];
      p.colors[3 * (size_t)v + 2] = (float)t[2];
    }
    if (p.with_uv) {  // rgbd_util.cc:192-195
      p.uv[2 * (size_t)v + 0] = u;
      p.uv[2 * (size_t)v + 1] = 1.0f - w;
    }
  }
}

// D2M-K4: vertex ids / first face ids of every grid point (CTA-local scan + the scanned CTA sums),
// and everything that belongs to the vertex alone: position, uv, nearest colour, point-cloud texel.
__global__ void __launch_bounds__(256) d2m_emit_vertices_kernel(const D2MParams p) {
  const int G = p.GX * p.GY;
  const int first = d2m_first(p);
  // The CTA owns the same kD2MBlock consecutive grid points as in d2m_count_kernel, but walks them as
  // kD2MItems slabs of 256 with lane = consecutive grid point: loads of depth and stores of the per-point
  // images are coalesced, and the compacted vertex ids of a warp are (nearly) consecutive, so the
  // vertex / uv / colour stores of a warp land in adjacent sectors.  One CTA scan per slab, running carry.
  unsigned long long carry = p.block_sums[blockIdx.x];
#pragma unroll 1
  for (int k = 0; k < kD2MItems; k++) {
    const int g = blockIdx.x * kD2MBlock + k * 256 + (int)threadIdx.x;
    unsigned tris;
    const unsigned long long c = d2m_contrib(p, g, G, first, tris);
    unsigned long long total;
    const unsigned long long run = carry + d2m_block_scan(c, total);
    carry += total;
    if (g >= G) continue;
    const int gy = d2m_row(p, g), gx = g - gy * p.GX;
    const float d = d2m_depth(p, gx, gy);
    d2m_emit_image(p, g, gx, gy, c != 0ull, d);
    if (c != 0ull) d2m_emit_vertex(p, g, gx, gy, d, run, tris);
  }
}

// D2M-K2..4 in ONE pass (decoupled look-back, Merrill & Garland): a CTA takes the next chunk of kD2MBlock grid
// points by ticket (so every predecessor chunk is already running or done: the spin below cannot deadlock), scans
// its own contributions, publishes its aggregate, sums the predecessors' aggregates back to the nearest published
// inclusive prefix (warp 0, 32 chunks per step), publishes its own inclusive prefix and emits.  The depth image
// is read once instead of twice and the count / scan launches are gone.
// state word: flag << 62 (0 = nothing yet, 1 = aggregate, 2 = inclusive prefix) | faces << 31 | vertices
// (both < 2^29: the host bounds W * H below 2^28).  64-bit stores are single-copy atomic, so no fence is needed
// between value and flag.
constexpr unsigned long long kD2MAggregate = 1ull << 62, kD2MInclusive = 2ull << 62;
constexpr unsigned long long kD2MStateValue = (1ull << 62) - 1ull;
__device__ __forceinline__ unsigned long long d2m_pack31(unsigned long long c) {  // faces << 32 | vertices -> << 31
  return (c & 0x7FFFFFFFull) | ((c >> 32) << 31);
}
__device__ __forceinline__ unsigned long long d2m_unpack31(unsigned long long s) {
  return (s & 0x7FFFFFFFull) | (((s >> 31) & 0x7FFFFFFFull) << 32);
}
__device__ __forceinline__ unsigned long long d2m_ld_state(const unsigned long long* q) {
  return *(const volatile unsigned long long*)q;
}

// next slab of the CTA: grid point g + 256 from (gx, gy) without a division (rows of >= 256 points)
__device__ __forceinline__ void d2m_step256(const D2MParams& p, int g, int& gx, int& gy) {
  if (p.GX >= 256) {
    gx += 256;
    if (gx >= p.GX) {
      gx -= p.GX;
      gy++;
    }
  } else {
    gy = d2m_row(p, g + 256);
    gx = (g + 256) - gy * p.GX;
  }
}

__global__ void __launch_bounds__(256) d2m_scan_emit_kernel(const D2MParams p, int nblocks) {
  __shared__ int s_chunk;
  __shared__ unsigned s_part[32];  // [slab * 8 + warp]: vertices | faces << 16 of the warp's 32 points of the slab
  __shared__ unsigned long long s_prefix;
  d2m_pdl_launch_dependents();
  d2m_pdl_wait();  // `first` of d2m_first_kernel
  if (threadIdx.x == 0) s_chunk = atomicAdd(p.ticket, 1);
  __syncthreads();
  const int chunk = s_chunk;
  const int G = p.GX * p.GY;
  const int first = d2m_first(p);
  const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5, lt = (1u << lane) - 1u;
  const int g0 = chunk * kD2MBlock + (int)threadIdx.x;
  const int gy0 = d2m_row(p, g0), gx0 = g0 - gy0 * p.GX;
  // Local scan.  A point contributes 1 vertex and 0..2 faces, so the warp-level prefix is three ballots and the
  // CTA-level one a 32-entry scan of 16-bit pairs (<= 1024 vertices, <= 2048 faces per chunk).
  float d[kD2MItems];
  unsigned tris[kD2MItems], wex[kD2MItems];
  unsigned valid = 0u;
  {
    int gx = gx0, gy = gy0;
#pragma unroll
    for (int k = 0; k < kD2MItems; k++) {  // lane = consecutive grid point, as in d2m_emit_vertices_kernel
      const int g = g0 + k * 256;
      bool v = false;
      unsigned t = 0u;
      float dd = 0.f;
      if (g < G) {
        dd = d2m_depth(p, gx, gy);
        v = d2m_valid(dd);
        if (v) t = d2m_tris(p, gx, gy, dd, first);
      }
      const unsigned bv = __ballot_sync(0xffffffffu, v), b0 = __ballot_sync(0xffffffffu, t & 1u),
                     b1 = __ballot_sync(0xffffffffu, t & 2u);
      wex[k] = (unsigned)__popc(bv & lt) | ((unsigned)(__popc(b0 & lt) + __popc(b1 & lt)) << 16);
      if (lane == 0u) s_part[k * 8 + warp] = (unsigned)__popc(bv) | ((unsigned)(__popc(b0) + __popc(b1)) << 16);
      d[k] = dd;
      tris[k] = t;
      if (v) valid |= 1u << k;
      d2m_step256(p, g, gx, gy);
    }
  }
  __syncthreads();
  const unsigned part = s_part[lane];
  unsigned incl = part;
#pragma unroll
  for (int s = 1; s < 32; s <<= 1) {
    const unsigned o = __shfl_up_sync(0xffffffffu, incl, s);
    if (lane >= (unsigned)s) incl += o;
  }
  const unsigned tot16 = __shfl_sync(0xffffffffu, incl, 31);
  const unsigned long long total = (unsigned long long)(tot16 & 0xFFFFu) | ((unsigned long long)(tot16 >> 16) << 32);
  unsigned long long excl[kD2MItems];
#pragma unroll
  for (int k = 0; k < kD2MItems; k++) {
    const unsigned e = __shfl_sync(0xffffffffu, incl - part, k * 8 + (int)warp) + wex[k];
    excl[k] = (unsigned long long)(e & 0xFFFFu) | ((unsigned long long)(e >> 16) << 32);
  }
  // The aggregate goes out first; the prefix-free part of the emission runs while the predecessors publish theirs.
  // (Warp 0 going to the look-back BEFORE its share of that work, so that its inclusive prefix is out sooner, measured
  //  slower: 0.094 -> 0.102 ms per call — its stores then sit on the critical path behind the spin.)
  if (threadIdx.x == 0 && chunk > 0) *(volatile unsigned long long*)(p.state + chunk) = kD2MAggregate | d2m_pack31(total);
  {
    int gx = gx0, gy = gy0;
#pragma unroll
    for (int k = 0; k < kD2MItems; k++) {
      const int g = g0 + k * 256;
      if (g < G) d2m_emit_image(p, g, gx, gy, (valid >> k) & 1u, d[k]);
      d2m_step256(p, g, gx, gy);
    }
  }
  if (threadIdx.x < 32u) {
    unsigned long long prefix = 0ull;  // packed31
    if (chunk > 0) {
      for (int base = chunk - 1;; base -= 32) {
        const int idx = base - (int)lane;
        unsigned long long st = idx >= 0 ? d2m_ld_state(p.state + idx) : kD2MInclusive;
        while (__any_sync(0xffffffffu, (st >> 62) == 0ull))
          if ((st >> 62) == 0ull) st = d2m_ld_state(p.state + idx);
        const unsigned inc = __ballot_sync(0xffffffffu, (st >> 62) == 2ull);
        const int stop = inc ? __ffs(inc) - 1 : 32;  // nearest inclusive prefix: lanes past it are not needed
        unsigned long long v = ((int)lane <= stop) ? (st & kD2MStateValue) : 0ull;
#pragma unroll
        for (int s = 16; s > 0; s >>= 1) v += __shfl_xor_sync(0xffffffffu, v, s);
        prefix += v;
        if (inc) break;
      }
    }
    if (lane == 0u) {
      *(volatile unsigned long long*)(p.state + chunk) = kD2MInclusive | (prefix + d2m_pack31(total));
      s_prefix = d2m_unpack31(prefix);
      if (chunk == nblocks - 1) p.totals[0] = d2m_unpack31(prefix) + total;
    }
  }
  __syncthreads();
  const unsigned long long prefix = s_prefix;
  {
    int gx = gx0, gy = gy0;
#pragma unroll
    for (int k = 0; k < kD2MItems; k++) {
      const int g = g0 + k * 256;
      if ((valid >> k) & 1u) d2m_emit_vertex(p, g, gx, gy, d[k], prefix + excl[k], tris[k]);
      d2m_step256(p, g, gx, gy);
    }
  }
}

// Vector3f::normalized(): x*x + (y*y + z*z), true divisions (Eigen Dot.h:121-145)
__device__ __forceinline__ void d2m_normalize(float (&v)[3]) {
  const float z = v[0] * v[0] + (v[1] * v[1] + v[2] * v[2]);
  if (z > 0.0f) {
    const float s = sqrtf(z);
    v[0] = v[0] / s;
    v[1] = v[1] / s;
    v[2] = v[2] / s;
  }
}

// D2M-K5: the (up to two) triangles of every vertex grid point and their face normals
// (Mesh::CalcFaceNormal, src/mesh.cc:339-349).
__global__ void __launch_bounds__(256) d2m_emit_faces_kernel(const D2MParams p) {
  d2m_pdl_launch_dependents();
  d2m_pdl_wait();
  const int G = p.GX * p.GY;
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= G) return;
  const unsigned fr = p.fref[g];
  const unsigned tris = fr >> 30;
  if (tris == 0u) return;
  unsigned f = fr & 0x3FFFFFFFu;
  const int cur = p.vid[g], ul = p.vid[g - p.GX - 1];
  const float* vc = p.vertices + 3 * (size_t)cur;
  const float* vu = p.vertices + 3 * (size_t)ul;
  // the diagonal current - upper_left is an edge of both triangles: normalised once
  float dg[3] = {vc[0] - vu[0], vc[1] - vu[1], vc[2] - vu[2]};
  d2m_normalize(dg);
#pragma unroll
  for (int t = 0; t < 2; t++) {
    if (!(tris & (1u << t))) continue;
    // t = 0: (upper_left, current, upper); t = 1: (upper_left, left, current)
    const int other = t == 0 ? p.vid[g - p.GX] : p.vid[g - 1];
    const float* vo = p.vertices + 3 * (size_t)other;
    p.faces[3 * (size_t)f + 0] = ul;
    p.faces[3 * (size_t)f + 1] = t == 0 ? cur : other;
    p.faces[3 * (size_t)f + 2] = t == 0 ? other : cur;
    float e[3] = {vo[0] - vu[0], vo[1] - vu[1], vo[2] - vu[2]};
    d2m_normalize(e);
    const float* a = t == 0 ? dg : e;  // (v1 - v0).normalized()
    const float* b = t == 0 ? e : dg;  // (v2 - v0).normalized()
    float n[3] = {a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0]};
    d2m_normalize(n);
    p.face_normals[3 * (size_t)f + 0] = n[0];
    p.face_normals[3 * (size_t)f + 1] = n[1];
    p.face_normals[3 * (size_t)f + 2] = n[2];
    f++;
  }
}

// D2M-K6: vertex normals (Mesh::CalcNormal, src/mesh.cc:310-336): sum of the adjacent face normals
// in face order, divided by their number, normalised; zero for unreferenced vertices.
__global__ void __launch_bounds__(256) d2m_vertex_normals_kernel(const D2MParams p) {
  d2m_pdl_launch_dependents();
  d2m_pdl_wait();
  const int G = p.GX * p.GY;
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= G) return;
  const int v = p.vid[g];
  const int gy = d2m_row(p, g), gx = g - gy * p.GX;
  if (p.zero_fill && p.normal_image) {  // steps (1, 1): the zeros of Init(normal, .., 0.0f), as in d2m_emit_point
    const int x = gx + 1, y = gy + 1;
    if (gx == 0) d2m_zero3(p.normal_image + 3 * ((size_t)y * p.W));
    if (gy == 0) d2m_zero3(p.normal_image + 3 * (size_t)x);
    if (g == 0) d2m_zero3(p.normal_image);
    if (v < 0) d2m_zero3(p.normal_image + 3 * ((size_t)y * p.W + x));
  }
  if (v < 0) return;
  float s[3] = {0.f, 0.f, 0.f};
  int cnt = 0;
  auto add = [&](unsigned f) {
    s[0] += p.face_normals[3 * (size_t)f + 0];
    s[1] += p.face_normals[3 * (size_t)f + 1];
    s[2] += p.face_normals[3 * (size_t)f + 2];
    cnt++;
  };
  auto visit = [&](int q, unsigned want) {  // the triangles `want` (bit 0 = first, bit 1 = second) of grid point q
    const unsigned fr = p.fref[q];
    const unsigned tris = fr >> 30;
    unsigned f = fr & 0x3FFFFFFFu;
    if (tris & 1u) {
      if (want & 1u) add(f);
      f++;
    }
    if ((tris & 2u) && (want & 2u)) add(f);
  };
  visit(g, 3u);                                              // current in both of its own triangles
  if (gx + 1 < p.GX) visit(g + 1, 2u);                       // `left` of the right neighbour's second
  if (gy + 1 < p.GY) visit(g + p.GX, 1u);                    // `upper` of the lower neighbour's first
  if (gx + 1 < p.GX && gy + 1 < p.GY) visit(g + p.GX + 1, 3u);  // `upper_left` of the lower-right's two
  if (cnt > 0) {
    const float c = (float)cnt;
    s[0] = s[0] / c;
    s[1] = s[1] / c;
    s[2] = s[2] / c;
    d2m_normalize(s);
  } else {
    s[0] = s[1] = s[2] = 0.f;
  }
  p.normals[3 * (size_t)v + 0] = s[0];
  p.normals[3 * (size_t)v + 1] = s[1];
  p.normals[3 * (size_t)v + 2] = s[2];
  if (p.normal_image) {  // rgbd_util.cc:268-277
    float* q = p.normal_image + 3 * ((size_t)((gy + 1) * p.ys) * p.W + (size_t)((gx + 1) * p.xs));
    q[0] = s[0]; q[1] = s[1]; q[2] = s[2];
  }
}

}  // namespace ugu_b200
