// sm_100a kernels for UV-space attribute rasterization:
//   ugu::RasterizeVertexAttributeToTexture / RasterizeFaceAttributeToTexture / GenerateUvMask
//   (reference include/ugu/util/raster_util.h:78-268, src/util/raster_util.cc:9-27).
//
// The reference walks the faces in ascending index and lets the LAST writer of a texel win (with
// num_threads > 1 it is racy; the sequential order, num_threads = 1, is the one reproduced here).
// "Last writer in face order" == "highest covering face index", which is order independent, so the
// device path is: (1) per-face bbox -> 8x4-texel tile jobs appended to a queue (warp prefix sum, one
// atomicAdd per warp), (2) a persistent, globally load-balanced tile kernel that evaluates the three
// edge functions per texel and does atomicMax(face) on a W*H winner buffer, (3) a resolve kernel that
// recomputes the winner's weights bit-identically, interpolates the attribute and writes the packed
// texture (+ mask), re-clearing the winner buffer for the next call.
//
// Float discipline as in raster_kernels.cuh: --fmad=false, single IEEE operations in the reference's
// association order, denormals kept (no -ftz).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "raster_kernels.cuh"

namespace ugu_b200 {

struct UvRasterParams {
  int W, H, F;
  int per_face;           // 1: attrs[f] (RasterizeFaceAttributeToTexture), 0: attrs[attr_idx[f][k]]
  int texture_f32;        // 0: 3 x u8 (Image3b), 1: 3 x f32 (Image3f)
  int keep_texture;       // 1: uncovered texels keep their contents (reference width/height <= 0 form)
  int keep_mask;          // 1: uncovered mask texels keep their contents (the reference never clears it)
  const float* attrs;     // AoS, 3 floats per attribute
  const int* attr_idx;    // AoS, 3 ints per face (unused when per_face)
  const float* uv;        // AoS, 2 floats per uv
  const int* uv_idx;      // AoS, 3 ints per face
  int* win;               // W*H winner face index, -1 = none (left all -1 by the resolve kernel)
  unsigned long long* counter;  // packed queue counter: [63:38] entries, [37:0] tile jobs
  int* q_face;            // F
  unsigned long long* q_job;    // F: first tile job of each queue entry
  float4* rec;            // F x 64 B per-face records written by uv_setup_kernel, read by the resolve:
                          // {x0 y0 x1 y1} {x2 y2 inv_area a0.x} {a0.y a0.z a1.x a1.y} {a1.z a2.x a2.y a2.z}
  void* texture;          // W*H*3 u8 or f32
  uint8_t* mask;          // W*H or null
};

// static_cast<int>(float) as the x86-64 reference compiles it (cvttss2si): truncation; out of range
// and NaN give INT_MIN.
__device__ __forceinline__ int cvtt_x86(float f) {
  if (!(f > -2147483904.0f && f < 2147483648.0f)) return (int)0x80000000;
  return __float2int_rz(f);
}

constexpr int kUvInlineMax = 64;     // bbox texels up to which a face is rasterised inside uv_setup_kernel
constexpr int kUvJobsPerChunk = 64;  // tile jobs a warp claims at a time in uv_tile_kernel

struct UvTri {
  float x0, y0, x1, y1, x2, y2;
  float inv_area;
};

// target_tri of raster_util.h:209-217 (U2X / V2Y with flip, :18-21) and area / inv_area of :83-88
__device__ __forceinline__ void uv_tri_setup(const UvRasterParams& p, int f, UvTri& t) {
  const int i0 = __ldg(p.uv_idx + 3 * (size_t)f + 0);
  const int i1 = __ldg(p.uv_idx + 3 * (size_t)f + 1);
  const int i2 = __ldg(p.uv_idx + 3 * (size_t)f + 2);
  const float2 a = __ldg(reinterpret_cast<const float2*>(p.uv) + i0);
  const float2 b = __ldg(reinterpret_cast<const float2*>(p.uv) + i1);
  const float2 c = __ldg(reinterpret_cast<const float2*>(p.uv) + i2);
  const float w = (float)p.W, h = (float)p.H;
  t.x0 = a.x * w - 0.5f; t.y0 = (1.f - a.y) * h - 0.5f;
  t.x1 = b.x * w - 0.5f; t.y1 = (1.f - b.y) * h - 0.5f;
  t.x2 = c.x * w - 0.5f; t.y2 = (1.f - c.y) * h - 0.5f;
  // EdgeFunction(t0, t1, t2) = (t2.x - t0.x)*(t1.y - t0.y) - (t2.y - t0.y)*(t1.x - t0.x)
  float area = (t.x2 - t.x0) * (t.y1 - t.y0) - (t.y2 - t.y0) * (t.x1 - t.x0);
  if (fabsf(area) < 1.17549435e-38f) area = area > 0.f ? 1.17549435e-38f : -1.17549435e-38f;
  t.inv_area = 1.0f / area;
}

// bbox of raster_util.h:90-104: (int)(min - 1) .. (int)(max + 1), each clamped into the image
__device__ __forceinline__ void uv_tri_bounds(const UvRasterParams& p, const UvTri& t, int& bx0, int& by0,
                                              int& bx1, int& by1) {
  const float xmin = fminf(fminf(t.x0, t.x1), t.x2), xmax = fmaxf(fmaxf(t.x0, t.x1), t.x2);
  const float ymin = fminf(fminf(t.y0, t.y1), t.y2), ymax = fmaxf(fmaxf(t.y0, t.y1), t.y2);
  bx0 = max(0, min(cvtt_x86(xmin - 1.f), p.W - 1));
  bx1 = max(0, min(cvtt_x86(xmax + 1.f), p.W - 1));
  by0 = max(0, min(cvtt_x86(ymin - 1.f), p.H - 1));
  by1 = max(0, min(cvtt_x86(ymax + 1.f), p.H - 1));
}

// w_k of raster_util.h:121-129: EdgeFunction(t1,t2,p), (t2,t0,p), (t0,t1,p), each times inv_area
__device__ __forceinline__ void uv_weights(const UvTri& t, float px, float py, float& w0, float& w1,
                                           float& w2) {
  w0 = (px - t.x1) * (t.y2 - t.y1) - (py - t.y1) * (t.x2 - t.x1);
  w1 = (px - t.x2) * (t.y0 - t.y2) - (py - t.y2) * (t.x0 - t.x2);
  w2 = (px - t.x0) * (t.y1 - t.y0) - (py - t.y0) * (t.x1 - t.x0);
  w0 = w0 * t.inv_area;
  w1 = w1 * t.inv_area;
  w2 = w2 * t.inv_area;
}

// ---------------------------------------------------------------------------------------------
// UV-K1: per-face setup.  Lane = face.  Faces whose bbox holds <= kUvInlineMax texels are rasterised
// here by their own lane; larger ones become tile jobs: the warp's entries are appended to the queue
// with one atomicAdd on the packed (entries, jobs) counter, so q_job is monotone over queue slots.
// ---------------------------------------------------------------------------------------------
// (The three kernels of a call form a programmatic-dependent-launch chain: each lets its successor be scheduled as it
//  starts and waits for its predecessor before it touches anything.)
__global__ void __launch_bounds__(256) uv_setup_kernel(const UvRasterParams p) {
  pdl_launch_dependents();
  pdl_wait();
  const int f = blockIdx.x * blockDim.x + threadIdx.x;
  const unsigned lane = threadIdx.x & 31u;
  unsigned long long mine = 0ull;
  if (f < p.F) {
    UvTri t;
    uv_tri_setup(p, f, t);
    {  // per-face record for the resolve: the projected triangle and the three corner attributes
      int a0 = f, a1 = f, a2 = f;
      if (!p.per_face) {
        a0 = __ldg(p.attr_idx + 3 * (size_t)f + 0);
        a1 = __ldg(p.attr_idx + 3 * (size_t)f + 1);
        a2 = __ldg(p.attr_idx + 3 * (size_t)f + 2);
      }
      const float* q0 = p.attrs + 3 * (size_t)a0;
      const float* q1 = p.attrs + 3 * (size_t)a1;
      const float* q2 = p.attrs + 3 * (size_t)a2;
      float4* r = p.rec + 4 * (size_t)f;
      r[0] = make_float4(t.x0, t.y0, t.x1, t.y1);
      r[1] = make_float4(t.x2, t.y2, t.inv_area, __ldg(q0 + 0));
      r[2] = make_float4(__ldg(q0 + 1), __ldg(q0 + 2), __ldg(q1 + 0), __ldg(q1 + 1));
      r[3] = make_float4(__ldg(q1 + 2), __ldg(q2 + 0), __ldg(q2 + 1), __ldg(q2 + 2));
    }
    int x0, y0, x1, y1;
    uv_tri_bounds(p, t, x0, y0, x1, y1);
    if (x1 >= x0 && y1 >= y0) {
      const int nx = x1 - x0 + 1;
      const long long ns = (long long)nx * (y1 - y0 + 1);
      if (ns <= (long long)kUvInlineMax) {
        // small bbox: walked by this lane (a warp covers 32 faces per step; meshes whose uv
        // triangles are of similar size keep the lanes busy), ~25 instructions per texel instead of
        // a queue entry + an 8x4 tile job per face
        int x = x0, y = y0;
        for (int k = 0; k < (int)ns; k++) {
          float w0, w1, w2;
          uv_weights(t, (float)x, (float)y, w0, w1, w2);
          if (!(w0 < 0.f || w1 < 0.f || w2 < 0.f)) atomicMax(p.win + ((size_t)y * p.W + x), f);
          if (++x > x1) {
            x = x0;
            ++y;
          }
        }
      } else {
        const unsigned long long njobs = (unsigned long long)((x1 - x0 + kJobW) / kJobW) *
                                         (unsigned long long)((y1 - y0 + kJobH) / kJobH);
        mine = (1ull << kQueueJobBits) | njobs;
      }
    }
  }
  if (!__any_sync(0xffffffffu, mine != 0ull)) return;
  unsigned long long incl = mine;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const unsigned long long o = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= (unsigned)d) incl += o;
  }
  const unsigned long long total = __shfl_sync(0xffffffffu, incl, 31);
  unsigned long long qbase = 0ull;
  if (lane == 31u) qbase = atomicAdd(p.counter, total);
  qbase = __shfl_sync(0xffffffffu, qbase, 31);
  if (mine != 0ull) {
    const unsigned long long excl = qbase + incl - mine;
    const unsigned qslot = (unsigned)(excl >> kQueueJobBits);
    p.q_face[qslot] = f;
    p.q_job[qslot] = excl & kQueueJobMask;
  }
}

// ---------------------------------------------------------------------------------------------
// UV-K2: tile jobs.  Persistent grid; a warp claims interleaved chunks of the global job space,
// finds the owning queue entry with a warp-cooperative 32-ary search and walks entries; a job is an
// 8x4 texel block of the face's bbox, one lane per texel.  Covered texels (w0, w1, w2 all >= 0 after
// the multiplication by inv_area, raster_util.h:127-135) vote with atomicMax(face).
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) uv_tile_kernel(const UvRasterParams p) {
  pdl_launch_dependents();
  pdl_wait();
  const unsigned long long packed = p.counter[0];
  const unsigned long long T = packed & kQueueJobMask;
  const unsigned n = (unsigned)(packed >> kQueueJobBits);
  if (T == 0ull) return;
  const unsigned lane = threadIdx.x & 31u;
  const unsigned long long nwarps = (unsigned long long)gridDim.x * (blockDim.x >> 5);
  const unsigned long long gw = (unsigned long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const unsigned long long rounds =
      (T + nwarps * (unsigned long long)kUvJobsPerChunk - 1) / (nwarps * (unsigned long long)kUvJobsPerChunk);
  unsigned long long per = (T + nwarps * rounds - 1) / (nwarps * rounds);
  per = per < 2ull ? 2ull : per;
  const unsigned long long nchunks = (T + per - 1) / per;
  const int lx = (int)(lane & (kJobW - 1)), ly = (int)(lane >> 3);

  for (unsigned long long chunk = gw; chunk < nchunks; chunk += nwarps) {
    const unsigned long long jb = chunk * per;
    const unsigned long long je = (jb + per < T) ? jb + per : T;
    unsigned lo = 0, hi = n;  // largest e with q_job[e] <= jb (q_job of the first entry is 0)
    while (hi - lo > 1u) {
      const unsigned span = hi - lo;
      const unsigned step = (span + 31u) / 32u;
      const unsigned idx = lo + lane * step;
      const bool pred = (idx < hi) && (p.q_job[idx] <= jb);
      const unsigned cnt = (unsigned)__popc(__ballot_sync(0xffffffffu, pred));
      const unsigned nlo = lo + (cnt - 1u) * step;
      const unsigned nhi = lo + cnt * step;
      lo = nlo;
      hi = (nhi < hi) ? nhi : hi;
    }
    unsigned e = lo;
    unsigned long long j = jb;
    while (j < je) {
      // Batch of up to 31 queue entries, lane = entry: the dependent q_face -> uv_idx -> uv loads and
      // the triangle setup of the whole batch are in flight together (one latency chain per batch
      // instead of one per entry); lane i + 1 supplies the job range end of entry i.
      const unsigned me = e + lane;
      const unsigned long long js_l = (me < n) ? p.q_job[me] : T;
      const bool mine = (me < n) && (js_l < je) && (lane < 31u);
      UvTri t_l;
      int bx0 = 0, by0 = 0, bx1 = -1, by1 = -1, f_l = -1;
      if (mine) {
        f_l = p.q_face[me];
        uv_tri_setup(p, f_l, t_l);
        uv_tri_bounds(p, t_l, bx0, by0, bx1, by1);
      }
      const int nb = __popc(__ballot_sync(0xffffffffu, mine));  // a prefix of the lanes: q_job is monotone
      for (int i = 0; i < nb; i++) {
        const int f = __shfl_sync(0xffffffffu, f_l, i);
        const unsigned long long js = __shfl_sync(0xffffffffu, js_l, i);
        const unsigned long long jn = __shfl_sync(0xffffffffu, js_l, i + 1);
        const unsigned long long jend = (jn < je) ? jn : je;
        UvTri t;
        t.x0 = __shfl_sync(0xffffffffu, t_l.x0, i);
        t.y0 = __shfl_sync(0xffffffffu, t_l.y0, i);
        t.x1 = __shfl_sync(0xffffffffu, t_l.x1, i);
        t.y1 = __shfl_sync(0xffffffffu, t_l.y1, i);
        t.x2 = __shfl_sync(0xffffffffu, t_l.x2, i);
        t.y2 = __shfl_sync(0xffffffffu, t_l.y2, i);
        t.inv_area = __shfl_sync(0xffffffffu, t_l.inv_area, i);
        const int x0 = __shfl_sync(0xffffffffu, bx0, i), y0 = __shfl_sync(0xffffffffu, by0, i);
        const int x1 = __shfl_sync(0xffffffffu, bx1, i), y1 = __shfl_sync(0xffffffffu, by1, i);
        const unsigned ntx = (unsigned)((x1 - x0 + kJobW) / kJobW);
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
            uv_weights(t, (float)x, (float)y, w0, w1, w2);
            if (!(w0 < 0.f || w1 < 0.f || w2 < 0.f)) atomicMax(p.win + ((size_t)y * p.W + x), f);
          }
        }
        j = jend;
      }
      e += (unsigned)nb;
    }
  }
}

// ---------------------------------------------------------------------------------------------
// UV-K3: resolve.  4 consecutive texels per thread: read + re-clear the winners (one 16-byte access
// each way), recompute the winner's weights with the expressions of UV-K2, interpolate
// (w0*c0 + w1*c1) + w2*c2 (raster_util.h:138-139), convert (saturate_cast, :106-115: std::round for
// integer texels, then the truncating cast) and store 12 / 48 packed bytes + 4 mask bytes.
// ---------------------------------------------------------------------------------------------
// (uint8_t)std::round(c) through cvttss2si.  std::round = half away from zero: truncate (one FRND), the exact
// remainder decides the step — the same integer as roundf's generic sequence for every |c| < 2^23, which covers the
// whole range whose low byte can be non-zero... values beyond are integers already (round is the identity) up to
// the conversion's range, and out-of-range / NaN inputs give the low byte of INT_MIN = 0.
__device__ __forceinline__ unsigned uv_to_u8(float c) {
  if (!(c > -2147483904.0f && c < 2147483648.0f)) return 0u;
  const float t = truncf(c);
  const float fr = c - t;  // exact
  const int step = (fr >= 0.5f) ? 1 : ((fr <= -0.5f) ? -1 : 0);
  return (unsigned)(__float2int_rz(t) + step) & 0xFFu;
}

constexpr int kUvGroupsPerThread = 4;  // 4-texel groups per thread: 64 bytes of winners in flight per thread

// One group of 4 consecutive texels starting at `base` whose winners are f[0..3].
template <bool RUNS>
__device__ __forceinline__ void uv_resolve_group(const UvRasterParams& p, size_t P, size_t base, bool full,
                                                 const int (&f)[4]) {
  if (full && (f[0] & f[1] & f[2] & f[3]) < 0) {
    // nothing covers these four texels (most of an atlas's gutters): the cleared values, or nothing at all
    if (!p.keep_texture) {
      if (p.texture_f32) {
        float4* o = reinterpret_cast<float4*>(reinterpret_cast<float*>(p.texture) + base * 3);
        o[0] = o[1] = o[2] = make_float4(0.f, 0.f, 0.f, 0.f);
      } else {
        unsigned* o = reinterpret_cast<unsigned*>(reinterpret_cast<uint8_t*>(p.texture) + base * 3);
        o[0] = o[1] = o[2] = 0u;
      }
    }
    if (p.mask && !p.keep_mask) *reinterpret_cast<unsigned*>(p.mask + base) = 0u;
    return;
  }
  float c[4][3];
  bool cov[4];
  if (RUNS) {
    // Long runs of one winner (few large faces): the (up to 4) distinct records of the group are requested together
    // as L1 prefetches, which hold no registers, and then ONE record is live at a time — a texel whose winner equals
    // its left neighbour's keeps the registers, another winner reloads them from the prefetched lines.  64 registers
    // and three CTAs per SM instead of 123 and two: vc2tex 0.157 -> 0.129 ms.  (Meshes of texel-sized faces change
    // winner at every texel and want the four loads in flight together: posmap 0.110 -> 0.116 ms with this form;
    // a two-record pipeline lost on both.)
#pragma unroll
    for (int k = 0; k < 4; k++) {
      cov[k] = f[k] >= 0;
      if (cov[k] && (k == 0 || f[k] != f[k - 1])) {
        const float4* q = p.rec + 4 * (size_t)f[k];
        asm volatile("prefetch.global.L1 [%0];" ::"l"(q));
      }
    }
    int y = (int)((unsigned)base / (unsigned)p.W), x = (int)((unsigned)base - (unsigned)y * (unsigned)p.W);
    float4 r0 = make_float4(0.f, 0.f, 0.f, 0.f), r1 = r0, r2 = r0, r3 = r0;
#pragma unroll
    for (int k = 0; k < 4; k++) {
      c[k][0] = c[k][1] = c[k][2] = 0.f;
      if (cov[k]) {
        if (k == 0 || f[k] != f[k - 1]) {
          const float4* q = p.rec + 4 * (size_t)f[k];
          r0 = q[0];  // plain loads: written by uv_setup_kernel of this call
          r1 = q[1];
          r2 = q[2];
          r3 = q[3];
        }
        UvTri t;
        t.x0 = r0.x; t.y0 = r0.y; t.x1 = r0.z; t.y1 = r0.w;
        t.x2 = r1.x; t.y2 = r1.y; t.inv_area = r1.z;
        float w0, w1, w2;
        uv_weights(t, (float)x, (float)y, w0, w1, w2);
        c[k][0] = (w0 * r1.w + w1 * r2.z) + w2 * r3.y;
        c[k][1] = (w0 * r2.x + w1 * r2.w) + w2 * r3.z;
        c[k][2] = (w0 * r2.y + w1 * r3.x) + w2 * r3.w;
      }
      if (++x == p.W) {
        x = 0;
        ++y;
      }
    }
  } else {
    // One dependent level after the winners: the 64-byte face records of the (up to 4) distinct
    // winners are requested together; a texel whose winner equals its left neighbour's reuses the
    // neighbour's registers.
    float4 r[4][4];
  #pragma unroll
    for (int k = 0; k < 4; k++) {
      cov[k] = f[k] >= 0;
      const bool need = cov[k] && (k == 0 || f[k] != f[k - 1]);
      if (need) {
        const float4* q = p.rec + 4 * (size_t)f[k];
        r[k][0] = q[0];  // plain loads: written by uv_setup_kernel of this call
        r[k][1] = q[1];
        r[k][2] = q[2];
        r[k][3] = q[3];
      }
    }
    int y = (int)((unsigned)base / (unsigned)p.W), x = (int)((unsigned)base - (unsigned)y * (unsigned)p.W);
  #pragma unroll
    for (int k = 0; k < 4; k++) {
      if (k > 0 && cov[k] && f[k] == f[k - 1]) {
        r[k][0] = r[k - 1][0];
        r[k][1] = r[k - 1][1];
        r[k][2] = r[k - 1][2];
        r[k][3] = r[k - 1][3];
      }
      c[k][0] = c[k][1] = c[k][2] = 0.f;
      if (cov[k]) {
        UvTri t;
        t.x0 = r[k][0].x; t.y0 = r[k][0].y; t.x1 = r[k][0].z; t.y1 = r[k][0].w;
        t.x2 = r[k][1].x; t.y2 = r[k][1].y; t.inv_area = r[k][1].z;
        float w0, w1, w2;
        uv_weights(t, (float)x, (float)y, w0, w1, w2);
        c[k][0] = (w0 * r[k][1].w + w1 * r[k][2].z) + w2 * r[k][3].y;
        c[k][1] = (w0 * r[k][2].x + w1 * r[k][2].w) + w2 * r[k][3].z;
        c[k][2] = (w0 * r[k][2].y + w1 * r[k][3].x) + w2 * r[k][3].w;
      }
      if (++x == p.W) {
        x = 0;
        ++y;
      }
    }
  }
  const bool any = cov[0] || cov[1] || cov[2] || cov[3];
  if (full && !p.keep_texture) {
    if (p.texture_f32) {
      float4* o = reinterpret_cast<float4*>(reinterpret_cast<float*>(p.texture) + base * 3);
      o[0] = make_float4(c[0][0], c[0][1], c[0][2], c[1][0]);
      o[1] = make_float4(c[1][1], c[1][2], c[2][0], c[2][1]);
      o[2] = make_float4(c[2][2], c[3][0], c[3][1], c[3][2]);
    } else {
      unsigned b[12];
#pragma unroll
      for (int k = 0; k < 4; k++)
#pragma unroll
        for (int q = 0; q < 3; q++) b[k * 3 + q] = cov[k] ? uv_to_u8(c[k][q]) : 0u;
      unsigned* o = reinterpret_cast<unsigned*>(reinterpret_cast<uint8_t*>(p.texture) + base * 3);
      o[0] = b[0] | (b[1] << 8) | (b[2] << 16) | (b[3] << 24);
      o[1] = b[4] | (b[5] << 8) | (b[6] << 16) | (b[7] << 24);
      o[2] = b[8] | (b[9] << 8) | (b[10] << 16) | (b[11] << 24);
    }
  } else if (any || !p.keep_texture) {
    for (int k = 0; k < 4 && base + k < P; k++) {
      if (!cov[k] && p.keep_texture) continue;
      for (int q = 0; q < 3; q++) {
        if (p.texture_f32)
          reinterpret_cast<float*>(p.texture)[(base + k) * 3 + q] = c[k][q];
        else
          reinterpret_cast<uint8_t*>(p.texture)[(base + k) * 3 + q] = cov[k] ? (uint8_t)uv_to_u8(c[k][q]) : 0;
      }
    }
  }
  if (p.mask) {
    if (full && !p.keep_mask) {
      *reinterpret_cast<unsigned*>(p.mask + base) =
          (cov[0] ? 0xFFu : 0u) | (cov[1] ? 0xFF00u : 0u) | (cov[2] ? 0xFF0000u : 0u) | (cov[3] ? 0xFF000000u : 0u);
    } else {
      for (int k = 0; k < 4 && base + k < P; k++) {
        if (cov[k]) p.mask[base + k] = 255;
        else if (!p.keep_mask) p.mask[base + k] = 0;
      }
    }
  }
}

// A CTA owns 256 * 16 consecutive texels; group j of thread t is the 4 texels at
// cta_base + j * 1024 + 4 t (every load / store instruction of a warp is contiguous).  All winner
// loads of a thread are issued before the first is consumed: with the L2 cold, a single 16-byte
// request per thread leaves too few bytes in flight to cover the DRAM latency.
template <bool RUNS>
__global__ void __launch_bounds__(256, RUNS ? 3 : 2) uv_resolve_kernel(const UvRasterParams p) {
  pdl_launch_dependents();
  pdl_wait();
  const size_t P = (size_t)p.W * p.H;
  if (blockIdx.x == 0 && threadIdx.x == 0) p.counter[0] = 0ull;  // queue reset for the next call
  const size_t cta_base = (size_t)blockIdx.x * (256u * 4u * kUvGroupsPerThread);
  int f[kUvGroupsPerThread][4];
  size_t base[kUvGroupsPerThread];
#pragma unroll
  for (int j = 0; j < kUvGroupsPerThread; j++) {
    base[j] = cta_base + (size_t)j * 1024u + (size_t)threadIdx.x * 4u;
    f[j][0] = f[j][1] = f[j][2] = f[j][3] = -1;
    if (base[j] + 3 < P) {
      const int4 w = *reinterpret_cast<const int4*>(p.win + base[j]);
      f[j][0] = w.x; f[j][1] = w.y; f[j][2] = w.z; f[j][3] = w.w;
    } else {
      for (int k = 0; k < 4 && base[j] + k < P; k++) f[j][k] = p.win[base[j] + k];
    }
  }
#pragma unroll
  for (int j = 0; j < kUvGroupsPerThread; j++) {
    if (base[j] >= P) continue;
    const bool full = base[j] + 3 < P;
    if (full) {
      *reinterpret_cast<int4*>(p.win + base[j]) = make_int4(-1, -1, -1, -1);
    } else {
      for (int k = 0; k < 4 && base[j] + k < P; k++) p.win[base[j] + k] = -1;
    }
    uv_resolve_group<RUNS>(p, P, base[j], full, f[j]);
  }
}

// UV-K3, lane-per-texel form: a warp owns 32 consecutive texels (chunk), walks chunks with a grid stride, and has
// the next chunk's winners in flight while it works.  One face record (16 registers) per thread instead of four, so
// the kernel runs at 6 CTAs per SM instead of 2 — it is a dependent chain winner -> record -> weights -> store, and
// occupancy is what hides it.  Texels of a chunk are staged in shared memory and leave as full-width vector stores
// (96 B of u8 / 384 B of f32 per chunk, 32 mask bytes); chunks nothing covers write their zeros at once.  The
// in/out forms (keep_texture / keep_mask) and a ragged last chunk store per texel.
__global__ void __launch_bounds__(256, 6) uv_resolve_lane_kernel(const UvRasterParams p) {
  __shared__ __align__(16) float s_f[8][96];
  __shared__ __align__(16) unsigned char s_b[8][96];
  pdl_launch_dependents();
  pdl_wait();
  const unsigned P = (unsigned)p.W * (unsigned)p.H;  // W * H < 2^31 (validated by the host)
  if (blockIdx.x == 0 && threadIdx.x == 0) p.counter[0] = 0ull;  // queue reset for the next call
  const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
  const unsigned nchunks = (P + 31u) >> 5, nwarps = gridDim.x * 8u;
  unsigned chunk = blockIdx.x * 8u + warp;
  int f_next = -1;
  if (chunk < nchunks && chunk * 32u + lane < P) f_next = p.win[chunk * 32u + lane];
  for (; chunk < nchunks; chunk += nwarps) {
    const int f = f_next;
    const unsigned base = chunk * 32u, t = base + lane;
    f_next = -1;
    if (chunk + nwarps < nchunks && t + nwarps * 32u < P) f_next = p.win[t + nwarps * 32u];
    const bool full = base + 32u <= P;
    const bool cov = f >= 0;
    const unsigned any = __ballot_sync(0xffffffffu, cov);
    const bool vec = full && !p.keep_texture;
    if (any == 0u) {
      if (vec) {
        if (p.texture_f32) {
          if (lane < 24u) reinterpret_cast<float4*>(reinterpret_cast<float*>(p.texture) + (size_t)base * 3)[lane] =
              make_float4(0.f, 0.f, 0.f, 0.f);
        } else if (lane < 6u) {
          reinterpret_cast<uint4*>(reinterpret_cast<uint8_t*>(p.texture) + (size_t)base * 3)[lane] = make_uint4(0u, 0u, 0u, 0u);
        }
      } else if (!p.keep_texture && t < P) {
        for (int q = 0; q < 3; q++) {
          if (p.texture_f32) reinterpret_cast<float*>(p.texture)[(size_t)t * 3 + q] = 0.f;
          else reinterpret_cast<uint8_t*>(p.texture)[(size_t)t * 3 + q] = 0;
        }
      }
      if (p.mask && !p.keep_mask) {
        if (full) {
          if (lane < 8u) reinterpret_cast<unsigned*>(p.mask + base)[lane] = 0u;
        } else if (t < P) {
          p.mask[t] = 0;
        }
      }
      continue;
    }
    if (cov) p.win[t] = -1;  // leave the winner buffer clear for the next call
    float c0 = 0.f, c1 = 0.f, c2 = 0.f;
    if (cov) {
      const float4* q = p.rec + 4 * (size_t)f;  // plain loads: written by uv_setup_kernel of this call
      const float4 r0 = q[0], r1 = q[1], r2 = q[2], r3 = q[3];
      const unsigned y = t / (unsigned)p.W, x = t - y * (unsigned)p.W;
      UvTri tr;
      tr.x0 = r0.x; tr.y0 = r0.y; tr.x1 = r0.z; tr.y1 = r0.w;
      tr.x2 = r1.x; tr.y2 = r1.y; tr.inv_area = r1.z;
      float w0, w1, w2;
      uv_weights(tr, (float)x, (float)y, w0, w1, w2);
      c0 = (w0 * r1.w + w1 * r2.z) + w2 * r3.y;
      c1 = (w0 * r2.x + w1 * r2.w) + w2 * r3.z;
      c2 = (w0 * r2.y + w1 * r3.x) + w2 * r3.w;
    }
    if (vec) {
      if (p.texture_f32) {
        s_f[warp][lane * 3u + 0u] = c0;
        s_f[warp][lane * 3u + 1u] = c1;
        s_f[warp][lane * 3u + 2u] = c2;
        __syncwarp();
        if (lane < 24u) reinterpret_cast<float4*>(reinterpret_cast<float*>(p.texture) + (size_t)base * 3)[lane] =
            reinterpret_cast<const float4*>(s_f[warp])[lane];
      } else {
        s_b[warp][lane * 3u + 0u] = cov ? (unsigned char)uv_to_u8(c0) : (unsigned char)0;
        s_b[warp][lane * 3u + 1u] = cov ? (unsigned char)uv_to_u8(c1) : (unsigned char)0;
        s_b[warp][lane * 3u + 2u] = cov ? (unsigned char)uv_to_u8(c2) : (unsigned char)0;
        __syncwarp();
        if (lane < 6u) reinterpret_cast<uint4*>(reinterpret_cast<uint8_t*>(p.texture) + (size_t)base * 3)[lane] =
            reinterpret_cast<const uint4*>(s_b[warp])[lane];
      }
      __syncwarp();  // the staging rows are rewritten by the next chunk
    } else if (t < P && (cov || !p.keep_texture)) {
      if (p.texture_f32) {
        float* o = reinterpret_cast<float*>(p.texture) + (size_t)t * 3;
        o[0] = c0; o[1] = c1; o[2] = c2;
      } else {
        uint8_t* o = reinterpret_cast<uint8_t*>(p.texture) + (size_t)t * 3;
        o[0] = cov ? (uint8_t)uv_to_u8(c0) : 0;
        o[1] = cov ? (uint8_t)uv_to_u8(c1) : 0;
        o[2] = cov ? (uint8_t)uv_to_u8(c2) : 0;
      }
    }
    if (p.mask) {
      if (full && !p.keep_mask) {
        if (lane < 8u) {
          const unsigned nb = (any >> (lane * 4u)) & 15u;
          reinterpret_cast<unsigned*>(p.mask + base)[lane] = ((nb * 0x00204081u) & 0x01010101u) * 0xFFu;
        }
      } else if (t < P) {
        if (cov) p.mask[t] = 255;
        else if (!p.keep_mask) p.mask[t] = 0;
      }
    }
  }
}

__global__ void fill_i32(int* __restrict__ p, size_t n, int v) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}

}  // namespace ugu_b200
