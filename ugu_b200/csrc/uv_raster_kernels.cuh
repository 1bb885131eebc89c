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
  void* texture;          // W*H*3 u8 or f32
  uint8_t* mask;          // W*H or null
};

// static_cast<int>(float) as the x86-64 reference compiles it (cvttss2si): truncation; out of range
// and NaN give INT_MIN.
__device__ __forceinline__ int cvtt_x86(float f) {
  if (!(f > -2147483904.0f && f < 2147483648.0f)) return (int)0x80000000;
  return __float2int_rz(f);
}

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
// UV-K1: per-face bbox -> tile jobs.  Lane = face; the warp's entries are appended to the queue with
// one atomicAdd on the packed (entries, jobs) counter, so q_job is monotone over queue slots.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) uv_setup_kernel(const UvRasterParams p) {
  const int f = blockIdx.x * blockDim.x + threadIdx.x;
  const unsigned lane = threadIdx.x & 31u;
  unsigned long long mine = 0ull;
  if (f < p.F) {
    UvTri t;
    uv_tri_setup(p, f, t);
    int x0, y0, x1, y1;
    uv_tri_bounds(p, t, x0, y0, x1, y1);
    if (x1 >= x0 && y1 >= y0) {
      const unsigned long long njobs = (unsigned long long)((x1 - x0 + kJobW) / kJobW) *
                                       (unsigned long long)((y1 - y0 + kJobH) / kJobH);
      mine = (1ull << kQueueJobBits) | njobs;
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
  const unsigned long long packed = p.counter[0];
  const unsigned long long T = packed & kQueueJobMask;
  const unsigned n = (unsigned)(packed >> kQueueJobBits);
  if (T == 0ull) return;
  const unsigned lane = threadIdx.x & 31u;
  const unsigned long long nwarps = (unsigned long long)gridDim.x * (blockDim.x >> 5);
  const unsigned long long gw = (unsigned long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const unsigned long long rounds =
      (T + nwarps * (unsigned long long)kJobsPerChunk - 1) / (nwarps * (unsigned long long)kJobsPerChunk);
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
      const int f = p.q_face[e];
      const unsigned long long js = p.q_job[e];
      const unsigned long long jn = (e + 1u < n) ? p.q_job[e + 1u] : T;
      const unsigned long long jend = (jn < je) ? jn : je;
      UvTri t;
      uv_tri_setup(p, f, t);
      int x0, y0, x1, y1;
      uv_tri_bounds(p, t, x0, y0, x1, y1);
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
      ++e;
    }
  }
}

// ---------------------------------------------------------------------------------------------
// UV-K3: resolve.  4 consecutive texels per thread: read + re-clear the winners (one 16-byte access
// each way), recompute the winner's weights with the expressions of UV-K2, interpolate
// (w0*c0 + w1*c1) + w2*c2 (raster_util.h:138-139), convert (saturate_cast, :106-115: std::round for
// integer texels, then the truncating cast) and store 12 / 48 packed bytes + 4 mask bytes.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ bool uv_shade(const UvRasterParams& p, int f, int x, int y, float& c0,
                                         float& c1, float& c2) {
  c0 = c1 = c2 = 0.f;
  if (f < 0) return false;
  UvTri t;
  uv_tri_setup(p, f, t);
  int a0 = f, a1 = f, a2 = f;
  if (!p.per_face) {
    a0 = __ldg(p.attr_idx + 3 * (size_t)f + 0);
    a1 = __ldg(p.attr_idx + 3 * (size_t)f + 1);
    a2 = __ldg(p.attr_idx + 3 * (size_t)f + 2);
  }
  const float* q0 = p.attrs + 3 * (size_t)a0;
  const float* q1 = p.attrs + 3 * (size_t)a1;
  const float* q2 = p.attrs + 3 * (size_t)a2;
  float w0, w1, w2;
  uv_weights(t, (float)x, (float)y, w0, w1, w2);
  c0 = (w0 * __ldg(q0 + 0) + w1 * __ldg(q1 + 0)) + w2 * __ldg(q2 + 0);
  c1 = (w0 * __ldg(q0 + 1) + w1 * __ldg(q1 + 1)) + w2 * __ldg(q2 + 1);
  c2 = (w0 * __ldg(q0 + 2) + w1 * __ldg(q1 + 2)) + w2 * __ldg(q2 + 2);
  return true;
}

__device__ __forceinline__ unsigned uv_to_u8(float c) {  // (uint8_t)std::round(c) through cvttss2si
  return (unsigned)cvtt_x86(roundf(c)) & 0xFFu;
}

__global__ void __launch_bounds__(256) uv_resolve_kernel(const UvRasterParams p) {
  const size_t P = (size_t)p.W * p.H;
  const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t == 0) p.counter[0] = 0ull;  // queue reset for the next call
  const size_t base = t * 4;
  if (base >= P) return;
  int f[4] = {-1, -1, -1, -1};
  const bool full = base + 3 < P;
  if (full) {
    const int4 w = *reinterpret_cast<const int4*>(p.win + base);
    f[0] = w.x; f[1] = w.y; f[2] = w.z; f[3] = w.w;
    *reinterpret_cast<int4*>(p.win + base) = make_int4(-1, -1, -1, -1);
  } else {
    for (int k = 0; k < 4 && base + k < P; k++) {
      f[k] = p.win[base + k];
      p.win[base + k] = -1;
    }
  }
  float c[4][3];
  bool cov[4];
  int y = (int)(base / (size_t)p.W), x = (int)(base - (size_t)y * p.W);
#pragma unroll
  for (int k = 0; k < 4; k++) {
    cov[k] = uv_shade(p, f[k], x, y, c[k][0], c[k][1], c[k][2]);
    if (++x == p.W) {
      x = 0;
      ++y;
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

__global__ void fill_i32(int* __restrict__ p, size_t n, int v) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}

}  // namespace ugu_b200
