// Raytracer parity mode: ugu::Raytracer::Impl::Render (src/renderer/cpu/raytracer.cc:
// 117-225) — one primary ray per pixel, nearest hit from BvhNanort::Intersect(ray, test_all = false)
// (include/ugu/accel/bvh_nanort.h:109-152), back-face test AFTER the nearest hit, depth from the hit
// point, (u, v) from the intersector — as sm_100a kernels WITHOUT a BVH.
//
// Why no BVH: nanort's traversal (third_party/nanort/nanort.h:2485-2552) only prunes boxes against the
// current nearest t, so what it returns is min over the triangles of the watertight intersector's t
// (TriangleIntersector::Intersect, nanort.h:1050-1148).  Primary rays of a pinhole camera are exactly the
// pixel grid, so the candidate triangles of a ray are the triangles whose screen projection covers its
// pixel: this is the rasterizer's loop structure (triangle -> bbox pixels) with nanort's intersector as
// the per-sample test and a 64-bit atomicMin on (t, ~face) as the nearest-hit reduction.  The intersector
// and the ray set-up (PrepareTraversal, nanort.h:1160-1199) are evaluated with the reference's single
// float operations in its order (--fmad=false), the U/V/W == 0 edge case in double like nanort.
//
// Candidates must be a superset of the hits: the bbox of the projected vertices is widened by a bound on
// the rounding error of the projection and of the intersector (it grows as a vertex approaches the camera
// plane), and the integer points inside it are the candidates; a triangle with a vertex at or behind the camera plane is tested against every pixel.
//
// Equal t at a pixel (duplicate / coincident triangles): nanort keeps the last triangle VISITED, which
// depends on its BVH; here the HIGHEST face index wins (what nanort does when the tied triangles share
// a leaf in index order, and what tests/golden/raytrace_kat_shared_edge_tie_* pin).
#pragma once

#include "raster_kernels.cuh"

namespace ugu_b200 {

constexpr int kRtInLaneMax = 64;   // bbox samples up to which a face is traced by its own lane
constexpr int kRtJob = 512;        // candidate pixels per job of a queued (large-bbox) face

struct RtParams {
  float org[3];          // PinholeCamera::org_ray_w: c2w translation as float (camera.h:431-436)
  float t_min, t_max;    // bvh_nanort.h:119-121: 1e-4, 1e30
  float depth_scale;     // RendererCpuOption::depth_scale (raytracer.cc:171)
  float coord_scale;     // max |world coordinate| of the mesh and the ray origin (error bound of the bbox)
  int inlane_max;        // candidate-bbox size up to which RT-K2 traces a face itself (<= kRtInLaneMax)
  float4* ray_a;         // per pixel: dir.x dir.y dir.z Sz
  float4* ray_b;         // per pixel: Sx Sy (kx | ky << 2 | kz << 4) -
};

__device__ __forceinline__ float rt_sel(float a0, float a1, float a2, int k) {
  return k == 0 ? a0 : (k == 1 ? a1 : a2);
}

// The ray of pixel (x, y) (PinholeCamera::ray_w, camera.h:438-448) and nanort's per-ray constants
// (TriangleIntersector::PrepareTraversal, nanort.h:1160-1199): ra = dir.xyz, Sz; rb = Sx, Sy, k-pack.
__device__ __forceinline__ void rt_make_ray(const ViewParams& vp, const FrameDev& fr, unsigned x, unsigned y,
                                            float4& ra, float4& rb) {
  float d[3];
  ray_world(vp, fr.ray_dx[x], fr.ray_dy[y], d[0], d[1], d[2]);
  int kz = 0;
  float absDir = fabsf(d[0]);
  if (absDir < fabsf(d[1])) {
    kz = 1;
    absDir = fabsf(d[1]);
  }
  if (absDir < fabsf(d[2])) {
    kz = 2;
    absDir = fabsf(d[2]);
  }
  int kx = kz + 1;
  if (kx == 3) kx = 0;
  int ky = kx + 1;
  if (ky == 3) ky = 0;
  const float dz = rt_sel(d[0], d[1], d[2], kz);
  if (dz < 0.0f) {
    const int t = kx;
    kx = ky;
    ky = t;
  }
  const float Sx = rt_sel(d[0], d[1], d[2], kx) / dz;
  const float Sy = rt_sel(d[0], d[1], d[2], ky) / dz;
  const float Sz = 1.0f / dz;
  ra = make_float4(d[0], d[1], d[2], Sz);
  rb = make_float4(Sx, Sy, __int_as_float(kx | (ky << 2) | (kz << 4)), 0.f);
}

// RT-K1: the ray table read by the candidate tests of RT-K2 / RT-K3 (32 bytes per pixel, L2-resident at
// the BASELINE image sizes).
// (The five kernels of a ray-cast view form a programmatic-dependent-launch chain like the rasterizer's: each lets its
//  successor be scheduled as it starts and waits for its predecessor before it touches anything.)
__global__ void __launch_bounds__(256) rt_ray_kernel(const ViewParams vp, const FrameDev fr, const RtParams rp) {
  pdl_launch_dependents();
  pdl_wait();
  const unsigned P = (unsigned)vp.W * (unsigned)vp.H;
  const unsigned p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= P) return;
  const unsigned y = p / (unsigned)vp.W, x = p - y * (unsigned)vp.W;
  float4 ra, rb;
  rt_make_ray(vp, fr, x, y, ra, rb);
  rp.ray_a[p] = ra;
  rp.ray_b[p] = rb;
}

// TriangleIntersector::Intersect (nanort.h:1050-1148) for one (ray, triangle) pair without the running
// nearest-t cut (the caller reduces with atomicMin).  A = p0 - org etc. are passed in (they do not
// depend on the pixel).  Returns false for a miss.
__device__ __forceinline__ bool rt_intersect(const float4& rb, float Sz, const float (&A)[3], const float (&B)[3],
                                             const float (&C)[3], float& tt, float& u, float& v) {
  const int kk = __float_as_int(rb.z);
  const int kx = kk & 3, ky = (kk >> 2) & 3, kz = kk >> 4;
  const float Akz = rt_sel(A[0], A[1], A[2], kz), Bkz = rt_sel(B[0], B[1], B[2], kz),
              Ckz = rt_sel(C[0], C[1], C[2], kz);
  const float Ax = rt_sel(A[0], A[1], A[2], kx) - rb.x * Akz;
  const float Ay = rt_sel(A[0], A[1], A[2], ky) - rb.y * Akz;
  const float Bx = rt_sel(B[0], B[1], B[2], kx) - rb.x * Bkz;
  const float By = rt_sel(B[0], B[1], B[2], ky) - rb.y * Bkz;
  const float Cx = rt_sel(C[0], C[1], C[2], kx) - rb.x * Ckz;
  const float Cy = rt_sel(C[0], C[1], C[2], ky) - rb.y * Ckz;
  float U = Cx * By - Cy * Bx;
  float V = Ax * Cy - Ay * Cx;
  float W = Bx * Ay - By * Ax;
  if (U == 0.0f || V == 0.0f || W == 0.0f) {  // nanort.h:1088-1101: edges in double
    U = (float)((double)Cx * (double)By - (double)Cy * (double)Bx);
    V = (float)((double)Ax * (double)Cy - (double)Ay * (double)Cx);
    W = (float)((double)Bx * (double)Ay - (double)By * (double)Ax);
  }
  if ((U < 0.0f || V < 0.0f || W < 0.0f) && (U > 0.0f || V > 0.0f || W > 0.0f)) return false;
  const float det = U + V + W;
  if (det == 0.0f) return false;
  const float Az = Sz * Akz, Bz = Sz * Bkz, Cz = Sz * Ckz;
  const float D = U * Az + V * Bz + W * Cz;
  const float rcpDet = 1.0f / det;
  tt = D * rcpDet;
  u = V * rcpDet;
  v = W * rcpDet;
  return true;
}

// One candidate (pixel, face): intersect and fold into the nearest-hit buffer.  Key = t bits in the
// high word (t >= t_min > 0, so the unsigned order is the float order), ~face in the low word (equal t:
// the highest face index wins); kVisEmpty (all ones) is above every valid key (t < 1e30).
__device__ __forceinline__ void rt_sample(const RtParams& rp, const FrameDev& fr, unsigned pix, int face,
                                          const float (&A)[3], const float (&B)[3], const float (&C)[3]) {
  const float4 ra = rp.ray_a[pix];
  const float4 rb = rp.ray_b[pix];
  float tt, u, v;
  if (!rt_intersect(rb, ra.w, A, B, C, tt, u, v)) return;
  if (!(tt >= rp.t_min) || !(tt < rp.t_max)) return;  // nanort.h:1129-1135, 2548 (NaN: off contract, dropped)
  const unsigned long long key =
      ((unsigned long long)__float_as_uint(tt) << 32) | (unsigned long long)(0xFFFFFFFFu - (unsigned)face);
  if (key < fr.vis[pix]) atomicMin(fr.vis + pix, key);
}

// Conservative pixel bbox of a face's hits.  Returns 0 = no candidate pixel, 1 = bbox, 2 = whole image.
__device__ __forceinline__ int rt_bounds(const ViewParams& vp, const RtParams& rp, const float4& a,
                                         const float4& b, const float4& c, int& x0, int& y0, int& x1, int& y1) {
  const float zmin = fminf(fminf(a.z, b.z), c.z), zmax = fmaxf(fmaxf(a.z, b.z), c.z);
  if (!(zmax > 0.0f)) return 0;  // wholly at / behind the camera plane: no hit with t >= t_min > 0
  x0 = 0; y0 = 0; x1 = vp.W - 1; y1 = vp.H - 1;
  if (!(zmin > 0.0f)) return 2;
  const float xmin = fminf(fminf(a.x, b.x), c.x), xmax = fmaxf(fmaxf(a.x, b.x), c.x);
  const float ymin = fminf(fminf(a.y, b.y), c.y), ymax = fmaxf(fmaxf(a.y, b.y), c.y);
  // rounding error of a projected coordinate X (pixels from the principal point): the projection's own
  // few ulps plus the cancellation error of the camera-space z (~1e-6 * coord_scale) relative to z
  const float ext = fmaxf(fmaxf(fabsf(xmin - vp.cx), fabsf(xmax - vp.cx)), fmaxf(fabsf(ymin - vp.cy), fabsf(ymax - vp.cy)));
  // (pixel (x, y)'s ray passes exactly through image point (x, y): only rounding separates "the ray hits"
  //  from "the integer point lies in the projected triangle", so the margin is an error bound, not a
  //  pixel footprint: 0.02 px floor + 4x the estimated relative error)
  const float margin = 0.02f + ext * (1.6e-5f + 4e-6f * rp.coord_scale / zmin);
  if (!(margin < 64.0f)) return 2;
  if (xmin - margin > vp.wm1 || xmax + margin < 0.0f || ymin - margin > vp.hm1 || ymax + margin < 0.0f) return 0;
  x0 = max(0, sat_int(ceilf(xmin - margin)));
  x1 = min(vp.W - 1, sat_int(floorf(xmax + margin)));
  y0 = max(0, sat_int(ceilf(ymin - margin)));
  y1 = min(vp.H - 1, sat_int(floorf(ymax + margin)));
  return (x1 >= x0 && y1 >= y0) ? 1 : 0;
}

__device__ __forceinline__ void rt_load_face(const MeshDev& m, const RtParams& rp, int f, int& i0, int& i1, int& i2,
                                             float (&A)[3], float (&B)[3], float (&C)[3]) {
  i0 = __ldg(m.idx0 + f);
  i1 = __ldg(m.idx1 + f);
  i2 = __ldg(m.idx2 + f);
  const float4 p0 = __ldg(m.pos4 + i0), p1 = __ldg(m.pos4 + i1), p2 = __ldg(m.pos4 + i2);
  A[0] = p0.x - rp.org[0]; A[1] = p0.y - rp.org[1]; A[2] = p0.z - rp.org[2];
  B[0] = p1.x - rp.org[0]; B[1] = p1.y - rp.org[1]; B[2] = p1.z - rp.org[2];
  C[0] = p2.x - rp.org[0]; C[1] = p2.y - rp.org[1]; C[2] = p2.z - rp.org[2];
}

// RT-K2: a warp per 32 faces.  Phase A (lane = face): gather, candidate bbox; faces whose bbox holds more
// than kRtInLaneMax pixels are appended to the job queue (one atomicAdd per warp) for RT-K3, the others are
// compacted into a per-warp shared-memory table.  Phase B (lane = candidate pixel): the candidates of
// all tabled faces form one index space [0, T) that the warp walks 32 at a time, so lanes stay busy
// whatever the mix of bbox sizes (the owner of a candidate is found from a REDUX.OR bitmask of the
// faces' start offsets, as in raster_kernel).
struct __align__(16) RtWarpFaces {
  float4 a[32];   // A.xyz, B.x        (A = p0 - org, ...)
  float4 b[32];   // B.yz, C.xy
  float4 c[32];   // C.z, (x0 | y0 << 16), face, -
  int2 meta[32];  // first candidate of the face in the warp's index space; nx | ceil(65536 / nx) << 9
};

__global__ void __launch_bounds__(256) rt_raster_kernel(const ViewParams vp, const MeshDev m, const FrameDev fr,
                                                        const RtParams rp) {
  __shared__ RtWarpFaces sh[8];
  RtWarpFaces& S = sh[threadIdx.x >> 5];
  pdl_launch_dependents();
  pdl_wait();
  // faces are walked in RASTER ORDER (raster_api.cu BuildFaceOrder: a warp's 32 faces are neighbours in space,
  // so are their positions, projected vertices and nearest-hit pixels); `f` is the face id the outputs carry
  unsigned blk = blockIdx.x;
  bool deferred = false;
  if (fr.order) {  // occlusion ordering (raster_kernels.cuh): camera-facing blocks of 256 faces first
    blk = (unsigned)fr.order[blockIdx.x];
    deferred = (unsigned long long)blockIdx.x >= fr.counters[3];
  }
  const int rs = (int)(blk * blockDim.x + threadIdx.x);
  int f = rs;
  const unsigned lane = threadIdx.x & 31u;
  const unsigned lt_mask = (1u << lane) - 1u;
  int x0 = 0, y0 = 0, x1 = -1, y1 = -1, kind = 0;
  float A[3], B[3], C[3];
  if (rs < m.F) {
    const int i0 = __ldg(m.r_idx0 + rs), i1 = __ldg(m.r_idx1 + rs), i2 = __ldg(m.r_idx2 + rs);
    if (m.r_face) f = __ldg(m.r_face + rs);
    const float4 p0 = __ldg(m.pos4 + i0), p1 = __ldg(m.pos4 + i1), p2 = __ldg(m.pos4 + i2);
    A[0] = p0.x - rp.org[0]; A[1] = p0.y - rp.org[1]; A[2] = p0.z - rp.org[2];
    B[0] = p1.x - rp.org[0]; B[1] = p1.y - rp.org[1]; B[2] = p1.z - rp.org[2];
    C[0] = p2.x - rp.org[0]; C[1] = p2.y - rp.org[1]; C[2] = p2.z - rp.org[2];
    const float4 qa = fr.proj[i0], qb = fr.proj[i1], qc = fr.proj[i2];
    kind = rt_bounds(vp, rp, qa, qb, qc, x0, y0, x1, y1);
    if (deferred && kind == 1 && x1 - x0 < 4 && y1 - y0 < 4) {
      // A deferred face whose candidate pixels all hold a nearer hit already cannot win any of them.  nanort's
      // t = (U Az + V Bz + W Cz) / (U + V + W) with U, V, W of one sign is a convex combination (up to 8 ulp) of
      // the ray parameters Az.. at which the ray crosses the kz-planes of the vertices, whatever the shape of
      // the triangle; and for a ray within angle alpha of the direction to a vertex at distance d that parameter
      // is >= d (cos alpha - sqrt(3) sin alpha) (kz = dominant axis of the unit direction) >= 0.9826 d for
      // alpha <= 0.01.  alpha <= (largest pixel distance between a candidate pixel and a projected vertex) / f.
      const float xmin = fminf(fminf(qa.x, qb.x), qc.x), xmax = fmaxf(fmaxf(qa.x, qb.x), qc.x);
      const float ymin = fminf(fminf(qa.y, qb.y), qc.y), ymax = fmaxf(fmaxf(qa.y, qb.y), qc.y);
      const float dx = fmaxf((float)x1 - xmin, xmax - (float)x0), dy = fmaxf((float)y1 - ymin, ymax - (float)y0);
      const float alpha = sqrtf(dx * dx + dy * dy) / fminf(vp.fx, vp.fy);
      const float d2 = fminf(fminf(A[0] * A[0] + A[1] * A[1] + A[2] * A[2], B[0] * B[0] + B[1] * B[1] + B[2] * B[2]),
                             C[0] * C[0] + C[1] * C[1] + C[2] * C[2]);
      if (alpha <= 0.0095f && d2 > 1e-20f && d2 < 1e30f) {
        const unsigned tbits = __float_as_uint(0.975f * sqrtf(d2));
        if (bbox_occluded(fr, vp.W, x0, y0, x1 - x0 + 1, y1 - y0 + 1, tbits)) kind = 0;
      }
    }
  }
  const int nx = x1 - x0 + 1, ny = y1 - y0 + 1;
  const bool large = kind != 0 && (long long)nx * ny > (long long)rp.inlane_max;
  const unsigned lb = __ballot_sync(0xffffffffu, large);
  if (lb) {  // queue: packed (entries, jobs) counter as in raster_kernel — one atomicAdd per warp hands every
             // queued face its entry slot and its first job in the global job space
    const unsigned long long njobs = large ? ((unsigned long long)nx * (unsigned long long)ny + kRtJob - 1) / kRtJob : 0ull;
    const unsigned long long mine = large ? ((1ull << kQueueJobBits) | njobs) : 0ull;
    unsigned long long incl = mine;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const unsigned long long o = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= (unsigned)d) incl += o;
    }
    const unsigned long long total = __shfl_sync(0xffffffffu, incl, 31);
    unsigned long long qbase = 0ull;
    if (lane == 31u) qbase = atomicAdd(fr.counters, total);
    qbase = __shfl_sync(0xffffffffu, qbase, 31);
    if (large) {
      const unsigned long long excl = qbase + incl - mine;
      const unsigned qslot = (unsigned)(excl >> kQueueJobBits);
      fr.q_face[qslot] = f;
      fr.q_job[qslot] = excl & kQueueJobMask;
    }
  }
  const int cnt = (kind != 0 && !large) ? nx * ny : 0;
  const unsigned rec = __ballot_sync(0xffffffffu, cnt > 0);
  if (rec == 0u) return;
  int incl = cnt;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const int o = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= (unsigned)d) incl += o;
  }
  const int T = __shfl_sync(0xffffffffu, incl, 31);
  if (cnt > 0) {
    const int slot = __popc(rec & lt_mask);
    S.a[slot] = make_float4(A[0], A[1], A[2], B[0]);
    S.b[slot] = make_float4(B[1], B[2], C[0], C[1]);
    S.c[slot] = make_float4(C[2], __int_as_float(x0 | (y0 << 16)), __int_as_float(f), 0.f);
    const int recip = (int)(65536.0f / small_int_to_float(nx));  // ceil(65536 / nx), nx <= 64
    S.meta[slot] = make_int2(incl - cnt, nx | ((recip + ((recip * nx) < 65536 ? 1 : 0)) << 9));
  }
  __syncwarp();
  const int nslots = __popc(rec);
  const int st = ((int)lane < nslots) ? S.meta[lane].x : 0x7fffffff;
  const unsigned le_mask = lt_mask | (1u << lane);
  int started = 0;
  for (int sb = 0; sb < T; sb += 32) {
    const unsigned rr = (unsigned)(st - sb);
    const unsigned starts = __reduce_or_sync(0xffffffffu, rr < 32u ? (1u << rr) : 0u);
    const int s = sb + (int)lane;
    if (s < T) {
      const int slot = started + __popc(starts & le_mask) - 1;
      const int2 mt = S.meta[slot];
      const float4 ta = S.a[slot], tb = S.b[slot], tc = S.c[slot];
      const int k = s - mt.x;
      const int w = mt.y & 511;
      const int dy = (k * (mt.y >> 9)) >> 16;  // k / w, exact for k, w <= 256
      const int dx = k - dy * w;
      const unsigned xy = __float_as_uint(tc.y);
      const unsigned x = (xy & 0xFFFFu) + (unsigned)dx, y = (xy >> 16) + (unsigned)dy;
      const float FA[3] = {ta.x, ta.y, ta.z}, FB[3] = {ta.w, tb.x, tb.y}, FC[3] = {tb.z, tb.w, tc.x};
      rt_sample(rp, fr, y * (unsigned)vp.W + x, __float_as_int(tc.z), FA, FB, FC);
    }
    started += __popc(starts);
  }
}

// RT-K3: queued faces.  Their candidate pixels are cut into jobs of kRtJob pixels that form one global job
// space [0, T) (the queue orders entries by first job), so the work is balanced whatever the mix of bbox
// sizes: every warp of a resident grid takes a contiguous range of jobs, finds the entry that owns its first
// job with one binary search and walks forward from there; lanes stride over a job's pixels.
__global__ void __launch_bounds__(256) rt_large_kernel(const ViewParams vp, const MeshDev m, const FrameDev fr,
                                                       const RtParams rp) {
  pdl_launch_dependents();
  pdl_wait();
  const unsigned long long packed = fr.counters[0];
  const unsigned long long T = packed & kQueueJobMask;
  const unsigned n = (unsigned)(packed >> kQueueJobBits);
  if (T == 0ull) return;
  const unsigned lane = threadIdx.x & 31u;
  const unsigned long long nwarps = (unsigned long long)gridDim.x * (blockDim.x >> 5);
  const unsigned long long gw = (unsigned long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const unsigned long long per = (T + nwarps - 1) / nwarps;
  unsigned long long j = gw * per;
  const unsigned long long jend = (j + per < T) ? j + per : T;
  if (j >= jend) return;
  // last entry e with q_job[e] <= j: warp-cooperative 32-ary search (q_job is ascending, q_job[0] == 0) —
  // log32(n) dependent loads instead of log2(n)
  unsigned lo = 0u, hi = n;
  while (hi - lo > 1u) {
    const unsigned step = (hi - lo + 31u) / 32u;
    const unsigned idx = lo + (lane + 1u) * step;
    const bool le = idx < hi && fr.q_job[idx] <= j;
    const unsigned c = (unsigned)__popc(__ballot_sync(0xffffffffu, le));  // the pivots <= j are a prefix
    lo += c * step;
    hi = min(hi, lo + step);
  }
  unsigned e = lo;
  while (j < jend) {
    const unsigned long long first = fr.q_job[e];
    const unsigned long long next = (e + 1u < n) ? fr.q_job[e + 1u] : T;  // jobs of entry e: [first, next)
    const int f = fr.q_face[e];
    int i0, i1, i2, x0, y0, x1, y1;
    float A[3], B[3], C[3];
    rt_load_face(m, rp, f, i0, i1, i2, A, B, C);
    const int kind = rt_bounds(vp, rp, fr.proj[i0], fr.proj[i1], fr.proj[i2], x0, y0, x1, y1);
    const unsigned nx = (unsigned)(x1 - x0 + 1);
    const unsigned long long ns = kind ? (unsigned long long)nx * (unsigned long long)(y1 - y0 + 1) : 0ull;
    const unsigned long long jstop = next < jend ? next : jend;
    for (; j < jstop; j++) {
      // ns <= W*H < 2^31 (validated by the host): 32-bit pixel arithmetic
      const unsigned k0 = (unsigned)(j - first) * (unsigned)kRtJob;
      const unsigned k1 = min(k0 + (unsigned)kRtJob, (unsigned)ns);
      for (unsigned k = k0 + lane; k < k1; k += 32u) {
        const unsigned r = k / nx, dx = k - r * nx;
        rt_sample(rp, fr, ((unsigned)y0 + r) * (unsigned)vp.W + (unsigned)x0 + dx, f, A, B, C);
      }
    }
    e++;
  }
}

// PixelShader::Process (include/ugu/renderer/cpu/pixel_shader.h:115-140,222-422) as raytracer.cc:205-211
// calls it: colourizer on (u, v), then the shader with light_dir == ray_w.  Same expressions as
// resolve_pixel (raster_kernels.cuh), selected at run time.  Returns r | g << 8 | b << 16.
__device__ __forceinline__ unsigned rt_shade(const MeshDev& m, const ShadeParams& sp, int fid, const int4& vr,
                                             float u, float v, float r0, float r1, float r2, float nwx,
                                             float nwy, float nwz) {
  unsigned cr = 255u, cg = 255u, cb = 255u;  // DiffuseDefaultColorizer
  const float k0 = 1.0f - u - v;
  if (sp.diffuse_color == 2) {  // DiffuseVertexColorColorizer (pixel_shader.h:234-256)
    const float4 c0 = __ldg(m.col4 + vr.x), c1 = __ldg(m.col4 + vr.y), c2 = __ldg(m.col4 + vr.z);
    cr = f2u8((k0 * c0.x + u * c1.x) + v * c2.x);
    cg = f2u8((k0 * c0.y + u * c1.y) + v * c2.y);
    cb = f2u8((k0 * c0.z + u * c1.z) + v * c2.z);
  } else if (sp.diffuse_color == 1) {  // texture colourizers (pixel_shader.h:260-359)
    const int4 ui = __ldg(m.uvidx + fid);
    const float2 t0 = __ldg(m.uv2 + ui.x), t1 = __ldg(m.uv2 + ui.y), t2 = __ldg(m.uv2 + ui.z);
    const int mat = vr.w;
    const uchar4* tex = m.tex[mat];
    const int2 ts = m.tex_size[mat];
    const float iu = (k0 * t0.x + u * t1.x) + v * t2.x;
    const float iv = (k0 * t0.y + u * t1.y) + v * t2.y;
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
  if (sp.diffuse_shading != 0) {
    float coeff;
    if (sp.diffuse_shading == 1) {  // DiffuseLambertianShader (pixel_shader.h:367-386)
      coeff = -dot3(r0, r1, r2, nwx, nwy, nwz);
      if (coeff < 0.0f) coeff = 0.0f;
    } else {  // DiffuseOrenNayarShader (pixel_shader.h:390-422)
      const float dot_light = -dot3(r0, r1, r2, nwx, nwy, nwz);
      const float theta_i = acosf(dot_light);
      const float blx = (-nwx) * dot_light - r0;
      const float bly = (-nwy) * dot_light - r1;
      const float blz = (-nwz) * dot_light - r2;
      const float phi = fmaxf(0.0f, dot3(blx, bly, blz, blx, bly, blz));
      coeff = fmaxf(0.0f, dot_light) * (sp.on_A + (sp.on_B * phi * sinf(theta_i) * tanf(theta_i)));
    }
    cr = f2u8(coeff * (float)cr);
    cg = f2u8(coeff * (float)cg);
    cb = f2u8(coeff * (float)cb);
  }
  return cr | (cg << 8) | (cb << 16);
}

// RT-K4: raytracer.cc:152-211 for the nearest hit of every pixel, fused with the clears of
// util_private.cc:72-87 and the reset of the nearest-hit buffer / queue counter for the next call.
// (t, u, v) are recomputed from the winner with the same intersector (bit-identical).
__global__ void __launch_bounds__(256) rt_resolve_kernel(const ViewParams vp, const MeshDev m, const FrameDev fr,
                                                         const RtParams rp, const ShadeParams sp) {
  __shared__ __align__(16) float s_n[8][96];
  __shared__ __align__(16) unsigned char s_c[8][96];
  pdl_launch_dependents();
  pdl_wait();
  const unsigned P = (unsigned)vp.W * (unsigned)vp.H;
  const unsigned p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p == 0u) {
    fr.counters[4] = fr.counters[0];  // host snapshot: packed (queued faces, jobs)
    fr.counters[5] = 0ull;
    fr.counters[0] = 0ull;
    fr.counters[3] = 0ull;  // block lists of the occlusion ordering (vertex kernel)
    fr.counters[7] = 0ull;
  }
  const unsigned long long key = p < P ? fr.vis[p] : kVisEmpty;
  unsigned color = 0u;
  float depth = 0.0f, nx = 0.0f, ny = 0.0f, nz = 0.0f;
  bool mask = false;
  int face_id = -1;
  if (key != kVisEmpty) {
    fr.vis[p] = kVisEmpty;
    const int fid = (int)(0xFFFFFFFFu - (unsigned)(key & 0xFFFFFFFFull));
    const int4 vr = __ldg(m.frec + 2 * (size_t)fid);  // issued first: the head of the gather chain
    // the winner's ray is recomputed (same function as the table: identical bits) instead of read back:
    // hits are a minority of the pixels and the table is 32 bytes per pixel
    float4 ra, rb;
    {
      const unsigned y = (vp.W == 1) ? p : (__umulhi(p, vp.w_magic) >> vp.w_shift);
      rt_make_ray(vp, fr, p - y * (unsigned)vp.W, y, ra, rb);
    }
    bool keep = true;
    if (sp.backface_culling)  // raytracer.cc:155-160: face normal . ray > 0 -> the pixel stays empty
      keep = !(dot3(__ldg(m.fn_x + fid), __ldg(m.fn_y + fid), __ldg(m.fn_z + fid), ra.x, ra.y, ra.z) > 0.0f);
    if (keep) {
      float A[3], B[3], C[3], t, u, v;
      const float4 p0 = __ldg(m.pos4 + vr.x), p1 = __ldg(m.pos4 + vr.y), p2 = __ldg(m.pos4 + vr.z);
      A[0] = p0.x - rp.org[0]; A[1] = p0.y - rp.org[1]; A[2] = p0.z - rp.org[2];
      B[0] = p1.x - rp.org[0]; B[1] = p1.y - rp.org[1]; B[2] = p1.z - rp.org[2];
      C[0] = p2.x - rp.org[0]; C[1] = p2.y - rp.org[1]; C[2] = p2.z - rp.org[2];
      (void)rt_intersect(rb, ra.w, A, B, C, t, u, v);
      face_id = fid;
      mask = true;
      {  // raytracer.cc:167-172
        const float hx = rp.org[0] + ra.x * t, hy = rp.org[1] + ra.y * t, hz = rp.org[2] + ra.z * t;
        depth = (dot3(vp.w2c_R[6], vp.w2c_R[7], vp.w2c_R[8], hx, hy, hz) + vp.w2c_t[2]) * rp.depth_scale;
      }
      float nwx = 0.f, nwy = 0.f, nwz = 0.f;
      const bool need_n = sp.normal != nullptr || (sp.color != nullptr && sp.diffuse_shading != 0);
      if (need_n) {
        if (sp.shading_normal == 0) {  // raytracer.cc:176-177
          nwx = __ldg(m.fn_x + fid); nwy = __ldg(m.fn_y + fid); nwz = __ldg(m.fn_z + fid);
        } else {  // raytracer.cc:178-186
          const int4 ni = __ldg(m.frec + 2 * (size_t)fid + 1);
          const float4 n0 = __ldg(m.nrm4 + ni.x), n1 = __ldg(m.nrm4 + ni.y), n2 = __ldg(m.nrm4 + ni.z);
          const float k0 = 1.0f - u - v;
          nwx = (k0 * n0.x + u * n1.x) + v * n2.x;
          nwy = (k0 * n0.y + u * n1.y) + v * n2.y;
          nwz = (k0 * n0.z + u * n1.z) + v * n2.z;
        }
        nx = dot3(vp.w2c_R[0], vp.w2c_R[1], vp.w2c_R[2], nwx, nwy, nwz);
        ny = dot3(vp.w2c_R[3], vp.w2c_R[4], vp.w2c_R[5], nwx, nwy, nwz);
        nz = dot3(vp.w2c_R[6], vp.w2c_R[7], vp.w2c_R[8], nwx, nwy, nwz);
      }
      if (sp.color) color = rt_shade(m, sp, fid, vr, u, v, ra.x, ra.y, ra.z, nwx, nwy, nwz);
    }
  }
  // stores: a warp owns 32 consecutive pixels; colour / normal are staged through shared memory so that
  // every global store is a full 16-byte vector (as resolve_kernel does)
  const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
  const unsigned wbase = p - lane;
  if (wbase + 32u <= P) {
    if (sp.depth) sp.depth[p] = depth;
    if (sp.face_id) sp.face_id[p] = face_id;
    if (sp.mask) {
      const unsigned bal = __ballot_sync(0xffffffffu, mask);
      if (lane < 8u) {
        const unsigned nb = (bal >> (lane * 4u)) & 15u;  // nibble -> 4 bytes of 0x00 / 0xFF
        *reinterpret_cast<unsigned*>(sp.mask + wbase + lane * 4u) = ((nb * 0x00204081u) & 0x01010101u) * 0xFFu;
      }
    }
    if (sp.normal) {
      s_n[warp][lane * 3u + 0u] = nx;
      s_n[warp][lane * 3u + 1u] = ny;
      s_n[warp][lane * 3u + 2u] = nz;
    }
    if (sp.color) {
      s_c[warp][lane * 3u + 0u] = (unsigned char)(color & 0xFFu);
      s_c[warp][lane * 3u + 1u] = (unsigned char)((color >> 8) & 0xFFu);
      s_c[warp][lane * 3u + 2u] = (unsigned char)((color >> 16) & 0xFFu);
    }
    __syncwarp();
    if (sp.normal && lane < 24u)
      reinterpret_cast<float4*>(sp.normal + (size_t)wbase * 3u)[lane] = reinterpret_cast<const float4*>(s_n[warp])[lane];
    if (sp.color && lane < 6u)
      reinterpret_cast<uint4*>(sp.color + (size_t)wbase * 3u)[lane] = reinterpret_cast<const uint4*>(s_c[warp])[lane];
  } else if (p < P) {  // ragged last chunk of the image
    if (sp.color) {
      sp.color[(size_t)p * 3 + 0] = (uint8_t)(color & 0xFFu);
      sp.color[(size_t)p * 3 + 1] = (uint8_t)((color >> 8) & 0xFFu);
      sp.color[(size_t)p * 3 + 2] = (uint8_t)((color >> 16) & 0xFFu);
    }
    if (sp.depth) sp.depth[p] = depth;
    if (sp.normal) {
      sp.normal[(size_t)p * 3 + 0] = nx;
      sp.normal[(size_t)p * 3 + 1] = ny;
      sp.normal[(size_t)p * 3 + 2] = nz;
    }
    if (sp.mask) sp.mask[p] = mask ? 255 : 0;
    if (sp.face_id) sp.face_id[p] = face_id;
  }
}

}  // namespace ugu_b200
