// sm_100a kernels of the face-visibility consumer: per-face statistics over a
// batch of rendered face_id maps — the rasterizer-side counterpart of what the reference's
// texturing pipeline collects per keyframe with BVH rays (ugu::FaceInfo / FaceInfoPerKeyframe,
// include/ugu/texturing/visibility_tester.h:135-181; FaceInfo::CalcStat's max_area_id,
// src/texturing/visibility_tester.cc:185-218).  For every face f:
//   pixel_count[f]  pixels owned by f, summed over the views
//   view_count[f]   views in which f owns at least one pixel
//   best[f]         max over views of (pixels of f in the view) << 32 | (0xFFFFFFFF - view id):
//                   the view that sees the most of f, ties to the lowest view id
// All three combine across ranks with SUM / SUM / MAX, so a multi-GPU job all-reduces F-sized
// vectors instead of gathering every face_id map.
//
// Two passes per batch of <= kVisSlots views (grid.y = view slot): pass 1 histograms the map into a
// per-slot scratch counter; pass 2 walks the map again and the first thread to swap a face's counter
// back to zero folds it into the accumulators — which also leaves the scratch clean.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace ugu_b200 {

constexpr int kVisSlots = 8;

struct VisParams {
  const int* face_id;            // views * P
  unsigned long long P;          // pixels per view
  int F;
  int first_view_id;
  unsigned* tmp;                 // kVisSlots * F scratch, all zero between batches
  unsigned* pixel_count;         // F
  unsigned* view_count;          // F
  unsigned long long* best;      // F
};

template <bool VEC>
__device__ __forceinline__ int vis_load4(const int* map, unsigned long long P, unsigned long long i, int (&f)[4]) {
  if (VEC) {
    const int4 w = __ldg(reinterpret_cast<const int4*>(map + i));
    f[0] = w.x; f[1] = w.y; f[2] = w.z; f[3] = w.w;
    return 4;
  }
  int n = 0;
  for (; n < 4 && i + n < P; n++) f[n] = __ldg(map + i + n);
  return n;
}

template <bool VEC>
__global__ void __launch_bounds__(256) vis_count_kernel(const VisParams p, int view0) {
  const unsigned long long i = ((unsigned long long)blockIdx.x * blockDim.x + threadIdx.x) * 4ull;
  if (i >= p.P) return;
  const int* map = p.face_id + (unsigned long long)(view0 + blockIdx.y) * p.P;
  unsigned* tmp = p.tmp + (size_t)blockIdx.y * p.F;
  int f[4];
  const int n = vis_load4<VEC>(map, p.P, i, f);
  int run = -1;
  unsigned cnt = 0;
  for (int k = 0; k < n; k++) {  // run-length merge: neighbouring pixels mostly share their face
    const int id = ((unsigned)f[k] < (unsigned)p.F) ? f[k] : -1;
    if (id != run) {
      if (run >= 0) atomicAdd(tmp + run, cnt);
      run = id;
      cnt = 0;
    }
    cnt++;
  }
  if (run >= 0) atomicAdd(tmp + run, cnt);
}

template <bool VEC>
__global__ void __launch_bounds__(256) vis_fold_kernel(const VisParams p, int view0) {
  const unsigned long long i = ((unsigned long long)blockIdx.x * blockDim.x + threadIdx.x) * 4ull;
  if (i >= p.P) return;
  const int* map = p.face_id + (unsigned long long)(view0 + blockIdx.y) * p.P;
  unsigned* tmp = p.tmp + (size_t)blockIdx.y * p.F;
  const unsigned view_id = (unsigned)(p.first_view_id + view0 + (int)blockIdx.y);
  int f[4];
  const int n = vis_load4<VEC>(map, p.P, i, f);
  int prev = -1;
  for (int k = 0; k < n; k++) {
    const int id = ((unsigned)f[k] < (unsigned)p.F) ? f[k] : -1;
    if (id < 0 || id == prev) continue;
    prev = id;
    if (__ldcg(tmp + id) == 0u) continue;          // already folded by another thread
    const unsigned c = atomicExch(tmp + id, 0u);   // exactly one thread per (face, view) gets c != 0
    if (c == 0u) continue;
    atomicAdd(p.pixel_count + id, c);
    atomicAdd(p.view_count + id, 1u);
    atomicMax(p.best + id, ((unsigned long long)c << 32) | (unsigned long long)(0xFFFFFFFFu - view_id));
  }
}

}  // namespace ugu_b200
