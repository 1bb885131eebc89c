// ugu_b200_face_visibility_static: the face-visibility fold of visibility_kernels.cuh for callers
// that cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
// No counterpart in the reference.
//
// The per-slot histogram scratch that ugu_b200_face_visibility_device keeps in a grown device
// buffer lives here in a file-scope __device__ array of kStaticVisWords words (16 MiB), resident on
// every GPU that loads the library until the process exits.  The entry point calls no cudaMalloc,
// cudaFree or allocating library routine: it resolves the array's device address once per device
// (cudaGetSymbolAddress, with that device current), clears the part it uses on the caller's stream
// and launches the two kernels.  All calls on a device share the array: calls that overlap on the
// GPU (different streams, or concurrent host threads on one device) race.
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <mutex>

#include "ugu_b200_raster.h"
#include "visibility_kernels.cuh"

namespace {

using namespace ugu_b200;

constexpr size_t kStaticVisWords = (size_t)UGU_B200_STATIC_VIS_MAX_FACES;  // 4 Mi words = 16 MiB
constexpr int kMaxDevices = 64;

// In host code this name is a host-side stand-in, NOT a device address: only the pointer returned
// by cudaGetSymbolAddress may be handed to a kernel.
static __device__ __align__(256) unsigned g_static_vis_scratch[kStaticVisWords];

std::mutex g_mu;
unsigned* g_scratch_ptr[kMaxDevices] = {};  // per device, resolved with that device current

int Err(char* err, size_t err_len, int code, const char* fmt, ...) {
  if (err && err_len) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(err, err_len, fmt, ap);
    va_end(ap);
  }
  return code;
}

}  // namespace

extern "C" int ugu_b200_face_visibility_static(int32_t device, const int32_t* face_id_dev, int32_t num_views,
                                               int64_t pixels_per_view, int32_t first_view_id,
                                               int32_t num_faces, uint32_t* pixel_count_dev,
                                               uint32_t* view_count_dev, uint64_t* best_dev, void* stream,
                                               char* err, size_t err_len) {
  if (err && err_len) err[0] = 0;
  if (!face_id_dev || !pixel_count_dev || !view_count_dev || !best_dev)
    return Err(err, err_len, UGU_B200_E_INVALID, "NULL argument");
  if (num_views < 0 || pixels_per_view <= 0 || num_faces <= 0 || first_view_id < 0 ||
      pixels_per_view > (1ll << 40) || (long long)first_view_id + num_views > 0x7fffffffll)
    return Err(err, err_len, UGU_B200_E_INVALID, "bad sizes");
  if ((size_t)num_faces > kStaticVisWords)
    return Err(err, err_len, UGU_B200_E_INVALID,
               "num_faces %d exceeds the static scratch limit UGU_B200_STATIC_VIS_MAX_FACES = %zu; use "
               "ugu_b200_face_visibility_device",
               num_faces, kStaticVisWords);
  if (device < 0 || device >= kMaxDevices) return Err(err, err_len, UGU_B200_E_INVALID, "bad device %d", device);
  if (num_views == 0) return UGU_B200_OK;
  cudaError_t e = cudaSetDevice(device);
  if (e != cudaSuccess)
    return Err(err, err_len, UGU_B200_E_CUDA, "cudaSetDevice(%d): %s", device, cudaGetErrorString(e));
  unsigned* scratch = nullptr;
  {
    std::lock_guard<std::mutex> lock(g_mu);
    if (!g_scratch_ptr[device]) {
      void* p = nullptr;
      e = cudaGetSymbolAddress(&p, g_static_vis_scratch);
      if (e != cudaSuccess || !p)
        return Err(err, err_len, UGU_B200_E_CUDA, "cudaGetSymbolAddress(static scratch) on device %d: %s", device,
                   cudaGetErrorString(e));
      g_scratch_ptr[device] = (unsigned*)p;
    }
    scratch = g_scratch_ptr[device];
  }
  cudaStream_t st = (cudaStream_t)stream;
  // as many view slots as fit the array for this face count (8 up to 512 Ki faces, 1 at the limit)
  const int slots = (int)std::min<size_t>((size_t)kVisSlots, kStaticVisWords / (size_t)num_faces);
  // the array keeps what the last call (any face count, possibly failed) left: start from zero
  e = cudaMemsetAsync(scratch, 0, (size_t)slots * (size_t)num_faces * 4, st);
  if (e != cudaSuccess) return Err(err, err_len, UGU_B200_E_CUDA, "cudaMemsetAsync: %s", cudaGetErrorString(e));
  VisParams p;
  p.face_id = face_id_dev;
  p.P = (unsigned long long)pixels_per_view;
  p.F = num_faces;
  p.first_view_id = first_view_id;
  p.tmp = scratch;
  p.pixel_count = pixel_count_dev;
  p.view_count = view_count_dev;
  p.best = (unsigned long long*)best_dev;
  const bool vec = (pixels_per_view % 4 == 0) && (((uintptr_t)face_id_dev & 15u) == 0);
  const size_t quads = (size_t)((pixels_per_view + 3) / 4);
  const unsigned gx = (unsigned)((quads + 255) / 256);
  for (int v0 = 0; v0 < num_views; v0 += slots) {
    const dim3 grid(gx, (unsigned)std::min(slots, num_views - v0));
    if (vec) {
      vis_count_kernel<true><<<grid, 256, 0, st>>>(p, v0);
      vis_fold_kernel<true><<<grid, 256, 0, st>>>(p, v0);
    } else {
      vis_count_kernel<false><<<grid, 256, 0, st>>>(p, v0);
      vis_fold_kernel<false><<<grid, 256, 0, st>>>(p, v0);
    }
  }
  e = cudaGetLastError();
  if (e != cudaSuccess) return Err(err, err_len, UGU_B200_E_CUDA, "kernel launch: %s", cudaGetErrorString(e));
  return UGU_B200_OK;
}
