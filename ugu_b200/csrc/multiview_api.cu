// Multi-view dispatch over several GPUs of one box from ONE host process, behind the C-ABI (SURVEY.md 8b last
// row, 8e): what a C++ host calls instead of the reference's sequential orbit loop
// (examples/ex09_renderer_realtime.cc:125-182: set_c2w -> Render per view on one renderer).
//
// Views are independent given a replicated read-only mesh, so rendering has NO data-path collective: device d
// renders its shard of the view list (contiguous or round robin) with the batched launch sets of its own
// ugu_b200_rasterizer, driven by its own host thread.  The one exchange step is the optional gather of the
// per-view face_id maps into a view-ordered buffer on a root device: every chunk of views is sent as soon as it
// is rendered (double-buffered staging on the sender, second stream) straight to its final place, so the
// transfer hides under the rendering of the next chunk.  Transport: NCCL send/recv over NVLink (libnccl.so.2 is
// dlopen-ed, the library does not link it) or cudaMemcpyPeerAsync.  The per-face visibility fold reduces 16 bytes
// per face instead (SUM / SUM / MAX: ncclReduce, or peer copies + a combine kernel).
//
// This file uses only the public C-ABI of the rasterizer (ugu_b200_raster.h) and the CUDA runtime.
#include <cuda_runtime.h>
#include <dlfcn.h>

#include <algorithm>
#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "ugu_b200_raster.h"

namespace {

// ---- the few NCCL entry points used, resolved at run time ---------------------------------------------------------
typedef void* NcclComm;
struct NcclApi {
  void* lib = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
  int (*CommInitAll)(NcclComm*, int, const int*) = nullptr;
  int (*CommDestroy)(NcclComm) = nullptr;
  int (*GroupStart)() = nullptr;
  int (*GroupEnd)() = nullptr;
  int (*Send)(const void*, size_t, int, int, NcclComm, cudaStream_t) = nullptr;
  int (*Recv)(void*, size_t, int, int, NcclComm, cudaStream_t) = nullptr;
  int (*Reduce)(const void*, void*, size_t, int, int, int, NcclComm, cudaStream_t) = nullptr;
  bool ok = false;
};
constexpr int kNcclInt8 = 0, kNcclUint32 = 3, kNcclUint64 = 5;  // ncclDataType_t (nccl.h, stable across 2.x)
constexpr int kNcclSum = 0, kNcclMax = 2;                        // ncclRedOp_t

NcclApi LoadNccl() {
  NcclApi a;
  for (const char* name : {"libnccl.so.2", "libnccl.so"}) {
    a.lib = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
    if (a.lib) break;
  }
  if (!a.lib) return a;
#define UGU_NCCL_SYM(field, sym) *(void**)(&a.field) = dlsym(a.lib, sym)
  UGU_NCCL_SYM(GetErrorString, "ncclGetErrorString");
  UGU_NCCL_SYM(CommInitAll, "ncclCommInitAll");
  UGU_NCCL_SYM(CommDestroy, "ncclCommDestroy");
  UGU_NCCL_SYM(GroupStart, "ncclGroupStart");
  UGU_NCCL_SYM(GroupEnd, "ncclGroupEnd");
  UGU_NCCL_SYM(Send, "ncclSend");
  UGU_NCCL_SYM(Recv, "ncclRecv");
  UGU_NCCL_SYM(Reduce, "ncclReduce");
#undef UGU_NCCL_SYM
  a.ok = a.GetErrorString && a.CommInitAll && a.CommDestroy && a.GroupStart && a.GroupEnd && a.Send && a.Recv &&
         a.Reduce;
  return a;
}

// dst accumulators (+)= src accumulators: SUM / SUM / MAX
__global__ void combine_visibility_kernel(uint32_t* __restrict__ pc, uint32_t* __restrict__ vc,
                                          unsigned long long* __restrict__ best, const uint32_t* __restrict__ pc_s,
                                          const uint32_t* __restrict__ vc_s,
                                          const unsigned long long* __restrict__ best_s, size_t n) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    pc[i] += pc_s[i];
    vc[i] += vc_s[i];
    const unsigned long long a = best[i], b = best_s[i];
    best[i] = a > b ? a : b;
  }
}

struct Buf {
  void* p = nullptr;
  size_t bytes = 0;
};

struct Dev {
  int device = 0;
  ugu_b200_rasterizer* r = nullptr;
  cudaStream_t compute = nullptr, copy = nullptr;
  cudaEvent_t rendered[2] = {nullptr, nullptr}, sent[2] = {nullptr, nullptr};
  Buf stage[2];  // face_id maps of one chunk each
  Buf acc;       // visibility accumulators: best (u64 F), pixel_count (u32 F), view_count (u32 F)
  Buf acc_in;    // root only: a peer's accumulators
  NcclComm comm = nullptr;
  int rc = UGU_B200_OK;
  std::string err;
};

}  // namespace

struct ugu_b200_multiview {
  std::vector<Dev> devs;
  NcclApi nccl;
  int transport = UGU_B200_TRANSPORT_PEER;
  int num_faces = 0;
  Buf gathered;  // internal destination of the gather on the last root (when the caller gives none)
  int gathered_dev = -1;
  std::string error;
};

namespace {

thread_local std::string g_mv_error;

int MvFail(ugu_b200_multiview* mv, int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  if (mv) mv->error = buf;
  g_mv_error = buf;
  return code;
}

#define MV_CU(d, expr)                                                                  \
  do {                                                                                  \
    cudaError_t e__ = (expr);                                                           \
    if (e__ != cudaSuccess) {                                                           \
      (d).err = std::string(#expr) + " failed: " + cudaGetErrorString(e__);             \
      return (d).rc = (e__ == cudaErrorMemoryAllocation ? UGU_B200_E_NOMEM : UGU_B200_E_CUDA); \
    }                                                                                   \
  } while (0)

int GrowBuf(Dev& d, Buf* b, size_t bytes) {
  if (bytes <= b->bytes) return UGU_B200_OK;
  if (b->p) MV_CU(d, cudaFree(b->p));
  b->p = nullptr;
  b->bytes = 0;
  MV_CU(d, cudaMalloc(&b->p, bytes));
  b->bytes = bytes;
  return UGU_B200_OK;
}

// first error of the devices -> the handle's error string
int Collect(ugu_b200_multiview* mv) {
  for (size_t i = 0; i < mv->devs.size(); i++) {
    Dev& d = mv->devs[i];
    if (d.rc != UGU_B200_OK) {
      const int rc = d.rc;
      MvFail(mv, rc, "device %d (slot %zu): %s", d.device, i, d.err.empty() ? ugu_b200_last_error(d.r) : d.err.c_str());
      for (Dev& e : mv->devs) {
        e.rc = UGU_B200_OK;
        e.err.clear();
      }
      return rc;
    }
  }
  return UGU_B200_OK;
}

template <typename Fn>
int OnEveryDevice(ugu_b200_multiview* mv, Fn fn) {
  const int n = (int)mv->devs.size();
  if (n == 1) {
    fn(0);
    return Collect(mv);
  }
  std::vector<std::thread> th;
  th.reserve((size_t)n);
  for (int i = 0; i < n; i++) th.emplace_back([&, i] { fn(i); });
  for (auto& t : th) t.join();
  return Collect(mv);
}

// global view indices of slot `rank` (ugu_b200_partition)
std::vector<int32_t> Shard(int32_t num_views, int32_t world, int32_t rank, int32_t mode) {
  std::vector<int32_t> idx;
  if (mode == UGU_B200_PARTITION_ROUND_ROBIN) {
    for (int32_t i = rank; i < num_views; i += world) idx.push_back(i);
  } else {
    const int32_t base = num_views / world, extra = num_views % world;
    const int32_t start = rank * base + std::min(rank, extra);
    for (int32_t i = 0; i < base + (rank < extra ? 1 : 0); i++) idx.push_back(start + i);
  }
  return idx;
}

int CheckViews(ugu_b200_multiview* mv, int32_t num_views, const ugu_b200_camera* cams, int32_t partition) {
  if (!mv) return MvFail(nullptr, UGU_B200_E_INVALID, "multiview handle is NULL");
  if (num_views <= 0 || !cams) return MvFail(mv, UGU_B200_E_INVALID, "no views");
  if (partition != UGU_B200_PARTITION_CONTIGUOUS && partition != UGU_B200_PARTITION_ROUND_ROBIN)
    return MvFail(mv, UGU_B200_E_INVALID, "unknown partition mode %d", partition);
  for (int32_t i = 1; i < num_views; i++)
    if (cams[i].width != cams[0].width || cams[i].height != cams[0].height)
      return MvFail(mv, UGU_B200_E_INVALID, "all views of a batch must share one image size");
  return UGU_B200_OK;
}

}  // namespace

extern "C" {

int ugu_b200_partition(int32_t num_views, int32_t world, int32_t rank, int32_t mode, int32_t* indices,
                       int32_t* count) {
  if (num_views < 0 || world < 1 || rank < 0 || rank >= world || !count ||
      (mode != UGU_B200_PARTITION_CONTIGUOUS && mode != UGU_B200_PARTITION_ROUND_ROBIN))
    return MvFail(nullptr, UGU_B200_E_INVALID, "bad partition arguments");
  const std::vector<int32_t> idx = Shard(num_views, world, rank, mode);
  *count = (int32_t)idx.size();
  if (indices) std::copy(idx.begin(), idx.end(), indices);
  return UGU_B200_OK;
}

int ugu_b200_multiview_create(const int32_t* devices, int32_t num_devices, int32_t transport,
                              ugu_b200_multiview** out) {
  if (!out) return MvFail(nullptr, UGU_B200_E_INVALID, "out is NULL");
  *out = nullptr;
  if (num_devices <= 0 || !devices) return MvFail(nullptr, UGU_B200_E_INVALID, "no devices");
  if (transport < UGU_B200_TRANSPORT_AUTO || transport > UGU_B200_TRANSPORT_NCCL)
    return MvFail(nullptr, UGU_B200_E_INVALID, "unknown transport %d", transport);
  ugu_b200_multiview* mv = new ugu_b200_multiview;
  mv->devs.resize((size_t)num_devices);
  auto fail = [&](int rc) {
    std::string msg = g_mv_error;
    ugu_b200_multiview_destroy(mv);
    g_mv_error = msg;
    return rc;
  };
  bool distinct = true;
  for (int i = 0; i < num_devices; i++)
    for (int j = 0; j < i; j++) distinct = distinct && devices[i] != devices[j];
  for (int i = 0; i < num_devices; i++) {
    Dev& d = mv->devs[(size_t)i];
    d.device = devices[i];
    int rc = ugu_b200_rasterizer_create(d.device, &d.r);
    if (rc) return fail(MvFail(nullptr, rc, "device %d: %s", d.device, ugu_b200_last_error(nullptr)));
    cudaError_t e = cudaSetDevice(d.device);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&d.compute, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&d.copy, cudaStreamNonBlocking);
    for (int k = 0; k < 2 && e == cudaSuccess; k++) {
      e = cudaEventCreateWithFlags(&d.rendered[k], cudaEventDisableTiming);
      if (e == cudaSuccess) e = cudaEventCreateWithFlags(&d.sent[k], cudaEventDisableTiming);
    }
    if (e != cudaSuccess) return fail(MvFail(nullptr, UGU_B200_E_CUDA, "device %d: %s", d.device, cudaGetErrorString(e)));
  }
  // peer access between every pair of distinct devices (unsupported pairs fall back to staged copies)
  for (int i = 0; i < num_devices; i++)
    for (int j = 0; j < num_devices; j++) {
      if (devices[i] == devices[j]) continue;
      int can = 0;
      if (cudaDeviceCanAccessPeer(&can, devices[i], devices[j]) == cudaSuccess && can) {
        cudaSetDevice(devices[i]);
        const cudaError_t e = cudaDeviceEnablePeerAccess(devices[j], 0);
        if (e != cudaSuccess) cudaGetLastError();  // already enabled / not supported
      }
    }
  if (const char* e = getenv("UGU_B200_MV_TRANSPORT")) {
    if (!strcmp(e, "peer")) transport = UGU_B200_TRANSPORT_PEER;
    if (!strcmp(e, "nccl")) transport = UGU_B200_TRANSPORT_NCCL;
  }
  mv->transport = UGU_B200_TRANSPORT_PEER;
  if (transport != UGU_B200_TRANSPORT_PEER && num_devices > 1) {
    // NCCL needs one communicator rank per DISTINCT device
    if (distinct) mv->nccl = LoadNccl();
    if (distinct && mv->nccl.ok) {
      std::vector<NcclComm> comms((size_t)num_devices, nullptr);
      std::vector<int> devs(devices, devices + num_devices);
      const int rc = mv->nccl.CommInitAll(comms.data(), num_devices, devs.data());
      if (rc == 0) {
        for (int i = 0; i < num_devices; i++) mv->devs[(size_t)i].comm = comms[(size_t)i];
        mv->transport = UGU_B200_TRANSPORT_NCCL;
      } else if (transport == UGU_B200_TRANSPORT_NCCL) {
        return fail(MvFail(nullptr, UGU_B200_E_CUDA, "ncclCommInitAll failed: %s", mv->nccl.GetErrorString(rc)));
      }
    } else if (transport == UGU_B200_TRANSPORT_NCCL) {
      return fail(MvFail(nullptr, UGU_B200_E_INVALID,
                         distinct ? "libnccl.so.2 could not be loaded" : "NCCL needs distinct devices"));
    }
  }
  *out = mv;
  return UGU_B200_OK;
}

void ugu_b200_multiview_destroy(ugu_b200_multiview* mv) {
  if (!mv) return;
  for (Dev& d : mv->devs) {
    cudaSetDevice(d.device);
    if (d.compute) cudaStreamSynchronize(d.compute);
    if (d.copy) cudaStreamSynchronize(d.copy);
    if (d.comm && mv->nccl.ok) mv->nccl.CommDestroy(d.comm);
    for (Buf* b : {&d.stage[0], &d.stage[1], &d.acc, &d.acc_in})
      if (b->p) cudaFree(b->p);
    for (int k = 0; k < 2; k++) {
      if (d.rendered[k]) cudaEventDestroy(d.rendered[k]);
      if (d.sent[k]) cudaEventDestroy(d.sent[k]);
    }
    if (d.compute) cudaStreamDestroy(d.compute);
    if (d.copy) cudaStreamDestroy(d.copy);
    if (d.r) ugu_b200_rasterizer_destroy(d.r);
  }
  if (mv->gathered.p && mv->gathered_dev >= 0) {
    cudaSetDevice(mv->gathered_dev);
    cudaFree(mv->gathered.p);
  }
  delete mv;
}

int32_t ugu_b200_multiview_size(const ugu_b200_multiview* mv) { return mv ? (int32_t)mv->devs.size() : 0; }

int32_t ugu_b200_multiview_transport(const ugu_b200_multiview* mv) { return mv ? mv->transport : 0; }

ugu_b200_rasterizer* ugu_b200_multiview_rasterizer(ugu_b200_multiview* mv, int32_t i) {
  return (mv && i >= 0 && (size_t)i < mv->devs.size()) ? mv->devs[(size_t)i].r : nullptr;
}

const char* ugu_b200_multiview_last_error(const ugu_b200_multiview* mv) {
  return mv ? mv->error.c_str() : g_mv_error.c_str();
}

int ugu_b200_multiview_set_option(ugu_b200_multiview* mv, const ugu_b200_option* opt) {
  if (!mv || !opt) return MvFail(mv, UGU_B200_E_INVALID, "NULL argument");
  for (Dev& d : mv->devs)
    if ((d.rc = ugu_b200_rasterizer_set_option(d.r, opt))) break;
  return Collect(mv);
}

// PrepareMesh on every device: the mesh is replicated (each device repacks its own copy from the host arrays)
int ugu_b200_multiview_upload_mesh(ugu_b200_multiview* mv, const ugu_b200_mesh_desc* mesh) {
  if (!mv || !mesh) return MvFail(mv, UGU_B200_E_INVALID, "NULL argument");
  mv->num_faces = mesh->num_faces;
  return OnEveryDevice(mv, [&](int i) {
    Dev& d = mv->devs[(size_t)i];
    d.rc = ugu_b200_rasterizer_upload_mesh(d.r, mesh);
  });
}

// The ex09 loop over `num_views` cameras, sharded: view i lands at element offset i * W*H of every non-NULL
// HOST output, whatever device rendered it.
int ugu_b200_multiview_render_batch(ugu_b200_multiview* mv, int32_t num_views, const ugu_b200_camera* cams,
                                    int32_t partition, const ugu_b200_outputs* host_out) {
  int rc = CheckViews(mv, num_views, cams, partition);
  if (rc) return rc;
  if (!host_out) return MvFail(mv, UGU_B200_E_INVALID, "all arguments are nullptr. nothing to do");
  const int world = (int)mv->devs.size();
  const size_t P = (size_t)cams[0].width * cams[0].height;
  return OnEveryDevice(mv, [&](int i) {
    Dev& d = mv->devs[(size_t)i];
    const std::vector<int32_t> idx = Shard(num_views, world, i, partition);
    if (idx.empty()) return;
    // runs of consecutive global views -> one batch call each (contiguous: one run; round robin: one per view)
    size_t a = 0;
    while (a < idx.size() && d.rc == UGU_B200_OK) {
      size_t b = a + 1;
      while (b < idx.size() && idx[b] == idx[b - 1] + 1) b++;
      const size_t first = (size_t)idx[a];
      ugu_b200_outputs o{};
      if (host_out->color) o.color = host_out->color + first * P * 3;
      if (host_out->depth) o.depth = host_out->depth + first * P;
      if (host_out->normal) o.normal = host_out->normal + first * P * 3;
      if (host_out->mask) o.mask = host_out->mask + first * P;
      if (host_out->face_id) o.face_id = host_out->face_id + first * P;
      d.rc = ugu_b200_rasterizer_render_batch(d.r, (int32_t)(b - a), cams + first, &o, 0, nullptr);
      a = b;
    }
  });
}

// Render the face_id map of every view (sharded) and gather the maps, in view order, on device slot `root`.
int ugu_b200_multiview_gather_face_ids(ugu_b200_multiview* mv, int32_t num_views, const ugu_b200_camera* cams,
                                       int32_t partition, int32_t root, int32_t chunk_views,
                                       int32_t* face_id_root_dev, int32_t* face_id_host) {
  int rc = CheckViews(mv, num_views, cams, partition);
  if (rc) return rc;
  const int world = (int)mv->devs.size();
  if (root < 0 || root >= world) return MvFail(mv, UGU_B200_E_INVALID, "root %d out of range", root);
  if (((uintptr_t)face_id_root_dev) & 15u) return MvFail(mv, UGU_B200_E_INVALID, "device outputs must be 16-byte aligned");
  const size_t P = (size_t)cams[0].width * cams[0].height;
  const int chunk = std::max(1, std::min(chunk_views > 0 ? chunk_views : 16, num_views));
  Dev& R = mv->devs[(size_t)root];
  int32_t* dst = face_id_root_dev;
  if (!dst) {  // internal destination
    if (mv->gathered_dev != R.device && mv->gathered.p) {
      cudaSetDevice(mv->gathered_dev);
      cudaFree(mv->gathered.p);
      mv->gathered = Buf{};
    }
    cudaSetDevice(R.device);
    mv->gathered_dev = R.device;
    if (GrowBuf(R, &mv->gathered, (size_t)num_views * P * 4 + 16)) return Collect(mv);
    dst = (int32_t*)mv->gathered.p;
  }
  const bool nccl = mv->transport == UGU_B200_TRANSPORT_NCCL;
  std::vector<std::vector<int32_t>> shards((size_t)world);
  for (int i = 0; i < world; i++) shards[(size_t)i] = Shard(num_views, world, i, partition);

  rc = OnEveryDevice(mv, [&](int i) -> void {
    Dev& d = mv->devs[(size_t)i];
    auto body = [&]() -> int {
      MV_CU(d, cudaSetDevice(d.device));
      const std::vector<int32_t>& idx = shards[(size_t)i];
      const bool in_place = i == root && partition == UGU_B200_PARTITION_CONTIGUOUS;
      if (!in_place)
        for (int k = 0; k < 2; k++)
          if (GrowBuf(d, &d.stage[k], (size_t)chunk * P * 4 + 16)) return d.rc;
      int slot = 0;
      bool pending[2] = {false, false};
      for (size_t a = 0; a < idx.size(); a += (size_t)chunk, slot ^= 1) {
        const int k = (int)std::min<size_t>((size_t)chunk, idx.size() - a);
        ugu_b200_outputs o{};
        // the chunk's cameras: consecutive for a contiguous shard, strided otherwise
        std::vector<ugu_b200_camera> strided;
        const ugu_b200_camera* cc = cams + idx[a];
        if (partition != UGU_B200_PARTITION_CONTIGUOUS) {
          for (int v = 0; v < k; v++) strided.push_back(cams[idx[a + (size_t)v]]);
          cc = strided.data();
        }
        if (in_place) {
          o.face_id = dst + (size_t)idx[a] * P;
          if ((d.rc = ugu_b200_rasterizer_render_batch(d.r, k, cc, &o, 1, d.compute))) return d.rc;
          continue;
        }
        if (pending[slot]) MV_CU(d, cudaStreamWaitEvent(d.compute, d.sent[slot], 0));  // staging slot free again?
        o.face_id = (int32_t*)d.stage[slot].p;
        if ((d.rc = ugu_b200_rasterizer_render_batch(d.r, k, cc, &o, 1, d.compute))) return d.rc;
        MV_CU(d, cudaEventRecord(d.rendered[slot], d.compute));
        MV_CU(d, cudaStreamWaitEvent(d.copy, d.rendered[slot], 0));
        if (nccl && i != root) {
          const int e = mv->nccl.Send(d.stage[slot].p, (size_t)k * P * 4, kNcclInt8, root, d.comm, d.copy);
          if (e) {
            d.err = std::string("ncclSend failed: ") + mv->nccl.GetErrorString(e);
            return d.rc = UGU_B200_E_CUDA;
          }
        } else {
          // straight to the final place on the root (runs of consecutive views in one copy)
          int v = 0;
          while (v < k) {
            int w = v + 1;
            while (w < k && idx[a + (size_t)w] == idx[a + (size_t)w - 1] + 1) w++;
            MV_CU(d, cudaMemcpyPeerAsync(dst + (size_t)idx[a + (size_t)v] * P, R.device,
                                         (const int32_t*)d.stage[slot].p + (size_t)v * P, d.device,
                                         (size_t)(w - v) * P * 4, d.copy));
            v = w;
          }
        }
        MV_CU(d, cudaEventRecord(d.sent[slot], d.copy));
        pending[slot] = true;
      }
      if (nccl && i == root) {
        // the receives, chunk by chunk in the order the peers send them, each straight to its final place
        // (contiguous shards: a chunk is one run of views; round robin: one receive per view)
        size_t most = 0;
        for (int p = 0; p < world; p++) most = std::max(most, shards[(size_t)p].size());
        for (size_t a = 0; a < most; a += (size_t)chunk) {
          bool any = false;
          for (int p = 0; p < world; p++) {
            if (p == root || a >= shards[(size_t)p].size()) continue;
            if (!any) mv->nccl.GroupStart();
            any = true;
            const std::vector<int32_t>& pidx = shards[(size_t)p];
            const int k = (int)std::min<size_t>((size_t)chunk, pidx.size() - a);
            if (partition == UGU_B200_PARTITION_CONTIGUOUS) {
              mv->nccl.Recv(dst + (size_t)pidx[a] * P, (size_t)k * P * 4, kNcclInt8, p, d.comm, d.copy);
            } else {
              // a strided chunk arrives as one message: land it in the peer-sized staging area of the root's
              // copy stream, then scatter (device-to-device) to the views' places
              // (kept simple: one receive per view would need k matching sends)
              Buf& in = d.acc_in;
              if (GrowBuf(d, &in, (size_t)chunk * P * 4 * (size_t)world + 16)) {
                mv->nccl.GroupEnd();
                return d.rc;
              }
              int32_t* land = (int32_t*)in.p + (size_t)p * (size_t)chunk * P;
              mv->nccl.Recv(land, (size_t)k * P * 4, kNcclInt8, p, d.comm, d.copy);
            }
          }
          if (any) {
            const int e = mv->nccl.GroupEnd();
            if (e) {
              d.err = std::string("ncclRecv failed: ") + mv->nccl.GetErrorString(e);
              return d.rc = UGU_B200_E_CUDA;
            }
            if (partition != UGU_B200_PARTITION_CONTIGUOUS) {
              for (int p = 0; p < world; p++) {
                if (p == root || a >= shards[(size_t)p].size()) continue;
                const std::vector<int32_t>& pidx = shards[(size_t)p];
                const int k = (int)std::min<size_t>((size_t)chunk, pidx.size() - a);
                const int32_t* land = (const int32_t*)d.acc_in.p + (size_t)p * (size_t)chunk * P;
                for (int v = 0; v < k; v++)
                  MV_CU(d, cudaMemcpyAsync(dst + (size_t)pidx[a + (size_t)v] * P, land + (size_t)v * P, P * 4,
                                           cudaMemcpyDeviceToDevice, d.copy));
              }
            }
          }
        }
      }
      MV_CU(d, cudaStreamSynchronize(d.compute));
      MV_CU(d, cudaStreamSynchronize(d.copy));
      return UGU_B200_OK;
    };
    body();
  });
  if (rc) return rc;
  if (face_id_host) {
    cudaSetDevice(R.device);
    const cudaError_t e = cudaMemcpy(face_id_host, dst, (size_t)num_views * P * 4, cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) return MvFail(mv, UGU_B200_E_CUDA, "cudaMemcpy failed: %s", cudaGetErrorString(e));
  }
  return UGU_B200_OK;
}

const int32_t* ugu_b200_multiview_gathered(const ugu_b200_multiview* mv) {
  return mv ? (const int32_t*)mv->gathered.p : nullptr;
}

// Render the face_id maps (sharded), fold them per device into per-face visibility statistics
// (ugu_b200_face_visibility_device) and reduce the accumulators to `root`: SUM / SUM / MAX of 16 bytes per face.
int ugu_b200_multiview_reduce_face_visibility(ugu_b200_multiview* mv, int32_t num_views,
                                              const ugu_b200_camera* cams, int32_t partition, int32_t root,
                                              int32_t chunk_views, uint32_t* pixel_count_host,
                                              uint32_t* view_count_host, int32_t* best_view_host,
                                              uint32_t* best_count_host) {
  int rc = CheckViews(mv, num_views, cams, partition);
  if (rc) return rc;
  const int world = (int)mv->devs.size();
  if (root < 0 || root >= world) return MvFail(mv, UGU_B200_E_INVALID, "root %d out of range", root);
  if (mv->num_faces <= 0) return MvFail(mv, UGU_B200_E_STATE, "mesh has not been initialized");
  const size_t F = (size_t)mv->num_faces;
  const size_t P = (size_t)cams[0].width * cams[0].height;
  const int chunk = std::max(1, std::min(chunk_views > 0 ? chunk_views : 16, num_views));
  const bool nccl = mv->transport == UGU_B200_TRANSPORT_NCCL;
  const size_t acc_bytes = F * 16;
  rc = OnEveryDevice(mv, [&](int i) -> void {
    Dev& d = mv->devs[(size_t)i];
    auto body = [&]() -> int {
      MV_CU(d, cudaSetDevice(d.device));
      if (GrowBuf(d, &d.acc, acc_bytes + 16) || GrowBuf(d, &d.stage[0], (size_t)chunk * P * 4 + 16)) return d.rc;
      unsigned long long* best = (unsigned long long*)d.acc.p;
      uint32_t* pc = (uint32_t*)(best + F);
      uint32_t* vc = pc + F;
      MV_CU(d, cudaMemsetAsync(d.acc.p, 0, acc_bytes, d.compute));
      const std::vector<int32_t> idx = Shard(num_views, world, i, partition);
      for (size_t a = 0; a < idx.size(); a += (size_t)chunk) {
        const int k = (int)std::min<size_t>((size_t)chunk, idx.size() - a);
        std::vector<ugu_b200_camera> strided;
        const ugu_b200_camera* cc = cams + idx[a];
        if (partition != UGU_B200_PARTITION_CONTIGUOUS) {
          for (int v = 0; v < k; v++) strided.push_back(cams[idx[a + (size_t)v]]);
          cc = strided.data();
        }
        ugu_b200_outputs o{};
        o.face_id = (int32_t*)d.stage[0].p;
        if ((d.rc = ugu_b200_rasterizer_render_batch(d.r, k, cc, &o, 1, d.compute))) return d.rc;
        if (partition == UGU_B200_PARTITION_CONTIGUOUS) {
          if ((d.rc = ugu_b200_face_visibility_device(d.r, o.face_id, k, (int64_t)P, idx[a], (int32_t)F, pc, vc,
                                                      (uint64_t*)best, d.compute)))
            return d.rc;
        } else {  // strided global view ids: one fold per view
          for (int v = 0; v < k; v++)
            if ((d.rc = ugu_b200_face_visibility_device(d.r, o.face_id + (size_t)v * P, 1, (int64_t)P,
                                                        idx[a + (size_t)v], (int32_t)F, pc, vc, (uint64_t*)best,
                                                        d.compute)))
              return d.rc;
        }
      }
      if (nccl) {
        int e = mv->nccl.GroupStart();
        if (!e) e = mv->nccl.Reduce(pc, pc, F, kNcclUint32, kNcclSum, root, d.comm, d.compute);
        if (!e) e = mv->nccl.Reduce(vc, vc, F, kNcclUint32, kNcclSum, root, d.comm, d.compute);
        if (!e) e = mv->nccl.Reduce(best, best, F, kNcclUint64, kNcclMax, root, d.comm, d.compute);
        const int e2 = mv->nccl.GroupEnd();
        if (e || e2) {
          d.err = std::string("ncclReduce failed: ") + mv->nccl.GetErrorString(e ? e : e2);
          return d.rc = UGU_B200_E_CUDA;
        }
      }
      MV_CU(d, cudaStreamSynchronize(d.compute));
      return UGU_B200_OK;
    };
    body();
  });
  if (rc) return rc;
  Dev& R = mv->devs[(size_t)root];
  auto finish = [&]() -> int {
    MV_CU(R, cudaSetDevice(R.device));
    unsigned long long* best = (unsigned long long*)R.acc.p;
    uint32_t* pc = (uint32_t*)(best + F);
    uint32_t* vc = pc + F;
    if (!nccl) {  // peer copies + combine on the root
      if (GrowBuf(R, &R.acc_in, acc_bytes + 16)) return R.rc;
      for (int p = 0; p < world; p++) {
        if (p == root) continue;
        Dev& S = mv->devs[(size_t)p];
        MV_CU(R, cudaMemcpyPeerAsync(R.acc_in.p, R.device, S.acc.p, S.device, acc_bytes, R.compute));
        const unsigned long long* bs = (const unsigned long long*)R.acc_in.p;
        const uint32_t* pcs = (const uint32_t*)(bs + F);
        combine_visibility_kernel<<<(unsigned)((F + 255) / 256), 256, 0, R.compute>>>(pc, vc, best, pcs, pcs + F, bs, F);
        MV_CU(R, cudaGetLastError());
      }
    }
    std::vector<unsigned long long> hb;
    if (best_view_host || best_count_host) hb.resize(F);
    if (pixel_count_host) MV_CU(R, cudaMemcpyAsync(pixel_count_host, pc, F * 4, cudaMemcpyDeviceToHost, R.compute));
    if (view_count_host) MV_CU(R, cudaMemcpyAsync(view_count_host, vc, F * 4, cudaMemcpyDeviceToHost, R.compute));
    if (!hb.empty()) MV_CU(R, cudaMemcpyAsync(hb.data(), best, F * 8, cudaMemcpyDeviceToHost, R.compute));
    MV_CU(R, cudaStreamSynchronize(R.compute));
    for (size_t f = 0; f < hb.size(); f++) {
      const unsigned long long b = hb[f];
      if (best_view_host) best_view_host[f] = b ? (int32_t)(0xFFFFFFFFu - (uint32_t)(b & 0xFFFFFFFFull)) : -1;
      if (best_count_host) best_count_host[f] = (uint32_t)(b >> 32);
    }
    return UGU_B200_OK;
  };
  finish();
  return Collect(mv);
}

}  // extern "C"
