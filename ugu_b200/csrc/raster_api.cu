// C-ABI implementation (include/ugu_b200_raster.h) over the kernels in raster_kernels.cuh.
// One rasterizer handle = one CUDA device + one stream + the device-resident SoA mesh + per-view
// scratch (projected vertices, visibility buffer, large-triangle queue).  No CPU fallback anywhere:
// without a usable sm_100 device every computing entry point returns UGU_B200_E_CUDA.
#include <cuda_runtime.h>

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <utility>
#include <vector>

#include "host_copy_pool.h"
#include "raster_kernels.cuh"
#include "uv_raster_kernels.cuh"
#include "visibility_kernels.cuh"
#include "visibility_tester_kernels.cuh"
#include "depth2mesh_kernels.cuh"
#include "gbuffer_kernels.cuh"
#include "raytrace_kernels.cuh"
#include "ugu_b200_raster.h"

using namespace ugu_b200;

namespace {

thread_local std::string g_last_error;

struct DevBuf {
  void* p = nullptr;
  size_t bytes = 0;
};

constexpr int kEventPool = 512;           // timed (stage, view) slots before a drain
constexpr size_t kStageChunk = 1u << 20;  // pinned staging chunk for pageable destinations (grows so that a
                                          // call never needs more than kStageChunksMax events)
constexpr size_t kStageChunksMax = 512;

}  // namespace

struct ugu_b200_rasterizer {
  int device = 0;
  cudaStream_t stream = nullptr;       // compute stream (own)
  cudaStream_t copy_stream = nullptr;  // D2H stream for batch overlap
  cudaEvent_t ev_band[2] = {nullptr, nullptr};  // host-form Render: the resolve bands' completion (BandPlan)
  int bands = 0;                       // UGU_B200_BANDS: 1 = host-form Render resolves in two bands (off: it loses)
  std::string error;
  int deferred_status = UGU_B200_OK;

  ugu_b200_option opt{};
  bool mesh_ready = false;
  MeshDev mesh{};
  bool has_nidx = false, has_uvidx = false, has_matids = false;
  std::vector<bool> tex_present;
  std::vector<void*> mesh_allocs;  // everything owned by the current mesh

  // per-view scratch
  DevBuf proj, ray_dx, ray_dy, vis, q_face, q_job, counters, order, tile_cnt, tile_bin;
  size_t tile_slab = 0, tile_slabs = 0;  // tiles per slab / slabs known to hold zero counts
  size_t vis_pixels = 0;  // pixels per visibility slab
  size_t vis_slabs = 0;   // slabs known to hold EMPTY
  int max_batch = 0;      // views per launch set of render_batch (0 = adaptive, UGU_B200_BATCH overrides)
  unsigned long long hazard_seen = 0;  // device hazard total already reported (see SnapshotCounters)
  cudaEvent_t ev_scratch = nullptr;    // recorded after every launch set: cross-stream ordering of the scratch
  cudaStream_t last_stream = nullptr;  // stream of the last launch set that used the scratch
  bool scratch_used = false;
  int coop_max = 0;  // bbox samples up to which a triangle is rasterised inside raster_kernel; larger
                     // ones go to the tile-job queue.  0 = adaptive (see RenderView)
  bool use_pdl = true;  // programmatic dependent launch of the per-view kernels (UGU_B200_PDL=0 disables)
  int tiles = -1;       // per-tile triangle lists for large triangles (tile_kernel): -1 = by mesh size, 0 / 1 forced (UGU_B200_TILES)
  int fold_binned = -1; // resolve_kernel owns the tile-job queue, no binned_kernel launch: -1 = by mesh size (UGU_B200_FOLD_BINNED)
  int two_pass = -1;    // occlusion ordering of raster_kernel: -1 = by mesh size, 0 / 1 forced (UGU_B200_TWOPASS)

  // internal device outputs for the host-destination paths (two sets for batch overlap)
  DevBuf out_color[2], out_depth[2], out_normal[2], out_mask[2], out_fid[2];
  cudaEvent_t ev_render[2] = {nullptr, nullptr};
  cudaEvent_t ev_copied[2] = {nullptr, nullptr};
  bool copied_pending[2] = {false, false};

  // UV-space rasterization scratch (uv_raster_kernels.cuh)
  DevBuf uv_win, uv_counter, uv_q_face, uv_q_job, uv_rec, uv_in, uv_tex, uv_mask;
  size_t uv_win_clean = 0;  // texels of uv_win known to hold -1

  // face-visibility consumer scratch (visibility_kernels.cuh)
  DevBuf vis_tmp, vis_in, vis_acc;
  int vis_tmp_faces = 0;  // faces per slot the scratch was zeroed for

  // keyframe visibility pass (visibility_tester_kernels.cuh): mesh copy, accumulators, record / image scratch
  DevBuf vt_pos, vt_nrm, vt_idx, vt_acc, vt_owner, vt_vvis, vt_vrec, vt_fvis, vt_frec, vt_img, vt_zbuf;
  int vt_V = 0, vt_F = 0;
  size_t vt_owner_clean = 0;  // elements of vt_owner known to hold kVtNoOwner

  // Depth2Mesh scratch + the device-resident result of the last call (depth2mesh_kernels.cuh)
  DevBuf d2m_in_depth, d2m_in_color, d2m_small, d2m_block_sums, d2m_vid, d2m_fref, d2m_vertices, d2m_faces,
      d2m_normals, d2m_face_normals, d2m_uv, d2m_colors, d2m_pc, d2m_nimg;
  unsigned long long* d2m_totals_host = nullptr;  // pinned: packed (faces << 32 | vertices)
  cudaStream_t d2m_stream = nullptr;               // stream of the last call
  int d2m_w = 0, d2m_h = 0, d2m_has_uv = 0, d2m_has_colors = 0;
  bool d2m_valid = false;

  DevBuf gbuf_arena;  // device outputs of the host-form G-buffer call (gbuffer_kernels.cuh)
  DevBuf rt_rays;     // per-pixel ray tables of the Raytracer parity mode (raytrace_kernels.cuh), 32 B / pixel
  float mesh_max_abs = 0.f;  // max |vertex coordinate| of the uploaded mesh

  // pinned staging for pageable destinations
  void* staging = nullptr;
  size_t staging_bytes = 0;
  std::vector<cudaEvent_t> chunk_events;
  // answers of cudaPointerGetAttributes for the last destinations seen (callers render into the same images
  // over and over).  A stale answer costs speed, never correctness: cudaMemcpyAsync accepts pageable memory
  // and the staged path accepts pinned memory.
  struct PinEntry {
    const void* p;
    bool pinned;
  };
  PinEntry pin_cache[16] = {};
  int pin_next = 0;
  HostCopyPool* copy_pool = nullptr;  // created by the first pageable destination
  int copy_threads = 0;               // host threads per pageable copy, the caller included (UGU_B200_COPY_THREADS)
  unsigned long long* counters_host = nullptr;  // pinned snapshot of the device counters
  bool counters_mapped = false;                 // the resolve kernel of a single view writes it directly ...
  bool host_snapshot = false;                   // ... during the host form of Render (0.7 us per view: not otherwise)

  // profiling
  bool profiling = false;
  uint32_t stage_mask = 0;
  std::vector<cudaEvent_t> ev_pool;  // 2 per slot
  std::vector<int> ev_stage;         // stage of each used slot
  int ev_used = 0;
  ugu_b200_stats stats{};
};

namespace {

int Fail(ugu_b200_rasterizer* r, int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  if (r) r->error = buf;
  g_last_error = buf;
  return code;
}

#define CU_TRY(r, expr)                                                                     \
  do {                                                                                      \
    cudaError_t e__ = (expr);                                                               \
    if (e__ != cudaSuccess)                                                                 \
      return Fail((r), e__ == cudaErrorMemoryAllocation ? UGU_B200_E_NOMEM : UGU_B200_E_CUDA, \
                  "%s failed: %s", #expr, cudaGetErrorString(e__));                         \
  } while (0)

int Grow(ugu_b200_rasterizer* r, DevBuf* b, size_t bytes) {
  if (bytes <= b->bytes) return UGU_B200_OK;
  if (b->p) CU_TRY(r, cudaFree(b->p));
  b->p = nullptr;
  b->bytes = 0;
  CU_TRY(r, cudaMalloc(&b->p, bytes));
  b->bytes = bytes;
  return UGU_B200_OK;
}

void FreeBuf(DevBuf* b) {
  if (b->p) cudaFree(b->p);
  b->p = nullptr;
  b->bytes = 0;
}

void FreeMesh(ugu_b200_rasterizer* r) {
  for (void* p : r->mesh_allocs) cudaFree(p);
  r->mesh_allocs.clear();
  r->mesh = MeshDev{};
  r->mesh_ready = false;
  r->has_nidx = r->has_uvidx = r->has_matids = false;
  r->tex_present.clear();
}

template <typename T>
int MeshAlloc(ugu_b200_rasterizer* r, T** out, size_t count) {
  void* p = nullptr;
  CU_TRY(r, cudaMalloc(&p, std::max<size_t>(count, 1) * sizeof(T)));
  r->mesh_allocs.push_back(p);
  *out = static_cast<T*>(p);
  return UGU_B200_OK;
}

inline float __int_as_float_host(int32_t v) {
  float f;
  memcpy(&f, &v, 4);
  return f;
}

inline unsigned Blocks(size_t n, unsigned threads) { return (unsigned)((n + threads - 1) / threads); }

bool IndicesInRange(const int32_t* idx, size_t count, int32_t limit) {
  int32_t lo = 0, hi = -1;
  for (size_t i = 0; i < count; i++) {
    lo = std::min(lo, idx[i]);
    hi = std::max(hi, idx[i]);
  }
  return lo >= 0 && hi < limit;
}

// First-use reordering of an indexed attribute stream (positions by vertex_indices, normals by
// normal_indices, uvs by uv_indices): element k of the new array is the k-th distinct element the
// face list references, unreferenced ones follow.  Faces keep their order (face ids are outputs);
// vertex / normal / uv numbering is internal, so this is invisible in the results.  It turns the
// per-face gathers of consecutive faces into near-contiguous accesses.
struct Reorder {
  std::vector<int32_t> new_of_old;  // old index -> new index
  std::vector<int32_t> idx;         // remapped 3F indices
  bool active = false;
};

// `face_order` (optional, slot -> face): the order in which the faces are walked when numbers are handed out
// (the raster order, see BuildFaceOrder); the remapped index list itself stays in face order.
void BuildReorder(const int32_t* idx, size_t n_idx, int count, Reorder* out,
                  const std::vector<int32_t>* face_order = nullptr) {
  out->new_of_old.assign((size_t)count, -1);
  out->idx.resize(n_idx);
  int32_t next = 0;
  const size_t nf = n_idx / 3;
  for (size_t s = 0; s < nf; s++) {
    const size_t f = face_order ? (size_t)(*face_order)[s] : s;
    for (int k = 0; k < 3; k++) {
      int32_t& slot = out->new_of_old[(size_t)idx[3 * f + k]];
      if (slot < 0) slot = next++;
    }
  }
  for (size_t i = 0; i < n_idx; i++) out->idx[i] = out->new_of_old[(size_t)idx[i]];
  for (int v = 0; v < count; v++)
    if (out->new_of_old[(size_t)v] < 0) out->new_of_old[(size_t)v] = next++;
  out->active = true;
}

// Raster order of the faces (slot -> face).  Face ids are outputs and keep their meaning; the ORDER in which
// raster_kernel walks the faces is free (the depth test is an order-independent atomicMin), and mesh files
// come in any order (a subdivided icosphere's consecutive faces lie all over the sphere).  Faces are sorted
// by (dominant axis and sign of the geometric normal, Morton code of the centroid): a warp's 32 consecutive
// faces then form a compact patch of similar orientation — its projected vertices, its visibility-buffer
// pixels and its attribute records are neighbours in memory, and the patch as a whole faces towards or away
// from the camera, which is what the two-pass occlusion ordering (raster_kernels.cuh) classifies.
uint32_t MortonSpread10(uint32_t v) {
  v &= 1023u;
  v = (v | (v << 16)) & 0x030000FFu;
  v = (v | (v << 8)) & 0x0300F00Fu;
  v = (v | (v << 4)) & 0x030C30C3u;
  v = (v | (v << 2)) & 0x09249249u;
  return v;
}

void BuildFaceOrder(const float* vtx, const int32_t* idx, int F, std::vector<int32_t>* slot_face) {
  float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
  std::vector<float> cen((size_t)F * 3);
  std::vector<uint8_t> bucket((size_t)F);
  for (int f = 0; f < F; f++) {
    const float* a = vtx + 3 * (size_t)idx[3 * (size_t)f];
    const float* b = vtx + 3 * (size_t)idx[3 * (size_t)f + 1];
    const float* c = vtx + 3 * (size_t)idx[3 * (size_t)f + 2];
    const float u[3] = {b[0] - a[0], b[1] - a[1], b[2] - a[2]}, v[3] = {c[0] - a[0], c[1] - a[1], c[2] - a[2]};
    const float n[3] = {u[1] * v[2] - u[2] * v[1], u[2] * v[0] - u[0] * v[2], u[0] * v[1] - u[1] * v[0]};
    int ax = 0;
    if (fabsf(n[1]) > fabsf(n[ax])) ax = 1;
    if (fabsf(n[2]) > fabsf(n[ax])) ax = 2;
    bucket[(size_t)f] = (uint8_t)(2 * ax + (n[ax] < 0.f ? 1 : 0));
    for (int k = 0; k < 3; k++) {
      const float m = (a[k] + b[k] + c[k]) * (1.0f / 3.0f);
      cen[3 * (size_t)f + k] = m;
      if (m == m && fabsf(m) < FLT_MAX) {  // finite
        lo[k] = std::min(lo[k], m);
        hi[k] = std::max(hi[k], m);
      }
    }
  }
  float scale[3];
  for (int k = 0; k < 3; k++) scale[k] = hi[k] > lo[k] ? 1023.0f / (hi[k] - lo[k]) : 0.f;
  std::vector<uint64_t> key((size_t)F), tmp((size_t)F);
  for (int f = 0; f < F; f++) {
    uint32_t q[3];
    for (int k = 0; k < 3; k++) {
      const float t = (cen[3 * (size_t)f + k] - lo[k]) * scale[k];
      q[k] = (t >= 0.f && t <= 1023.f) ? (uint32_t)t : 0u;  // NaN / inf -> 0
    }
    const uint64_t code = ((uint64_t)bucket[(size_t)f] << 30) | (MortonSpread10(q[0]) | (MortonSpread10(q[1]) << 1) |
                                                                   (MortonSpread10(q[2]) << 2));
    key[(size_t)f] = (code << 31) | (uint64_t)f;  // F < 2^26 (validated)
  }
  // LSD radix sort on the 33 code bits (3 passes of 11 bits): stable, so equal codes keep face order
  for (int pass = 0; pass < 3; pass++) {
    const int shift = 31 + 11 * pass;
    std::vector<size_t> count(2049, 0);
    for (int f = 0; f < F; f++) count[((key[(size_t)f] >> shift) & 2047u) + 1]++;
    for (int i = 0; i < 2048; i++) count[i + 1] += count[i];
    for (int f = 0; f < F; f++) tmp[count[(key[(size_t)f] >> shift) & 2047u]++] = key[(size_t)f];
    key.swap(tmp);
  }
  slot_face->resize((size_t)F);
  for (int f = 0; f < F; f++) (*slot_face)[(size_t)f] = (int32_t)(key[(size_t)f] & 0x7FFFFFFFu);
}

void PermuteRows(const float* src, int count, int width, const Reorder& ro, std::vector<float>* dst) {
  dst->resize((size_t)count * width);
  for (int v = 0; v < count; v++) {
    const size_t to = (size_t)ro.new_of_old[(size_t)v] * width;
    for (int k = 0; k < width; k++) (*dst)[to + k] = src[(size_t)v * width + k];
  }
}

// ---- profiling ---------------------------------------------------------------------------------
int DrainEvents(ugu_b200_rasterizer* r) {
  for (int s = 0; s < r->ev_used; s++) {
    CU_TRY(r, cudaEventSynchronize(r->ev_pool[2 * s + 1]));
    float ms = 0.f;
    CU_TRY(r, cudaEventElapsedTime(&ms, r->ev_pool[2 * s], r->ev_pool[2 * s + 1]));
    r->stats.stage_ms[r->ev_stage[s]] += ms;
  }
  r->ev_used = 0;
  return UGU_B200_OK;
}

struct StageTimer {
  ugu_b200_rasterizer* r;
  cudaStream_t st;
  int slot = -1;
  StageTimer(ugu_b200_rasterizer* r_, cudaStream_t st_, int stage) : r(r_), st(st_) {
    if (!r->profiling || !(r->stage_mask & (1u << stage))) return;
    if (r->ev_used >= kEventPool) DrainEvents(r);
    slot = r->ev_used++;
    r->ev_stage[slot] = stage;
    cudaEventRecord(r->ev_pool[2 * slot], st);
  }
  void Stop() {
    if (slot >= 0) cudaEventRecord(r->ev_pool[2 * slot + 1], st);
    slot = -1;
  }
  ~StageTimer() { Stop(); }
};

// ---- validation (mirrors ValidateAndInitBeforeRender, util_private.cc:12-67) --------------------
int Validate(ugu_b200_rasterizer* r, const ugu_b200_camera* cam, const ugu_b200_outputs* out) {
  if (cam == nullptr) return Fail(r, UGU_B200_E_INVALID, "camera has not been set");
  if (!r->mesh_ready) return Fail(r, UGU_B200_E_STATE, "mesh has not been initialized");
  const ugu_b200_option& o = r->opt;
  const MeshDev& m = r->mesh;
  if (o.backface_culling && m.num_face_normals == 0)
    return Fail(r, UGU_B200_E_INVALID, "specified back-face culling but face normal is empty.");
  if (o.diffuse_color == UGU_B200_COLOR_TEXTURE && m.num_materials == 0)
    return Fail(r, UGU_B200_E_INVALID, "specified texture as diffuse color but texture is empty.");
  if (o.diffuse_color == UGU_B200_COLOR_TEXTURE) {
    for (int i = 0; i < m.num_materials; i++)
      if (!r->tex_present[i])
        return Fail(r, UGU_B200_E_INVALID,
                    "specified texture as diffuse color but %d th texture is empty.", i);
    if (m.num_uv == 0 || !r->has_uvidx || !r->has_matids)
      return Fail(r, UGU_B200_E_INVALID,
                  "specified texture as diffuse color but uv / uv_indices / material_ids are empty.");
  }
  if (o.diffuse_color == UGU_B200_COLOR_VERTEX && m.num_colors == 0)
    return Fail(r, UGU_B200_E_INVALID,
                "specified vertex color as diffuse color but vertex color is empty.");
  if (o.shading_normal == UGU_B200_NORMAL_FACE && m.num_face_normals == 0)
    return Fail(r, UGU_B200_E_INVALID,
                "specified face normal as shading normal but face normal is empty.");
  if (o.shading_normal == UGU_B200_NORMAL_VERTEX && (m.num_normals == 0 || !r->has_nidx))
    return Fail(r, UGU_B200_E_INVALID,
                "specified vertex normal as shading normal but vertex normal is empty.");
  if (out == nullptr ||
      (!out->color && !out->depth && !out->normal && !out->mask && !out->face_id))
    return Fail(r, UGU_B200_E_INVALID, "all arguments are nullptr. nothing to do");
  if (cam->width <= 0 || cam->height <= 0 || cam->width > 65535 || cam->height > 65535 ||
      (long long)cam->width * cam->height >= (1ll << 31))
    return Fail(r, UGU_B200_E_INVALID, "camera size %dx%d is invalid", cam->width, cam->height);
  if (o.diffuse_color < 0 || o.diffuse_color > 2 || o.interp < 0 || o.interp > 1 ||
      o.shading_normal < 0 || o.shading_normal > 1 || o.diffuse_shading < 0 || o.diffuse_shading > 2)
    return Fail(r, UGU_B200_E_INVALID, "option enumerator out of range");
  return UGU_B200_OK;
}

// ---- per-view scratch --------------------------------------------------------------------------
// Slabs for `B` views of W x H: ray tables, visibility buffers (EMPTY between calls), projected vertices and
// the large-triangle queues.  A (re)sized visibility area is filled with EMPTY here; afterwards resolve_kernel
// leaves every slab it consumed EMPTY again.
int EnsureFrame(ugu_b200_rasterizer* r, int W, int H, int B, cudaStream_t st) {
  const size_t P = (size_t)W * H;
  const size_t Vpad = std::max<size_t>(((size_t)r->mesh.V + 3) & ~size_t(3), 4);
  const size_t Fq = std::max<size_t>((size_t)r->mesh.F, 1);
  int rc;
  if ((rc = Grow(r, &r->ray_dx, (size_t)B * W * 4))) return rc;
  if ((rc = Grow(r, &r->ray_dy, (size_t)B * H * 4))) return rc;
  if ((rc = Grow(r, &r->proj, (size_t)B * Vpad * sizeof(float4)))) return rc;
  if ((rc = Grow(r, &r->q_face, (size_t)B * Fq * 4))) return rc;
  if ((rc = Grow(r, &r->q_job, (size_t)B * Fq * 8))) return rc;
  if ((rc = Grow(r, &r->order, (size_t)B * (((size_t)r->mesh.F + 255) / 256 + 1) * 4))) return rc;
  {
    const size_t NT = (size_t)((W + kTile - 1) / kTile) * ((H + kTile - 1) / kTile);
    if (NT != r->tile_slab || (size_t)B > r->tile_slabs) {
      if ((rc = Grow(r, &r->tile_cnt, (size_t)B * NT * 4))) return rc;
      if ((rc = Grow(r, &r->tile_bin, (size_t)B * NT * kTileCap * 4))) return rc;
      CU_TRY(r, cudaMemsetAsync(r->tile_cnt.p, 0, (size_t)B * NT * 4, st));
      r->tile_slab = NT;
      r->tile_slabs = (size_t)B;
    }
  }
  if (P != r->vis_pixels || (size_t)B > r->vis_slabs) {
    // the slabs of one size are contiguous, so a new size or more slabs re-fills the whole area
    if ((rc = Grow(r, &r->vis, (size_t)B * P * 8))) return rc;
    const size_t n = (size_t)B * P;
    fill_u64<<<Blocks(n, 256), 256, 0, st>>>((unsigned long long*)r->vis.p, n, kVisEmpty);
    r->stats.kernel_launches++;
    CU_TRY(r, cudaGetLastError());
    r->vis_pixels = P;
    r->vis_slabs = (size_t)B;
  }
  return UGU_B200_OK;
}

void MakeViewParams(const ugu_b200_camera* cam, ViewParams* vp) {
  vp->W = cam->width;
  vp->H = cam->height;
  vp->fx = cam->fx;
  vp->fy = cam->fy;
  vp->cx = cam->cx;
  vp->cy = cam->cy;
  vp->wm1 = (float)(cam->width - 1);
  vp->hm1 = (float)(cam->height - 1);
  {  // facing_backface(): 64 u rho with rho = max row 1-norm of c2w_R (a hair of slack for its own rounding)
    float rho = 0.f;
    for (int i = 0; i < 3; i++)
      rho = std::max(rho, std::fabs(cam->c2w_R[3 * i]) + std::fabs(cam->c2w_R[3 * i + 1]) +
                              std::fabs(cam->c2w_R[3 * i + 2]));
    vp->facing_eps = 64.0f * 5.9604645e-8f * rho * 1.0001f;
    vp->inv_fx = 1.0f / cam->fx;
    vp->inv_fy = 1.0f / cam->fy;
  }
  // p / W for p < 2^31 as a multiply-high (Granlund-Montgomery, N = 31): l = ceil(log2 W),
  // magic = ceil(2^(31 + l) / W) < 2^32, quotient = umulhi(p, magic) >> (l - 1)
  vp->w_magic = 0;
  vp->w_shift = 0;
  if (cam->width > 1) {
    unsigned l = 0;
    while ((1ull << l) < (unsigned long long)cam->width) l++;
    const unsigned long long num = 1ull << (31 + l);
    vp->w_magic = (unsigned)((num + (unsigned long long)cam->width - 1) / (unsigned long long)cam->width);
    vp->w_shift = l - 1;
  }
  memcpy(vp->w2c_R, cam->w2c_R, sizeof(vp->w2c_R));
  memcpy(vp->w2c_t, cam->w2c_t, sizeof(vp->w2c_t));
  memcpy(vp->c2w_R, cam->c2w_R, sizeof(vp->c2w_R));
  for (int i = 0; i < 3; i++)  // camera centre (ordering heuristic only: exactness is irrelevant)
    vp->eye[i] = -(cam->c2w_R[3 * i] * cam->w2c_t[0] + cam->c2w_R[3 * i + 1] * cam->w2c_t[1] +
                   cam->c2w_R[3 * i + 2] * cam->w2c_t[2]);
}

int g_num_sms = 148;
int g_binned_ctas_per_sm = 8;
constexpr int kTwoPassMinFaces = 200000;
int g_ordered_minb = 5;  // CTAs/SM of the ordered raster kernel (UGU_B200_RASTER_MINB: 5 = 48 registers, 6 = 40 + spills)

// Launch with programmatic stream serialization (PDL): the kernel may be scheduled into the tail of
// its predecessor on the stream; it synchronises on the predecessor's results itself (pdl_wait()).
template <typename... KArgs, typename... Args>
cudaError_t LaunchView(void (*kernel)(KArgs...), dim3 grid, unsigned block, cudaStream_t st, bool pdl,
                       Args&&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = dim3(block);
  cfg.dynamicSmemBytes = 0;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
}

// Cross-stream ordering of the handle's single set of scratch buffers (proj / vis / queues / counters / ray
// tables ...): every launch set records an event; a call on a DIFFERENT stream first waits for it.
int AcquireScratch(ugu_b200_rasterizer* r, cudaStream_t st) {
  if (r->scratch_used && r->last_stream != st) CU_TRY(r, cudaStreamWaitEvent(st, r->ev_scratch, 0));
  return UGU_B200_OK;
}
int ReleaseScratch(ugu_b200_rasterizer* r, cudaStream_t st) {
  CU_TRY(r, cudaEventRecord(r->ev_scratch, st));
  r->last_stream = st;
  r->scratch_used = true;
  return UGU_B200_OK;
}

FrameDev MakeFrame(const ugu_b200_rasterizer* r, int W, int H) {
  FrameDev fr;
  fr.proj = (float4*)r->proj.p;
  fr.ray_dx = (float*)r->ray_dx.p;
  fr.ray_dy = (float*)r->ray_dy.p;
  fr.vis = (unsigned long long*)r->vis.p;
  fr.counters = (unsigned long long*)r->counters.p;
  fr.host_counters = (r->counters_mapped && r->host_snapshot) ? r->counters_host : nullptr;
  fr.q_face = (int*)r->q_face.p;
  fr.q_job = (unsigned long long*)r->q_job.p;
  fr.coop_max = 0;
  fr.order = nullptr;
  fr.tile_cnt = nullptr;
  fr.tile_bin = nullptr;
  fr.tiles_x = (W + kTile - 1) / kTile;
  fr.s_tiles = (size_t)((W + kTile - 1) / kTile) * ((H + kTile - 1) / kTile);
  fr.s_order = ((size_t)r->mesh.F + 255) / 256;
  fr.s_proj = std::max<size_t>(((size_t)r->mesh.V + 3) & ~size_t(3), 4);
  fr.s_ray_dx = (size_t)W;
  fr.s_ray_dy = (size_t)H;
  fr.s_vis = (size_t)W * H;
  fr.s_q = std::max<size_t>((size_t)r->mesh.F, 1);
  return fr;
}

ShadeParams MakeShade(const ugu_b200_rasterizer* r, const ugu_b200_outputs* dout, size_t view_pixels,
                      unsigned first_view = 0) {
  ShadeParams sp;
  sp.chunk_begin = sp.chunk_end = 0u;
  sp.diffuse_color = r->opt.diffuse_color;
  sp.interp = r->opt.interp;
  sp.shading_normal = r->opt.shading_normal;
  sp.diffuse_shading = r->opt.diffuse_shading;
  sp.backface_culling = r->opt.backface_culling;
  {  // OrenNayarParam (pixel_shader.h:143-148)
    const float sigma = r->opt.oren_nayar_sigma;
    const float s2 = sigma * sigma;
    sp.on_A = 1.0f - (0.5f * s2 / (s2 + 0.33f));
    sp.on_B = 0.45f * s2 / (s2 + 0.09f);
  }
  sp.color = dout->color;
  sp.depth = dout->depth;
  sp.normal = dout->normal;
  sp.mask = dout->mask;
  sp.face_id = dout->face_id;
  sp.view_pixels = view_pixels;
  sp.first_view = first_view;
  sp.fold_binned = 0;
  return sp;
}

// Views per launch set for a batch of `n` views: as many as the scratch budget allows, at most kMaxBatch.
int BatchViews(const ugu_b200_rasterizer* r, int n, int W, int H) {
  if (n <= 1) return 1;
  const size_t P = (size_t)W * H;
  const size_t per_view = (((size_t)r->mesh.V + 3) & ~size_t(3)) * 16 + P * 8 + (size_t)r->mesh.F * 12 + ((size_t)W + H) * 4;
  const size_t budget = (size_t)3 << 29;  // 1.5 GiB of per-view scratch
  int b = (int)std::max<size_t>(1, std::min<size_t>(kMaxBatch, budget / std::max<size_t>(per_view, 1)));
  if (r->max_batch > 0) b = std::min(b, r->max_batch);
  return std::min(b, n);
}

// One launch set (vertex, raster, tile jobs, resolve) for `n` <= kMaxBatch views of one image size into
// device outputs on stream `st`; view v writes at element offset v * W*H of every requested output.
// Resolve of a single view in bands (host-form Render): band b covers the 32-pixel chunks [split[b-1], split[b])
// (split[-1] = 0, the last band runs to the end of the image) and records ev[b] when its outputs are complete.
struct BandPlan {
  int n = 1;
  unsigned split[3] = {0, 0, 0};
  cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
};

template <int NV>
int LaunchViewSet(ugu_b200_rasterizer* r, const ugu_b200_camera* cams, int n, const ugu_b200_outputs* dout,
                  unsigned first_view, cudaStream_t st, const GBufParams* gbuf, const BandPlan* bands = nullptr) {
  const int W = cams[0].width, H = cams[0].height;
  int rc = EnsureFrame(r, W, H, n, st);
  if (rc) return rc;
  ViewBatch<NV> vb;
  for (int v = 0; v < NV; v++) MakeViewParams(&cams[std::min(v, n - 1)], &vb.v[v]);
  FrameDev fr = MakeFrame(r, W, H);
  // Triangles up to coop_max bbox samples are rasterised inside raster_kernel (one warp per 32
  // faces); larger ones go to the globally load-balanced tile-job queue.  The in-kernel path is the
  // cheaper one per triangle but only as parallel as the face count: with fewer raster warps than
  // the machine holds (F < ~600k) mid-size triangles are better spread over the whole GPU by the
  // queue (measured: config 4, 10M faces, 1.17 -> 0.93 ms with 32; spot, 5.9k faces, 33 -> 38 us).
  fr.coop_max = r->coop_max > 0 ? r->coop_max : (r->mesh.F >= 4 * 32 * 32 * g_num_sms ? 32 : 16);
  // Occlusion ordering of raster_kernel (raster_kernels.cuh): camera-facing blocks first, the others after them
  // with the triangles dropped that are provably behind what is already stored.  Pays once the face warps
  // outnumber what the machine holds at a time (the ordered kernel cannot prefetch its indices under the
  // predecessor's tail).
  // Per-tile triangle lists (tile_kernel) for the triangles too large for the cooperative path (the tile-job queue
  // stays for what overlaps more than 3 x 3 tiles or finds a list full).  OFF unless UGU_B200_TILES=1: measured
  // against the tile-job queue it loses on the configs it was built for (spot 1024^2, 15.3 M warp
  // instructions + 2.5 M left in the queue against 11.8 M; 35.7 + 9.9 us against 27.8 us) — a 16 x 16 tile lists ~30
  // of spot's triangles while ~2 cover any one pixel, so walking every listed triangle at every pixel costs more
  // than walking each triangle's own bbox.
  const bool use_tiles = r->tiles > 0;
  if (use_tiles) {
    fr.tile_cnt = (int*)r->tile_cnt.p;
    fr.tile_bin = (int*)r->tile_bin.p;
  }
  const bool two_launch = r->mesh.grp != nullptr && r->two_pass == 2;
  if (!two_launch && r->mesh.grp != nullptr && (r->two_pass >= 0 ? r->two_pass != 0 : r->mesh.F >= kTwoPassMinFaces))
    fr.order = (int*)r->order.p;
  const MeshDev& m = r->mesh;
  const size_t P = (size_t)W * H;
  ShadeParams sp = MakeShade(r, dout, P, first_view);
  // Meshes of small triangles rarely queue anything: no binned_kernel launch (it would read one counter and exit,
  // ~3 us of every view); the resolve kernel rasterises whatever was queued behind a grid barrier (raster_kernels.cuh).
  // Not with the G-buffer kernel in between, which needs the finished visibility buffer.
  // Measured (profiles/r03b_fold_binned_ab.jsonl): config 2 74.1 -> 72.9 us, config 5 38.9 -> 37.7 us per single view;
  // a batch amortises the launch over its views (0.4 us per view) and keeps it.
  // (not the texture instantiations: they sit at the register limit, and carrying the tile-job code costs their hot
  //  loop more than the launch — config 4 741 -> 756 us per view)
  const bool fold_binned = gbuf == nullptr && r->tiles <= 0 && bands == nullptr &&
                           (r->fold_binned >= 0 ? r->fold_binned != 0
                                                : (NV == 1 && r->mesh.F >= kTwoPassMinFaces &&
                                                   !(dout->color != nullptr && r->opt.diffuse_color == UGU_B200_COLOR_TEXTURE)));
  sp.fold_binned = fold_binned ? 1 : 0;
  const bool pdl = r->use_pdl;
  const unsigned ny = (unsigned)n;
  // Fused projection (block-local vertex tables, upload_mesh): raster_kernel projects its own vertices, so no vertex
  // kernel runs — only the block lists of the occlusion ordering are built ahead of it, by classify_kernel.
  const bool fused = m.lv != nullptr;
  {
    StageTimer t(r, st, UGU_B200_STAGE_VERTEX);
    if (!fused) {
      const size_t threads = std::max<size_t>(std::max<size_t>((size_t)m.V, (size_t)W + H), ((size_t)m.F + 255) / 256);
      CU_TRY(r, LaunchView(vertex_kernel_b<NV>, dim3(Blocks(threads, 256)), 256, st, pdl, vb, n, m, fr));
    } else if (fr.order) {
      CU_TRY(r, LaunchView(classify_kernel<NV>, dim3(Blocks(((size_t)m.F + 255) / 256, 256)), 256, st, pdl, vb, n, m, fr));
    }
  }
  if (m.F > 0) {
    StageTimer t(r, st, UGU_B200_STAGE_RASTER);
    // 6 resident CTAs per SM (40 registers, no spills) measured 2-4 % faster than 4 on every config
    const dim3 grid(Blocks((size_t)m.F, 8 * 32), ny);
    if (two_launch) {
      CU_TRY(r, LaunchView(raster_kernel<1, 6, NV, 2>, grid, 256, st, pdl, vb, m, fr));
      CU_TRY(r, LaunchView(raster_kernel<1, 5, NV, 3>, grid, 256, st, pdl, vb, m, fr));
      r->stats.kernel_launches++;
    } else if (fr.order && use_tiles) CU_TRY(r, LaunchView(raster_kernel<1, 4, NV, 1, true>, grid, 256, st, pdl, vb, m, fr));
    else if (fr.order && g_ordered_minb == 6) CU_TRY(r, LaunchView(raster_kernel<1, 6, NV, 1>, grid, 256, st, pdl, vb, m, fr));
    else if (fr.order) CU_TRY(r, LaunchView(raster_kernel<1, 5, NV, 1>, grid, 256, st, pdl, vb, m, fr));
    else if (use_tiles) CU_TRY(r, LaunchView(raster_kernel<1, 4, NV, 0, true>, grid, 256, st, pdl, vb, m, fr));
    else CU_TRY(r, LaunchView(raster_kernel<1, 6, NV, 0>, grid, 256, st, pdl, vb, m, fr));
  }
  if (m.F > 0 && !fold_binned) {
    StageTimer t(r, st, UGU_B200_STAGE_BINNED);
    if (use_tiles && !two_launch) {
      CU_TRY(r, LaunchView(tile_kernel<NV>, dim3((unsigned)fr.s_tiles, ny), 256, st, pdl, vb, m, fr));
      r->stats.kernel_launches++;
    }
    // persistent: exactly one resident wave, shared by the views of the batch
    const unsigned resident = (unsigned)(g_num_sms * g_binned_ctas_per_sm);
    CU_TRY(r, LaunchView(binned_kernel<NV>, dim3(std::max(1u, (resident + ny - 1) / ny), ny), 128, st, pdl, vb, m, fr));
  }
  if (gbuf) {
    // G-buffer fields from the finished visibility buffer, before the resolve kernel consumes and
    // clears it (both wait for their predecessor's completion before they touch anything)
    GBufParams gp = *gbuf;
    gp.backface_culling = sp.backface_culling;
    gp.have_normals = (m.num_normals > 0 && r->has_nidx) ? 1 : 0;
    gp.have_uv = (m.num_uv > 0 && r->has_uvidx) ? 1 : 0;
    const unsigned chunks = (unsigned)((P + 31) / 32);
    const unsigned grid = std::max(1u, std::min((unsigned)(g_num_sms * 8), (chunks + 7) / 8));
    CU_TRY(r, LaunchView(gbuffer_kernel, dim3(grid), 256, st, pdl, vb.v[0], m, fr, gp));
    r->stats.kernel_launches++;
  }
  const int nbands = (NV == 1 && bands) ? bands->n : 1;
  StageTimer resolve_timer(r, st, UGU_B200_STAGE_RESOLVE);
  for (int band = 0; band < nbands; band++) {
    // persistent: enough CTAs to fill the machine, never more than there are 32-pixel chunks
    const unsigned all_chunks = (unsigned)((P + 31) / 32);
    sp.chunk_begin = band > 0 ? bands->split[band - 1] : 0u;
    sp.chunk_end = band + 1 < nbands ? bands->split[band] : 0u;
    const unsigned chunks = (sp.chunk_end ? sp.chunk_end : all_chunks) - sp.chunk_begin;
    const bool need_color = sp.color != nullptr;
    const bool need_nrm = sp.normal != nullptr || (need_color && sp.diffuse_shading != 0);
    const int color_mode = need_color ? sp.diffuse_color : 0;
    const int nmode = !need_nrm ? 0 : (sp.shading_normal == 0 ? 1 : (m.vattr ? 3 : 2));
    // (folded tile jobs: the grid barrier needs every CTA resident at once — round down, never up)
    const unsigned per_view = fold_binned ? (unsigned)(g_num_sms * 4) / ny : ((unsigned)(g_num_sms * 4) + ny - 1) / ny;
    const unsigned gx = std::max(1u, std::min(per_view, (chunks + 7) / 8));
    const dim3 grid(gx, ny);
    // (the one-chunk-deeper variant <.., true>, which also prefetches the projected vertices, measured
    //  slower on every config: its extra live registers spill at 4 CTAs/SM — see DESIGN.md)
#define UGU_RESOLVE_CASE(C, N)                                                                      \
  case (C)*4 + (N):                                                                                 \
    if (fold_binned) CU_TRY(r, LaunchView(resolve_kernel<C, N, 4, false, NV, true>, grid, 256, st, pdl, vb, m, fr, sp)); \
    else CU_TRY(r, LaunchView(resolve_kernel<C, N, 4, false, NV, false>, grid, 256, st, pdl, vb, m, fr, sp)); \
    break;
    switch (color_mode * 4 + nmode) {
      UGU_RESOLVE_CASE(0, 0)
      UGU_RESOLVE_CASE(0, 1)
      UGU_RESOLVE_CASE(0, 2)
      UGU_RESOLVE_CASE(0, 3)
      UGU_RESOLVE_CASE(1, 0)
      UGU_RESOLVE_CASE(1, 1)
      UGU_RESOLVE_CASE(1, 2)
      UGU_RESOLVE_CASE(1, 3)
      UGU_RESOLVE_CASE(2, 0)
      UGU_RESOLVE_CASE(2, 1)
      UGU_RESOLVE_CASE(2, 2)
      UGU_RESOLVE_CASE(2, 3)
      default:
        return Fail(r, UGU_B200_E_INVALID, "option enumerator out of range");
    }
#undef UGU_RESOLVE_CASE
    if (nbands > 1) CU_TRY(r, cudaEventRecord(bands->ev[band], st));
  }
  resolve_timer.Stop();
  CU_TRY(r, cudaGetLastError());
  r->stats.kernel_launches += ((m.F > 0) ? (fold_binned ? 3 : 4) : 2) - ((fused && !fr.order) ? 1 : 0) + (nbands - 1);
  r->stats.views += (uint64_t)n;
  if (r->profiling) r->stats.timed_views += (uint64_t)n;
  r->stats.h2d_bytes += sizeof(ViewParams) * (size_t)n;
  return UGU_B200_OK;
}

// `n` views of one image size into device outputs (view i at element offset i * W*H), cut into launch sets
// of at most `batch` views.
int RenderViews(ugu_b200_rasterizer* r, const ugu_b200_camera* cams, int n, const ugu_b200_outputs* dout,
                cudaStream_t st, int batch) {
  int rc = AcquireScratch(r, st);
  if (rc) return rc;
  for (int i = 0; i < n; i += batch) {
    const int k = std::min(batch, n - i);
    if (k == 1 && i == 0) rc = LaunchViewSet<1>(r, cams, 1, dout, 0u, st, nullptr);
    else rc = LaunchViewSet<kMaxBatch>(r, cams + i, k, dout, (unsigned)i, st, nullptr);
    if (rc) return rc;
  }
  return ReleaseScratch(r, st);
}

// One view (optionally with the G-buffer kernel between the tile-job and the resolve kernel).
int RenderView(ugu_b200_rasterizer* r, const ugu_b200_camera* cam, const ugu_b200_outputs* dout,
               cudaStream_t st, const GBufParams* gbuf = nullptr, const BandPlan* bands = nullptr) {
  int rc = AcquireScratch(r, st);
  if (rc) return rc;
  if ((rc = LaunchViewSet<1>(r, cam, 1, dout, 0u, st, gbuf, bands))) return rc;
  return ReleaseScratch(r, st);
}

bool IsPinnedHost(const void* p) {
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return a.type == cudaMemoryTypeHost;
}

bool DestPinned(ugu_b200_rasterizer* r, const void* p) {
  for (const auto& e : r->pin_cache)
    if (e.p == p) return e.pinned;
  const bool pinned = IsPinnedHost(p);
  r->pin_cache[r->pin_next] = {p, pinned};
  r->pin_next = (r->pin_next + 1) % 16;
  return pinned;
}

int EnsureInternalOutputs(ugu_b200_rasterizer* r, int set, const ugu_b200_outputs* want, size_t P,
                          ugu_b200_outputs* dout) {
  int rc;
  memset(dout, 0, sizeof(*dout));
  if (want->color) {
    if ((rc = Grow(r, &r->out_color[set], (P * 3 + 15) & ~size_t(15)))) return rc;
    dout->color = (uint8_t*)r->out_color[set].p;
  }
  if (want->depth) {
    if ((rc = Grow(r, &r->out_depth[set], P * 4 + 16))) return rc;
    dout->depth = (float*)r->out_depth[set].p;
  }
  if (want->normal) {
    if ((rc = Grow(r, &r->out_normal[set], P * 12 + 16))) return rc;
    dout->normal = (float*)r->out_normal[set].p;
  }
  if (want->mask) {
    if ((rc = Grow(r, &r->out_mask[set], P + 16))) return rc;
    dout->mask = (uint8_t*)r->out_mask[set].p;
  }
  if (want->face_id) {
    if ((rc = Grow(r, &r->out_fid[set], P * 4 + 16))) return rc;
    dout->face_id = (int32_t*)r->out_fid[set].p;
  }
  return UGU_B200_OK;
}

struct CopySeg {
  void* dst;
  const void* src;
  size_t bytes;
};

// `count` consecutive views starting at host view index `view` <- the first `count` view slabs of `dev`
int CollectSegs(const ugu_b200_outputs* host, const ugu_b200_outputs* dev, size_t P, size_t view, size_t count,
                std::vector<CopySeg>* segs) {
  const size_t n = P * count;
  if (host->color) segs->push_back({host->color + view * P * 3, dev->color, n * 3});
  if (host->depth) segs->push_back({host->depth + view * P, dev->depth, n * 4});
  if (host->normal) segs->push_back({host->normal + view * P * 3, dev->normal, n * 12});
  if (host->mask) segs->push_back({host->mask + view * P, dev->mask, n});
  if (host->face_id) segs->push_back({host->face_id + view * P, dev->face_id, n * 4});
  return UGU_B200_OK;
}

// Device -> host copy of the segments on `st`.  Pinned destinations: one async DMA each.
// Pageable destinations: chunked through the pinned staging buffer, the host memcpy of chunk k
// overlapping the DMA of chunk k+1.  Returns after everything has landed in the destinations.
int CopyOut(ugu_b200_rasterizer* r, const std::vector<CopySeg>& segs, cudaStream_t st) {
  std::vector<CopySeg> pageable;
  size_t pageable_bytes = 0;
  for (const CopySeg& s : segs) {
    r->stats.d2h_bytes += s.bytes;
    if (DestPinned(r, s.dst)) {
      CU_TRY(r, cudaMemcpyAsync(s.dst, s.src, s.bytes, cudaMemcpyDeviceToHost, st));
    } else {
      pageable.push_back(s);
      pageable_bytes += s.bytes;
    }
  }
  if (!pageable.empty()) {
    if (r->staging_bytes < pageable_bytes) {
      if (r->staging) cudaFreeHost(r->staging);
      r->staging = nullptr;
      r->staging_bytes = 0;
      CU_TRY(r, cudaHostAlloc(&r->staging, pageable_bytes, cudaHostAllocDefault));
      r->staging_bytes = pageable_bytes;
    }
    if (!r->copy_pool) {
      int n = r->copy_threads;
      if (n <= 0) {  // 3/8 of the host threads, at most 6 (measured on a 16-thread B200 host, config 2's 49.8 MB per
                     // view: 1 thread 6.5 ms, 2 3.3, 3 2.2, 4 1.57, 6 1.37, 8 1.40; pinned destinations 1.01)
        const unsigned hw = std::thread::hardware_concurrency();
        n = (int)std::min(6u, std::max(1u, hw * 3u / 8u));
      }
      r->copy_threads = n;
      r->copy_pool = new HostCopyPool(r->device, n - 1);
    }
    const size_t chunk_bytes =
        std::max(kStageChunk, ((pageable_bytes / kStageChunksMax) + 4095) & ~size_t(4095));
    std::vector<HostChunk> chunks;
    size_t off = 0;
    for (const CopySeg& s : pageable) {
      for (size_t o = 0; o < s.bytes; o += chunk_bytes) {
        const size_t n = std::min(chunk_bytes, s.bytes - o);
        char* stage = (char*)r->staging + off;
        CU_TRY(r, cudaMemcpyAsync(stage, (const char*)s.src + o, n, cudaMemcpyDeviceToHost, st));
        if (chunks.size() >= r->chunk_events.size()) {
          cudaEvent_t e;
          CU_TRY(r, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
          r->chunk_events.push_back(e);
        }
        CU_TRY(r, cudaEventRecord(r->chunk_events[chunks.size()], st));
        chunks.push_back({(char*)s.dst + o, stage, n, r->chunk_events[chunks.size()]});
        off += n;
      }
    }
    CU_TRY(r, r->copy_pool->Run(chunks));
  }
  CU_TRY(r, cudaStreamSynchronize(st));
  return UGU_B200_OK;
}

int SnapshotCounters(ugu_b200_rasterizer* r) {
  // slab 0 words [4..6]: the last single view's (queue, hazards) as saved by resolve_kernel, and the
  // running hazard total over every view of every batch (reported as the delta since the last snapshot)
  const unsigned long long packed = r->counters_host[0];
  r->stats.large_triangles = packed >> kQueueJobBits;
  r->stats.tile_jobs = packed & kQueueJobMask;
  const unsigned long long total = r->counters_host[2];
  r->stats.hazard_fragments = total - r->hazard_seen;
  r->hazard_seen = total;
  if (r->stats.hazard_fragments >= kQueueOverflowFlag)
    return Fail(r, UGU_B200_E_INVALID, "tile-job queue overflow: more than 2^%d tile jobs in one view", kQueueJobBits);
  return UGU_B200_OK;
}

}  // namespace

// =================================================================================================
extern "C" {

int ugu_b200_abi_version(void) { return UGU_B200_ABI_VERSION; }

int ugu_b200_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  int usable = 0;
  for (int i = 0; i < n; i++) {
    cudaDeviceProp p;
    if (cudaGetDeviceProperties(&p, i) == cudaSuccess && p.major == 10) usable++;
  }
  return usable;
}

const char* ugu_b200_last_error(const ugu_b200_rasterizer* r) {
  return r ? r->error.c_str() : g_last_error.c_str();
}

int ugu_b200_rasterizer_create(int device, ugu_b200_rasterizer** out) {
  if (out == nullptr) return Fail(nullptr, UGU_B200_E_INVALID, "out is NULL");
  *out = nullptr;
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0) {
    cudaGetLastError();
    return Fail(nullptr, UGU_B200_E_CUDA,
                "no CUDA device available (%s); ugu_b200 has no CPU fallback",
                e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
  }
  if (device < 0 || device >= n) return Fail(nullptr, UGU_B200_E_INVALID, "device %d out of range", device);
  cudaDeviceProp prop;
  CU_TRY(nullptr, cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10)
    return Fail(nullptr, UGU_B200_E_CUDA,
                "device %d is sm_%d%d; this library carries sm_100a code only", device, prop.major,
                prop.minor);
  CU_TRY(nullptr, cudaSetDevice(device));
  g_num_sms = prop.multiProcessorCount;
  {
    int n_ctas = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n_ctas, binned_kernel<1>, 128, 0) == cudaSuccess && n_ctas > 0)
      g_binned_ctas_per_sm = n_ctas;
    if (const char* e = getenv("UGU_B200_BINNED_CTAS")) g_binned_ctas_per_sm = std::max(1, atoi(e));
  }
  ugu_b200_rasterizer* r = new ugu_b200_rasterizer;
  r->device = device;
  r->opt.diffuse_color = UGU_B200_COLOR_NONE;
  r->opt.interp = UGU_B200_INTERP_BILINEAR;
  r->opt.shading_normal = UGU_B200_NORMAL_VERTEX;
  r->opt.diffuse_shading = UGU_B200_SHADING_NONE;
  r->opt.backface_culling = 1;
  r->opt.depth_scale = 1.0f;
  r->opt.oren_nayar_sigma = 0.3f;
  if (const char* e = getenv("UGU_B200_COOP_MAX")) {  // tuning knob, clamped to the kernel's limit
    r->coop_max = std::max(1, std::min(atoi(e), kCoopMaxSamplesLimit));
  }
  if (const char* e = getenv("UGU_B200_PDL")) r->use_pdl = atoi(e) != 0;
  if (const char* e = getenv("UGU_B200_BANDS")) r->bands = atoi(e);
  if (const char* e = getenv("UGU_B200_TWOPASS")) r->two_pass = std::max(0, std::min(2, atoi(e)));
  if (const char* e = getenv("UGU_B200_FOLD_BINNED")) r->fold_binned = atoi(e) != 0 ? 1 : 0;
  if (const char* e = getenv("UGU_B200_TILES")) r->tiles = atoi(e) != 0 ? 1 : 0;
  if (const char* e = getenv("UGU_B200_RASTER_MINB")) g_ordered_minb = atoi(e) == 6 ? 6 : 5;
  if (const char* e = getenv("UGU_B200_COPY_THREADS")) r->copy_threads = std::max(1, std::min(atoi(e), 64));
  if (const char* e = getenv("UGU_B200_BATCH")) r->max_batch = std::max(1, std::min(atoi(e), (int)kMaxBatch));
  auto fail = [&](cudaError_t err, const char* what) {
    int rc = Fail(nullptr, UGU_B200_E_CUDA, "%s failed: %s", what, cudaGetErrorString(err));
    ugu_b200_rasterizer_destroy(r);
    return rc;
  };
  if ((e = cudaStreamCreateWithFlags(&r->stream, cudaStreamNonBlocking)) != cudaSuccess)
    return fail(e, "cudaStreamCreate");
  if ((e = cudaStreamCreateWithFlags(&r->copy_stream, cudaStreamNonBlocking)) != cudaSuccess)
    return fail(e, "cudaStreamCreate");
  for (int i = 0; i < 2; i++) {
    if ((e = cudaEventCreateWithFlags(&r->ev_render[i], cudaEventDisableTiming)) != cudaSuccess)
      return fail(e, "cudaEventCreate");
    if ((e = cudaEventCreateWithFlags(&r->ev_copied[i], cudaEventDisableTiming)) != cudaSuccess)
      return fail(e, "cudaEventCreate");
  }
  r->ev_pool.resize(2 * kEventPool, nullptr);
  r->ev_stage.resize(kEventPool, 0);
  for (auto& ev : r->ev_pool)
    if ((e = cudaEventCreate(&ev)) != cudaSuccess) return fail(e, "cudaEventCreate");
  const size_t counter_bytes = (size_t)kMaxBatch * kCounterSlab * 8;
  if ((e = cudaMalloc(&r->counters.p, counter_bytes)) != cudaSuccess) return fail(e, "cudaMalloc");
  r->counters.bytes = counter_bytes;
  if ((e = cudaMemset(r->counters.p, 0, counter_bytes)) != cudaSuccess) return fail(e, "cudaMemset");
  if ((e = cudaEventCreateWithFlags(&r->ev_scratch, cudaEventDisableTiming)) != cudaSuccess)
    return fail(e, "cudaEventCreate");
  if ((e = cudaHostAlloc((void**)&r->counters_host, 64, cudaHostAllocDefault)) != cudaSuccess)
    return fail(e, "cudaHostAlloc");
  memset(r->counters_host, 0, 64);
  {  // unified addressing: the pinned block is reachable from the device under the same pointer
    int uva = 0;
    cudaDeviceGetAttribute(&uva, cudaDevAttrUnifiedAddressing, device);
    r->counters_mapped = uva != 0 && !(getenv("UGU_B200_MAPPED_COUNTERS") && atoi(getenv("UGU_B200_MAPPED_COUNTERS")) == 0);
  }
  *out = r;
  return UGU_B200_OK;
}

void ugu_b200_rasterizer_destroy(ugu_b200_rasterizer* r) {
  if (!r) return;
  cudaSetDevice(r->device);
  if (r->stream) cudaStreamSynchronize(r->stream);
  if (r->copy_stream) cudaStreamSynchronize(r->copy_stream);
  FreeMesh(r);
  DevBuf* bufs[] = {&r->proj, &r->ray_dx, &r->ray_dy, &r->vis, &r->q_face, &r->q_job, &r->counters, &r->order, &r->tile_cnt, &r->tile_bin,
                    &r->uv_win, &r->uv_counter, &r->uv_q_face, &r->uv_q_job, &r->uv_rec, &r->uv_in, &r->uv_tex, &r->uv_mask,
                    &r->vis_tmp, &r->vis_in, &r->vis_acc,
                    &r->vt_pos, &r->vt_nrm, &r->vt_idx, &r->vt_acc, &r->vt_owner, &r->vt_vvis, &r->vt_vrec,
                    &r->vt_fvis, &r->vt_frec, &r->vt_img, &r->vt_zbuf,
                    &r->d2m_in_depth, &r->d2m_in_color, &r->d2m_small, &r->d2m_block_sums, &r->d2m_vid,
                    &r->d2m_fref, &r->d2m_vertices, &r->d2m_faces, &r->d2m_normals, &r->d2m_face_normals,
                    &r->d2m_uv, &r->d2m_colors, &r->d2m_pc, &r->d2m_nimg};
  if (r->d2m_totals_host) cudaFreeHost(r->d2m_totals_host);
  for (DevBuf* b : bufs) FreeBuf(b);
  for (int i = 0; i < 2; i++) {
    FreeBuf(&r->out_color[i]);
    FreeBuf(&r->out_depth[i]);
    FreeBuf(&r->out_normal[i]);
    FreeBuf(&r->out_mask[i]);
    FreeBuf(&r->out_fid[i]);
    if (r->ev_render[i]) cudaEventDestroy(r->ev_render[i]);
    if (r->ev_copied[i]) cudaEventDestroy(r->ev_copied[i]);
  }
  FreeBuf(&r->gbuf_arena);
  FreeBuf(&r->rt_rays);
  for (auto ev : r->ev_pool)
    if (ev) cudaEventDestroy(ev);
  for (auto ev : r->chunk_events) cudaEventDestroy(ev);
  delete r->copy_pool;
  if (r->staging) cudaFreeHost(r->staging);
  if (r->counters_host) cudaFreeHost(r->counters_host);
  if (r->ev_scratch) cudaEventDestroy(r->ev_scratch);
  if (r->stream) cudaStreamDestroy(r->stream);
  if (r->copy_stream) cudaStreamDestroy(r->copy_stream);
  for (cudaEvent_t& e : r->ev_band)
    if (e) cudaEventDestroy(e);
  delete r;
}

int ugu_b200_rasterizer_set_option(ugu_b200_rasterizer* r, const ugu_b200_option* opt) {
  if (!r || !opt) return Fail(r, UGU_B200_E_INVALID, "NULL argument");
  r->opt = *opt;
  return UGU_B200_OK;
}

int ugu_b200_rasterizer_upload_mesh(ugu_b200_rasterizer* r, const ugu_b200_mesh_desc* d) {
  if (!r || !d) return Fail(r, UGU_B200_E_INVALID, "NULL argument");
  CU_TRY(r, cudaSetDevice(r->device));
  CU_TRY(r, cudaStreamSynchronize(r->stream));
  FreeMesh(r);
  const int V = d->num_vertices, F = d->num_faces;
  if (V < 0 || F < 0 || d->num_normals < 0 || d->num_colors < 0 || d->num_face_normals < 0 ||
      d->num_uv < 0 || d->num_materials < 0)
    return Fail(r, UGU_B200_E_INVALID, "negative element count");
  if ((unsigned long long)F >= (1ull << (64 - kQueueJobBits)))
    return Fail(r, UGU_B200_E_INVALID, "more than %llu faces are not supported",
                (1ull << (64 - kQueueJobBits)) - 1);
  if ((V > 0 && !d->vertices) || (F > 0 && !d->vertex_indices))
    return Fail(r, UGU_B200_E_INVALID, "vertices / vertex_indices are NULL");
  if (F > 0 && !IndicesInRange(d->vertex_indices, (size_t)F * 3, V))
    return Fail(r, UGU_B200_E_INVALID, "vertex_indices reference a vertex outside [0, %d)", V);
  if (d->num_face_normals > 0 && d->num_face_normals < F)
    return Fail(r, UGU_B200_E_INVALID, "face_normals has %d entries for %d faces", d->num_face_normals, F);
  if (d->num_colors > 0 && d->num_colors < V)
    return Fail(r, UGU_B200_E_INVALID, "vertex_colors has %d entries for %d vertices", d->num_colors, V);
  if (d->normal_indices && F > 0 && !IndicesInRange(d->normal_indices, (size_t)F * 3, d->num_normals))
    return Fail(r, UGU_B200_E_INVALID, "normal_indices reference a normal outside [0, %d)", d->num_normals);
  if (d->uv_indices && F > 0 && !IndicesInRange(d->uv_indices, (size_t)F * 3, d->num_uv))
    return Fail(r, UGU_B200_E_INVALID, "uv_indices reference a uv outside [0, %d)", d->num_uv);
  if (d->material_ids && F > 0 && d->num_materials > 0 &&
      !IndicesInRange(d->material_ids, (size_t)F, d->num_materials))
    return Fail(r, UGU_B200_E_INVALID, "material_ids reference a material outside [0, %d)", d->num_materials);

  // host-side first-use reordering (UGU_B200_REORDER=0 disables it, for A/B measurements)
  const bool reorder = !(getenv("UGU_B200_REORDER") && atoi(getenv("UGU_B200_REORDER")) == 0);
  Reorder ro_v, ro_n, ro_uv;
  std::vector<float> pos_p, col_p, nrm_p, uv_p;
  const float* h_vertices = d->vertices;
  {
    float mx = 0.f;
    for (size_t i = 0; i < (size_t)V * 3; i++) {
      const float a = fabsf(d->vertices[i]);
      if (a > mx) mx = a;  // NaNs are skipped
    }
    r->mesh_max_abs = mx;
  }
  const float* h_colors = d->vertex_colors;
  const float* h_normals = d->normals;
  const float* h_uv = d->uv;
  const int32_t* h_vidx = d->vertex_indices;
  const int32_t* h_nidx = d->normal_indices;
  const int32_t* h_uvidx = d->uv_indices;
  // raster order of the faces (UGU_B200_FACE_ORDER=0: the order of the file)
  std::vector<int32_t> slot_face;
  if (F > 0 && V > 0 && !(getenv("UGU_B200_FACE_ORDER") && atoi(getenv("UGU_B200_FACE_ORDER")) == 0))
    BuildFaceOrder(d->vertices, d->vertex_indices, F, &slot_face);
  const std::vector<int32_t>* order = slot_face.empty() ? nullptr : &slot_face;
  if (reorder && F > 0 && V > 0) {
    BuildReorder(d->vertex_indices, (size_t)F * 3, V, &ro_v, order);
    PermuteRows(d->vertices, V, 3, ro_v, &pos_p);
    h_vertices = pos_p.data();
    h_vidx = ro_v.idx.data();
    if (d->num_colors > 0 && d->vertex_colors) {
      PermuteRows(d->vertex_colors, V, 3, ro_v, &col_p);  // num_colors >= V was validated
      h_colors = col_p.data();
    }
    if (d->normal_indices && d->num_normals > 0 && d->normals) {
      BuildReorder(d->normal_indices, (size_t)F * 3, d->num_normals, &ro_n, order);
      PermuteRows(d->normals, d->num_normals, 3, ro_n, &nrm_p);
      h_normals = nrm_p.data();
      h_nidx = ro_n.idx.data();
    }
    if (d->uv_indices && d->num_uv > 0 && d->uv) {
      BuildReorder(d->uv_indices, (size_t)F * 3, d->num_uv, &ro_uv, order);
      PermuteRows(d->uv, d->num_uv, 2, ro_uv, &uv_p);
      h_uv = uv_p.data();
      h_uvidx = ro_uv.idx.data();
    }
  }

  MeshDev m{};
  m.V = V;
  m.F = F;
  m.num_normals = d->num_normals;
  m.num_colors = d->num_colors;
  m.num_face_normals = d->num_face_normals > 0 ? F : 0;
  m.num_uv = d->num_uv;
  m.num_materials = d->num_materials;
  cudaStream_t st = r->stream;
  int rc;

  // staging area on the device for the AoS host arrays (freed at the end)
  std::vector<void*> temps;
  auto upload = [&](const void* host, size_t bytes, void** dev) -> int {
    CU_TRY(r, cudaMalloc(dev, std::max<size_t>(bytes, 16)));
    temps.push_back(*dev);
    CU_TRY(r, cudaMemcpyAsync(*dev, host, bytes, cudaMemcpyHostToDevice, st));
    return UGU_B200_OK;
  };
  auto cleanup = [&](int code) {
    cudaStreamSynchronize(st);
    for (void* p : temps) cudaFree(p);
    if (code != UGU_B200_OK) FreeMesh(r);
    return code;
  };
#define UP_TRY(expr)                   \
  do {                                 \
    if ((rc = (expr))) return cleanup(rc); \
  } while (0)

  float *px = nullptr, *py = nullptr, *pz = nullptr;
  // planes padded to a multiple of 4 so the float4 loads of vertex_kernel stay in bounds
  const size_t Vpad = ((size_t)V + 3) & ~size_t(3);
  UP_TRY(MeshAlloc(r, &px, Vpad));
  UP_TRY(MeshAlloc(r, &py, Vpad));
  UP_TRY(MeshAlloc(r, &pz, Vpad));
  if (V > 0) {
    void* t;
    UP_TRY(upload(h_vertices, (size_t)V * 12, &t));
    repack_vec3_to_planes<<<Blocks(V, 256), 256, 0, st>>>((const float*)t, V, px, py, pz);
    r->stats.kernel_launches++;
    float4* p4 = nullptr;
    UP_TRY(MeshAlloc(r, &p4, V));
    repack_vec3_to_float4<<<Blocks(V, 256), 256, 0, st>>>((const float*)t, V, p4);
    r->stats.kernel_launches++;
    m.pos4 = p4;
  }
  m.pos_x = px; m.pos_y = py; m.pos_z = pz;

  int *i0 = nullptr, *i1 = nullptr, *i2 = nullptr;
  UP_TRY(MeshAlloc(r, &i0, F));
  UP_TRY(MeshAlloc(r, &i1, F));
  UP_TRY(MeshAlloc(r, &i2, F));
  void* vidx_aos = nullptr;
  if (F > 0) {
    UP_TRY(upload(h_vidx, (size_t)F * 12, &vidx_aos));
    repack_ivec3_to_planes<<<Blocks(F, 256), 256, 0, st>>>((const int*)vidx_aos, F, i0, i1, i2);
    r->stats.kernel_launches++;
  }
  m.idx0 = i0; m.idx1 = i1; m.idx2 = i2;

  if (m.num_face_normals > 0) {
    float *fx = nullptr, *fy = nullptr, *fz = nullptr;
    UP_TRY(MeshAlloc(r, &fx, F));
    UP_TRY(MeshAlloc(r, &fy, F));
    UP_TRY(MeshAlloc(r, &fz, F));
    void* t;
    UP_TRY(upload(d->face_normals, (size_t)F * 12, &t));
    repack_vec3_to_planes<<<Blocks(F, 256), 256, 0, st>>>((const float*)t, F, fx, fy, fz);
    r->stats.kernel_launches++;
    m.fn_x = fx; m.fn_y = fy; m.fn_z = fz;
  }
  // the streams of raster_kernel in raster order, and the group records of its two-pass ordering
  m.r_idx0 = i0; m.r_idx1 = i1; m.r_idx2 = i2;
  m.r_fn_x = m.fn_x; m.r_fn_y = m.fn_y; m.r_fn_z = m.fn_z;
  if (order) {
    int* planes[4] = {nullptr, nullptr, nullptr, nullptr};
    for (auto& pl : planes) UP_TRY(MeshAlloc(r, &pl, F));
    UP_TRY([&]() -> int {
      CU_TRY(r, cudaMemcpyAsync(planes[3], slot_face.data(), (size_t)F * 4, cudaMemcpyHostToDevice, st));
      return UGU_B200_OK;
    }());
    gather_planes<<<Blocks(F, 256), 256, 0, st>>>(planes[3], F, i0, i1, i2, planes[0], planes[1], planes[2]);
    r->stats.kernel_launches++;
    m.r_idx0 = planes[0]; m.r_idx1 = planes[1]; m.r_idx2 = planes[2]; m.r_face = planes[3];
    if (m.num_face_normals > 0) {
      float* fp[3] = {nullptr, nullptr, nullptr};
      for (auto& pl : fp) UP_TRY(MeshAlloc(r, &pl, F));
      gather_planes<<<Blocks(F, 256), 256, 0, st>>>(planes[3], F, m.fn_x, m.fn_y, m.fn_z, fp[0], fp[1], fp[2]);
      r->stats.kernel_launches++;
      m.r_fn_x = fp[0]; m.r_fn_y = fp[1]; m.r_fn_z = fp[2];
    }
  }
  if (F > 0 && V > 0) {
    float4* grp = nullptr;
    const size_t blocks = ((size_t)F + 255) / 256;
    UP_TRY(MeshAlloc(r, &grp, blocks));
    block_records_kernel<<<(unsigned)blocks, 256, 0, st>>>(m.pos4, m.r_idx0, m.r_idx1, m.r_idx2, F, grp);
    r->stats.kernel_launches++;
    m.grp = grp;
  }
  // Block-local vertex tables of the fused projection (raster_kernel): per 256-face block of the raster order its
  // <= 256 distinct vertices {x, y, z, global id} and per face three 8-bit indices into them, so the block projects
  // its own vertices into shared memory and no vertex kernel runs.  A block with more distinct vertices (a soup of
  // unshared triangles) gets lidx = 0xFFFFFFFF and its lanes project their own corners.  UGU_B200_FUSED=0: off.
  std::vector<float4> h_lv;
  std::vector<uint32_t> h_lidx;
  if (F > 0 && V > 0 && !(getenv("UGU_B200_FUSED") && atoi(getenv("UGU_B200_FUSED")) == 0)) {
    const size_t blocks = ((size_t)F + 255) / 256;
    h_lv.assign(blocks * 256, make_float4(0.f, 0.f, 0.f, __int_as_float_host(-1)));
    h_lidx.assign((size_t)F, 0xFFFFFFFFu);
    std::vector<int32_t> stamp((size_t)V, -1);
    std::vector<uint16_t> local((size_t)V, 0);
    for (size_t b = 0; b < blocks; b++) {
      const size_t s0 = b * 256, s1 = std::min(s0 + 256, (size_t)F);
      int n = 0;
      for (size_t s = s0; s < s1; s++) {
        const int32_t* tri = h_vidx + (size_t)(order ? slot_face[s] : (int32_t)s) * 3;
        for (int k = 0; k < 3; k++)
          if (stamp[tri[k]] != (int32_t)b) {
            stamp[tri[k]] = (int32_t)b;
            local[tri[k]] = (uint16_t)n++;
          }
      }
      if (n > 256) continue;  // lidx stays 0xFFFFFFFF, lv stays empty
      for (size_t s = s0; s < s1; s++) {
        const int32_t* tri = h_vidx + (size_t)(order ? slot_face[s] : (int32_t)s) * 3;
        uint32_t packed = 0;
        for (int k = 0; k < 3; k++) {
          const int32_t v = tri[k];
          const uint32_t l = local[v];
          packed |= l << (8 * k);
          h_lv[b * 256 + l] = make_float4(h_vertices[(size_t)v * 3], h_vertices[(size_t)v * 3 + 1],
                                          h_vertices[(size_t)v * 3 + 2], __int_as_float_host(v));
        }
        h_lidx[s] = packed;
      }
    }
    float4* lv = nullptr;
    unsigned* lidx = nullptr;
    UP_TRY(MeshAlloc(r, &lv, h_lv.size()));
    UP_TRY(MeshAlloc(r, &lidx, h_lidx.size()));
    UP_TRY([&]() -> int {
      CU_TRY(r, cudaMemcpyAsync(lv, h_lv.data(), h_lv.size() * 16, cudaMemcpyHostToDevice, st));
      CU_TRY(r, cudaMemcpyAsync(lidx, h_lidx.data(), h_lidx.size() * 4, cudaMemcpyHostToDevice, st));
      return UGU_B200_OK;
    }());
    m.lv = lv;
    m.lidx = lidx;
  }
  void* nrm_aos = nullptr;
  void* col_aos = nullptr;
  if (d->num_normals > 0 && d->normals) {
    float4* n4 = nullptr;
    UP_TRY(MeshAlloc(r, &n4, d->num_normals));
    void* t;
    UP_TRY(upload(h_normals, (size_t)d->num_normals * 12, &t));
    nrm_aos = t;
    repack_vec3_to_float4<<<Blocks(d->num_normals, 256), 256, 0, st>>>((const float*)t, d->num_normals, n4);
    r->stats.kernel_launches++;
    m.nrm4 = n4;
  } else {
    m.num_normals = 0;
  }
  if (d->num_colors > 0 && d->vertex_colors) {
    float4* c4 = nullptr;
    const int nc = V;  // colours are indexed by vertex (validated: num_colors >= V)
    UP_TRY(MeshAlloc(r, &c4, nc));
    void* t;
    UP_TRY(upload(h_colors, (size_t)nc * 12, &t));
    col_aos = t;
    repack_vec3_to_float4<<<Blocks(nc, 256), 256, 0, st>>>((const float*)t, nc, c4);
    r->stats.kernel_launches++;
    m.col4 = c4;
  } else {
    m.num_colors = 0;
  }
  // Merged per-vertex attribute records when normals are indexed exactly like vertices (meshes whose
  // normals come from CalcNormal, mesh.cc): one 32-byte gather per corner instead of two.
  if (nrm_aos && d->normal_indices && F > 0 && d->num_normals == V &&
      !(getenv("UGU_B200_MERGE_VATTR") && atoi(getenv("UGU_B200_MERGE_VATTR")) == 0) &&
      memcmp(d->normal_indices, d->vertex_indices, (size_t)F * 12) == 0) {
    float4* va = nullptr;
    UP_TRY(MeshAlloc(r, &va, (size_t)V * 2));
    build_vertex_attr<<<Blocks(V, 256), 256, 0, st>>>((const float*)nrm_aos, (const float*)col_aos, V, va);
    r->stats.kernel_launches++;
    m.vattr = va;
  }
  {  // per-face records for the resolve kernel: one 32-byte sector per winning fragment
    int4* rec = nullptr;
    UP_TRY(MeshAlloc(r, &rec, (size_t)std::max(F, 1) * 2));
    if (F > 0) {
      void* nidx_aos = nullptr;
      void* mat_dev = nullptr;
      if (d->normal_indices) {
        UP_TRY(upload(h_nidx, (size_t)F * 12, &nidx_aos));
        r->has_nidx = true;
      }
      if (d->material_ids) {
        UP_TRY(upload(d->material_ids, (size_t)F * 4, &mat_dev));
        r->has_matids = true;
      }
      build_face_records<<<Blocks(F, 256), 256, 0, st>>>((const int*)vidx_aos, (const int*)nidx_aos,
                                                         (const int*)mat_dev, F, rec);
      r->stats.kernel_launches++;
    }
    m.frec = rec;
  }
  if (d->uv_indices && F > 0) {
    int4* n = nullptr;
    UP_TRY(MeshAlloc(r, &n, F));
    void* t;
    UP_TRY(upload(h_uvidx, (size_t)F * 12, &t));
    repack_ivec3_to_int4<<<Blocks(F, 256), 256, 0, st>>>((const int*)t, F, n);
    r->stats.kernel_launches++;
    m.uvidx = n;
    r->has_uvidx = true;
  }
  if (d->num_uv > 0 && d->uv) {
    float2* u = nullptr;
    UP_TRY(MeshAlloc(r, &u, d->num_uv));
    if ((rc = [&]() -> int {
          CU_TRY(r, cudaMemcpyAsync(u, h_uv, (size_t)d->num_uv * 8, cudaMemcpyHostToDevice, st));
          return UGU_B200_OK;
        }()))
      return cleanup(rc);
    m.uv2 = u;
  } else {
    m.num_uv = 0;
  }
  r->tex_present.assign(d->num_materials, false);
  if (d->num_materials > 0) {
    std::vector<const uchar4*> tex_ptrs(d->num_materials, nullptr);
    std::vector<int2> tex_sizes(d->num_materials, make_int2(0, 0));
    for (int i = 0; i < d->num_materials; i++) {
      const bool present = d->tex_data && d->tex_data[i] && d->tex_w[i] > 0 && d->tex_h[i] > 0;
      r->tex_present[i] = present;
      if (!present) continue;
      const size_t n = (size_t)d->tex_w[i] * d->tex_h[i];
      uchar4* t4 = nullptr;
      UP_TRY(MeshAlloc(r, &t4, n));
      void* t;
      UP_TRY(upload(d->tex_data[i], n * 3, &t));
      repack_rgb_to_rgba<<<Blocks(n, 256), 256, 0, st>>>((const uint8_t*)t, n, t4);
      r->stats.kernel_launches++;
      tex_ptrs[i] = t4;
      tex_sizes[i] = make_int2(d->tex_w[i], d->tex_h[i]);
    }
    const uchar4** dev_ptrs = nullptr;
    int2* dev_sizes = nullptr;
    UP_TRY(MeshAlloc(r, &dev_ptrs, d->num_materials));
    UP_TRY(MeshAlloc(r, &dev_sizes, d->num_materials));
    if ((rc = [&]() -> int {
          CU_TRY(r, cudaMemcpyAsync(dev_ptrs, tex_ptrs.data(), sizeof(void*) * d->num_materials,
                                    cudaMemcpyHostToDevice, st));
          CU_TRY(r, cudaMemcpyAsync(dev_sizes, tex_sizes.data(), sizeof(int2) * d->num_materials,
                                    cudaMemcpyHostToDevice, st));
          CU_TRY(r, cudaStreamSynchronize(st));  // tex_ptrs / tex_sizes are stack-owned
          return UGU_B200_OK;
        }()))
      return cleanup(rc);
    m.tex = dev_ptrs;
    m.tex_size = dev_sizes;
  }
#undef UP_TRY
  if ((rc = [&]() -> int {
        CU_TRY(r, cudaGetLastError());
        CU_TRY(r, cudaStreamSynchronize(st));
        return UGU_B200_OK;
      }()))
    return cleanup(rc);
  // per-view scratch that depends on the mesh size
  if ((rc = Grow(r, &r->proj, std::max<size_t>(Vpad, 4) * sizeof(float4)))) return cleanup(rc);
  if ((rc = Grow(r, &r->q_face, std::max<size_t>(F, 1) * 4))) return cleanup(rc);
  if ((rc = Grow(r, &r->q_job, std::max<size_t>(F, 1) * 8))) return cleanup(rc);
  r->scratch_used = false;  // the stream was synchronised above: nothing in flight uses the scratch
  r->mesh = m;
  r->mesh_ready = true;
  return cleanup(UGU_B200_OK);
}

int ugu_b200_rasterizer_render_device(ugu_b200_rasterizer* r, const ugu_b200_camera* cam,
                                      const ugu_b200_outputs* dev_out, void* stream) {
  if (!r) return Fail(nullptr, UGU_B200_E_INVALID, "rasterizer is NULL");
  int rc = Validate(r, cam, dev_out);
  if (rc) return rc;
  const uintptr_t align = (uintptr_t)dev_out->color | (uintptr_t)dev_out->depth |
                          (uintptr_t)dev_out->normal | (uintptr_t)dev_out->mask |
                          (uintptr_t)dev_out->face_id;
  if (align & 15u) return Fail(r, UGU_B200_E_INVALID, "device outputs must be 16-byte aligned");
  CU_TRY(r, cudaSetDevice(r->device));
  cudaStream_t st = stream ? (cudaStream_t)stream : r->stream;
  return RenderView(r, cam, dev_out, st);
}

int ugu_b200_rasterizer_render(ugu_b200_rasterizer* r, const ugu_b200_camera* cam,
                               const ugu_b200_outputs* host_out) {
  if (!r) return Fail(nullptr, UGU_B200_E_INVALID, "rasterizer is NULL");
  int rc = Validate(r, cam, host_out);
  if (rc) return rc;
  CU_TRY(r, cudaSetDevice(r->device));
  const size_t P = (size_t)cam->width * cam->height;
  ugu_b200_outputs dout;
  if ((rc = EnsureInternalOutputs(r, 0, host_out, P, &dout))) return rc;
  std::vector<CopySeg> segs;
  CollectSegs(host_out, &dout, P, 0, 1, &segs);
  // The D2H of a view (24 B per pixel with all five outputs) is 12x its kernels, and nothing can be copied before
  // the raster kernel is done — but the resolve need not be: it runs in two bands (the first eighth of the image,
  // then the rest), and the copy stream starts on band 0 while band 1 is resolved.  Pinned destinations only (the
  // pageable path stages through its own ring).  MEASURED AND OFF BY DEFAULT (UGU_B200_BANDS=1 selects it): ten DMAs
  // and two cross-stream waits cost more than the 25 us of resolve they hide — config 2 e2e 1.024 vs 1.005 ms
  // (profiles/r05e_banded_render_ab.jsonl).
  // (UGU_B200_BANDS=2 bands every image of at least 64 chunks: the tests' small fixtures)
  bool banded = r->bands != 0 && r->mesh.F > 0 && P >= (r->bands >= 2 ? size_t(2048) : (size_t(1) << 18));
  for (const CopySeg& s : segs) banded = banded && DestPinned(r, s.dst);
  if (banded) {
    BandPlan plan;
    plan.n = 2;
    plan.split[0] = (unsigned)(((P + 31) / 32 / 8 + 7) & ~size_t(7));
    for (int b = 0; b < 2; b++) {
      if (!r->ev_band[b]) CU_TRY(r, cudaEventCreateWithFlags(&r->ev_band[b], cudaEventDisableTiming));
      plan.ev[b] = r->ev_band[b];
    }
    if ((rc = RenderView(r, cam, &dout, r->stream, nullptr, &plan))) return rc;
    CU_TRY(r, cudaMemcpyAsync(r->counters_host, (char*)r->counters.p + 32, 24, cudaMemcpyDeviceToHost,
                              r->stream));
    const size_t cut = std::min(P, (size_t)plan.split[0] * 32);  // first pixel of band 1
    for (int b = 0; b < 2; b++) {
      CU_TRY(r, cudaStreamWaitEvent(r->copy_stream, plan.ev[b], 0));
      const size_t p0 = b ? cut : 0, p1 = b ? P : cut;
      for (const CopySeg& s : segs) {
        const size_t bpp = s.bytes / P;  // 3, 4, 12, 1, 4
        if (p1 > p0)
          CU_TRY(r, cudaMemcpyAsync((char*)s.dst + p0 * bpp, (const char*)s.src + p0 * bpp, (p1 - p0) * bpp,
                                    cudaMemcpyDeviceToHost, r->copy_stream));
      }
    }
    for (const CopySeg& s : segs) r->stats.d2h_bytes += s.bytes;
    CU_TRY(r, cudaStreamSynchronize(r->copy_stream));
    CU_TRY(r, cudaStreamSynchronize(r->stream));
    return SnapshotCounters(r);
  }
  r->host_snapshot = true;
  rc = RenderView(r, cam, &dout, r->stream);
  r->host_snapshot = false;
  if (rc) return rc;
  if (!r->counters_mapped)
    CU_TRY(r, cudaMemcpyAsync(r->counters_host, (char*)r->counters.p + 32, 24, cudaMemcpyDeviceToHost,
                              r->stream));
  if ((rc = CopyOut(r, segs, r->stream))) return rc;
  return SnapshotCounters(r);
}

namespace {
// Shared front end of the two G-buffer entry points: validation and the kernel parameter block.
int PrepareGbuf(ugu_b200_rasterizer* r, const ugu_b200_camera* cam, const ugu_b200_gbuffer_option* opt,
                const ugu_b200_gbuffer* out, GBufParams* gp) {
  if (!opt) return Fail(r, UGU_B200_E_INVALID, "gbuffer option is NULL");
  if (!out || (!out->color && !out->normal_cam && !out->normal_wld && !out->pos_cam && !out->pos_wld &&
               !out->depth_01 && !out->stencil && !out->face_id && !out->uv && !out->bary && !out->geo_id))
    return Fail(r, UGU_B200_E_INVALID, "all arguments are nullptr. nothing to do");
  // the Render preconditions (util_private.cc:12-90); without a colour request only the camera, the
  // mesh and back-face culling matter, so validate as a depth-only Render
  ugu_b200_outputs probe{};
  float dummy = 0.f;
  if (out->color) probe.color = out->color;
  else probe.depth = &dummy;
  ugu_b200_option saved = r->opt;
  if (!out->color) {
    r->opt.diffuse_color = UGU_B200_COLOR_NONE;
    r->opt.diffuse_shading = 0;
    r->opt.shading_normal = (r->mesh.num_face_normals > 0) ? UGU_B200_NORMAL_FACE : r->opt.shading_normal;
  }
  const int rc = Validate(r, cam, &probe);
  r->opt = saved;
  if (rc) return rc;
  if (!(opt->z_near > 0.0f) || !(opt->z_far > opt->z_near))
    return Fail(r, UGU_B200_E_INVALID, "gbuffer needs 0 < z_near < z_far (got %g, %g)", opt->z_near, opt->z_far);
  memset(gp, 0, sizeof(*gp));
  gp->normal_cam = out->normal_cam;
  gp->normal_wld = out->normal_wld;
  gp->pos_cam = out->pos_cam;
  gp->pos_wld = out->pos_wld;
  gp->depth_01 = out->depth_01;
  gp->stencil = out->stencil;
  gp->face_id = out->face_id;
  gp->uv = out->uv;
  gp->bary = out->bary;
  gp->geo_id = out->geo_id;
  // GetProjectionMatrixOpenGlForPinhole (camera_util.cc:254-255), single float operations
  const volatile float zn = opt->z_near, zf = opt->z_far;
  const volatile float sum = zf + zn, diff = zf - zn, two_f = 2.f * zf;
  const volatile float two_fn = two_f * zn;
  gp->A = -sum / diff;
  gp->B = -two_fn / diff;
  gp->flat_normal = opt->flat_normal ? 1 : 0;
  gp->geo_id_value = opt->geo_id;
  return UGU_B200_OK;
}
}  // namespace

int ugu_b200_rasterizer_render_gbuffer_device(ugu_b200_rasterizer* r, const ugu_b200_camera* cam,
                                              const ugu_b200_gbuffer_option* opt,
                                              const ugu_b200_gbuffer* dev_out, void* stream) {
  if (!r) return Fail(nullptr, UGU_B200_E_INVALID, "rasterizer is NULL");
  GBufParams gp{};
  int rc = PrepareGbuf(r, cam, opt, dev_out, &gp);
  if (rc) return rc;
  const uintptr_t align = (uintptr_t)dev_out->color | (uintptr_t)gp.normal_cam | (uintptr_t)gp.normal_wld |
                          (uintptr_t)gp.pos_cam | (uintptr_t)gp.pos_wld | (uintptr_t)gp.depth_01 |
                          (uintptr_t)gp.stencil | (uintptr_t)gp.face_id | (uintptr_t)gp.uv |
                          (uintptr_t)gp.bary | (uintptr_t)gp.geo_id;
  if (align & 15u) return Fail(r, UGU_B200_E_INVALID, "device outputs must be 16-byte aligned");
  CU_TRY(r, cudaSetDevice(r->device));
  cudaStream_t st = stream ? (cudaStream_t)stream : r->stream;
  ugu_b200_outputs dout{};
  dout.color = dev_out->color;
  return RenderView(r, cam, &dout, st, &gp);
}

int ugu_b200_rasterizer_render_gbuffer(ugu_b200_rasterizer* r, const ugu_b200_camera* cam,
                                       const ugu_b200_gbuffer_option* opt, const ugu_b200_gbuffer* host_out) {
  if (!r) return Fail(nullptr, UGU_B200_E_INVALID, "rasterizer is NULL");
  GBufParams gp{};
  int rc = PrepareGbuf(r, cam, opt, host_out, &gp);
  if (rc) return rc;
  CU_TRY(r, cudaSetDevice(r->device));
  const size_t P = (size_t)cam->width * cam->height;
  // one device arena for the requested fields (each slice 256-byte aligned)
  struct Field { void* host; void** dev; size_t bytes; };
  ugu_b200_outputs dout{};
  Field fields[] = {
      {host_out->color, (void**)&dout.color, P * 3},        {host_out->normal_cam, (void**)&gp.normal_cam, P * 12},
      {host_out->normal_wld, (void**)&gp.normal_wld, P * 12}, {host_out->pos_cam, (void**)&gp.pos_cam, P * 12},
      {host_out->pos_wld, (void**)&gp.pos_wld, P * 12},     {host_out->depth_01, (void**)&gp.depth_01, P * 4},
      {host_out->stencil, (void**)&gp.stencil, P},          {host_out->face_id, (void**)&gp.face_id, P * 4},
      {host_out->uv, (void**)&gp.uv, P * 12},               {host_out->bary, (void**)&gp.bary, P * 12},
      {host_out->geo_id, (void**)&gp.geo_id, P * 4}};
  size_t total = 0;
  for (const Field& f : fields)
    if (f.host) total += (f.bytes + 255) & ~size_t(255);
  if ((rc = Grow(r, &r->gbuf_arena, total + 256))) return rc;
  std::vector<CopySeg> segs;
  size_t off = 0;
  for (const Field& f : fields) {
    if (!f.host) continue;
    *f.dev = (char*)r->gbuf_arena.p + off;
    segs.push_back({f.host, *f.dev, f.bytes});
    off += (f.bytes + 255) & ~size_t(255);
  }
  if ((rc = RenderView(r, cam, &dout, r->stream, &gp))) return rc;
  CU_TRY(r, cudaMemcpyAsync(r->counters_host, (char*)r->counters.p + 32, 24, cudaMemcpyDeviceToHost,
                            r->stream));
  if ((rc = CopyOut(r, segs, r->stream))) return rc;
  return SnapshotCounters(r);
}

namespace {
// The five kernels of one Raytracer-mode view into device outputs `dout` on stream `st`.
int RaytraceView(ugu_b200_rasterizer* r, const ugu_b200_camera* cam, const float* org_w,
                 const ugu_b200_outputs* dout, cudaStream_t st) {
  int rc = AcquireScratch(r, st);
  if (rc) return rc;
  if ((rc = EnsureFrame(r, cam->width, cam->height, 1, st))) return rc;
  const size_t P = (size_t)cam->width * cam->height;
  if ((rc = Grow(r, &r->rt_rays, P * 32))) return rc;
  ViewBatch<1> vb;
  MakeViewParams(cam, &vb.v[0]);
  const ViewParams& vp = vb.v[0];
  FrameDev fr = MakeFrame(r, cam->width, cam->height);
  const MeshDev& m = r->mesh;
  RtParams rp;
  rp.org[0] = org_w[0];
  rp.org[1] = org_w[1];
  rp.org[2] = org_w[2];
  rp.t_min = 0.0001f;  // bvh_nanort.h:119-121
  rp.t_max = 1.0e+30f;
  rp.depth_scale = r->opt.depth_scale;
  rp.coord_scale = fmaxf(r->mesh_max_abs, fmaxf(fabsf(org_w[0]), fmaxf(fabsf(org_w[1]), fabsf(org_w[2]))));
  // like coop_max of the rasterizer: with fewer face warps than the machine holds, mid-size bboxes are better
  // spread over the whole GPU by the job queue than walked by their own warp
  rp.inlane_max = (r->mesh.F >= 4 * 32 * 32 * g_num_sms) ? kRtInLaneMax : 16;
  rp.ray_a = (float4*)r->rt_rays.p;
  rp.ray_b = rp.ray_a + P;
  const ShadeParams sp = MakeShade(r, dout, P);
  const bool pdl = r->use_pdl;
  // occlusion ordering of rt_raster_kernel, same policy as the rasterizer's
  if (m.grp != nullptr && (r->two_pass >= 0 ? r->two_pass != 0 : m.F >= kTwoPassMinFaces)) fr.order = (int*)r->order.p;
  const size_t vthreads = std::max<size_t>(std::max<size_t>((size_t)m.V, (size_t)vp.W + vp.H), ((size_t)m.F + 255) / 256);
  CU_TRY(r, LaunchView(vertex_kernel_b<1>, Blocks(vthreads, 256), 256, st, pdl, vb, 1, m, fr));
  CU_TRY(r, LaunchView(rt_ray_kernel, Blocks(P, 256), 256, st, pdl, vp, fr, rp));
  int launches = 3;
  if (m.F > 0) {
    CU_TRY(r, LaunchView(rt_raster_kernel, Blocks((size_t)m.F, 256), 256, st, pdl, vp, m, fr, rp));
    // resident grid: 8 CTAs of 8 warps per SM share the global job space
    CU_TRY(r, LaunchView(rt_large_kernel, (unsigned)(g_num_sms * 8), 256, st, pdl, vp, m, fr, rp));
    launches += 2;
  }
  CU_TRY(r, LaunchView(rt_resolve_kernel, Blocks(P, 256), 256, st, pdl, vp, m, fr, rp, sp));
  CU_TRY(r, cudaGetLastError());
  r->stats.kernel_launches += launches;
  r->stats.views++;
  r->stats.h2d_bytes += sizeof(ViewParams) + sizeof(RtParams);
  return ReleaseScratch(r, st);
}
}  // namespace

int ugu_b200_rasterizer_raytrace_device(ugu_b200_rasterizer* r, const ugu_b200_camera* cam, const float* org_w,
                                        const ugu_b200_outputs* dev_out, void* stream) {
  if (!r) return Fail(nullptr, UGU_B200_E_INVALID, "rasterizer is NULL");
  int rc = Validate(r, cam, dev_out);
  if (rc) return rc;
  if (!org_w) return Fail(r, UGU_B200_E_INVALID, "ray origin is NULL");
  const uintptr_t align = (uintptr_t)dev_out->color | (uintptr_t)dev_out->depth | (uintptr_t)dev_out->normal |
                          (uintptr_t)dev_out->mask | (uintptr_t)dev_out->face_id;
  if (align & 15u) return Fail(r, UGU_B200_E_INVALID, "device outputs must be 16-byte aligned");
  CU_TRY(r, cudaSetDevice(r->device));
  return RaytraceView(r, cam, org_w, dev_out, stream ? (cudaStream_t)stream : r->stream);
}

int ugu_b200_rasterizer_raytrace(ugu_b200_rasterizer* r, const ugu_b200_camera* cam, const float* org_w,
                                 const ugu_b200_outputs* host_out) {
  if (!r) return Fail(nullptr, UGU_B200_E_INVALID, "rasterizer is NULL");
  int rc = Validate(r, cam, host_out);
  if (rc) return rc;
  if (!org_w) return Fail(r, UGU_B200_E_INVALID, "ray origin is NULL");
  CU_TRY(r, cudaSetDevice(r->device));
  const size_t P = (size_t)cam->width * cam->height;
  ugu_b200_outputs dout;
  if ((rc = EnsureInternalOutputs(r, 0, host_out, P, &dout))) return rc;
  if ((rc = RaytraceView(r, cam, org_w, &dout, r->stream))) return rc;
  CU_TRY(r, cudaMemcpyAsync(r->counters_host, (char*)r->counters.p + 32, 24, cudaMemcpyDeviceToHost,
                            r->stream));
  std::vector<CopySeg> segs;
  CollectSegs(host_out, &dout, P, 0, 1, &segs);
  if ((rc = CopyOut(r, segs, r->stream))) return rc;
  return SnapshotCounters(r);
}

int ugu_b200_rasterizer_render_batch(ugu_b200_rasterizer* r, int32_t num_views,
                                     const ugu_b200_camera* cams, const ugu_b200_outputs* out,
                                     int32_t outputs_on_device, void* stream) {
  if (!r) return Fail(nullptr, UGU_B200_E_INVALID, "rasterizer is NULL");
  if (num_views <= 0 || !cams) return Fail(r, UGU_B200_E_INVALID, "no views");
  int rc;
  for (int i = 0; i < num_views; i++) {
    if ((rc = Validate(r, &cams[i], out))) return rc;
    if (cams[i].width != cams[0].width || cams[i].height != cams[0].height)
      return Fail(r, UGU_B200_E_INVALID, "all views of a batch must share one image size");
  }
  CU_TRY(r, cudaSetDevice(r->device));
  const int W = cams[0].width, H = cams[0].height;
  const size_t P = (size_t)W * H;
  const int batch = BatchViews(r, num_views, W, H);
  if (outputs_on_device) {
    // only the bases must be 16-byte aligned: a view slab that is not (odd W*H) is written with scalar stores
    const uintptr_t align = (uintptr_t)out->color | (uintptr_t)out->depth | (uintptr_t)out->normal |
                            (uintptr_t)out->mask | (uintptr_t)out->face_id;
    if (align & 15u) return Fail(r, UGU_B200_E_INVALID, "device outputs must be 16-byte aligned");
    cudaStream_t st = stream ? (cudaStream_t)stream : r->stream;
    return RenderViews(r, cams, num_views, out, st, batch);
  }
  // Host destinations: launch sets of `batch` views render into one of two internal output sets on the
  // compute stream; pinned destinations are drained by DMA on the copy stream while the next set renders
  // (one copy per output and set: the views of a set are contiguous on both sides), pageable ones through
  // the staging ring.
  bool all_pinned = true;
  {
    std::vector<CopySeg> probe;
    ugu_b200_outputs none{};
    CollectSegs(out, &none, P, 0, 1, &probe);
    for (const CopySeg& s : probe) all_pinned = all_pinned && DestPinned(r, s.dst);
  }
  int set = 0;
  for (int i = 0; i < num_views; i += batch, set ^= (all_pinned ? 1 : 0)) {
    const int k = std::min(batch, num_views - i);
    ugu_b200_outputs dout;
    if ((rc = EnsureInternalOutputs(r, set, out, P * (size_t)batch, &dout))) return rc;
    if (all_pinned && r->copied_pending[set]) CU_TRY(r, cudaStreamWaitEvent(r->stream, r->ev_copied[set], 0));
    if ((rc = RenderViews(r, cams + i, k, &dout, r->stream, batch))) return rc;
    std::vector<CopySeg> segs;
    CollectSegs(out, &dout, P, (size_t)i, (size_t)k, &segs);
    if (all_pinned) {
      CU_TRY(r, cudaEventRecord(r->ev_render[set], r->stream));
      CU_TRY(r, cudaStreamWaitEvent(r->copy_stream, r->ev_render[set], 0));
      for (const CopySeg& s : segs) {
        r->stats.d2h_bytes += s.bytes;
        CU_TRY(r, cudaMemcpyAsync(s.dst, s.src, s.bytes, cudaMemcpyDeviceToHost, r->copy_stream));
      }
      CU_TRY(r, cudaEventRecord(r->ev_copied[set], r->copy_stream));
      r->copied_pending[set] = true;
    } else {
      if ((rc = CopyOut(r, segs, r->stream))) return rc;
    }
  }
  CU_TRY(r, cudaMemcpyAsync(r->counters_host, (char*)r->counters.p + 32, 24, cudaMemcpyDeviceToHost,
                            r->stream));
  CU_TRY(r, cudaStreamSynchronize(r->stream));
  CU_TRY(r, cudaStreamSynchronize(r->copy_stream));
  r->copied_pending[0] = r->copied_pending[1] = false;
  if ((rc = SnapshotCounters(r))) return rc;
  return UGU_B200_OK;
}

uint64_t ugu_b200_rasterizer_last_hazards(const ugu_b200_rasterizer* r) {
  return r ? r->stats.hazard_fragments : 0;
}

int ugu_b200_rasterizer_sync(ugu_b200_rasterizer* r) {
  if (!r) return Fail(nullptr, UGU_B200_E_INVALID, "rasterizer is NULL");
  CU_TRY(r, cudaSetDevice(r->device));
  CU_TRY(r, cudaStreamSynchronize(r->stream));
  CU_TRY(r, cudaStreamSynchronize(r->copy_stream));
  return UGU_B200_OK;
}

int ugu_b200_depth_to_point_cloud_device(ugu_b200_rasterizer* r, const ugu_b200_camera* cam,
                                         const float* depth_dev, float* point_cloud_dev,
                                         int32_t gl_coord, int32_t zero_invalid, void* stream) {
  if (!r) return Fail(nullptr, UGU_B200_E_INVALID, "rasterizer is NULL");
  if (!cam || !depth_dev || !point_cloud_dev) return Fail(r, UGU_B200_E_INVALID, "NULL argument");
  if (cam->width <= 0 || cam->height <= 0 || cam->width > 65535 || cam->height > 65535 ||
      (long long)cam->width * cam->height >= (1ll << 31))
    return Fail(r, UGU_B200_E_INVALID, "camera size %dx%d is invalid", cam->width, cam->height);
  if (((uintptr_t)depth_dev | (uintptr_t)point_cloud_dev) & 15u)
    return Fail(r, UGU_B200_E_INVALID, "device buffers must be 16-byte aligned");
  CU_TRY(r, cudaSetDevice(r->device));
  cudaStream_t st = stream ? (cudaStream_t)stream : r->stream;
  ViewParams vp;
  MakeViewParams(cam, &vp);
  const size_t P = (size_t)cam->width * cam->height;
  depth_to_point_cloud_kernel<<<Blocks((P + 3) / 4, 256), 256, 0, st>>>(vp, depth_dev, point_cloud_dev,
                                                                         gl_coord, zero_invalid);
  CU_TRY(r, cudaGetLastError());
  r->stats.kernel_launches++;
  return UGU_B200_OK;
}

int ugu_b200_depth_to_point_cloud(ugu_b200_rasterizer* r, const ugu_b200_camera* cam,
                                  const float* depth_host, float* point_cloud_host, int32_t gl_coord,
                                  int32_t zero_invalid) {
  if (!r) return Fail(nullptr, UGU_B200_E_INVALID, "rasterizer is NULL");
  if (!cam || !depth_host || !point_cloud_host) return Fail(r, UGU_B200_E_INVALID, "NULL argument");
  if (cam->width <= 0 || cam->height <= 0)
    return Fail(r, UGU_B200_E_INVALID, "camera size %dx%d is invalid", cam->width, cam->height);
  CU_TRY(r, cudaSetDevice(r->device));
  const size_t P = (size_t)cam->width * cam->height;
  int rc;
  // the depth / normal output buffers of set 1 double as scratch (set 0 belongs to Render)
  if ((rc = Grow(r, &r->out_depth[1], P * 4 + 16))) return rc;
  if ((rc = Grow(r, &r->out_normal[1], P * 12 + 16))) return rc;
  float* d_dev = (float*)r->out_depth[1].p;
  float* pc_dev = (float*)r->out_normal[1].p;
  CU_TRY(r, cudaMemcpyAsync(d_dev, depth_host, P * 4, cudaMemcpyHostToDevice, r->stream));
  r->stats.h2d_bytes += P * 4;
  if (!zero_invalid) {  // in/out semantics: the kernel must see what the image held
    CU_TRY(r, cudaMemcpyAsync(pc_dev, point_cloud_host, P * 12, cudaMemcpyHostToDevice, r->stream));
    r->stats.h2d_bytes += P * 12;
  }
  if ((rc = ugu_b200_depth_to_point_cloud_device(r, cam, d_dev, pc_dev, gl_coord, zero_invalid, r->stream)))
    return rc;
  std::vector<CopySeg> segs;
  segs.push_back({point_cloud_host, pc_dev, P * 12});
  return CopyOut(r, segs, r->stream);
}

int ugu_b200_uv_raster_device(ugu_b200_rasterizer* r, const ugu_b200_uv_raster_desc* d,
                              void* texture_dev, uint8_t* mask_dev, void* stream) {
  if (!r) return Fail(nullptr, UGU_B200_E_INVALID, "rasterizer is NULL");
  if (!d || !texture_dev) return Fail(r, UGU_B200_E_INVALID, "NULL argument");
  // raster_util.h:148-151,238-240: the reference returns false
  if (d->num_faces <= 0) return Fail(r, UGU_B200_E_INVALID, "no faces");
  if (d->per_face && d->num_attrs != d->num_faces)
    return Fail(r, UGU_B200_E_INVALID, "face attribute count %d != face count %d", d->num_attrs, d->num_faces);
  if (!d->attrs || !d->uv || !d->uv_indices || (!d->per_face && !d->attr_indices))
    return Fail(r, UGU_B200_E_INVALID, "NULL array");
  if (d->width <= 0 || d->height <= 0 || (long long)d->width * d->height >= (1ll << 31))
    return Fail(r, UGU_B200_E_INVALID, "texture size %dx%d is invalid", d->width, d->height);
  if (d->num_faces >= (1 << (64 - kQueueJobBits)))
    return Fail(r, UGU_B200_E_INVALID, "more than 2^26 - 1 faces");
  if (d->texture_type != 0 && d->texture_type != 1) return Fail(r, UGU_B200_E_INVALID, "bad texture_type");
  if (((uintptr_t)texture_dev | (uintptr_t)mask_dev | (uintptr_t)d->attrs | (uintptr_t)d->uv |
       (uintptr_t)d->uv_indices | (uintptr_t)d->attr_indices) & 15u)
    return Fail(r, UGU_B200_E_INVALID, "device buffers must be 16-byte aligned");
  CU_TRY(r, cudaSetDevice(r->device));
  cudaStream_t st = stream ? (cudaStream_t)stream : r->stream;
  const size_t P = (size_t)d->width * d->height;
  int rc;
  if (r->uv_win.bytes < P * 4 + 16) r->uv_win_clean = 0;
  if ((rc = Grow(r, &r->uv_win, P * 4 + 16))) return rc;
  if ((rc = Grow(r, &r->uv_q_face, (size_t)d->num_faces * 4))) return rc;
  if ((rc = Grow(r, &r->uv_q_job, (size_t)d->num_faces * 8))) return rc;
  if ((rc = Grow(r, &r->uv_rec, (size_t)d->num_faces * 64))) return rc;
  if (!r->uv_counter.p) {
    if ((rc = Grow(r, &r->uv_counter, 16))) return rc;
    CU_TRY(r, cudaMemsetAsync(r->uv_counter.p, 0, 16, st));
  }
  if (r->uv_win_clean < P) {  // the resolve kernel leaves [0, P) at -1 for the next call
    fill_i32<<<Blocks(P, 256), 256, 0, st>>>((int*)r->uv_win.p, P, -1);
    r->uv_win_clean = P;
    r->stats.kernel_launches++;
  }
  UvRasterParams p;
  p.W = d->width;
  p.H = d->height;
  p.F = d->num_faces;
  p.per_face = d->per_face ? 1 : 0;
  p.texture_f32 = d->texture_type;
  p.keep_texture = d->keep_texture ? 1 : 0;
  p.keep_mask = d->keep_mask ? 1 : 0;
  p.attrs = d->attrs;
  p.attr_idx = d->attr_indices;
  p.uv = d->uv;
  p.uv_idx = d->uv_indices;
  p.win = (int*)r->uv_win.p;
  p.counter = (unsigned long long*)r->uv_counter.p;
  p.q_face = (int*)r->uv_q_face.p;
  p.q_job = (unsigned long long*)r->uv_q_job.p;
  p.rec = (float4*)r->uv_rec.p;
  p.texture = texture_dev;
  p.mask = mask_dev;
  const bool pdl = r->use_pdl;
  CU_TRY(r, LaunchView(uv_setup_kernel, dim3(Blocks((size_t)p.F, 256)), 256, st, pdl, p));
  CU_TRY(r, LaunchView(uv_tile_kernel, dim3((unsigned)(g_num_sms * g_binned_ctas_per_sm)), 128, st, pdl, p));
  // Long runs of one winner (>= 64 texels per face on average: vc2tex) take the 16-texels-per-thread kernel with one
  // live record; meshes of texel-sized faces (posmap) the lane-per-texel kernel (one record per thread, 6 CTAs per
  // SM).  Measured both ways (profiles/r05p_uvraster_bench_lane*.jsonl): posmap 0.108 -> 0.094 ms with the lane
  // kernel, vc2tex 0.119 -> 0.147 ms.  UGU_B200_UV_RESOLVE=0: the first form (four records per thread) for both.
  // (1 / 2 force the lane / the one-live-record kernel: the tests run every fixture through each)
  const int uv_resolve_mode = getenv("UGU_B200_UV_RESOLVE") ? atoi(getenv("UGU_B200_UV_RESOLVE")) : -1;
  const bool long_runs = uv_resolve_mode == 1 ? false
                         : (uv_resolve_mode == 2 ? true : P >= (size_t)64 * (size_t)std::max(d->num_faces, 1));
  if (uv_resolve_mode == 0 && !long_runs) {
    CU_TRY(r, LaunchView(uv_resolve_kernel<false>, dim3(Blocks(P, 256 * 4 * kUvGroupsPerThread)), 256, st, pdl, p));
  } else if (long_runs) {
    CU_TRY(r, LaunchView(uv_resolve_kernel<true>, dim3(Blocks(P, 256 * 4 * kUvGroupsPerThread)), 256, st, pdl, p));
  } else {
    const unsigned chunks = (unsigned)((P + 31) / 32);
    const unsigned grid = std::max(1u, std::min((unsigned)(g_num_sms * 6), (chunks + 7) / 8));
    CU_TRY(r, LaunchView(uv_resolve_lane_kernel, dim3(grid), 256, st, pdl, p));
  }
  CU_TRY(r, cudaGetLastError());
  r->stats.kernel_launches += 3;
  return UGU_B200_OK;
}

int ugu_b200_uv_raster(ugu_b200_rasterizer* r, const ugu_b200_uv_raster_desc* d, void* texture_host,
                       uint8_t* mask_host) {
  if (!r) return Fail(nullptr, UGU_B200_E_INVALID, "rasterizer is NULL");
  if (!d || !texture_host) return Fail(r, UGU_B200_E_INVALID, "NULL argument");
  if (d->num_faces <= 0) return Fail(r, UGU_B200_E_INVALID, "no faces");
  if (d->per_face && d->num_attrs != d->num_faces)
    return Fail(r, UGU_B200_E_INVALID, "face attribute count %d != face count %d", d->num_attrs, d->num_faces);
  if (d->num_attrs <= 0 || d->num_uv <= 0 || !d->attrs || !d->uv || !d->uv_indices ||
      (!d->per_face && !d->attr_indices))
    return Fail(r, UGU_B200_E_INVALID, "empty or NULL array");
  if (d->width <= 0 || d->height <= 0 || (long long)d->width * d->height >= (1ll << 31))
    return Fail(r, UGU_B200_E_INVALID, "texture size %dx%d is invalid", d->width, d->height);
  const size_t nf3 = (size_t)d->num_faces * 3;
  if (!IndicesInRange(d->uv_indices, nf3, d->num_uv))
    return Fail(r, UGU_B200_E_INVALID, "uv index out of range");
  if (!d->per_face && !IndicesInRange(d->attr_indices, nf3, d->num_attrs))
    return Fail(r, UGU_B200_E_INVALID, "attribute index out of range");
  CU_TRY(r, cudaSetDevice(r->device));
  auto up16 = [](size_t b) { return (b + 15) & ~(size_t)15; };
  const size_t b_attr = up16((size_t)d->num_attrs * 12), b_uv = up16((size_t)d->num_uv * 8);
  const size_t b_idx = up16(nf3 * 4);
  const size_t P = (size_t)d->width * d->height;
  const size_t b_tex = P * (d->texture_type ? 12 : 3);
  int rc;
  if ((rc = Grow(r, &r->uv_in, b_attr + b_uv + 2 * b_idx))) return rc;
  if ((rc = Grow(r, &r->uv_tex, up16(b_tex)))) return rc;
  if (mask_host && (rc = Grow(r, &r->uv_mask, up16(P)))) return rc;
  char* base = (char*)r->uv_in.p;
  ugu_b200_uv_raster_desc dd = *d;
  dd.attrs = (const float*)base;
  dd.uv = (const float*)(base + b_attr);
  dd.uv_indices = (const int32_t*)(base + b_attr + b_uv);
  dd.attr_indices = d->per_face ? nullptr : (const int32_t*)(base + b_attr + b_uv + b_idx);
  cudaStream_t st = r->stream;
  CU_TRY(r, cudaMemcpyAsync((void*)dd.attrs, d->attrs, (size_t)d->num_attrs * 12, cudaMemcpyHostToDevice, st));
  CU_TRY(r, cudaMemcpyAsync((void*)dd.uv, d->uv, (size_t)d->num_uv * 8, cudaMemcpyHostToDevice, st));
  CU_TRY(r, cudaMemcpyAsync((void*)dd.uv_indices, d->uv_indices, nf3 * 4, cudaMemcpyHostToDevice, st));
  r->stats.h2d_bytes += (size_t)d->num_attrs * 12 + (size_t)d->num_uv * 8 + nf3 * 4;
  if (!d->per_face) {
    CU_TRY(r, cudaMemcpyAsync((void*)dd.attr_indices, d->attr_indices, nf3 * 4, cudaMemcpyHostToDevice, st));
    r->stats.h2d_bytes += nf3 * 4;
  }
  if (d->keep_texture) {
    CU_TRY(r, cudaMemcpyAsync(r->uv_tex.p, texture_host, b_tex, cudaMemcpyHostToDevice, st));
    r->stats.h2d_bytes += b_tex;
  }
  if (mask_host && d->keep_mask) {
    CU_TRY(r, cudaMemcpyAsync(r->uv_mask.p, mask_host, P, cudaMemcpyHostToDevice, st));
    r->stats.h2d_bytes += P;
  }
  if ((rc = ugu_b200_uv_raster_device(r, &dd, r->uv_tex.p, mask_host ? (uint8_t*)r->uv_mask.p : nullptr, st)))
    return rc;
  std::vector<CopySeg> segs;
  segs.push_back({texture_host, r->uv_tex.p, b_tex});
  if (mask_host) segs.push_back({mask_host, r->uv_mask.p, P});
  return CopyOut(r, segs, st);
}

int ugu_b200_uv_mask(ugu_b200_rasterizer* r, int32_t num_faces, int32_t num_uv, const float* uv,
                     const int32_t* uv_indices, int32_t width, int32_t height, int32_t color,
                     uint8_t* mask_host) {
  if (!r) return Fail(nullptr, UGU_B200_E_INVALID, "rasterizer is NULL");
  if (!mask_host) return Fail(r, UGU_B200_E_INVALID, "NULL argument");
  if (width <= 0 || height <= 0) return Fail(r, UGU_B200_E_INVALID, "mask size %dx%d is invalid", width, height);
  // raster_util.cc:13-26: one constant attribute {c, c, c} rasterised as Image3b; channel 0 is the
  // mask.  (uint8_t)round((w0*c + w1*c) + w2*c) depends on the weights, so the texture path runs
  // as it is and the first channel is extracted on the host.
  const float c = (float)(uint8_t)color;
  const float attr[3] = {c, c, c};
  std::vector<int32_t> zero_idx((size_t)std::max(num_faces, 0) * 3, 0);
  std::vector<uint8_t> tex((size_t)width * height * 3);
  ugu_b200_uv_raster_desc d{};
  d.num_attrs = 1;
  d.num_faces = num_faces;
  d.num_uv = num_uv;
  d.attrs = attr;
  d.attr_indices = zero_idx.data();
  d.uv = uv;
  d.uv_indices = uv_indices;
  d.width = width;
  d.height = height;
  const int rc = ugu_b200_uv_raster(r, &d, tex.data(), nullptr);
  if (rc) return rc;
  const size_t P = (size_t)width * height;
  for (size_t i = 0; i < P; i++) mask_host[i] = tex[i * 3];
  return UGU_B200_OK;
}

int ugu_b200_face_visibility_device(ugu_b200_rasterizer* r, const int32_t* face_id_dev,
                                    int32_t num_views, int64_t pixels_per_view, int32_t first_view_id,
                                    int32_t num_faces, uint32_t* pixel_count_dev,
                                    uint32_t* view_count_dev, uint64_t* best_dev, void* stream) {
  if (!r) return Fail(nullptr, UGU_B200_E_INVALID, "rasterizer is NULL");
  if (!face_id_dev || !pixel_count_dev || !view_count_dev || !best_dev)
    return Fail(r, UGU_B200_E_INVALID, "NULL argument");
  if (num_views < 0 || pixels_per_view <= 0 || num_faces <= 0 || first_view_id < 0 ||
      pixels_per_view > (1ll << 40) || (long long)first_view_id + num_views > 0x7fffffffll)
    return Fail(r, UGU_B200_E_INVALID, "bad sizes");
  if (num_views == 0) return UGU_B200_OK;
  CU_TRY(r, cudaSetDevice(r->device));
  cudaStream_t st = stream ? (cudaStream_t)stream : r->stream;
  int rc;
  const size_t tmp_bytes = (size_t)kVisSlots * (size_t)num_faces * 4;
  if (r->vis_tmp.bytes < tmp_bytes || r->vis_tmp_faces != num_faces) {
    // pass 2 leaves the scratch zero, but only in the slot layout (stride = num_faces) it ran with
    if ((rc = Grow(r, &r->vis_tmp, tmp_bytes))) return rc;
    CU_TRY(r, cudaMemsetAsync(r->vis_tmp.p, 0, r->vis_tmp.bytes, st));
    r->vis_tmp_faces = num_faces;
  }
  VisParams p;
  p.face_id = face_id_dev;
  p.P = (unsigned long long)pixels_per_view;
  p.F = num_faces;
  p.first_view_id = first_view_id;
  p.tmp = (unsigned*)r->vis_tmp.p;
  p.pixel_count = pixel_count_dev;
  p.view_count = view_count_dev;
  p.best = (unsigned long long*)best_dev;
  const bool vec = (pixels_per_view % 4 == 0) && (((uintptr_t)face_id_dev & 15u) == 0);
  const unsigned gx = Blocks((size_t)((pixels_per_view + 3) / 4), 256);
  for (int v0 = 0; v0 < num_views; v0 += kVisSlots) {
    const dim3 grid(gx, (unsigned)std::min(kVisSlots, num_views - v0));
    if (vec) {
      vis_count_kernel<true><<<grid, 256, 0, st>>>(p, v0);
      vis_fold_kernel<true><<<grid, 256, 0, st>>>(p, v0);
    } else {
      vis_count_kernel<false><<<grid, 256, 0, st>>>(p, v0);
      vis_fold_kernel<false><<<grid, 256, 0, st>>>(p, v0);
    }
    r->stats.kernel_launches += 2;
  }
  CU_TRY(r, cudaGetLastError());
  return UGU_B200_OK;
}

int ugu_b200_face_visibility(ugu_b200_rasterizer* r, const int32_t* face_id_host, int32_t num_views,
                             int64_t pixels_per_view, int32_t first_view_id, int32_t num_faces,
                             uint32_t* pixel_count_host, uint32_t* view_count_host,
                             int32_t* best_view_host, uint32_t* best_count_host) {
  if (!r) return Fail(nullptr, UGU_B200_E_INVALID, "rasterizer is NULL");
  if (!face_id_host && num_views > 0) return Fail(r, UGU_B200_E_INVALID, "NULL argument");
  if (num_views < 0 || pixels_per_view <= 0 || num_faces <= 0 || first_view_id < 0)
    return Fail(r, UGU_B200_E_INVALID, "bad sizes");
  CU_TRY(r, cudaSetDevice(r->device));
  cudaStream_t st = r->stream;
  const size_t map_bytes = (size_t)num_views * (size_t)pixels_per_view * 4;
  const size_t F = (size_t)num_faces;
  int rc;
  if ((rc = Grow(r, &r->vis_in, map_bytes + 16))) return rc;
  if ((rc = Grow(r, &r->vis_acc, F * 16))) return rc;
  CU_TRY(r, cudaMemsetAsync(r->vis_acc.p, 0, F * 16, st));
  if (map_bytes) CU_TRY(r, cudaMemcpyAsync(r->vis_in.p, face_id_host, map_bytes, cudaMemcpyHostToDevice, st));
  r->stats.h2d_bytes += map_bytes;
  unsigned long long* best = (unsigned long long*)r->vis_acc.p;
  uint32_t* pix = (uint32_t*)(best + F);
  uint32_t* views = pix + F;
  if ((rc = ugu_b200_face_visibility_device(r, (const int32_t*)r->vis_in.p, num_views, pixels_per_view,
                                            first_view_id, num_faces, pix, views, (uint64_t*)best, st)))
    return rc;
  std::vector<unsigned long long> hb;
  if (best_view_host || best_count_host) hb.resize(F);
  std::vector<CopySeg> segs;
  if (pixel_count_host) segs.push_back({pixel_count_host, pix, F * 4});
  if (view_count_host) segs.push_back({view_count_host, views, F * 4});
  if (!hb.empty()) segs.push_back({hb.data(), best, F * 8});
  if ((rc = CopyOut(r, segs, st))) return rc;
  for (size_t f = 0; f < hb.size(); f++) {
    const unsigned long long b = hb[f];
    if (best_view_host) best_view_host[f] = b ? (int32_t)(0xFFFFFFFFu - (uint32_t)(b & 0xFFFFFFFFull)) : -1;
    if (best_count_host) best_count_host[f] = (uint32_t)(b >> 32);
  }
  return UGU_B200_OK;
}

// ---- keyframe visibility pass (ugu::VisibilityTester) -----------------------------------------------------------
namespace {

// device layout of the accumulators inside vt_acc (one allocation, 16-byte aligned slices)
struct VtAccLayout {
  size_t f_count, f_max_area, f_min_angle, f_min_dist, v_count, v_sum_color, v_sum_ad, v_min_angle, v_min_dist,
      v_min_colors, total;
};
VtAccLayout VtLayout(size_t V, size_t F) {
  VtAccLayout l{};
  size_t off = 0;
  auto take = [&](size_t bytes) {
    const size_t at = off;
    off += (bytes + 255) & ~size_t(255);
    return at;
  };
  l.f_count = take(F * 4);
  l.f_max_area = take(F * 8);
  l.f_min_angle = take(F * 8);
  l.f_min_dist = take(F * 8);
  l.v_count = take(V * 4);
  l.v_sum_color = take(V * 16);
  l.v_sum_ad = take(V * 8);
  l.v_min_angle = take(V * 8);
  l.v_min_dist = take(V * 8);
  l.v_min_colors = take(V * 8);
  l.total = off;
  return l;
}
VtAcc VtAccPointers(ugu_b200_rasterizer* r) {
  const VtAccLayout l = VtLayout((size_t)r->vt_V, (size_t)r->vt_F);
  char* b = (char*)r->vt_acc.p;
  VtAcc a;
  a.f_count = (unsigned*)(b + l.f_count);
  a.f_max_area = (unsigned long long*)(b + l.f_max_area);
  a.f_min_angle = (unsigned long long*)(b + l.f_min_angle);
  a.f_min_dist = (unsigned long long*)(b + l.f_min_dist);
  a.v_count = (unsigned*)(b + l.v_count);
  a.v_sum_color = (float4*)(b + l.v_sum_color);
  a.v_sum_ad = (float2*)(b + l.v_sum_ad);
  a.v_min_angle = (unsigned long long*)(b + l.v_min_angle);
  a.v_min_dist = (unsigned long long*)(b + l.v_min_dist);
  a.v_min_colors = (uint2*)(b + l.v_min_colors);
  return a;
}

int VtResetAcc(ugu_b200_rasterizer* r, cudaStream_t st) {
  const size_t V = (size_t)r->vt_V, F = (size_t)r->vt_F;
  const VtAccLayout l = VtLayout(V, F);
  CU_TRY(r, cudaMemsetAsync(r->vt_acc.p, 0, l.total, st));
  const VtAcc a = VtAccPointers(r);
  vt_fill_u64<<<Blocks(F, 256), 256, 0, st>>>(a.f_min_angle, F, ~0ull);
  vt_fill_u64<<<Blocks(F, 256), 256, 0, st>>>(a.f_min_dist, F, ~0ull);
  vt_fill_u64<<<Blocks(V, 256), 256, 0, st>>>(a.v_min_angle, V, ~0ull);
  vt_fill_u64<<<Blocks(V, 256), 256, 0, st>>>(a.v_min_dist, V, ~0ull);
  r->stats.kernel_launches += 4;
  CU_TRY(r, cudaGetLastError());
  return UGU_B200_OK;
}

int VtValidate(ugu_b200_rasterizer* r, int K, const ugu_b200_keyframe* kfs, int first_order,
               const ugu_b200_vis_option* opt) {
  if (r->vt_V == 0) return Fail(r, UGU_B200_E_STATE, "mesh has not been initialized");  // :283-286
  if (K <= 0 || !kfs) return Fail(r, UGU_B200_E_INVALID, "keyframe has not been set");  // :279-282
  if (!opt) return Fail(r, UGU_B200_E_INVALID, "option is NULL");
  if (first_order < 0 || (long long)first_order + K > 0x7fffffffll) return Fail(r, UGU_B200_E_INVALID, "bad order");
  if (opt->interp < 0 || opt->interp > 1) return Fail(r, UGU_B200_E_INVALID, "option enumerator out of range");
  for (int k = 0; k < K; k++) {
    const ugu_b200_keyframe& kf = kfs[k];
    if (kf.cam.width <= 0 || kf.cam.height <= 0 || (long long)kf.cam.width * kf.cam.height >= (1ll << 31))
      return Fail(r, UGU_B200_E_INVALID, "keyframe %d: camera size %dx%d is invalid", k, kf.cam.width, kf.cam.height);
    if (!kf.color) return Fail(r, UGU_B200_E_INVALID, "keyframe color is empty");  // :288-291
    if (opt->use_mask && !kf.mask) return Fail(r, UGU_B200_E_INVALID, "use_mask is true but keyframe mask is empty");
    if (opt->use_depth && !kf.depth) return Fail(r, UGU_B200_E_INVALID, "use_depth is true but keyframe depth is empty");
    if (opt->self_occlusion && !kf.zbuf && !r->mesh_ready)
      return Fail(r, UGU_B200_E_STATE, "self_occlusion without a zbuf needs the mesh uploaded to the rasterizer");
  }
  return UGU_B200_OK;
}

// Core: keyframes with DEVICE images; `rec` = DEVICE record arrays (any NULL).  Launch sets of kVtMaxKeyframes.
int VtRun(ugu_b200_rasterizer* r, int K, const ugu_b200_keyframe* kfs, int first_order, const ugu_b200_vis_option* o,
          const ugu_b200_vis_records* rec, cudaStream_t st, bool keep_face_rows = false) {
  const size_t V = (size_t)r->vt_V, F = (size_t)r->vt_F;
  int rc;
  // internal rows for the record arrays the caller does not want
  if (!rec || !rec->vertex_visible) if ((rc = Grow(r, &r->vt_vvis, kVtMaxKeyframes * V))) return rc;
  if (!rec || !rec->vertex) if ((rc = Grow(r, &r->vt_vrec, kVtMaxKeyframes * V * sizeof(VtVertexRec)))) return rc;
  if (!rec || !rec->face_visible) if ((rc = Grow(r, &r->vt_fvis, kVtMaxKeyframes * F))) return rc;
  if (!rec || !rec->face) if ((rc = Grow(r, &r->vt_frec, kVtMaxKeyframes * F * sizeof(VtFaceRec)))) return rc;
  if (o->facewise) {
    const size_t n = kVtMaxKeyframes * V;
    if ((rc = Grow(r, &r->vt_owner, n * 4))) return rc;
    if (r->vt_owner_clean < n) {
      vt_fill_u32<<<Blocks(n, 256), 256, 0, st>>>((unsigned*)r->vt_owner.p, n, kVtNoOwner);
      r->stats.kernel_launches++;
      r->vt_owner_clean = n;
    }
  }
  VtOption vo;
  vo.backface_culling_angle_rad_th = o->backface_culling_angle_rad_th;
  vo.use_mask = o->use_mask;
  vo.use_depth = o->use_depth;
  vo.interp = o->interp;
  vo.max_viewing_angle = o->max_viewing_angle;
  vo.max_distance = o->max_distance_from_camera_to_vertex;
  vo.max_depth_difference = o->max_depth_difference;
  vo.eps = o->eps;
  VtMesh m;
  m.V = r->vt_V;
  m.F = r->vt_F;
  m.pos = (const float*)r->vt_pos.p;
  m.nrm = (const float*)r->vt_nrm.p;
  m.idx = (const int*)r->vt_idx.p;
  const VtAcc acc = VtAccPointers(r);
  for (int k0 = 0; k0 < K; k0 += kVtMaxKeyframes) {
    const int nk = std::min((int)kVtMaxKeyframes, K - k0);
    // self depth maps the caller did not supply: one depth-only batch render with back-face culling off
    const float* zbuf[kVtMaxKeyframes] = {};
    if (o->self_occlusion) {
      std::vector<ugu_b200_camera> cams;
      std::vector<int> slot;
      for (int k = 0; k < nk; k++) {
        zbuf[k] = kfs[k0 + k].zbuf;
        if (!zbuf[k]) {
          cams.push_back(kfs[k0 + k].cam);
          slot.push_back(k);
        }
      }
      if (!cams.empty()) {
        const ugu_b200_option saved = r->opt;
        r->opt.diffuse_color = UGU_B200_COLOR_NONE;
        r->opt.diffuse_shading = UGU_B200_SHADING_NONE;
        r->opt.backface_culling = 0;
        size_t off = 0;
        std::vector<size_t> offs;
        for (const ugu_b200_camera& c : cams) {
          offs.push_back(off);
          off += (((size_t)c.width * c.height * 4) + 255) & ~size_t(255);
        }
        rc = Grow(r, &r->vt_zbuf, off + 256);
        for (size_t i = 0; i < cams.size() && !rc; i++) {
          ugu_b200_outputs d{};
          d.depth = (float*)((char*)r->vt_zbuf.p + offs[i]);
          // consecutive cameras of one size render as one launch set
          size_t j = i + 1;
          const size_t P = (size_t)cams[i].width * cams[i].height;
          while (j < cams.size() && cams[j].width == cams[i].width && cams[j].height == cams[i].height &&
                 (P * 4) % 256 == 0)
            j++;
          // (no Render validation: a depth-only pass needs positions and indices, nothing the options name)
          if (cams[i].width > 65535 || cams[i].height > 65535)
            rc = Fail(r, UGU_B200_E_INVALID, "camera size %dx%d is invalid", cams[i].width, cams[i].height);
          else
            rc = RenderViews(r, &cams[i], (int)(j - i), &d, st, std::min((int)(j - i), (int)kMaxBatch));
          i = j - 1;
        }
        r->opt = saved;
        if (rc) return rc;
        for (size_t i = 0; i < slot.size(); i++) zbuf[slot[i]] = (const float*)((char*)r->vt_zbuf.p + offs[i]);
      }
    }
    VtBatch kb;
    memset(&kb, 0, sizeof(kb));
    for (int k = 0; k < nk; k++) {
      const ugu_b200_keyframe& s = kfs[k0 + k];
      VtKeyframe& d = kb.k[k];
      d.cam.W = s.cam.width;
      d.cam.H = s.cam.height;
      d.cam.fx = s.cam.fx;
      d.cam.fy = s.cam.fy;
      d.cam.cx = s.cam.cx;
      d.cam.cy = s.cam.cy;
      memcpy(d.cam.w2c, s.w2c, sizeof(d.cam.w2c));
      memcpy(d.cam.c2w_R, s.cam.c2w_R, sizeof(d.cam.c2w_R));
      d.color = s.color;
      d.depth = s.depth;
      d.mask = s.mask;
      d.zbuf = zbuf[k];
      d.order = (unsigned)(first_order + k0 + k);
    }
    VtOut out;
    out.row0 = 0;
    out.v_visible = (rec && rec->vertex_visible) ? rec->vertex_visible + (size_t)k0 * V : (uint8_t*)r->vt_vvis.p;
    out.v_rec = (rec && rec->vertex) ? (VtVertexRec*)rec->vertex + (size_t)k0 * V : (VtVertexRec*)r->vt_vrec.p;
    out.f_visible = (rec && rec->face_visible) ? rec->face_visible + (size_t)k0 * F : (uint8_t*)r->vt_fvis.p;
    out.f_rec = (rec && rec->face) ? (VtFaceRec*)rec->face + (size_t)k0 * F : (VtFaceRec*)r->vt_frec.p;
    out.owner = (unsigned*)r->vt_owner.p;
    // no dense face records asked for: the face kernels fold into the accumulators themselves
    const bool fold_faces = !keep_face_rows && !(rec && (rec->face || rec->face_visible)) &&
                            !(getenv("UGU_B200_VT_FOLD") && atoi(getenv("UGU_B200_VT_FOLD")) == 0);
    out.fold_count = fold_faces ? acc.f_count : nullptr;
    out.fold_max_area = acc.f_max_area;
    out.fold_min_angle = acc.f_min_angle;
    out.fold_min_dist = acc.f_min_dist;
    const dim3 gv(Blocks(V, 256), (unsigned)nk), gf(Blocks(F, 256), (unsigned)nk);
    if (o->facewise) {
      vt_facewise_kernel<<<gf, 256, 0, st>>>(kb, vo, m, out);
      vt_owner_kernel<<<gv, 256, 0, st>>>(kb, vo, m, out);
    } else {
      vt_vertex_kernel<<<gv, 256, 0, st>>>(kb, vo, m, out);
      vt_face_kernel<<<gf, 256, 0, st>>>(kb, m, out);
    }
    if (!fold_faces) vt_fold_face_kernel<<<Blocks(F, 256), 256, 0, st>>>(kb, nk, m, out, acc);
    vt_fold_vertex_kernel<<<Blocks(V, 256), 256, 0, st>>>(kb, nk, m, out, acc);
    r->stats.kernel_launches += fold_faces ? 3 : 4;
    CU_TRY(r, cudaGetLastError());
  }
  return UGU_B200_OK;
}

}  // namespace

int ugu_b200_visibility_set_data(ugu_b200_rasterizer* r, int32_t V, const float* vertices, const float* normals,
                                 int32_t F, const int32_t* indices) {
  if (!r) return Fail(nullptr, UGU_B200_E_INVALID, "rasterizer is NULL");
  r->vt_V = r->vt_F = 0;
  if (V <= 0 || F <= 0 || !vertices || !normals || !indices) return Fail(r, UGU_B200_E_INVALID, "mesh is empty");  // :237-240
  if (!IndicesInRange(indices, (size_t)F * 3, V))
    return Fail(r, UGU_B200_E_INVALID, "indices reference a vertex outside [0, %d)", V);
  if ((unsigned long long)F >= (1ull << 30)) return Fail(r, UGU_B200_E_INVALID, "too many faces");
  CU_TRY(r, cudaSetDevice(r->device));
  cudaStream_t st = r->stream;
  int rc;
  if ((rc = Grow(r, &r->vt_pos, (size_t)V * 12))) return rc;
  if ((rc = Grow(r, &r->vt_nrm, (size_t)V * 12))) return rc;
  if ((rc = Grow(r, &r->vt_idx, (size_t)F * 12))) return rc;
  if ((rc = Grow(r, &r->vt_acc, VtLayout((size_t)V, (size_t)F).total))) return rc;
  CU_TRY(r, cudaMemcpyAsync(r->vt_pos.p, vertices, (size_t)V * 12, cudaMemcpyHostToDevice, st));
  CU_TRY(r, cudaMemcpyAsync(r->vt_nrm.p, normals, (size_t)V * 12, cudaMemcpyHostToDevice, st));
  CU_TRY(r, cudaMemcpyAsync(r->vt_idx.p, indices, (size_t)F * 12, cudaMemcpyHostToDevice, st));
  r->stats.h2d_bytes += (size_t)V * 24 + (size_t)F * 12;
  r->vt_V = V;
  r->vt_F = F;
  r->vt_owner_clean = 0;
  if ((rc = VtResetAcc(r, st))) return rc;
  CU_TRY(r, cudaStreamSynchronize(st));  // the host arrays may go away
  return UGU_B200_OK;
}

int ugu_b200_visibility_reset(ugu_b200_rasterizer* r) {
  if (!r) return Fail(nullptr, UGU_B200_E_INVALID, "rasterizer is NULL");
  if (r->vt_V == 0) return Fail(r, UGU_B200_E_STATE, "mesh has not been initialized");
  CU_TRY(r, cudaSetDevice(r->device));
  return VtResetAcc(r, r->stream);
}

int ugu_b200_visibility_test_device(ugu_b200_rasterizer* r, int32_t K, const ugu_b200_keyframe* kfs,
                                    int32_t first_order, const ugu_b200_vis_option* opt,
                                    const ugu_b200_vis_records* rec, void* stream) {
  if (!r) return Fail(nullptr, UGU_B200_E_INVALID, "rasterizer is NULL");
  int rc = VtValidate(r, K, kfs, first_order, opt);
  if (rc) return rc;
  if (rec && ((uintptr_t)rec->vertex & 15u)) return Fail(r, UGU_B200_E_INVALID, "vertex records must be 16-byte aligned");
  CU_TRY(r, cudaSetDevice(r->device));
  cudaStream_t st = stream ? (cudaStream_t)stream : r->stream;
  if ((rc = AcquireScratch(r, st))) return rc;
  if ((rc = VtRun(r, K, kfs, first_order, opt, rec, st))) return rc;
  return ReleaseScratch(r, st);
}

int ugu_b200_visibility_test(ugu_b200_rasterizer* r, int32_t K, const ugu_b200_keyframe* kfs, int32_t first_order,
                             const ugu_b200_vis_option* opt, const ugu_b200_vis_records* rec) {
  if (!r) return Fail(nullptr, UGU_B200_E_INVALID, "rasterizer is NULL");
  int rc = VtValidate(r, K, kfs, first_order, opt);
  if (rc) return rc;
  CU_TRY(r, cudaSetDevice(r->device));
  cudaStream_t st = r->stream;
  if ((rc = AcquireScratch(r, st))) return rc;
  const size_t V = (size_t)r->vt_V, F = (size_t)r->vt_F;
  const bool want = rec && (rec->vertex_visible || rec->vertex || rec->face_visible || rec->face);
  // keyframes go through the device in launch sets: upload the set's images, run, copy its records back
  for (int k0 = 0; k0 < K; k0 += kVtMaxKeyframes) {
    const int nk = std::min((int)kVtMaxKeyframes, K - k0);
    size_t bytes = 0;
    for (int k = 0; k < nk; k++) {
      const size_t P = (size_t)kfs[k0 + k].cam.width * kfs[k0 + k].cam.height;
      bytes += ((P * 3 + 255) & ~size_t(255)) + ((P * 4 + 255) & ~size_t(255)) * 2 + ((P + 255) & ~size_t(255));
    }
    if ((rc = Grow(r, &r->vt_img, bytes))) return rc;
    std::vector<ugu_b200_keyframe> dk(kfs + k0, kfs + k0 + nk);
    size_t off = 0;
    auto up = [&](const void* host, size_t n) -> const void* {
      if (!host) return nullptr;
      void* d = (char*)r->vt_img.p + off;
      off += (n + 255) & ~size_t(255);
      if (cudaMemcpyAsync(d, host, n, cudaMemcpyHostToDevice, st) != cudaSuccess) return nullptr;
      r->stats.h2d_bytes += n;
      return d;
    };
    for (int k = 0; k < nk; k++) {
      const size_t P = (size_t)dk[k].cam.width * dk[k].cam.height;
      dk[k].color = (const uint8_t*)up(kfs[k0 + k].color, P * 3);
      dk[k].depth = (const float*)up(opt->use_depth ? kfs[k0 + k].depth : nullptr, P * 4);
      dk[k].mask = (const uint8_t*)up(opt->use_mask ? kfs[k0 + k].mask : nullptr, P);
      dk[k].zbuf = (const float*)up(opt->self_occlusion ? kfs[k0 + k].zbuf : nullptr, P * 4);
    }
    CU_TRY(r, cudaGetLastError());
    // (the rows of the face records are copied out of the internal buffers below: keep them when asked for)
    if ((rc = VtRun(r, nk, dk.data(), first_order + k0, opt, nullptr, st, want && (rec->face_visible || rec->face))))
      return rc;
    if (want) {
      std::vector<CopySeg> segs;
      if (rec->vertex_visible) segs.push_back({rec->vertex_visible + (size_t)k0 * V, r->vt_vvis.p, (size_t)nk * V});
      if (rec->vertex) segs.push_back({rec->vertex + (size_t)k0 * V, r->vt_vrec.p, (size_t)nk * V * sizeof(VtVertexRec)});
      if (rec->face_visible) segs.push_back({rec->face_visible + (size_t)k0 * F, r->vt_fvis.p, (size_t)nk * F});
      if (rec->face) segs.push_back({rec->face + (size_t)k0 * F, r->vt_frec.p, (size_t)nk * F * sizeof(VtFaceRec)});
      if ((rc = CopyOut(r, segs, st))) return rc;
    } else {
      CU_TRY(r, cudaStreamSynchronize(st));  // the image scratch is reused by the next set
    }
  }
  return ReleaseScratch(r, st);
}

int ugu_b200_visibility_accumulators(ugu_b200_rasterizer* r, ugu_b200_vis_accumulators* out) {
  if (!r || !out) return Fail(r, UGU_B200_E_INVALID, "NULL argument");
  if (r->vt_V == 0) return Fail(r, UGU_B200_E_STATE, "mesh has not been initialized");
  const VtAcc a = VtAccPointers(r);
  out->face_count = a.f_count;
  out->face_max_area = (uint64_t*)a.f_max_area;
  out->face_min_angle = (uint64_t*)a.f_min_angle;
  out->face_min_dist = (uint64_t*)a.f_min_dist;
  out->vertex_count = a.v_count;
  out->vertex_sum_color = (float*)a.v_sum_color;
  out->vertex_sum_angle_dist = (float*)a.v_sum_ad;
  out->vertex_min_angle = (uint64_t*)a.v_min_angle;
  out->vertex_min_dist = (uint64_t*)a.v_min_dist;
  out->vertex_min_colors = (uint32_t*)a.v_min_colors;
  return UGU_B200_OK;
}

int ugu_b200_visibility_stats(ugu_b200_rasterizer* r, const ugu_b200_vis_stats* o) {
  if (!r || !o) return Fail(r, UGU_B200_E_INVALID, "NULL argument");
  if (r->vt_V == 0) return Fail(r, UGU_B200_E_STATE, "mesh has not been initialized");
  CU_TRY(r, cudaSetDevice(r->device));
  const size_t V = (size_t)r->vt_V, F = (size_t)r->vt_F;
  const VtAccLayout l = VtLayout(V, F);
  std::vector<char> host(l.total);
  if (r->scratch_used) CU_TRY(r, cudaEventSynchronize(r->ev_scratch));
  CU_TRY(r, cudaMemcpy(host.data(), r->vt_acc.p, l.total, cudaMemcpyDeviceToHost));
  r->stats.d2h_bytes += l.total;
  const unsigned* fc = (const unsigned*)(host.data() + l.f_count);
  const unsigned long long* fa = (const unsigned long long*)(host.data() + l.f_max_area);
  const unsigned long long* fg = (const unsigned long long*)(host.data() + l.f_min_angle);
  const unsigned long long* fd = (const unsigned long long*)(host.data() + l.f_min_dist);
  auto f32 = [](unsigned long long key) {
    const unsigned u = (unsigned)(key >> 32);
    float f;
    memcpy(&f, &u, 4);
    return f;
  };
  for (size_t f = 0; f < F; f++) {
    const bool any = fc[f] != 0;
    if (o->face_count) o->face_count[f] = fc[f];
    if (o->face_max_area_order) o->face_max_area_order[f] = any ? (int32_t)(0xFFFFFFFFu - (unsigned)fa[f]) : -1;
    if (o->face_min_viewing_angle_order) o->face_min_viewing_angle_order[f] = any ? (int32_t)(unsigned)fg[f] : -1;
    if (o->face_min_distance_order) o->face_min_distance_order[f] = any ? (int32_t)(unsigned)fd[f] : -1;
    if (o->face_max_area) o->face_max_area[f] = any ? f32(fa[f]) : 0.f;
    if (o->face_min_viewing_angle) o->face_min_viewing_angle[f] = any ? f32(fg[f]) : 0.f;
    if (o->face_min_distance) o->face_min_distance[f] = any ? f32(fd[f]) : 0.f;
  }
  const unsigned* vc = (const unsigned*)(host.data() + l.v_count);
  const float* sc = (const float*)(host.data() + l.v_sum_color);
  const float* sad = (const float*)(host.data() + l.v_sum_ad);
  const unsigned long long* vg = (const unsigned long long*)(host.data() + l.v_min_angle);
  const unsigned long long* vd = (const unsigned long long*)(host.data() + l.v_min_dist);
  const unsigned* mc = (const unsigned*)(host.data() + l.v_min_colors);
  for (size_t v = 0; v < V; v++) {
    const bool any = vc[v] != 0;
    const float n = (float)vc[v];
    if (o->vertex_count) o->vertex_count[v] = vc[v];
    if (o->vertex_min_viewing_angle_order) o->vertex_min_viewing_angle_order[v] = any ? (int32_t)(unsigned)vg[v] : -1;
    if (o->vertex_min_distance_order) o->vertex_min_distance_order[v] = any ? (int32_t)(unsigned)vd[v] : -1;
    for (int c = 0; c < 3; c++) {
      // mean_color /= static_cast<float>(size) (visibility_tester.cc:80)
      if (o->vertex_mean_color) o->vertex_mean_color[v * 3 + c] = any ? sc[v * 4 + c] / n : 0.f;
      if (o->vertex_min_viewing_angle_color)
        o->vertex_min_viewing_angle_color[v * 3 + c] = any ? (float)((mc[v * 2] >> (8 * c)) & 255u) : 0.f;
      if (o->vertex_min_distance_color)
        o->vertex_min_distance_color[v * 3 + c] = any ? (float)((mc[v * 2 + 1] >> (8 * c)) & 255u) : 0.f;
    }
    if (o->vertex_mean_viewing_angle) o->vertex_mean_viewing_angle[v] = any ? sad[v * 2] / n : 0.f;
    if (o->vertex_mean_distance) o->vertex_mean_distance[v] = any ? sad[v * 2 + 1] / n : 0.f;
  }
  return UGU_B200_OK;
}

namespace {
// `cc` != NULL: the two-camera overload (rgbd_util.h:39-48; Depth2MeshImpl, rgbd_util.cc:283-466)
int D2MDevice(ugu_b200_rasterizer* r, const ugu_b200_camera* cam, const ugu_b200_color_camera* cc,
              const float* depth_dev, const uint8_t* color_dev, int32_t color_width, int32_t color_height,
              int32_t with_vertex_color, float max_connect_z_diff, int32_t x_step, int32_t y_step,
              int32_t gl_coord, void* stream) {
  if (!r) return Fail(nullptr, UGU_B200_E_INVALID, "rasterizer is NULL");
  if (!cam || !depth_dev) return Fail(r, UGU_B200_E_INVALID, "NULL argument");
  if (cc && !color_dev) return Fail(r, UGU_B200_E_INVALID, "the two-camera form needs the colour image");
  r->d2m_valid = false;
  // rgbd_util.cc:100-112 (the reference's messages)
  if (max_connect_z_diff < 0) return Fail(r, UGU_B200_E_INVALID, "Depth2Mesh max_connect_z_diff must be positive %f", max_connect_z_diff);
  if (x_step < 1) return Fail(r, UGU_B200_E_INVALID, "Depth2Mesh x_step must be positive %d", x_step);
  if (y_step < 1) return Fail(r, UGU_B200_E_INVALID, "Depth2Mesh y_step must be positive %d", y_step);
  const int W = cam->width, H = cam->height;
  if (W <= 0 || H <= 0 || (long long)W * H >= (1ll << 28))
    return Fail(r, UGU_B200_E_INVALID, "depth size %dx%d is invalid", W, H);
  if (color_dev) {  // rgbd_util.cc:121-136
    if (color_width <= 0 || color_height <= 0) return Fail(r, UGU_B200_E_INVALID, "colour size is invalid");
    const float da = (float)W / (float)H, ca = (float)color_width / (float)color_height;
    if (!cc && std::fabs(da - ca) > 0.01f)  // (the two-camera overload does not check)
      return Fail(r, UGU_B200_E_INVALID, "Depth2Mesh depth aspect ratio %f and color aspect ratio %f are very different", da, ca);
  }
  CU_TRY(r, cudaSetDevice(r->device));
  cudaStream_t st = stream ? (cudaStream_t)stream : r->stream;
  D2MParams p{};
  p.W = W; p.H = H; p.xs = x_step; p.ys = y_step;
  p.GX = (W - 1) / x_step;
  p.GY = (H - 1) / y_step;
  p.fx = cam->fx; p.fy = cam->fy; p.cx = cam->cx; p.cy = cam->cy;
  p.th = max_connect_z_diff;
  p.gl = gl_coord ? 1 : 0;
  p.with_uv = color_dev ? 1 : 0;
  p.with_color = (color_dev && with_vertex_color) ? 1 : 0;
  p.cw = color_width; p.ch = color_height;
  p.gx_magic = 0;
  p.gx_shift = 0;
  if (p.GX > 1) {  // Granlund-Montgomery, N = 31, as MakeViewParams does for W
    unsigned l = 0;
    while ((1ull << l) < (unsigned long long)p.GX) l++;
    p.gx_magic = (unsigned)(((1ull << (31 + l)) + (unsigned long long)p.GX - 1) / (unsigned long long)p.GX);
    p.gx_shift = l - 1;
  }
  p.two_cam = cc ? 1 : 0;
  if (cc) {
    for (int i = 0; i < 9; i++) p.d2c[i] = cc->depth2color_R[i];
    for (int i = 0; i < 3; i++) p.d2c[9 + i] = cc->depth2color_t[i];
    p.cfx = cc->fx; p.cfy = cc->fy; p.ccx = cc->cx; p.ccy = cc->cy;
  }
  p.depth = depth_dev;
  p.color = color_dev;
  const size_t G = (size_t)p.GX * (size_t)p.GY, P = (size_t)W * H;
  const int nblocks = (int)((G + kD2MBlock - 1) / kD2MBlock);
  int rc;
  // one block of small words, zeroed by ONE memset per call: [0] first (inverted), [16] totals, [32] ticket, [64..] the
  // look-back state of the single-pass kernel
  if ((rc = Grow(r, &r->d2m_small, 64 + (size_t)std::max(nblocks, 1) * 8))) return rc;
  if ((rc = Grow(r, &r->d2m_block_sums, (size_t)std::max(nblocks, 1) * 8))) return rc;
  if ((rc = Grow(r, &r->d2m_vid, std::max<size_t>(G, 1) * 4))) return rc;
  if ((rc = Grow(r, &r->d2m_fref, std::max<size_t>(G, 1) * 4))) return rc;
  if ((rc = Grow(r, &r->d2m_vertices, std::max<size_t>(G, 1) * 12))) return rc;
  if ((rc = Grow(r, &r->d2m_normals, std::max<size_t>(G, 1) * 12))) return rc;
  if ((rc = Grow(r, &r->d2m_faces, std::max<size_t>(G, 1) * 24))) return rc;
  if ((rc = Grow(r, &r->d2m_face_normals, std::max<size_t>(G, 1) * 24))) return rc;
  if (p.with_uv && (rc = Grow(r, &r->d2m_uv, std::max<size_t>(G, 1) * 8))) return rc;
  if (p.with_color && (rc = Grow(r, &r->d2m_colors, std::max<size_t>(G, 1) * 12))) return rc;
  if ((rc = Grow(r, &r->d2m_pc, P * 12))) return rc;
  if ((rc = Grow(r, &r->d2m_nimg, P * 12))) return rc;
  if (!r->d2m_totals_host) CU_TRY(r, cudaHostAlloc((void**)&r->d2m_totals_host, 16, cudaHostAllocDefault));
  p.first = (int*)r->d2m_small.p;
  p.totals = (unsigned long long*)((char*)r->d2m_small.p + 16);
  p.block_sums = (unsigned long long*)r->d2m_block_sums.p;
  p.vid = (int*)r->d2m_vid.p;
  p.fref = (unsigned*)r->d2m_fref.p;
  p.vertices = (float*)r->d2m_vertices.p;
  p.faces = (int*)r->d2m_faces.p;
  p.normals = (float*)r->d2m_normals.p;
  p.face_normals = (float*)r->d2m_face_normals.p;
  p.uv = (float*)r->d2m_uv.p;
  p.colors = (float*)r->d2m_colors.p;
  p.point_cloud = (float*)r->d2m_pc.p;
  p.normal_image = (float*)r->d2m_nimg.p;
  // Single pass (d2m_scan_emit_kernel: count + scan + vertex emission behind a decoupled look-back) with the three
  // kernels behind d2m_first_kernel chained by programmatic dependent launch; UGU_B200_D2M_FUSED=0 selects the
  // three-kernel count / scan / emit form (A/B).  With steps (1, 1) every pixel has a grid point beside it and the
  // kernels write the zeros of the two organized images themselves instead of two 12 B/pixel memsets.
  static const bool single_pass = !(getenv("UGU_B200_D2M_FUSED") && atoi(getenv("UGU_B200_D2M_FUSED")) == 0);
  p.state = (unsigned long long*)((char*)r->d2m_small.p + 64);
  p.ticket = (int*)((char*)r->d2m_small.p + 32);
  p.zero_fill = (single_pass && x_step == 1 && y_step == 1 && G > 0) ? 1 : 0;
  CU_TRY(r, cudaMemsetAsync(r->d2m_small.p, 0, 64 + (single_pass ? (size_t)std::max(nblocks, 1) * 8 : 0), st));
  if (!p.zero_fill) {
    CU_TRY(r, cudaMemsetAsync(r->d2m_pc.p, 0, P * 12, st));            // ugu::Init(point_cloud, .., 0.0f)
    CU_TRY(r, cudaMemsetAsync(r->d2m_nimg.p, 0, P * 12, st));
  }
  if (G > 0) {
    const unsigned first_grid = (unsigned)std::min<size_t>((size_t)nblocks, (size_t)g_num_sms * 8);
    if (single_pass) {
      d2m_first_kernel<<<first_grid, 256, 0, st>>>(p);
      const bool pdl = r->use_pdl;
      CU_TRY(r, LaunchView(d2m_scan_emit_kernel, dim3((unsigned)nblocks), 256, st, pdl, p, nblocks));
      // (faces + vertex normals fused over 2-D tiles with the face normals in shared memory: built, bit-exact, no
      //  faster — 45.5 us against 25 + 19)
      CU_TRY(r, LaunchView(d2m_emit_faces_kernel, dim3(Blocks(G, 256)), 256, st, pdl, p));
      CU_TRY(r, LaunchView(d2m_vertex_normals_kernel, dim3(Blocks(G, 256)), 256, st, pdl, p));
      r->stats.kernel_launches += 4;
    } else {
      d2m_first_kernel<<<first_grid, 256, 0, st>>>(p);
      d2m_count_kernel<<<(unsigned)nblocks, 256, 0, st>>>(p);
      d2m_scan_kernel<<<1, 256, 0, st>>>(p, nblocks);
      d2m_emit_vertices_kernel<<<(unsigned)nblocks, 256, 0, st>>>(p);
      d2m_emit_faces_kernel<<<Blocks(G, 256), 256, 0, st>>>(p);
      d2m_vertex_normals_kernel<<<Blocks(G, 256), 256, 0, st>>>(p);
      r->stats.kernel_launches += 6;
    }
  }
  CU_TRY(r, cudaGetLastError());
  CU_TRY(r, cudaMemcpyAsync(r->d2m_totals_host, p.totals, 8, cudaMemcpyDeviceToHost, st));
  r->d2m_stream = st;
  r->d2m_w = W;
  r->d2m_h = H;
  r->d2m_has_uv = p.with_uv;
  r->d2m_has_colors = p.with_color;
  r->d2m_valid = true;
  return UGU_B200_OK;
}
}  // namespace

int ugu_b200_depth_to_mesh_device(ugu_b200_rasterizer* r, const ugu_b200_camera* cam,
                                  const float* depth_dev, const uint8_t* color_dev, int32_t color_width,
                                  int32_t color_height, int32_t with_vertex_color,
                                  float max_connect_z_diff, int32_t x_step, int32_t y_step,
                                  int32_t gl_coord, void* stream) {
  return D2MDevice(r, cam, nullptr, depth_dev, color_dev, color_width, color_height, with_vertex_color,
                   max_connect_z_diff, x_step, y_step, gl_coord, stream);
}

int ugu_b200_depth_to_mesh_two_camera_device(ugu_b200_rasterizer* r, const ugu_b200_camera* depth_cam,
                                             const ugu_b200_color_camera* color_cam, const float* depth_dev,
                                             const uint8_t* color_dev, int32_t color_width, int32_t color_height,
                                             int32_t with_vertex_color, float max_connect_z_diff, int32_t x_step,
                                             int32_t y_step, int32_t gl_coord, void* stream) {
  if (r && !color_cam) return Fail(r, UGU_B200_E_INVALID, "NULL argument");
  return D2MDevice(r, depth_cam, color_cam, depth_dev, color_dev, color_width, color_height, with_vertex_color,
                   max_connect_z_diff, x_step, y_step, gl_coord, stream);
}

namespace {
int D2MCounts(ugu_b200_rasterizer* r, ugu_b200_depth2mesh_counts* c) {
  if (!r->d2m_valid) return Fail(r, UGU_B200_E_STATE, "no Depth2Mesh result to fetch");
  CU_TRY(r, cudaSetDevice(r->device));
  CU_TRY(r, cudaStreamSynchronize(r->d2m_stream));
  const unsigned long long t = r->d2m_totals_host[0];
  c->num_vertices = (int32_t)(t & 0xFFFFFFFFull);
  c->num_faces = (int32_t)(t >> 32);
  c->has_normals = c->num_faces > 0 ? 1 : 0;
  c->has_uv = r->d2m_has_uv;
  c->has_colors = r->d2m_has_colors;
  c->reserved = 0;
  return UGU_B200_OK;
}
}  // namespace

namespace {
int D2MHost(ugu_b200_rasterizer* r, const ugu_b200_camera* cam, const ugu_b200_color_camera* cc,
            const float* depth_host, const uint8_t* color_host, int32_t color_width, int32_t color_height,
            int32_t with_vertex_color, float max_connect_z_diff, int32_t x_step, int32_t y_step, int32_t gl_coord,
            ugu_b200_depth2mesh_counts* counts) {
  if (!r) return Fail(nullptr, UGU_B200_E_INVALID, "rasterizer is NULL");
  if (!cam || !depth_host || !counts) return Fail(r, UGU_B200_E_INVALID, "NULL argument");
  if (cam->width <= 0 || cam->height <= 0) return Fail(r, UGU_B200_E_INVALID, "depth size is invalid");
  if (color_host && (color_width <= 0 || color_height <= 0)) return Fail(r, UGU_B200_E_INVALID, "colour size is invalid");
  CU_TRY(r, cudaSetDevice(r->device));
  const size_t P = (size_t)cam->width * cam->height;
  int rc;
  if ((rc = Grow(r, &r->d2m_in_depth, P * 4))) return rc;
  CU_TRY(r, cudaMemcpyAsync(r->d2m_in_depth.p, depth_host, P * 4, cudaMemcpyHostToDevice, r->stream));
  r->stats.h2d_bytes += P * 4;
  const uint8_t* cdev = nullptr;
  if (color_host) {
    const size_t cb = (size_t)color_width * color_height * 3;
    if ((rc = Grow(r, &r->d2m_in_color, cb))) return rc;
    CU_TRY(r, cudaMemcpyAsync(r->d2m_in_color.p, color_host, cb, cudaMemcpyHostToDevice, r->stream));
    r->stats.h2d_bytes += cb;
    cdev = (const uint8_t*)r->d2m_in_color.p;
  }
  if ((rc = D2MDevice(r, cam, cc, (const float*)r->d2m_in_depth.p, cdev, color_width, color_height,
                      with_vertex_color, max_connect_z_diff, x_step, y_step, gl_coord, r->stream)))
    return rc;
  return D2MCounts(r, counts);
}
}  // namespace

int ugu_b200_depth_to_mesh(ugu_b200_rasterizer* r, const ugu_b200_camera* cam, const float* depth_host,
                           const uint8_t* color_host, int32_t color_width, int32_t color_height,
                           int32_t with_vertex_color, float max_connect_z_diff, int32_t x_step,
                           int32_t y_step, int32_t gl_coord, ugu_b200_depth2mesh_counts* counts) {
  return D2MHost(r, cam, nullptr, depth_host, color_host, color_width, color_height, with_vertex_color,
                 max_connect_z_diff, x_step, y_step, gl_coord, counts);
}

int ugu_b200_depth_to_mesh_two_camera(ugu_b200_rasterizer* r, const ugu_b200_camera* depth_cam,
                                      const ugu_b200_color_camera* color_cam, const float* depth_host,
                                      const uint8_t* color_host, int32_t color_width, int32_t color_height,
                                      int32_t with_vertex_color, float max_connect_z_diff, int32_t x_step,
                                      int32_t y_step, int32_t gl_coord, ugu_b200_depth2mesh_counts* counts) {
  if (r && (!color_cam || !color_host)) return Fail(r, UGU_B200_E_INVALID, "NULL argument");
  return D2MHost(r, depth_cam, color_cam, depth_host, color_host, color_width, color_height, with_vertex_color,
                 max_connect_z_diff, x_step, y_step, gl_coord, counts);
}

int ugu_b200_depth_to_mesh_fetch(ugu_b200_rasterizer* r, ugu_b200_depth2mesh_counts* counts,
                                 float* vertices, int32_t* faces, float* normals, float* face_normals,
                                 float* uv, float* colors, float* point_cloud, float* normal_image) {
  if (!r) return Fail(nullptr, UGU_B200_E_INVALID, "rasterizer is NULL");
  ugu_b200_depth2mesh_counts c;
  int rc = D2MCounts(r, &c);
  if (rc) return rc;
  if (counts) *counts = c;
  const size_t nv = (size_t)c.num_vertices, nf = (size_t)c.num_faces, P = (size_t)r->d2m_w * r->d2m_h;
  std::vector<CopySeg> segs;
  if (vertices && nv) segs.push_back({vertices, r->d2m_vertices.p, nv * 12});
  if (faces && nf) segs.push_back({faces, r->d2m_faces.p, nf * 12});
  if (normals && nv && c.has_normals) segs.push_back({normals, r->d2m_normals.p, nv * 12});
  if (face_normals && nf) segs.push_back({face_normals, r->d2m_face_normals.p, nf * 12});
  if (uv && nv && c.has_uv) segs.push_back({uv, r->d2m_uv.p, nv * 8});
  if (colors && nv && c.has_colors) segs.push_back({colors, r->d2m_colors.p, nv * 12});
  if (point_cloud) segs.push_back({point_cloud, r->d2m_pc.p, P * 12});
  if (normal_image) segs.push_back({normal_image, r->d2m_nimg.p, P * 12});
  return CopyOut(r, segs, r->d2m_stream);
}

int ugu_b200_depth_to_u16(const float* depth, uint16_t* depth_w, size_t count) {
  if (!depth || !depth_w) return Fail(nullptr, UGU_B200_E_INVALID, "NULL argument");
  // ConvertTo(f_depth, depth) with scale 1 (image_proc.h:453-460): static_cast<uint16_t>(float),
  // compiled on x86-64 as cvttss2si + low 16 bits.
  for (size_t i = 0; i < count; i++) {
    const float f = depth[i];
    int32_t v = (f > -2147483904.0f && f < 2147483648.0f) ? (int32_t)f : INT32_MIN;
    depth_w[i] = (uint16_t)(uint32_t)v;
  }
  return UGU_B200_OK;
}

int ugu_b200_rasterizer_set_profiling(ugu_b200_rasterizer* r, int32_t enable, uint32_t stage_mask) {
  if (!r) return Fail(nullptr, UGU_B200_E_INVALID, "rasterizer is NULL");
  CU_TRY(r, cudaSetDevice(r->device));
  int rc = DrainEvents(r);
  if (rc) return rc;
  r->profiling = enable != 0;
  r->stage_mask = stage_mask;
  return UGU_B200_OK;
}

int ugu_b200_rasterizer_get_stats(ugu_b200_rasterizer* r, ugu_b200_stats* out) {
  if (!r || !out) return Fail(r, UGU_B200_E_INVALID, "NULL argument");
  CU_TRY(r, cudaSetDevice(r->device));
  int rc = DrainEvents(r);
  if (rc) return rc;
  *out = r->stats;
  return UGU_B200_OK;
}

int ugu_b200_rasterizer_reset_stats(ugu_b200_rasterizer* r) {
  if (!r) return Fail(nullptr, UGU_B200_E_INVALID, "rasterizer is NULL");
  CU_TRY(r, cudaSetDevice(r->device));
  int rc = DrainEvents(r);
  if (rc) return rc;
  r->stats = ugu_b200_stats{};
  return UGU_B200_OK;
}

void* ugu_b200_host_alloc(size_t bytes) {
  void* p = nullptr;
  if (cudaHostAlloc(&p, std::max<size_t>(bytes, 1), cudaHostAllocPortable) != cudaSuccess) {
    Fail(nullptr, UGU_B200_E_NOMEM, "cudaHostAlloc(%zu) failed: %s", bytes,
         cudaGetErrorString(cudaGetLastError()));
    return nullptr;
  }
  return p;
}

void ugu_b200_host_free(void* p) {
  if (p) cudaFreeHost(p);
}

}  // extern "C"
