// G-buffer attribute kernel: the fields of ugu::GBuffer (include/ugu/renderer/base.h:
// 34-77) that the reference's OpenGL renderer reads back in RendererGl::ReadGbuf
// (src/renderer/gl/renderer.cc:501-650) from the outputs of its shaders
// (src/shader/glsl/vert/gbuf.vert:42-56, geom/dummy.geom:26-80, frag/gbuf.frag:21-44) — computed here
// from the winners of THIS rasterizer's visibility buffer, i.e. with the CPU Rasterizer's sampling
// convention (integer pixel centres, inclusive edges, rasterizer.cc:194-237), not GL's.  The GL path
// itself cannot run on this box (no graphics stack), so the definition below is a restatement of the
// shader / read-back math with a fixed float op order; the oracle is oracle_gbuffer
// (oracle/raster_oracle.c), parity with the reference's GL output is UNPINNED.
//
// Per pixel with a winner (face f, depth z, perspective-correct weights w0 w1 w2 of rasterizer.cc:
// 208-224; a culled back-face winner counts as no winner when backface_culling):
//   ndc      = (A * (-z) + B) / z        A = -(far + near) / (far - near), B = -(2 far near) / (far - near)
//                                        (GetProjectionMatrixOpenGlForPinhole, camera_util.cc:254-257)
//   d        = 0.5 * ndc + 0.5           window depth, glDepthRange(0, 1)
//   stencil  = (d <= 0 || 1 <= d) ? 0 : 255,  depth_01 = clamp(d, 0, 1)          renderer.cc:531-541
//   -- everything below only where stencil == 255 (renderer.cc:552,574,...), zeros elsewhere --
//   u = w1, v = w2, k0 = 1 - u - v       the (1,0) / (0,1) corner attribute of dummy.geom:69-75
//   bary     = (u, v, 1 - u - v)                                                 renderer.cc:637-640
//   pos_wld  = (k0 * P0 + u * P1) + v * P2        gbuf.vert:43,45 with model = identity
//   pos_cam  = w2c_t + w2c_R * pos_wld                                           renderer.cc:561
//   normal_wld = normalize((k0 * N0 + u * N1) + v * N2)   gbuf.frag:25, through normal_indices; or, with
//                flat_normal, normalize(cross(normalize(P1 - P0), normalize(P2 - P0)))  dummy.geom:58-59
//   normal_cam = w2c_R * normal_wld                                              renderer.cc:583
//   uv       = (s, t, 1 - s - t), (s, t) = (k0 * T0 + u * T1) + v * T2           renderer.cc:641-644
//   face_id  = f + 1   (gbuf.frag:36; 0 = background)     geo_id = the caller's id (gbuf.frag:37)
// Background: depth_01 = 1 (the cleared GL depth), stencil 0, the rest 0.
// A winner whose window depth falls outside (0, 1) is ERASED from the visibility buffer, so the colour
// image the resolve kernel writes afterwards shows background there too (GL clips such fragments).
#pragma once

#include "raster_kernels.cuh"

namespace ugu_b200 {

struct GBufParams {
  float* normal_cam;
  float* normal_wld;
  float* pos_cam;
  float* pos_wld;
  float* depth_01;
  uint8_t* stencil;
  int* face_id;
  float* uv;
  float* bary;
  int* geo_id;
  float A, B;             // projection terms, computed on the host from (near, far)
  int flat_normal;        // RendererGl::m_flat_normal
  int geo_id_value;
  int backface_culling;
  int have_normals;       // vertex normals + normal_indices uploaded
  int have_uv;            // uv + uv_indices uploaded
};

__device__ __forceinline__ void normalize3(float& x, float& y, float& z) {
  const float n2 = x * x + (y * y + z * z);
  if (n2 > 0.0f) {
    const float s = sqrtf(n2);
    x = x / s;
    y = y / s;
    z = z / s;
  }
}

struct GBufPixel {
  float ncx, ncy, ncz, nwx, nwy, nwz, pcx, pcy, pcz, pwx, pwy, pwz, uvx, uvy, uvz, bx, by, bz;
  float d01;
  int face_id, geo_id;
  bool stencil;
};

// Returns true when the pixel's winner must be erased (window depth outside (0, 1)).
__device__ __forceinline__ bool gbuffer_pixel(const ViewParams& vp, const MeshDev& m, const FrameDev& fr,
                                              const GBufParams& gp, unsigned long long key, int x, int y,
                                              GBufPixel& o) {
  o.ncx = o.ncy = o.ncz = o.nwx = o.nwy = o.nwz = 0.f;
  o.pcx = o.pcy = o.pcz = o.pwx = o.pwy = o.pwz = 0.f;
  o.uvx = o.uvy = o.uvz = o.bx = o.by = o.bz = 0.f;
  o.d01 = 1.0f;
  o.face_id = 0;
  o.geo_id = 0;
  o.stencil = false;
  if (key == kVisEmpty) return false;
  const unsigned low = (unsigned)(key & 0xFFFFFFFFull);
  if (gp.backface_culling && (low & 1u)) return false;  // rasterizer.cc:248-252: no winner
  const int fid = (int)(low >> 1);
  const float z = __uint_as_float((unsigned)(key >> 32));
  const float ndc = (gp.A * (-z) + gp.B) / z;
  const float d = 0.5f * ndc + 0.5f;
  o.d01 = fminf(fmaxf(d, 0.0f), 1.0f);
  if (d <= 0.0f || 1.0f <= d) return true;
  o.stencil = true;
  o.face_id = fid + 1;
  o.geo_id = gp.geo_id_value;

  const int4 vr = __ldg(m.frec + 2 * (size_t)fid);
  const float4 p0 = fr.proj[vr.x], p1 = fr.proj[vr.y], p2 = fr.proj[vr.z];
  const float4 q0 = __ldg(m.pos4 + vr.x), q1 = __ldg(m.pos4 + vr.y), q2 = __ldg(m.pos4 + vr.z);
  const float P0x = q0.x, P0y = q0.y, P0z = q0.z;
  const float P1x = q1.x, P1y = q1.y, P1z = q1.z;
  const float P2x = q2.x, P2y = q2.y, P2z = q2.z;
  float4 n0 = make_float4(0.f, 0.f, 0.f, 0.f), n1 = n0, n2 = n0;
  const bool vnormal = !gp.flat_normal && gp.have_normals;
  if (vnormal) {
    const int4 ni = __ldg(m.frec + 2 * (size_t)fid + 1);
    n0 = __ldg(m.nrm4 + ni.x);
    n1 = __ldg(m.nrm4 + ni.y);
    n2 = __ldg(m.nrm4 + ni.z);
  }
  float2 t0 = make_float2(0.f, 0.f), t1 = t0, t2 = t0;
  if (gp.have_uv) {
    const int4 ui = __ldg(m.uvidx + fid);
    t0 = __ldg(m.uv2 + ui.x);
    t1 = __ldg(m.uv2 + ui.y);
    t2 = __ldg(m.uv2 + ui.z);
  }
  float w0, w1, w2;  // recomputed bit-identically to the raster kernels (no weight image)
  {
    Tri t;
    tri_setup(t, p0, p1, p2);
    tri_edges(t, (float)x, (float)y, w0, w1, w2);
    (void)tri_depth(t, w0, w1, w2);
  }
  const float u = w1, v = w2;
  const float k0 = 1.0f - u - v;
  o.bx = u;
  o.by = v;
  o.bz = k0;
  o.pwx = (k0 * P0x + u * P1x) + v * P2x;
  o.pwy = (k0 * P0y + u * P1y) + v * P2y;
  o.pwz = (k0 * P0z + u * P1z) + v * P2z;
  o.pcx = vp.w2c_t[0] + dot3(vp.w2c_R[0], vp.w2c_R[1], vp.w2c_R[2], o.pwx, o.pwy, o.pwz);
  o.pcy = vp.w2c_t[1] + dot3(vp.w2c_R[3], vp.w2c_R[4], vp.w2c_R[5], o.pwx, o.pwy, o.pwz);
  o.pcz = vp.w2c_t[2] + dot3(vp.w2c_R[6], vp.w2c_R[7], vp.w2c_R[8], o.pwx, o.pwy, o.pwz);
  float nx = 0.f, ny = 0.f, nz = 0.f;
  if (gp.flat_normal) {
    float ax = P1x - P0x, ay = P1y - P0y, az = P1z - P0z;
    float bx = P2x - P0x, by = P2y - P0y, bz = P2z - P0z;
    normalize3(ax, ay, az);
    normalize3(bx, by, bz);
    nx = ay * bz - az * by;
    ny = az * bx - ax * bz;
    nz = ax * by - ay * bx;
    normalize3(nx, ny, nz);
  } else if (vnormal) {
    nx = (k0 * n0.x + u * n1.x) + v * n2.x;
    ny = (k0 * n0.y + u * n1.y) + v * n2.y;
    nz = (k0 * n0.z + u * n1.z) + v * n2.z;
    normalize3(nx, ny, nz);
  }
  o.nwx = nx;
  o.nwy = ny;
  o.nwz = nz;
  o.ncx = dot3(vp.w2c_R[0], vp.w2c_R[1], vp.w2c_R[2], nx, ny, nz);
  o.ncy = dot3(vp.w2c_R[3], vp.w2c_R[4], vp.w2c_R[5], nx, ny, nz);
  o.ncz = dot3(vp.w2c_R[6], vp.w2c_R[7], vp.w2c_R[8], nx, ny, nz);
  if (gp.have_uv) {
    o.uvx = (k0 * t0.x + u * t1.x) + v * t2.x;
    o.uvy = (k0 * t0.y + u * t1.y) + v * t2.y;
    o.uvz = 1.0f - o.uvx - o.uvy;
  }
  return false;
}

// One warp per 32-pixel chunk, grid-stride.  Scalars go out as coalesced 4-byte stores (stencil: one
// word per 4 pixels from a ballot); the six 3-float fields are staged through shared memory so that
// every global store is a full 16-byte vector (24 lanes x float4 = the chunk's 384 bytes per field).
// HBM-bound streaming: 85 output bytes + 8 key bytes per pixel, plus the winners' gathers.
__global__ void __launch_bounds__(256) gbuffer_kernel(const ViewParams vp, const MeshDev m, const FrameDev fr,
                                                      const GBufParams gp) {
  __shared__ __align__(16) float s_v[8][6][96];
  pdl_launch_dependents();  // a link of the view's PDL chain: the resolve kernel behind it waits for its completion
  pdl_wait();
  const unsigned P = (unsigned)vp.W * (unsigned)vp.H;
  const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
  const unsigned nchunks = (P + 31u) >> 5;
  const unsigned nwarps = gridDim.x * 8u;
  float* const out3[6] = {gp.normal_cam, gp.normal_wld, gp.pos_cam, gp.pos_wld, gp.uv, gp.bary};
  for (unsigned chunk = blockIdx.x * 8u + warp; chunk < nchunks; chunk += nwarps) {
    const unsigned wbase = chunk * 32u;
    const unsigned p = wbase + lane;
    const unsigned long long key = p < P ? fr.vis[p] : kVisEmpty;
    const unsigned y = (vp.W == 1) ? p : (__umulhi(p, vp.w_magic) >> vp.w_shift);
    const unsigned x = p - y * (unsigned)vp.W;
    const bool full = wbase + 32u <= P;
    if (full && __all_sync(0xffffffffu, key == kVisEmpty)) {  // background chunk
      if (gp.depth_01) gp.depth_01[p] = 1.0f;
      if (gp.face_id) gp.face_id[p] = 0;
      if (gp.geo_id) gp.geo_id[p] = 0;
      if (gp.stencil && lane < 8u) *reinterpret_cast<unsigned*>(gp.stencil + wbase + lane * 4u) = 0u;
      if (lane < 24u) {
#pragma unroll
        for (int k = 0; k < 6; k++)
          if (out3[k]) reinterpret_cast<float4*>(out3[k] + (size_t)wbase * 3u)[lane] = make_float4(0.f, 0.f, 0.f, 0.f);
      }
      continue;
    }
    GBufPixel o;
    if (gbuffer_pixel(vp, m, fr, gp, key, (int)x, (int)y, o)) fr.vis[p] = kVisEmpty;
    const float f3[6][3] = {{o.ncx, o.ncy, o.ncz}, {o.nwx, o.nwy, o.nwz}, {o.pcx, o.pcy, o.pcz},
                            {o.pwx, o.pwy, o.pwz}, {o.uvx, o.uvy, o.uvz}, {o.bx, o.by, o.bz}};
    if (full) {
      if (gp.depth_01) gp.depth_01[p] = o.d01;
      if (gp.face_id) gp.face_id[p] = o.face_id;
      if (gp.geo_id) gp.geo_id[p] = o.geo_id;
      if (gp.stencil) {
        const unsigned bal = __ballot_sync(0xffffffffu, o.stencil);
        if (lane < 8u) {
          const unsigned nb = (bal >> (lane * 4u)) & 15u;  // nibble -> 4 bytes of 0x00 / 0xFF
          *reinterpret_cast<unsigned*>(gp.stencil + wbase + lane * 4u) = ((nb * 0x00204081u) & 0x01010101u) * 0xFFu;
        }
      }
#pragma unroll
      for (int k = 0; k < 6; k++) {
        if (out3[k]) {
          s_v[warp][k][lane * 3u + 0u] = f3[k][0];
          s_v[warp][k][lane * 3u + 1u] = f3[k][1];
          s_v[warp][k][lane * 3u + 2u] = f3[k][2];
        }
      }
      __syncwarp();
      if (lane < 24u) {
#pragma unroll
        for (int k = 0; k < 6; k++)
          if (out3[k])
            reinterpret_cast<float4*>(out3[k] + (size_t)wbase * 3u)[lane] =
                reinterpret_cast<const float4*>(s_v[warp][k])[lane];
      }
      __syncwarp();
    } else if (p < P) {  // ragged last chunk of the image
      if (gp.depth_01) gp.depth_01[p] = o.d01;
      if (gp.face_id) gp.face_id[p] = o.face_id;
      if (gp.geo_id) gp.geo_id[p] = o.geo_id;
      if (gp.stencil) gp.stencil[p] = o.stencil ? 255 : 0;
#pragma unroll
      for (int k = 0; k < 6; k++)
        if (out3[k]) {
          out3[k][(size_t)p * 3 + 0] = f3[k][0];
          out3[k][(size_t)p * 3 + 1] = f3[k][1];
          out3[k][(size_t)p * 3 + 2] = f3[k][2];
        }
    }
  }
}

}  // namespace ugu_b200
