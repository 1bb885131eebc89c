/*
 * TEST INFRASTRUCTURE — NOT PART OF THE PRODUCT.
 *
 * Plain-C, dependency-free CPU restatement ("port") of the reference software rasterizer
 * ugu::Rasterizer::Impl::Render (src/renderer/cpu/rasterizer.cc:105-307) together with the pieces of
 * ugu::PinholeCamera (include/ugu/camera.h:380-387,438-448,482-513) and the pixel shaders
 * (include/ugu/renderer/cpu/pixel_shader.h:143-148,222-422) that it calls.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load
 * this library, and only as the checker.  Nothing under ugu_b200/ links, imports or executes it.
 *
 * Parity status: PINNED.  The reference ships no tests or golden vectors, so this
 * restatement is pinned differentially: tests/test_oracle_vs_reference.py compares it bit-for-bit
 * with the compiled, unmodified reference (oracle/_ref/libugu_ref.so, built by oracle/Makefile from
 * /root/reference) whenever that library is present, and tests/golden/ holds outputs of the compiled
 * reference that this file must reproduce bit-for-bit everywhere (CPU suite).
 *
 * Float discipline: every expression below is evaluated in IEEE binary32 with round-to-nearest
 * single operations in exactly the association order of the compiled reference (Eigen 3.4.0
 * fixed-size reductions associate a0*b0 + (a1*b1 + a2*b2), third_party/eigen/Eigen/src/Core/
 * Redux.h:99-113).  Compile with -ffp-contract=off and without -ffast-math (oracle/Makefile).
 */

#include <float.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct OracleMesh {
  int32_t num_vertices;
  int32_t num_normals;
  int32_t num_colors;
  int32_t num_faces;
  int32_t num_face_normals;
  int32_t num_uv;
  int32_t num_materials;
  int32_t pad_;
  const float* vertices;          /* num_vertices * 3 */
  const float* normals;           /* num_normals * 3 */
  const float* vertex_colors;     /* num_colors * 3, 0..255 as float */
  const int32_t* vertex_indices;  /* num_faces * 3 */
  const int32_t* normal_indices;  /* num_faces * 3 or NULL */
  const float* face_normals;      /* num_face_normals * 3 */
  const float* uv;                /* num_uv * 2 */
  const int32_t* uv_indices;      /* num_faces * 3 or NULL */
  const int32_t* material_ids;    /* num_faces or NULL */
  const uint8_t* const* tex_data; /* per material RGB8 row-major */
  const int32_t* tex_w;
  const int32_t* tex_h;
} OracleMesh;

/* Float camera constants, derived on the host exactly as rasterizer.cc:118-119 and
 * camera.h:317-331 do (row-major 3x3). */
typedef struct OracleCamera {
  int32_t width, height;
  float fx, fy, cx, cy;
  float w2c_R[9];
  float w2c_t[3];
  float c2w_R[9];
} OracleCamera;

typedef struct OracleOption {
  int32_t diffuse_color;    /* 0 none, 1 texture, 2 vertex     (renderer/base.h:11-15) */
  int32_t interp;           /* 0 nearest, 1 bilinear           (common.h:50-53) */
  int32_t shading_normal;   /* 0 face, 1 vertex                (renderer/base.h:19-22) */
  int32_t diffuse_shading;  /* 0 none, 1 Lambertian, 2 Oren-Nayar (renderer/base.h:26-32) */
  int32_t backface_culling;
  float depth_scale;        /* ignored by the Rasterizer */
  float oren_nayar_sigma;   /* the reference never copies it, so callers pass 0.3f (renderer.h:30-37) */
} OracleOption;

/* Eigen fixed-size 3-vector reduction order. */
static inline float dot3(const float* a, const float* b) {
  return a[0] * b[0] + (a[1] * b[1] + a[2] * b[2]);
}

/* Row-major 3x3 times vector, each row reduced like dot3. */
static inline void mat3_mul(const float* M, const float* v, float* out) {
  float r0 = M[0] * v[0] + (M[1] * v[1] + M[2] * v[2]);
  float r1 = M[3] * v[0] + (M[4] * v[1] + M[5] * v[2]);
  float r2 = M[6] * v[0] + (M[7] * v[1] + M[8] * v[2]);
  out[0] = r0;
  out[1] = r1;
  out[2] = r2;
}

/* rasterizer.cc:25-28 */
static inline float edge_function(const float* a, const float* b, const float* c) {
  return (c[0] - a[0]) * (b[1] - a[1]) - (c[1] - a[1]) * (b[0] - a[0]);
}

/* camera.h:438-448 (ray_c then rotation by c2w_R_f_), evaluated per pixel instead of through the
 * lazily built table of camera.h:482-513 — the table entries are these very values. */
static inline void ray_w(const OracleCamera* cam, int x, int y, float* out) {
  float d[3];
  d[0] = ((float)x - cam->cx) / cam->fx;
  d[1] = ((float)y - cam->cy) / cam->fy;
  d[2] = 1.0f;
  float n2 = d[0] * d[0] + (d[1] * d[1] + d[2] * d[2]);
  if (n2 > 0.0f) {
    float s = sqrtf(n2);
    d[0] = d[0] / s;
    d[1] = d[1] / s;
    d[2] = d[2] / s;
  }
  mat3_mul(cam->c2w_R, d, out);
}

/* static_cast<unsigned char>(float) as x86-64 gcc compiles it: cvttss2si to int32 (0x80000000 when
 * out of range or NaN), then the low byte. */
static inline uint8_t f2u8(float f) {
  if (!(f > -2147483904.0f && f < 2147483648.0f)) return 0;
  return (uint8_t)(int32_t)f;
}

/* Image3b::at<Vec3b>(y, x) without bounds checks (image.h:415-427): a linear address.  Reads that
 * fall outside the buffer are undefined behaviour in the reference; here they return 0, which is
 * what the reference effectively uses whenever the tap's weight is 0 (uv on the last row/col). */
static inline void texel(const uint8_t* tex, int w, int h, int x, int y, float* rgb) {
  int64_t off = ((int64_t)y * w + x) * 3;
  int64_t last = (int64_t)w * h * 3 - 3;
  if (off < 0 || off > last) {
    rgb[0] = rgb[1] = rgb[2] = 0.0f;
    return;
  }
  rgb[0] = (float)tex[off];
  rgb[1] = (float)tex[off + 1];
  rgb[2] = (float)tex[off + 2];
}

static inline int32_t sat_i32(float f) {
  /* (int32_t)float with the out-of-range cases pinned (UB in the reference; unreachable for
   * on-contract geometry). */
  if (!(f > -2147483904.0f)) return INT32_MIN;
  if (!(f < 2147483648.0f)) return INT32_MAX;
  return (int32_t)f;
}

/* The colourizer + shader of one pixel: PixelShader::Process (include/ugu/renderer/cpu/pixel_shader.h:
 * 115-140,222-422) as both CPU renderers call it (rasterizer.cc:291-298, raytracer.cc:205-211), with
 * (u, v) the barycentrics of vertices 1 and 2, `ray` the world ray (== light direction), `n_w` the world
 * shading normal. */
static void shade_pixel(const OracleMesh* mesh, const OracleOption* opt, float on_A, float on_B, int32_t fid,
                        float u, float v, const float* ray, const float* n_w, uint8_t* c) {
  /* colourizer: pixel_shader.h:222-359 */
  if (opt->diffuse_color == 2) {
    const int32_t* f = &mesh->vertex_indices[(size_t)fid * 3];
    const float* c0 = &mesh->vertex_colors[(size_t)f[0] * 3];
    const float* c1 = &mesh->vertex_colors[(size_t)f[1] * 3];
    const float* c2 = &mesh->vertex_colors[(size_t)f[2] * 3];
    const float k0 = 1.0f - u - v;
    for (int k = 0; k < 3; k++) c[k] = f2u8((k0 * c0[k] + u * c1[k]) + v * c2[k]);
  } else if (opt->diffuse_color == 1) {
    const int32_t* ui = &mesh->uv_indices[(size_t)fid * 3];
    const float* t0 = &mesh->uv[(size_t)ui[0] * 2];
    const float* t1 = &mesh->uv[(size_t)ui[1] * 2];
    const float* t2 = &mesh->uv[(size_t)ui[2] * 2];
    const int mat = mesh->material_ids[fid];
    const uint8_t* tex = mesh->tex_data[mat];
    const int tw = mesh->tex_w[mat], th = mesh->tex_h[mat];
    const float k0 = 1.0f - u - v;
    const float iu = (k0 * t0[0] + u * t1[0]) + v * t2[0];
    const float iv = (k0 * t0[1] + u * t1[1]) + v * t2[1];
    const float tx = iu * (float)(tw - 1);
    const float ty = (1.0f - iv) * (float)(th - 1);
    if (opt->interp == 0) {
      float rgb[3];
      texel(tex, tw, th, sat_i32(roundf(tx)), sat_i32(roundf(ty)), rgb);
      for (int k = 0; k < 3; k++) c[k] = f2u8(rgb[k]);
    } else {
      const int xm = sat_i32(floorf(tx)), ym = sat_i32(floorf(ty));
      const float lu = tx - (float)xm, lv = ty - (float)ym;
      float a[3], b[3], cc[3], dd[3];
      texel(tex, tw, th, xm, ym, a);
      texel(tex, tw, th, xm + 1, ym, b);
      texel(tex, tw, th, xm, ym + 1, cc);
      texel(tex, tw, th, xm + 1, ym + 1, dd);
      for (int k = 0; k < 3; k++) {
        float val = (1.0f - lu) * (1.0f - lv) * a[k] + lu * (1.0f - lv) * b[k] +
                    (1.0f - lu) * lv * cc[k] + lu * lv * dd[k];
        c[k] = f2u8(val);
      }
    }
  } else {
    c[0] = c[1] = c[2] = 255;
  }
  /* shader: pixel_shader.h:363-422; light_dir == ray_w (rasterizer.cc:293) */
  if (opt->diffuse_shading == 1) {
    float coeff = -dot3(ray, n_w);
    if (coeff < 0.0f) coeff = 0.0f;
    for (int k = 0; k < 3; k++) c[k] = f2u8(coeff * (float)c[k]);
  } else if (opt->diffuse_shading == 2) {
    float dot_light = -dot3(ray, n_w);
    float theta_i = acosf(dot_light);
    float dot_ray = -dot3(ray, n_w);
    float theta_r = acosf(dot_ray);
    float bl[3], br[3];
    for (int k = 0; k < 3; k++) {
      bl[k] = (-n_w[k]) * dot_light - ray[k];
      br[k] = (-n_w[k]) * dot_light - ray[k];
    }
    float phi_diff_cos = fmaxf(0.0f, dot3(bl, br));
    float alpha = (theta_i < theta_r) ? theta_r : theta_i; /* std::max(theta_i, theta_r) */
    float beta = (theta_r < theta_i) ? theta_r : theta_i;  /* std::min(theta_i, theta_r) */
    float coeff = (dot_light > 0.0f ? dot_light : 0.0f) *
                  (on_A + (on_B * phi_diff_cos * sinf(alpha) * tanf(beta)));
    for (int k = 0; k < 3; k++) c[k] = f2u8(coeff * (float)c[k]);
  }
}

/* Raw z-buffer state after rasterizer.cc:162-241 and BEFORE the cull / resolve loop (:244-301), for
 * oracle_gbuffer: winner face (-1 = none), its depth, weights and back-face flag per pixel. */
typedef struct RawWinners {
  int32_t* fid;      /* P */
  float* depth;      /* P */
  float* weight;     /* P * 3 */
  uint8_t* backface; /* P */
  const uint8_t* erase; /* P or NULL: winners to erase before the resolve loop (set by a 2nd pass) */
} RawWinners;

static int render_impl(const OracleMesh* mesh, const OracleCamera* cam, const OracleOption* opt,
                       uint8_t* color, float* depth, float* normal, uint8_t* mask, int32_t* face_id,
                       RawWinners* raw);

int oracle_render(const OracleMesh* mesh, const OracleCamera* cam, const OracleOption* opt,
                  uint8_t* color, float* depth, float* normal, uint8_t* mask, int32_t* face_id) {
  return render_impl(mesh, cam, opt, color, depth, normal, mask, face_id, NULL);
}

static int render_impl(const OracleMesh* mesh, const OracleCamera* cam, const OracleOption* opt,
                       uint8_t* color, float* depth, float* normal, uint8_t* mask, int32_t* face_id,
                       RawWinners* raw) {
  const int W = cam->width, H = cam->height;
  const size_t P = (size_t)W * (size_t)H;
  const int V = mesh->num_vertices, F = mesh->num_faces;
  if (W <= 0 || H <= 0) return 2;

  /* util_private.cc:19-67 — the validations that depend on mesh/option content. */
  if (opt->backface_culling && mesh->num_face_normals == 0) return 1;
  if (opt->diffuse_color == 1 && mesh->num_materials == 0) return 1;
  if (opt->diffuse_color == 1) {
    for (int i = 0; i < mesh->num_materials; i++)
      if (mesh->tex_data[i] == NULL || mesh->tex_w[i] < 1 || mesh->tex_h[i] < 1) return 1;
  }
  if (opt->diffuse_color == 2 && mesh->num_colors == 0) return 1;
  if (opt->shading_normal == 0 && mesh->num_face_normals == 0) return 1;
  if (opt->shading_normal == 1 && mesh->num_normals == 0) return 1;
  if (!color && !depth && !normal && !mask && !face_id) return 1;

  /* util_private.cc:72-87 */
  if (color) memset(color, 0, P * 3);
  if (normal) memset(normal, 0, P * 12);
  if (mask) memset(mask, 0, P);

  /* pixel_shader.h:143-148 */
  const float sigma = opt->oren_nayar_sigma;
  const float sigma_2 = sigma * sigma;
  const float on_A = 1.0f - (0.5f * sigma_2 / (sigma_2 + 0.33f));
  const float on_B = 0.45f * sigma_2 / (sigma_2 + 0.09f);

  /* rasterizer.cc:125-136 + camera.h:380-387 */
  float* img = (float*)malloc((size_t)(V > 0 ? V : 1) * 3 * sizeof(float));
  float* depth_ = depth ? depth : (float*)malloc(P * sizeof(float));
  int32_t* fid_ = face_id ? face_id : (int32_t*)malloc(P * sizeof(int32_t));
  uint8_t* backface = (uint8_t*)calloc(P, 1);
  float* weight = (float*)calloc(P * 3, sizeof(float));
  if (!img || !depth_ || !fid_ || !backface || !weight) return 3;
  for (int i = 0; i < V; i++) {
    float p[3];
    mat3_mul(cam->w2c_R, &mesh->vertices[(size_t)i * 3], p);
    p[0] = p[0] + cam->w2c_t[0];
    p[1] = p[1] + cam->w2c_t[1];
    p[2] = p[2] + cam->w2c_t[2];
    img[(size_t)i * 3 + 0] = cam->fx / p[2] * p[0] + cam->cx;
    img[(size_t)i * 3 + 1] = cam->fy / p[2] * p[1] + cam->cy;
    img[(size_t)i * 3 + 2] = p[2];
  }

  /* rasterizer.cc:138-159 */
  for (size_t i = 0; i < P; i++) {
    depth_[i] = 0.0f;
    fid_[i] = -1;
  }

  /* rasterizer.cc:162-241 — z-buffer in ascending face order */
  for (int i = 0; i < F; i++) {
    const int32_t* face = &mesh->vertex_indices[(size_t)i * 3];
    const float* v0 = &img[(size_t)face[0] * 3];
    const float* v1 = &img[(size_t)face[1] * 3];
    const float* v2 = &img[(size_t)face[2] * 3];
    if (v0[2] < 0.0f || v1[2] < 0.0f || v2[2] < 0.0f) continue;

    float xmin = fminf(fminf(v0[0], v1[0]), v2[0]);
    float ymin = fminf(fminf(v0[1], v1[1]), v2[1]);
    float xmax = fmaxf(fmaxf(v0[0], v1[0]), v2[0]);
    float ymax = fmaxf(fmaxf(v0[1], v1[1]), v2[1]);
    if (xmin > (float)(W - 1) || xmax < 0.0f || ymin > (float)(H - 1) || ymax < 0.0f) continue;

    int32_t x0 = sat_i32(ceilf(xmin));
    if (x0 < 0) x0 = 0;
    int32_t x1 = sat_i32(floorf(xmax));
    if (x1 > W - 1) x1 = W - 1;
    int32_t y0 = sat_i32(ceilf(ymin));
    if (y0 < 0) y0 = 0;
    int32_t y1 = sat_i32(floorf(ymax));
    if (y1 > H - 1) y1 = H - 1;

    float area = edge_function(v0, v1, v2);
    if (fabsf(area) < FLT_MIN) continue;

    const float* fn = mesh->num_face_normals > i ? &mesh->face_normals[(size_t)i * 3] : NULL;
    for (int32_t y = y0; y <= y1; ++y) {
      for (int32_t x = x0; x <= x1; ++x) {
        float ray[3];
        ray_w(cam, x, y, ray);
        int bf = fn ? (dot3(fn, ray) > 0.0f) : 0;
        float ps[3] = {(float)x, (float)y, 0.0f};
        float w0 = edge_function(v1, v2, ps);
        float w1 = edge_function(v2, v0, ps);
        float w2 = edge_function(v0, v1, ps);
        if ((!bf && (w0 >= 0 && w1 >= 0 && w2 >= 0)) || (bf && (w0 <= 0 && w1 <= 0 && w2 <= 0))) {
          w0 /= area;
          w1 /= area;
          w2 /= area;
          w0 /= v0[2];
          w1 /= v1[2];
          w2 /= v2[2];
          float z = 1.0f / (w0 + w1 + w2);
          w0 = w0 * z;
          w1 = w1 * z;
          w2 = w2 * z;
          size_t pi = (size_t)y * W + x;
          float d = depth_[pi];
          if (d < FLT_MIN || z < d) {
            depth_[pi] = z;
            fid_[pi] = i;
            weight[pi * 3 + 0] = w0;
            weight[pi * 3 + 1] = w1;
            weight[pi * 3 + 2] = w2;
            backface[pi] = bf ? 255 : 0;
          }
        }
      }
    }
  }

  if (raw) {
    if (raw->fid) memcpy(raw->fid, fid_, P * sizeof(int32_t));
    if (raw->depth) memcpy(raw->depth, depth_, P * sizeof(float));
    if (raw->weight) memcpy(raw->weight, weight, P * 3 * sizeof(float));
    if (raw->backface) memcpy(raw->backface, backface, P);
    if (raw->erase)
      for (size_t i = 0; i < P; i++)
        if (raw->erase[i]) {
          depth_[i] = 0.0f;
          fid_[i] = -1;
          backface[i] = 0;
        }
  }

  /* rasterizer.cc:244-301 */
  for (int y = 0; y < H; y++) {
    for (int x = 0; x < W; x++) {
      size_t pi = (size_t)y * W + x;
      if (opt->backface_culling && backface[pi] == 255) {
        depth_[pi] = 0.0f;
        fid_[pi] = -1;
        continue;
      }
      int32_t fid = fid_[pi];
      if (fid > 0) { /* sic: face 0 is never shaded */
        float ray[3];
        ray_w(cam, x, y, ray);
        float w0 = weight[pi * 3 + 0], w1 = weight[pi * 3 + 1], w2 = weight[pi * 3 + 2];
        if (mask) mask[pi] = 255;

        float n_w[3] = {0.0f, 0.0f, 0.0f};
        if (opt->shading_normal == 0) {
          memcpy(n_w, &mesh->face_normals[(size_t)fid * 3], 12);
        } else if (opt->shading_normal == 1) {
          const int32_t* ni = &mesh->normal_indices[(size_t)fid * 3];
          const float* n0 = &mesh->normals[(size_t)ni[0] * 3];
          const float* n1 = &mesh->normals[(size_t)ni[1] * 3];
          const float* n2 = &mesh->normals[(size_t)ni[2] * 3];
          for (int k = 0; k < 3; k++) n_w[k] = (w0 * n0[k] + w1 * n1[k]) + w2 * n2[k];
        }
        if (normal) mat3_mul(cam->w2c_R, n_w, &normal[pi * 3]);

        if (color) shade_pixel(mesh, opt, on_A, on_B, fid, w1, w2, ray, n_w, &color[pi * 3]);
      }
    }
  }

  free(img);
  free(backface);
  free(weight);
  if (!depth) free(depth_);
  if (!face_id) free(fid_);
  return 0;
}

/* ugu::Depth2PointCloud(const Image1f&, const Camera&, Image3f*, bool) — src/util/rgbd_util.cc:472-513
 * with PinholeCamera::Unproject, include/ugu/camera.h:406-413:
 *   camera_p = ((x - cx) * d / fx, (y - cy) * d / fy, d), y and z negated when gl_coord.
 * Pixels with d < FLT_MIN are left untouched (pc is in/out), as the reference does. */
int oracle_depth2pointcloud(const OracleCamera* cam, const float* depth, float* pc, int gl_coord) {
  if (!cam || !depth || !pc) return 1;
  for (int y = 0; y < cam->height; y++) {
    for (int x = 0; x < cam->width; x++) {
      const size_t p = (size_t)y * (size_t)cam->width + (size_t)x;
      const float d = depth[p];
      if (d < FLT_MIN) continue;
      float cx_ = ((float)x - cam->cx) * d / cam->fx;
      float cy_ = ((float)y - cam->cy) * d / cam->fy;
      float cz_ = d;
      if (gl_coord) {
        cy_ = -cy_;
        cz_ = -cz_;
      }
      pc[p * 3 + 0] = cx_;
      pc[p * 3 + 1] = cy_;
      pc[p * 3 + 2] = cz_;
    }
  }
  return 0;
}

/* static_cast<int>(float) as x86-64 compiles it (cvttss2si): truncation, 0x80000000 when out of range / NaN */
static int32_t cvtt_x86(float f) {
  if (!(f > -2147483904.0f && f < 2147483648.0f)) return INT32_MIN;
  return (int32_t)f;
}

static float edge2(const float* a, const float* b, const float* c) {
  /* EdgeFunction, include/ugu/util/raster_util.h:31-34 */
  return (c[0] - a[0]) * (b[1] - a[1]) - (c[1] - a[1]) * (b[0] - a[0]);
}

static float min3f(float a, float b, float c) { /* std::min({a,b,c}): first smallest by operator< */
  float m = a;
  if (b < m) m = b;
  if (c < m) m = c;
  return m;
}
static float max3f(float a, float b, float c) { /* std::max({a,b,c}) */
  float m = a;
  if (m < b) m = b;
  if (m < c) m = c;
  return m;
}
static int clampi(int v, int hi) { /* std::max(0, std::min(v, hi)) */
  int t = v < hi ? v : hi;
  return t > 0 ? t : 0;
}

/* ugu::RasterizeVertexAttributeToTexture / RasterizeFaceAttributeToTexture with num_threads = 1
 * (include/ugu/util/raster_util.h:139-268) and RasterizeTriangle (:78-137); U2X / V2Y (:18-21).
 * Same argument meaning as ugu_ref_uv_raster in oracle/ref_shim.cc.  Faces in ascending index, the
 * last writer of a pixel wins. */
int oracle_uv_raster(int32_t num_attrs, int32_t num_faces, int32_t num_uv, int32_t per_face,
                     const float* attrs, const int32_t* attr_indices, const float* uv,
                     const int32_t* uv_indices, int32_t width, int32_t height, int32_t texture_type,
                     int32_t keep_texture, void* texture, uint8_t* mask) {
  (void)num_uv;
  if (per_face) {
    if (num_attrs == 0 || num_attrs != num_faces) return 1; /* raster_util.h:238-240 */
  } else {
    if (num_faces == 0) return 1; /* :148-151 (equal sizes are implied by the flat arguments) */
  }
  if (width <= 0 || height <= 0) return 1;
  const size_t P = (size_t)width * (size_t)height;
  uint8_t* tb = (uint8_t*)texture;
  float* tf = (float*)texture;
  if (!keep_texture) memset(texture, 0, texture_type == 0 ? P * 3 : P * 12);
  for (int32_t f = 0; f < num_faces; f++) {
    const int32_t* ui = uv_indices + (size_t)f * 3;
    float tri[3][2];
    const float* col[3];
    for (int k = 0; k < 3; k++) {
      const float* q = uv + (size_t)ui[k] * 2;
      tri[k][0] = q[0] * (float)width - 0.5f;             /* U2X */
      tri[k][1] = (1.f - q[1]) * (float)height - 0.5f;    /* V2Y, flip */
      col[k] = per_face ? attrs + (size_t)f * 3 : attrs + (size_t)attr_indices[(size_t)f * 3 + k] * 3;
    }
    float area = edge2(tri[0], tri[1], tri[2]);
    if (fabsf(area) < FLT_MIN) area = area > 0 ? FLT_MIN : -FLT_MIN;
    const float inv_area = 1.0f / area;
    const int xmin = clampi(cvtt_x86(min3f(tri[0][0], tri[1][0], tri[2][0]) - 1), width - 1);
    const int xmax = clampi(cvtt_x86(max3f(tri[0][0], tri[1][0], tri[2][0]) + 1), width - 1);
    const int ymin = clampi(cvtt_x86(min3f(tri[0][1], tri[1][1], tri[2][1]) - 1), height - 1);
    const int ymax = clampi(cvtt_x86(max3f(tri[0][1], tri[1][1], tri[2][1]) + 1), height - 1);
    for (int y = ymin; y <= ymax; y++) {
      for (int x = xmin; x <= xmax; x++) {
        const float p[2] = {(float)x, (float)y};
        float w0 = edge2(tri[1], tri[2], p);
        float w1 = edge2(tri[2], tri[0], p);
        float w2 = edge2(tri[0], tri[1], p);
        w0 *= inv_area;
        w1 *= inv_area;
        w2 *= inv_area;
        if (w0 < 0 || w1 < 0 || w2 < 0) continue;
        const size_t pix = (size_t)y * (size_t)width + (size_t)x;
        for (int k = 0; k < 3; k++) {
          const float c = (w0 * col[0][k] + w1 * col[1][k]) + w2 * col[2][k];
          if (texture_type == 0)
            tb[pix * 3 + k] = (uint8_t)(cvtt_x86(roundf(c)) & 0xFF); /* saturate_cast: std::round, then cast */
          else
            tf[pix * 3 + k] = c;
        }
        if (mask) mask[pix] = 255;
      }
    }
  }
  return 0;
}

/* ugu::GenerateUvMask(uvs, uv_indices, Image1b&, uint8_t color, ...) — src/util/raster_util.cc:9-27:
 * one constant attribute rasterised into an Image3b, channel 0 copied out. */
int oracle_uv_mask(int32_t num_faces, int32_t num_uv, const float* uv, const int32_t* uv_indices,
                   int32_t width, int32_t height, int32_t color, uint8_t* mask) {
  if (width <= 0 || height <= 0) return 1;
  const size_t P = (size_t)width * (size_t)height;
  const float c = (float)(uint8_t)color;
  const float attr[3] = {c, c, c};
  int32_t* zero_idx = (int32_t*)calloc((size_t)(num_faces > 0 ? num_faces : 1) * 3, sizeof(int32_t));
  uint8_t* tex = (uint8_t*)malloc(P * 3);
  if (!zero_idx || !tex) {
    free(zero_idx);
    free(tex);
    return 1;
  }
  const int rc = oracle_uv_raster(1, num_faces, num_uv, 0, attr, zero_idx, uv, uv_indices, width, height,
                                  0, 0, tex, NULL);
  if (rc == 0)
    for (size_t i = 0; i < P; i++) mask[i] = tex[i * 3];
  free(zero_idx);
  free(tex);
  return rc;
}

/* Eigen Vector3f::normalized() / normalize() (third_party/eigen/Eigen/src/Core/Dot.h:121-145):
 * z = squaredNorm() = x*x + (y*y + z*z); if (z > 0) v / sqrt(z) (true divisions), else unchanged. */
static void normalize3(float* v) {
  const float z = v[0] * v[0] + (v[1] * v[1] + v[2] * v[2]);
  if (z > 0.0f) {
    const float s = sqrtf(z);
    v[0] = v[0] / s;
    v[1] = v[1] / s;
    v[2] = v[2] / s;
  }
}

/* Mesh::CalcFaceNormal (src/mesh.cc:339-349) for one face */
static void face_normal3(const float* a, const float* b, const float* c, float* n) {
  float v1[3] = {b[0] - a[0], b[1] - a[1], b[2] - a[2]};
  float v2[3] = {c[0] - a[0], c[1] - a[1], c[2] - a[2]};
  normalize3(v1);
  normalize3(v2);
  n[0] = v1[1] * v2[2] - v1[2] * v2[1];
  n[1] = v1[2] * v2[0] - v1[0] * v2[2];
  n[2] = v1[0] * v2[1] - v1[1] * v2[0];
  normalize3(n);
}

/* ugu::Depth2Mesh — Depth2MeshImpl, src/util/rgbd_util.cc:94-282, with PinholeCamera::Unproject
 * (include/ugu/camera.h:406-413), X2U / Y2V / U2X / V2Y (include/ugu/util/raster_util.h:18-28) and
 * Mesh::CalcNormal (src/mesh.cc:297-337).  color == NULL: the depth-only overload (no uv, no colours);
 * else the textured overload (uv always, vertex colours when with_vertex_color).
 * Caller-sized outputs: vertices / normals / colors hold up to G*3 floats, uv G*2, faces /
 * face_normals 2*G*3 with G = ((W-1)/x_step) * ((H-1)/y_step) grid points; point_cloud and
 * normal_image are W*H*3 (zero-filled here like ugu::Init).  counts[0..1] = Nv, Nf.  Any output
 * except counts may be NULL.  Returns 1 where the reference returns false. */
/* two_cam != NULL: the two-camera overload (rgbd_util.h:39-48; Depth2MeshImpl, rgbd_util.cc:283-466) — uv and
 * nearest colour from the projection of depth2color * p into the colour camera (PinholeCamera::Project,
 * camera.h:380-387).  two_cam = {R row-major 9, t 3, fx, fy, cx, cy} of depth2color / the colour camera; the
 * Eigen::Affine3f product evaluates ((R_i0 x + R_i1 y) + R_i2 z) + t_i (pinned against the compiled reference).
 * No aspect-ratio check in that overload; a colour pixel outside the image is undefined there and here. */
static int depth2mesh_impl(const OracleCamera* cam, const float* two_cam, const float* depth, const uint8_t* color,
                           int32_t cw, int32_t ch, int32_t with_vertex_color, float max_connect_z_diff,
                           int32_t x_step, int32_t y_step, int32_t gl_coord, float* vertices, int32_t* faces,
                           float* normals, float* face_normals, float* uv, float* colors, float* point_cloud,
                           float* normal_image, int32_t* counts) {
  if (!cam || !depth || !counts) return 1;
  if (max_connect_z_diff < 0 || x_step < 1 || y_step < 1) return 1; /* rgbd_util.cc:100-112 */
  const int W = cam->width, H = cam->height;
  if (color && !two_cam) { /* rgbd_util.cc:121-136 */
    const float da = (float)W / (float)H, ca = (float)cw / (float)ch;
    if (fabsf(da - ca) > 0.01f) return 1;
  }
  const size_t P = (size_t)W * (size_t)H;
  int32_t* added = (int32_t*)malloc(P * sizeof(int32_t));
  float* vbuf = (float*)malloc((P ? P : 1) * 3 * sizeof(float));
  int32_t* fbuf = (int32_t*)malloc((P ? P : 1) * 6 * sizeof(int32_t));
  int32_t* vxy = (int32_t*)malloc((P ? P : 1) * 2 * sizeof(int32_t));
  if (!added || !vbuf || !fbuf || !vxy) {
    free(added); free(vbuf); free(fbuf); free(vxy);
    return 1;
  }
  for (size_t i = 0; i < P; i++) added[i] = -1;
  int32_t nv = 0, nf = 0;
  for (int y = y_step; y < H; y += y_step) {
    for (int x = x_step; x < W; x += x_step) {
      const float d = depth[(size_t)y * W + x];
      if (d < FLT_MIN) continue;
      float px = ((float)x - cam->cx) * d / cam->fx;
      float py = ((float)y - cam->cy) * d / cam->fy;
      float pz = d;
      if (gl_coord) {
        py = -py;
        pz = -pz;
      }
      vbuf[nv * 3 + 0] = px;
      vbuf[nv * 3 + 1] = py;
      vbuf[nv * 3 + 2] = pz;
      vxy[nv * 2 + 0] = x;
      vxy[nv * 2 + 1] = y;
      float u = ((float)x + 0.5f) / (float)W;            /* X2U */
      float v = ((float)y + 0.5f) / (float)H;            /* Y2V, flip = false */
      if (two_cam) { /* rgbd_util.cc:345-350 */
        const float* R = two_cam;
        const float* t = two_cam + 9;
        const float qx = ((R[0] * px + R[1] * py) + R[2] * pz) + t[0];
        const float qy = ((R[3] * px + R[4] * py) + R[5] * pz) + t[1];
        const float qz = ((R[6] * px + R[7] * py) + R[8] * pz) + t[2];
        const float ix = two_cam[12] / qz * qx + two_cam[14];
        const float iy = two_cam[13] / qz * qy + two_cam[15];
        u = (ix + 0.5f) / (float)cw;
        v = (iy + 0.5f) / (float)ch;
      }
      if (color && with_vertex_color && colors) {
        const int cx_ = (int)roundf(u * (float)cw - 0.5f); /* U2X */
        const int cy_ = (int)roundf(v * (float)ch - 0.5f); /* V2Y, flip = false */
        const uint8_t* t = color + ((size_t)cy_ * cw + cx_) * 3;
        colors[nv * 3 + 0] = (float)t[0];
        colors[nv * 3 + 1] = (float)t[1];
        colors[nv * 3 + 2] = (float)t[2];
      }
      if (color && uv) {
        uv[nv * 2 + 0] = u;
        uv[nv * 2 + 1] = 1.0f - v;
      }
      added[(size_t)y * W + x] = nv;
      const int32_t ul = added[(size_t)(y - y_step) * W + (x - x_step)];
      const int32_t up = added[(size_t)(y - y_step) * W + x];
      const int32_t lf = added[(size_t)y * W + (x - x_step)];
      const float ul_diff = fabsf(depth[(size_t)(y - y_step) * W + (x - x_step)] - d);
      const float up_diff = fabsf(depth[(size_t)(y - y_step) * W + x] - d);
      const float lf_diff = fabsf(depth[(size_t)y * W + (x - x_step)] - d);
      /* sic: `> 0`, so vertex 0 never joins a face (rgbd_util.cc:222-234) */
      if (ul > 0 && up > 0 && ul_diff < max_connect_z_diff && up_diff < max_connect_z_diff) {
        fbuf[nf * 3 + 0] = ul; fbuf[nf * 3 + 1] = nv; fbuf[nf * 3 + 2] = up;
        nf++;
      }
      if (ul > 0 && lf > 0 && ul_diff < max_connect_z_diff && lf_diff < max_connect_z_diff) {
        fbuf[nf * 3 + 0] = ul; fbuf[nf * 3 + 1] = lf; fbuf[nf * 3 + 2] = nv;
        nf++;
      }
      nv++;
    }
  }
  /* Mesh::CalcNormal: returns early (no normals at all) when there are no faces (mesh.cc:298-300) */
  float* nsum = (float*)calloc((size_t)(nv ? nv : 1) * 3, sizeof(float));
  int32_t* ncnt = (int32_t*)calloc((size_t)(nv ? nv : 1), sizeof(int32_t));
  const int have_normals = nf > 0;
  for (int32_t i = 0; i < nf; i++) {
    float n[3];
    const int32_t* f = fbuf + i * 3;
    face_normal3(vbuf + f[0] * 3, vbuf + f[1] * 3, vbuf + f[2] * 3, n);
    if (face_normals) memcpy(face_normals + i * 3, n, sizeof(n));
    for (int j = 0; j < 3; j++) {
      for (int k = 0; k < 3; k++) nsum[f[j] * 3 + k] += n[k];
      ncnt[f[j]]++;
    }
  }
  for (int32_t i = 0; i < nv; i++) {
    if (ncnt[i] > 0) {
      for (int k = 0; k < 3; k++) nsum[i * 3 + k] /= (float)ncnt[i];
      normalize3(nsum + i * 3);
    } else {
      nsum[i * 3] = nsum[i * 3 + 1] = nsum[i * 3 + 2] = 0.0f;
    }
  }
  if (vertices) memcpy(vertices, vbuf, (size_t)nv * 3 * sizeof(float));
  if (faces) memcpy(faces, fbuf, (size_t)nf * 3 * sizeof(int32_t));
  if (normals && have_normals) memcpy(normals, nsum, (size_t)nv * 3 * sizeof(float));
  if (point_cloud) memset(point_cloud, 0, P * 3 * sizeof(float));
  if (normal_image) memset(normal_image, 0, P * 3 * sizeof(float));
  for (int32_t i = 0; i < nv; i++) {
    const size_t pix = (size_t)vxy[i * 2 + 1] * W + vxy[i * 2];
    if (point_cloud) memcpy(point_cloud + pix * 3, vbuf + i * 3, 3 * sizeof(float));
    if (normal_image && have_normals) memcpy(normal_image + pix * 3, nsum + i * 3, 3 * sizeof(float));
  }
  counts[0] = nv;
  counts[1] = nf;
  counts[2] = have_normals;
  free(added); free(vbuf); free(fbuf); free(vxy); free(nsum); free(ncnt);
  return 0;
}

int oracle_depth2mesh(const OracleCamera* cam, const float* depth, const uint8_t* color, int32_t cw,
                      int32_t ch, int32_t with_vertex_color, float max_connect_z_diff, int32_t x_step,
                      int32_t y_step, int32_t gl_coord, float* vertices, int32_t* faces, float* normals,
                      float* face_normals, float* uv, float* colors, float* point_cloud,
                      float* normal_image, int32_t* counts) {
  return depth2mesh_impl(cam, NULL, depth, color, cw, ch, with_vertex_color, max_connect_z_diff, x_step, y_step,
                         gl_coord, vertices, faces, normals, face_normals, uv, colors, point_cloud, normal_image,
                         counts);
}

int oracle_depth2mesh2(const OracleCamera* cam, const float* two_cam16, const float* depth, const uint8_t* color,
                       int32_t cw, int32_t ch, int32_t with_vertex_color, float max_connect_z_diff,
                       int32_t x_step, int32_t y_step, int32_t gl_coord, float* vertices, int32_t* faces,
                       float* normals, float* face_normals, float* uv, float* colors, float* point_cloud,
                       float* normal_image, int32_t* counts) {
  if (!two_cam16 || !color) return 1;
  return depth2mesh_impl(cam, two_cam16, depth, color, cw, ch, with_vertex_color, max_connect_z_diff, x_step,
                         y_step, gl_coord, vertices, faces, normals, face_normals, uv, colors, point_cloud,
                         normal_image, counts);
}

/* ---------------------------------------------------------------------------------------------------
 * G-buffer fields of ugu::GBuffer (include/ugu/renderer/base.h:34-77) as RendererGl::ReadGbuf derives
 * them (src/renderer/gl/renderer.cc:501-650) from the gbuf shaders (src/shader/glsl/vert/gbuf.vert:42-56,
 * geom/dummy.geom:26-80, frag/gbuf.frag:21-44), evaluated on the winners of the CPU Rasterizer above.
 *
 * PARITY UNPINNED: the reference computes these on an OpenGL pipeline, which cannot run on this box
 * (no graphics stack), and GL's sampling rule (pixel centres at +0.5, top-left fill, clipping) differs
 * from the Rasterizer's anyway.  This is a restatement of the shader / read-back arithmetic with a
 * fixed single-operation float order; its inputs (winner, depth, weights) ARE reference-pinned.
 * ------------------------------------------------------------------------------------------------- */
typedef struct OracleGBuffer {
  uint8_t* color;
  float* normal_cam;
  float* normal_wld;
  float* pos_cam;
  float* pos_wld;
  float* depth_01;
  uint8_t* stencil;
  int32_t* face_id;
  float* uv;
  float* bary;
  int32_t* geo_id;
} OracleGBuffer;

int oracle_gbuffer(const OracleMesh* mesh, const OracleCamera* cam, const OracleOption* opt, float z_near,
                   float z_far, int32_t flat_normal, int32_t geo_id, const OracleGBuffer* out) {
  const int W = cam->width, H = cam->height;
  if (W <= 0 || H <= 0) return 2;
  if (!(z_near > 0.0f) || !(z_far > z_near)) return 1;
  const size_t P = (size_t)W * (size_t)H;
  RawWinners raw;
  raw.fid = (int32_t*)malloc(P * 4);
  raw.depth = (float*)malloc(P * 4);
  raw.weight = (float*)malloc(P * 12);
  raw.backface = (uint8_t*)malloc(P);
  raw.erase = NULL;
  uint8_t* erase = (uint8_t*)calloc(P, 1);
  int32_t* tmp_fid = (int32_t*)malloc(P * 4);
  if (!raw.fid || !raw.depth || !raw.weight || !raw.backface || !erase || !tmp_fid) return 3;
  OracleOption o1 = *opt; /* pass 1: z-buffer only; colour options must not trip the validations */
  o1.diffuse_color = 0;
  o1.diffuse_shading = 0;
  if (mesh->num_face_normals > 0) o1.shading_normal = 0;
  int rc = render_impl(mesh, cam, &o1, NULL, NULL, NULL, NULL, tmp_fid, &raw);
  if (rc) goto done;

  {
    /* GetProjectionMatrixOpenGlForPinhole, src/util/camera_util.cc:254-255 */
    const volatile float sum = z_far + z_near, diff = z_far - z_near, two_f = 2.f * z_far;
    const volatile float two_fn = two_f * z_near;
    const float A = -sum / diff;
    const float B = -two_fn / diff;
    const int have_normals = mesh->num_normals > 0 && mesh->normal_indices != NULL;
    const int have_uv = mesh->num_uv > 0 && mesh->uv_indices != NULL;
    float* f3[6] = {out->normal_cam, out->normal_wld, out->pos_cam, out->pos_wld, out->uv, out->bary};
    for (int k = 0; k < 6; k++)
      if (f3[k]) memset(f3[k], 0, P * 12);
    if (out->stencil) memset(out->stencil, 0, P);
    if (out->face_id) memset(out->face_id, 0, P * 4);
    if (out->geo_id) memset(out->geo_id, 0, P * 4);
    for (size_t pi = 0; pi < P; pi++) {
      if (out->depth_01) out->depth_01[pi] = 1.0f; /* the cleared GL depth */
      const int32_t fid = raw.fid[pi];
      if (fid < 0) continue;
      if (opt->backface_culling && raw.backface[pi] == 255) continue; /* rasterizer.cc:248-252 */
      const float z = raw.depth[pi];
      const float ndc = (A * (-z) + B) / z;
      const float d = 0.5f * ndc + 0.5f;
      if (out->depth_01) out->depth_01[pi] = fminf(fmaxf(d, 0.0f), 1.0f); /* renderer.cc:541 */
      if (d <= 0.0f || 1.0f <= d) { /* renderer.cc:534-538; GL clips the fragment */
        erase[pi] = 1;
        continue;
      }
      if (out->stencil) out->stencil[pi] = 255;
      if (out->face_id) out->face_id[pi] = fid + 1; /* gbuf.frag:36, renderer.cc:621 */
      if (out->geo_id) out->geo_id[pi] = geo_id;    /* gbuf.frag:37, renderer.cc:615-620 */
      const float u = raw.weight[pi * 3 + 1], v = raw.weight[pi * 3 + 2];
      const float k0 = 1.0f - u - v;
      if (out->bary) { /* dummy.geom:69-75, renderer.cc:637-640 */
        out->bary[pi * 3 + 0] = u;
        out->bary[pi * 3 + 1] = v;
        out->bary[pi * 3 + 2] = k0;
      }
      const int32_t* vi = &mesh->vertex_indices[(size_t)fid * 3];
      const float* P0 = &mesh->vertices[(size_t)vi[0] * 3];
      const float* P1 = &mesh->vertices[(size_t)vi[1] * 3];
      const float* P2 = &mesh->vertices[(size_t)vi[2] * 3];
      float pw[3], pc[3];
      for (int k = 0; k < 3; k++) pw[k] = (k0 * P0[k] + u * P1[k]) + v * P2[k]; /* gbuf.vert:43,45 */
      mat3_mul(cam->w2c_R, pw, pc);
      for (int k = 0; k < 3; k++) pc[k] = cam->w2c_t[k] + pc[k]; /* renderer.cc:561 */
      if (out->pos_wld) memcpy(&out->pos_wld[pi * 3], pw, 12);
      if (out->pos_cam) memcpy(&out->pos_cam[pi * 3], pc, 12);
      float n[3] = {0.0f, 0.0f, 0.0f}, nc[3];
      if (flat_normal) { /* dummy.geom:58-59 */
        float a[3], b[3];
        for (int k = 0; k < 3; k++) {
          a[k] = P1[k] - P0[k];
          b[k] = P2[k] - P0[k];
        }
        normalize3(a);
        normalize3(b);
        n[0] = a[1] * b[2] - a[2] * b[1];
        n[1] = a[2] * b[0] - a[0] * b[2];
        n[2] = a[0] * b[1] - a[1] * b[0];
        normalize3(n);
      } else if (have_normals) { /* gbuf.frag:25 */
        const int32_t* ni = &mesh->normal_indices[(size_t)fid * 3];
        const float* N0 = &mesh->normals[(size_t)ni[0] * 3];
        const float* N1 = &mesh->normals[(size_t)ni[1] * 3];
        const float* N2 = &mesh->normals[(size_t)ni[2] * 3];
        for (int k = 0; k < 3; k++) n[k] = (k0 * N0[k] + u * N1[k]) + v * N2[k];
        normalize3(n);
      }
      mat3_mul(cam->w2c_R, n, nc); /* renderer.cc:583 */
      if (out->normal_wld) memcpy(&out->normal_wld[pi * 3], n, 12);
      if (out->normal_cam) memcpy(&out->normal_cam[pi * 3], nc, 12);
      if (have_uv && out->uv) { /* renderer.cc:641-644 */
        const int32_t* ui = &mesh->uv_indices[(size_t)fid * 3];
        const float* t0 = &mesh->uv[(size_t)ui[0] * 2];
        const float* t1 = &mesh->uv[(size_t)ui[1] * 2];
        const float* t2 = &mesh->uv[(size_t)ui[2] * 2];
        const float s = (k0 * t0[0] + u * t1[0]) + v * t2[0];
        const float t = (k0 * t0[1] + u * t1[1]) + v * t2[1];
        out->uv[pi * 3 + 0] = s;
        out->uv[pi * 3 + 1] = t;
        out->uv[pi * 3 + 2] = 1.0f - s - t;
      }
    }
  }
  if (out->color) { /* pass 2: the Rasterizer's colour with the clipped winners erased */
    RawWinners r2;
    memset(&r2, 0, sizeof(r2));
    r2.erase = erase;
    rc = render_impl(mesh, cam, opt, out->color, NULL, NULL, NULL, NULL, &r2);
  }
done:
  free(raw.fid);
  free(raw.depth);
  free(raw.weight);
  free(raw.backface);
  free(erase);
  free(tmp_fid);
  return rc;
}

/* ---------------------------------------------------------------------------------------------------
 * ugu::Raytracer::Impl::Render (src/renderer/cpu/raytracer.cc:117-225) with the nanort nearest-hit query
 * of BvhNanort::Intersect(ray, test_all = false) (include/ugu/accel/bvh_nanort.h:109-152).
 *
 * nanort (third_party/nanort/nanort.h, vendored in the reference tree) finds the hit with a BVH; the
 * value it returns is min over the triangles of the watertight intersector's t
 * (TriangleIntersector::Intersect / PrepareTraversal, nanort.h:1050-1199), because the traversal only
 * prunes boxes against the current nearest t (nanort.h:2485-2552).  This restatement therefore evaluates
 * the same intersector arithmetic on EVERY triangle (brute force) — usable on small cases only.
 * Traversal-order dependence that cannot be restated without the BVH: two triangles with bit-identical t
 * at a pixel (nanort keeps the LAST one visited, `tt > t` rejects; here the HIGHEST face index wins, which
 * is what nanort does too whenever the tied triangles share a leaf in index order), and
 * the non-conservative lower slab bound of IntersectRayAABB (nanort.h:2281-2321).  PINNED differentially:
 * tests/golden/raytrace_*.npz (compiled reference Raytracer) and the live reference where oracle/_ref exists.
 * ------------------------------------------------------------------------------------------------- */
typedef struct RtRay {
  float org[3], dir[3];
  float Sx, Sy, Sz;
  int kx, ky, kz;
} RtRay;

static void rt_prepare(RtRay* r) { /* nanort.h:1160-1199 */
  r->kz = 0;
  float absDir = fabsf(r->dir[0]);
  if (absDir < fabsf(r->dir[1])) {
    r->kz = 1;
    absDir = fabsf(r->dir[1]);
  }
  if (absDir < fabsf(r->dir[2])) {
    r->kz = 2;
    absDir = fabsf(r->dir[2]);
  }
  r->kx = r->kz + 1;
  if (r->kx == 3) r->kx = 0;
  r->ky = r->kx + 1;
  if (r->ky == 3) r->ky = 0;
  if (r->dir[r->kz] < 0.0f) {
    const int t = r->kx;
    r->kx = r->ky;
    r->ky = t;
  }
  r->Sx = r->dir[r->kx] / r->dir[r->kz];
  r->Sy = r->dir[r->ky] / r->dir[r->kz];
  r->Sz = 1.0f / r->dir[r->kz];
}

/* nanort.h:1050-1148; *t is the current nearest distance (in/out).  Returns 1 on an accepted hit. */
static int rt_intersect(const RtRay* r, const float* p0, const float* p1, const float* p2, float t_min,
                        float* t, float* u, float* v) {
  float A[3], B[3], Cc[3];
  for (int k = 0; k < 3; k++) {
    A[k] = p0[k] - r->org[k];
    B[k] = p1[k] - r->org[k];
    Cc[k] = p2[k] - r->org[k];
  }
  const float Ax = A[r->kx] - r->Sx * A[r->kz];
  const float Ay = A[r->ky] - r->Sy * A[r->kz];
  const float Bx = B[r->kx] - r->Sx * B[r->kz];
  const float By = B[r->ky] - r->Sy * B[r->kz];
  const float Cx = Cc[r->kx] - r->Sx * Cc[r->kz];
  const float Cy = Cc[r->ky] - r->Sy * Cc[r->kz];
  float U = Cx * By - Cy * Bx;
  float V = Ax * Cy - Ay * Cx;
  float W = Bx * Ay - By * Ax;
  if (U == 0.0f || V == 0.0f || W == 0.0f) {
    const double CxBy = (double)Cx * (double)By, CyBx = (double)Cy * (double)Bx;
    U = (float)(CxBy - CyBx);
    const double AxCy = (double)Ax * (double)Cy, AyCx = (double)Ay * (double)Cx;
    V = (float)(AxCy - AyCx);
    const double BxAy = (double)Bx * (double)Ay, ByAx = (double)By * (double)Ax;
    W = (float)(BxAy - ByAx);
  }
  if (U < 0.0f || V < 0.0f || W < 0.0f) {
    if (U > 0.0f || V > 0.0f || W > 0.0f) return 0; /* cull_back_face is false (BVHTraceOptions default) */
  }
  const float det = U + V + W;
  if (det == 0.0f) return 0;
  const float Az = r->Sz * A[r->kz];
  const float Bz = r->Sz * B[r->kz];
  const float Cz = r->Sz * Cc[r->kz];
  const float D = U * Az + V * Bz + W * Cz;
  const float rcpDet = 1.0f / det;
  const float tt = D * rcpDet;
  if (tt > *t) return 0;
  if (tt < t_min) return 0;
  *t = tt;
  *u = V * rcpDet;
  *v = W * rcpDet;
  return 1;
}

/* org_w = PinholeCamera::org_ray_w = c2w translation cast to float (camera.h:431-436). */
int oracle_raytrace(const OracleMesh* mesh, const OracleCamera* cam, const float* org_w,
                    const OracleOption* opt, uint8_t* color, float* depth, float* normal, uint8_t* mask,
                    int32_t* face_id) {
  const int W = cam->width, H = cam->height;
  const size_t P = (size_t)W * (size_t)H;
  const int F = mesh->num_faces;
  if (W <= 0 || H <= 0) return 2;
  /* util_private.cc:19-67 (shared with the Rasterizer) */
  if (opt->backface_culling && mesh->num_face_normals == 0) return 1;
  if (opt->diffuse_color == 1 && mesh->num_materials == 0) return 1;
  if (opt->diffuse_color == 1) {
    for (int i = 0; i < mesh->num_materials; i++)
      if (mesh->tex_data[i] == NULL || mesh->tex_w[i] < 1 || mesh->tex_h[i] < 1) return 1;
  }
  if (opt->diffuse_color == 2 && mesh->num_colors == 0) return 1;
  if (opt->shading_normal == 0 && mesh->num_face_normals == 0) return 1;
  if (opt->shading_normal == 1 && mesh->num_normals == 0) return 1;
  if (!color && !depth && !normal && !mask && !face_id) return 1;
  if (color) memset(color, 0, P * 3);
  if (depth) memset(depth, 0, P * 4);
  if (normal) memset(normal, 0, P * 12);
  if (mask) memset(mask, 0, P);
  if (face_id)
    for (size_t i = 0; i < P; i++) face_id[i] = -1;
  const float sigma = opt->oren_nayar_sigma, sigma_2 = sigma * sigma;
  const float on_A = 1.0f - (0.5f * sigma_2 / (sigma_2 + 0.33f));
  const float on_B = 0.45f * sigma_2 / (sigma_2 + 0.09f);

  for (int y = 0; y < H; y++) {
    for (int x = 0; x < W; x++) {
      RtRay r;
      ray_w(cam, x, y, r.dir);
      memcpy(r.org, org_w, 12);
      rt_prepare(&r);
      float t = 1.0e+30f, u = 0.0f, v = 0.0f; /* bvh_nanort.h:119-121: min_t 1e-4, max_t 1e30 */
      int32_t fid = -1;
      for (int i = 0; i < F; i++) {
        const int32_t* f = &mesh->vertex_indices[(size_t)i * 3];
        float tl = t, ul, vl;
        if (rt_intersect(&r, &mesh->vertices[(size_t)f[0] * 3], &mesh->vertices[(size_t)f[1] * 3],
                         &mesh->vertices[(size_t)f[2] * 3], 0.0001f, &tl, &ul, &vl)) {
          t = tl;
          u = ul;
          v = vl;
          fid = i;
        }
      }
      if (fid < 0 || !(t < 1.0e+30f)) continue; /* nanort.h:2548 */
      const float* ray = r.dir;
      if (opt->backface_culling && dot3(&mesh->face_normals[(size_t)fid * 3], ray) > 0.0f) continue;
      const size_t pi = (size_t)y * W + x;
      if (face_id) face_id[pi] = fid;
      if (mask) mask[pi] = 255;
      if (depth) { /* raytracer.cc:167-172 */
        float hw[3], hc[3];
        for (int k = 0; k < 3; k++) hw[k] = r.org[k] + ray[k] * t;
        mat3_mul(cam->w2c_R, hw, hc);
        depth[pi] = (hc[2] + cam->w2c_t[2]) * opt->depth_scale;
      }
      float n_w[3] = {0.0f, 0.0f, 0.0f};
      if (opt->shading_normal == 0) {
        memcpy(n_w, &mesh->face_normals[(size_t)fid * 3], 12);
      } else if (opt->shading_normal == 1) { /* raytracer.cc:178-186 */
        const int32_t* ni = &mesh->normal_indices[(size_t)fid * 3];
        const float* n0 = &mesh->normals[(size_t)ni[0] * 3];
        const float* n1 = &mesh->normals[(size_t)ni[1] * 3];
        const float* n2 = &mesh->normals[(size_t)ni[2] * 3];
        const float k0 = 1.0f - u - v;
        for (int k = 0; k < 3; k++) n_w[k] = (k0 * n0[k] + u * n1[k]) + v * n2[k];
      }
      if (normal) mat3_mul(cam->w2c_R, n_w, &normal[pi * 3]);
      if (color) shade_pixel(mesh, opt, on_A, on_B, fid, u, v, ray, n_w, &color[pi * 3]);
    }
  }
  return 0;
}


/* ================================================================================================
 * Visibility pass: ugu::VisibilityTester's per-keyframe vertex / face tests
 * (src/texturing/visibility_tester.cc:416-529 TestVertex, 531-549 TestVertices, 551-604 TestFaces,
 * 606-668 TestFacewiseVertices) and the folds of FaceInfo::CalcStat (:185-218) / VertexInfo::CalcStat
 * (:66-152, the min / mean subset), restated with ONE substitution: the reference asks a BVH for the
 * nearest hit of the ray through the vertex's sub-pixel projection (:457-483); this restatement — like the
 * CUDA consumer it checks — reads `zbuf`, the mesh's own depth map rendered from the keyframe camera by
 * the rasterizer (back-face culling off), at the nearest pixel:
 *     t_v = |camera_p|                      distance of the (offset) vertex along its own ray
 *     t_s = zbuf(nn) * (t_v / camera_p.z)   distance of the surface seen at that pixel along that ray
 *     occluded  iff  zbuf(nn) >= FLT_MIN  and  t_v - t_s > eps        (reference: |hit - vertex| > eps)
 *     distance  =  t_v                                               (reference: the BVH hit's t)
 *     depth check uses camera_p.z                                     (reference: the hit's camera z)
 * Everything else follows the reference expression by expression.  Parity status: PINNED on the pairs
 * both methods call visible — tests/test_visibility_tester.py compares with fixtures generated by the
 * compiled reference tester (projected_pos, viewing_angle, colour, intensity, area bit-exact; distance to
 * float tolerance) and states the size of the disagreement set (sub-pixel rays vs the z-buffer at
 * silhouettes).
 * ================================================================================================ */
typedef struct OracleVisCamera {
  int32_t width, height;
  float fx, fy, cx, cy;
  float w2c[12];   /* row-major 3x4: camera->w2c().cast<float>() */
  float c2w_R[9];  /* c2w_R_f_ (camera.h:321) */
  float c2w_t[3];  /* c2w_t_f_ (camera.h:322) = org_ray_w */
} OracleVisCamera;

typedef struct OracleVisOption {
  float backface_culling_angle_rad_th;
  int32_t use_mask, use_depth, interp;
  float max_viewing_angle, max_distance_from_camera_to_vertex, max_depth_difference, eps;
  int32_t facewise;
  int32_t reserved;
} OracleVisOption;

typedef struct OracleVisVertex {
  float color[3];
  float intensity;
  float projected[2];
  float viewing_angle;
  float distance;
} OracleVisVertex;

/* image_util.h:156-193 (T = Vec3b) */
static void vis_bilinear(float x, float y, const uint8_t* img, int cols, int rows, float* out) {
  int x0 = (int)floorf(x), y0 = (int)floorf(y);
  int x1 = x0 + 1, y1 = y0 + 1;
  if (x0 < 0) x0 = 0;
  if (y0 < 0) y0 = 0;
  if (cols <= x1) x1 = cols - 1;
  if (rows <= y1) y1 = rows - 1;
  const float lu = x - (float)x0, lv = y - (float)y0;
  for (int i = 0; i < 3; i++) {
    const float c = (1.0f - lu) * (1.0f - lv) * (float)img[((size_t)y0 * cols + x0) * 3 + i] +
                    lu * (1.0f - lv) * (float)img[((size_t)y0 * cols + x1) * 3 + i] +
                    (1.0f - lu) * lv * (float)img[((size_t)y1 * cols + x0) * 3 + i] +
                    lu * lv * (float)img[((size_t)y1 * cols + x1) * 3 + i];
    out[i] = (float)f2u8(c);  /* static_cast<unsigned char>(colorf), then widened back to float */
  }
}

/* visibility_tester.cc:416-529 with the substitution above.  world_p = vertex + offset. */
static int vis_test_vertex(const OracleVisCamera* cam, const OracleVisOption* opt, const float* vtx,
                           const float* nrm, const float* offset, const uint8_t* color, const float* depth,
                           const uint8_t* mask, const float* zbuf, OracleVisVertex* out) {
  float wp[3] = {vtx[0] + offset[0], vtx[1] + offset[1], vtx[2] + offset[2]};
  /* Affine3f * Vector3f (Eigen/src/Geometry/Transform.h, transform_right_product_impl): res = translation;
   * res.noalias() += linear * v, the 3x3 * 3x1 product accumulated column by column — the association that
   * reproduces the compiled reference bit for bit is ((m0 x + m1 y) + m2 z) + t (probed on 3045 projections). */
  float cp[3];
  for (int r = 0; r < 3; r++)
    cp[r] = ((cam->w2c[r * 4 + 0] * wp[0] + cam->w2c[r * 4 + 1] * wp[1]) + cam->w2c[r * 4 + 2] * wp[2]) + cam->w2c[r * 4 + 3];
  /* PinholeCamera::Project (camera.h:380-387) */
  const float ix = cam->fx / cp[2] * cp[0] + cam->cx;
  const float iy = cam->fy / cp[2] * cp[1] + cam->cy;
  const float d = cp[2];
  if (d < 0 || ix < 0 || (float)(cam->width - 1) < ix || iy < 0 || (float)(cam->height - 1) < iy) return 0;
  const int nn_x = (int)roundf(ix), nn_y = (int)roundf(iy);
  const size_t nn = (size_t)nn_y * cam->width + nn_x;
  if (opt->use_mask && mask[nn] != 255) return 0;
  if (opt->use_depth && depth[nn] < FLT_EPSILON) return 0;
  /* ray_w(float, float): camera.h:438-448 */
  float dir[3] = {(ix - cam->cx) / cam->fx, (iy - cam->cy) / cam->fy, 1.0f};
  const float n2 = dir[0] * dir[0] + (dir[1] * dir[1] + dir[2] * dir[2]);
  if (n2 > 0.0f) {
    const float s = sqrtf(n2);
    dir[0] = dir[0] / s;
    dir[1] = dir[1] / s;
    dir[2] = dir[2] / s;
  }
  float ray[3];
  mat3_mul(cam->c2w_R, dir, ray);
  const float dot = -dot3(nrm, ray);
  const float viewing_angle = acosf(dot);
  if (opt->backface_culling_angle_rad_th < viewing_angle) return 0;
  /* the BVH query of :457-483, replaced by the z-buffer */
  const float t_v = sqrtf(cp[0] * cp[0] + (cp[1] * cp[1] + cp[2] * cp[2]));
  if (zbuf) {
    const float z_s = zbuf[nn];
    if (z_s >= FLT_MIN) {
      const float t_s = z_s * (t_v / cp[2]);
      if (t_v - t_s > opt->eps) return 0;
    }
  }
  if (opt->use_depth) {
    const float diff = fabsf(cp[2] - depth[nn]);
    if (opt->max_depth_difference < diff) return 0;
  }
  out->projected[0] = ix;
  out->projected[1] = iy;
  out->viewing_angle = viewing_angle;
  out->distance = t_v;
  if (out->distance > opt->max_distance_from_camera_to_vertex || out->viewing_angle > opt->max_viewing_angle) return 0;
  if (opt->interp == 0) {
    for (int i = 0; i < 3; i++) out->color[i] = (float)color[nn * 3 + i];
  } else {
    vis_bilinear(ix, iy, color, cam->width, cam->height, out->color);
  }
  /* Color2Gray (image_util.h:228-244): double arithmetic, cast to float */
  out->intensity = (float)(0.2989 * out->color[0] + 0.5870 * out->color[1] + 0.1140 * out->color[2]);
  return 1;
}

/* One keyframe.  Outputs are dense: v_visible[V] (0/1) + v_rec[V]; f_visible[F] + f_area / f_angle /
 * f_distance [F] + f_tri [F*6]; rows of invisible elements are left untouched. */
int oracle_visibility_keyframe(int32_t V, const float* vertices, const float* normals, int32_t F,
                               const int32_t* indices, const OracleVisCamera* cam, const OracleVisOption* opt,
                               const uint8_t* color, const float* depth, const uint8_t* mask, const float* zbuf,
                               uint8_t* v_visible, OracleVisVertex* v_rec, uint8_t* f_visible, float* f_area,
                               float* f_angle, float* f_distance, float* f_tri) {
  if (!color || (opt->use_mask && !mask) || (opt->use_depth && !depth)) return 1;
  const float zero[3] = {0.f, 0.f, 0.f};
  if (!opt->facewise) {
    for (int32_t v = 0; v < V; v++) {  /* TestVertices */
      OracleVisVertex r;
      if (vis_test_vertex(cam, opt, vertices + 3 * (size_t)v, normals + 3 * (size_t)v, zero, color, depth, mask, zbuf, &r)) {
        v_visible[v] = 1;
        v_rec[v] = r;
      }
    }
  }
  for (int32_t i = 0; i < F; i++) {  /* TestFaces / TestFacewiseVertices */
    float angle = 0.0f, dist = 0.0f, tri[6];
    int ok = 1;
    float center[3] = {0.f, 0.f, 0.f};
    if (opt->facewise) {  /* face_centers_ (PrepareData, :259-266): setZero, += each corner, /= 3 */
      for (int j = 0; j < 3; j++)
        for (int k = 0; k < 3; k++) center[k] += vertices[3 * (size_t)indices[3 * (size_t)i + j] + k];
      for (int k = 0; k < 3; k++) center[k] /= 3;
    }
    for (int j = 0; j < 3 && ok; j++) {
      const int32_t vid = indices[3 * (size_t)i + j];
      const float* vtx = vertices + 3 * (size_t)vid;
      OracleVisVertex r;
      if (opt->facewise) {
        float off[3];
        for (int k = 0; k < 3; k++) off[k] = (center[k] - vtx[k]) * 1e-3f;
        ok = vis_test_vertex(cam, opt, vtx, normals + 3 * (size_t)vid, off, color, depth, mask, zbuf, &r);
        if (ok && !v_visible[vid]) {  /* "Update only once per vertex", sequential order */
          v_visible[vid] = 1;
          v_rec[vid] = r;
        }
      } else {
        ok = v_visible[vid];
        if (ok) r = v_rec[vid];
      }
      if (!ok) break;
      angle += r.viewing_angle;
      dist += r.distance;
      tri[2 * j] = r.projected[0];
      tri[2 * j + 1] = r.projected[1];
    }
    if (!ok) continue;
    f_visible[i] = 1;
    f_angle[i] = angle / 3;
    f_distance[i] = dist / 3;
    f_area[i] = 0.5f * fabsf(edge_function(tri, tri + 2, tri + 4));
    for (int k = 0; k < 6; k++) f_tri[6 * (size_t)i + k] = tri[k];
  }
  return 0;
}

int oracle_abi_version(void) { return 7; }
