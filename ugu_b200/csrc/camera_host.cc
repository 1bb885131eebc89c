// Host-side camera constants for the C-ABI (ugu_b200_camera_from_c2w).
//
// Mirrors what the reference derives on the host before its vertex loop:
//   w2c      = c2w.inverse()                               (include/ugu/camera.h:317-319)
//   w2c_R    = w2c.rotation().cast<float>()                (src/renderer/cpu/rasterizer.cc:118)
//   w2c_t    = w2c.translation().cast<float>()             (rasterizer.cc:119)
//   c2w_R_f  = c2w.matrix().block<3,3>(0,0).cast<float>()  (camera.h:321)
// Eigen's Affine3d::rotation() is the SVD polar factor U * diag(1,1,±1) * V^T of the linear part
// (third_party/eigen/Eigen/src/Geometry/Transform.h:1046-1055,1076-1082).  Everything here is
// double precision and only the final values are rounded to float, so the result equals Eigen's
// except when a double-rounding difference (~1e-16) straddles a float rounding boundary; the
// golden fixtures in tests/golden pin that it does not for the cameras the tests use.  The C++
// adapter (ugu_b200/cpp/rasterizer_cuda.cc) does not use this file: it takes the constants from the
// ugu::Camera object itself.
#include <cmath>
#include <cstdint>
#include <cstring>

#include "ugu_b200_raster.h"

namespace {

struct M3 {
  double m[3][3];
};

M3 Mul(const M3& a, const M3& b) {
  M3 r;
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) r.m[i][j] = a.m[i][0] * b.m[0][j] + a.m[i][1] * b.m[1][j] + a.m[i][2] * b.m[2][j];
  return r;
}

M3 Transpose(const M3& a) {
  M3 r;
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) r.m[i][j] = a.m[j][i];
  return r;
}

double Det(const M3& a) {
  return a.m[0][0] * (a.m[1][1] * a.m[2][2] - a.m[1][2] * a.m[2][1]) -
         a.m[0][1] * (a.m[1][0] * a.m[2][2] - a.m[1][2] * a.m[2][0]) +
         a.m[0][2] * (a.m[1][0] * a.m[2][1] - a.m[1][1] * a.m[2][0]);
}

// Cofactor inverse, as Eigen does for fixed 3x3 (Eigen/src/LU/InverseImpl.h, compute_inverse<.,.,3>).
bool Inverse(const M3& a, M3* out) {
  const double c00 = a.m[1][1] * a.m[2][2] - a.m[1][2] * a.m[2][1];
  const double c10 = a.m[1][2] * a.m[2][0] - a.m[1][0] * a.m[2][2];
  const double c20 = a.m[1][0] * a.m[2][1] - a.m[1][1] * a.m[2][0];
  const double det = a.m[0][0] * c00 + a.m[0][1] * c10 + a.m[0][2] * c20;
  if (det == 0.0 || !std::isfinite(det)) return false;
  const double inv = 1.0 / det;
  out->m[0][0] = c00 * inv;
  out->m[1][0] = c10 * inv;
  out->m[2][0] = c20 * inv;
  out->m[0][1] = (a.m[0][2] * a.m[2][1] - a.m[0][1] * a.m[2][2]) * inv;
  out->m[1][1] = (a.m[0][0] * a.m[2][2] - a.m[0][2] * a.m[2][0]) * inv;
  out->m[2][1] = (a.m[0][1] * a.m[2][0] - a.m[0][0] * a.m[2][1]) * inv;
  out->m[0][2] = (a.m[0][1] * a.m[1][2] - a.m[0][2] * a.m[1][1]) * inv;
  out->m[1][2] = (a.m[0][2] * a.m[1][0] - a.m[0][0] * a.m[1][2]) * inv;
  out->m[2][2] = (a.m[0][0] * a.m[1][1] - a.m[0][1] * a.m[1][0]) * inv;
  return true;
}

// Symmetric 3x3 eigen-decomposition by cyclic Jacobi rotations: S = V * diag(w) * V^T.
void JacobiEigen(M3 s, M3* v, double w[3]) {
  M3 q = {{{1, 0, 0}, {0, 1, 0}, {0, 0, 1}}};
  for (int sweep = 0; sweep < 64; sweep++) {
    const double off = std::fabs(s.m[0][1]) + std::fabs(s.m[0][2]) + std::fabs(s.m[1][2]);
    if (off < 1e-300) break;
    for (int p = 0; p < 2; p++) {
      for (int r = p + 1; r < 3; r++) {
        if (s.m[p][r] == 0.0) continue;
        const double theta = (s.m[r][r] - s.m[p][p]) / (2.0 * s.m[p][r]);
        const double t = (theta >= 0 ? 1.0 : -1.0) / (std::fabs(theta) + std::sqrt(theta * theta + 1.0));
        const double c = 1.0 / std::sqrt(t * t + 1.0), sn = t * c;
        M3 j = {{{1, 0, 0}, {0, 1, 0}, {0, 0, 1}}};
        j.m[p][p] = c;
        j.m[r][r] = c;
        j.m[p][r] = sn;
        j.m[r][p] = -sn;
        s = Mul(Mul(Transpose(j), s), j);
        q = Mul(q, j);
      }
    }
  }
  *v = q;
  for (int i = 0; i < 3; i++) w[i] = s.m[i][i];
}

// Polar rotation factor of `a` with Eigen's sign convention (proper rotation, the smallest
// singular direction flipped when det < 0).
bool PolarRotation(const M3& a, M3* rot) {
  M3 v;
  double w[3];
  JacobiEigen(Mul(Transpose(a), a), &v, w);
  // sort singular values descending (Eigen's JacobiSVD order), permuting V's columns
  int order[3] = {0, 1, 2};
  for (int i = 0; i < 2; i++)
    for (int j = i + 1; j < 3; j++)
      if (w[order[j]] > w[order[i]]) {
        int t = order[i];
        order[i] = order[j];
        order[j] = t;
      }
  M3 vs, u;
  for (int k = 0; k < 3; k++) {
    const double sv = std::sqrt(w[order[k]] > 0 ? w[order[k]] : 0.0);
    if (!(sv > 0.0)) return false;
    for (int i = 0; i < 3; i++) vs.m[i][k] = v.m[i][order[k]];
    for (int i = 0; i < 3; i++)
      u.m[i][k] = (a.m[i][0] * vs.m[0][k] + a.m[i][1] * vs.m[1][k] + a.m[i][2] * vs.m[2][k]) / sv;
  }
  M3 r = Mul(u, Transpose(vs));
  if (Det(r) < 0) {
    for (int i = 0; i < 3; i++) u.m[i][2] = -u.m[i][2];
    r = Mul(u, Transpose(vs));
  }
  // One Newton step toward the orthogonal factor removes the O(eps * cond) error of forming A^T A.
  M3 rinv;
  if (Inverse(r, &rinv)) {
    const M3 rit = Transpose(rinv);
    for (int i = 0; i < 3; i++)
      for (int j = 0; j < 3; j++) r.m[i][j] = 0.5 * (r.m[i][j] + rit.m[i][j]);
  }
  *rot = r;
  return true;
}

}  // namespace

extern "C" int ugu_b200_camera_from_c2w(const double* c2w16, int32_t width, int32_t height, float fx,
                                        float fy, float cx, float cy, ugu_b200_camera* out) {
  if (c2w16 == nullptr || out == nullptr) return UGU_B200_E_INVALID;
  M3 lin;
  double t[3];
  for (int i = 0; i < 3; i++) {
    for (int j = 0; j < 3; j++) lin.m[i][j] = c2w16[i * 4 + j];
    t[i] = c2w16[i * 4 + 3];
  }
  M3 inv;
  if (!Inverse(lin, &inv)) return UGU_B200_E_INVALID;
  M3 rot;
  if (!PolarRotation(inv, &rot)) return UGU_B200_E_INVALID;
  std::memset(out, 0, sizeof(*out));
  out->width = width;
  out->height = height;
  out->fx = fx;
  out->fy = fy;
  out->cx = cx;
  out->cy = cy;
  for (int i = 0; i < 3; i++) {
    for (int j = 0; j < 3; j++) {
      out->w2c_R[i * 3 + j] = static_cast<float>(rot.m[i][j]);
      out->c2w_R[i * 3 + j] = static_cast<float>(lin.m[i][j]);
    }
    // Affine inverse translation: (-inv) * t   (Eigen/src/Geometry/Transform.h, inverse(Affine): unary minus on
    // the matrix, so a zero sum keeps Eigen's sign of zero)
    out->w2c_t[i] = static_cast<float>((-inv.m[i][0]) * t[0] + (-inv.m[i][1]) * t[1] + (-inv.m[i][2]) * t[2]);
  }
  return UGU_B200_OK;
}

extern "C" int ugu_b200_keyframe_camera_from_c2w(const double* c2w16, int32_t width, int32_t height, float fx,
                                                 float fy, float cx, float cy, ugu_b200_keyframe* kf) {
  if (c2w16 == nullptr || kf == nullptr) return UGU_B200_E_INVALID;
  const int rc = ugu_b200_camera_from_c2w(c2w16, width, height, fx, fy, cx, cy, &kf->cam);
  if (rc != UGU_B200_OK) return rc;
  M3 lin, inv;
  double t[3];
  for (int i = 0; i < 3; i++) {
    for (int j = 0; j < 3; j++) lin.m[i][j] = c2w16[i * 4 + j];
    t[i] = c2w16[i * 4 + 3];
  }
  if (!Inverse(lin, &inv)) return UGU_B200_E_INVALID;
  // w2c_ = c2w_.inverse() (Affine: [L^-1 | -(L^-1 t)]), then .cast<float>() (visibility_tester.cc:422)
  for (int i = 0; i < 3; i++) {
    for (int j = 0; j < 3; j++) kf->w2c[i * 4 + j] = static_cast<float>(inv.m[i][j]);
    kf->w2c[i * 4 + 3] = static_cast<float>((-inv.m[i][0]) * t[0] + (-inv.m[i][1]) * t[1] + (-inv.m[i][2]) * t[2]);
    kf->c2w_t[i] = static_cast<float>(t[i]);
  }
  return UGU_B200_OK;
}
