// ORACLE — TEST INFRASTRUCTURE ONLY (see small_linalg.hpp header).
//
// SE(3) pieces the reference takes from GTSAM (un-vendored; CMakeLists.txt:25 unpinned,
// extern/svn_ndt/CMakeLists.txt:11 says 4.0).  Restated from GTSAM's published formulas with the
// 4.2 constants; Pose3::retract is pinned to the FULL exponential map (GTSAM_POSE3_EXPMAP=ON, the
// 4.2 default) -- not determinable from the tree, so SVN PARITY IS UNPINNED against a reference
// binary.  Used at svn_ndt_impl.hpp:222,237 (Logmap/between), :715,854,870 (retract),
// :765-767 (rpy/translation), :761 (matrix), :867,878,914 (Logmap).
#pragma once
#include <cmath>
#include <cstring>
#include <limits>

namespace oracle {

struct Pose3 {
  double R[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};  // row-major
  double t[3] = {0, 0, 0};
};

inline void m3mul(const double* a, const double* b, double* o) {
  double r[9];
  for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) r[3 * i + j] = (a[3 * i] * b[j] + a[3 * i + 1] * b[3 + j]) + a[3 * i + 2] * b[6 + j];
  std::memcpy(o, r, sizeof(r));
}
inline void m3vec(const double* a, const double* v, double* o) {
  double r[3];
  for (int i = 0; i < 3; ++i) r[i] = (a[3 * i] * v[0] + a[3 * i + 1] * v[1]) + a[3 * i + 2] * v[2];
  std::memcpy(o, r, sizeof(r));
}
inline void skew(const double* w, double* W) { W[0] = 0; W[1] = -w[2]; W[2] = w[1]; W[3] = w[2]; W[4] = 0; W[5] = -w[0]; W[6] = -w[1]; W[7] = w[0]; W[8] = 0; }
inline void cross3(const double* a, const double* b, double* o) { double r[3] = {a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0]}; std::memcpy(o, r, sizeof(r)); }

// gtsam::so3::ExpmapFunctor::expmap
inline void so3_expmap(const double* w, double* R) {
  const double theta2 = (w[0] * w[0] + w[1] * w[1]) + w[2] * w[2];
  double W[9]; skew(w, W);
  if (theta2 <= std::numeric_limits<double>::epsilon()) {
    double WW[9]; m3mul(W, W, WW);
    for (int i = 0; i < 9; ++i) R[i] = ((i % 4 == 0) ? 1.0 : 0.0) + W[i] + 0.5 * WW[i];
    return;
  }
  const double theta = std::sqrt(theta2), sin_theta = std::sin(theta), s2 = std::sin(theta / 2.0);
  const double one_minus_cos = 2.0 * s2 * s2;
  double K[9], KK[9];
  for (int i = 0; i < 9; ++i) K[i] = W[i] / theta;
  m3mul(K, K, KK);
  for (int i = 0; i < 9; ++i) R[i] = ((i % 4 == 0) ? 1.0 : 0.0) + sin_theta * K[i] + one_minus_cos * KK[i];
}

// gtsam::SO3::Logmap
inline void so3_logmap(const double* R, double* w) {
  const double R11 = R[0], R12 = R[1], R13 = R[2], R21 = R[3], R22 = R[4], R23 = R[5], R31 = R[6], R32 = R[7], R33 = R[8];
  const double tr = (R11 + R22) + R33;
  if (tr + 1.0 < 1e-3) {  // theta ~ pi
    double v[3], r;
    if (R33 > R22 && R33 > R11) { r = std::sqrt(2.0 * (1.0 + R33)); v[0] = R13 / r; v[1] = R23 / r; v[2] = (1.0 + R33) / r; }
    else if (R22 > R11) { r = std::sqrt(2.0 * (1.0 + R22)); v[0] = R12 / r; v[1] = (1.0 + R22) / r; v[2] = R32 / r; }
    else { r = std::sqrt(2.0 * (1.0 + R11)); v[0] = (1.0 + R11) / r; v[1] = R21 / r; v[2] = R31 / r; }
    const double n = std::sqrt((v[0] * v[0] + v[1] * v[1]) + v[2] * v[2]);
    for (int i = 0; i < 3; ++i) w[i] = M_PI * v[i] / n;
    return;
  }
  double magnitude;
  const double tr_3 = tr - 3.0;
  if (tr_3 < -1e-6) { const double theta = std::acos((tr - 1.0) / 2.0); magnitude = theta / (2.0 * std::sin(theta)); }
  else magnitude = 0.5 - tr_3 / 12.0 + tr_3 * tr_3 / 60.0;
  w[0] = magnitude * (R32 - R23); w[1] = magnitude * (R13 - R31); w[2] = magnitude * (R21 - R12);
}

// gtsam::Pose3::Expmap, xi = (omega, v)
inline Pose3 pose_expmap(const double* xi) {
  Pose3 p;
  const double* w = xi; const double* v = xi + 3;
  so3_expmap(w, p.R);
  const double theta2 = (w[0] * w[0] + w[1] * w[1]) + w[2] * w[2];
  if (theta2 > std::numeric_limits<double>::epsilon()) {
    const double wv = (w[0] * v[0] + w[1] * v[1]) + w[2] * v[2];
    double t_par[3] = {w[0] * wv, w[1] * wv, w[2] * wv}, wxv[3], Rwxv[3];
    cross3(w, v, wxv);
    m3vec(p.R, wxv, Rwxv);
    for (int i = 0; i < 3; ++i) p.t[i] = (wxv[i] - Rwxv[i] + t_par[i]) / theta2;
  } else {
    for (int i = 0; i < 3; ++i) p.t[i] = v[i];
  }
  return p;
}

// gtsam::Pose3::Logmap -> (omega, u)
inline void pose_logmap(const Pose3& p, double* xi) {
  double w[3]; so3_logmap(p.R, w);
  const double t = std::sqrt((w[0] * w[0] + w[1] * w[1]) + w[2] * w[2]);
  if (t < 1e-10) { for (int i = 0; i < 3; ++i) { xi[i] = w[i]; xi[3 + i] = p.t[i]; } return; }
  double wn[3] = {w[0] / t, w[1] / t, w[2] / t}, W[9], WT[3], WWT[3];
  skew(wn, W);
  const double Tan = std::tan(0.5 * t);
  m3vec(W, p.t, WT);
  m3vec(W, WT, WWT);
  for (int i = 0; i < 3; ++i) { xi[i] = w[i]; xi[3 + i] = p.t[i] - (0.5 * t) * WT[i] + (1.0 - t / (2.0 * Tan)) * WWT[i]; }
}

inline Pose3 pose_compose(const Pose3& a, const Pose3& b) {
  Pose3 o; m3mul(a.R, b.R, o.R);
  double Rt[3]; m3vec(a.R, b.t, Rt);
  for (int i = 0; i < 3; ++i) o.t[i] = a.t[i] + Rt[i];
  return o;
}
// a.between(b) = a^-1 * b
inline Pose3 pose_between(const Pose3& a, const Pose3& b) {
  double Rt[9];
  for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) Rt[3 * i + j] = a.R[3 * j + i];
  Pose3 o; m3mul(Rt, b.R, o.R);
  double d[3] = {b.t[0] - a.t[0], b.t[1] - a.t[1], b.t[2] - a.t[2]};
  m3vec(Rt, d, o.t);
  return o;
}
inline Pose3 pose_retract(const Pose3& p, const double* xi) { return pose_compose(p, pose_expmap(xi)); }

// gtsam::Rot3::rpy() via RQ (gtsam/geometry/Rot3.cpp RQ())
inline void rot_rpy(const double* A, double* rpy) {
  const double x = -std::atan2(-A[7], A[8]);
  { }
  const double cx = std::cos(-x), sx = std::sin(-x);
  const double Qx[9] = {1, 0, 0, 0, cx, -sx, 0, sx, cx};
  double B[9]; m3mul(A, Qx, B);
  const double y = -std::atan2(B[6], B[8]);
  const double cy = std::cos(-y), sy = std::sin(-y);
  const double Qy[9] = {cy, 0, sy, 0, 1, 0, -sy, 0, cy};
  double C[9]; m3mul(B, Qy, C);
  const double z = -std::atan2(-C[3], C[4]);
  rpy[0] = x; rpy[1] = y; rpy[2] = z;
}

inline void pose_from_matrix16(const double* M, Pose3& p) {
  for (int r = 0; r < 3; ++r) { for (int c = 0; c < 3; ++c) p.R[3 * r + c] = M[4 * r + c]; p.t[r] = M[4 * r + 3]; }
}
inline void pose_to_matrix16(const Pose3& p, double* M) {
  for (int r = 0; r < 3; ++r) { for (int c = 0; c < 3; ++c) M[4 * r + c] = p.R[3 * r + c]; M[4 * r + 3] = p.t[r]; }
  M[12] = 0; M[13] = 0; M[14] = 0; M[15] = 1;
}

}  // namespace oracle
