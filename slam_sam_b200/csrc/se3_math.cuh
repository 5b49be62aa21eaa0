// SE(3) helpers for the SVN path (host + device).  The reference takes these from GTSAM
// (svn_ndt_impl.hpp:222,237,715,765-767,854,867-870,878,914); restated from the published formulas
// with GTSAM 4.2 constants and the full exponential map as the retraction.
#pragma once
#include "newton_math.cuh"

namespace ndtb {

struct Pose {
  double R[9];
  double t[3];
};

NM_HD void d33_mul(const double* a, const double* b, double* o) {
  double r[9];
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) r[3 * i + j] = (a[3 * i] * b[j] + a[3 * i + 1] * b[3 + j]) + a[3 * i + 2] * b[6 + j];
  for (int i = 0; i < 9; ++i) o[i] = r[i];
}
NM_HD void d33_vec(const double* a, const double* v, double* o) {
  double r[3];
  for (int i = 0; i < 3; ++i) r[i] = (a[3 * i] * v[0] + a[3 * i + 1] * v[1]) + a[3 * i + 2] * v[2];
  o[0] = r[0]; o[1] = r[1]; o[2] = r[2];
}
NM_HD void hat3(const double* w, double* W) {
  W[0] = 0; W[1] = -w[2]; W[2] = w[1];
  W[3] = w[2]; W[4] = 0; W[5] = -w[0];
  W[6] = -w[1]; W[7] = w[0]; W[8] = 0;
}

NM_HD void so3_exp(const double* w, double* R) {
  const double th2 = (w[0] * w[0] + w[1] * w[1]) + w[2] * w[2];
  double W[9];
  hat3(w, W);
  if (th2 <= DBL_EPSILON) {
    double WW[9];
    d33_mul(W, W, WW);
    for (int i = 0; i < 9; ++i) R[i] = ((i % 4 == 0) ? 1.0 : 0.0) + W[i] + 0.5 * WW[i];
    return;
  }
  const double th = sqrt(th2), s = sin(th), s2 = sin(th / 2.0);
  const double omc = 2.0 * s2 * s2;
  double K[9], KK[9];
  for (int i = 0; i < 9; ++i) K[i] = W[i] / th;
  d33_mul(K, K, KK);
  for (int i = 0; i < 9; ++i) R[i] = ((i % 4 == 0) ? 1.0 : 0.0) + s * K[i] + omc * KK[i];
}

NM_HD void so3_log(const double* R, double* w) {
  const double tr = (R[0] + R[4]) + R[8];
  if (tr + 1.0 < 1e-3) {
    double v[3], r;
    if (R[8] > R[4] && R[8] > R[0]) { r = sqrt(2.0 * (1.0 + R[8])); v[0] = R[2] / r; v[1] = R[5] / r; v[2] = (1.0 + R[8]) / r; }
    else if (R[4] > R[0]) { r = sqrt(2.0 * (1.0 + R[4])); v[0] = R[1] / r; v[1] = (1.0 + R[4]) / r; v[2] = R[7] / r; }
    else { r = sqrt(2.0 * (1.0 + R[0])); v[0] = (1.0 + R[0]) / r; v[1] = R[3] / r; v[2] = R[6] / r; }
    const double n = sqrt((v[0] * v[0] + v[1] * v[1]) + v[2] * v[2]);
    for (int i = 0; i < 3; ++i) w[i] = M_PI * v[i] / n;
    return;
  }
  double mag;
  const double tr3 = tr - 3.0;
  if (tr3 < -1e-6) { const double th = acos((tr - 1.0) / 2.0); mag = th / (2.0 * sin(th)); }
  else mag = 0.5 - tr3 / 12.0 + tr3 * tr3 / 60.0;
  w[0] = mag * (R[7] - R[5]); w[1] = mag * (R[2] - R[6]); w[2] = mag * (R[3] - R[1]);
}

NM_HD Pose pose_expmap(const double* xi) {
  Pose p;
  const double* w = xi; const double* v = xi + 3;
  so3_exp(w, p.R);
  const double th2 = (w[0] * w[0] + w[1] * w[1]) + w[2] * w[2];
  if (th2 > DBL_EPSILON) {
    const double wv = (w[0] * v[0] + w[1] * v[1]) + w[2] * v[2];
    const double c[3] = {w[1] * v[2] - w[2] * v[1], w[2] * v[0] - w[0] * v[2], w[0] * v[1] - w[1] * v[0]};
    double Rc[3];
    d33_vec(p.R, c, Rc);
    for (int i = 0; i < 3; ++i) p.t[i] = (c[i] - Rc[i] + w[i] * wv) / th2;
  } else {
    p.t[0] = v[0]; p.t[1] = v[1]; p.t[2] = v[2];
  }
  return p;
}

NM_HD void pose_logmap(const Pose& p, double* xi) {
  double w[3];
  so3_log(p.R, w);
  const double t = sqrt((w[0] * w[0] + w[1] * w[1]) + w[2] * w[2]);
  if (t < 1e-10) { for (int i = 0; i < 3; ++i) { xi[i] = w[i]; xi[3 + i] = p.t[i]; } return; }
  const double wn[3] = {w[0] / t, w[1] / t, w[2] / t};
  double W[9], WT[3], WWT[3];
  hat3(wn, W);
  const double Tan = tan(0.5 * t);
  d33_vec(W, p.t, WT);
  d33_vec(W, WT, WWT);
  for (int i = 0; i < 3; ++i) { xi[i] = w[i]; xi[3 + i] = p.t[i] - (0.5 * t) * WT[i] + (1.0 - t / (2.0 * Tan)) * WWT[i]; }
}

NM_HD Pose pose_compose(const Pose& a, const Pose& b) {
  Pose o;
  d33_mul(a.R, b.R, o.R);
  double Rt[3];
  d33_vec(a.R, b.t, Rt);
  for (int i = 0; i < 3; ++i) o.t[i] = a.t[i] + Rt[i];
  return o;
}
NM_HD Pose pose_between(const Pose& a, const Pose& b) {  // a^-1 * b
  double At[9];
  for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) At[3 * i + j] = a.R[3 * j + i];
  Pose o;
  d33_mul(At, b.R, o.R);
  const double d[3] = {b.t[0] - a.t[0], b.t[1] - a.t[1], b.t[2] - a.t[2]};
  d33_vec(At, d, o.t);
  return o;
}
NM_HD Pose pose_retract(const Pose& p, const double* xi) { return pose_compose(p, pose_expmap(xi)); }

// gtsam::Rot3::rpy() (RQ decomposition)
NM_HD void rot_rpy(const double* A, double* rpy) {
  const double x = -atan2(-A[7], A[8]);
  const double cx = cos(-x), sx = sin(-x);
  const double Qx[9] = {1, 0, 0, 0, cx, -sx, 0, sx, cx};
  double B[9];
  d33_mul(A, Qx, B);
  const double y = -atan2(B[6], B[8]);
  const double cy = cos(-y), sy = sin(-y);
  const double Qy[9] = {cy, 0, sy, 0, 1, 0, -sy, 0, cy};
  double C[9];
  d33_mul(B, Qy, C);
  const double z = -atan2(-C[3], C[4]);
  rpy[0] = x; rpy[1] = y; rpy[2] = z;
}

NM_HD void pose_from16(const double* M, Pose& p) {
  for (int r = 0; r < 3; ++r) { p.R[3 * r] = M[4 * r]; p.R[3 * r + 1] = M[4 * r + 1]; p.R[3 * r + 2] = M[4 * r + 2]; p.t[r] = M[4 * r + 3]; }
}
NM_HD void pose_to16(const Pose& p, double* M) {
  for (int r = 0; r < 3; ++r) { M[4 * r] = p.R[3 * r]; M[4 * r + 1] = p.R[3 * r + 1]; M[4 * r + 2] = p.R[3 * r + 2]; M[4 * r + 3] = p.t[r]; }
  M[12] = 0; M[13] = 0; M[14] = 0; M[15] = 1;
}

// Diagonal-pivoted LDL^T solve of a symmetric 6x6 (Eigen::LDLT), false = NumericalIssue.
NM_HD bool ldlt6_solve(const double* A, const double* b, double* x) {
  double m[36];
  for (int i = 0; i < 36; ++i) m[i] = A[i];
  int tr[6];
  bool ok = true;
  for (int k = 0; k < 6; ++k) {
    int piv = k;
    double best = fabs(m[7 * k]);
    for (int i = k + 1; i < 6; ++i) if (fabs(m[7 * i]) > best) { best = fabs(m[7 * i]); piv = i; }
    tr[k] = piv;
    if (piv != k) {
      double t;
      for (int j = 0; j < k; ++j) { t = m[6 * k + j]; m[6 * k + j] = m[6 * piv + j]; m[6 * piv + j] = t; }
      for (int i = piv + 1; i < 6; ++i) { t = m[6 * i + k]; m[6 * i + k] = m[6 * i + piv]; m[6 * i + piv] = t; }
      t = m[7 * k]; m[7 * k] = m[7 * piv]; m[7 * piv] = t;
      for (int i = k + 1; i < piv; ++i) { t = m[6 * i + k]; m[6 * i + k] = m[6 * piv + i]; m[6 * piv + i] = t; }
    }
    double temp[6];
    for (int j = 0; j < k; ++j) temp[j] = m[7 * j] * m[6 * k + j];
    double akk = m[7 * k];
    if (k > 0) { double s = 0.0; for (int j = 0; j < k; ++j) s = s + m[6 * k + j] * temp[j]; akk = akk - s; }
    m[7 * k] = akk;
    if (k > 0)
      for (int i = k + 1; i < 6; ++i) { double s = 0.0; for (int j = 0; j < k; ++j) s = s + m[6 * i + j] * temp[j]; m[6 * i + k] = m[6 * i + k] - s; }
    const bool valid = fabs(akk) > 0.0;
    if (k == 0 && !valid) { for (int j = 0; j < 6; ++j) tr[j] = j; for (int i = 1; i < 6; ++i) if (m[6 * i] != 0.0) ok = false; break; }
    for (int i = k + 1; i < 6; ++i) { if (valid) m[6 * i + k] = m[6 * i + k] / akk; else if (m[6 * i + k] != 0.0) ok = false; }
  }
  double y[6];
  for (int i = 0; i < 6; ++i) y[i] = b[i];
  for (int k = 0; k < 6; ++k) { const double t = y[k]; y[k] = y[tr[k]]; y[tr[k]] = t; }
  for (int i = 0; i < 6; ++i) { double s = y[i]; for (int j = 0; j < i; ++j) s = s - m[6 * i + j] * y[j]; y[i] = s; }
  const double tol = 1.0 / DBL_MAX;
  for (int i = 0; i < 6; ++i) { if (fabs(m[7 * i]) > tol) y[i] = y[i] / m[7 * i]; else y[i] = 0.0; }
  for (int i = 5; i >= 0; --i) { double s = y[i]; for (int j = i + 1; j < 6; ++j) s = s - m[6 * j + i] * y[j]; y[i] = s; }
  for (int k = 5; k >= 0; --k) { const double t = y[k]; y[k] = y[tr[k]]; y[tr[k]] = t; }
  for (int i = 0; i < 6; ++i) x[i] = y[i];
  return ok;
}

// Host-only: floor the eigenvalues of a symmetric 6x6 at `floor_v` and recompose if any was raised
// (svn_ndt_impl.hpp:932-945), cyclic Jacobi.
inline void floor_eigenvalues6(double* cov, double floor_v) {
  double a[36], v[36];
  for (int i = 0; i < 6; ++i) for (int j = 0; j < 6; ++j) { a[6 * i + j] = (i >= j) ? cov[6 * i + j] : cov[6 * j + i]; v[6 * i + j] = (i == j) ? 1.0 : 0.0; }
  for (int sweep = 0; sweep < 100; ++sweep) {
    double off = 0.0;
    for (int i = 0; i < 6; ++i) for (int j = i + 1; j < 6; ++j) off += fabs(a[6 * i + j]);
    if (off == 0.0) break;
    for (int p = 0; p < 5; ++p)
      for (int q = p + 1; q < 6; ++q) {
        const double apq = a[6 * p + q];
        if (apq == 0.0) continue;
        const double g = 100.0 * fabs(apq);
        if (sweep > 3 && fabs(a[7 * p]) + g == fabs(a[7 * p]) && fabs(a[7 * q]) + g == fabs(a[7 * q])) { a[6 * p + q] = a[6 * q + p] = 0.0; continue; }
        const double hd = a[7 * q] - a[7 * p];
        double t;
        if (fabs(hd) + g == fabs(hd)) t = apq / hd;
        else { const double th = 0.5 * hd / apq; t = 1.0 / (fabs(th) + sqrt(1.0 + th * th)); if (th < 0.0) t = -t; }
        const double c = 1.0 / sqrt(1.0 + t * t), s = t * c, tau = s / (1.0 + c), hh = t * apq;
        a[7 * p] -= hh; a[7 * q] += hh; a[6 * p + q] = a[6 * q + p] = 0.0;
        for (int r = 0; r < 6; ++r) {
          if (r != p && r != q) {
            const double arp = a[6 * r + p], arq = a[6 * r + q];
            const double nrp = arp - s * (arq + arp * tau), nrq = arq + s * (arp - arq * tau);
            a[6 * r + p] = a[6 * p + r] = nrp; a[6 * r + q] = a[6 * q + r] = nrq;
          }
          const double vrp = v[6 * r + p], vrq = v[6 * r + q];
          v[6 * r + p] = vrp - s * (vrq + vrp * tau); v[6 * r + q] = vrq + s * (vrp - vrq * tau);
        }
      }
  }
  bool need = false;
  double ev[6];
  for (int i = 0; i < 6; ++i) { ev[i] = a[7 * i]; if (ev[i] < floor_v) { ev[i] = floor_v; need = true; } }
  if (!need) return;
  for (int i = 0; i < 6; ++i)
    for (int j = 0; j < 6; ++j) { double s = 0; for (int k = 0; k < 6; ++k) s += (v[6 * i + k] * ev[k]) * v[6 * j + k]; cov[6 * i + j] = s; }
}

}  // namespace ndtb
