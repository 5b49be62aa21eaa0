// ORACLE — TEST INFRASTRUCTURE ONLY.  Not part of the product path.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// build, load or call anything under oracle/.
//
// Small dense linear algebra used by the CPU restatement of the reference's NDT path.
// The reference gets these from Eigen3 (un-vendored, unpinned: /root/reference/CMakeLists.txt:22;
// PCL >= 1.14 implies Eigen 3.4.0 on the reference's platform).  Eigen is absent here, so each
// routine restates the published algorithm and PINS an explicit evaluation order:
//   * all products are "coefficient = sum_k a(i,k)*b(k,j)", k ascending, left-to-right adds,
//     no FMA contraction (build with -ffp-contract=off) -- Eigen's scalar (non-vectorised) path;
//   * 3x3 inverse = cofactor/determinant form (Eigen compute_inverse_size3 helper);
//   * 3x3 self-adjoint eigen solver = cyclic Jacobi (Rutishauser), lower triangle read, ascending
//     eigenvalues -- Eigen uses tridiagonal QL for `compute()`; both are backward stable, the
//     difference is O(eps) and only matters when eigenvalues are inflated (PARITY UNPINNED there);
//   * 6x6 SVD solve = two-sided Jacobi (what Eigen::JacobiSVD does for square real matrices);
//   * 6x6 LDLT = diagonal-pivoted Bunch-Kaufman-free LDL^T as in Eigen::LDLT.
#pragma once
#include <cmath>
#include <cstring>
#include <algorithm>
#include <limits>

namespace oracle {

// ---------------------------------------------------------------- 3x3 (row-major double[9])
inline void mat3_mul(const double* a, const double* b, double* out) {
  double r[9];
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) {
      double s = a[3 * i + 0] * b[0 + j];
      s = s + a[3 * i + 1] * b[3 + j];
      s = s + a[3 * i + 2] * b[6 + j];
      r[3 * i + j] = s;
    }
  std::memcpy(out, r, sizeof(r));
}

// Cofactor inverse; returns false if det == 0 (result then holds inf/nan like Eigen would).
inline void mat3_inverse(const double* m, double* inv) {
  const double c00 = m[4] * m[8] - m[5] * m[7];
  const double c10 = m[5] * m[6] - m[3] * m[8];
  const double c20 = m[3] * m[7] - m[4] * m[6];
  const double det = (m[0] * c00 + m[1] * c10) + m[2] * c20;
  const double invdet = 1.0 / det;
  double r[9];
  r[0] = c00 * invdet;
  r[3] = c10 * invdet;
  r[6] = c20 * invdet;
  r[1] = (m[2] * m[7] - m[1] * m[8]) * invdet;
  r[4] = (m[0] * m[8] - m[2] * m[6]) * invdet;
  r[7] = (m[1] * m[6] - m[0] * m[7]) * invdet;
  r[2] = (m[1] * m[5] - m[2] * m[4]) * invdet;
  r[5] = (m[2] * m[3] - m[0] * m[5]) * invdet;
  r[8] = (m[0] * m[4] - m[1] * m[3]) * invdet;
  std::memcpy(inv, r, sizeof(r));
}

// Cyclic Jacobi for a symmetric 3x3 (reads the LOWER triangle, like SelfAdjointEigenSolver).
// evals ascending; evecs row-major with eigenvectors in COLUMNS.
inline void eig3_sym(const double* A, double* evals, double* evecs) {
  double a[3][3];
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) a[i][j] = (i >= j) ? A[3 * i + j] : A[3 * j + i];
  double v[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
  for (int sweep = 0; sweep < 60; ++sweep) {
    const double off = (std::fabs(a[0][1]) + std::fabs(a[0][2])) + std::fabs(a[1][2]);
    if (off == 0.0) break;
    for (int p = 0; p < 2; ++p)
      for (int q = p + 1; q < 3; ++q) {
        const double apq = a[p][q];
        if (apq == 0.0) continue;
        const double g = 100.0 * std::fabs(apq);
        if (sweep > 3 && std::fabs(a[p][p]) + g == std::fabs(a[p][p]) &&
            std::fabs(a[q][q]) + g == std::fabs(a[q][q])) {
          a[p][q] = 0.0;
          a[q][p] = 0.0;
          continue;
        }
        const double h = a[q][q] - a[p][p];
        double t;
        if (std::fabs(h) + g == std::fabs(h)) {
          t = apq / h;
        } else {
          const double theta = 0.5 * h / apq;
          t = 1.0 / (std::fabs(theta) + std::sqrt(1.0 + theta * theta));
          if (theta < 0.0) t = -t;
        }
        const double c = 1.0 / std::sqrt(1.0 + t * t);
        const double s = t * c;
        const double tau = s / (1.0 + c);
        const double hh = t * apq;
        a[p][p] = a[p][p] - hh;
        a[q][q] = a[q][q] + hh;
        a[p][q] = 0.0;
        a[q][p] = 0.0;
        const int r = 3 - p - q;  // the remaining index
        {
          const double arp = a[r][p], arq = a[r][q];
          const double nrp = arp - s * (arq + arp * tau);
          const double nrq = arq + s * (arp - arq * tau);
          a[r][p] = nrp; a[p][r] = nrp;
          a[r][q] = nrq; a[q][r] = nrq;
        }
        for (int k = 0; k < 3; ++k) {
          const double vkp = v[k][p], vkq = v[k][q];
          v[k][p] = vkp - s * (vkq + vkp * tau);
          v[k][q] = vkq + s * (vkp - vkq * tau);
        }
      }
  }
  int idx[3] = {0, 1, 2};
  double d[3] = {a[0][0], a[1][1], a[2][2]};
  // stable 3-element insertion sort, ascending
  for (int i = 1; i < 3; ++i)
    for (int j = i; j > 0 && d[idx[j]] < d[idx[j - 1]]; --j) std::swap(idx[j], idx[j - 1]);
  for (int c = 0; c < 3; ++c) {
    evals[c] = d[idx[c]];
    for (int r = 0; r < 3; ++r) evecs[3 * r + c] = v[r][idx[c]];
  }
}

// ---------------------------------------------------------------- 6x6 (row-major double[36])
struct JacobiRot { double c, s; };

// 2x2 symmetric Jacobi (Eigen JacobiRotation::makeJacobi(x, y, z), real case).
inline JacobiRot make_jacobi(double x, double y, double z) {
  JacobiRot j;
  const double deno = 2.0 * std::fabs(y);
  if (deno < std::numeric_limits<double>::min()) { j.c = 1.0; j.s = 0.0; return j; }
  const double tau = (x - z) / deno;
  const double w = std::sqrt(tau * tau + 1.0);
  const double t = (tau > 0.0) ? 1.0 / (tau + w) : 1.0 / (tau - w);
  const double sign_t = t > 0.0 ? 1.0 : -1.0;
  const double n = 1.0 / std::sqrt(t * t + 1.0);
  j.s = -sign_t * (y / std::fabs(y)) * std::fabs(t) * n;
  j.c = n;
  return j;
}

// rows p,q of a 6x6: x' = c x + s y ; y' = -s x + c y
inline void rot_rows(double* m, int p, int q, JacobiRot j) {
  for (int i = 0; i < 6; ++i) {
    const double x = m[6 * p + i], y = m[6 * q + i];
    m[6 * p + i] = j.c * x + j.s * y;
    m[6 * q + i] = -j.s * x + j.c * y;
  }
}
// columns p,q with the TRANSPOSED rotation (Eigen applyOnTheRight): x' = c x - s y ; y' = s x + c y
inline void rot_cols(double* m, int p, int q, JacobiRot j) {
  for (int i = 0; i < 6; ++i) {
    const double x = m[6 * i + p], y = m[6 * i + q];
    m[6 * i + p] = j.c * x - j.s * y;
    m[6 * i + q] = j.s * x + j.c * y;
  }
}

// Two-sided Jacobi SVD of a real 6x6: A = U diag(S) V^T, S descending, then x = V S^+ U^T b with
// Eigen's default rank threshold (eps * 6 relative to S[0]).  Follows ndt_omp_impl.hpp:142-145.
inline void svd6_solve(const double* A, const double* b, double* x, double* S_out = nullptr) {
  double W[36], U[36], V[36];
  double scale = 0.0;
  for (int i = 0; i < 36; ++i) scale = std::max(scale, std::fabs(A[i]));
  if (!(scale > 0.0) || !std::isfinite(scale)) scale = (scale == 0.0) ? 1.0 : scale;
  for (int i = 0; i < 36; ++i) { W[i] = A[i] / scale; U[i] = 0.0; V[i] = 0.0; }
  for (int i = 0; i < 6; ++i) { U[7 * i] = 1.0; V[7 * i] = 1.0; }
  const double precision = 2.0 * std::numeric_limits<double>::epsilon();
  const double consider_zero = std::numeric_limits<double>::min();
  double max_diag = 0.0;
  for (int i = 0; i < 6; ++i) max_diag = std::max(max_diag, std::fabs(W[7 * i]));
  bool finished = false;
  int guard = 0;
  while (!finished && guard++ < 200) {
    finished = true;
    for (int p = 1; p < 6; ++p)
      for (int q = 0; q < p; ++q) {
        const double threshold = std::max(consider_zero, precision * max_diag);
        if (std::fabs(W[6 * p + q]) > threshold || std::fabs(W[6 * q + p]) > threshold) {
          finished = false;
          // real_2x2_jacobi_svd on [[W(p,p), W(p,q)], [W(q,p), W(q,q)]]
          double m00 = W[7 * p], m01 = W[6 * p + q], m10 = W[6 * q + p], m11 = W[7 * q];
          JacobiRot rot1;
          const double t = m00 + m11, d = m10 - m01;
          if (std::fabs(d) < consider_zero) { rot1.c = 1.0; rot1.s = 0.0; }
          else { const double u = t / d; const double tmp = std::sqrt(1.0 + u * u); rot1.s = 1.0 / tmp; rot1.c = u / tmp; }
          // m.applyOnTheLeft(0,1,rot1)
          const double n00 = rot1.c * m00 + rot1.s * m10, n01 = rot1.c * m01 + rot1.s * m11;
          const double n11 = -rot1.s * m01 + rot1.c * m11;
          const JacobiRot jr = make_jacobi(n00, n01, n11);
          // j_left = rot1 * jr^T ; jr^T = (c, -s)
          JacobiRot jl;
          jl.c = rot1.c * jr.c - rot1.s * (-jr.s);
          jl.s = rot1.c * (-jr.s) + rot1.s * jr.c;
          rot_rows(W, p, q, jl);
          // U.applyOnTheRight(p,q,jl^T): rot_cols applies the transpose of its argument, and
          // (jl^T)^T = jl, so pass jl^T = (c,-s).
          rot_cols(U, p, q, JacobiRot{jl.c, -jl.s});
          rot_cols(W, p, q, jr);
          rot_cols(V, p, q, jr);
          max_diag = std::max(max_diag, std::max(std::fabs(W[7 * p]), std::fabs(W[7 * q])));
        }
      }
  }
  double S[6];
  for (int i = 0; i < 6; ++i) {
    const double a = W[7 * i];
    S[i] = std::fabs(a);
    if (a < 0.0) for (int r = 0; r < 6; ++r) U[6 * r + i] = -U[6 * r + i];
  }
  for (int i = 0; i < 6; ++i) S[i] *= scale;
  for (int i = 0; i < 6; ++i) {  // selection sort, descending
    int pos = i;
    for (int k = i + 1; k < 6; ++k) if (S[k] > S[pos]) pos = k;
    if (S[pos] == 0.0) break;
    if (pos != i) {
      std::swap(S[i], S[pos]);
      for (int r = 0; r < 6; ++r) { std::swap(U[6 * r + i], U[6 * r + pos]); std::swap(V[6 * r + i], V[6 * r + pos]); }
    }
  }
  if (S_out) std::memcpy(S_out, S, sizeof(S));
  const double thr = std::max(S[0] * (std::numeric_limits<double>::epsilon() * 6.0), std::numeric_limits<double>::min());
  int rank = 6;
  while (rank > 0 && !(S[rank - 1] >= thr)) --rank;  // NaN-safe: NaN counts as below threshold
  if (S[0] != S[0]) rank = 6;                        // all-NaN input propagates NaN like Eigen
  double tmp[6];
  for (int i = 0; i < rank; ++i) {
    double s = 0.0;
    for (int r = 0; r < 6; ++r) s = s + U[6 * r + i] * b[r];
    tmp[i] = (1.0 / S[i]) * s;
  }
  for (int r = 0; r < 6; ++r) {
    double s = 0.0;
    for (int i = 0; i < rank; ++i) s = s + V[6 * r + i] * tmp[i];
    x[r] = s;
  }
}

// Diagonal-pivoted LDL^T solve (Eigen::LDLT<Matrix6d>, svn_ndt_impl.hpp:833-838).
// Returns false when Eigen would report NumericalIssue.
inline bool ldlt6_solve(const double* A, const double* b, double* x) {
  const int n = 6;
  double m[36];
  std::memcpy(m, A, sizeof(m));
  int tr[6];
  bool ok = true;
  for (int k = 0; k < n; ++k) {
    int piv = k; double best = std::fabs(m[7 * k]);
    for (int i = k + 1; i < n; ++i) if (std::fabs(m[7 * i]) > best) { best = std::fabs(m[7 * i]); piv = i; }
    tr[k] = piv;
    if (piv != k) {  // symmetric swap on the lower triangle
      for (int j = 0; j < k; ++j) std::swap(m[6 * k + j], m[6 * piv + j]);
      for (int i = piv + 1; i < n; ++i) std::swap(m[6 * i + k], m[6 * i + piv]);
      std::swap(m[7 * k], m[7 * piv]);
      for (int i = k + 1; i < piv; ++i) std::swap(m[6 * i + k], m[6 * piv + i]);
    }
    double temp[6];
    for (int j = 0; j < k; ++j) temp[j] = m[7 * j] * m[6 * k + j];
    double akk = m[7 * k];
    if (k > 0) { double s = 0.0; for (int j = 0; j < k; ++j) s = s + m[6 * k + j] * temp[j]; akk = akk - s; }
    m[7 * k] = akk;
    for (int i = k + 1; i < n; ++i) {
      if (k > 0) { double s = 0.0; for (int j = 0; j < k; ++j) s = s + m[6 * i + j] * temp[j]; m[6 * i + k] = m[6 * i + k] - s; }
    }
    const bool valid = std::fabs(akk) > 0.0;
    if (k == 0 && !valid) { for (int j = 0; j < n; ++j) tr[j] = j; for (int i = 1; i < n; ++i) if (m[6 * i] != 0.0) ok = false; break; }
    for (int i = k + 1; i < n; ++i) { if (valid) m[6 * i + k] = m[6 * i + k] / akk; else if (m[6 * i + k] != 0.0) ok = false; }
  }
  double y[6];
  std::memcpy(y, b, sizeof(y));
  for (int k = 0; k < n; ++k) std::swap(y[k], y[tr[k]]);
  for (int i = 0; i < n; ++i) { double s = y[i]; for (int j = 0; j < i; ++j) s = s - m[6 * i + j] * y[j]; y[i] = s; }
  const double tol = 1.0 / std::numeric_limits<double>::max();
  for (int i = 0; i < n; ++i) { if (std::fabs(m[7 * i]) > tol) y[i] = y[i] / m[7 * i]; else y[i] = 0.0; }
  for (int i = n - 1; i >= 0; --i) { double s = y[i]; for (int j = i + 1; j < n; ++j) s = s - m[6 * j + i] * y[j]; y[i] = s; }
  for (int k = n - 1; k >= 0; --k) std::swap(y[k], y[tr[k]]);
  std::memcpy(x, y, sizeof(y));
  return ok;
}

// Symmetric 6x6 Jacobi eigen-decomposition (ascending) for the SVN final-covariance floor
// (svn_ndt_impl.hpp:932-945).  evecs row-major, eigenvectors in columns.
inline void eig6_sym(const double* A, double* evals, double* evecs) {
  double a[36], v[36];
  for (int i = 0; i < 6; ++i) for (int j = 0; j < 6; ++j) { a[6 * i + j] = (i >= j) ? A[6 * i + j] : A[6 * j + i]; v[6 * i + j] = (i == j); }
  for (int sweep = 0; sweep < 100; ++sweep) {
    double off = 0.0;
    for (int i = 0; i < 6; ++i) for (int j = i + 1; j < 6; ++j) off += std::fabs(a[6 * i + j]);
    if (off == 0.0) break;
    for (int p = 0; p < 5; ++p) for (int q = p + 1; q < 6; ++q) {
      const double apq = a[6 * p + q];
      if (apq == 0.0) continue;
      const double g = 100.0 * std::fabs(apq);
      if (sweep > 3 && std::fabs(a[7 * p]) + g == std::fabs(a[7 * p]) && std::fabs(a[7 * q]) + g == std::fabs(a[7 * q])) { a[6 * p + q] = a[6 * q + p] = 0.0; continue; }
      const double h = a[7 * q] - a[7 * p];
      double t;
      if (std::fabs(h) + g == std::fabs(h)) t = apq / h;
      else { const double theta = 0.5 * h / apq; t = 1.0 / (std::fabs(theta) + std::sqrt(1.0 + theta * theta)); if (theta < 0.0) t = -t; }
      const double c = 1.0 / std::sqrt(1.0 + t * t), s = t * c, tau = s / (1.0 + c), hh = t * apq;
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
  int idx[6] = {0, 1, 2, 3, 4, 5};
  for (int i = 1; i < 6; ++i) for (int j = i; j > 0 && a[7 * idx[j]] < a[7 * idx[j - 1]]; --j) std::swap(idx[j], idx[j - 1]);
  for (int c = 0; c < 6; ++c) { evals[c] = a[7 * idx[c]]; for (int r = 0; r < 6; ++r) evecs[6 * r + c] = v[6 * r + idx[c]]; }
}

}  // namespace oracle
