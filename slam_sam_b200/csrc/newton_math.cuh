// Small-matrix maths of the optimiser loops, usable from host code and from device epilogues.
//   pose <-> float 4x4      ndt_omp_impl.hpp:117-120, 167-174, 969-975; ndt_omp.h:205-211
//   gauss constants         ndt_omp_impl.hpp:95-99 ; svn_ndt_impl.hpp:80-131
//   angle tables            ndt_omp_impl.hpp:396-538 ; svn_ndt_impl.hpp:255-334
//   Newton direction        Eigen::JacobiSVD<6x6>::solve (ndt_omp_impl.hpp:142-145), restated as a
//                           two-sided Jacobi SVD with Eigen's rank threshold
//   Stein update solve      Eigen::LDLT<6x6> (svn_ndt_impl.hpp:833-838)
// Everything is FP64 (or float where the reference is float) with explicit evaluation order; the
// translation unit is built with -fmad=false / -ffp-contract=off.
#pragma once
#include <math.h>
#include <float.h>
#include <string.h>
#include <cuda_runtime.h>

#define NM_HD __host__ __device__ inline

namespace ndtb {

#ifdef __CUDA_ARCH__
// libm's sinf/cosf/atan2f are (nearly) correctly rounded; CUDA's float versions are not.  Going
// through FP64 and rounding once reproduces the host bits except for ~1e-8 of arguments.
NM_HD float nm_sinf(float x) { return static_cast<float>(sin(static_cast<double>(x))); }
NM_HD float nm_cosf(float x) { return static_cast<float>(cos(static_cast<double>(x))); }
NM_HD float nm_atan2f(float y, float x) { return static_cast<float>(atan2(static_cast<double>(y), static_cast<double>(x))); }
#else
NM_HD float nm_sinf(float x) { return sinf(x); }
NM_HD float nm_cosf(float x) { return cosf(x); }
NM_HD float nm_atan2f(float y, float x) { return atan2f(y, x); }
#endif

// ---- sincosf as the host's libm computes it ---------------------------------------------------------------------
// pose_to_T (below) is Eigen's AngleAxis<float>: std::sin / std::cos of a float, which gcc turns into ONE sincosf call.  glibc's
// sincosf (sysdeps/ieee754/flt-32/s_sincosf.c, 2.28 and later; the x86-64 build selected on CPUs with FMA + AVX2) is a pair of
// double-precision polynomials -- within 0.56 ulp, i.e. NOT the correctly rounded value for ~1 % of the arguments -- so a device
// step that must land on the host's bits restates it operation by operation: the same products, the same fused multiply-adds
// (IEEE operations, identical on both machines), the same two coefficient tables, the same quadrant logic.  Arguments of
// magnitude >= 120 take glibc's large-argument reduction, which is not restated: the function then returns false.
// nm_sincosf_port_matches_libm() checks the restatement against the running libm; the device Newton step is only used when
// it holds (another libm build, e.g. the non-FMA one, would simply switch it off).
NM_HD double nm_fma(double a, double b, double c) {
#ifdef __CUDA_ARCH__
  return __fma_rn(a, b, c);
#else
  return fma(a, b, c);
#endif
}
NM_HD bool nm_sincosf_glibc(float xf, float* sp, float* cp) {
  // c0, c1, c2, c3, c4, s1, s2, s3 of the two quadrant tables (the second one holds -cos)
  const double C[2][5] = {{0x1p0, -0x1.ffffffd0c621cp-2, 0x1.55553e1068f19p-5, -0x1.6c087e89a359dp-10, 0x1.99343027bf8c3p-16},
                          {-0x1p0, 0x1.ffffffd0c621cp-2, -0x1.55553e1068f19p-5, 0x1.6c087e89a359dp-10, -0x1.99343027bf8c3p-16}};
  const double S1 = -0x1.555545995a603p-3, S2 = 0x1.1107605230bc4p-7, S3 = -0x1.994eb3774cf24p-13;
  const double HPI_INV = 0x1.45f306dc9c883p+23 /* 2/pi * 2^24 */, HPI = 0x1.921fb54442d18p+0;
  unsigned int bits;
#ifdef __CUDA_ARCH__
  bits = __float_as_uint(xf);
#else
  memcpy(&bits, &xf, 4);
#endif
  const unsigned int top = (bits >> 20) & 0x7ffu;
  const double x = static_cast<double>(xf);
  double xs, x2;
  int n = 0;
  if (top <= 0x3f3u) {            // |x| < 0.75
    if (top <= 0x397u) { *sp = xf; *cp = 1.0f; return true; }   // |x| < 2^-12
    xs = x; x2 = x * x;
  } else if (top <= 0x42eu) {     // |x| < 120
    const double r = x * HPI_INV;
    n = (static_cast<int>(r) + 0x800000) >> 24;
    const double xr = nm_fma(-static_cast<double>(n), HPI, x);
    const double sign = ((n & 3) == 1 || (n & 3) == 2) ? -1.0 : 1.0;
    xs = xr * sign; x2 = xr * xr;
  } else {
    return false;
  }
  const double* c = C[(n >> 1) & 1];
  const double x3 = x2 * xs, x4 = x2 * x2;
  const double sa = nm_fma(x2, S3, S2), ca = nm_fma(x2, c[4], c[3]);
  const double x5 = x2 * x3, x6 = x2 * x4;
  const double c01 = nm_fma(x2, c[1], c[0]);
  const double sb = nm_fma(x3, S1, xs), cb = nm_fma(x4, c[2], c01);
  const float sv = static_cast<float>(nm_fma(sa, x5, sb)), cv = static_cast<float>(nm_fma(ca, x6, cb));
  if (n & 1) { *sp = cv; *cp = sv; } else { *sp = sv; *cp = cv; }
  return true;
}

// ---- gauss constants ------------------------------------------------------------------------
NM_HD void gauss_pclomp(double outlier_ratio, float resolution, double* d1, double* d2, double* d3) {
  const double c1 = 10 * (1 - outlier_ratio);
  const double c2 = outlier_ratio / pow(static_cast<double>(resolution), 3);
  *d3 = -log(c2);
  *d1 = -log(c1 + c2) - *d3;
  *d2 = -2 * log((-log(c1 * exp(-0.5) + c2) - *d3) / *d1);
}

NM_HD void gauss_svn(double outlier_ratio, float resolution, double* d1, double* d2, double* d3) {
  if (resolution <= 1e-6f) { *d1 = 1.0; *d2 = 1.0; *d3 = 0.0; return; }
  const double eps = 1e-9;
  double c1 = 10.0 * (1.0 - outlier_ratio);
  double c2 = outlier_ratio / pow(static_cast<double>(resolution), 3);
  if (c1 <= eps) c1 = eps;
  if (c2 <= eps) c2 = eps;
  *d3 = -log(c2);
  *d1 = -log(c1 + c2) - *d3;
  *d2 = 1.0;
  if (!(fabs(*d1) < eps)) {
    const double inner = c1 * exp(-0.5) + c2;
    if (inner > eps) {
      const double outer = (-log(inner) - *d3) / *d1;
      if (outer > eps) *d2 = -2.0 * log(outer);
    }
  }
  if (!isfinite(*d1) || !isfinite(*d2) || !isfinite(*d3)) { *d1 = 1.0; *d2 = 1.0; *d3 = 0.0; }
}

// ---- angle tables: jang[8][3], hang[15][3] as float ---------------------------------------------
NM_HD void angle_tables(const double* p, double snap, float* jang, float* hang) {
  double cx = 1.0, sx = 0.0, cy = 1.0, sy = 0.0, cz = 1.0, sz = 0.0;
  if (!(fabs(p[3]) < snap)) { cx = cos(p[3]); sx = sin(p[3]); }
  if (!(fabs(p[4]) < snap)) { cy = cos(p[4]); sy = sin(p[4]); }
  if (!(fabs(p[5]) < snap)) { cz = cos(p[5]); sz = sin(p[5]); }
  const double j[24] = {(-sx * sz + cx * sy * cz), (-sx * cz - cx * sy * sz), (-cx * cy),
                        (cx * sz + sx * sy * cz),  (cx * cz - sx * sy * sz),  (-sx * cy),
                        (-sy * cz),                sy * sz,                   cy,
                        sx * cy * cz,              (-sx * cy * sz),           sx * sy,
                        (-cx * cy * cz),           cx * cy * sz,              (-cx * sy),
                        (-cy * sz),                (-cy * cz),                0,
                        (cx * cz - sx * sy * sz),  (-cx * sz - sx * sy * cz), 0,
                        (sx * cz + cx * sy * sz),  (cx * sy * cz - sx * sz),  0};
  const double h[45] = {(-cx * sz - sx * sy * cz), (-cx * cz + sx * sy * sz), sx * cy,
                        (-sx * sz + cx * sy * cz), (-cx * sy * sz - sx * cz), (-cx * cy),
                        (cx * cy * cz),            (-cx * cy * sz),           (cx * sy),
                        (sx * cy * cz),            (-sx * cy * sz),           (sx * sy),
                        (-sx * cz - cx * sy * sz), (sx * sz - cx * sy * cz),  0,
                        (cx * cz - sx * sy * sz),  (-sx * sy * cz - cx * sz), 0,
                        (-cy * cz),                (cy * sz),                 (sy),
                        (-sx * sy * cz),           (sx * sy * sz),            (sx * cy),
                        (cx * sy * cz),            (-cx * sy * sz),           (-cx * cy),
                        (sy * sz),                 (sy * cz),                 0,
                        (-sx * cy * sz),           (-sx * cy * cz),           0,
                        (cx * cy * sz),            (cx * cy * cz),            0,
                        (-cy * cz),                (cy * sz),                 0,
                        (-cx * sz - sx * sy * cz), (-cx * cz + sx * sy * sz), 0,
                        (-sx * sz + cx * sy * cz), (-cx * sy * sz - sx * cz), 0};
  for (int i = 0; i < 24; ++i) jang[i] = static_cast<float>(j[i]);
  for (int i = 0; i < 45; ++i) hang[i] = static_cast<float>(h[i]);
}

// ---- float rotation helpers ----------------------------------------------------------------------
NM_HD void f33_mul(const float* a, const float* b, float* o) {
  float r[9];
  for (int i = 0; i < 3; ++i)
    for (int k = 0; k < 3; ++k) {
      float s = a[3 * i] * b[k];
      s = s + a[3 * i + 1] * b[3 + k];
      s = s + a[3 * i + 2] * b[6 + k];
      r[3 * i + k] = s;
    }
  for (int i = 0; i < 9; ++i) o[i] = r[i];
}

// Eigen::AngleAxis<float>(angle, unit axis).toRotationMatrix()
NM_HD void f_angle_axis(float angle, int axis, float* R) {
  float u[3] = {0.0f, 0.0f, 0.0f};
  u[axis] = 1.0f;
  const float s = nm_sinf(angle), c = nm_cosf(angle);
  const float su[3] = {s * u[0], s * u[1], s * u[2]};
  const float cu[3] = {(1.0f - c) * u[0], (1.0f - c) * u[1], (1.0f - c) * u[2]};
  float t = cu[0] * u[1];
  R[1] = t - su[2]; R[3] = t + su[2];
  t = cu[0] * u[2];
  R[2] = t + su[1]; R[6] = t - su[1];
  t = cu[1] * u[2];
  R[5] = t - su[0]; R[7] = t + su[0];
  R[0] = cu[0] * u[0] + c; R[4] = cu[1] * u[1] + c; R[8] = cu[2] * u[2] + c;
}

// T = Translation(p0,p1,p2) * Rx(p3) * Ry(p4) * Rz(p5), float, row-major 4x4
NM_HD void pose_to_T(const double* p, float* T) {
  float A[9], B[9], C[9], R[9];
  f_angle_axis(static_cast<float>(p[3]), 0, A);
  f_angle_axis(static_cast<float>(p[4]), 1, B);
  f_angle_axis(static_cast<float>(p[5]), 2, C);
  f33_mul(A, B, R);
  f33_mul(R, C, R);
  for (int r = 0; r < 3; ++r) {
    T[4 * r] = R[3 * r]; T[4 * r + 1] = R[3 * r + 1]; T[4 * r + 2] = R[3 * r + 2];
    T[4 * r + 3] = static_cast<float>(p[r]);
  }
  T[12] = 0.0f; T[13] = 0.0f; T[14] = 0.0f; T[15] = 1.0f;
}

// Eigen 3.4 eulerAngles(0,1,2) of the rotation block of a float 4x4 (first angle folded to [0,pi])
NM_HD void euler_012(const float* T, float* e) {
  const float r00 = T[0], r01 = T[1], r02 = T[2], r10 = T[4], r11 = T[5], r12 = T[6], r20 = T[8], r21 = T[9], r22 = T[10];
  float a0 = nm_atan2f(r12, r22);
  const float c2 = sqrtf(r00 * r00 + r01 * r01);
  float a1;
  if (a0 > 0.0f) { a0 -= static_cast<float>(M_PI); a1 = nm_atan2f(-r02, -c2); }
  else a1 = nm_atan2f(-r02, c2);
  const float s1 = nm_sinf(a0), c1 = nm_cosf(a0);
  const float a2 = nm_atan2f(s1 * r20 - c1 * r10, c1 * r11 - s1 * r21);
  e[0] = -a0; e[1] = -a1; e[2] = -a2;
}

// ---- 6x6 two-sided Jacobi SVD solve ---------------------------------------------------------------
struct Rot2 { double c, s; };

NM_HD Rot2 sym_jacobi(double x, double y, double z) {
  Rot2 j;
  const double deno = 2.0 * fabs(y);
  if (deno < DBL_MIN) { j.c = 1.0; j.s = 0.0; return j; }
  const double tau = (x - z) / deno;
  const double w = sqrt(tau * tau + 1.0);
  const double t = (tau > 0.0) ? 1.0 / (tau + w) : 1.0 / (tau - w);
  const double sign_t = t > 0.0 ? 1.0 : -1.0;
  const double n = 1.0 / sqrt(t * t + 1.0);
  j.s = -sign_t * (y / fabs(y)) * fabs(t) * n;
  j.c = n;
  return j;
}

NM_HD void rows_rot(double* m, int p, int q, double c, double s) {  // x' = c x + s y ; y' = -s x + c y
  for (int i = 0; i < 6; ++i) {
    const double x = m[6 * p + i], y = m[6 * q + i];
    m[6 * p + i] = c * x + s * y;
    m[6 * q + i] = -s * x + c * y;
  }
}
NM_HD void cols_rot(double* m, int p, int q, double c, double s) {  // x' = c x - s y ; y' = s x + c y
  for (int i = 0; i < 6; ++i) {
    const double x = m[6 * i + p], y = m[6 * i + q];
    m[6 * i + p] = c * x - s * y;
    m[6 * i + q] = s * x + c * y;
  }
}

NM_HD void svd6_solve(const double* A, const double* b, double* x) {
  double W[36], U[36], V[36];
  double scale = 0.0;
  for (int i = 0; i < 36; ++i) scale = fmax(scale, fabs(A[i]));
  if (scale == 0.0) scale = 1.0;
  for (int i = 0; i < 36; ++i) { W[i] = A[i] / scale; U[i] = 0.0; V[i] = 0.0; }
  for (int i = 0; i < 6; ++i) { U[7 * i] = 1.0; V[7 * i] = 1.0; }
  const double precision = 2.0 * DBL_EPSILON;
  double max_diag = 0.0;
  for (int i = 0; i < 6; ++i) max_diag = fmax(max_diag, fabs(W[7 * i]));
  bool finished = false;
  for (int guard = 0; !finished && guard < 200; ++guard) {
    finished = true;
    for (int p = 1; p < 6; ++p)
      for (int q = 0; q < p; ++q) {
        const double thr = fmax(DBL_MIN, precision * max_diag);
        if (fabs(W[6 * p + q]) > thr || fabs(W[6 * q + p]) > thr) {
          finished = false;
          const double m00 = W[7 * p], m01 = W[6 * p + q], m10 = W[6 * q + p], m11 = W[7 * q];
          double rc = 1.0, rs = 0.0;
          const double t = m00 + m11, d = m10 - m01;
          if (!(fabs(d) < DBL_MIN)) { const double u = t / d; const double tmp = sqrt(1.0 + u * u); rs = 1.0 / tmp; rc = u / tmp; }
          const double n00 = rc * m00 + rs * m10, n01 = rc * m01 + rs * m11, n11 = -rs * m01 + rc * m11;
          const Rot2 jr = sym_jacobi(n00, n01, n11);
          const double lc = rc * jr.c + rs * jr.s;
          const double ls = rs * jr.c - rc * jr.s;
          rows_rot(W, p, q, lc, ls);
          cols_rot(U, p, q, lc, -ls);
          cols_rot(W, p, q, jr.c, jr.s);
          cols_rot(V, p, q, jr.c, jr.s);
          max_diag = fmax(max_diag, fmax(fabs(W[7 * p]), fabs(W[7 * q])));
        }
      }
  }
  double S[6];
  for (int i = 0; i < 6; ++i) {
    const double a = W[7 * i];
    S[i] = fabs(a) * scale;
    if (a < 0.0) for (int r = 0; r < 6; ++r) U[6 * r + i] = -U[6 * r + i];
  }
  for (int i = 0; i < 6; ++i) {
    int pos = i;
    for (int k = i + 1; k < 6; ++k) if (S[k] > S[pos]) pos = k;
    if (S[pos] == 0.0) break;
    if (pos != i) {
      double t = S[i]; S[i] = S[pos]; S[pos] = t;
      for (int r = 0; r < 6; ++r) {
        t = U[6 * r + i]; U[6 * r + i] = U[6 * r + pos]; U[6 * r + pos] = t;
        t = V[6 * r + i]; V[6 * r + i] = V[6 * r + pos]; V[6 * r + pos] = t;
      }
    }
  }
  const double thr = fmax(S[0] * (DBL_EPSILON * 6.0), DBL_MIN);
  int rank = 6;
  while (rank > 0 && !(S[rank - 1] >= thr)) --rank;
  if (S[0] != S[0]) rank = 6;
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

// ---- More-Thuente line search pieces (ndt_omp.h:473-489, ndt_omp_impl.hpp:805-911) -------------------------
NM_HD double mt_psi(double a, double f_a, double f_0, double g_0, double mu) { return f_a - f_0 - mu * g_0 * a; }
NM_HD double mt_dpsi(double g_a, double g_0, double mu) { return g_a - mu * g_0; }

// Update Algorithm / Modified Update Algorithm: returns true when the interval has converged.
NM_HD bool mt_update_interval(double& a_l, double& f_l, double& g_l, double& a_u, double& f_u, double& g_u, double a_t, double f_t, double g_t) {
  if (f_t > f_l) { a_u = a_t; f_u = f_t; g_u = g_t; return false; }                                   // case U1 / a
  if (g_t * (a_l - a_t) > 0) { a_l = a_t; f_l = f_t; g_l = g_t; return false; }                       // case U2 / b
  if (g_t * (a_l - a_t) < 0) { a_u = a_l; f_u = f_l; g_u = g_l; a_l = a_t; f_l = f_t; g_l = g_t; return false; }  // case U3 / c
  return true;
}

// Trial value selection: cubic / quadratic / secant minimisers of the four cases.
NM_HD double mt_trial_value(double a_l, double f_l, double g_l, double a_u, double f_u, double g_u, double a_t, double f_t, double g_t) {
  if (f_t > f_l) {
    const double z = 3 * (f_t - f_l) / (a_t - a_l) - g_t - g_l;
    const double w = sqrt(z * z - g_t * g_l);
    const double a_c = a_l + (a_t - a_l) * (w - g_l - z) / (g_t - g_l + 2 * w);
    const double a_q = a_l - 0.5 * (a_l - a_t) * g_l / (g_l - (f_l - f_t) / (a_l - a_t));
    if (fabs(a_c - a_l) < fabs(a_q - a_l)) return a_c;
    return 0.5 * (a_q + a_c);
  }
  if (g_t * g_l < 0) {
    const double z = 3 * (f_t - f_l) / (a_t - a_l) - g_t - g_l;
    const double w = sqrt(z * z - g_t * g_l);
    const double a_c = a_l + (a_t - a_l) * (w - g_l - z) / (g_t - g_l + 2 * w);
    const double a_s = a_l - (a_l - a_t) / (g_l - g_t) * g_l;
    if (fabs(a_c - a_t) >= fabs(a_s - a_t)) return a_c;
    return a_s;
  }
  if (fabs(g_t) <= fabs(g_l)) {
    const double z = 3 * (f_t - f_l) / (a_t - a_l) - g_t - g_l;
    const double w = sqrt(z * z - g_t * g_l);
    const double a_c = a_l + (a_t - a_l) * (w - g_l - z) / (g_t - g_l + 2 * w);
    const double a_s = a_l - (a_l - a_t) / (g_l - g_t) * g_l;
    const double a_t_next = (fabs(a_c - a_t) < fabs(a_s - a_t)) ? a_c : a_s;
    const double lim = a_t + 0.66 * (a_u - a_t);
    if (a_t > a_l) return (a_t_next < lim) ? a_t_next : lim;  // std::min(lim, a_t_next)
    return (lim < a_t_next) ? a_t_next : lim;                  // std::max(lim, a_t_next)
  }
  const double z = 3 * (f_t - f_u) / (a_t - a_u) - g_t - g_u;
  const double w = sqrt(z * z - g_t * g_u);
  return a_u + (a_t - a_u) * (w - g_u - z) / (g_t - g_u + 2 * w);
}

// Newton direction for a well-conditioned H: Gaussian elimination with complete pivoting plus one step of
// iterative refinement.  Returns false (x untouched) unless the pivots certify cond(H) <~ 1/min_ratio; the caller
// then takes svd6_solve, whose rank truncation is what the reference's JacobiSVD::solve does for a (nearly)
// singular H (ndt_omp_impl.hpp:142-145).  For a full-rank H both give H^-1 b: they agree to ~cond(H) * 1e-16
// relative, i.e. <= ~1e-10 at min_ratio = 1e-6, far inside the 1e-9 per-iteration parity bar -- at 1/30 of the
// host time of the Jacobi sweeps, which sit on the critical path of every Newton iteration.
NM_HD bool gecp6_solve(const double* A, const double* b, double* x, double min_ratio) {
  double M[36];  // P A Q = L U in place (unit L below the diagonal)
  int rowperm[6], colperm[6];
  for (int i = 0; i < 36; ++i) M[i] = A[i];
  for (int i = 0; i < 6; ++i) { rowperm[i] = i; colperm[i] = i; }
  double first_pivot = 0.0;
  for (int k = 0; k < 6; ++k) {
    int pr = k, pc = k;
    double best = -1.0;
    for (int i = k; i < 6; ++i)
      for (int j = k; j < 6; ++j) {
        const double a = fabs(M[6 * i + j]);
        if (a > best) { best = a; pr = i; pc = j; }
      }
    if (k == 0) first_pivot = best;
    if (!(best >= first_pivot * min_ratio) || !(best > 0.0) || best > DBL_MAX) return false;  // also catches NaN
    if (pr != k) {
      for (int j = 0; j < 6; ++j) { const double t = M[6 * k + j]; M[6 * k + j] = M[6 * pr + j]; M[6 * pr + j] = t; }
      const int t = rowperm[k]; rowperm[k] = rowperm[pr]; rowperm[pr] = t;
    }
    if (pc != k) {
      for (int i = 0; i < 6; ++i) { const double t = M[6 * i + k]; M[6 * i + k] = M[6 * i + pc]; M[6 * i + pc] = t; }
      const int t = colperm[k]; colperm[k] = colperm[pc]; colperm[pc] = t;
    }
    const double inv = 1.0 / M[7 * k];
    for (int i = k + 1; i < 6; ++i) {
      const double f = M[6 * i + k] * inv;
      M[6 * i + k] = f;
      for (int j = k + 1; j < 6; ++j) M[6 * i + j] = M[6 * i + j] - f * M[6 * k + j];
    }
  }
  double sol[6] = {0, 0, 0, 0, 0, 0};
  double rhs[6];
  for (int i = 0; i < 6; ++i) rhs[i] = b[i];
  for (int pass = 0; pass < 2; ++pass) {  // pass 0: A x0 = b ; pass 1: one refinement step A d = b - A x0
    double y[6];
    for (int i = 0; i < 6; ++i) {
      double sum = rhs[rowperm[i]];
      for (int j = 0; j < i; ++j) sum = sum - M[6 * i + j] * y[j];
      y[i] = sum;
    }
    for (int i = 5; i >= 0; --i) {
      double sum = y[i];
      for (int j = i + 1; j < 6; ++j) sum = sum - M[6 * i + j] * y[j];
      y[i] = sum / M[7 * i];
    }
    for (int i = 0; i < 6; ++i) sol[colperm[i]] = sol[colperm[i]] + y[i];
    if (pass == 0) {
      for (int i = 0; i < 6; ++i) {
        double r = b[i];
        for (int j = 0; j < 6; ++j) r = r - A[6 * i + j] * sol[j];
        rhs[i] = r;
      }
    }
  }
  for (int i = 0; i < 6; ++i) if (!(fabs(sol[i]) <= DBL_MAX)) return false;
  for (int i = 0; i < 6; ++i) x[i] = sol[i];
  return true;
}

}  // namespace ndtb
