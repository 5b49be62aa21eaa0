// ORACLE — TEST INFRASTRUCTURE ONLY (see small_linalg.hpp header).
//
// CPU restatement of pclomp::NormalDistributionsTransform, following
//   extern/ndt_omp/include/pclomp/ndt_omp_impl.hpp:51-81 (ctor defaults), :85-202
//   (computeTransformation), :217-392 (computeDerivatives), :396-538 (computeAngleDerivatives),
//   :542-587 (computePointDerivatives, float overload), :640-697 (updateDerivatives),
//   :701-801 (computeHessian/updateHessian), :805-1079 (More-Thuente), :1169-1281 (score helpers)
//   and extern/ndt_omp/include/pclomp/ndt_omp.h:205-222 (pose -> matrix).
//
// Parity: pinned to the reference's published output (README fitness of the bundled scan pair, six digits,
// tests/test_real_scan_pair.py); not bit-pinned (Eigen/PCL are absent, no reference binary).  Evaluation order:
//   * every float dot/matrix product: k ascending, left-to-right adds, no FMA, padded 4th lane
//     (always 0) skipped -- bit-identical to Eigen's scalar path for finite inputs;
//   * exp at ndt_omp_impl.hpp:655 = float(exp(double(arg)));
//   * pcl::transformPointCloud (PCL SSE2 Transformer): x' = m00*x + (m01*y + (m02*z + m03));
//   * Eigen 3.4 eulerAngles(0,1,2) incl. its [0,pi] first-angle canonicalisation;
//   * AngleAxis<float>::toRotationMatrix + (Rx*Ry)*Rz float products, sinf/cosf from libm.
#pragma once
#include <vector>
#include <cmath>
#include <cstring>
#include <omp.h>
#include "voxel_grid.hpp"

namespace oracle {

enum SearchMethod { KDTREE = 0, DIRECT26 = 1, DIRECT7 = 2, DIRECT1 = 3 };  // ndt_struct.hpp order

struct AngleTables {
  float j_ang[8][3];
  float h_ang[15][3];
};

// ndt_omp_impl.hpp:396-538 (snap 10e-5) / svn_ndt_impl.hpp:255-334 (snap 1e-7)
inline void compute_angle_derivatives(const double* p, double snap, AngleTables& t) {
  double cx, cy, cz, sx, sy, sz;
  if (std::fabs(p[3]) < snap) { cx = 1.0; sx = 0.0; } else { cx = std::cos(p[3]); sx = std::sin(p[3]); }
  if (std::fabs(p[4]) < snap) { cy = 1.0; sy = 0.0; } else { cy = std::cos(p[4]); sy = std::sin(p[4]); }
  if (std::fabs(p[5]) < snap) { cz = 1.0; sz = 0.0; } else { cz = std::cos(p[5]); sz = std::sin(p[5]); }
  const double j[8][3] = {
      {(-sx * sz + cx * sy * cz), (-sx * cz - cx * sy * sz), (-cx * cy)},
      {(cx * sz + sx * sy * cz), (cx * cz - sx * sy * sz), (-sx * cy)},
      {(-sy * cz), sy * sz, cy},
      {sx * cy * cz, (-sx * cy * sz), sx * sy},
      {(-cx * cy * cz), cx * cy * sz, (-cx * sy)},
      {(-cy * sz), (-cy * cz), 0},
      {(cx * cz - sx * sy * sz), (-cx * sz - sx * sy * cz), 0},
      {(sx * cz + cx * sy * sz), (cx * sy * cz - sx * sz), 0}};
  const double h[15][3] = {
      {(-cx * sz - sx * sy * cz), (-cx * cz + sx * sy * sz), sx * cy},      // a2
      {(-sx * sz + cx * sy * cz), (-cx * sy * sz - sx * cz), (-cx * cy)},   // a3
      {(cx * cy * cz), (-cx * cy * sz), (cx * sy)},                         // b2
      {(sx * cy * cz), (-sx * cy * sz), (sx * sy)},                         // b3
      {(-sx * cz - cx * sy * sz), (sx * sz - cx * sy * cz), 0},             // c2
      {(cx * cz - sx * sy * sz), (-sx * sy * cz - cx * sz), 0},             // c3
      {(-cy * cz), (cy * sz), (sy)},                                        // d1
      {(-sx * sy * cz), (sx * sy * sz), (sx * cy)},                         // d2
      {(cx * sy * cz), (-cx * sy * sz), (-cx * cy)},                        // d3
      {(sy * sz), (sy * cz), 0},                                            // e1
      {(-sx * cy * sz), (-sx * cy * cz), 0},                                // e2
      {(cx * cy * sz), (cx * cy * cz), 0},                                  // e3
      {(-cy * cz), (cy * sz), 0},                                           // f1
      {(-cx * sz - sx * sy * cz), (-cx * cz + sx * sy * sz), 0},            // f2
      {(-sx * sz + cx * sy * cz), (-cx * sy * sz - sx * cz), 0}};           // f3
  for (int r = 0; r < 8; ++r) for (int c = 0; c < 3; ++c) t.j_ang[r][c] = static_cast<float>(j[r][c]);
  for (int r = 0; r < 15; ++r) for (int c = 0; c < 3; ++c) t.h_ang[r][c] = static_cast<float>(h[r][c]);
}

inline float dot3f(const float* a, const float* b) { float s = a[0] * b[0]; s = s + a[1] * b[1]; s = s + a[2] * b[2]; return s; }

// Dense 4x6 Jp and 24x6 Hp exactly as the reference lays them out (ndt_omp_impl.hpp:542-587).
struct PointDeriv {
  float Jp[4][6];
  float Hp[24][6];
};

inline void compute_point_derivatives(const AngleTables& t, const double* x, PointDeriv& d, bool compute_hessian) {
  const float x4[3] = {static_cast<float>(x[0]), static_cast<float>(x[1]), static_cast<float>(x[2])};
  float xj[8];
  for (int r = 0; r < 8; ++r) xj[r] = dot3f(t.j_ang[r], x4);
  d.Jp[1][3] = xj[0]; d.Jp[2][3] = xj[1];
  d.Jp[0][4] = xj[2]; d.Jp[1][4] = xj[3]; d.Jp[2][4] = xj[4];
  d.Jp[0][5] = xj[5]; d.Jp[1][5] = xj[6]; d.Jp[2][5] = xj[7];
  if (compute_hessian) {
    float xh[15];
    for (int r = 0; r < 15; ++r) xh[r] = dot3f(t.h_ang[r], x4);
    const float a[4] = {0, xh[0], xh[1], 0}, b[4] = {0, xh[2], xh[3], 0}, c[4] = {0, xh[4], xh[5], 0};
    const float dd[4] = {xh[6], xh[7], xh[8], 0}, e[4] = {xh[9], xh[10], xh[11], 0}, f[4] = {xh[12], xh[13], xh[14], 0};
    for (int k = 0; k < 4; ++k) {
      d.Hp[12 + k][3] = a[k]; d.Hp[16 + k][3] = b[k]; d.Hp[20 + k][3] = c[k];
      d.Hp[12 + k][4] = b[k]; d.Hp[16 + k][4] = dd[k]; d.Hp[20 + k][4] = e[k];
      d.Hp[12 + k][5] = c[k]; d.Hp[16 + k][5] = e[k]; d.Hp[20 + k][5] = f[k];
    }
  }
}

inline void init_point_deriv(PointDeriv& d) {
  std::memset(&d, 0, sizeof(d));
  d.Jp[0][0] = d.Jp[1][1] = d.Jp[2][2] = 1.0f;
}

// ndt_omp_impl.hpp:640-697.  g[6], H[36] (row-major) are FP64 accumulators.
inline double update_derivatives_pclomp(double* g, double* H, const PointDeriv& d, const double* x_trans,
                                        const double* c_inv, double gauss_d1, double gauss_d2_d, bool compute_hessian) {
  const float x4[3] = {static_cast<float>(x_trans[0]), static_cast<float>(x_trans[1]), static_cast<float>(x_trans[2])};
  float C[3][3];
  for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) C[r][c] = static_cast<float>(c_inv[3 * r + c]);
  const float gauss_d2 = static_cast<float>(gauss_d2_d);
  float xC[3];  // x_trans4 * c_inv4 (row vector times matrix)
  for (int c = 0; c < 3; ++c) { float s = x4[0] * C[0][c]; s = s + x4[1] * C[1][c]; s = s + x4[2] * C[2][c]; xC[c] = s; }
  const float q = dot3f(x4, xC);
  float e_x_cov_x = static_cast<float>(std::exp(static_cast<double>(-gauss_d2 * q * 0.5f)));
  const float score_inc = static_cast<float>(-gauss_d1 * static_cast<double>(e_x_cov_x));
  e_x_cov_x = gauss_d2 * e_x_cov_x;
  if (e_x_cov_x > 1 || e_x_cov_x < 0 || e_x_cov_x != e_x_cov_x) return 0;
  e_x_cov_x = static_cast<float>(static_cast<double>(e_x_cov_x) * gauss_d1);
  float CJ[3][6];  // c_inv4 * point_gradient4 (4th row is zero)
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 6; ++c) { float s = C[r][0] * d.Jp[0][c]; s = s + C[r][1] * d.Jp[1][c]; s = s + C[r][2] * d.Jp[2][c]; CJ[r][c] = s; }
  float xCJ[6];
  for (int c = 0; c < 6; ++c) { float s = x4[0] * CJ[0][c]; s = s + x4[1] * CJ[1][c]; s = s + x4[2] * CJ[2][c]; xCJ[c] = s; }
  for (int c = 0; c < 6; ++c) g[c] += static_cast<double>(e_x_cov_x * xCJ[c]);
  if (compute_hessian) {
    float JCJ[6][6];  // point_gradient4^T * (c_inv4 * point_gradient4)
    for (int r = 0; r < 6; ++r)
      for (int c = 0; c < 6; ++c) { float s = d.Jp[0][r] * CJ[0][c]; s = s + d.Jp[1][r] * CJ[1][c]; s = s + d.Jp[2][r] * CJ[2][c]; JCJ[r][c] = s; }
    for (int i = 0; i < 6; ++i) {
      float xCH[6];
      for (int j = 0; j < 6; ++j) { float s = xC[0] * d.Hp[4 * i + 0][j]; s = s + xC[1] * d.Hp[4 * i + 1][j]; s = s + xC[2] * d.Hp[4 * i + 2][j]; xCH[j] = s; }
      for (int j = 0; j < 6; ++j)
        H[6 * i + j] += static_cast<double>(e_x_cov_x * (-gauss_d2 * xCJ[i] * xCJ[j] + xCH[j] + JCJ[j][i]));
    }
  }
  return static_cast<double>(score_inc);
}

// FP64 variant used by computeHessian after a line search (ndt_omp_impl.hpp:774-801) and the
// double point-derivative overload (:591-636).
inline void update_hessian_f64(double* H, const AngleTables& t_unused, const double jd[8][3], const double hd[15][3],
                               const double* x, const double* x_trans, const double* c_inv, double d1, double d2) {
  (void)t_unused;
  double Jp[3][6] = {{1, 0, 0, 0, 0, 0}, {0, 1, 0, 0, 0, 0}, {0, 0, 1, 0, 0, 0}};
  double Hp[18][6]; std::memset(Hp, 0, sizeof(Hp));
  auto dot = [&](const double* a) { return (x[0] * a[0] + x[1] * a[1]) + x[2] * a[2]; };
  Jp[1][3] = dot(jd[0]); Jp[2][3] = dot(jd[1]); Jp[0][4] = dot(jd[2]); Jp[1][4] = dot(jd[3]); Jp[2][4] = dot(jd[4]);
  Jp[0][5] = dot(jd[5]); Jp[1][5] = dot(jd[6]); Jp[2][5] = dot(jd[7]);
  const double a[3] = {0, dot(hd[0]), dot(hd[1])}, b[3] = {0, dot(hd[2]), dot(hd[3])}, c[3] = {0, dot(hd[4]), dot(hd[5])};
  const double dd[3] = {dot(hd[6]), dot(hd[7]), dot(hd[8])}, e[3] = {dot(hd[9]), dot(hd[10]), dot(hd[11])}, f[3] = {dot(hd[12]), dot(hd[13]), dot(hd[14])};
  for (int k = 0; k < 3; ++k) {
    Hp[9 + k][3] = a[k]; Hp[12 + k][3] = b[k]; Hp[15 + k][3] = c[k];
    Hp[9 + k][4] = b[k]; Hp[12 + k][4] = dd[k]; Hp[15 + k][4] = e[k];
    Hp[9 + k][5] = c[k]; Hp[12 + k][5] = e[k]; Hp[15 + k][5] = f[k];
  }
  auto Cv = [&](const double* v, double* out) { for (int r = 0; r < 3; ++r) out[r] = (c_inv[3 * r] * v[0] + c_inv[3 * r + 1] * v[1]) + c_inv[3 * r + 2] * v[2]; };
  auto d3 = [&](const double* u, const double* v) { return (u[0] * v[0] + u[1] * v[1]) + u[2] * v[2]; };
  double Cx[3]; Cv(x_trans, Cx);
  double e_x = d2 * std::exp(-d2 * d3(x_trans, Cx) / 2);
  if (e_x > 1 || e_x < 0 || e_x != e_x) return;
  e_x *= d1;
  for (int i = 0; i < 6; ++i) {
    const double ci[3] = {Jp[0][i], Jp[1][i], Jp[2][i]};
    double cov_dxd_pi[3]; Cv(ci, cov_dxd_pi);
    for (int j = 0; j < 6; ++j) {
      const double cj[3] = {Jp[0][j], Jp[1][j], Jp[2][j]};
      double Ccj[3]; Cv(cj, Ccj);
      const double hij[3] = {Hp[3 * i][j], Hp[3 * i + 1][j], Hp[3 * i + 2][j]};
      double Ch[3]; Cv(hij, Ch);
      H[6 * i + j] += e_x * (-d2 * d3(x_trans, cov_dxd_pi) * d3(x_trans, Ccj) + d3(x_trans, Ch) + d3(cj, cov_dxd_pi));
    }
  }
}

// ---------------------------------------------------------------- pose <-> matrix helpers (float)
// Column-major 4x4 would be Eigen's storage; we keep ROW-major float[16] everywhere in the ABI.
inline void mat3f_mul(const float* a, const float* b, float* out) {
  float r[9];
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) { float s = a[3 * i] * b[j]; s = s + a[3 * i + 1] * b[3 + j]; s = s + a[3 * i + 2] * b[6 + j]; r[3 * i + j] = s; }
  std::memcpy(out, r, sizeof(r));
}

// Eigen::AngleAxis<float>(angle, Unit{X,Y,Z}).toRotationMatrix()
inline void angle_axis_f(float angle, int axis, float* R) {
  float ax[3] = {0, 0, 0}; ax[axis] = 1.0f;
  const float s = std::sin(angle), c = std::cos(angle);
  const float sin_axis[3] = {s * ax[0], s * ax[1], s * ax[2]};
  const float cos1_axis[3] = {(1.0f - c) * ax[0], (1.0f - c) * ax[1], (1.0f - c) * ax[2]};
  float tmp;
  tmp = cos1_axis[0] * ax[1]; R[1] = tmp - sin_axis[2]; R[3] = tmp + sin_axis[2];
  tmp = cos1_axis[0] * ax[2]; R[2] = tmp + sin_axis[1]; R[6] = tmp - sin_axis[1];
  tmp = cos1_axis[1] * ax[2]; R[5] = tmp - sin_axis[0]; R[7] = tmp + sin_axis[0];
  R[0] = cos1_axis[0] * ax[0] + c; R[4] = cos1_axis[1] * ax[1] + c; R[8] = cos1_axis[2] * ax[2] + c;
}

// Translation(x,y,z) * Rx * Ry * Rz in float (ndt_omp_impl.hpp:167-174, 969-975).
inline void pose_to_matrix_f(const double* p, float* T) {
  float Rx[9], Ry[9], Rz[9], R[9];
  angle_axis_f(static_cast<float>(p[3]), 0, Rx);
  angle_axis_f(static_cast<float>(p[4]), 1, Ry);
  angle_axis_f(static_cast<float>(p[5]), 2, Rz);
  mat3f_mul(Rx, Ry, R);
  mat3f_mul(R, Rz, R);
  for (int r = 0; r < 3; ++r) { for (int c = 0; c < 3; ++c) T[4 * r + c] = R[3 * r + c]; T[4 * r + 3] = static_cast<float>(p[r]); }
  T[12] = 0; T[13] = 0; T[14] = 0; T[15] = 1;
}

// Eigen 3.4 MatrixBase::eulerAngles(0,1,2) on the float rotation block (ndt_omp_impl.hpp:118).
inline void euler_angles_012_f(const float* T, float* res) {
  auto m = [&](int r, int c) { return T[4 * r + c]; };
  const int i = 0, j = 1, k = 2;  // odd = 0
  res[0] = std::atan2(m(j, k), m(k, k));
  const float c2 = std::sqrt(m(i, i) * m(i, i) + m(i, j) * m(i, j));
  if (res[0] > 0.0f) {
    res[0] -= static_cast<float>(M_PI);
    res[1] = std::atan2(-m(i, k), -c2);
  } else {
    res[1] = std::atan2(-m(i, k), c2);
  }
  const float s1 = std::sin(res[0]), c1 = std::cos(res[0]);
  res[2] = std::atan2(s1 * m(k, i) - c1 * m(j, i), c1 * m(j, j) - s1 * m(k, j));
  res[0] = -res[0]; res[1] = -res[1]; res[2] = -res[2];
}

// pcl::transformPointCloud with a Matrix4f (PCL SSE2 Transformer::se3 association).
inline void transform_point_f(const float* T, const float* p, float* o) {
  for (int r = 0; r < 3; ++r) o[r] = T[4 * r] * p[0] + (T[4 * r + 1] * p[1] + (T[4 * r + 2] * p[2] + T[4 * r + 3]));
}

struct NdtParams {
  double trans_epsilon = 0.1;     // ndt_omp_impl.hpp:63-70
  double step_size = 0.1;
  float resolution = 1.0f;
  int max_iterations = 35;
  int search_method = DIRECT7;
  int num_threads = 1;
  float regularization_scale_factor = 0.0f;
  bool use_line_search = false;
};

struct NdtResult {
  float final_transformation[16];
  double hessian[36];
  double trans_probability = 0, nvtl = 0;
  int iterations = 0;
  bool converged = false;
  std::vector<float> transformation_array;  // 16 floats per entry
  // per-evaluation trace for parity tests: p(6), score, g(6), H(36), nvtl  = 50 doubles
  std::vector<double> trace;
};

struct PclompNdt {
  NdtParams params;
  double outlier_ratio = 0.55;
  VoxelGrid grid;
  std::vector<float> source;  // xyz packed, 3 floats per point
  bool has_reg_pose = false;
  float reg_translation[3] = {0, 0, 0};
  double gauss_d1 = 0, gauss_d2 = 0, gauss_d3 = 0;
  double nvtl = 0;
  bool hessian_stored = false;
  double stored_hessian[36];
  AngleTables tables;
  double jd[8][3], hd[15][3];

  PclompNdt() { grid.flavour = FLAVOUR_PCLOMP; update_gauss(); std::memset(stored_hessian, 0, sizeof(stored_hessian)); }

  void update_gauss() {  // ndt_omp_impl.hpp:95-99
    const double c1 = 10 * (1 - outlier_ratio);
    const double c2 = outlier_ratio / std::pow(static_cast<double>(params.resolution), 3);
    gauss_d3 = -std::log(c2);
    gauss_d1 = -std::log(c1 + c2) - gauss_d3;
    gauss_d2 = -2 * std::log((-std::log(c1 * std::exp(-0.5) + c2) - gauss_d3) / gauss_d1);
  }

  int set_target(const float* xyz, size_t n, size_t stride_floats, bool is_dense) {
    grid.set_leaf_size(params.resolution);
    const int rc = grid.build(xyz, n, stride_floats, is_dense);
    grid.prepare_radius_search();
    return rc;
  }
  // The neighbour search switch of ndt_omp_impl.hpp:279-294 (same at :734-749 for computeHessian).
  int find_neighbours(const float* xt, const Leaf** nb) const {
    switch (params.search_method) {
      case KDTREE: {
        std::vector<std::pair<float, const Leaf*>> hits;
        grid.radius_search(xt, static_cast<double>(params.resolution), false, hits);
        std::stable_sort(hits.begin(), hits.end(), [](const std::pair<float, const Leaf*>& a, const std::pair<float, const Leaf*>& b) { return a.first < b.first; });
        int n = 0;
        for (auto& h : hits) if (n < MAX_NB) nb[n++] = h.second;
        return n;
      }
      case DIRECT26: return grid.neighbours26(xt[0], xt[1], xt[2], nb);
      case DIRECT1: return grid.neighbours(xt[0], xt[1], xt[2], 1, nb);
      default: return grid.neighbours(xt[0], xt[1], xt[2], 7, nb);
    }
  }
  static constexpr int MAX_NB = 128;
  void set_source(const float* xyz, size_t n, size_t stride_floats) {
    source.resize(3 * n);
    for (size_t i = 0; i < n; ++i) for (int a = 0; a < 3; ++a) source[3 * i + a] = xyz[i * stride_floats + a];
  }

  // ndt_omp_impl.hpp:217-392.  trans = transformed cloud (3 floats per point).
  double compute_derivatives(double* g, double* H, const std::vector<float>& trans, const double* p, bool compute_hessian) {
    const size_t N = source.size() / 3;
    std::memset(g, 0, 6 * sizeof(double));
    std::memset(H, 0, 36 * sizeof(double));
    std::vector<double> scores(N, 0.0), nearest(N, 0.0), grads(6 * N, 0.0), hess(36 * N, 0.0);
    std::vector<size_t> found(N, 0);
    std::vector<int> counts(N, 0);
    compute_angle_derivatives(p, 10e-5, tables);
    const int nt = std::max(1, params.num_threads);
#pragma omp parallel for num_threads(nt) schedule(guided, 8)
    for (long idx = 0; idx < static_cast<long>(N); ++idx) {
      PointDeriv pd; init_point_deriv(pd);
      const float* xt = &trans[3 * idx];
      const Leaf* nb[MAX_NB];
      const int nn = find_neighbours(xt, nb);
      double sum_score = 0, nearest_pt = 0;
      double* gp = &grads[6 * idx];
      double* hp = &hess[36 * idx];
      for (int k = 0; k < nn; ++k) {
        const double x[3] = {source[3 * idx], source[3 * idx + 1], source[3 * idx + 2]};
        const double x_trans[3] = {static_cast<double>(xt[0]) - nb[k]->mean[0], static_cast<double>(xt[1]) - nb[k]->mean[1],
                                   static_cast<double>(xt[2]) - nb[k]->mean[2]};
        compute_point_derivatives(tables, x, pd, compute_hessian);
        const double s = update_derivatives_pclomp(gp, hp, pd, x_trans, nb[k]->icov, gauss_d1, gauss_d2, compute_hessian);
        sum_score += s;
        if (s > nearest_pt) nearest_pt = s;
      }
      if (nn > 0) ++found[idx];
      scores[idx] = sum_score; nearest[idx] = nearest_pt; counts[idx] += nn;
    }
    double score = 0, nearest_sum = 0; size_t found_num = 0; int total_count = 0;
    for (size_t i = 0; i < N; ++i) {
      score += scores[i]; nearest_sum += nearest[i]; found_num += found[i];
      for (int c = 0; c < 6; ++c) g[c] += grads[6 * i + c];
      for (int c = 0; c < 36; ++c) H[c] += hess[36 * i + c];
      total_count += counts[i];
    }
    if (has_reg_pose) {  // ndt_omp_impl.hpp:349-382
      const float dx = reg_translation[0] - static_cast<float>(p[0]);
      const float dy = reg_translation[1] - static_cast<float>(p[1]);
      const float sin_yaw = static_cast<float>(std::sin(p[5])), cos_yaw = static_cast<float>(std::cos(p[5]));
      const float longitudinal = dy * sin_yaw + dx * cos_yaw;
      const float w = static_cast<float>(total_count);
      const float sf = params.regularization_scale_factor;
      const float reg_score = -sf * w * longitudinal * longitudinal;
      score += reg_score;
      g[0] += static_cast<double>(sf * w * 2.0f * cos_yaw * longitudinal);
      g[1] += static_cast<double>(sf * w * 2.0f * sin_yaw * longitudinal);
      const double h00 = -sf * w * 2.0f * cos_yaw * cos_yaw, h01 = -sf * w * 2.0f * cos_yaw * sin_yaw, h11 = -sf * w * 2.0f * sin_yaw * sin_yaw;
      H[0] += h00; H[1] += h01; H[7] += h11; H[6] += h01;
    }
    nvtl = found_num ? nearest_sum / static_cast<double>(found_num) : 0.0;
    return score;
  }

  void transform_cloud(const float* T, std::vector<float>& out) const {
    const size_t N = source.size() / 3;
    out.resize(3 * N);
    for (size_t i = 0; i < N; ++i) transform_point_f(T, &source[3 * i], &out[3 * i]);
  }

  void align(const float* guess, NdtResult& res);
  double step_length_mt(const double* x, double* step_dir, double step_init, double step_max, double step_min, double& score,
                        double* g, double* H, std::vector<float>& trans, float* final_T, NdtResult& res);
  void compute_hessian_f64(double* H, const std::vector<float>& trans);
};

}  // namespace oracle
