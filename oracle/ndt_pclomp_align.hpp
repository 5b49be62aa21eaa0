// ORACLE — TEST INFRASTRUCTURE ONLY (see small_linalg.hpp header).
//
// Newton loop + More-Thuente step of pclomp NDT:
//   extern/ndt_omp/include/pclomp/ndt_omp_impl.hpp:85-202 (computeTransformation),
//   :701-770 (computeHessian), :805-911 (updateIntervalMT / trialValueSelectionMT),
//   :915-1079 (computeStepLengthMT); ndt_omp.h:255-276 (auxiliaryFunction_*MT).
// pcl::Registration::align (PCL, not in tree): output := copy of input, final_transformation_ := I,
// then computeTransformation(output, guess).
#pragma once
#include "ndt_pclomp.hpp"

namespace oracle {

inline double aux_psi_mt(double a, double f_a, double f_0, double g_0, double mu) { return f_a - f_0 - mu * g_0 * a; }
inline double aux_dpsi_mt(double g_a, double g_0, double mu) { return g_a - mu * g_0; }

inline bool update_interval_mt(double& a_l, double& f_l, double& g_l, double& a_u, double& f_u, double& g_u, double a_t, double f_t, double g_t) {
  if (f_t > f_l) { a_u = a_t; f_u = f_t; g_u = g_t; return false; }
  if (g_t * (a_l - a_t) > 0) { a_l = a_t; f_l = f_t; g_l = g_t; return false; }
  if (g_t * (a_l - a_t) < 0) { a_u = a_l; f_u = f_l; g_u = g_l; a_l = a_t; f_l = f_t; g_l = g_t; return false; }
  return true;
}

inline double trial_value_selection_mt(double a_l, double f_l, double g_l, double a_u, double f_u, double g_u, double a_t, double f_t, double g_t) {
  if (f_t > f_l) {
    const double z = 3 * (f_t - f_l) / (a_t - a_l) - g_t - g_l;
    const double w = std::sqrt(z * z - g_t * g_l);
    const double a_c = a_l + (a_t - a_l) * (w - g_l - z) / (g_t - g_l + 2 * w);
    const double a_q = a_l - 0.5 * (a_l - a_t) * g_l / (g_l - (f_l - f_t) / (a_l - a_t));
    if (std::fabs(a_c - a_l) < std::fabs(a_q - a_l)) return a_c;
    return 0.5 * (a_q + a_c);
  }
  if (g_t * g_l < 0) {
    const double z = 3 * (f_t - f_l) / (a_t - a_l) - g_t - g_l;
    const double w = std::sqrt(z * z - g_t * g_l);
    const double a_c = a_l + (a_t - a_l) * (w - g_l - z) / (g_t - g_l + 2 * w);
    const double a_s = a_l - (a_l - a_t) / (g_l - g_t) * g_l;
    if (std::fabs(a_c - a_t) >= std::fabs(a_s - a_t)) return a_c;
    return a_s;
  }
  if (std::fabs(g_t) <= std::fabs(g_l)) {
    const double z = 3 * (f_t - f_l) / (a_t - a_l) - g_t - g_l;
    const double w = std::sqrt(z * z - g_t * g_l);
    const double a_c = a_l + (a_t - a_l) * (w - g_l - z) / (g_t - g_l + 2 * w);
    const double a_s = a_l - (a_l - a_t) / (g_l - g_t) * g_l;
    const double a_t_next = (std::fabs(a_c - a_t) < std::fabs(a_s - a_t)) ? a_c : a_s;
    if (a_t > a_l) return std::min(a_t + 0.66 * (a_u - a_t), a_t_next);
    return std::max(a_t + 0.66 * (a_u - a_t), a_t_next);
  }
  const double z = 3 * (f_t - f_u) / (a_t - a_u) - g_t - g_u;
  const double w = std::sqrt(z * z - g_t * g_u);
  return a_u + (a_t - a_u) * (w - g_u - z) / (g_t - g_u + 2 * w);
}

// FP64 angle vectors (the reference keeps them as Vector3d members, ndt_omp_impl.hpp:433-440,469-488)
inline void angle_vectors_f64(const double* p, double jd[8][3], double hd[15][3]) {
  double cx, cy, cz, sx, sy, sz;
  const double snap = 10e-5;
  if (std::fabs(p[3]) < snap) { cx = 1.0; sx = 0.0; } else { cx = std::cos(p[3]); sx = std::sin(p[3]); }
  if (std::fabs(p[4]) < snap) { cy = 1.0; sy = 0.0; } else { cy = std::cos(p[4]); sy = std::sin(p[4]); }
  if (std::fabs(p[5]) < snap) { cz = 1.0; sz = 0.0; } else { cz = std::cos(p[5]); sz = std::sin(p[5]); }
  const double j[8][3] = {{(-sx * sz + cx * sy * cz), (-sx * cz - cx * sy * sz), (-cx * cy)}, {(cx * sz + sx * sy * cz), (cx * cz - sx * sy * sz), (-sx * cy)},
                          {(-sy * cz), sy * sz, cy}, {sx * cy * cz, (-sx * cy * sz), sx * sy}, {(-cx * cy * cz), cx * cy * sz, (-cx * sy)},
                          {(-cy * sz), (-cy * cz), 0}, {(cx * cz - sx * sy * sz), (-cx * sz - sx * sy * cz), 0}, {(sx * cz + cx * sy * sz), (cx * sy * cz - sx * sz), 0}};
  const double h[15][3] = {{(-cx * sz - sx * sy * cz), (-cx * cz + sx * sy * sz), sx * cy}, {(-sx * sz + cx * sy * cz), (-cx * sy * sz - sx * cz), (-cx * cy)},
                           {(cx * cy * cz), (-cx * cy * sz), (cx * sy)}, {(sx * cy * cz), (-sx * cy * sz), (sx * sy)},
                           {(-sx * cz - cx * sy * sz), (sx * sz - cx * sy * cz), 0}, {(cx * cz - sx * sy * sz), (-sx * sy * cz - cx * sz), 0},
                           {(-cy * cz), (cy * sz), (sy)}, {(-sx * sy * cz), (sx * sy * sz), (sx * cy)}, {(cx * sy * cz), (-cx * sy * sz), (-cx * cy)},
                           {(sy * sz), (sy * cz), 0}, {(-sx * cy * sz), (-sx * cy * cz), 0}, {(cx * cy * sz), (cx * cy * cz), 0},
                           {(-cy * cz), (cy * sz), 0}, {(-cx * sz - sx * sy * cz), (-cx * cz + sx * sy * sz), 0}, {(-sx * sz + cx * sy * cz), (-cx * sy * sz - sx * cz), 0}};
  std::memcpy(jd, j, sizeof(j));
  std::memcpy(hd, h, sizeof(h));
}

inline void PclompNdt::compute_hessian_f64(double* H, const std::vector<float>& trans) {
  // ndt_omp_impl.hpp:701-770 -- serial, FP64, uses the angle vectors of the LAST computeDerivatives.
  std::memset(H, 0, 36 * sizeof(double));
  const size_t N = source.size() / 3;
  for (size_t idx = 0; idx < N; ++idx) {
    const float* xt = &trans[3 * idx];
    const Leaf* nb[MAX_NB];
    const int nn = find_neighbours(xt, nb);
    for (int k = 0; k < nn; ++k) {
      const double x[3] = {source[3 * idx], source[3 * idx + 1], source[3 * idx + 2]};
      const double x_trans[3] = {static_cast<double>(xt[0]) - nb[k]->mean[0], static_cast<double>(xt[1]) - nb[k]->mean[1], static_cast<double>(xt[2]) - nb[k]->mean[2]};
      update_hessian_f64(H, tables, jd, hd, x, x_trans, nb[k]->icov, gauss_d1, gauss_d2);
    }
  }
}

inline void push_trace(NdtResult& res, const double* p, double score, const double* g, const double* H, double nvtl) {
  for (int i = 0; i < 6; ++i) res.trace.push_back(p[i]);
  res.trace.push_back(score);
  for (int i = 0; i < 6; ++i) res.trace.push_back(g[i]);
  for (int i = 0; i < 36; ++i) res.trace.push_back(H[i]);
  res.trace.push_back(nvtl);
}

inline double PclompNdt::step_length_mt(const double* x, double* step_dir, double step_init, double step_max, double step_min, double& score,
                                        double* g, double* H, std::vector<float>& trans, float* final_T, NdtResult& res) {
  const double phi_0 = -score;
  double d_phi_0 = 0;
  for (int i = 0; i < 6; ++i) d_phi_0 += g[i] * step_dir[i];
  d_phi_0 = -d_phi_0;
  double x_t[6];
  if (d_phi_0 >= 0) {
    if (d_phi_0 == 0) return 0;
    d_phi_0 *= -1;
    for (int i = 0; i < 6; ++i) step_dir[i] *= -1;
  }
  const int max_step_iterations = 10;
  int step_iterations = 0;
  const double mu = 1.e-4, nu = 0.9;
  double a_l = 0, a_u = 0;
  double f_l = aux_psi_mt(a_l, phi_0, phi_0, d_phi_0, mu), g_l = aux_dpsi_mt(d_phi_0, d_phi_0, mu);
  double f_u = aux_psi_mt(a_u, phi_0, phi_0, d_phi_0, mu), g_u = aux_dpsi_mt(d_phi_0, d_phi_0, mu);
  bool interval_converged = (step_max - step_min) < 0, open_interval = true;
  double a_t = step_init;
  a_t = std::min(a_t, step_max);
  a_t = std::max(a_t, step_min);
  for (int i = 0; i < 6; ++i) x_t[i] = x[i] + step_dir[i] * a_t;
  pose_to_matrix_f(x_t, final_T);
  transform_cloud(final_T, trans);
  score = compute_derivatives(g, H, trans, x_t, true);
  angle_vectors_f64(x_t, jd, hd);
  push_trace(res, x_t, score, g, H, nvtl);
  if (params.use_line_search) {
    double phi_t = -score, d_phi_t = 0;
    for (int i = 0; i < 6; ++i) d_phi_t += g[i] * step_dir[i];
    d_phi_t = -d_phi_t;
    double psi_t = aux_psi_mt(a_t, phi_t, phi_0, d_phi_0, mu), d_psi_t = aux_dpsi_mt(d_phi_t, d_phi_0, mu);
    while (!interval_converged && step_iterations < max_step_iterations && !(psi_t <= 0 && d_phi_t <= -nu * d_phi_0)) {
      if (open_interval) a_t = trial_value_selection_mt(a_l, f_l, g_l, a_u, f_u, g_u, a_t, psi_t, d_psi_t);
      else a_t = trial_value_selection_mt(a_l, f_l, g_l, a_u, f_u, g_u, a_t, phi_t, d_phi_t);
      a_t = std::min(a_t, step_max);
      a_t = std::max(a_t, step_min);
      for (int i = 0; i < 6; ++i) x_t[i] = x[i] + step_dir[i] * a_t;
      pose_to_matrix_f(x_t, final_T);
      transform_cloud(final_T, trans);
      score = compute_derivatives(g, H, trans, x_t, false);
      angle_vectors_f64(x_t, jd, hd);
      phi_t = -score; d_phi_t = 0;
      for (int i = 0; i < 6; ++i) d_phi_t += g[i] * step_dir[i];
      d_phi_t = -d_phi_t;
      psi_t = aux_psi_mt(a_t, phi_t, phi_0, d_phi_0, mu);
      d_psi_t = aux_dpsi_mt(d_phi_t, d_phi_0, mu);
      if (open_interval && (psi_t <= 0 && d_psi_t >= 0)) {
        open_interval = false;
        f_l = f_l + phi_0 - mu * d_phi_0 * a_l; g_l = g_l + mu * d_phi_0;
        f_u = f_u + phi_0 - mu * d_phi_0 * a_u; g_u = g_u + mu * d_phi_0;
      }
      if (open_interval) interval_converged = update_interval_mt(a_l, f_l, g_l, a_u, f_u, g_u, a_t, psi_t, d_psi_t);
      else interval_converged = update_interval_mt(a_l, f_l, g_l, a_u, f_u, g_u, a_t, phi_t, d_phi_t);
      step_iterations++;
    }
  }
  if (step_iterations) {
    compute_hessian_f64(H, trans);
    push_trace(res, x_t, score, g, H, nvtl);  // extra trace row: the FP64 Hessian-only pass (parity hook, not in the reference)
  }
  return a_t;
}

inline void PclompNdt::align(const float* guess, NdtResult& res) {
  const size_t N = source.size() / 3;
  res = NdtResult();
  int nr_iterations = 0;
  bool converged = false;
  update_gauss();
  float final_T[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
  std::vector<float> trans(source);  // pcl::Registration::align: output = *input_
  bool guess_is_identity = true;
  for (int i = 0; i < 16; ++i) if (guess[i] != final_T[i]) guess_is_identity = false;
  if (!guess_is_identity) { std::memcpy(final_T, guess, sizeof(final_T)); transform_cloud(guess, trans); }
  res.transformation_array.insert(res.transformation_array.end(), final_T, final_T + 16);
  double p[6], delta_p[6], g[6], H[36];
  float eul[3];
  euler_angles_012_f(final_T, eul);
  p[0] = final_T[3]; p[1] = final_T[7]; p[2] = final_T[11]; p[3] = eul[0]; p[4] = eul[1]; p[5] = eul[2];
  double score = compute_derivatives(g, H, trans, p, true);
  angle_vectors_f64(p, jd, hd);
  push_trace(res, p, score, g, H, nvtl);
  hessian_stored = false;
  while (!converged) {
    double neg_g[6];
    for (int i = 0; i < 6; ++i) neg_g[i] = -g[i];
    svd6_solve(H, neg_g, delta_p);
    double nrm2 = 0;
    for (int i = 0; i < 6; ++i) nrm2 += delta_p[i] * delta_p[i];
    double delta_p_norm = std::sqrt(nrm2);
    if (delta_p_norm == 0 || delta_p_norm != delta_p_norm) {
      res.trans_probability = N ? score / static_cast<double>(N) : 0.0;
      converged = delta_p_norm == delta_p_norm;
      break;  // early return: hessian_ not stored (ndt_omp_impl.hpp:150-159)
    }
    for (int i = 0; i < 6; ++i) delta_p[i] = delta_p[i] / delta_p_norm;
    delta_p_norm = step_length_mt(p, delta_p, delta_p_norm, params.step_size, params.trans_epsilon / 2, score, g, H, trans, final_T, res);
    for (int i = 0; i < 6; ++i) delta_p[i] *= delta_p_norm;
    res.transformation_array.insert(res.transformation_array.end(), final_T, final_T + 16);
    for (int i = 0; i < 6; ++i) p[i] = p[i] + delta_p[i];
    if (nr_iterations > params.max_iterations || (nr_iterations && (std::fabs(delta_p_norm) < params.trans_epsilon))) converged = true;
    nr_iterations++;
    if (converged) {
      res.trans_probability = N ? score / static_cast<double>(N) : 0.0;
      std::memcpy(stored_hessian, H, sizeof(stored_hessian));
      hessian_stored = true;
    }
  }
  std::memcpy(res.final_transformation, final_T, sizeof(final_T));
  std::memcpy(res.hessian, stored_hessian, sizeof(stored_hessian));
  res.nvtl = nvtl;
  res.iterations = nr_iterations;
  res.converged = converged;
}

}  // namespace oracle
