// ORACLE — TEST INFRASTRUCTURE ONLY (see small_linalg.hpp header).
//
// CPU restatement of svn_ndt::SvnNormalDistributionsTransform, following
//   extern/svn_ndt/include/svn_ndt_impl.hpp:60-131 (ctor, updateNdtConstants), :214-244 (RBF kernel),
//   :255-396 (angle / point derivatives), :401-513 (updateDerivatives), :518-668
//   (computeParticleDerivatives), :675-964 (align).
// Deviations that make it a deterministic oracle (documented in DESIGN.md):
//   * initial particles are supplied by the caller or drawn from std::mt19937_64(seed) with the
//     reference's sigmas (:710); the reference seeds from the wall clock (:712);
//   * stage 1 computes the angle tables per particle (the reference races on members, :527).
#pragma once
#include <vector>
#include <algorithm>
#include <random>
#include "ndt_pclomp.hpp"
#include "se3.hpp"

namespace oracle {

// svn_ndt_impl.hpp:401-513
inline double update_derivatives_svn(double* g, double* H, const PointDeriv& d, const double* x_trans, const double* c_inv,
                                     double gauss_d1, double gauss_d2, bool compute_hessian, bool use_gn) {
  const float x4[3] = {static_cast<float>(x_trans[0]), static_cast<float>(x_trans[1]), static_cast<float>(x_trans[2])};
  float C[3][3];
  for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) C[r][c] = static_cast<float>(c_inv[3 * r + c]);
  double Cx[3];
  for (int r = 0; r < 3; ++r) Cx[r] = (c_inv[3 * r] * x_trans[0] + c_inv[3 * r + 1] * x_trans[1]) + c_inv[3 * r + 2] * x_trans[2];
  double mahal_sq = (x_trans[0] * Cx[0] + x_trans[1] * Cx[1]) + x_trans[2] * Cx[2];
  if (!std::isfinite(mahal_sq) || mahal_sq < -1e-9) return 0.0;
  if (mahal_sq < 0.0) mahal_sq = 0.0;
  const double exponent_arg = gauss_d2 * mahal_sq * 0.5;
  if (exponent_arg > 50.0) return 0.0;
  const double exp_term = std::exp(-exponent_arg);
  const double score_inc = -gauss_d1 * exp_term;
  const double factor = gauss_d1 * gauss_d2 * exp_term;
  if (!std::isfinite(factor) || std::fabs(factor) < 1e-15) return score_inc;
  float CJ[3][6];
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 6; ++c) { float s = C[r][0] * d.Jp[0][c]; s = s + C[r][1] * d.Jp[1][c]; s = s + C[r][2] * d.Jp[2][c]; CJ[r][c] = s; }
  float gc[6];
  for (int c = 0; c < 6; ++c) { float s = x4[0] * CJ[0][c]; s = s + x4[1] * CJ[1][c]; s = s + x4[2] * CJ[2][c]; gc[c] = s; }
  double grad_inc[6]; bool finite = true;
  for (int c = 0; c < 6; ++c) { grad_inc[c] = factor * static_cast<double>(gc[c]); if (!std::isfinite(grad_inc[c])) finite = false; }
  if (finite) for (int c = 0; c < 6; ++c) g[c] += grad_inc[c];
  if (compute_hessian) {
    double hc[36];
    for (int r = 0; r < 6; ++r)
      for (int c = 0; c < 6; ++c) { float s = d.Jp[0][r] * CJ[0][c]; s = s + d.Jp[1][r] * CJ[1][c]; s = s + d.Jp[2][r] * CJ[2][c]; hc[6 * r + c] = static_cast<double>(s); }
    if (!use_gn) {
      float xC[3];
      for (int c = 0; c < 3; ++c) { float s = x4[0] * C[0][c]; s = s + x4[1] * C[1][c]; s = s + x4[2] * C[2][c]; xC[c] = s; }
      double term3[36];
      for (int i = 0; i < 6; ++i)
        for (int j = i; j < 6; ++j) { float s = xC[0] * d.Hp[4 * i][j]; s = s + xC[1] * d.Hp[4 * i + 1][j]; s = s + xC[2] * d.Hp[4 * i + 2][j]; term3[6 * i + j] = static_cast<double>(s); }
      for (int i = 0; i < 6; ++i) for (int j = 0; j < i; ++j) term3[6 * i + j] = term3[6 * j + i];
      for (int i = 0; i < 6; ++i)
        for (int j = 0; j < 6; ++j) {
          const double term1 = -gauss_d2 * (static_cast<double>(gc[i]) * static_cast<double>(gc[j]));
          hc[6 * i + j] = (term1 + hc[6 * i + j]) + term3[6 * i + j];
        }
    }
    bool hf = true;
    for (int i = 0; i < 36; ++i) { hc[i] = hc[i] * factor; if (!std::isfinite(hc[i])) hf = false; }
    if (hf) for (int i = 0; i < 36; ++i) H[i] += hc[i];
  }
  return score_inc;
}

struct SvnResult {
  Pose3 final_pose;
  double final_covariance[36];
  bool converged = false;
  int iterations = 0;
};

struct SvnNdt {
  VoxelGrid grid;
  float resolution = 1.0f;
  double outlier_ratio = 0.55;
  int search_method = DIRECT7;
  int num_threads = 1;
  int K = 30, max_iter = 50;
  double kernel_h = 1.0, step_size = 1.0, stop_thresh = 1e-4;
  bool use_gn = true;
  double gauss_d1 = 0, gauss_d2 = 0, gauss_d3 = 0;
  std::vector<float> source;

  SvnNdt() { grid.flavour = FLAVOUR_SVN; grid.min_points_per_voxel = 6; update_constants(); }

  void update_constants() {  // svn_ndt_impl.hpp:80-131
    if (resolution <= 1e-6f) { gauss_d1 = 1.0; gauss_d2 = 1.0; gauss_d3 = 0.0; return; }
    double c1 = 10.0 * (1.0 - outlier_ratio);
    double c2 = outlier_ratio / std::pow(static_cast<double>(resolution), 3);
    const double eps = 1e-9;
    if (c1 <= eps) c1 = eps;
    if (c2 <= eps) c2 = eps;
    gauss_d3 = -std::log(c2);
    gauss_d1 = -std::log(c1 + c2) - gauss_d3;
    if (std::fabs(gauss_d1) < eps) gauss_d2 = 1.0;
    else {
      const double inner = c1 * std::exp(-0.5) + c2;
      if (inner <= eps) gauss_d2 = 1.0;
      else {
        const double outer = (-std::log(inner) - gauss_d3) / gauss_d1;
        gauss_d2 = (outer <= eps) ? 1.0 : -2.0 * std::log(outer);
      }
    }
    if (!std::isfinite(gauss_d1) || !std::isfinite(gauss_d2) || !std::isfinite(gauss_d3)) { gauss_d1 = 1.0; gauss_d2 = 1.0; gauss_d3 = 0.0; }
  }
  int set_target(const float* xyz, size_t n, size_t sf) {
    grid.set_leaf_size(resolution);
    const int rc = grid.build(xyz, n, sf, true);
    grid.prepare_radius_search();   // filter(search_method_ == KDTREE), svn_ndt_impl.hpp:149-150
    update_constants();
    return rc;
  }
  void set_source(const float* xyz, size_t n, size_t sf) {
    source.resize(3 * n);
    for (size_t i = 0; i < n; ++i) for (int a = 0; a < 3; ++a) source[3 * i + a] = xyz[i * sf + a];
  }

  // svn_ndt_impl.hpp:518-668.  trans: transformed cloud; p: (x,y,z,r,p,y).
  double particle_derivatives(double* g, double* H, const std::vector<float>& trans, const double* p) const {
    const size_t N = trans.size() / 3;
    std::memset(g, 0, 48); std::memset(H, 0, 288);
    if (N == 0) return 0.0;
    AngleTables tab;
    compute_angle_derivatives(p, 1e-7, tab);
    std::vector<double> scores(N, 0.0), grads(6 * N, 0.0), hess(36 * N, 0.0);
    const int n_probe = (search_method == DIRECT7) ? 7 : 1;
    for (size_t idx = 0; idx < N; ++idx) {
      const float* xt = &trans[3 * idx];
      if (!std::isfinite(xt[0]) || !std::isfinite(xt[1]) || !std::isfinite(xt[2])) continue;
      const Leaf* nb[64];
      int nn = 0;
      if (search_method == KDTREE) {   // svn_ndt_impl.hpp:576-579: radiusSearch(x_trans, resolution), results sorted by distance
        std::vector<std::pair<float, const Leaf*>> hits;
        grid.radius_search(xt, static_cast<double>(grid.leaf_size), false, hits);
        std::stable_sort(hits.begin(), hits.end(), [](const std::pair<float, const Leaf*>& a, const std::pair<float, const Leaf*>& b) { return a.first < b.first; });
        for (auto& hh : hits) if (nn < 64) nb[nn++] = hh.second;
      } else {
        nn = grid.neighbours(xt[0], xt[1], xt[2], n_probe, nb);
      }
      if (nn == 0) continue;
      const double x[3] = {source[3 * idx], source[3 * idx + 1], source[3 * idx + 2]};
      PointDeriv pd; init_point_deriv(pd);
      compute_point_derivatives(tab, x, pd, true);
      double sc = 0;
      for (int k = 0; k < nn; ++k) {
        const double xr[3] = {static_cast<double>(xt[0]) - nb[k]->mean[0], static_cast<double>(xt[1]) - nb[k]->mean[1], static_cast<double>(xt[2]) - nb[k]->mean[2]};
        sc += update_derivatives_svn(&grads[6 * idx], &hess[36 * idx], pd, xr, nb[k]->icov, gauss_d1, gauss_d2, true, use_gn);
      }
      scores[idx] = sc;
    }
    double total = 0;
    for (size_t i = 0; i < N; ++i) {
      total += scores[i];
      for (int c = 0; c < 6; ++c) g[c] += grads[6 * i + c];
      for (int c = 0; c < 36; ++c) H[c] += hess[36 * i + c];
    }
    for (int i = 0; i < 6; ++i) H[7 * i] += 1e-6;
    bool gf = true, hf = true;
    for (int i = 0; i < 6; ++i) if (!std::isfinite(g[i])) gf = false;
    for (int i = 0; i < 36; ++i) if (!std::isfinite(H[i])) hf = false;
    if (!gf) std::memset(g, 0, 48);
    if (!hf) for (int i = 0; i < 36; ++i) H[i] = (i % 7 == 0) ? 1.0 : 0.0;
    return total;
  }

  // Stage 1 for one particle (svn_ndt_impl.hpp:759-781)
  double stage1(const Pose3& P, double* g, double* H) const {
    float T[16];
    for (int r = 0; r < 3; ++r) { for (int c = 0; c < 3; ++c) T[4 * r + c] = static_cast<float>(P.R[3 * r + c]); T[4 * r + 3] = static_cast<float>(P.t[r]); }
    T[12] = T[13] = T[14] = 0; T[15] = 1;
    const size_t N = source.size() / 3;
    std::vector<float> trans(3 * N);
    for (size_t i = 0; i < N; ++i) transform_point_f(T, &source[3 * i], &trans[3 * i]);
    double p[6], rpy[3];
    rot_rpy(P.R, rpy);
    p[0] = P.t[0]; p[1] = P.t[1]; p[2] = P.t[2]; p[3] = rpy[0]; p[4] = rpy[1]; p[5] = rpy[2];
    return particle_derivatives(g, H, trans, p);
  }

  // Stage 2 (svn_ndt_impl.hpp:789-839): grads/hess in NDT order -> updates in GTSAM order.
  void stage2(const std::vector<Pose3>& P, const std::vector<double>& grads, const std::vector<double>& hess, std::vector<double>& updates) const {
    const int Kn = static_cast<int>(P.size());
    updates.assign(6 * Kn, 0.0);
    static const int perm[6] = {3, 4, 5, 0, 1, 2};
#pragma omp parallel for num_threads(num_threads) schedule(dynamic)
    for (int k = 0; k < Kn; ++k) {
      double phi[6] = {0, 0, 0, 0, 0, 0}, Ht[36];
      std::memset(Ht, 0, sizeof(Ht));
      for (int l = 0; l < Kn; ++l) {
        double k_val, k_grad[6];
        if (kernel_h <= 1e-12) { k_val = (l == k) ? 1.0 : 0.0; std::memset(k_grad, 0, sizeof(k_grad)); }
        else {
          double xi[6]; pose_logmap(pose_between(P[l], P[k]), xi);
          double sq = 0; for (int i = 0; i < 6; ++i) sq += xi[i] * xi[i];
          k_val = std::exp(-sq / kernel_h);
          const double f = k_val * (-2.0 / kernel_h);
          for (int i = 0; i < 6; ++i) k_grad[i] = f * xi[i];
        }
        bool fin = std::isfinite(k_val);
        for (int i = 0; i < 6; ++i) if (!std::isfinite(k_grad[i])) fin = false;
        if (!fin) continue;
        double gl[6], hl[36]; bool gfin = true, hfin = true;
        for (int i = 0; i < 6; ++i) { gl[i] = grads[6 * l + perm[i]]; if (!std::isfinite(gl[i])) gfin = false; }
        for (int i = 0; i < 6; ++i) for (int j = 0; j < 6; ++j) { hl[6 * i + j] = hess[36 * l + 6 * perm[i] + perm[j]]; if (!std::isfinite(hl[6 * i + j])) hfin = false; }
        if (gfin) for (int i = 0; i < 6; ++i) phi[i] += k_val * gl[i];
        for (int i = 0; i < 6; ++i) phi[i] += k_grad[i];
        if (hfin) { const double kk = k_val * k_val; for (int i = 0; i < 36; ++i) Ht[i] += kk * hl[i]; }
        for (int i = 0; i < 6; ++i) for (int j = 0; j < 6; ++j) Ht[6 * i + j] += k_grad[i] * k_grad[j];
      }
      for (int i = 0; i < 6; ++i) phi[i] /= static_cast<double>(Kn);
      for (int i = 0; i < 36; ++i) Ht[i] /= static_cast<double>(Kn);
      for (int i = 0; i < 6; ++i) Ht[7 * i] += 1e-6;
      bool hfin = true;
      for (int i = 0; i < 36; ++i) if (!std::isfinite(Ht[i])) hfin = false;
      double nphi[6], upd[6];
      for (int i = 0; i < 6; ++i) nphi[i] = -phi[i];
      const bool ok = ldlt6_solve(Ht, nphi, upd);
      bool ufin = true;
      for (int i = 0; i < 6; ++i) if (!std::isfinite(upd[i])) ufin = false;
      if (ok && hfin && ufin) for (int i = 0; i < 6; ++i) updates[6 * k + i] = upd[i];
    }
  }

  // One full iteration (stages 1-3).  Returns per-particle scores/grads/hess/updates for parity.
  void step(std::vector<Pose3>& P, std::vector<double>& scores, std::vector<double>& grads, std::vector<double>& hess, std::vector<double>& updates) const {
    const int Kn = static_cast<int>(P.size());
    scores.assign(Kn, 0.0); grads.assign(6 * Kn, 0.0); hess.assign(36 * Kn, 0.0);
#pragma omp parallel for num_threads(num_threads) schedule(dynamic, 1)
    for (int k = 0; k < Kn; ++k) {
      scores[k] = stage1(P[k], &grads[6 * k], &hess[36 * k]);
      bool gf = true, hf = true;
      for (int i = 0; i < 6; ++i) if (!std::isfinite(grads[6 * k + i])) gf = false;
      for (int i = 0; i < 36; ++i) if (!std::isfinite(hess[36 * k + i])) hf = false;
      if (!gf) for (int i = 0; i < 6; ++i) grads[6 * k + i] = 0;
      if (!hf) for (int i = 0; i < 36; ++i) hess[36 * k + i] = (i % 7 == 0) ? 1.0 : 0.0;
    }
    stage2(P, grads, hess, updates);
    for (int k = 0; k < Kn; ++k) {  // stage 3 (:848-855)
      double su[6]; bool fin = true;
      for (int i = 0; i < 6; ++i) { su[i] = step_size * updates[6 * k + i]; if (!std::isfinite(su[i])) fin = false; }
      if (!fin) continue;
      P[k] = pose_retract(P[k], su);
    }
  }

  static void sample_particles(const Pose3& prior, int Kn, unsigned long long seed, std::vector<Pose3>& P) {
    static const double sig[6] = {0.01, 0.01, 0.02, 0.05, 0.05, 0.05};  // svn_ndt_impl.hpp:710 (rot, trans)
    std::mt19937_64 gen(seed);
    P.resize(Kn);
    for (int k = 0; k < Kn; ++k) {
      double xi[6];
      for (int i = 0; i < 6; ++i) { std::normal_distribution<double> d(0.0, sig[i]); xi[i] = d(gen); }
      P[k] = pose_retract(prior, xi);
    }
  }

  // svn_ndt_impl.hpp:675-964.  `P` holds the initial particles on entry and the final ones on exit.
  SvnResult align(const Pose3& prior, std::vector<Pose3>& P) const {
    SvnResult res;
    const int Kn = static_cast<int>(P.size());
    res.final_pose = prior;
    for (int i = 0; i < 36; ++i) res.final_covariance[i] = (i % 7 == 0) ? 1.0 : 0.0;
    if (!grid.built || grid.leaves_robin.size() == 0 || source.empty() || Kn <= 0) return res;
    Pose3 mean_cur = prior, mean_last;
    std::vector<double> scores, grads, hess, updates;
    for (int iter = 0; iter < max_iter; ++iter) {
      mean_last = mean_cur;
      step(P, scores, grads, hess, updates);
      double mean_xi[6] = {0, 0, 0, 0, 0, 0};
      for (int k = 0; k < Kn; ++k) { double xi[6]; pose_logmap(pose_between(prior, P[k]), xi); for (int i = 0; i < 6; ++i) mean_xi[i] += xi[i]; }
      for (int i = 0; i < 6; ++i) mean_xi[i] /= static_cast<double>(Kn);
      mean_cur = pose_retract(prior, mean_xi);
      res.iterations = iter + 1;
      double dxi[6]; pose_logmap(pose_between(mean_last, mean_cur), dxi);
      double nrm = 0; for (int i = 0; i < 6; ++i) nrm += dxi[i] * dxi[i];
      if (std::sqrt(nrm) < stop_thresh) { res.converged = true; break; }
    }
    res.final_pose = mean_cur;
    double* cov = res.final_covariance;
    if (Kn > 1) {
      std::memset(cov, 0, 288);
      std::vector<double> tv(6 * Kn);
      double m[6] = {0, 0, 0, 0, 0, 0};
      for (int k = 0; k < Kn; ++k) { pose_logmap(pose_between(res.final_pose, P[k]), &tv[6 * k]); for (int i = 0; i < 6; ++i) m[i] += tv[6 * k + i]; }
      for (int i = 0; i < 6; ++i) m[i] /= static_cast<double>(Kn);
      for (int k = 0; k < Kn; ++k) {
        double d[6]; for (int i = 0; i < 6; ++i) d[i] = tv[6 * k + i] - m[i];
        for (int i = 0; i < 6; ++i) for (int j = 0; j < 6; ++j) cov[6 * i + j] += d[i] * d[j];
      }
      for (int i = 0; i < 36; ++i) cov[i] /= static_cast<double>(Kn - 1);
    } else {
      static const double sig[6] = {0.01, 0.01, 0.02, 0.05, 0.05, 0.05};
      std::memset(cov, 0, 288);
      for (int i = 0; i < 6; ++i) cov[7 * i] = 1e-6 * (sig[i] * sig[i]);
    }
    double ev[6], V[36];
    eig6_sym(cov, ev, V);
    bool need = false;
    for (int i = 0; i < 6; ++i) if (ev[i] < 1e-9) { ev[i] = 1e-9; need = true; }
    if (need) {
      for (int i = 0; i < 6; ++i) for (int j = 0; j < 6; ++j) { double s = 0; for (int k = 0; k < 6; ++k) s += (V[6 * i + k] * ev[k]) * V[6 * j + k]; cov[6 * i + j] = s; }
    }
    return res;
  }
};

}  // namespace oracle
