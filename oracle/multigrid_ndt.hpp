// ORACLE — TEST INFRASTRUCTURE ONLY (see small_linalg.hpp header).
//
// CPU restatement of pclomp::MultiGridNormalDistributionsTransform (TIER IV's dynamic-map variant of ndt_omp):
//   extern/ndt_omp/include/multigrid_pclomp/multigrid_ndt_omp.h:137-165 (setInputTarget / addTarget / removeTarget /
//     createVoxelKdtree), :255-276 (auxiliary functions)
//   extern/ndt_omp/include/multigrid_pclomp/multigrid_ndt_omp_impl.hpp:180-207 (ctor defaults), :210-323
//     (computeTransformation), :338-494 (computeDerivatives, FP64), :498-586 (computeAngleDerivatives), :590-636
//     (computePointDerivatives, FP64), :640-697 (updateDerivatives, FP64), :701-801 (computeHessian / updateHessian),
//     :805-1052 (More-Thuente), :1054-1145 (score helpers)
//   extern/ndt_omp/include/multigrid_pclomp/multi_voxel_grid_covariance_omp_impl.hpp:157-176 (setInputCloudAndFilter),
//     :179-190 (removeCloud), :193-233 (createKdtree), :236-258 (radiusSearch), :291-362 (applyFilter), :365-439
//     (getLeafID / updateLeaf / computeLeafParams).
//
// Differences from pclomp::NormalDistributionsTransform that this file carries:
//   * the target is a SET of voxel grids keyed by a string id, each with its own bounding box; a grid keeps only the
//     leaves with >= min_points points (leaves flagged -1 by the eigenvalue test stay, with icov = 0);
//   * the only neighbour search is a radius search (radius = resolution) over the leaves' float centroids
//     (float running sum / float n, NOT float(mean)), all grids in one kd-tree built by createVoxelKdtree();
//   * score, gradient and Hessian pair maths are FP64 end to end (the float SSE tables of pclomp are gone);
//   * the loop ends when nr_iterations >= max_iterations (pclomp: > max_iterations, tested before the increment) and
//     hessian_ is stored after the loop.
// Evaluation order: every dot / matrix product k ascending, left-to-right (Eigen's vectorised 4-lane reductions may
// associate differently: O(1 ulp) in FP64, far below the 1e-9 parity tolerance).  Per-thread accumulators are summed
// thread 0..T-1 (multigrid_ndt_omp_impl.hpp:437-444); with num_threads = 1 this is the reference's serial order.
#pragma once
#include <string>
#include "ndt_pclomp_align.hpp"

namespace oracle {

struct MultigridNdt {
  NdtParams params;              // search_method is ignored: KDTREE only (multigrid_ndt_omp_impl.hpp:197)
  double outlier_ratio = 0.55;
  int min_points_per_voxel = 6;  // multi_voxel_grid_covariance_omp.h:268
  double min_covar_eigvalue_mult = 0.01;
  std::map<std::string, VoxelGrid> grids;   // sid_to_iid_ + grid_list_; createKdtree walks it in id order (:203-213)
  std::vector<const VoxelGrid*> tree;       // the grids the current kd-tree was built over
  bool tree_dirty = false;
  std::vector<float> source;
  bool has_reg_pose = false;
  float reg_translation[3] = {0, 0, 0};
  double gauss_d1 = 0, gauss_d2 = 0, gauss_d3 = 0;
  double nvtl = 0;
  double stored_hessian[36];
  double jd[8][3], hd[15][3];

  MultigridNdt() { update_gauss(); std::memset(stored_hessian, 0, sizeof(stored_hessian)); }

  void update_gauss() {  // multigrid_ndt_omp_impl.hpp:218-222
    const double c1 = 10 * (1 - outlier_ratio);
    const double c2 = outlier_ratio / std::pow(static_cast<double>(params.resolution), 3);
    gauss_d3 = -std::log(c2);
    gauss_d1 = -std::log(c1 + c2) - gauss_d3;
    gauss_d2 = -2 * std::log((-std::log(c1 * std::exp(-0.5) + c2) - gauss_d3) / gauss_d1);
  }

  // addTarget (multigrid_ndt_omp.h:150-155) -> setInputCloudAndFilter -> applyFilter.  An existing id is replaced
  // (the old node is orphaned by sid_to_iid_[id] = new index and dropped at the next createKdtree).
  int add_target(const float* xyz, size_t n, size_t stride_floats, bool is_dense, const std::string& id) {
    grids.erase(id);
    VoxelGrid& g = grids[id];
    g.flavour = FLAVOUR_PCLOMP;
    g.min_points_per_voxel = min_points_per_voxel;
    g.min_covar_eigvalue_mult = min_covar_eigvalue_mult;
    g.set_leaf_size(params.resolution);
    const int rc = g.build(xyz, n, stride_floats, is_dense);
    // applyFilter :344-348 drops under-populated leaves; they are never searchable either way
    for (auto it = g.leaves_map.begin(); it != g.leaves_map.end();)
      if (it->second.nr_points >= 0 && it->second.nr_points < min_points_per_voxel) it = g.leaves_map.erase(it); else ++it;
    g.prepare_radius_search();
    tree_dirty = true;
    return rc;
  }
  void remove_target(const std::string& id) { if (grids.erase(id)) tree_dirty = true; }
  void create_kdtree() {
    tree.clear();
    for (auto& kv : grids) tree.push_back(&kv.second);
    tree_dirty = false;
  }
  size_t leaf_count() const { size_t c = 0; for (auto* g : tree) c += g->leaves_map.size(); return c; }

  void set_source(const float* xyz, size_t n, size_t stride_floats) {
    source.resize(3 * n);
    for (size_t i = 0; i < n; ++i) for (int a = 0; a < 3; ++a) source[3 * i + a] = xyz[i * stride_floats + a];
  }

  // radiusSearch (multi_voxel_grid_covariance_omp_impl.hpp:236-258): FLANN returns the hits sorted by distance
  void radius_search(const float* q, std::vector<std::pair<float, const Leaf*>>& hits) const {
    hits.clear();
    for (auto* g : tree) g->radius_search(q, static_cast<double>(params.resolution), true, hits);
    std::stable_sort(hits.begin(), hits.end(), [](const std::pair<float, const Leaf*>& a, const std::pair<float, const Leaf*>& b) { return a.first < b.first; });
  }

  // computePointDerivatives, FP64 (multigrid_ndt_omp_impl.hpp:590-636)
  struct PointDeriv64 { double Jp[3][6]; double Hp[18][6]; };
  void point_derivatives(const double* x, PointDeriv64& d, bool compute_hessian) const {
    auto dot = [&](const double* a) { double s = a[0] * x[0]; s = s + a[1] * x[1]; s = s + a[2] * x[2]; return s; };
    d.Jp[1][3] = dot(jd[0]); d.Jp[2][3] = dot(jd[1]);
    d.Jp[0][4] = dot(jd[2]); d.Jp[1][4] = dot(jd[3]); d.Jp[2][4] = dot(jd[4]);
    d.Jp[0][5] = dot(jd[5]); d.Jp[1][5] = dot(jd[6]); d.Jp[2][5] = dot(jd[7]);
    if (compute_hessian) {
      const double a[3] = {0, dot(hd[0]), dot(hd[1])}, b[3] = {0, dot(hd[2]), dot(hd[3])}, c[3] = {0, dot(hd[4]), dot(hd[5])};
      const double dd[3] = {dot(hd[6]), dot(hd[7]), dot(hd[8])}, e[3] = {dot(hd[9]), dot(hd[10]), dot(hd[11])}, f[3] = {dot(hd[12]), dot(hd[13]), dot(hd[14])};
      for (int k = 0; k < 3; ++k) {
        d.Hp[9 + k][3] = a[k]; d.Hp[12 + k][3] = b[k]; d.Hp[15 + k][3] = c[k];
        d.Hp[9 + k][4] = b[k]; d.Hp[12 + k][4] = dd[k]; d.Hp[15 + k][4] = e[k];
        d.Hp[9 + k][5] = c[k]; d.Hp[12 + k][5] = e[k]; d.Hp[15 + k][5] = f[k];
      }
    }
  }
  static void init_point_deriv(PointDeriv64& d) {
    std::memset(&d, 0, sizeof(d));
    d.Jp[0][0] = d.Jp[1][1] = d.Jp[2][2] = 1.0;
  }

  // updateDerivatives, FP64 (multigrid_ndt_omp_impl.hpp:640-697); the zero 4th lane of the reference is skipped
  double update_derivatives(double* g, double* H, const PointDeriv64& d, const double* x, const double* C, bool compute_hessian) const {
    double xC[3];
    for (int c = 0; c < 3; ++c) { double s = x[0] * C[c]; s = s + x[1] * C[3 + c]; s = s + x[2] * C[6 + c]; xC[c] = s; }
    double q = x[0] * xC[0]; q = q + x[1] * xC[1]; q = q + x[2] * xC[2];
    double e_x_cov_x = std::exp(-gauss_d2 * q * 0.5);
    const double score_inc = -gauss_d1 * e_x_cov_x;
    e_x_cov_x = gauss_d2 * e_x_cov_x;
    if (e_x_cov_x > 1 || e_x_cov_x < 0 || e_x_cov_x != e_x_cov_x) return 0;
    e_x_cov_x *= gauss_d1;
    double CJ[3][6];
    for (int r = 0; r < 3; ++r)
      for (int c = 0; c < 6; ++c) { double s = C[3 * r] * d.Jp[0][c]; s = s + C[3 * r + 1] * d.Jp[1][c]; s = s + C[3 * r + 2] * d.Jp[2][c]; CJ[r][c] = s; }
    double xCJ[6];
    for (int c = 0; c < 6; ++c) { double s = x[0] * CJ[0][c]; s = s + x[1] * CJ[1][c]; s = s + x[2] * CJ[2][c]; xCJ[c] = s; }
    for (int c = 0; c < 6; ++c) g[c] += e_x_cov_x * xCJ[c];
    if (compute_hessian) {
      double JCJ[6][6];
      for (int r = 0; r < 6; ++r)
        for (int c = 0; c < 6; ++c) { double s = d.Jp[0][r] * CJ[0][c]; s = s + d.Jp[1][r] * CJ[1][c]; s = s + d.Jp[2][r] * CJ[2][c]; JCJ[r][c] = s; }
      for (int i = 0; i < 6; ++i) {
        double xCH[6];
        for (int j = 0; j < 6; ++j) { double s = xC[0] * d.Hp[3 * i][j]; s = s + xC[1] * d.Hp[3 * i + 1][j]; s = s + xC[2] * d.Hp[3 * i + 2][j]; xCH[j] = s; }
        for (int j = 0; j < 6; ++j) H[6 * i + j] += e_x_cov_x * (-gauss_d2 * xCJ[i] * xCJ[j] + xCH[j] + JCJ[j][i]);
      }
    }
    return score_inc;
  }

  // computeDerivatives (multigrid_ndt_omp_impl.hpp:338-494).  trans = transformed cloud (3 floats per point).
  double compute_derivatives(double* g, double* H, const std::vector<float>& trans, const double* p, bool compute_hessian) {
    const size_t N = source.size() / 3;
    std::memset(g, 0, 6 * sizeof(double));
    std::memset(H, 0, 36 * sizeof(double));
    const int nt = std::max(1, params.num_threads);
    std::vector<double> scores(nt, 0.0), nearest(nt, 0.0), grads(6 * nt, 0.0), hess(36 * nt, 0.0);
    std::vector<size_t> found(nt, 0);
    std::vector<long> counts(nt, 0);
    angle_vectors_f64(p, jd, hd);  // computeAngleDerivatives :498-586 (FP64 tables, snap 10e-5)
#pragma omp parallel num_threads(nt)
    {
      const int tid = omp_get_thread_num();
      const size_t lo = N * tid / nt, hi = N * (tid + 1) / nt;  // contiguous chunks: thread t sees its points in order
      PointDeriv64 pd; init_point_deriv(pd);
      std::vector<std::pair<float, const Leaf*>> hits;
      for (size_t idx = lo; idx < hi; ++idx) {
        const float* xt = &trans[3 * idx];
        radius_search(xt, hits);
        if (hits.empty()) continue;
        const double x[3] = {source[3 * idx], source[3 * idx + 1], source[3 * idx + 2]};
        point_derivatives(x, pd, compute_hessian);
        double sum_score = 0, nearest_pt = 0;
        for (auto& h : hits) {
          const double xr[3] = {static_cast<double>(xt[0]) - h.second->mean[0], static_cast<double>(xt[1]) - h.second->mean[1],
                                static_cast<double>(xt[2]) - h.second->mean[2]};
          const double s = update_derivatives(&grads[6 * tid], &hess[36 * tid], pd, xr, h.second->icov, compute_hessian);
          sum_score += s;
          if (s > nearest_pt) nearest_pt = s;
        }
        ++found[tid];
        scores[tid] += sum_score; nearest[tid] += nearest_pt; counts[tid] += static_cast<long>(hits.size());
      }
    }
    double score = 0, nearest_sum = 0; size_t found_num = 0; long total_count = 0;
    for (int t = 0; t < nt; ++t) {
      score += scores[t]; nearest_sum += nearest[t]; found_num += found[t];
      for (int c = 0; c < 6; ++c) g[c] += grads[6 * t + c];
      for (int c = 0; c < 36; ++c) H[c] += hess[36 * t + c];
      total_count += counts[t];
    }
    if (has_reg_pose) {  // multigrid_ndt_omp_impl.hpp:446-480
      const float dx = reg_translation[0] - static_cast<float>(p[0]);
      const float dy = reg_translation[1] - static_cast<float>(p[1]);
      const float sin_yaw = static_cast<float>(std::sin(p[5])), cos_yaw = static_cast<float>(std::cos(p[5]));
      const float longitudinal = dy * sin_yaw + dx * cos_yaw;
      const float w = static_cast<float>(static_cast<int>(total_count));
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

  // computeHessian / updateHessian (multigrid_ndt_omp_impl.hpp:701-801): FP64, angle tables of the last computeDerivatives
  void compute_hessian(double* H, const std::vector<float>& trans) {
    std::memset(H, 0, 36 * sizeof(double));
    const size_t N = source.size() / 3;
    AngleTables unused{};
    std::vector<std::pair<float, const Leaf*>> hits;
    for (size_t idx = 0; idx < N; ++idx) {
      const float* xt = &trans[3 * idx];
      radius_search(xt, hits);
      for (auto& h : hits) {
        const double x[3] = {source[3 * idx], source[3 * idx + 1], source[3 * idx + 2]};
        const double xr[3] = {static_cast<double>(xt[0]) - h.second->mean[0], static_cast<double>(xt[1]) - h.second->mean[1],
                              static_cast<double>(xt[2]) - h.second->mean[2]};
        update_hessian_f64(H, unused, jd, hd, x, xr, h.second->icov, gauss_d1, gauss_d2);
      }
    }
  }

  void transform_cloud(const float* T, std::vector<float>& out) const {
    const size_t N = source.size() / 3;
    out.resize(3 * N);
    for (size_t i = 0; i < N; ++i) transform_point_f(T, &source[3 * i], &out[3 * i]);
  }

  // computeStepLengthMT (multigrid_ndt_omp_impl.hpp:893-1052): identical control flow to pclomp's
  double step_length_mt(const double* x, double* step_dir, double step_init, double step_max, double step_min, double& score, double* g,
                        double* H, std::vector<float>& trans, float* final_T, NdtResult& res) {
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
      compute_hessian(H, trans);
      push_trace(res, x_t, score, g, H, nvtl);  // extra trace row (parity hook, not in the reference)
    }
    return a_t;
  }

  // pcl::Registration::align + computeTransformation (multigrid_ndt_omp_impl.hpp:210-323)
  void align(const float* guess, NdtResult& res) {
    const size_t N = source.size() / 3;
    res = NdtResult();
    int nr_iterations = 0;
    bool converged = false;
    update_gauss();
    float final_T[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
    std::vector<float> trans(source);
    bool guess_is_identity = true;
    for (int i = 0; i < 16; ++i) if (guess[i] != final_T[i]) guess_is_identity = false;
    if (!guess_is_identity) { std::memcpy(final_T, guess, sizeof(final_T)); transform_cloud(guess, trans); }
    res.transformation_array.insert(res.transformation_array.end(), final_T, final_T + 16);
    double p[6], delta_p[6], g[6], H[36];
    float eul[3];
    euler_angles_012_f(final_T, eul);
    p[0] = final_T[3]; p[1] = final_T[7]; p[2] = final_T[11]; p[3] = eul[0]; p[4] = eul[1]; p[5] = eul[2];
    double score = compute_derivatives(g, H, trans, p, true);
    push_trace(res, p, score, g, H, nvtl);
    bool early_return = false;
    while (!converged) {
      double neg_g[6];
      for (int i = 0; i < 6; ++i) neg_g[i] = -g[i];
      svd6_solve(H, neg_g, delta_p);
      double nrm2 = 0;
      for (int i = 0; i < 6; ++i) nrm2 += delta_p[i] * delta_p[i];
      double delta_p_norm = std::sqrt(nrm2);
      if (delta_p_norm == 0 || delta_p_norm != delta_p_norm) {  // :256-265: early return, hessian_ not stored
        converged = delta_p_norm == delta_p_norm;
        early_return = true;
        break;
      }
      for (int i = 0; i < 6; ++i) delta_p[i] = delta_p[i] / delta_p_norm;
      delta_p_norm = step_length_mt(p, delta_p, delta_p_norm, params.step_size, params.trans_epsilon / 2, score, g, H, trans, final_T, res);
      for (int i = 0; i < 6; ++i) delta_p[i] *= delta_p_norm;
      res.transformation_array.insert(res.transformation_array.end(), final_T, final_T + 16);
      for (int i = 0; i < 6; ++i) p[i] = p[i] + delta_p[i];
      nr_iterations++;
      if (nr_iterations >= params.max_iterations || (nr_iterations && (std::fabs(delta_p_norm) < params.trans_epsilon))) converged = true;  // :300-304
    }
    res.trans_probability = N ? score / static_cast<double>(N) : 0.0;
    if (!early_return) std::memcpy(stored_hessian, H, sizeof(stored_hessian));  // :316
    std::memcpy(res.final_transformation, final_T, sizeof(final_T));
    std::memcpy(res.hessian, stored_hessian, sizeof(stored_hessian));
    res.nvtl = nvtl;
    res.iterations = nr_iterations;
    res.converged = converged;
  }
};

}  // namespace oracle
