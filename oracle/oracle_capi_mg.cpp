// ORACLE — TEST INFRASTRUCTURE ONLY.  C entry points (ctypes) over the multigrid NDT restatement
// (multigrid_ndt.hpp).  Only tests/ may load this library.
#include "multigrid_ndt.hpp"

using namespace oracle;

extern "C" {

void* orc_mg_create() { return new MultigridNdt(); }
void orc_mg_destroy(void* h) { delete static_cast<MultigridNdt*>(h); }
void orc_mg_set_params(void* h, float resolution, double step_size, double trans_epsilon, int max_iterations, int num_threads, float reg_scale,
                       int use_line_search, double outlier_ratio, int min_pts, double eig_mult) {
  MultigridNdt* n = static_cast<MultigridNdt*>(h);
  n->params.resolution = resolution; n->params.step_size = step_size; n->params.trans_epsilon = trans_epsilon;
  n->params.max_iterations = max_iterations; n->params.search_method = KDTREE; n->params.num_threads = num_threads;
  n->params.regularization_scale_factor = reg_scale; n->params.use_line_search = use_line_search != 0;
  n->outlier_ratio = outlier_ratio; n->min_points_per_voxel = min_pts < 3 ? 3 : min_pts /* multi_voxel_grid_covariance_omp.h setMinPointPerVoxel */; n->min_covar_eigvalue_mult = eig_mult;
  n->update_gauss();
}
int orc_mg_add_target(void* h, const float* xyz, size_t n, size_t stride_floats, int is_dense, const char* id) {
  return static_cast<MultigridNdt*>(h)->add_target(xyz, n, stride_floats, is_dense != 0, id);
}
void orc_mg_remove_target(void* h, const char* id) { static_cast<MultigridNdt*>(h)->remove_target(id); }
void orc_mg_create_kdtree(void* h) { static_cast<MultigridNdt*>(h)->create_kdtree(); }
long orc_mg_leaf_count(void* h) { return static_cast<long>(static_cast<MultigridNdt*>(h)->leaf_count()); }
// Leaves of the current tree in tree order (grid id order, ascending linear index inside a grid).
long orc_mg_export(void* h, long cap, int* n_pts, double* mean, double* icov, float* centroid) {
  MultigridNdt* n = static_cast<MultigridNdt*>(h);
  long k = 0;
  for (auto* g : n->tree)
    for (auto& kv : g->leaves_map) {
      if (k < cap) {
        n_pts[k] = kv.second.nr_points;
        std::memcpy(mean + 3 * k, kv.second.mean, 24);
        std::memcpy(icov + 9 * k, kv.second.icov, 72);
        std::memcpy(centroid + 3 * k, kv.second.centroid, 12);
      }
      ++k;
    }
  return k;
}
// Radius search for one point: the hit leaves' means, nearest first.  Returns the count.
int orc_mg_radius_search(void* h, float x, float y, float z, int cap, double* means_out, float* dist_out) {
  MultigridNdt* n = static_cast<MultigridNdt*>(h);
  std::vector<std::pair<float, const Leaf*>> hits;
  const float q[3] = {x, y, z};
  n->radius_search(q, hits);
  int k = 0;
  for (auto& hh : hits) { if (k < cap) { std::memcpy(means_out + 3 * k, hh.second->mean, 24); dist_out[k] = hh.first; } ++k; }
  return k;
}
void orc_mg_set_source(void* h, const float* xyz, size_t n, size_t stride_floats) { static_cast<MultigridNdt*>(h)->set_source(xyz, n, stride_floats); }
void orc_mg_set_regularization_pose(void* h, const float* T) {
  MultigridNdt* n = static_cast<MultigridNdt*>(h);
  n->has_reg_pose = (T != nullptr);
  if (T) { n->reg_translation[0] = T[3]; n->reg_translation[1] = T[7]; n->reg_translation[2] = T[11]; }
}
// One evaluation at pose p, cloud transformed with T(p) as computeStepLengthMT does (:946-958).
double orc_mg_compute_derivatives(void* h, const double* p, int compute_hessian, double* g, double* H, double* nvtl) {
  MultigridNdt* n = static_cast<MultigridNdt*>(h);
  n->update_gauss();
  float T[16]; pose_to_matrix_f(p, T);
  std::vector<float> trans; n->transform_cloud(T, trans);
  const double s = n->compute_derivatives(g, H, trans, p, compute_hessian != 0);
  if (nvtl) *nvtl = n->nvtl;
  return s;
}
// calculateTransformationProbability + calculateNearestVoxelTransformationLikelihood (multigrid_ndt_omp_impl.hpp:1054-1145)
void orc_mg_calculate_scores(void* h, const float* T, double* tp, double* nvtl_out) {
  MultigridNdt* n = static_cast<MultigridNdt*>(h);
  n->update_gauss();
  std::vector<float> trans; n->transform_cloud(T, trans);
  const size_t N = trans.size() / 3;
  double score = 0, nearest = 0; size_t found = 0;
  std::vector<std::pair<float, const Leaf*>> hits;
  for (size_t i = 0; i < N; ++i) {
    const float* xt = &trans[3 * i];
    n->radius_search(xt, hits);
    if (hits.empty()) continue;
    double tmp = 0, best = 0;
    for (auto& hh : hits) {
      const double x[3] = {static_cast<double>(xt[0]) - hh.second->mean[0], static_cast<double>(xt[1]) - hh.second->mean[1], static_cast<double>(xt[2]) - hh.second->mean[2]};
      const double* C = hh.second->icov;
      double Cx[3];
      for (int r = 0; r < 3; ++r) Cx[r] = (C[3 * r] * x[0] + C[3 * r + 1] * x[1]) + C[3 * r + 2] * x[2];
      const double q = (x[0] * Cx[0] + x[1] * Cx[1]) + x[2] * Cx[2];
      const double s = -n->gauss_d1 * std::exp(-n->gauss_d2 * q / 2.0);
      tmp += s;
      if (s > best) best = s;
    }
    score += tmp; nearest += best; ++found;
  }
  *tp = N ? score / static_cast<double>(N) : 0.0;
  *nvtl_out = found ? nearest / static_cast<double>(found) : 0.0;
}
int orc_mg_align(void* h, const float* guess, float* final_T, double* hessian, double* trans_probability, double* nvtl, int* iterations,
                 int* converged, double* trace, int trace_cap, int* n_trace, float* tarray, int tarray_cap, int* n_tarray) {
  MultigridNdt* n = static_cast<MultigridNdt*>(h);
  if (n->tree_dirty) return -1;  // the reference would search a stale kd-tree (dangling leaf pointers after removeTarget)
  NdtResult r;
  n->align(guess, r);
  std::memcpy(final_T, r.final_transformation, sizeof(r.final_transformation));
  if (hessian) std::memcpy(hessian, r.hessian, sizeof(r.hessian));
  if (trans_probability) *trans_probability = r.trans_probability;
  if (nvtl) *nvtl = r.nvtl;
  *iterations = r.iterations; *converged = r.converged ? 1 : 0;
  const int nt = static_cast<int>(r.trace.size() / 50);
  if (n_trace) *n_trace = nt;
  if (trace) std::memcpy(trace, r.trace.data(), sizeof(double) * 50 * std::min(nt, trace_cap));
  const int na = static_cast<int>(r.transformation_array.size() / 16);
  if (n_tarray) *n_tarray = na;
  if (tarray) std::memcpy(tarray, r.transformation_array.data(), sizeof(float) * 16 * std::min(na, tarray_cap));
  return 0;
}

}  // extern "C"
