// ORACLE — TEST INFRASTRUCTURE ONLY.  C entry points (ctypes) over the CPU restatement.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// load this library.  Not bit-pinned against a reference binary (PCL/Eigen/GTSAM absent); pinned to the README fitness values and
// against the reference's own test criteria (tests/test_oracle_*.py).
#include <set>
#include <array>
#include <map>
#include <cstdio>
#include "ndt_pclomp_align.hpp"
#include "svn_ndt.hpp"

using namespace oracle;

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <limits>
#include <utility>
#include <vector>

extern "C" {

// ------------------------------------------------------------------ small-linalg probes (tests)
void orc_eig3_sym(const double* A, double* evals, double* evecs) { eig3_sym(A, evals, evecs); }
void orc_mat3_inverse(const double* A, double* inv) { mat3_inverse(A, inv); }
void orc_svd6_solve(const double* A, const double* b, double* x, double* S) { svd6_solve(A, b, x, S); }
int orc_ldlt6_solve(const double* A, const double* b, double* x) { return ldlt6_solve(A, b, x) ? 1 : 0; }
void orc_eig6_sym(const double* A, double* evals, double* evecs) { eig6_sym(A, evals, evecs); }
void orc_pose_to_matrix_f(const double* p, float* T) { pose_to_matrix_f(p, T); }
void orc_euler_angles_012_f(const float* T, float* e) { euler_angles_012_f(T, e); }

// ------------------------------------------------------------------ voxel grid (both flavours)
void* orc_grid_create(int flavour, float leaf, int min_pts, double eig_mult) {
  VoxelGrid* g = new VoxelGrid();
  g->flavour = flavour; g->set_leaf_size(leaf); g->min_points_per_voxel = min_pts < 3 ? 3 : min_pts /* setMinPointPerVoxel, voxel_grid_covariance_omp.h:214-224 */; g->min_covar_eigvalue_mult = eig_mult;
  return g;
}
void orc_grid_destroy(void* h) { delete static_cast<VoxelGrid*>(h); }
int orc_grid_build(void* h, const float* xyz, size_t n, size_t stride_floats, int is_dense) {
  return static_cast<VoxelGrid*>(h)->build(xyz, n, stride_floats, is_dense != 0);
}
void orc_grid_meta(void* h, int* min_b, int* max_b, int* div_b) {
  VoxelGrid* g = static_cast<VoxelGrid*>(h);
  for (int a = 0; a < 3; ++a) { min_b[a] = g->min_b[a]; max_b[a] = g->max_b[a]; div_b[a] = g->div_b[a]; }
}
size_t orc_grid_count(void* h) {
  VoxelGrid* g = static_cast<VoxelGrid*>(h);
  return g->flavour == FLAVOUR_PCLOMP ? g->leaves_map.size() : g->leaves_robin.size();
}
// Arrays sized by orc_grid_count: idx[V], n[V], mean[3V], cov[9V], icov[9V], evecs[9V], evals[3V]
void orc_grid_export(void* h, long long* idx, int* n, double* mean, double* cov, double* icov, double* evecs, double* evals) {
  VoxelGrid* g = static_cast<VoxelGrid*>(h);
  std::vector<size_t> keys; std::vector<Leaf> leaves;
  g->export_leaves(keys, leaves);
  for (size_t v = 0; v < keys.size(); ++v) {
    idx[v] = static_cast<long long>(keys[v]); n[v] = leaves[v].nr_points;
    std::memcpy(mean + 3 * v, leaves[v].mean, 24); std::memcpy(cov + 9 * v, leaves[v].cov, 72);
    std::memcpy(icov + 9 * v, leaves[v].icov, 72); std::memcpy(evecs + 9 * v, leaves[v].evecs, 72);
    std::memcpy(evals + 3 * v, leaves[v].evals, 24);
  }
}
// Neighbour probe for one point: writes linear-free identifiers (the leaf means) -- returns count.
int orc_grid_neighbours(void* h, float x, float y, float z, int n_probe, double* means_out) {
  VoxelGrid* g = static_cast<VoxelGrid*>(h);
  const Leaf* nb[7];
  const int nn = g->neighbours(x, y, z, n_probe, nb);
  for (int k = 0; k < nn; ++k) std::memcpy(means_out + 3 * k, nb[k]->mean, 24);
  return nn;
}

// ------------------------------------------------------------------ pclomp NDT
void* orc_ndt_create() { return new PclompNdt(); }
void orc_ndt_destroy(void* h) { delete static_cast<PclompNdt*>(h); }
void* orc_ndt_grid(void* h) { return &static_cast<PclompNdt*>(h)->grid; }
void orc_ndt_set_params(void* h, float resolution, double step_size, double trans_epsilon, int max_iterations, int search_method,
                        int num_threads, float reg_scale, int use_line_search, double outlier_ratio, int min_pts, double eig_mult) {
  PclompNdt* n = static_cast<PclompNdt*>(h);
  n->params.resolution = resolution; n->params.step_size = step_size; n->params.trans_epsilon = trans_epsilon;
  n->params.max_iterations = max_iterations; n->params.search_method = search_method; n->params.num_threads = num_threads;
  n->params.regularization_scale_factor = reg_scale; n->params.use_line_search = use_line_search != 0;
  n->outlier_ratio = outlier_ratio; n->grid.min_points_per_voxel = min_pts < 3 ? 3 : min_pts /* voxel_grid_covariance_omp.h:214-224 */; n->grid.min_covar_eigvalue_mult = eig_mult;
  n->update_gauss();
}
void orc_ndt_gauss(void* h, double* d) { PclompNdt* n = static_cast<PclompNdt*>(h); d[0] = n->gauss_d1; d[1] = n->gauss_d2; d[2] = n->gauss_d3; }
int orc_ndt_set_target(void* h, const float* xyz, size_t n, size_t stride_floats, int is_dense) {
  return static_cast<PclompNdt*>(h)->set_target(xyz, n, stride_floats, is_dense != 0);
}
void orc_ndt_set_source(void* h, const float* xyz, size_t n, size_t stride_floats) { static_cast<PclompNdt*>(h)->set_source(xyz, n, stride_floats); }
void orc_ndt_set_regularization_pose(void* h, const float* T) {
  PclompNdt* n = static_cast<PclompNdt*>(h);
  n->has_reg_pose = (T != nullptr);
  if (T) { n->reg_translation[0] = T[3]; n->reg_translation[1] = T[7]; n->reg_translation[2] = T[11]; }
}
// One evaluation at pose p: cloud transformed with T(p) as computeStepLengthMT does (:969-978).
double orc_ndt_compute_derivatives(void* h, const double* p, int compute_hessian, double* g, double* H, double* nvtl) {
  PclompNdt* n = static_cast<PclompNdt*>(h);
  n->update_gauss();
  float T[16]; pose_to_matrix_f(p, T);
  std::vector<float> trans; n->transform_cloud(T, trans);
  const double s = n->compute_derivatives(g, H, trans, p, compute_hessian != 0);
  if (nvtl) *nvtl = n->nvtl;
  return s;
}
// Same but with an explicit float 4x4 for the cloud (the initial-guess evaluation, :101-135).
double orc_ndt_compute_derivatives_T(void* h, const float* T, const double* p, int compute_hessian, double* g, double* H, double* nvtl) {
  PclompNdt* n = static_cast<PclompNdt*>(h);
  n->update_gauss();
  std::vector<float> trans; n->transform_cloud(T, trans);
  const double s = n->compute_derivatives(g, H, trans, p, compute_hessian != 0);
  if (nvtl) *nvtl = n->nvtl;
  return s;
}
// calculateTransformationProbability + calculateNearestVoxelTransformationLikelihood
// (ndt_omp_impl.hpp:1169-1281): serial, FP64 exp.
void orc_ndt_calculate_scores(void* h, const float* T, double* tp, double* nvtl_out) {
  PclompNdt* n = static_cast<PclompNdt*>(h);
  n->update_gauss();
  std::vector<float> trans; n->transform_cloud(T, trans);
  const size_t N = trans.size() / 3;
  double score = 0, nearest = 0; size_t found = 0;
  for (size_t i = 0; i < N; ++i) {
    const float* xt = &trans[3 * i];
    const Leaf* nb[PclompNdt::MAX_NB];
    const int nn = n->find_neighbours(xt, nb);
    double best = 0;
    for (int k = 0; k < nn; ++k) {
      const double x[3] = {static_cast<double>(xt[0]) - nb[k]->mean[0], static_cast<double>(xt[1]) - nb[k]->mean[1], static_cast<double>(xt[2]) - nb[k]->mean[2]};
      const double* C = nb[k]->icov;
      double Cx[3];
      for (int r = 0; r < 3; ++r) Cx[r] = (C[3 * r] * x[0] + C[3 * r + 1] * x[1]) + C[3 * r + 2] * x[2];
      const double q = (x[0] * Cx[0] + x[1] * Cx[1]) + x[2] * Cx[2];
      const double s = -n->gauss_d1 * std::exp(-n->gauss_d2 * q / 2);
      score += s;
      if (s > best) best = s;
    }
    if (nn > 0) { ++found; nearest += best; }
  }
  *tp = N ? score / static_cast<double>(N) : 0.0;
  *nvtl_out = found ? nearest / static_cast<double>(found) : 0.0;
}
// calculateScore (ndt_omp_impl.hpp:1083-1167) from cleared maps (cleanScores, ndt_omp.h:236-240): returns score / N; getScoreMap()
// = per-voxel mean score keyed by getLeafIndex(cell mean) (voxel_grid_covariance_omp.h:340-349); getEmptyVoxels() = the voxel ids
// floor(x / resolution) of the points without neighbours, ordered by EigenCmp (ndt_omp.h:57-65).  nomap_points_num_ is keyed by an
// uninitialised variable in the reference (:1124-1127) and is not restated.
double orc_ndt_calculate_score(void* h, const float* T, size_t map_cap, long long* vox_idx, double* vox_mean_score, size_t* n_vox,
                               size_t empty_cap, int* empty_ijk, size_t* n_empty) {
  PclompNdt* n = static_cast<PclompNdt*>(h);
  n->update_gauss();
  std::vector<float> trans; n->transform_cloud(T, trans);
  const size_t N = trans.size() / 3;
  double score = 0;
  std::map<size_t, double> score_map; std::map<size_t, size_t> num;
  std::set<std::array<int, 3>> empty;
  const VoxelGrid& g = n->grid;
  for (size_t i = 0; i < N; ++i) {
    const float* xt = &trans[3 * i];
    const Leaf* nb[PclompNdt::MAX_NB];
    const int nn = n->find_neighbours(xt, nb);
    if (nn == 0) {
      std::array<int, 3> v;
      for (int a = 0; a < 3; ++a) v[a] = static_cast<int>(std::floor(xt[a] / n->params.resolution));
      empty.insert(v);
      continue;
    }
    for (int k = 0; k < nn; ++k) {
      const double x[3] = {static_cast<double>(xt[0]) - nb[k]->mean[0], static_cast<double>(xt[1]) - nb[k]->mean[1], static_cast<double>(xt[2]) - nb[k]->mean[2]};
      const double* C = nb[k]->icov;
      double Cx[3];
      for (int r = 0; r < 3; ++r) Cx[r] = (C[3 * r] * x[0] + C[3 * r + 1] * x[1]) + C[3 * r + 2] * x[2];
      const double q = (x[0] * Cx[0] + x[1] * Cx[1]) + x[2] * Cx[2];
      const double s = -n->gauss_d1 * std::exp(-n->gauss_d2 * q / 2);
      score += s;
      int ijk[3];
      for (int a = 0; a < 3; ++a) ijk[a] = static_cast<int>(std::floor(nb[k]->mean[a] * static_cast<double>(g.inverse_leaf_size)) - g.min_b[a]);
      const size_t vi = static_cast<size_t>(ijk[0] * g.divb_mul[0] + ijk[1] * g.divb_mul[1] + ijk[2] * g.divb_mul[2]);
      score_map[vi] += s; num[vi] += 1;
    }
  }
  size_t c = 0;
  for (auto& kv : score_map) { if (c < map_cap) { vox_idx[c] = static_cast<long long>(kv.first); vox_mean_score[c] = kv.second / static_cast<double>(num[kv.first]); } ++c; }
  if (n_vox) *n_vox = c;
  c = 0;
  for (auto& v : empty) { if (c < empty_cap) { empty_ijk[3 * c] = v[0]; empty_ijk[3 * c + 1] = v[1]; empty_ijk[3 * c + 2] = v[2]; } ++c; }
  if (n_empty) *n_empty = c;
  return N ? score / static_cast<double>(N) : 0.0;
}
// extractNdtData + writeNdtDataToFiles (include/pipeline.hpp:163-274): ndt_ellipsoids.txt = mean, evecs row-major, evals, point count;
// ndt_voxels.txt = getLeafCenter(index) (voxel_grid_covariance_omp.h:352-367), resolution; std::fixed << std::setprecision(8).
// The reference compares getPointCount() (int) with a size_t min_points, so leaves flagged nr_points_ = -1 pass the test and are
// written with the count static_cast<size_t>(-1).
long orc_ndt_write_files(void* h, const char* ellipsoid_path, const char* voxel_path) {
  PclompNdt* n = static_cast<PclompNdt*>(h);
  const VoxelGrid& g = n->grid;
  FILE* fe = std::fopen(ellipsoid_path, "w"); FILE* fv = std::fopen(voxel_path, "w");
  if (!fe || !fv) { if (fe) std::fclose(fe); if (fv) std::fclose(fv); return -1; }
  long written = 0;
  const size_t min_points = static_cast<size_t>(g.min_points_per_voxel);
  for (auto& kv : g.leaves_map) {
    const Leaf& l = kv.second;
    if (!(static_cast<size_t>(l.nr_points) >= min_points)) continue;
    std::fprintf(fe, "%.8f %.8f %.8f ", l.mean[0], l.mean[1], l.mean[2]);
    for (int i = 0; i < 9; ++i) std::fprintf(fe, "%.8f ", l.evecs[i]);
    std::fprintf(fe, "%.8f %.8f %.8f %zu\n", l.evals[0], l.evals[1], l.evals[2], static_cast<size_t>(l.nr_points));
    const size_t index = kv.first;
    const int ijk2 = static_cast<int>(index / g.divb_mul[2]);
    const int ijk1 = static_cast<int>((index % g.divb_mul[2]) / g.divb_mul[1]);
    const int ijk0 = static_cast<int>(((index % g.divb_mul[2]) % g.divb_mul[1]) / g.divb_mul[0]);
    const double inv = static_cast<double>(g.inverse_leaf_size);
    std::fprintf(fv, "%.8f %.8f %.8f %.8f\n", (static_cast<double>(ijk0 + g.min_b[0]) + 0.5) / inv, (static_cast<double>(ijk1 + g.min_b[1]) + 0.5) / inv,
                 (static_cast<double>(ijk2 + g.min_b[2]) + 0.5) / inv, static_cast<double>(n->params.resolution));
    ++written;
  }
  std::fclose(fe); std::fclose(fv);
  return written;
}
// trace: 50 doubles per evaluation (p6, score, g6, H36, nvtl); tarray: 16 floats per entry.
int orc_ndt_align(void* h, const float* guess, float* final_T, double* hessian, double* trans_probability, double* nvtl, int* iterations,
                  int* converged, double* trace, int trace_cap, int* n_trace, float* tarray, int tarray_cap, int* n_tarray) {
  PclompNdt* n = static_cast<PclompNdt*>(h);
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

// Sliding-window target assembly (run/pipeline_ligo_tc.cpp:519-526): for every keyframe, in order,
//   pcl::transformPointCloud(*archive.points, *cur, Tb2m.matrix());   *target += *cur;
// pcl::transformPointCloud<PointT, double> is PCL code (>= 1.14, common/impl/transforms.hpp, not vendored in the
// reference): generic Transformer<double>::se3 -- out[r] = float(T(r,0)*x + T(r,1)*y + T(r,2)*z + T(r,3)) evaluated
// left to right in double, out[3] = 1 (the reference is built without -mavx, so no AVX specialisation).
// clouds: n_keyframes pointers to float arrays with `stride` floats per point; poses: n x 16 row-major doubles;
// out: packed x,y,z,w floats for sum(n_points) points.
void orc_assemble_window(int n_keyframes, const float* const* clouds, const size_t* n_points, size_t stride, const double* poses16, float* out) {
  size_t o = 0;
  for (int k = 0; k < n_keyframes; ++k) {
    const double* T = poses16 + 16 * static_cast<size_t>(k);
    for (size_t i = 0; i < n_points[k]; ++i, ++o) {
      const float* s = clouds[k] + stride * i;
      const double p[3] = {s[0], s[1], s[2]};
      out[4 * o + 0] = static_cast<float>(T[0] * p[0] + T[1] * p[1] + T[2] * p[2] + T[3]);
      out[4 * o + 1] = static_cast<float>(T[4] * p[0] + T[5] * p[1] + T[6] * p[2] + T[7]);
      out[4 * o + 2] = static_cast<float>(T[8] * p[0] + T[9] * p[1] + T[10] * p[2] + T[11]);
      out[4 * o + 3] = 1.0f;
    }
  }
}

// pcl::VoxelGrid<PointT>::applyFilter with downsample_all_data_ = true, min_points_per_voxel_ = 0 (PCL >= 1.14,
// filters/impl/voxel_grid.hpp -- not vendored in the reference; call sites extern/ndt_omp/apps/align.cpp:65-74,
// run/pipeline_ins_map_distribution.cpp:324-341): getMinMax3D over the finite points, min_b = floor(min_p * inv_leaf),
// idx = sum_k int(floor(p_k * inv_leaf) - float(min_b_k)) * divb_mul_k, points ordered by idx (PCL sorts with an
// unstable integer sort; here: stable, i.e. input order inside a leaf), one output per occupied leaf in ascending idx
// = float32 sum / float(count) (CentroidPoint / AccumulatorXYZ).  Returns the number of output points, -2 on index overflow.
long orc_voxelgrid_filter(const float* xyz, size_t n, size_t stride, float leaf, float* out_xyzw) {
  const float inv = 1.0f / leaf;
  float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
  std::vector<size_t> finite;
  finite.reserve(n);
  for (size_t i = 0; i < n; ++i) {
    const float* p = xyz + stride * i;
    if (!std::isfinite(p[0]) || !std::isfinite(p[1]) || !std::isfinite(p[2])) continue;
    finite.push_back(i);
    for (int k = 0; k < 3; ++k) { mn[k] = std::min(mn[k], p[k]); mx[k] = std::max(mx[k], p[k]); }
  }
  if (finite.empty()) return 0;
  const int64_t dx = static_cast<int64_t>((mx[0] - mn[0]) * inv) + 1, dy = static_cast<int64_t>((mx[1] - mn[1]) * inv) + 1,
                dz = static_cast<int64_t>((mx[2] - mn[2]) * inv) + 1;
  if (dx * dy * dz > static_cast<int64_t>(std::numeric_limits<int32_t>::max())) return -2;
  int min_b[3], div_b[3];
  for (int k = 0; k < 3; ++k) {
    min_b[k] = static_cast<int>(std::floor(mn[k] * inv));
    div_b[k] = static_cast<int>(std::floor(mx[k] * inv)) - min_b[k] + 1;
  }
  const int mul[3] = {1, div_b[0], div_b[0] * div_b[1]};
  std::vector<std::pair<unsigned int, size_t>> iv;
  iv.reserve(finite.size());
  for (size_t i : finite) {
    const float* p = xyz + stride * i;
    int idx = 0;
    for (int k = 0; k < 3; ++k) idx += static_cast<int>(std::floor(p[k] * inv) - static_cast<float>(min_b[k])) * mul[k];
    iv.emplace_back(static_cast<unsigned int>(idx), i);
  }
  std::stable_sort(iv.begin(), iv.end(), [](const std::pair<unsigned int, size_t>& a, const std::pair<unsigned int, size_t>& b) { return a.first < b.first; });
  long m = 0;
  for (size_t a = 0; a < iv.size();) {
    size_t b = a;
    float s[3] = {0.0f, 0.0f, 0.0f};
    while (b < iv.size() && iv[b].first == iv[a].first) {
      const float* p = xyz + stride * iv[b].second;
      s[0] = s[0] + p[0]; s[1] = s[1] + p[1]; s[2] = s[2] + p[2];
      ++b;
    }
    const float cnt = static_cast<float>(b - a);
    out_xyzw[4 * m + 0] = s[0] / cnt; out_xyzw[4 * m + 1] = s[1] / cnt; out_xyzw[4 * m + 2] = s[2] / cnt; out_xyzw[4 * m + 3] = 1.0f;
    ++m;
    a = b;
  }
  return m;
}

}  // extern "C"
