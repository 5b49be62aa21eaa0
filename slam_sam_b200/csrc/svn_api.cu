// C ABI, SVN-NDT part (svn_ndt::SvnNormalDistributionsTransform, extern/svn_ndt/include/svn_ndt.h:100-182,
// svn_ndt_impl.hpp:675-964): stage 1 = particle x point derivative kernel (derivatives.cu),
// stage 2/3 = K x K Stein kernel (svn_stein.cu).
#include "handle.cuh"
#include "newton_math.cuh"
#include "se3_math.cuh"
#include "svn_stein.cuh"
#include <cstring>
#include <random>
#include <vector>
#include <algorithm>

using namespace ndtb;

#define CKH(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { h->last_error = std::string(#x) + ": " + cudaGetErrorString(e_); return NDTB_ERR_CUDA; } } while (0)

namespace {

int check_svn(ndtb_handle_s* h) {
  if (!h) return NDTB_ERR_INVALID;
  cudaSetDevice(h->device);
  if (h->params.grid_flavour != NDTB_FLAVOUR_SVN) return set_error(h, NDTB_ERR_INVALID, "svn entry point needs grid_flavour = NDTB_FLAVOUR_SVN");
  if (!h->grid.built || h->n_source == 0) return set_error(h, NDTB_ERR_NOT_READY, "target and source must be set");
  return NDTB_OK;
}

// Stage 1 for particles [k0, k0+cnt) step `stride` (svn_ndt_impl.hpp:759-781): per-particle float
// transform (particles[k].matrix().cast<float>()), p_k = (translation, rpy), launch, read back.
int stage1(ndtb_handle_s* h, int K, const double* poses, int first, int stride, double* rec43 /* K x 43, only own rows written */) {
  const int n = static_cast<int>(h->n_source);
  int rc = ensure_eval_buffers(h, K, n);
  if (rc != NDTB_OK) return rc;
  std::vector<int> mine;
  for (int k = first; k < K; k += stride) mine.push_back(k);
  if (mine.empty()) return NDTB_OK;
  // own particles are packed into slots 0..m-1
  for (size_t s = 0; s < mine.size(); ++s) {
    const double* M = poses + 16 * mine[s];
    float T[16];
    for (int i = 0; i < 16; ++i) T[i] = static_cast<float>(M[i]);
    double p[6], rpy[3], R[9];
    for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) R[3 * r + c] = M[4 * r + c];
    rot_rpy(R, rpy);
    p[0] = M[3]; p[1] = M[7]; p[2] = M[11]; p[3] = rpy[0]; p[4] = rpy[1]; p[5] = rpy[2];
    EvalState& st = h->h_states[s];
    std::memset(&st, 0, sizeof(st));
    std::memcpy(st.T, T, 12 * sizeof(float));
    angle_tables(p, 1e-7, st.jang, st.hang);
    gauss_svn(h->params.outlier_ratio, h->params.resolution, &st.d1, &st.d2, &st.d3);
    std::memcpy(st.p, p, sizeof(p));
  }
  const int m = static_cast<int>(mine.size());
  CKH(cudaMemcpyAsync(h->eval.d_states, h->h_states, sizeof(EvalState) * m, cudaMemcpyHostToDevice, h->stream));
  const int n_probe = (h->params.search_method == NDTB_DIRECT7) ? 7 : 1;
  cudaEventRecord(h->ev[2], h->stream);
  launch_eval(FLAVOUR_SVN, n_probe, true, h->svn.use_gauss_newton != 0, h->d_source, n, h->grid.meta, h->grid.recs, h->grid.table,
              h->grid.c_mean, h->grid.c_icov, h->eval, m, 0, h->stream);
  cudaEventRecord(h->ev[3], h->stream);
  CKH(cudaMemcpyAsync(h->h_results, h->eval.d_results, sizeof(double) * RED_N * m, cudaMemcpyDeviceToHost, h->stream));
  CKH(cudaStreamSynchronize(h->stream));
  CKH(cudaGetLastError());
  float ms = 0;
  cudaEventElapsedTime(&ms, h->ev[2], h->ev[3]);
  h->eval_ms_total += ms; h->eval_launches += 1; h->launches += 1;
  for (int s = 0; s < m; ++s) {
    const double* r = h->h_results + static_cast<size_t>(s) * RED_N;
    double* o = rec43 + 43 * static_cast<size_t>(mine[s]);
    o[0] = r[RED_SCORE];
    std::memcpy(o + 1, r + RED_G, 48);
    std::memcpy(o + 7, r + RED_H, 288);
    // svn_ndt_impl.hpp:650-663: H += 1e-6 I, NaN guards
    for (int i = 0; i < 6; ++i) o[7 + 7 * i] += 1e-6;
    bool gf = true, hf = true;
    for (int i = 0; i < 6; ++i) if (!std::isfinite(o[1 + i])) gf = false;
    for (int i = 0; i < 36; ++i) if (!std::isfinite(o[7 + i])) hf = false;
    if (!gf) for (int i = 0; i < 6; ++i) o[1 + i] = 0.0;
    if (!hf) for (int i = 0; i < 36; ++i) o[7 + i] = (i % 7 == 0) ? 1.0 : 0.0;
    if (s == 0) { h->last_found = r[RED_FOUND]; h->last_hits = n ? r[RED_COUNT] / n : 0.0; }
  }
  return NDTB_OK;
}

void unpack(int K, const double* rec43, double* scores, double* grads, double* hess) {
  for (int k = 0; k < K; ++k) {
    if (scores) scores[k] = rec43[43 * k];
    if (grads) std::memcpy(grads + 6 * k, rec43 + 43 * k + 1, 48);
    if (hess) std::memcpy(hess + 36 * k, rec43 + 43 * k + 7, 288);
  }
}

}  // namespace

extern "C" {

int ndtb_svn_default_params(ndtb_svn_params* p) {
  if (!p) return NDTB_ERR_INVALID;
  p->particle_count = 30;     // svn_ndt_impl.hpp:65
  p->max_iterations = 50;
  p->kernel_bandwidth = 1.0;
  p->step_size = 1.0;         // :69
  p->stop_threshold = 1e-4;   // :70
  p->use_gauss_newton = 1;    // :71
  return NDTB_OK;
}

int ndtb_svn_set_params(ndtb_handle h, const ndtb_svn_params* p) {
  if (!h || !p) return NDTB_ERR_INVALID;
  h->svn = *p;
  return NDTB_OK;
}

int ndtb_svn_particle_derivatives(ndtb_handle h, int K, const double* poses, double* scores, double* grads, double* hess) {
  int rc = check_svn(h);
  if (rc != NDTB_OK) return rc;
  if (K <= 0 || !poses) return set_error(h, NDTB_ERR_INVALID, "K must be positive");
  std::vector<double> rec(43 * static_cast<size_t>(K));
  rc = stage1(h, K, poses, 0, 1, rec.data());
  if (rc != NDTB_OK) return rc;
  unpack(K, rec.data(), scores, grads, hess);
  return NDTB_OK;
}

int ndtb_svn_stage1_local(ndtb_handle h, int K, const double* poses, int rank, int world, double* records43) {
  int rc = check_svn(h);
  if (rc != NDTB_OK) return rc;
  if (K <= 0 || !poses || !records43 || world <= 0 || rank < 0 || rank >= world) return set_error(h, NDTB_ERR_INVALID, "bad arguments");
  return stage1(h, K, poses, rank, world, records43);
}

int ndtb_svn_stage2(ndtb_handle h, int K, const double* poses, const double* grads, const double* hess, double* updates) {
  int rc = check_svn(h);
  if (rc != NDTB_OK) return rc;
  if (K <= 0 || !poses || !grads || !hess || !updates) return set_error(h, NDTB_ERR_INVALID, "bad arguments");
  return svn_stein_update(h, K, poses, grads, hess, updates, nullptr);
}

int ndtb_svn_finish_step(ndtb_handle h, int K, double* poses, const double* records43, double* updates) {
  int rc = check_svn(h);
  if (rc != NDTB_OK) return rc;
  if (K <= 0 || !poses || !records43) return set_error(h, NDTB_ERR_INVALID, "bad arguments");
  std::vector<double> g(6 * static_cast<size_t>(K)), H(36 * static_cast<size_t>(K)), u(6 * static_cast<size_t>(K));
  unpack(K, records43, nullptr, g.data(), H.data());
  std::vector<double> newp(16 * static_cast<size_t>(K));
  rc = svn_stein_update(h, K, poses, g.data(), H.data(), u.data(), newp.data());
  if (rc != NDTB_OK) return rc;
  std::memcpy(poses, newp.data(), sizeof(double) * 16 * K);
  if (updates) std::memcpy(updates, u.data(), sizeof(double) * 6 * K);
  return NDTB_OK;
}

int ndtb_svn_step(ndtb_handle h, int K, double* poses, double* scores, double* grads, double* hess, double* updates) {
  int rc = check_svn(h);
  if (rc != NDTB_OK) return rc;
  if (K <= 0 || !poses) return set_error(h, NDTB_ERR_INVALID, "bad arguments");
  std::vector<double> rec(43 * static_cast<size_t>(K));
  rc = stage1(h, K, poses, 0, 1, rec.data());
  if (rc != NDTB_OK) return rc;
  unpack(K, rec.data(), scores, grads, hess);
  return ndtb_svn_finish_step(h, K, poses, rec.data(), updates);
}

int ndtb_svn_mean_pose(int K, const double* poses, const double prior[16], double mean16[16]) {
  if (K <= 0 || !poses || !prior || !mean16) return NDTB_ERR_INVALID;
  Pose prior_p; pose_from16(prior, prior_p);
  double mean_xi[6] = {0, 0, 0, 0, 0, 0};
  for (int k = 0; k < K; ++k) {
    Pose pk; pose_from16(poses + 16 * k, pk);
    double xi[6]; pose_logmap(pose_between(prior_p, pk), xi);
    for (int i = 0; i < 6; ++i) mean_xi[i] += xi[i];
  }
  for (int i = 0; i < 6; ++i) mean_xi[i] /= static_cast<double>(K);
  pose_to16(pose_retract(prior_p, mean_xi), mean16);
  return NDTB_OK;
}

int ndtb_svn_pose_distance(const double a[16], const double b[16], double* norm) {
  if (!a || !b || !norm) return NDTB_ERR_INVALID;
  Pose pa, pb; pose_from16(a, pa); pose_from16(b, pb);
  double d[6]; pose_logmap(pose_between(pa, pb), d);
  double n2 = 0; for (int i = 0; i < 6; ++i) n2 += d[i] * d[i];
  *norm = sqrt(n2);
  return NDTB_OK;
}

int ndtb_svn_align(ndtb_handle h, const double prior[16], const double* init_particles, uint64_t seed, ndtb_svn_result* result,
                   double* particles_out) {
  if (!h || !prior || !result) return NDTB_ERR_INVALID;
  cudaSetDevice(h->device);
  std::memset(result, 0, sizeof(*result));
  std::memcpy(result->final_pose, prior, 128);
  for (int i = 0; i < 6; ++i) result->final_covariance[7 * i] = 1.0;
  const int K = h->svn.particle_count;
  // svn_ndt_impl.hpp:682-702: invalid inputs -> converged=false, pose=prior, cov=I
  if (h->params.grid_flavour != NDTB_FLAVOUR_SVN) return set_error(h, NDTB_ERR_INVALID, "svn entry point needs grid_flavour = NDTB_FLAVOUR_SVN");
  if (!h->grid.built || h->grid.n_valid == 0) return set_error(h, NDTB_ERR_NOT_READY, "target NDT grid is not initialized");
  if (h->n_source == 0) return set_error(h, NDTB_ERR_EMPTY, "input source cloud is empty");
  if (K <= 0) return set_error(h, NDTB_ERR_INVALID, "particle count must be positive");
  cudaEventRecord(h->ev[0], h->stream);
  h->eval_ms_total = 0; h->eval_launches = 0;
  Pose prior_p; pose_from16(prior, prior_p);
  std::vector<double> P(16 * static_cast<size_t>(K));
  if (init_particles) std::memcpy(P.data(), init_particles, sizeof(double) * 16 * K);
  else {
    static const double sig[6] = {0.01, 0.01, 0.02, 0.05, 0.05, 0.05};  // :710 (rot, trans)
    std::mt19937_64 gen(seed);
    for (int k = 0; k < K; ++k) {
      double xi[6];
      for (int i = 0; i < 6; ++i) { std::normal_distribution<double> d(0.0, sig[i]); xi[i] = d(gen); }
      pose_to16(pose_retract(prior_p, xi), P.data() + 16 * k);
    }
  }
  Pose mean_cur = prior_p, mean_last;
  std::vector<double> rec(43 * static_cast<size_t>(K));
  for (int iter = 0; iter < h->svn.max_iterations; ++iter) {
    mean_last = mean_cur;
    int rc = stage1(h, K, P.data(), 0, 1, rec.data());
    if (rc != NDTB_OK) return rc;
    rc = ndtb_svn_finish_step(h, K, P.data(), rec.data(), nullptr);
    if (rc != NDTB_OK) return rc;
    double mean_xi[6] = {0, 0, 0, 0, 0, 0};
    for (int k = 0; k < K; ++k) {
      Pose pk; pose_from16(P.data() + 16 * k, pk);
      double xi[6]; pose_logmap(pose_between(prior_p, pk), xi);
      for (int i = 0; i < 6; ++i) mean_xi[i] += xi[i];
    }
    for (int i = 0; i < 6; ++i) mean_xi[i] /= static_cast<double>(K);
    mean_cur = pose_retract(prior_p, mean_xi);
    result->iterations = iter + 1;
    double dxi[6]; pose_logmap(pose_between(mean_last, mean_cur), dxi);
    double nrm = 0; for (int i = 0; i < 6; ++i) nrm += dxi[i] * dxi[i];
    if (sqrt(nrm) < h->svn.stop_threshold) { result->converged = 1; break; }
  }
  pose_to16(mean_cur, result->final_pose);
  double* cov = result->final_covariance;
  std::memset(cov, 0, 288);
  if (K > 1) {
    std::vector<double> tv(6 * static_cast<size_t>(K));
    double m[6] = {0, 0, 0, 0, 0, 0};
    for (int k = 0; k < K; ++k) {
      Pose pk; pose_from16(P.data() + 16 * k, pk);
      pose_logmap(pose_between(mean_cur, pk), &tv[6 * k]);
      for (int i = 0; i < 6; ++i) m[i] += tv[6 * k + i];
    }
    for (int i = 0; i < 6; ++i) m[i] /= static_cast<double>(K);
    for (int k = 0; k < K; ++k) {
      double d[6]; for (int i = 0; i < 6; ++i) d[i] = tv[6 * k + i] - m[i];
      for (int i = 0; i < 6; ++i) for (int j = 0; j < 6; ++j) cov[6 * i + j] += d[i] * d[j];
    }
    for (int i = 0; i < 36; ++i) cov[i] /= static_cast<double>(K - 1);
  } else {
    static const double sig[6] = {0.01, 0.01, 0.02, 0.05, 0.05, 0.05};
    for (int i = 0; i < 6; ++i) cov[7 * i] = 1e-6 * (sig[i] * sig[i]);
  }
  floor_eigenvalues6(cov, 1e-9);  // svn_ndt_impl.hpp:932-945
  if (particles_out) std::memcpy(particles_out, P.data(), sizeof(double) * 16 * K);
  cudaEventRecord(h->ev[1], h->stream);
  cudaEventSynchronize(h->ev[1]);
  cudaEventElapsedTime(&h->align_ms, h->ev[0], h->ev[1]);
  h->align_time_pending = false;
  return NDTB_OK;
}

}  // extern "C"
