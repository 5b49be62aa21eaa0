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

// Stage 1 for the particles k = first + s * stride (svn_ndt_impl.hpp:759-781), host-staged parity hook: the poses are uploaded,
// the device path of the align loop runs (state kernel + derivative kernel, records into this handle's window), and the
// records of the own particles are copied back.  rec43: K x 43, only own rows written.
int stage1(ndtb_handle_s* h, int K, const double* poses, int first, int stride, double* rec43) {
  int rc = svn_dev_upload_poses(h, K, poses);
  if (rc != NDTB_OK) return rc;
  CKH(cudaMemsetAsync(h->svnd.d_done, 0, sizeof(int), h->stream));
  unsigned long long stamp = 0;
  cudaEventRecord(h->ev[2], h->stream);
  rc = svn_dev_stage1(h, K, first, stride, false, &stamp);
  if (rc != NDTB_OK) return rc;
  cudaEventRecord(h->ev[3], h->stream);
  const int n = static_cast<int>(h->n_source);
  std::vector<double> win(static_cast<size_t>(SVN_REC) * K);
  const double* d_rec = svn_window_recs(h->svnd.d_window) + static_cast<size_t>(stamp & 1ull) * ndtb_handle_s::SvnDev::CAPACITY * SVN_REC;
  CKH(cudaMemcpyAsync(win.data(), d_rec, sizeof(double) * SVN_REC * K, cudaMemcpyDeviceToHost, h->stream));
  double r0[RED_N];
  CKH(cudaMemcpyAsync(r0, h->eval.d_results, sizeof(r0), cudaMemcpyDeviceToHost, h->stream));
  CKH(cudaStreamSynchronize(h->stream));
  float ms = 0;
  cudaEventElapsedTime(&ms, h->ev[2], h->ev[3]);
  h->eval_ms_total += ms;
  for (int k = first; k < K; k += stride) std::memcpy(rec43 + SVN_REC * static_cast<size_t>(k), &win[SVN_REC * static_cast<size_t>(k)], sizeof(double) * SVN_REC);
  if (first < K) { h->last_found = r0[RED_FOUND]; h->last_hits = n ? r0[RED_COUNT] / n : 0.0; }
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

// align(source, prior) (svn_ndt_impl.hpp:675-964) with the whole loop on the device.  Per iteration three launches are queued --
// state kernel, derivative kernel (its last blocks store the records into the window of every rank), Stein kernel (acquires
// the records, updates the particles, mean pose + stop test) -- and the host only polls the pinned status slot of iteration
// i - 1 while iteration i is already queued; kernels queued after the stop test fired see the device `done` flag and return.
// With a communicator (ndtb_comm_*) the call is COLLECTIVE: every rank calls it with the same arguments, evaluates the
// particles k % world == rank and ends with bit-identical particles, mean pose and covariance.
static int svn_align_device(ndtb_handle h, const double prior[16], const double* init_particles, uint64_t seed, ndtb_svn_result* result,
                            double* particles_out) {
  std::memset(result, 0, sizeof(*result));
  std::memcpy(result->final_pose, prior, 128);
  for (int i = 0; i < 6; ++i) result->final_covariance[7 * i] = 1.0;
  const int K = h->svn.particle_count;
  // svn_ndt_impl.hpp:682-702: invalid inputs -> converged=false, pose=prior, cov=I
  if (h->params.grid_flavour != NDTB_FLAVOUR_SVN) return set_error(h, NDTB_ERR_INVALID, "svn entry point needs grid_flavour = NDTB_FLAVOUR_SVN");
  if (!h->grid.built || h->grid.n_valid == 0) return set_error(h, NDTB_ERR_NOT_READY, "target NDT grid is not initialized");
  if (h->n_source == 0) return set_error(h, NDTB_ERR_EMPTY, "input source cloud is empty");
  if (K <= 0) return set_error(h, NDTB_ERR_INVALID, "particle count must be positive");
  int rc = svn_dev_ensure(h);
  if (rc != NDTB_OK) return rc;
  ndtb_handle_s::SvnDev& d = h->svnd;
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
      for (int i = 0; i < 6; ++i) { std::normal_distribution<double> dd(0.0, sig[i]); xi[i] = dd(gen); }
      pose_to16(pose_retract(prior_p, xi), P.data() + 16 * k);
    }
  }
  rc = svn_dev_upload_poses(h, K, P.data());
  if (rc != NDTB_OK) return rc;
  // mean_prev of iteration 0 and the prior itself (d_mean: [0..15] / [16..31] ping-pong, [32..47] prior)
  double m48[48];
  std::memcpy(m48, prior, 128); std::memcpy(m48 + 16, prior, 128); std::memcpy(m48 + 32, prior, 128);
  CKH(cudaMemcpyAsync(d.d_mean, m48, sizeof(m48), cudaMemcpyHostToDevice, h->stream));
  CKH(cudaMemsetAsync(d.d_done, 0, sizeof(int), h->stream));
  CKH(cudaMemsetAsync(d.d_wait_ns, 0, sizeof(unsigned long long), h->stream));
  const int max_iter = h->svn.max_iterations;
  while (d.ev.size() < 3 * static_cast<size_t>(std::max(max_iter, 1))) { cudaEvent_t e; CKH(cudaEventCreate(&e)); d.ev.push_back(e); }
  // Work units of the loop: every particle is cut into `vsh` point shards (the point groups dealt out between them), the K * vsh
  // units go round-robin to the ranks -- an even load for any particle count, 100 particles over 8 GPUs included -- and the Stein
  // kernel adds a particle's shard records in shard order.  The shard count depends on K alone, not on the number of ranks, so the
  // sharded align and the single-GPU align add the same partial sums in the same order: bit-identical results.
  // vsh = the smallest power of two that makes K * vsh a multiple of 8 (an even split over 1, 2, 4 or 8 ranks): 2 for K = 100.
  int vsh = 1;
  while (vsh < 8 && (K * vsh) % 8 != 0) vsh <<= 1;
  while (vsh > 1 && K * vsh > ndtb_handle_s::SvnDev::CAPACITY) vsh >>= 1;
  const unsigned long long stamp0 = d.stamp;
  int iterations = 0, done = 0, queued = 0;
  auto wait_status = [&](int it) -> int {   // returns the done code of iteration `it` (0 running on, 1 converged, 2 peer time-out), < 0 on a CUDA error
    const unsigned long long key = 0x9E3779B97F4A7C15ull * (stamp0 + static_cast<unsigned long long>(it) + 1ull) | 1ull;
    volatile unsigned long long* sl = d.h_status + 2 * (it & 63);
    unsigned int spins = 0;
    for (;;) {
      const unsigned long long w0 = sl[0], w1 = sl[1];
      if ((w0 ^ w1) == key) return static_cast<int>(w0 >> 32);
      if ((++spins & 4095u) == 0) {
        const cudaError_t q = cudaStreamQuery(h->stream);
        if (q != cudaSuccess && q != cudaErrorNotReady) return -1;
        if (q == cudaSuccess) { const unsigned long long x0 = sl[0], x1 = sl[1]; if ((x0 ^ x1) == key) return static_cast<int>(x0 >> 32); return -1; }
      }
    }
  };
  for (int iter = 0; iter < max_iter; ++iter) {
    unsigned long long stamp = 0;
    cudaEventRecord(d.ev[3 * iter], h->stream);
    rc = svn_dev_stage1(h, K, d.rank, d.world, true, &stamp, vsh);
    if (rc != NDTB_OK) return rc;
    cudaEventRecord(d.ev[3 * iter + 1], h->stream);
    rc = svn_dev_stein(h, K, stamp, true, true, true, iter, vsh);
    if (rc != NDTB_OK) return rc;
    cudaEventRecord(d.ev[3 * iter + 2], h->stream);
    queued = iter + 1;
    if (iter >= 1) {   // iteration iter is queued: now look at the outcome of iter - 1
      done = wait_status(iter - 1);
      if (done < 0) { h->last_error = "svn align: kernel failure while waiting for an iteration's status"; return NDTB_ERR_CUDA; }
      if (done) { iterations = iter; break; }
    }
  }
  if (!done) {
    done = wait_status(queued - 1);
    if (done < 0) { h->last_error = "svn align: kernel failure while waiting for an iteration's status"; return NDTB_ERR_CUDA; }
    iterations = queued;
  }
  // a speculatively queued iteration returned at once (done flag) without touching particles or stamps on any rank: every rank
  // queued the same number of launches, so the stamps stay in lock-step; rewind the ping-pong and the stamp it consumed
  if (queued > iterations) {
    d.stamp = stamp0 + static_cast<unsigned long long>(iterations);
    if ((queued - iterations) & 1) d.cur ^= 1;
  }
  std::vector<double> mean16(16);
  CKH(cudaMemcpyAsync(P.data(), d.d_poses[d.cur], sizeof(double) * 16 * K, cudaMemcpyDeviceToHost, h->stream));
  CKH(cudaMemcpyAsync(mean16.data(), d.d_mean + 16 * (iterations & 1), 128, cudaMemcpyDeviceToHost, h->stream));
  unsigned long long wait_ns = 0;
  double r0[RED_N];
  CKH(cudaMemcpyAsync(&wait_ns, d.d_wait_ns, 8, cudaMemcpyDeviceToHost, h->stream));
  CKH(cudaMemcpyAsync(r0, h->eval.d_results, sizeof(r0), cudaMemcpyDeviceToHost, h->stream));
  cudaEventRecord(h->ev[1], h->stream);
  CKH(cudaStreamSynchronize(h->stream));
  if (done == 2) return set_error(h, NDTB_ERR_CUDA, "svn align: a peer rank did not deliver its particle records in time");
  // row 0 is one point shard of a particle (the shards interleave the point groups): scaled to the whole cloud
  h->last_found = r0[RED_FOUND] * vsh; h->last_hits = r0[RED_COUNT] * vsh / static_cast<double>(h->n_source);
  d.eval_ms = 0; d.stein_ms = 0;
  for (int it = 0; it < iterations; ++it) {
    float a = 0, b = 0;
    cudaEventElapsedTime(&a, d.ev[3 * it], d.ev[3 * it + 1]);
    cudaEventElapsedTime(&b, d.ev[3 * it + 1], d.ev[3 * it + 2]);
    d.eval_ms += a; d.stein_ms += b;
  }
  d.wait_ms = static_cast<float>(wait_ns * 1e-6);
  d.iterations = iterations;
  h->eval_ms_total = d.eval_ms;
  cudaEventElapsedTime(&h->align_ms, h->ev[0], h->ev[1]);
  h->align_time_pending = false;
  result->iterations = iterations;
  result->converged = (done == 1) ? 1 : 0;
  Pose mean_cur; pose_from16(mean16.data(), mean_cur);
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
      double dd[6]; for (int i = 0; i < 6; ++i) dd[i] = tv[6 * k + i] - m[i];
      for (int i = 0; i < 6; ++i) for (int j = 0; j < 6; ++j) cov[6 * i + j] += dd[i] * dd[j];
    }
    for (int i = 0; i < 36; ++i) cov[i] /= static_cast<double>(K - 1);
  } else {
    static const double sig[6] = {0.01, 0.01, 0.02, 0.05, 0.05, 0.05};
    for (int i = 0; i < 6; ++i) cov[7 * i] = 1e-6 * (sig[i] * sig[i]);
  }
  floor_eigenvalues6(cov, 1e-9);  // svn_ndt_impl.hpp:932-945
  if (particles_out) std::memcpy(particles_out, P.data(), sizeof(double) * 16 * K);
  return NDTB_OK;
}

int ndtb_svn_align(ndtb_handle h, const double prior[16], const double* init_particles, uint64_t seed, ndtb_svn_result* result,
                   double* particles_out) {
  if (!h || !prior || !result) return NDTB_ERR_INVALID;
  cudaSetDevice(h->device);
  return svn_align_device(h, prior, init_particles, seed, result, particles_out);
}

int ndtb_svn_align_sharded(ndtb_handle h, const double prior[16], const double* init_particles, uint64_t seed, ndtb_svn_result* result,
                           double* particles_out) {
  if (!h || !prior || !result) return NDTB_ERR_INVALID;
  if (h->svnd.world < 2) return set_error(h, NDTB_ERR_NOT_READY, "no communicator: call ndtb_comm_connect / ndtb_comm_init first");
  cudaSetDevice(h->device);
  return svn_align_device(h, prior, init_particles, seed, result, particles_out);
}

int ndtb_svn_get_timings(ndtb_handle h, float* eval_ms, float* stein_ms, float* exchange_wait_ms, int* iterations) {
  if (!h) return NDTB_ERR_INVALID;
  if (eval_ms) *eval_ms = h->svnd.eval_ms;
  if (stein_ms) *stein_ms = h->svnd.stein_ms;
  if (exchange_wait_ms) *exchange_wait_ms = h->svnd.wait_ms;
  if (iterations) *iterations = h->svnd.iterations;
  return NDTB_OK;
}

// ---- communicator: the record windows of all ranks, mapped into every rank ------------------------------------------
int ndtb_comm_window_handle(ndtb_handle h, void* ipc_handle, size_t capacity_bytes, size_t* handle_bytes) {
  if (handle_bytes) *handle_bytes = sizeof(cudaIpcMemHandle_t);
  if (!h) return NDTB_ERR_INVALID;
  if (!ipc_handle) return NDTB_OK;
  if (capacity_bytes < sizeof(cudaIpcMemHandle_t)) return set_error(h, NDTB_ERR_INVALID, "ipc handle buffer too small");
  cudaSetDevice(h->device);
  int rc = svn_dev_ensure(h);
  if (rc != NDTB_OK) return rc;
  cudaIpcMemHandle_t mh;
  CKH(cudaIpcGetMemHandle(&mh, h->svnd.d_window));
  std::memcpy(ipc_handle, &mh, sizeof(mh));
  return NDTB_OK;
}

static void comm_drop_peers(ndtb_handle_s* h) {
  ndtb_handle_s::SvnDev& d = h->svnd;
  for (int r = 0; r < 8; ++r) {
    if (d.peer_is_ipc[r] && d.peer_base[r]) cudaIpcCloseMemHandle(d.peer_base[r]);
    d.peer_is_ipc[r] = false; d.peer_base[r] = nullptr;
  }
  d.peer_base[0] = d.d_window; d.rank = 0; d.world = 1;
}

int ndtb_comm_connect(ndtb_handle h, int rank, int world, const void* all_ipc_handles) {
  if (!h || !all_ipc_handles) return NDTB_ERR_INVALID;
  if (world < 1 || world > SVN_MAX_RANKS || rank < 0 || rank >= world) return set_error(h, NDTB_ERR_INVALID, "rank / world out of range (world <= 8)");
  cudaSetDevice(h->device);
  int rc = svn_dev_ensure(h);
  if (rc != NDTB_OK) return rc;
  CKH(cudaStreamSynchronize(h->stream));
  comm_drop_peers(h);
  ndtb_handle_s::SvnDev& d = h->svnd;
  const cudaIpcMemHandle_t* hs = static_cast<const cudaIpcMemHandle_t*>(all_ipc_handles);
  for (int r = 0; r < world; ++r) {
    if (r == rank) { d.peer_base[r] = d.d_window; continue; }
    void* p = nullptr;
    const cudaError_t e = cudaIpcOpenMemHandle(&p, hs[r], cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) { comm_drop_peers(h); h->last_error = std::string("cudaIpcOpenMemHandle: ") + cudaGetErrorString(e); return NDTB_ERR_CUDA; }
    d.peer_base[r] = p; d.peer_is_ipc[r] = true;
  }
  d.rank = rank; d.world = world;
  d.stamp = 0;   // every rank restarts its stamps here; the caller's barrier after connect orders this against the first align
  CKH(cudaMemset(d.d_window, 0, SVN_WINDOW_FLAG_BYTES));
  return NDTB_OK;
}

int ndtb_comm_init(ndtb_handle* handles, int n) {
  if (!handles || n < 1 || n > SVN_MAX_RANKS) return NDTB_ERR_INVALID;
  for (int r = 0; r < n; ++r) {
    ndtb_handle_s* h = handles[r];
    if (!h) return NDTB_ERR_INVALID;
    cudaSetDevice(h->device);
    int rc = svn_dev_ensure(h);
    if (rc != NDTB_OK) return rc;
    CKH(cudaStreamSynchronize(h->stream));
    comm_drop_peers(h);
  }
  for (int r = 0; r < n; ++r) {
    ndtb_handle_s* h = handles[r];
    cudaSetDevice(h->device);
    for (int q = 0; q < n; ++q) {
      if (q == r || handles[q]->device == h->device) continue;
      int can = 0;
      CKH(cudaDeviceCanAccessPeer(&can, h->device, handles[q]->device));
      if (!can) return set_error(h, NDTB_ERR_CUDA, "no peer access between the devices of two handles");
      const cudaError_t e = cudaDeviceEnablePeerAccess(handles[q]->device, 0);
      if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { h->last_error = std::string("cudaDeviceEnablePeerAccess: ") + cudaGetErrorString(e); return NDTB_ERR_CUDA; }
      cudaGetLastError();
    }
    ndtb_handle_s::SvnDev& d = h->svnd;
    for (int q = 0; q < n; ++q) d.peer_base[q] = handles[q]->svnd.d_window;
    d.rank = r; d.world = n; d.stamp = 0;
    CKH(cudaMemset(d.d_window, 0, SVN_WINDOW_FLAG_BYTES));
  }
  return NDTB_OK;
}

int ndtb_comm_close(ndtb_handle h) {
  if (!h) return NDTB_ERR_INVALID;
  cudaSetDevice(h->device);
  cudaStreamSynchronize(h->stream);
  comm_drop_peers(h);
  return NDTB_OK;
}

int ndtb_comm_info(ndtb_handle h, int* rank, int* world) {
  if (!h) return NDTB_ERR_INVALID;
  if (rank) *rank = h->svnd.rank;
  if (world) *world = h->svnd.world;
  return NDTB_OK;
}

}  // extern "C"
