// Replay batches (BASELINE configs[3]): n independent scans against the current target map.
// One Newton loop is a ping-pong -- the GPU evaluates (~29 us), the host solves and sends the next pose (~8 us with
// the PCIe hops) -- so a single alignment leaves the GPU idle a fifth of the time.  A batch does not have to: this
// file runs up to THREE alignments interleaved on the handle's stream.  While the host post-processes scan A's evaluation,
// scan B's evaluation kernel is already running, and A's next kernel is queued behind it before B finishes.  The
// next scans are uploaded on a second stream meanwhile.  Every alignment performs exactly the operations of
// ndtb_align (same kernels, same host arithmetic, same order), so the results are bit-identical to the serial loop
// (tests/test_gpu_large_map.py::test_replay_batch_*).
#include "handle.cuh"
#include "newton_math.cuh"
#include <algorithm>
#include <cmath>
#include <cstring>
#include <deque>
#include <string>
#include <vector>

using namespace ndtb;

#define CKH(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { h->last_error = std::string(#x) + ": " + cudaGetErrorString(e_); return NDTB_ERR_CUDA; } } while (0)

namespace {

constexpr int SLOTS = 4;  // up to three alignments in flight + one slot whose scan is being uploaded
static const int IN_FLIGHT = getenv("NDTB_BATCH_IN_FLIGHT") ? atoi(getenv("NDTB_BATCH_IN_FLIGHT")) : 3;   // measured: 3692 / 4428 / 4860 scans/s with 1 / 2 / 3 in flight; 4 in flight: 3742

struct Job {
  int scan = -1;                 // index into the batch, -1: slot free
  float4* d_src = nullptr; size_t n = 0;
  cudaEvent_t uploaded = nullptr;
  double* results = nullptr;     // this slot's 46 self-validating result slots (pinned host memory)
  unsigned long long key = 0;    // key of the evaluation in flight
  // Newton loop state (the locals of ndtb_align)
  float final_T[16];
  double p[6], x_t[6], delta_p[6], g[6], H[36], score, nvtl, a_t;
  int nr_iterations, evaluations;
  bool first, converged;
};

}  // namespace

namespace ndtb {

static int launch_job_eval(ndtb_handle_s* h, Job& j, const double* pose, const float* T16) {
  EvalState st;
  fill_eval_state(h, st, pose, T16, 10e-5, false);
  j.key = 0x9E3779B97F4A7C15ull * static_cast<unsigned long long>(++h->flag_seq) | 1ull;
  const int n_probe = (h->params.search_method == NDTB_DIRECT1) ? 1 : 7;
  CKH(launch_eval_pclomp(n_probe, true, j.d_src, static_cast<int>(j.n), h->grid.meta, h->grid.recs, h->grid.table, st, h->eval, j.results, j.key,
                         h->stream));
  h->launches += 1;
  j.evaluations += 1;
  return NDTB_OK;
}

static int wait_job(ndtb_handle_s* h, Job& j, double* rbuf) {
  volatile unsigned long long* slots = reinterpret_cast<volatile unsigned long long*>(j.results);
  unsigned int spins = 0;
  for (int c = RED_N - 1; c >= 0; --c) {
    for (;;) {
      const unsigned long long w0 = slots[2 * c], w1 = slots[2 * c + 1];
      if ((w0 ^ w1) == j.key) { std::memcpy(&rbuf[c], &w0, 8); break; }
      if ((++spins & 4095u) == 0) {
        const cudaError_t q = cudaStreamQuery(h->stream);
        if (q != cudaSuccess && q != cudaErrorNotReady) { h->last_error = std::string("eval kernel: ") + cudaGetErrorString(q); return NDTB_ERR_CUDA; }
        if (q == cudaSuccess) {
          const unsigned long long x0 = slots[2 * c], x1 = slots[2 * c + 1];
          if ((x0 ^ x1) != j.key) { h->last_error = "eval kernel finished without publishing its result"; return NDTB_ERR_CUDA; }
        }
      }
    }
  }
  __sync_synchronize();
  return NDTB_OK;
}

// The body of ndtb_align's loop between two evaluations (use_line_search == 0), as a resumable step: consumes the
// evaluation that just arrived and either launches the next one (returns 1) or finishes the alignment (returns 0).
static int advance_job(ndtb_handle_s* h, Job& j, const double* rbuf, ndtb_result* res) {
  const ndtb_params& P = h->params;
  const size_t N = j.n;
  finish_eval(h, static_cast<int>(N), rbuf, j.first ? j.p : j.x_t, &j.score, j.g, j.H, &j.nvtl);
  if (!j.first) {
    // tail of the previous iteration (ndt_omp_impl.hpp:176-190)
    for (int i = 0; i < 6; ++i) j.delta_p[i] *= j.a_t;
    for (int i = 0; i < 6; ++i) j.p[i] = j.p[i] + j.delta_p[i];
    if (j.nr_iterations > P.max_iterations || (j.nr_iterations && (fabs(j.a_t) < P.trans_epsilon))) j.converged = true;
    j.nr_iterations++;
    if (j.converged) {
      res->transform_probability = N ? j.score / static_cast<double>(N) : 0.0;
      std::memcpy(res->hessian, j.H, sizeof(res->hessian));
      return 0;
    }
  }
  j.first = false;
  for (;;) {
    double neg_g[6];
    for (int i = 0; i < 6; ++i) neg_g[i] = -j.g[i];
    if (!(h->direct_solver && gecp6_solve(j.H, neg_g, j.delta_p, 1e-6))) svd6_solve(j.H, neg_g, j.delta_p);
    double nrm2 = 0;
    for (int i = 0; i < 6; ++i) nrm2 += j.delta_p[i] * j.delta_p[i];
    double delta_p_norm = sqrt(nrm2);
    if (delta_p_norm == 0 || delta_p_norm != delta_p_norm) {
      res->transform_probability = N ? j.score / static_cast<double>(N) : 0.0;
      j.converged = delta_p_norm == delta_p_norm;
      return 0;
    }
    for (int i = 0; i < 6; ++i) j.delta_p[i] = j.delta_p[i] / delta_p_norm;
    double d_phi_0 = 0;
    for (int i = 0; i < 6; ++i) d_phi_0 += j.g[i] * j.delta_p[i];
    d_phi_0 = -d_phi_0;
    if (d_phi_0 >= 0 && d_phi_0 == 0) {
      // zero step without an evaluation: the iteration completes on the host and the loop goes round again
      j.a_t = 0;
      for (int i = 0; i < 6; ++i) j.delta_p[i] *= j.a_t;
      for (int i = 0; i < 6; ++i) j.p[i] = j.p[i] + j.delta_p[i];
      if (j.nr_iterations > P.max_iterations || (j.nr_iterations && (fabs(j.a_t) < P.trans_epsilon))) j.converged = true;
      j.nr_iterations++;
      if (j.converged) {
        res->transform_probability = N ? j.score / static_cast<double>(N) : 0.0;
        std::memcpy(res->hessian, j.H, sizeof(res->hessian));
        return 0;
      }
      continue;
    }
    if (d_phi_0 >= 0) for (int i = 0; i < 6; ++i) j.delta_p[i] *= -1;
    j.a_t = delta_p_norm;
    j.a_t = std::min(j.a_t, P.step_size);
    j.a_t = std::max(j.a_t, P.trans_epsilon / 2);
    for (int i = 0; i < 6; ++i) j.x_t[i] = j.p[i] + j.delta_p[i] * j.a_t;
    pose_to_T(j.x_t, j.final_T);
    const int rc = launch_job_eval(h, j, j.x_t, j.final_T);
    return rc == NDTB_OK ? 1 : -rc - 1000;
  }
}

}  // namespace ndtb

extern "C" {

// Replay batch: n independent source scans against the current target map, results in order.
int ndtb_align_batch(ndtb_handle h, int n_scans, const float* const* xyz, const size_t* n_points, size_t stride_bytes, const float* guesses16,
                     ndtb_result* results) {
  if (!h || n_scans < 0 || !xyz || !n_points || !guesses16 || !results) return NDTB_ERR_INVALID;
  cudaSetDevice(h->device);
  const bool interleave = h->params.grid_flavour == NDTB_FLAVOUR_PCLOMP && !h->params.use_line_search && !h->kernel_events && h->batch_interleave &&
                          h->grid.built && n_scans > 1 && h->params.search_method != NDTB_DIRECT26 && h->params.search_method != NDTB_KDTREE;
  if (!interleave) {  // serial: one ndtb_set_source + ndtb_align per scan
    for (int i = 0; i < n_scans; ++i) {
      int rc = ndtb_set_source(h, xyz[i], n_points[i], stride_bytes);
      if (rc != NDTB_OK) return rc;
      rc = ndtb_align(h, guesses16 + 16 * i, results + i, nullptr, 0, nullptr);
      if (rc != NDTB_OK) return rc;
    }
    return NDTB_OK;
  }
  size_t n_max = 0;
  for (int i = 0; i < n_scans; ++i) {
    if (!xyz[i] || n_points[i] == 0) return set_error(h, NDTB_ERR_EMPTY, "empty source cloud in the batch");
    if (stride_bytes < sizeof(float4)) return set_error(h, NDTB_ERR_INVALID, "batch scans must be packed x,y,z,w floats or wider (stride >= 16)");
    n_max = std::max(n_max, n_points[i]);
  }
  int rc = ensure_eval_buffers(h, 1, static_cast<int>(n_max));
  if (rc != NDTB_OK) return rc;
  const size_t res_words = 2 * (RED_N + 2);
  if (!h->batch_copy_stream) CKH(cudaStreamCreateWithFlags(&h->batch_copy_stream, cudaStreamNonBlocking));
  if (!h->batch_h_results) {
    CKH(cudaMallocHost(&h->batch_h_results, sizeof(double) * res_words * SLOTS));
    std::memset(h->batch_h_results, 0, sizeof(double) * res_words * SLOTS);
  }
  cudaStream_t copy_stream = h->batch_copy_stream;
  Job jobs[SLOTS];
  for (int s = 0; s < SLOTS; ++s) {
    ndtb_handle_s::BatchSlot& bs = h->batch_slot[s];
    if (!bs.uploaded) CKH(cudaEventCreateWithFlags(&bs.uploaded, cudaEventDisableTiming));
    if (bs.alloc < n_max) {
      CKH(cudaStreamSynchronize(h->stream));
      cudaFree(bs.d_src); bs.d_src = nullptr; bs.alloc = 0;
      CKH(cudaMalloc(&bs.d_src, (n_max + n_max / 8 + 256) * sizeof(float4)));
      bs.alloc = n_max + n_max / 8 + 256;
    }
    jobs[s].d_src = bs.d_src; jobs[s].uploaded = bs.uploaded;
    jobs[s].results = h->batch_h_results + res_words * s;
  }
  auto cleanup = [&]() {  // nothing of this call may still be running when the caller's buffers and results are released
    cudaStreamSynchronize(copy_stream);
    cudaStreamSynchronize(h->stream);
  };
  int next = 0, done = 0;
  std::deque<int> staged;    // slots whose scan is uploaded (or uploading) but whose first evaluation is not launched yet
  std::deque<int> in_flight; // slots with an evaluation on the stream, in launch order
  auto upload = [&](int s) -> int {
    Job& j = jobs[s];
    j.scan = next; j.n = n_points[next];
    cudaError_t e = (stride_bytes == sizeof(float4))
                        ? cudaMemcpyAsync(j.d_src, xyz[next], j.n * sizeof(float4), cudaMemcpyHostToDevice, copy_stream)
                        : cudaMemcpy2DAsync(j.d_src, sizeof(float4), xyz[next], stride_bytes, sizeof(float4), j.n, cudaMemcpyHostToDevice, copy_stream);
    if (e == cudaSuccess) e = cudaEventRecord(j.uploaded, copy_stream);
    if (e != cudaSuccess) { h->last_error = std::string("batch upload: ") + cudaGetErrorString(e); return NDTB_ERR_CUDA; }
    ++next;
    staged.push_back(s);
    return NDTB_OK;
  };
  auto start = [&](int s) -> int {  // head of ndtb_align: guess -> p, first evaluation
    Job& j = jobs[s];
    ndtb_result* res = results + j.scan;
    std::memset(res, 0, sizeof(*res));
    const float* guess = guesses16 + 16 * static_cast<size_t>(j.scan);
    const float ident[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
    std::memcpy(j.final_T, ident, sizeof(ident));
    bool guess_is_identity = true;
    for (int i = 0; i < 16; ++i) if (guess[i] != ident[i]) guess_is_identity = false;
    if (!guess_is_identity) std::memcpy(j.final_T, guess, sizeof(j.final_T));
    float eul[3];
    euler_012(j.final_T, eul);
    j.p[0] = j.final_T[3]; j.p[1] = j.final_T[7]; j.p[2] = j.final_T[11]; j.p[3] = eul[0]; j.p[4] = eul[1]; j.p[5] = eul[2];
    j.nr_iterations = 0; j.evaluations = 0; j.first = true; j.converged = false; j.score = 0; j.nvtl = 0; j.a_t = 0;
    if (cudaStreamWaitEvent(h->stream, j.uploaded, 0) != cudaSuccess) { h->last_error = "batch: stream wait"; return NDTB_ERR_CUDA; }
    const int rc2 = launch_job_eval(h, j, j.p, j.final_T);
    if (rc2 == NDTB_OK) in_flight.push_back(s);
    return rc2;
  };
  auto free_slot = [&]() -> int { for (int s = 0; s < SLOTS; ++s) if (jobs[s].scan < 0) return s; return -1; };
  rc = NDTB_OK;
  while (done < n_scans && rc == NDTB_OK) {
    // keep the upload slot busy and three alignments in flight
    for (int s; next < n_scans && (s = free_slot()) >= 0 && rc == NDTB_OK;) rc = upload(s);
    while (rc == NDTB_OK && static_cast<int>(in_flight.size()) < std::min(std::max(IN_FLIGHT, 1), SLOTS - 1) && !staged.empty()) { const int s = staged.front(); staged.pop_front(); rc = start(s); }
    if (rc != NDTB_OK || in_flight.empty()) break;
    const int s = in_flight.front();
    in_flight.pop_front();
    Job& j = jobs[s];
    double rbuf[RED_N];
    rc = wait_job(h, j, rbuf);
    if (rc != NDTB_OK) break;
    ndtb_result* res = results + j.scan;
    const int more = advance_job(h, j, rbuf, res);
    if (more == 1) { in_flight.push_back(s); continue; }
    if (more < 0) { rc = NDTB_ERR_CUDA; break; }
    std::memcpy(res->pose, j.final_T, sizeof(j.final_T));
    res->nearest_voxel_transformation_likelihood = j.nvtl;
    res->iteration_num = j.nr_iterations;
    res->converged = j.converged ? 1 : 0;
    res->transformation_array_len = 0;
    res->evaluations = j.evaluations;
    j.scan = -1;
    ++done;
  }
  cleanup();
  if (rc == NDTB_OK && done != n_scans) rc = set_error(h, NDTB_ERR_CUDA, "batch scheduler stalled");
  return rc;
}

}  // extern "C"
