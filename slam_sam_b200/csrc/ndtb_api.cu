// C ABI of libndtb: lifetime, parameters, clouds, voxel-map access, pclomp NDT (include/ndtb.h).
// Host logic mirrors pclomp::NormalDistributionsTransform (ndt_omp.h:119-291, ndt_omp_impl.hpp:85-202)
// around the CUDA kernels; there is no CPU compute path.
#include "handle.cuh"
#include "newton_math.cuh"
#include <cstring>
#include <time.h>
#include <cstdio>
#include <cmath>
#include <algorithm>
#include <new>

using namespace ndtb;

namespace ndtb {

int set_error(ndtb_handle_s* h, int code, const char* what) {
  if (h) {
    h->last_error = what ? what : "";
    if (code == NDTB_ERR_CUDA) {
      const cudaError_t e = cudaGetLastError();
      h->last_error += std::string(" [cuda: ") + cudaGetErrorString(e) + "]";
    }
  }
  return code;
}

#define CKH(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { h->last_error = std::string(#x) + ": " + cudaGetErrorString(e_); return NDTB_ERR_CUDA; } } while (0)

int ensure_eval_buffers(ndtb_handle_s* h, int particles, int n_points) {
  const int blocks = eval_blocks_for(n_points);
  EvalBuffers& b = h->eval;
  if (particles > b.max_particles || blocks > b.max_blocks) {
    const int np = std::max(particles, b.max_particles), nb = std::max(blocks, b.max_blocks);
    cudaFree(b.d_states); cudaFree(b.d_partials); cudaFree(b.d_results); cudaFree(b.d_tickets);
    if (h->h_states) cudaFreeHost(h->h_states);
    if (h->h_results) cudaFreeHost(h->h_results);
    // nothing dangles if an allocation below fails: the capacity is only restored once all of them succeeded
    b.d_states = nullptr; b.d_partials = nullptr; b.d_results = nullptr; b.d_tickets = nullptr; h->h_states = nullptr; h->h_results = nullptr;
    b.max_particles = 0; b.max_blocks = 0;
    CKH(cudaMalloc(&b.d_states, sizeof(EvalState) * np));
    CKH(cudaMalloc(&b.d_partials, sizeof(double) * RED_N * static_cast<size_t>(np) * nb));
    CKH(cudaMalloc(&b.d_results, sizeof(double) * RED_N * np));
    CKH(cudaMalloc(&b.d_tickets, sizeof(unsigned int) * np));
    CKH(cudaMemset(b.d_tickets, 0, sizeof(unsigned int) * np));
    CKH(cudaMallocHost(&h->h_states, sizeof(EvalState) * np));
    CKH(cudaMallocHost(&h->h_results, sizeof(double) * std::max(RED_N * np, 2 * RED_N + 36)));
    std::memset(h->h_results, 0, sizeof(double) * std::max(RED_N * np, 2 * RED_N + 36));
    if (!h->h_flag) { CKH(cudaMallocHost(const_cast<unsigned int**>(&h->h_flag), 64)); *h->h_flag = 0; }
    b.max_particles = np; b.max_blocks = nb;
  }
  return NDTB_OK;
}

int upload_cloud(ndtb_handle_s* h, const float* xyz, size_t n, size_t stride_bytes, float4** d_buf, size_t* alloc, bool* owned) {
  if (!*owned) { *d_buf = nullptr; *alloc = 0; *owned = true; }
  if (n > *alloc) {
    cudaFree(*d_buf);
    *d_buf = nullptr;
    *alloc = 0;
    const size_t cap = n + n / 8 + 256;
    CKH(cudaMalloc(d_buf, cap * sizeof(float4)));
    *alloc = cap;
  }
  if (n == 0) return NDTB_OK;
  return copy_cloud_to_device(h, xyz, n, stride_bytes, *d_buf, h->stream, &h->d_raw, &h->raw_alloc);
}

// Point i of a strided host cloud (already on the device, byte for byte) -> packed float4: x, y, z and, for strides >= 16, the
// fourth float of the point (what the 16-byte rows of the former 2-D copy carried); 0 for packed xyz.
__global__ void __launch_bounds__(256) repack_cloud_kernel(const unsigned char* __restrict__ raw, size_t stride, size_t n, float4* __restrict__ out) {
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n; i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const float* p = reinterpret_cast<const float*>(raw + i * stride);
    float4 v;
    v.x = p[0]; v.y = p[1]; v.z = p[2];
    v.w = (stride >= sizeof(float4)) ? p[3] : 0.0f;
    out[i] = v;
  }
}

// out = T * p for every source point, rows r of the float 4x4: T[4r] x + (T[4r+1] y + (T[4r+2] z + T[4r+3]))  (the evaluation
// order of the host loop this kernel replaced; -fmad=false keeps the two roundings per term).
struct TransformT { float m[12]; };
__global__ void __launch_bounds__(256) transform_cloud_kernel(const float4* __restrict__ src, size_t n, const TransformT T, float* __restrict__ out_xyz) {
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n; i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const float4 p = __ldg(src + i);
#pragma unroll
    for (int r = 0; r < 3; ++r) out_xyz[3 * i + r] = T.m[4 * r] * p.x + (T.m[4 * r + 1] * p.y + (T.m[4 * r + 2] * p.z + T.m[4 * r + 3]));
  }
}

int copy_cloud_to_device(ndtb_handle_s* h, const float* xyz, size_t n, size_t stride_bytes, float4* d_dst, cudaStream_t stream, unsigned char** d_raw,
                         size_t* raw_alloc) {
  if (stride_bytes == sizeof(float4)) {
    CKH(cudaMemcpyAsync(d_dst, xyz, n * sizeof(float4), cudaMemcpyHostToDevice, stream));
    return NDTB_OK;
  }
  if (!(stride_bytes == 12 || stride_bytes >= sizeof(float4)) || (stride_bytes & 3u)) return set_error(h, NDTB_ERR_INVALID, "stride_bytes must be 12 or >= 16, and a multiple of 4");
  const size_t tail = stride_bytes >= sizeof(float4) ? sizeof(float4) : 12;   // bytes of the last point that are read
  const size_t bytes = (n - 1) * stride_bytes + tail;
  if (bytes > *raw_alloc) {
    CKH(cudaStreamSynchronize(stream));
    cudaFree(*d_raw);
    *d_raw = nullptr; *raw_alloc = 0;
    const size_t cap = bytes + bytes / 8 + 4096;
    CKH(cudaMalloc(d_raw, cap));
    *raw_alloc = cap;
  }
  CKH(cudaMemcpyAsync(*d_raw, xyz, bytes, cudaMemcpyHostToDevice, stream));
  const int blocks = static_cast<int>(std::min<size_t>((n + 255) / 256, 148 * 8));
  repack_cloud_kernel<<<blocks, 256, 0, stream>>>(*d_raw, stride_bytes, n, d_dst);
  CKH(cudaGetLastError());
  h->launches += 1;
  return NDTB_OK;
}

int rebuild_grid(ndtb_handle_s* h) {
  if (!h->d_target || h->n_target == 0) return NDTB_ERR_EMPTY;
  cudaEventRecord(h->ev_b[0], h->stream);
  // setMinPointPerVoxel forces values below 3 to 3 in both grids (voxel_grid_covariance_omp.h:214-224, voxel_grid_covariance.h:176-184)
  const int min_pts = std::max(3, h->params.min_points_per_voxel);
  const int rc = build_voxel_map(h->grid, h->d_target, h->n_target, h->params.resolution, min_pts, h->params.min_covar_eigvalue_mult,
                                 h->params.grid_flavour, h->params.is_dense, h->stream);
  // the build has already told the host it is done (polled result slots): the event pair is only read by ndtb_get_timings
  cudaEventRecord(h->ev_b[1], h->stream);
  h->build_time_pending = true;
  h->built_resolution = h->params.resolution; h->built_min_pts = h->params.min_points_per_voxel;
  h->built_flavour = h->params.grid_flavour; h->built_eig = h->params.min_covar_eigvalue_mult;
  if (rc == NDTB_ERR_CUDA) return set_error(h, rc, "voxel map build failed");
  if (rc == NDTB_ERR_OVERFLOW) return set_error(h, rc, "leaf size too small for the input dataset: integer indices would overflow");
  return rc;
}

// Fill EvalState `slot` for pose p (angle tables on the host with libm, like the oracle's host code).
void fill_eval_state(ndtb_handle_s* h, EvalState& s, const double* p, const float* T16, double snap, bool svn);
static void fill_state(ndtb_handle_s* h, int slot, const double* p, const float* T16, double snap, bool svn) {
  fill_eval_state(h, h->h_states[slot], p, T16, snap, svn);
}
void fill_eval_state(ndtb_handle_s* h, EvalState& s, const double* p, const float* T16, double snap, bool svn) {
  std::memset(&s, 0, sizeof(s));
  std::memcpy(s.T, T16, 12 * sizeof(float));
  s.one = 1.0f;
  angle_tables(p, snap, s.jang, s.hang);
  if (svn) gauss_svn(h->params.outlier_ratio, h->params.resolution, &s.d1, &s.d2, &s.d3);
  else gauss_pclomp(h->params.outlier_ratio, h->params.resolution, &s.d1, &s.d2, &s.d3);
  std::memcpy(s.p, p, 6 * sizeof(double));
}

// Sum the CUDA-event durations of the evaluations issued since the last call (needs the stream idle).
static int resolve_eval_events(ndtb_handle_s* h) {
  if (h->ev_used == 0) return NDTB_OK;
  CKH(cudaEventSynchronize(h->ev_pool[h->ev_used - 1]));
  for (size_t i = 0; i + 1 < h->ev_used; i += 2) {
    float ms = 0;
    cudaEventElapsedTime(&ms, h->ev_pool[i], h->ev_pool[i + 1]);
    h->eval_ms_total += ms;
  }
  h->ev_used = 0;
  return NDTB_OK;
}

// ---- alignment session (host side) ---------------------------------------------------------------------------
static void fill_session_cmd(ndtb_handle_s* h, SessionCmd& cmd, const EvalState* st, unsigned long long result_key, unsigned long long flags) {
  std::memset(&cmd, 0, sizeof(cmd));
  if (st) cmd.st = *st;
  cmd.result_key = result_key;
  cmd.flags = flags;
  cmd.key_seq = h->flag_seq;        // result_key == result_key_of(key_seq); picks the result-slot set of a speculating session
  if (flags & SESSION_SPECULATE) {  // what newton_step_warp needs: the constants of the loop and where the loop stands
    cmd.step_max = h->params.step_size; cmd.step_min = h->params.trans_epsilon / 2; cmd.trans_eps = h->params.trans_epsilon;
    cmd.max_iterations = h->params.max_iterations;
    cmd.first = h->spec_carry.first; cmd.a_t = h->spec_carry.a_t; cmd.nr_iterations = static_cast<unsigned long long>(h->spec_carry.nr_iterations);
  }
}

static void session_send(ndtb_handle_s* h, const SessionCmd& cmd) {
  unsigned long long words[SESSION_UNITS] = {0};
  std::memcpy(words, &cmd, sizeof(cmd));
  const unsigned long long key = session_key(++h->mail_seq);
  volatile unsigned long long* mb = reinterpret_cast<volatile unsigned long long*>(h->h_mailbox);
  for (int i = 0; i < SESSION_UNITS; ++i) { mb[2 * i] = words[i]; mb[2 * i + 1] = words[i] ^ key; }
  __sync_synchronize();
}

static inline void poll_backoff(unsigned int spins);

// The device Newton step restates the sincosf of the host's libm (nm_sincosf_glibc).  Checked once per process against the libm
// that is actually running -- through nm_sinf / nm_cosf, i.e. the very calls pose_to_T makes -- on 24 k arguments of every range;
// another libm build (no FMA, another version) fails this check on ~1 % of them and the step then stays on the host.
bool sincosf_port_matches_libm() {
  static const bool ok = [] {
    unsigned int state = 2463534242u;
    auto next = [&state] { state = state * 1664525u + 1013904223u; return state; };
    for (int i = 0; i < 24000; ++i) {
      const unsigned int r = next();
      const double u = static_cast<double>(next() >> 8) * (2.0 / 16777216.0) - 1.0;   // [-1, 1)
      volatile float x = (i % 3 == 0) ? static_cast<float>(u * 119.0) : (i % 3 == 1) ? static_cast<float>(ldexp(u, -static_cast<int>(r % 30u))) : static_cast<float>(u * 3.3);
      const float xa = x;
      const float s = nm_sinf(xa), c = nm_cosf(xa);
      float s2 = 0, c2 = 0;
      if (!nm_sincosf_glibc(xa, &s2, &c2)) return false;
      if (std::memcmp(&s, &s2, 4) != 0 || std::memcmp(&c, &c2, 4) != 0) return false;
    }
    return true;
  }();
  return ok;
}

// One self-validating unit {word, word ^ key} of the speculation record; false when the kernel has left the stream instead.
static bool spec_wait_unit(ndtb_handle_s* h, unsigned long long key_seq, int unit, unsigned long long key, unsigned long long* word) {
  volatile unsigned long long* rec = reinterpret_cast<volatile unsigned long long*>(h->h_spec + (key_seq & 1ull) * SESSION_UNITS);
  unsigned int spins = 0;
  for (;;) {
    const unsigned long long w0 = rec[2 * unit], w1 = rec[2 * unit + 1];
    if ((w0 ^ w1) == key) { *word = w0; return true; }
    poll_backoff(spins);
    if ((++spins & 4095u) == 0 && cudaStreamQuery(h->stream) != cudaErrorNotReady) {
      const unsigned long long x0 = rec[2 * unit], x1 = rec[2 * unit + 1];
      if ((x0 ^ x1) == key) { *word = x0; return true; }
      return false;
    }
  }
}

// The resident kernel runs on by itself (a chain of speculated evaluations nobody will read) or has been lost: wait until it has
// left -- at the latest one idle time-out after its chain ends -- and go on with one launch per evaluation.
static void session_drain(ndtb_handle_s* h) {
  cudaStreamSynchronize(h->stream);
  cudaGetLastError();
  h->session_open = false;
  h->spec_armed = false;
  h->pending_time_key = 0;
}

// Flow control of the device Newton step: the results of every evaluation up to key_seq have been read (the device lets the
// evaluation after next overwrite them only then, align_session_kernel).
static void spec_acknowledge(ndtb_handle_s* h, unsigned long long key_seq) {
  volatile unsigned long long* ack = reinterpret_cast<volatile unsigned long long*>(h->h_spec + SESSION_SPEC_ACK_UNIT);
  ack[0] = key_seq;   // one aligned 8-byte store
  __sync_synchronize();
}

enum { SPEC_RUNNING = 0, SPEC_WAITING = 1, SPEC_DRAINED = 2 };
// After an evaluation whose command carried SESSION_SPECULATE.  `key_seq` numbers the NEXT evaluation, `mine` is the command
// the host would send for it.  SPEC_RUNNING: the device has started exactly that evaluation.  SPEC_WAITING: it declined (the step
// needs the host) and polls the mailbox.  SPEC_DRAINED: it started something else, or its record never came -- session closed.
static int spec_check(ndtb_handle_s* h, unsigned long long key_seq, const SessionCmd* mine) {
  const unsigned long long key = result_key_of(key_seq);
  unsigned long long sw = 0;
  if (!spec_wait_unit(h, key_seq, SESSION_SPEC_STATUS_UNIT, key, &sw)) { h->spec_mismatches += 1; session_drain(h); return SPEC_DRAINED; }
  if ((sw & 0xffull) != 0) { h->spec_declined += 1; return SPEC_WAITING; }   // NEWTON_CONVERGED / NEWTON_HOST / NEWTON_FLOW
  bool match = mine != nullptr;
  unsigned long long theirs[SESSION_CMD_WORDS];
  for (int i = 0; i < SESSION_CMD_WORDS && match; ++i) match = spec_wait_unit(h, key_seq, i, key, &theirs[i]);
  if (match) match = std::memcmp(theirs, mine, sizeof(SessionCmd)) == 0;
  if (match && h->spec_fault_at && ++h->spec_checks == h->spec_fault_at) match = false;
  if (!match) { h->spec_mismatches += 1; session_drain(h); return SPEC_DRAINED; }
  h->mail_seq += 1;   // the device consumed a sequence number for the evaluation it started
  h->spec_accepted += 1;
  return SPEC_RUNNING;
}

// The duration slot of the previous session evaluation is published a little after its results: collect it late.
static void session_collect_time(ndtb_handle_s* h) {
  if (!h->pending_time_key) return;
  volatile unsigned long long* slots = h->pending_time_slots;
  for (int spin = 0; spin < 200000; ++spin) {
    const unsigned long long w0 = slots[2 * RED_N], w1 = slots[2 * RED_N + 1];
    if ((w0 ^ w1) == h->pending_time_key) { h->session_eval_ns += static_cast<double>(w0); break; }
  }
  h->pending_time_key = 0;
}

static void session_begin(ndtb_handle_s* h) {
  static const bool env_off = getenv("NDTB_NO_SESSION") != nullptr;
  h->session_eval_ns = 0;
  if (!h->use_sessions || h->kernel_events || env_off || h->session_open || h->params.use_line_search) return;
  if (h->params.search_method == NDTB_DIRECT26 || h->params.search_method == NDTB_KDTREE) return;  // served by the generic kernel, one launch per evaluation
  const int n = static_cast<int>(h->n_source);
  if (ensure_eval_buffers(h, 1, n) != NDTB_OK) return;
  if (!h->h_mailbox) {
    if (cudaMallocHost(&h->h_mailbox, SESSION_UNITS * sizeof(ulonglong2)) != cudaSuccess ||
        cudaMallocHost(&h->h_spec, SESSION_SPEC_AREA_UNITS * sizeof(ulonglong2)) != cudaSuccess ||
        cudaMalloc(&h->d_cmd, SESSION_DCMD_WORDS * sizeof(unsigned long long)) != cudaSuccess) { cudaGetLastError(); h->use_sessions = false; return; }
    std::memset(h->h_mailbox, 0, SESSION_UNITS * sizeof(ulonglong2));
    std::memset(h->h_spec, 0, SESSION_SPEC_AREA_UNITS * sizeof(ulonglong2));
  }
  // the device copy of the sequence number starts at the value the kernel starts from; time stamp and decline word at zero
  static_assert(SESSION_TSEEN_WORD == SESSION_SEQ_WORD + 1 && SESSION_DECLINE_WORD == SESSION_SEQ_WORD + 2, "one copy initialises the three words");
  h->seq_init[0] = h->mail_seq; h->seq_init[1] = 0; h->seq_init[2] = 0;
#ifdef NDTB_NEWTON_TIMING
  { static bool cleared = false; if (!cleared) { cudaMemset(h->d_cmd + 67, 0, 9 * sizeof(unsigned long long)); cleared = true; } }
#endif
  h->spec_armed = false;
  h->session_spec = h->device_newton && h->direct_solver && !h->has_reg_pose && sincosf_port_matches_libm();
  if (h->session_spec) spec_acknowledge(h, h->flag_seq);   // nothing older than this session is waiting to be read
  if (cudaMemcpyAsync(h->d_cmd + SESSION_SEQ_WORD, h->seq_init, sizeof(h->seq_init), cudaMemcpyHostToDevice, h->stream) != cudaSuccess) {
    cudaGetLastError(); h->use_sessions = false; return;
  }
  const int n_probe = (h->params.search_method == NDTB_DIRECT1) ? 1 : 7;
  const cudaError_t e = launch_align_session(n_probe, h->d_source, n, h->grid.meta, h->grid.recs, h->grid.table, h->h_mailbox, h->d_cmd, h->eval,
                                             h->h_results, h->session_spec ? h->h_spec : nullptr, h->mail_seq, static_cast<unsigned long long>(h->session_idle_ms * 1e6), h->stream);
  if (e != cudaSuccess) { cudaGetLastError(); h->use_sessions = false; return; }  // e.g. the CTAs cannot be co-resident: one-shot launches
  h->session_open = true;
  h->launches += 1;
}

static void session_end(ndtb_handle_s* h) {
  if (!h->session_open) return;
  session_collect_time(h);
  if (h->spec_armed) {
    // the last command allowed the device to go on: it must have declined (the loop is over), else nobody reads what it computes
    h->spec_armed = false;
    if (spec_check(h, static_cast<unsigned long long>(h->flag_seq) + 1, nullptr) != SPEC_WAITING) return;
  }
  SessionCmd quit;
  fill_session_cmd(h, quit, nullptr, 0, SESSION_QUIT);
  session_send(h, quit);
  h->session_open = false;
#ifdef NDTB_NEWTON_TIMING
  {
    static unsigned long long acc[9] = {0}, sessions = 0;
    unsigned long long w[9];
    cudaStreamSynchronize(h->stream);
    cudaMemcpy(w, h->d_cmd + 67, sizeof(w), cudaMemcpyDeviceToHost);
    cudaMemset(h->d_cmd + 67, 0, sizeof(w));
    for (int i = 0; i < 9; ++i) acc[i] += w[i];
    if ((++sessions % 100) == 0 && acc[8])
      fprintf(stderr, "[newton timing, %llu steps] ns/step: load+eliminate %.0f | tri-solves %.0f | step %.0f | sincos %.0f | tables %.0f | T %.0f | compose+release %.0f | barrier->start %.0f\n",
              acc[8], acc[0] / (double)acc[8], acc[1] / (double)acc[8], acc[2] / (double)acc[8], acc[3] / (double)acc[8], acc[4] / (double)acc[8],
              acc[5] / (double)acc[8], acc[6] / (double)acc[8], acc[7] / (double)acc[8]);
  }
#endif
}

void finish_eval(ndtb_handle_s* h, int n, const double* r, const double* p, double* score, double* g, double* H, double* nvtl);

// Bounded back-off of the host poll loops: a pipeline hint per probe (the sibling hyper-thread gets the core's issue slots), and
// once an answer is clearly late (> ~0.1 ms of spinning: a kernel queued behind other work, a debugger, a contended box) the
// thread yields its time slice instead of burning the core -- at most 50 us per probe, so the latency cost stays bounded.
static inline void poll_backoff(unsigned int spins) {
#if defined(__x86_64__) || defined(__i386__)
  __builtin_ia32_pause();
#endif
  if (spins > (1u << 15) && (spins & 255u) == 0) {
    struct timespec ts = {0, 50000};
    nanosleep(&ts, nullptr);
  }
}

// Result slots of the session evaluation numbered h->flag_seq: the handle's own set, or -- with the device Newton step -- one of
// the two sets of the speculation area.
static volatile unsigned long long* session_slots(ndtb_handle_s* h) {
  if (!h->session_spec) return reinterpret_cast<volatile unsigned long long*>(h->h_results);
  return reinterpret_cast<volatile unsigned long long*>(h->h_spec + SESSION_SPEC_RESULT_UNIT0 + (h->flag_seq & 1u) * SESSION_SPEC_RESULT_UNITS);
}

// One pclomp evaluation: upload state, launch, read back the 46-double result, apply the
// regulariser (ndt_omp_impl.hpp:349-382) and NVTL (:384-389) on the host.
static int eval_pclomp(ndtb_handle_s* h, const double* p, const float* T16, bool compute_hessian, double* score, double* g, double* H,
                       double* nvtl) {
  const int n = static_cast<int>(h->n_source);
  int rc = ensure_eval_buffers(h, 1, n);
  if (rc != NDTB_OK) return rc;
  fill_state(h, 0, p, T16, 10e-5, false);
  const int n_probe = (h->params.search_method == NDTB_DIRECT1) ? 1 : 7;
  // event pair from the pool: elapsed times are resolved by resolve_eval_events() after the loop
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  if (h->kernel_events) {
    if (h->ev_used + 2 > h->ev_pool.size()) {
      for (int i = 0; i < 2; ++i) { cudaEvent_t e; CKH(cudaEventCreate(&e)); h->ev_pool.push_back(e); }
    }
    e0 = h->ev_pool[h->ev_used]; e1 = h->ev_pool[h->ev_used + 1];
    h->ev_used += 2;
  }
  if (h->params.search_method == NDTB_DIRECT26 || h->params.search_method == NDTB_KDTREE) {
    // KDTREE: radius search over the leaf means (27 cells + float distance test).  DIRECT26: getNeighborhoodAtPoint with the 26 cells of pcl::getAllNeighborCellIndices (voxel_grid_covariance_omp_impl.hpp:366-411): off the
    // hot path north_star names, served by the generic two-phase kernel (state from device memory, results by one D2H copy)
    if (h->session_open) session_end(h);
    CKH(cudaMemcpyAsync(h->eval.d_states, h->h_states, sizeof(EvalState), cudaMemcpyHostToDevice, h->stream));
    if (e0) cudaEventRecord(e0, h->stream);
    launch_eval(FLAVOUR_PCLOMP, h->params.search_method == NDTB_KDTREE ? 27 : 26, compute_hessian, false, h->d_source, n, h->grid.meta, h->grid.recs, h->grid.table, h->grid.c_mean, h->grid.c_icov,
                h->eval, 1, 0, h->stream);
    CKH(cudaGetLastError());
    if (e1) cudaEventRecord(e1, h->stream);
    double r26[RED_N];
    CKH(cudaMemcpyAsync(h->h_results, h->eval.d_results, sizeof(double) * RED_N, cudaMemcpyDeviceToHost, h->stream));
    CKH(cudaStreamSynchronize(h->stream));
    std::memcpy(r26, h->h_results, sizeof(r26));
    h->launches += 1;
    h->eval_launches += 1;
    finish_eval(h, n, r26, p, score, g, H, nvtl);
    return NDTB_OK;
  }
  static unsigned long long* d_dbg = nullptr;  // development aid: NDTB_DEBUG_TIMING=1 prints in-kernel globaltimer stamps
  static const bool dbg_on = getenv("NDTB_DEBUG_TIMING") != nullptr;
  double rbuf[RED_N];
  for (int attempt = 0;; ++attempt) {
    const unsigned long long key = result_key_of(static_cast<unsigned long long>(++h->flag_seq));
    const bool via_session = h->session_open && compute_hessian && !dbg_on;
    if (via_session) {
      // one command into the resident kernel's mailbox; it answers through the same result slots.  With the device Newton step
      // the kernel may already have started this very evaluation (see align_session_kernel): then there is nothing to send.
      session_collect_time(h);
      const bool speculate = h->session_spec && h->spec_carry.valid && h->direct_solver && !h->has_reg_pose;
      SessionCmd cmd;
      fill_session_cmd(h, cmd, &h->h_states[0], key, speculate ? SESSION_SPECULATE : 0);
      int state = SPEC_WAITING;
      if (h->spec_armed) {
        h->spec_armed = false;
        state = spec_check(h, h->flag_seq, speculate ? &cmd : nullptr);
        if (state == SPEC_DRAINED) continue;   // the session is closed: this evaluation again, as a one-shot launch
      }
      if (state == SPEC_WAITING) session_send(h, cmd);
      // after this evaluation the device's step starts with the loop's convergence test: when that will fire, it declines
      const ndtb_handle_s::SpecCarry& c = h->spec_carry;
      const bool loop_ends = !c.first && (c.nr_iterations > h->params.max_iterations || (c.nr_iterations && std::fabs(c.a_t) < h->params.trans_epsilon));
      h->spec_armed = speculate && !loop_ends;
      h->pending_time_key = key;
      h->pending_time_slots = session_slots(h);
    } else {
      if (h->session_open) session_end(h);  // an evaluation the session kernel does not serve: it leaves, the launch queues behind it
      if (e0) cudaEventRecord(e0, h->stream);
      if (dbg_on && !d_dbg) { cudaMalloc(&d_dbg, 8 * 148 * sizeof(unsigned long long)); }
      if (dbg_on) cudaMemsetAsync(d_dbg, 0, 8 * 148 * sizeof(unsigned long long), h->stream);
      // state by value (kernel parameter); the kernel writes 46 self-validating slots {bits, bits ^ key} straight into
      // pinned host memory, and this thread polls them instead of synchronising the stream
      CKH(launch_eval_pclomp(n_probe, compute_hessian, h->d_source, n, h->grid.meta, h->grid.recs, h->grid.table, h->h_states[0], h->eval,
                             h->h_results, key, h->stream, dbg_on ? d_dbg : nullptr));
      if (e1) cudaEventRecord(e1, h->stream);
      h->launches += 1;
    }
    bool lost = false;
    volatile unsigned long long* slots = via_session ? session_slots(h) : reinterpret_cast<volatile unsigned long long*>(h->h_results);
    unsigned int spins = 0;
    for (int c = RED_N - 1; c >= 0 && !lost; --c) {
      for (;;) {
        const unsigned long long w0 = slots[2 * c], w1 = slots[2 * c + 1];
        if ((w0 ^ w1) == key) { std::memcpy(&rbuf[c], &w0, 8); break; }
        poll_backoff(spins);
        if ((++spins & 4095u) == 0) {  // a stream query every 4096 spins catches a failed launch
          const cudaError_t q = cudaStreamQuery(h->stream);
          if (q != cudaSuccess && q != cudaErrorNotReady) { h->session_open = false; h->last_error = std::string("eval kernel: ") + cudaGetErrorString(q); return NDTB_ERR_CUDA; }
          if (q == cudaSuccess) {
            const unsigned long long x0 = slots[2 * c], x1 = slots[2 * c + 1];
            if ((x0 ^ x1) == key) continue;
            if (via_session && attempt == 0) { lost = true; break; }  // the session kernel left on its idle timeout: one-shot launch
            h->last_error = "eval kernel finished without publishing its result";
            return NDTB_ERR_CUDA;
          }
        }
      }
    }
    if (!lost) {
      if (via_session && h->session_spec) spec_acknowledge(h, h->flag_seq);
      break;
    }
    h->session_open = false;
    h->pending_time_key = 0;
  }
  __sync_synchronize();
  if (dbg_on) {
    std::vector<unsigned long long> st(8 * 148);
    cudaStreamSynchronize(h->stream);
    cudaMemcpy(st.data(), d_dbg, st.size() * 8, cudaMemcpyDeviceToHost);
    const int nb = eval_blocks_pclomp(n, h->eval.free_sms);
    unsigned long long t0 = ~0ull;
    for (int b = 0; b < nb; ++b) t0 = std::min(t0, st[8 * b]);
    unsigned long long mx[6] = {0, 0, 0, 0, 0, 0}, mn[6] = {~0ull, ~0ull, ~0ull, ~0ull, ~0ull, ~0ull};
    for (int b = 0; b < nb; ++b) for (int k = 0; k < 6; ++k) if (st[8 * b + k]) { mx[k] = std::max(mx[k], st[8 * b + k] - t0); mn[k] = std::min(mn[k], st[8 * b + k] - t0); }
    fprintf(stderr, "[dbg n=%d blocks=%d] ns since first block start: start %llu..%llu | main done %llu..%llu | block reduced %llu..%llu | ticket %llu..%llu | final %llu | flag %llu\n",
            n, nb, mn[0], mx[0], mn[1], mx[1], mn[2], mx[2], mn[3], mx[3], mx[4], mx[5]);
  }
  h->eval_launches += 1;
  finish_eval(h, n, rbuf, p, score, g, H, nvtl);
  return NDTB_OK;
}

// Host epilogue of one evaluation: unpack the 46 reduced values, add the regulariser (ndt_omp_impl.hpp:349-382), NVTL (:384-389).
void finish_eval(ndtb_handle_s* h, int n, const double* r, const double* p, double* score, double* g, double* H, double* nvtl) {
  double sc = r[RED_SCORE];
  std::memcpy(g, r + RED_G, 6 * sizeof(double));
  std::memcpy(H, r + RED_H, 36 * sizeof(double));
  const double found = r[RED_FOUND], count = r[RED_COUNT];
  h->last_found = found;
  h->last_hits = n ? count / static_cast<double>(n) : 0.0;
  if (h->has_reg_pose) {
    const float dx = h->reg_translation[0] - static_cast<float>(p[0]);
    const float dy = h->reg_translation[1] - static_cast<float>(p[1]);
    const float sin_yaw = static_cast<float>(sin(p[5])), cos_yaw = static_cast<float>(cos(p[5]));
    const float longitudinal = dy * sin_yaw + dx * cos_yaw;
    const float w = static_cast<float>(static_cast<int>(count));
    const float sf = h->params.regularization_scale_factor;
    const float reg_score = -sf * w * longitudinal * longitudinal;
    sc += reg_score;
    g[0] += static_cast<double>(sf * w * 2.0f * cos_yaw * longitudinal);
    g[1] += static_cast<double>(sf * w * 2.0f * sin_yaw * longitudinal);
    const float h00 = -sf * w * 2.0f * cos_yaw * cos_yaw, h01 = -sf * w * 2.0f * cos_yaw * sin_yaw, h11 = -sf * w * 2.0f * sin_yaw * sin_yaw;
    H[0] += static_cast<double>(h00); H[1] += static_cast<double>(h01); H[7] += static_cast<double>(h11); H[6] += static_cast<double>(h01);
  }
  *score = sc;
  if (nvtl) *nvtl = found != 0 ? r[RED_NEAREST] / found : 0.0;
}

}  // namespace ndtb

extern "C" {

const char* ndtb_version(void) { return "ndtb 0.1 (sm_100a)"; }

int ndtb_default_params(int grid_flavour, ndtb_params* p) {
  if (!p) return NDTB_ERR_INVALID;
  std::memset(p, 0, sizeof(*p));
  p->resolution = 1.0f;            // ndt_omp_impl.hpp:65 / svn_ndt_impl.hpp:62
  p->step_size = 0.1;              // :64
  p->trans_epsilon = 0.1;          // :63
  p->max_iterations = 35;          // :66
  p->search_method = NDTB_DIRECT7; // :67
  p->outlier_ratio = 0.55;         // :54
  p->min_points_per_voxel = 6;     // voxel_grid_covariance_omp.h:195 / voxel_grid_covariance.h:153
  p->min_covar_eigvalue_mult = 0.01;
  p->regularization_scale_factor = 0.0f;
  p->use_line_search = 0;
  p->grid_flavour = grid_flavour;
  p->is_dense = 1;
  return NDTB_OK;
}

int ndtb_create(int cuda_device, ndtb_handle* out) {
  if (!out) return NDTB_ERR_INVALID;
  *out = nullptr;
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0 || cuda_device < 0 || cuda_device >= count) return NDTB_ERR_CUDA;
  if (cudaSetDevice(cuda_device) != cudaSuccess) return NDTB_ERR_CUDA;
  ndtb_handle_s* h = new (std::nothrow) ndtb_handle_s();
  if (!h) return NDTB_ERR_INVALID;
  h->device = cuda_device;
  ndtb_default_params(NDTB_FLAVOUR_PCLOMP, &h->params);
  ndtb_svn_default_params(&h->svn);
  if (cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking) != cudaSuccess) { delete h; return NDTB_ERR_CUDA; }
  for (int i = 0; i < 4; ++i) cudaEventCreate(&h->ev[i]);
  for (int i = 0; i < 2; ++i) cudaEventCreate(&h->ev_b[i]);
  *out = h;
  return NDTB_OK;
}

int ndtb_destroy(ndtb_handle h) {
  if (!h) return NDTB_ERR_INVALID;
  cudaSetDevice(h->device);
  cudaStreamSynchronize(h->stream);
  if (h->target_owned) cudaFree(h->d_target);
  if (h->source_owned) cudaFree(h->d_source);
  if (h->h_stage) cudaFreeHost(h->h_stage);
  cudaFree(h->d_raw);
  free_device_grid(h->grid);
  cudaFree(h->eval.d_states); cudaFree(h->eval.d_partials); cudaFree(h->eval.d_results); cudaFree(h->eval.d_tickets);
  if (h->h_states) cudaFreeHost(h->h_states);
  if (h->h_results) cudaFreeHost(h->h_results);
  if (h->h_flag) cudaFreeHost(const_cast<unsigned int*>(h->h_flag));
  if (h->h_mailbox) cudaFreeHost(h->h_mailbox);
  if (h->h_spec) cudaFreeHost(h->h_spec);
  cudaFree(h->d_cmd);
  for (auto& kv : h->keyframes) cudaFree(kv.second.d_points);
  free_device_grid(h->ds_grid);
  mg_free(h);
  cudaFree(h->d_ds_in); cudaFree(h->d_ds_out);
  for (auto& bs : h->batch_slot) { cudaFree(bs.d_src); if (bs.uploaded) cudaEventDestroy(bs.uploaded); }
  if (h->batch_copy_stream) cudaStreamDestroy(h->batch_copy_stream);
  if (h->batch_h_results) cudaFreeHost(h->batch_h_results);
  for (cudaEvent_t e : h->ev_pool) cudaEventDestroy(e);
  svn_dev_free(h);
  stage_free(h);
  for (int i = 0; i < 4; ++i) if (h->ev[i]) cudaEventDestroy(h->ev[i]);
  for (int i = 0; i < 2; ++i) if (h->ev_b[i]) cudaEventDestroy(h->ev_b[i]);
  cudaStreamDestroy(h->stream);
  delete h;
  return NDTB_OK;
}

const char* ndtb_last_error(ndtb_handle h) { return h ? h->last_error.c_str() : "null handle"; }

int ndtb_get_params(ndtb_handle h, ndtb_params* p) {
  if (!h || !p) return NDTB_ERR_INVALID;
  *p = h->params;
  return NDTB_OK;
}

int ndtb_set_params(ndtb_handle h, const ndtb_params* p) {
  if (!h || !p) return NDTB_ERR_INVALID;
  const bool d26_ok = (p->search_method == NDTB_DIRECT26 || p->search_method == NDTB_KDTREE) && p->grid_flavour == NDTB_FLAVOUR_PCLOMP && !p->use_line_search;
  const bool svn_kd_ok = p->search_method == NDTB_KDTREE && p->grid_flavour == NDTB_FLAVOUR_SVN;   // svn_ndt::NeighborSearchMethod has no DIRECT26 (svn_ndt.h:31-36)
  if (p->search_method != NDTB_DIRECT7 && p->search_method != NDTB_DIRECT1 && !d26_ok && !svn_kd_ok)
    return set_error(h, NDTB_ERR_INVALID, "DIRECT7 / DIRECT1 / KDTREE (both flavours) and DIRECT26 (pclomp flavour; it and pclomp KDTREE without line search) are provided");
  if (!(p->resolution > 0.0f)) return set_error(h, NDTB_ERR_INVALID, "resolution must be positive");
  if (p->grid_flavour != NDTB_FLAVOUR_PCLOMP && p->grid_flavour != NDTB_FLAVOUR_SVN) return set_error(h, NDTB_ERR_INVALID, "bad grid_flavour");
  cudaSetDevice(h->device);
  h->params = *p;
  // setMinPointPerVoxel: values below 3 become 3 in every grid flavour (voxel_grid_covariance_omp.h:214-224,
  // voxel_grid_covariance.h:176-184, multi_voxel_grid_covariance_omp.h), so getMinPointPerVoxel / ndtb_get_params report 3
  if (h->params.min_points_per_voxel < 3) h->params.min_points_per_voxel = 3;
  p = &h->params;
  // setResolution rebuilds only if the value changed and a target is set (ndt_omp.h:135-142)
  const bool changed = h->grid.built && (h->built_resolution != p->resolution || h->built_min_pts != p->min_points_per_voxel ||
                                         h->built_flavour != p->grid_flavour || h->built_eig != p->min_covar_eigvalue_mult);
  if (changed && h->d_target && h->n_target) return rebuild_grid(h);
  return NDTB_OK;
}

int ndtb_set_target(ndtb_handle h, const float* xyz, size_t n, size_t stride_bytes) {
  if (!h) return NDTB_ERR_INVALID;
  cudaSetDevice(h->device);
  if (!xyz || n == 0) { h->n_target = 0; h->grid.built = false; return set_error(h, NDTB_ERR_EMPTY, "empty target cloud"); }
  int rc = upload_cloud(h, xyz, n, stride_bytes, &h->d_target, &h->target_alloc, &h->target_owned);
  if (rc != NDTB_OK) return rc;
  h->n_target = n;
  return rebuild_grid(h);
}

int ndtb_set_target_device(ndtb_handle h, const void* d_xyzw, size_t n) {
  if (!h) return NDTB_ERR_INVALID;
  cudaSetDevice(h->device);
  if (!d_xyzw || n == 0) { h->n_target = 0; h->grid.built = false; return set_error(h, NDTB_ERR_EMPTY, "empty target cloud"); }
  if (h->target_owned) { cudaFree(h->d_target); h->target_alloc = 0; }
  h->d_target = const_cast<float4*>(static_cast<const float4*>(d_xyzw));
  h->target_owned = false;
  h->n_target = n;
  return rebuild_grid(h);
}

int ndtb_set_source(ndtb_handle h, const float* xyz, size_t n, size_t stride_bytes) {
  if (!h) return NDTB_ERR_INVALID;
  cudaSetDevice(h->device);
  if (!xyz || n == 0) { h->n_source = 0; return set_error(h, NDTB_ERR_EMPTY, "empty source cloud"); }
  int rc = upload_cloud(h, xyz, n, stride_bytes, &h->d_source, &h->source_alloc, &h->source_owned);
  if (rc != NDTB_OK) return rc;
  h->n_source = n;
  return NDTB_OK;
}

int ndtb_set_source_device(ndtb_handle h, const void* d_xyzw, size_t n) {
  if (!h) return NDTB_ERR_INVALID;
  cudaSetDevice(h->device);
  if (!d_xyzw || n == 0) { h->n_source = 0; return set_error(h, NDTB_ERR_EMPTY, "empty source cloud"); }
  if (h->source_owned) { cudaFree(h->d_source); h->source_alloc = 0; }
  h->d_source = const_cast<float4*>(static_cast<const float4*>(d_xyzw));
  h->source_owned = false;
  h->n_source = n;
  return NDTB_OK;
}

int ndtb_set_regularization_pose(ndtb_handle h, const float* T16) {
  if (!h) return NDTB_ERR_INVALID;
  h->has_reg_pose = (T16 != nullptr);
  if (T16) { h->reg_translation[0] = T16[3]; h->reg_translation[1] = T16[7]; h->reg_translation[2] = T16[11]; }
  return NDTB_OK;
}

int ndtb_compute_derivatives(ndtb_handle h, const double p[6], const float* T16, int compute_hessian, double* score, double gradient[6],
                             double hessian[36], double* nvtl) {
  if (!h || !p || !score || !gradient || !hessian) return NDTB_ERR_INVALID;
  cudaSetDevice(h->device);
  if (!h->grid.built || h->n_source == 0) return set_error(h, NDTB_ERR_NOT_READY, "target and source must be set");
  if (h->params.grid_flavour != NDTB_FLAVOUR_PCLOMP) return set_error(h, NDTB_ERR_INVALID, "pclomp entry point on an svn-flavour handle");
  float T[16];
  if (T16) std::memcpy(T, T16, sizeof(T)); else pose_to_T(p, T);
  const int rc = eval_pclomp(h, p, T, compute_hessian != 0, score, gradient, hessian, nvtl);
  if (rc != NDTB_OK) return rc;
  return resolve_eval_events(h);
}

// computeTransformation of pclomp::NormalDistributionsTransform (ndt_omp_impl.hpp:85-202, multigrid = false) and of
// pclomp::MultiGridNormalDistributionsTransform (multigrid_ndt_omp_impl.hpp:210-323, multigrid = true).  The two
// loops differ in the evaluation they call, in the iteration cap (`nr_iterations > max` tested before the increment
// vs `>= max` after it) and in when hessian_ is stored.
static int align_impl(ndtb_handle h, bool multigrid, const float guess[16], ndtb_result* res, float* transformation_array, int capacity, float* out_xyz) {
  if (!h || !guess || !res) return NDTB_ERR_INVALID;
  cudaSetDevice(h->device);
  std::memset(res, 0, sizeof(*res));
  if (multigrid) {
    if (!mg_ready(h) || h->n_source == 0) return set_error(h, NDTB_ERR_NOT_READY, "multigrid: targets + ndtb_mg_create_kdtree and a source must be set");
  } else {
    if (!h->grid.built || h->n_source == 0) return set_error(h, NDTB_ERR_NOT_READY, "target and source must be set");
  }
  if (h->params.grid_flavour != NDTB_FLAVOUR_PCLOMP) return set_error(h, NDTB_ERR_INVALID, "pclomp entry point on an svn-flavour handle");
  auto evaluate = [&](const double* pp, const float* T16, bool hess, double* sc, double* gg, double* HH, double* nv) {
    return multigrid ? mg_eval(h, pp, T16, hess, sc, gg, HH, nv) : eval_pclomp(h, pp, T16, hess, sc, gg, HH, nv);
  };
  cudaEventRecord(h->ev[0], h->stream);
  h->eval_ms_total = 0; h->eval_launches = 0;
  h->trace.clear();
  const size_t N = h->n_source;
  const ndtb_params& P = h->params;
  int nr_iterations = 0;
  bool converged = false;
  float final_T[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
  bool guess_is_identity = true;
  for (int i = 0; i < 16; ++i) if (guess[i] != final_T[i]) guess_is_identity = false;
  if (!guess_is_identity) std::memcpy(final_T, guess, sizeof(final_T));
  int n_arr = 0;
  auto push_T = [&]() { if (transformation_array && n_arr < capacity) std::memcpy(transformation_array + 16 * n_arr, final_T, sizeof(final_T)); ++n_arr; };
  push_T();
  double p[6], delta_p[6], g[6], H[36], score = 0, nvtl = 0;
  float eul[3];
  euler_012(final_T, eul);
  p[0] = final_T[3]; p[1] = final_T[7]; p[2] = final_T[11]; p[3] = eul[0]; p[4] = eul[1]; p[5] = eul[2];
  auto push_trace = [&](const double* pp) {
    h->trace.insert(h->trace.end(), pp, pp + 6); h->trace.push_back(score);
    h->trace.insert(h->trace.end(), g, g + 6); h->trace.insert(h->trace.end(), H, H + 36); h->trace.push_back(nvtl);
  };
  // every evaluation of the loop below is served by ONE resident kernel (closed on every way out of this function)
  struct SessionGuard {
    ndtb_handle_s* h;
    ~SessionGuard() { session_end(h); }
  } session_guard{h};
  if (!multigrid) session_begin(h);
  // where the loop stands, for the device Newton step of the resident kernel (eval_pclomp reads it; invalid outside this loop)
  struct CarryGuard {
    ndtb_handle_s* h;
    ~CarryGuard() { h->spec_carry.valid = false; }
  } carry_guard{h};
  h->spec_checks = 0;
  h->spec_carry.a_t = 0; h->spec_carry.nr_iterations = 0; h->spec_carry.first = 1; h->spec_carry.valid = !multigrid;
  int rc = evaluate(p, final_T, true, &score, g, H, &nvtl);
  if (rc != NDTB_OK) return rc;
  push_trace(p);
  while (!converged) {
    double neg_g[6];
    for (int i = 0; i < 6; ++i) neg_g[i] = -g[i];
    // H^-1 (-g): elimination with complete pivoting when the pivots certify a well-conditioned H, else the Jacobi SVD
    if (!(h->direct_solver && gecp6_solve(H, neg_g, delta_p, 1e-6))) svd6_solve(H, neg_g, delta_p);
    double nrm2 = 0;
    for (int i = 0; i < 6; ++i) nrm2 += delta_p[i] * delta_p[i];
    double delta_p_norm = sqrt(nrm2);
    if (delta_p_norm == 0 || delta_p_norm != delta_p_norm) {
      res->transform_probability = N ? score / static_cast<double>(N) : 0.0;
      converged = delta_p_norm == delta_p_norm;
      break;
    }
    for (int i = 0; i < 6; ++i) delta_p[i] = delta_p[i] / delta_p_norm;
    // computeStepLengthMT (ndt_omp_impl.hpp:915-1079)
    {
      const double phi_0 = -score;
      double d_phi_0 = 0;
      for (int i = 0; i < 6; ++i) d_phi_0 += g[i] * delta_p[i];
      d_phi_0 = -d_phi_0;
      double a_t;
      if (d_phi_0 >= 0 && d_phi_0 == 0) {
        a_t = 0;
      } else {
        if (d_phi_0 >= 0) { d_phi_0 *= -1; for (int i = 0; i < 6; ++i) delta_p[i] *= -1; }
        const int max_step_iterations = 10;
        int step_iterations = 0;
        const double mu = 1.e-4, nu = 0.9;
        const double step_max = P.step_size, step_min = P.trans_epsilon / 2;
        double a_l = 0, a_u = 0;
        double f_l = mt_psi(a_l, phi_0, phi_0, d_phi_0, mu), g_l = mt_dpsi(d_phi_0, d_phi_0, mu);
        double f_u = mt_psi(a_u, phi_0, phi_0, d_phi_0, mu), g_u = mt_dpsi(d_phi_0, d_phi_0, mu);
        bool interval_converged = (step_max - step_min) < 0, open_interval = true;
        a_t = delta_p_norm;
        a_t = std::min(a_t, step_max);
        a_t = std::max(a_t, step_min);
        double x_t[6];
        for (int i = 0; i < 6; ++i) x_t[i] = p[i] + delta_p[i] * a_t;
        pose_to_T(x_t, final_T);
        h->spec_carry.a_t = a_t; h->spec_carry.nr_iterations = nr_iterations; h->spec_carry.first = 0;
        rc = evaluate(x_t, final_T, true, &score, g, H, &nvtl);
        if (rc != NDTB_OK) return rc;
        push_trace(x_t);
        if (P.use_line_search) {
          h->spec_carry.valid = false;
          double phi_t = -score, d_phi_t = 0;
          for (int i = 0; i < 6; ++i) d_phi_t += g[i] * delta_p[i];
          d_phi_t = -d_phi_t;
          double psi_t = mt_psi(a_t, phi_t, phi_0, d_phi_0, mu), d_psi_t = mt_dpsi(d_phi_t, d_phi_0, mu);
          while (!interval_converged && step_iterations < max_step_iterations && !(psi_t <= 0 && d_phi_t <= -nu * d_phi_0)) {
            if (open_interval) a_t = mt_trial_value(a_l, f_l, g_l, a_u, f_u, g_u, a_t, psi_t, d_psi_t);
            else a_t = mt_trial_value(a_l, f_l, g_l, a_u, f_u, g_u, a_t, phi_t, d_phi_t);
            a_t = std::min(a_t, step_max);
            a_t = std::max(a_t, step_min);
            for (int i = 0; i < 6; ++i) x_t[i] = p[i] + delta_p[i] * a_t;
            pose_to_T(x_t, final_T);
            rc = evaluate(x_t, final_T, false, &score, g, H, &nvtl);  // score and gradient only (:1031)
            if (rc != NDTB_OK) return rc;
            phi_t = -score; d_phi_t = 0;
            for (int i = 0; i < 6; ++i) d_phi_t += g[i] * delta_p[i];
            d_phi_t = -d_phi_t;
            psi_t = mt_psi(a_t, phi_t, phi_0, d_phi_0, mu);
            d_psi_t = mt_dpsi(d_phi_t, d_phi_0, mu);
            if (open_interval && (psi_t <= 0 && d_psi_t >= 0)) {
              open_interval = false;
              f_l = f_l + phi_0 - mu * d_phi_0 * a_l; g_l = g_l + mu * d_phi_0;
              f_u = f_u + phi_0 - mu * d_phi_0 * a_u; g_u = g_u + mu * d_phi_0;
            }
            if (open_interval) interval_converged = mt_update_interval(a_l, f_l, g_l, a_u, f_u, g_u, a_t, psi_t, d_psi_t);
            else interval_converged = mt_update_interval(a_l, f_l, g_l, a_u, f_u, g_u, a_t, phi_t, d_phi_t);
            step_iterations++;
          }
        }
        if (step_iterations) {  // the loop evaluated without Hessians: computeHessian, all FP64 (:1073)
          if (multigrid) {  // computeHessian (multigrid_ndt_omp_impl.hpp:701-801): the FP64 pair maths again, Hessian kept
            double sc2, g2[6], nv2;
            const bool reg = h->has_reg_pose;
            h->has_reg_pose = false;  // computeHessian carries no regulariser
            rc = mg_eval(h, x_t, final_T, true, &sc2, g2, H, &nv2);
            h->has_reg_pose = reg;
          } else {
            rc = compute_hessian_f64(h, x_t, final_T, H);
          }
          if (rc != NDTB_OK) return rc;
          push_trace(x_t);
        }
      }
      delta_p_norm = a_t;
    }
    for (int i = 0; i < 6; ++i) delta_p[i] *= delta_p_norm;
    push_T();
    for (int i = 0; i < 6; ++i) p[i] = p[i] + delta_p[i];
    if (multigrid) {
      nr_iterations++;
      if (nr_iterations >= P.max_iterations || (nr_iterations && (fabs(delta_p_norm) < P.trans_epsilon))) converged = true;  // :300-304
    } else {
      if (nr_iterations > P.max_iterations || (nr_iterations && (fabs(delta_p_norm) < P.trans_epsilon))) converged = true;
      nr_iterations++;
    }
    if (converged) {
      res->transform_probability = N ? score / static_cast<double>(N) : 0.0;
      std::memcpy(res->hessian, H, sizeof(res->hessian));
    }
  }
  std::memcpy(res->pose, final_T, sizeof(final_T));
  res->nearest_voxel_transformation_likelihood = nvtl;
  res->iteration_num = nr_iterations;
  res->converged = converged ? 1 : 0;
  res->transformation_array_len = std::min(n_arr, capacity > 0 ? capacity : 0);
  res->evaluations = h->eval_launches;
  // the resident kernel is told to leave BEFORE anything else is queued on the stream: a copy queued behind it would
  // only complete once the kernel's idle time-out fires
  session_end(h);
  if (out_xyz) {
    // `output` of pcl::Registration::align: the source transformed by the final transformation.  Transformed on the device
    // (same float operations in the same order as the host loop it replaces), packed xyz, one D2H copy of 12 N bytes.
    const size_t bytes = N * 3 * sizeof(float);
    if (bytes > h->raw_alloc) {
      cudaFree(h->d_raw);
      h->d_raw = nullptr; h->raw_alloc = 0;
      CKH(cudaMalloc(&h->d_raw, bytes + bytes / 8 + 4096));
      h->raw_alloc = bytes + bytes / 8 + 4096;
    }
    TransformT Tk;
    std::memcpy(Tk.m, final_T, 12 * sizeof(float));
    transform_cloud_kernel<<<static_cast<int>(std::min<size_t>((N + 255) / 256, 148 * 8)), 256, 0, h->stream>>>(h->d_source, N, Tk, reinterpret_cast<float*>(h->d_raw));
    CKH(cudaGetLastError());
    h->launches += 1;
    CKH(cudaMemcpyAsync(out_xyz, h->d_raw, bytes, cudaMemcpyDeviceToHost, h->stream));
    CKH(cudaStreamSynchronize(h->stream));
  }
  cudaEventRecord(h->ev[1], h->stream);
  h->align_time_pending = true;  // resolved by ndtb_get_timings: the caller does not wait for the resident kernel to leave
  if (h->session_eval_ns > 0) h->eval_ms_total = static_cast<float>(h->session_eval_ns * 1e-6);  // device-side (globaltimer) durations
  return resolve_eval_events(h);
}

int ndtb_align(ndtb_handle h, const float guess[16], ndtb_result* res, float* transformation_array, int capacity, float* out_xyz) {
  return align_impl(h, false, guess, res, transformation_array, capacity, out_xyz);
}

int ndtb_mg_align(ndtb_handle h, const float guess[16], ndtb_result* res, float* transformation_array, int capacity, float* out_xyz) {
  return align_impl(h, true, guess, res, transformation_array, capacity, out_xyz);
}

int ndtb_get_trace(ndtb_handle h, double* trace, int capacity, int* count) {
  if (!h) return NDTB_ERR_INVALID;
  const int n = static_cast<int>(h->trace.size() / 50);
  if (count) *count = n;
  if (trace && capacity > 0) std::memcpy(trace, h->trace.data(), sizeof(double) * 50 * std::min(n, capacity));
  return NDTB_OK;
}

int ndtb_get_voxel_count(ndtb_handle h, size_t* count, int min_b[3], int max_b[3], int div_b[3]) {
  if (!h) return NDTB_ERR_INVALID;
  if (!h->grid.built) return set_error(h, NDTB_ERR_NOT_READY, "no voxel map");
  cudaSetDevice(h->device);
  if (count) {
    if (h->params.grid_flavour == NDTB_FLAVOUR_SVN) *count = h->grid.n_valid; else *count = h->grid.V;
  }
  for (int a = 0; a < 3; ++a) {
    if (min_b) min_b[a] = h->grid.meta.min_b[a];
    if (max_b) max_b[a] = h->grid.meta.max_b[a];
    if (div_b) div_b[a] = h->grid.meta.div_b[a];
  }
  return NDTB_OK;
}

int ndtb_get_voxels(ndtb_handle h, int64_t* idx, int32_t* nr_points, double* mean, double* cov, double* icov, double* evecs, double* evals) {
  if (!h) return NDTB_ERR_INVALID;
  if (!h->grid.built) return set_error(h, NDTB_ERR_NOT_READY, "no voxel map");
  cudaSetDevice(h->device);
  const size_t V = h->grid.V;
  std::vector<uint32_t> key(V); std::vector<int32_t> n(V);
  std::vector<double> m(3 * V), c(9 * V), ic(9 * V), ev(9 * V), el(3 * V);
  CKH(cudaMemcpy(key.data(), h->grid.c_key, V * 4, cudaMemcpyDeviceToHost));
  CKH(cudaMemcpy(n.data(), h->grid.c_n, V * 4, cudaMemcpyDeviceToHost));
  CKH(cudaMemcpy(m.data(), h->grid.c_mean, V * 24, cudaMemcpyDeviceToHost));
  CKH(cudaMemcpy(c.data(), h->grid.c_cov, V * 72, cudaMemcpyDeviceToHost));
  CKH(cudaMemcpy(ic.data(), h->grid.c_icov, V * 72, cudaMemcpyDeviceToHost));
  CKH(cudaMemcpy(ev.data(), h->grid.c_evecs, V * 72, cudaMemcpyDeviceToHost));
  CKH(cudaMemcpy(el.data(), h->grid.c_evals, V * 24, cudaMemcpyDeviceToHost));
  size_t o = 0;
  for (size_t v = 0; v < V; ++v) {
    if (h->params.grid_flavour == NDTB_FLAVOUR_SVN && n[v] == -2) continue;  // erased leaf
    if (idx) idx[o] = key[v];
    if (nr_points) nr_points[o] = n[v];
    if (mean) std::memcpy(mean + 3 * o, &m[3 * v], 24);
    if (cov) std::memcpy(cov + 9 * o, &c[9 * v], 72);
    if (icov) std::memcpy(icov + 9 * o, &ic[9 * v], 72);
    if (evecs) std::memcpy(evecs + 9 * o, &ev[9 * v], 72);
    if (evals) std::memcpy(evals + 3 * o, &el[3 * v], 24);
    ++o;
  }
  return NDTB_OK;
}

int ndtb_get_last_hits(ndtb_handle h, double* mean_hits_per_point, double* points_with_neighbours) {
  if (!h) return NDTB_ERR_INVALID;
  if (mean_hits_per_point) *mean_hits_per_point = h->last_hits;
  if (points_with_neighbours) *points_with_neighbours = h->last_found;
  return NDTB_OK;
}

int ndtb_get_timings(ndtb_handle h, float* build_ms, float* align_ms, float* eval_kernel_ms_total, int* eval_kernel_launches) {
  if (!h) return NDTB_ERR_INVALID;
  if (h->build_time_pending) {
    h->build_time_pending = false;
    if (cudaEventSynchronize(h->ev_b[1]) == cudaSuccess) cudaEventElapsedTime(&h->build_ms, h->ev_b[0], h->ev_b[1]);
  }
  if (h->align_time_pending) {
    h->align_time_pending = false;
    if (cudaEventSynchronize(h->ev[1]) == cudaSuccess) cudaEventElapsedTime(&h->align_ms, h->ev[0], h->ev[1]);
  }
  if (build_ms) *build_ms = h->build_ms;
  if (align_ms) *align_ms = h->align_ms;
  if (eval_kernel_ms_total) *eval_kernel_ms_total = h->eval_ms_total;
  if (eval_kernel_launches) *eval_kernel_launches = h->eval_launches;
  return NDTB_OK;
}

int ndtb_get_launch_count(ndtb_handle h, uint64_t* launches) {
  if (!h || !launches) return NDTB_ERR_INVALID;
  *launches = h->launches + h->grid.launches;
  return NDTB_OK;
}

int ndtb_time_eval_kernel(ndtb_handle h, const double p[6], const float* T16, int reps, float* mean_us) {
  if (!h || !p || !mean_us || reps < 1) return NDTB_ERR_INVALID;
  cudaSetDevice(h->device);
  if (!h->grid.built || h->n_source == 0) return set_error(h, NDTB_ERR_NOT_READY, "target and source must be set");
  if (h->params.grid_flavour != NDTB_FLAVOUR_PCLOMP) return set_error(h, NDTB_ERR_INVALID, "pclomp entry point on an svn-flavour handle");
  float Tloc[16];
  if (T16) std::memcpy(Tloc, T16, sizeof(Tloc)); else pose_to_T(p, Tloc);
  const int n = static_cast<int>(h->n_source);
  int rc = ensure_eval_buffers(h, 1, n);
  if (rc != NDTB_OK) return rc;
  fill_state(h, 0, p, Tloc, 10e-5, false);
  if (h->params.search_method == NDTB_DIRECT26 || h->params.search_method == NDTB_KDTREE) return set_error(h, NDTB_ERR_INVALID, "ndtb_time_eval_kernel times the DIRECT1 / DIRECT7 kernel");
  const int n_probe = (h->params.search_method == NDTB_DIRECT1) ? 1 : 7;
  for (int warm = 0; warm < 2; ++warm)
    CKH(launch_eval_pclomp(n_probe, true, h->d_source, n, h->grid.meta, h->grid.recs, h->grid.table, h->h_states[0], h->eval, h->h_results, 1ull, h->stream));
  CKH(cudaEventRecord(h->ev[2], h->stream));
  for (int r = 0; r < reps; ++r)
    CKH(launch_eval_pclomp(n_probe, true, h->d_source, n, h->grid.meta, h->grid.recs, h->grid.table, h->h_states[0], h->eval, h->h_results, 1ull, h->stream));
  CKH(cudaEventRecord(h->ev[3], h->stream));
  CKH(cudaEventSynchronize(h->ev[3]));
  float ms = 0;
  CKH(cudaEventElapsedTime(&ms, h->ev[2], h->ev[3]));
  h->launches += static_cast<uint64_t>(reps) + 2;
  *mean_us = ms * 1000.0f / static_cast<float>(reps);
  return NDTB_OK;
}

int ndtb_set_profiling(ndtb_handle h, int per_kernel_events) {
  if (!h) return NDTB_ERR_INVALID;
  h->kernel_events = per_kernel_events != 0;
  return NDTB_OK;
}

int ndtb_hessian_to_covariance(const double hessian[36], double covariance_gtsam[36]) {
  if (!hessian || !covariance_gtsam) return NDTB_ERR_INVALID;
  double A[36], inv[36];
  for (int i = 0; i < 36; ++i) A[i] = hessian[i];
  for (int i = 0; i < 6; ++i) A[7 * i] = A[7 * i] + 1e-6;
  for (int j = 0; j < 6; ++j) {
    double e[6] = {0, 0, 0, 0, 0, 0}, x[6];
    e[j] = 1.0;
    const bool ok = gecp6_solve(A, e, x, 0.0);
    for (int i = 0; i < 6; ++i) inv[6 * i + j] = ok ? -x[i] : NAN;
  }
  // reorderCovarianceForGTSAM: C_rr -> top-left, C_tt -> bottom-right, the two cross blocks stay where they are
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 3; ++c) {
      covariance_gtsam[6 * (3 + r) + (3 + c)] = inv[6 * r + c];
      covariance_gtsam[6 * r + c] = inv[6 * (3 + r) + (3 + c)];
      covariance_gtsam[6 * r + (3 + c)] = inv[6 * r + (3 + c)];
      covariance_gtsam[6 * (3 + r) + c] = inv[6 * (3 + r) + c];
    }
  return NDTB_OK;
}

int ndtb_newton_direction(const double H[36], const double b[6], int method, double x[6], int* used_direct) {
  if (!H || !b || !x) return NDTB_ERR_INVALID;
  const bool direct = method != 0 && gecp6_solve(H, b, x, 1e-6);
  if (!direct) svd6_solve(H, b, x);
  if (used_direct) *used_direct = direct ? 1 : 0;
  return NDTB_OK;
}

int ndtb_set_option(ndtb_handle h, const char* name, double value) {
  if (!h || !name) return NDTB_ERR_INVALID;
  const std::string k(name);
  if (k == "direct_solver") h->direct_solver = value != 0;           // 0: always the Jacobi SVD (Eigen::JacobiSVD restatement)
  else if (k == "sessions") h->use_sessions = value != 0;            // 0: one kernel launch per evaluation
  else if (k == "session_idle_timeout_ms") h->session_idle_ms = value > 0 ? value : 50.0;
  else if (k == "session_free_sms") {   // the one-CTA-per-SM evaluation kernels (resident session included) leave this many SMs free
    if (value < 0 || value > 140) return set_error(h, NDTB_ERR_INVALID, "session_free_sms must be in [0, 140]");
    if (h->session_open) return set_error(h, NDTB_ERR_NOT_READY, "an alignment session is running");
    h->eval.free_sms = static_cast<int>(value);
  }
  else if (k == "device_newton") h->device_newton = value != 0;       // 0: every Newton step of a session on the host (one PCIe round trip per evaluation)
  else if (k == "device_newton_fault_at") h->spec_fault_at = static_cast<int>(value);  // test hook: that comparison of every align "fails"
  else if (k == "batch_interleave") h->batch_interleave = value != 0;  // 0: ndtb_align_batch aligns one scan after the other
  else return set_error(h, NDTB_ERR_INVALID, "unknown option");
  return NDTB_OK;
}

int ndtb_get_session_stats(ndtb_handle h, unsigned long long stats[4]) {
  if (!stats) return NDTB_ERR_INVALID;
  stats[0] = h ? h->spec_accepted : 0; stats[1] = h ? h->spec_declined : 0; stats[2] = h ? h->spec_mismatches : 0;   // h may be NULL: stats[3] only
  stats[3] = sincosf_port_matches_libm() ? 1ull : 0ull;
  return NDTB_OK;
}

int ndtb_get_stream(ndtb_handle h, void** cuda_stream) {
  if (!h || !cuda_stream) return NDTB_ERR_INVALID;
  *cuda_stream = h->stream;
  return NDTB_OK;
}

}  // extern "C"
