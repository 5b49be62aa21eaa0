// SVN-NDT stages around the derivative kernel, all on the device (svn_ndt::SvnNormalDistributionsTransform::align,
// extern/svn_ndt/include/svn_ndt_impl.hpp:675-964):
//   svn_state_kernel : particle pose -> evaluation state of stage 1 (:761-767): float transform, p = (translation, rpy),
//                      angle tables with the 1e-7 snap (:255-334)
//   stein_kernel     : stages 2 and 3 as ONE kernel, grid = K blocks (particle k), threads stride over l
//       xi_lk = Logmap(P_l^-1 P_k); k = exp(-|xi|^2 / h); grad_k = k * (-2/h) * xi                                  :214-244
//       phi_k = (1/K) sum_l [k * Perm g_l + grad_k];  H_k = (1/K) sum_l [k^2 Perm H_l Perm + grad_k grad_k^T] + 1e-6 I   :795-815
//       update_k = LDLT(H_k).solve(-phi_k)  :833-838 ;  P_k <- P_k.retract(step * update_k)  :848-855
//     and, in the last block to finish, the mean pose prior.retract(mean_k Logmap(prior^-1 P_k)) and the stop test
//     || Logmap(mean_prev^-1 mean) || < threshold (:865-898), published to a pinned host slot.
//     The records {score, g, H} of particle l are read from this rank's record window once the particle's flag carries the
//     iteration's stamp: they were stored there -- by the last block of the derivative kernel of whichever GPU owns particle l,
//     over NVLink -- with a system-scope release (derivatives.cu, SvnExchange).
// All FP64.  The sum over l is a fixed-shape tree (strided partial sums, warp shuffles, warps in order): deterministic, it
// differs from the reference's serial sum only by FP64 rounding; every rank computes the same bits.
#include "svn_stein.cuh"
#include "se3_math.cuh"
#include <cstring>
#include <string>
#include <vector>

namespace ndtb {

constexpr int STEIN_THREADS = 128;
constexpr int STEIN_N = 42;  // phi(6) + H(36)
constexpr unsigned long long SVN_LOCAL_STAMP = 1ull << 63;
constexpr unsigned long long SVN_PEER_TIMEOUT_NS = 4000000000ull;  // a peer that does not deliver within 4 s is reported, not waited for

__device__ __forceinline__ unsigned long long svn_globaltimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// Stage-1 state of the rank's own particles: slot s <- particle k = first + s * stride (svn_ndt_impl.hpp:761-767, 255-334).
__global__ void __launch_bounds__(128) svn_state_kernel(int K, int first, int stride, const double* __restrict__ poses, EvalState* __restrict__ states,
                                                        double d1, double d2, double d3, const int* __restrict__ done) {
  if (done && *done) return;
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  const int k = first + s * stride;
  if (k >= K) return;
  const double* M = poses + 16 * k;
  EvalState st;
#pragma unroll
  for (int i = 0; i < 12; ++i) st.T[i] = static_cast<float>(M[i]);   // particles[k].matrix().cast<float>()
  double R[9], rpy[3], p[6];
#pragma unroll
  for (int r = 0; r < 3; ++r)
#pragma unroll
    for (int c = 0; c < 3; ++c) R[3 * r + c] = M[4 * r + c];
  rot_rpy(R, rpy);
  p[0] = M[3]; p[1] = M[7]; p[2] = M[11]; p[3] = rpy[0]; p[4] = rpy[1]; p[5] = rpy[2];
  angle_tables(p, 1e-7, st.jang, st.hang);
  st.one = 1.0f;
  st.d1 = d1; st.d2 = d2; st.d3 = d3;
#pragma unroll
  for (int i = 0; i < 6; ++i) st.p[i] = p[i];
  states[s] = st;
}

struct SteinMean {               // mean pose + stop test + status slot (all optional: prior == nullptr disables the tail)
  const double* prior;           // [16]
  const double* mean_prev;       // [16]
  double* mean_out;              // [16]
  double stop_threshold;
  int* done;                     // set to 1 when the stop test fires
  unsigned int* ticket;
  volatile unsigned long long* h_status;  // pinned host slots
  unsigned long long status_key;
  int iter;                      // 0-based iteration index
  unsigned long long* wait_ns;   // accumulated record wait of block 0
};

// Shared memory of one block: the records, kernel values and kernel gradients of a tile of STEIN_THREADS particles l.
struct SteinTile {
  double rec[STEIN_THREADS][STEIN_N];   // {g (GTSAM order), H (GTSAM order)} of particle l, guards applied
  double kg[STEIN_THREADS][6];          // grad_k of the RBF kernel
  double kv[STEIN_THREADS];             // k(P_l, P_k)
  unsigned char use[STEIN_THREADS];     // bit 0: pair finite, bit 1: gradient finite, bit 2: Hessian finite
};

__global__ void __launch_bounds__(STEIN_THREADS) stein_kernel(int K, const double* __restrict__ poses, const double* rec /* [K * vshards][43] window half */,
                                                              const unsigned long long* flags /* [K * vshards] or nullptr */, unsigned long long stamp,
                                                              double kernel_h, double step_size, double* __restrict__ updates,
                                                              double* __restrict__ new_poses, const SteinMean mm, int vshards) {
  if (mm.done && *mm.done) return;  // speculatively queued iteration after the align converged
  extern __shared__ __align__(16) unsigned char stein_raw[];   // sizeof(SteinTile): above the 48 KB static limit
  SteinTile& tl = *reinterpret_cast<SteinTile*>(stein_raw);
  __shared__ double s_acc[STEIN_N];
  double (*s_xi)[6] = tl.kg;                                    // the tail reuses the kernel-gradient rows
  __shared__ bool s_last;
  __shared__ int s_timeout;
  const int k = blockIdx.x;
  const int tid = threadIdx.x;
  if (tid == 0) s_timeout = 0;
  Pose Pk;
  pose_from16(poses + 16 * k, Pk);
  const unsigned long long t_wait0 = (flags && blockIdx.x == 0 && tid == 0) ? svn_globaltimer() : 0ull;
  unsigned long long waited = 0;
  // component t of {phi, H}: its own accumulator, summed over l IN PARTICLE ORDER like the reference's loop (:795-815)
  double acc = 0.0;
  const int ci = (tid < 6) ? tid : (tid - 6) / 6, cj = (tid < 6) ? 0 : (tid - 6) % 6;
  __syncthreads();
  for (int base = 0; base < K; base += STEIN_THREADS) {
    const int l = base + tid;
    // ---- phase 1, one thread per particle l: its record (shards added in order, guards), kernel value and gradient ---------
    if (l < K) {
      double gl[6], Hl[36];
#pragma unroll
      for (int i = 0; i < 6; ++i) gl[i] = 0.0;
#pragma unroll
      for (int i = 0; i < 36; ++i) Hl[i] = 0.0;
      for (int sh = 0; sh < vshards; ++sh) {
        const int row = l * vshards + sh;
        if (flags) {
          // acquire the record: stored by the derivative kernel of the GPU that owns the unit (possibly over NVLink)
          unsigned long long seen;
          unsigned int spins = 0;
          unsigned long long t0 = 0;
          for (;;) {
            asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(seen) : "l"(flags + row) : "memory");
            if (seen == stamp) break;
            if ((++spins & 1023u) == 0) {
              const unsigned long long now = svn_globaltimer();
              if (!t0) t0 = now;
              if (now - t0 > SVN_PEER_TIMEOUT_NS) { s_timeout = 1; break; }
            }
          }
        }
        const double* rl = rec + static_cast<size_t>(SVN_REC) * row;
        if (vshards == 1) {
#pragma unroll
          for (int i = 0; i < 6; ++i) gl[i] = __ldcg(rl + 1 + i);
#pragma unroll
          for (int i = 0; i < 36; ++i) Hl[i] = __ldcg(rl + 7 + i);
        } else {
#pragma unroll
          for (int i = 0; i < 6; ++i) gl[i] += __ldcg(rl + 1 + i);
#pragma unroll
          for (int i = 0; i < 36; ++i) Hl[i] += __ldcg(rl + 7 + i);
        }
      }
      if (t_wait0 && base == 0) waited = svn_globaltimer() - t_wait0;
      if (vshards > 1) {
        // particle l's record = the sum, in shard order, of its unit records; the guards of svn_ndt_impl.hpp:650-663 follow the sum
#pragma unroll
        for (int i = 0; i < 6; ++i) Hl[7 * i] += 1e-6;                     // H += 1e-6 I  (:650-653)
        bool gbad = false, hbad = false;
#pragma unroll
        for (int i = 0; i < 6; ++i) gbad = gbad || !isfinite(gl[i]);
#pragma unroll
        for (int i = 0; i < 36; ++i) hbad = hbad || !isfinite(Hl[i]);
        if (gbad) {                                                         // non-finite gradient -> zero (:656-659)
#pragma unroll
          for (int i = 0; i < 6; ++i) gl[i] = 0.0;
        }
        if (hbad) {                                                         // non-finite Hessian -> identity (:660-663)
#pragma unroll
          for (int i = 0; i < 36; ++i) Hl[i] = (i % 7 == 0) ? 1.0 : 0.0;
        }
      }
      bool gfin = true, hfin = true;
#pragma unroll
      for (int i = 0; i < 6; ++i) gfin = gfin && isfinite(gl[i]);
#pragma unroll
      for (int i = 0; i < 36; ++i) hfin = hfin && isfinite(Hl[i]);
      // NDT (t, r) -> GTSAM (r, t) order  svn_ndt_impl.hpp:731-735: component i of GTSAM is component (i + 3) % 6 of NDT
#pragma unroll
      for (int i = 0; i < 6; ++i) tl.rec[tid][i] = gl[(i + 3) % 6];
#pragma unroll
      for (int i = 0; i < 6; ++i)
#pragma unroll
        for (int j = 0; j < 6; ++j) tl.rec[tid][6 + 6 * i + j] = Hl[6 * ((i + 3) % 6) + (j + 3) % 6];
      double k_val, k_grad[6];
      if (kernel_h <= 1e-12) {
        k_val = (l == k) ? 1.0 : 0.0;
#pragma unroll
        for (int i = 0; i < 6; ++i) k_grad[i] = 0.0;
      } else {
        Pose Pl;
        pose_from16(poses + 16 * l, Pl);
        double xi[6];
        pose_logmap(pose_between(Pl, Pk), xi);                              // xi_lk = Logmap(P_l^-1 P_k)   :214-244
        double sq = 0.0;
#pragma unroll
        for (int i = 0; i < 6; ++i) sq += xi[i] * xi[i];
        k_val = exp(-sq / kernel_h);
        const double f = k_val * (-2.0 / kernel_h);
#pragma unroll
        for (int i = 0; i < 6; ++i) k_grad[i] = f * xi[i];
      }
      bool fin = isfinite(k_val);
#pragma unroll
      for (int i = 0; i < 6; ++i) fin = fin && isfinite(k_grad[i]);
      tl.kv[tid] = k_val;
#pragma unroll
      for (int i = 0; i < 6; ++i) tl.kg[tid][i] = k_grad[i];
      tl.use[tid] = static_cast<unsigned char>((fin ? 1 : 0) | (gfin ? 2 : 0) | (hfin ? 4 : 0));
    }
    __syncthreads();
    // ---- phase 2, one thread per component: the sum over the tile's particles, in particle order ---------------------------
    if (tid < STEIN_N) {
      const int cnt = min(STEIN_THREADS, K - base);
      for (int q = 0; q < cnt; ++q) {
        const unsigned char u = tl.use[q];
        if (!(u & 1)) continue;
        const double kv = tl.kv[q];
        if (tid < 6) {
          if (u & 2) acc += kv * tl.rec[q][tid];
          acc += tl.kg[q][tid];
        } else {
          if (u & 4) acc += (kv * kv) * tl.rec[q][tid];
          acc += tl.kg[q][ci] * tl.kg[q][cj];
        }
      }
    }
    __syncthreads();
  }
  if (tid < STEIN_N) s_acc[tid] = acc;
  __syncthreads();
  if (tid == 0) {
    if (t_wait0 && mm.wait_ns) *mm.wait_ns += waited;
    double phi[6], Ht[36];
    for (int i = 0; i < STEIN_N; ++i) {
      const double v = s_acc[i];
      if (i < 6) phi[i] = v / static_cast<double>(K); else Ht[i - 6] = v / static_cast<double>(K);
    }
    for (int i = 0; i < 6; ++i) Ht[7 * i] += 1e-6;
    bool hfin = true;
    for (int i = 0; i < 36; ++i) hfin = hfin && isfinite(Ht[i]);
    double nphi[6], upd[6];
    for (int i = 0; i < 6; ++i) nphi[i] = -phi[i];
    const bool ok = ldlt6_solve(Ht, nphi, upd);
    bool ufin = true;
    for (int i = 0; i < 6; ++i) ufin = ufin && isfinite(upd[i]);
    if (!(ok && hfin && ufin)) for (int i = 0; i < 6; ++i) upd[i] = 0.0;
    for (int i = 0; i < 6; ++i) updates[6 * k + i] = upd[i];
    if (new_poses) {
      double su[6];
      bool sfin = true;
      for (int i = 0; i < 6; ++i) { su[i] = step_size * upd[i]; sfin = sfin && isfinite(su[i]); }
      const Pose np = sfin ? pose_retract(Pk, su) : Pk;
      pose_to16(np, new_poses + 16 * k);
    }
  }
  if (!mm.prior) return;
  // ---- last block done: mean pose, stop test, status slot ------------------------------------------------------
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    unsigned int tk;
    asm volatile("atom.acq_rel.gpu.global.add.u32 %0, [%1], 1;" : "=r"(tk) : "l"(mm.ticket) : "memory");
    s_last = (tk == gridDim.x - 1);
    if (s_timeout) atomicExch(mm.done, 2);
  }
  __syncthreads();
  if (!s_last) return;
  Pose prior_p;
  pose_from16(mm.prior, prior_p);
  double mean_xi[6] = {0, 0, 0, 0, 0, 0};
  double mean_one = 0.0;
  for (int base = 0; base < K; base += STEIN_THREADS) {
    const int kk = base + threadIdx.x;
    if (kk < K) {
      Pose pk;
      double m16[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) m16[i] = __ldcg(new_poses + 16 * kk + i);
      pose_from16(m16, pk);
      double xi[6];
      pose_logmap(pose_between(prior_p, pk), xi);
#pragma unroll
      for (int i = 0; i < 6; ++i) s_xi[threadIdx.x][i] = xi[i];
    }
    __syncthreads();
    if (threadIdx.x < 6) {                // one thread per component, k order like the reference's loop (:865-870)
      const int cnt = min(STEIN_THREADS, K - base);
      for (int t = 0; t < cnt; ++t) mean_one += s_xi[t][threadIdx.x];
    }
    __syncthreads();
  }
  if (threadIdx.x < 6) s_xi[0][threadIdx.x] = mean_one;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int i = 0; i < 6; ++i) mean_xi[i] = s_xi[0][i] / static_cast<double>(K);
    const Pose mean_cur = pose_retract(prior_p, mean_xi);
    Pose mean_last;
    pose_from16(mm.mean_prev, mean_last);
    double dxi[6];
    pose_logmap(pose_between(mean_last, mean_cur), dxi);
    double nrm = 0;
    for (int i = 0; i < 6; ++i) nrm += dxi[i] * dxi[i];
    const bool converged = sqrt(nrm) < mm.stop_threshold;
    pose_to16(mean_cur, mm.mean_out);
    int d = atomicAdd(mm.done, 0);
    if (converged && d == 0) { *mm.done = 1; d = 1; }
    *mm.ticket = 0;
    __threadfence();   // device data (mean, done flag) is read after the kernel boundary; the host only reads the self-validating slot
    const unsigned long long word = static_cast<unsigned long long>(mm.iter + 1) | (static_cast<unsigned long long>(d) << 32);
    publish_slot(mm.h_status, mm.iter & 63, word, mm.status_key);
  }
}

// ------------------------------------------------------------------------------------------------- host side
#define CKS(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { h->last_error = std::string(#x) + ": " + cudaGetErrorString(e_); return NDTB_ERR_CUDA; } } while (0)

int svn_dev_ensure(ndtb_handle_s* h) {
  ndtb_handle_s::SvnDev& d = h->svnd;
  if (d.d_window) return NDTB_OK;
  constexpr int C = ndtb_handle_s::SvnDev::CAPACITY;
  CKS(cudaMalloc(&d.d_window, SVN_WINDOW_BYTES));
  CKS(cudaMemset(d.d_window, 0, SVN_WINDOW_BYTES));   // flags start at 0; stamps start at 1
  for (int i = 0; i < 2; ++i) CKS(cudaMalloc(&d.d_poses[i], sizeof(double) * 16 * C));
  CKS(cudaMalloc(&d.d_updates, sizeof(double) * 6 * C));
  CKS(cudaMalloc(&d.d_mean, sizeof(double) * 48));
  CKS(cudaMalloc(&d.d_done, sizeof(int)));
  CKS(cudaMalloc(&d.d_ticket, sizeof(unsigned int)));
  CKS(cudaMalloc(&d.d_wait_ns, sizeof(unsigned long long)));
  CKS(cudaMemset(d.d_done, 0, sizeof(int)));
  CKS(cudaMemset(d.d_ticket, 0, sizeof(unsigned int)));
  CKS(cudaMemset(d.d_wait_ns, 0, sizeof(unsigned long long)));
  CKS(cudaMallocHost(const_cast<unsigned long long**>(&d.h_status), 128 * sizeof(unsigned long long)));
  std::memset(const_cast<unsigned long long*>(d.h_status), 0, 128 * sizeof(unsigned long long));
  CKS(cudaDeviceSynchronize());   // the zeroed window is complete before its address can reach a peer
  d.peer_base[0] = d.d_window;
  d.rank = 0; d.world = 1;
  return NDTB_OK;
}

void svn_dev_free(ndtb_handle_s* h) {
  ndtb_handle_s::SvnDev& d = h->svnd;
  for (int r = 0; r < 8; ++r)
    if (d.peer_is_ipc[r] && d.peer_base[r]) cudaIpcCloseMemHandle(d.peer_base[r]);
  cudaFree(d.d_window); cudaFree(d.d_poses[0]); cudaFree(d.d_poses[1]); cudaFree(d.d_updates); cudaFree(d.d_mean);
  cudaFree(d.d_done); cudaFree(d.d_ticket); cudaFree(d.d_wait_ns);
  if (d.h_status) cudaFreeHost(const_cast<unsigned long long*>(d.h_status));
  for (cudaEvent_t e : d.ev) cudaEventDestroy(e);
  d = ndtb_handle_s::SvnDev();
}

int svn_dev_upload_poses(ndtb_handle_s* h, int K, const double* poses) {
  int rc = svn_dev_ensure(h);
  if (rc != NDTB_OK) return rc;
  if (K > ndtb_handle_s::SvnDev::CAPACITY) return set_error(h, NDTB_ERR_INVALID, "particle count exceeds the device window capacity (1024)");
  CKS(cudaMemcpyAsync(h->svnd.d_poses[h->svnd.cur], poses, sizeof(double) * 16 * K, cudaMemcpyHostToDevice, h->stream));
  return NDTB_OK;
}

int svn_dev_stage1(ndtb_handle_s* h, int K, int first, int stride, bool to_peers, unsigned long long* stamp_out, int vshards) {
  ndtb_handle_s::SvnDev& d = h->svnd;
  const int n = static_cast<int>(h->n_source);
  const int units = K * vshards;                                                  // vshards == 1: a unit is a particle
  const int m = (units > first) ? (units - first + stride - 1) / stride : 0;      // own units
  int rc = ensure_eval_buffers(h, std::max(vshards > 1 ? std::max(m, K) : m, 1), n);
  if (rc != NDTB_OK) return rc;
  // collective iterations count in d.stamp (all ranks in lock-step); local parity hooks take stamps from their own sequence with
  // the top bit set, so they can never satisfy -- or be mistaken for -- a flag a peer is waited on for
  const unsigned long long stamp = to_peers ? ++d.stamp : (SVN_LOCAL_STAMP | ++d.local_seq);
  if (stamp_out) *stamp_out = stamp;
  if (m == 0) return NDTB_OK;
  double d1, d2, d3;
  gauss_svn(h->params.outlier_ratio, h->params.resolution, &d1, &d2, &d3);
  if (vshards > 1) svn_state_kernel<<<(K + 127) / 128, 128, 0, h->stream>>>(K, 0, 1, d.d_poses[d.cur], h->eval.d_states, d1, d2, d3, d.d_done);  // every particle: units index the states by particle
  else svn_state_kernel<<<(m + 127) / 128, 128, 0, h->stream>>>(K, first, stride, d.d_poses[d.cur], h->eval.d_states, d1, d2, d3, d.d_done);
  SvnExchange x;
  x.vshards = vshards;
  if (vshards > 1) {
    // CTAs per unit: the units of one rank AT 8 RANKS should fill whole waves of 148 SMs x EVAL_MIN_BLOCKS CTAs (a last wave that
    // is half empty costs 4 % at 25 units x 222 CTAs).  Derived from K and the cloud size only -- never from the actual number of
    // ranks -- so every world size cuts the cloud into the same blocks and adds the same partial sums (bit-identical results).
    const int wave = 148 * EVAL_MIN_BLOCKS;
    const int b0 = std::max(1, eval_blocks_for(n) / vshards);
    const int u8 = (units + 7) / 8;
    const int waves = (u8 * b0) / wave;
    x.blocks_per_unit = waves > 0 ? std::min(b0, std::max(1, (waves * wave) / u8)) : b0;
  }
  x.world = stride; x.rank = first; x.capacity = ndtb_handle_s::SvnDev::CAPACITY;
  x.parity = static_cast<int>(stamp & 1ull); x.stamp = stamp; x.done = d.d_done;
  if (to_peers && d.world > 1) {
    x.n_dest = d.world;
    for (int r = 0; r < d.world; ++r) { x.rec[r] = svn_window_recs(d.peer_base[r]); x.flag[r] = svn_window_flags(d.peer_base[r]); }
  } else {
    x.n_dest = 1;
    x.rec[0] = svn_window_recs(d.d_window); x.flag[0] = svn_window_flags(d.d_window);
  }
  const int n_probe = (h->params.search_method == NDTB_DIRECT7) ? 7 : (h->params.search_method == NDTB_KDTREE ? 27 : 1);
  launch_eval(FLAVOUR_SVN, n_probe, true, h->svn.use_gauss_newton != 0, h->d_source, n, h->grid.meta, h->grid.recs, h->grid.table,
              h->grid.c_mean, h->grid.c_icov, h->eval, m, 0, h->stream, &x);
  CKS(cudaGetLastError());
  h->launches += 2; h->eval_launches += 1;
  return NDTB_OK;
}

int svn_dev_stein(ndtb_handle_s* h, int K, unsigned long long stamp, bool wait_flags, bool write_poses, bool with_mean, int iter, int vshards) {
  ndtb_handle_s::SvnDev& d = h->svnd;
  constexpr int C = ndtb_handle_s::SvnDev::CAPACITY;
  const int parity = static_cast<int>(stamp & 1ull);
  const double* rec = svn_window_recs(d.d_window) + static_cast<size_t>(parity) * C * SVN_REC;
  const unsigned long long* flags = wait_flags ? svn_window_flags(d.d_window) + static_cast<size_t>(parity) * C : nullptr;
  SteinMean mm{};
  if (with_mean) {
    mm.prior = d.d_mean + 32;
    mm.mean_prev = d.d_mean + 16 * (iter & 1);
    mm.mean_out = d.d_mean + 16 * ((iter + 1) & 1);
    mm.stop_threshold = h->svn.stop_threshold;
    mm.done = d.d_done; mm.ticket = d.d_ticket; mm.h_status = d.h_status;
    mm.status_key = 0x9E3779B97F4A7C15ull * stamp | 1ull;
    mm.iter = iter; mm.wait_ns = d.d_wait_ns;
  }
  static bool configured[MAX_DEVICES] = {};
  if (needs_func_setup(configured)) CKS(cudaFuncSetAttribute(stein_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(sizeof(SteinTile))));
  double* new_poses = write_poses ? d.d_poses[d.cur ^ 1] : nullptr;
  if (wait_flags) {
    // the kernel acquires every record through its flag, so it need not wait for the derivative kernel to END: launched as a
    // programmatic dependent (and without a griddepcontrol.wait of its own) its blocks are scheduled while that grid drains
    CKS(launch_pdl(stein_kernel, dim3(K), dim3(STEIN_THREADS), sizeof(SteinTile), h->stream, K, static_cast<const double*>(d.d_poses[d.cur]), rec, flags, stamp,
                   h->svn.kernel_bandwidth, h->svn.step_size, d.d_updates, new_poses, mm, vshards));
  } else {
    stein_kernel<<<K, STEIN_THREADS, sizeof(SteinTile), h->stream>>>(K, d.d_poses[d.cur], rec, flags, stamp, h->svn.kernel_bandwidth, h->svn.step_size,
                                                                     d.d_updates, new_poses, mm, vshards);
  }
  CKS(cudaGetLastError());
  if (write_poses) d.cur ^= 1;
  h->launches += 1;
  return NDTB_OK;
}

int svn_stein_update(ndtb_handle_s* h, int K, const double* poses, const double* grads, const double* hess, double* updates,
                     double* new_poses) {
  int rc = svn_dev_upload_poses(h, K, poses);
  if (rc != NDTB_OK) return rc;
  ndtb_handle_s::SvnDev& d = h->svnd;
  constexpr int C = ndtb_handle_s::SvnDev::CAPACITY;
  const unsigned long long stamp = SVN_LOCAL_STAMP | ++d.local_seq;
  std::vector<double> rec(static_cast<size_t>(SVN_REC) * K);
  for (int k = 0; k < K; ++k) {
    rec[SVN_REC * k] = 0.0;
    std::memcpy(&rec[SVN_REC * k + 1], grads + 6 * k, 48);
    std::memcpy(&rec[SVN_REC * k + 7], hess + 36 * k, 288);
  }
  double* d_rec = svn_window_recs(d.d_window) + static_cast<size_t>(stamp & 1ull) * C * SVN_REC;
  CKS(cudaMemcpyAsync(d_rec, rec.data(), sizeof(double) * SVN_REC * K, cudaMemcpyHostToDevice, h->stream));
  rc = svn_dev_stein(h, K, stamp, false, new_poses != nullptr, false, 0);
  if (rc != NDTB_OK) return rc;
  CKS(cudaMemcpyAsync(updates, d.d_updates, sizeof(double) * 6 * K, cudaMemcpyDeviceToHost, h->stream));
  if (new_poses) CKS(cudaMemcpyAsync(new_poses, d.d_poses[d.cur], sizeof(double) * 16 * K, cudaMemcpyDeviceToHost, h->stream));
  CKS(cudaStreamSynchronize(h->stream));
  return NDTB_OK;
}

}  // namespace ndtb
