// SVN stages 2 and 3 as ONE small kernel: grid = K blocks (particle k), threads stride over l.
//   xi_lk = Logmap(P_l^-1 P_k); k = exp(-|xi|^2 / h); grad_k = k * (-2/h) * xi      svn_ndt_impl.hpp:214-244
//   phi_k = (1/K) sum_l [k * Perm g_l + grad_k];  H_k = (1/K) sum_l [k^2 Perm H_l Perm + grad_k grad_k^T] + 1e-6 I   :795-815
//   update_k = LDLT(H_k).solve(-phi_k)  :833-838 ;  P_k <- P_k.retract(step * update_k)  :848-855
// All FP64.  The sum over l is a fixed-shape tree (strided partial sums, warp shuffles, warps in
// order) -- deterministic, differs from the reference's serial sum only by FP64 rounding.
#include "svn_stein.cuh"
#include "se3_math.cuh"
#include <string>

namespace ndtb {

constexpr int STEIN_THREADS = 128;
constexpr int STEIN_N = 42;  // phi(6) + H(36)

__global__ void __launch_bounds__(STEIN_THREADS) stein_kernel(int K, const double* __restrict__ poses, const double* __restrict__ grads,
                                                              const double* __restrict__ hess, double kernel_h, double step_size,
                                                              double* __restrict__ updates, double* __restrict__ new_poses) {
  __shared__ double s_red[STEIN_THREADS / 32][STEIN_N];
  const int k = blockIdx.x;
  Pose Pk;
  pose_from16(poses + 16 * k, Pk);
  double acc[STEIN_N];
#pragma unroll
  for (int i = 0; i < STEIN_N; ++i) acc[i] = 0.0;
  const int perm[6] = {3, 4, 5, 0, 1, 2};  // NDT (t, r) -> GTSAM (r, t)   svn_ndt_impl.hpp:731-735
  for (int l = threadIdx.x; l < K; l += blockDim.x) {
    double k_val, k_grad[6];
    if (kernel_h <= 1e-12) {
      k_val = (l == k) ? 1.0 : 0.0;
#pragma unroll
      for (int i = 0; i < 6; ++i) k_grad[i] = 0.0;
    } else {
      Pose Pl;
      pose_from16(poses + 16 * l, Pl);
      double xi[6];
      pose_logmap(pose_between(Pl, Pk), xi);
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
    if (!fin) continue;
    bool gfin = true, hfin = true;
#pragma unroll
    for (int i = 0; i < 6; ++i) gfin = gfin && isfinite(grads[6 * l + i]);
    for (int i = 0; i < 36; ++i) hfin = hfin && isfinite(hess[36 * l + i]);
#pragma unroll
    for (int i = 0; i < 6; ++i) {
      if (gfin) acc[i] += k_val * grads[6 * l + perm[i]];
      acc[i] += k_grad[i];
    }
    const double kk = k_val * k_val;
#pragma unroll
    for (int i = 0; i < 6; ++i)
#pragma unroll
      for (int j = 0; j < 6; ++j) {
        if (hfin) acc[6 + 6 * i + j] += kk * hess[36 * l + 6 * perm[i] + perm[j]];
        acc[6 + 6 * i + j] += k_grad[i] * k_grad[j];
      }
  }
#pragma unroll
  for (int i = 0; i < STEIN_N; ++i) {
    double v = acc[i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    acc[i] = v;
  }
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (lane == 0) {
#pragma unroll
    for (int i = 0; i < STEIN_N; ++i) s_red[w][i] = acc[i];
  }
  __syncthreads();
  if (threadIdx.x != 0) return;
  double phi[6], Ht[36];
  for (int i = 0; i < STEIN_N; ++i) {
    double v = s_red[0][i];
    for (int ww = 1; ww < STEIN_THREADS / 32; ++ww) v += s_red[ww][i];
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

int svn_stein_update(ndtb_handle_s* h, int K, const double* poses, const double* grads, const double* hess, double* updates,
                     double* new_poses) {
#define CKS(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { h->last_error = std::string(#x) + ": " + cudaGetErrorString(e_); return NDTB_ERR_CUDA; } } while (0)
  if (K > h->svn_alloc) {
    cudaFree(h->d_svn_in); cudaFree(h->d_svn_out);
    CKS(cudaMalloc(&h->d_svn_in, sizeof(double) * 58 * K));
    CKS(cudaMalloc(&h->d_svn_out, sizeof(double) * 22 * K));
    h->svn_alloc = K;
  }
  double* d_poses = h->d_svn_in;
  double* d_grads = d_poses + 16 * K;
  double* d_hess = d_grads + 6 * K;
  double* d_upd = h->d_svn_out;
  double* d_new = d_upd + 6 * K;
  CKS(cudaMemcpyAsync(d_poses, poses, sizeof(double) * 16 * K, cudaMemcpyHostToDevice, h->stream));
  CKS(cudaMemcpyAsync(d_grads, grads, sizeof(double) * 6 * K, cudaMemcpyHostToDevice, h->stream));
  CKS(cudaMemcpyAsync(d_hess, hess, sizeof(double) * 36 * K, cudaMemcpyHostToDevice, h->stream));
  stein_kernel<<<K, STEIN_THREADS, 0, h->stream>>>(K, d_poses, d_grads, d_hess, h->svn.kernel_bandwidth, h->svn.step_size, d_upd,
                                                   new_poses ? d_new : nullptr);
  h->launches += 1;
  CKS(cudaMemcpyAsync(updates, d_upd, sizeof(double) * 6 * K, cudaMemcpyDeviceToHost, h->stream));
  if (new_poses) CKS(cudaMemcpyAsync(new_poses, d_new, sizeof(double) * 16 * K, cudaMemcpyDeviceToHost, h->stream));
  CKS(cudaStreamSynchronize(h->stream));
  CKS(cudaGetLastError());
  return NDTB_OK;
#undef CKS
}

}  // namespace ndtb
