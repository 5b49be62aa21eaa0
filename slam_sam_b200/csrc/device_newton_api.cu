// Test hook of the warp-cooperative device Newton step (device_newton.cuh): one step on the device next to the SAME step computed
// by the host code of ndtb_align (newton_math.cuh), for bit-for-bit comparison.
#include "device_newton.cuh"
#include "../../include/ndtb.h"
#include <cstring>

namespace ndtb {

struct NewtonStepOut {
  int status, nr_iterations;
  double a_t;
  EvalState st;
};

__global__ void newton_step_test_kernel(const double* g, const double* H, EvalState st_in, NewtonCarry carry, NewtonParams prm, NewtonStepOut* out) {
  __shared__ NewtonScratch s;
  __shared__ EvalState st_out;
  const int status = newton_step_warp(g, H, st_in, carry, prm, &st_out, s);
  __syncwarp();
  if (threadIdx.x == 0) {
    out->status = status; out->nr_iterations = carry.nr_iterations; out->a_t = carry.a_t;
    if (status == NEWTON_CONTINUE) out->st = st_out;
  }
}

// The host's version of the same step: the statements of align_impl (ndtb_api.cu) between two evaluations, direct solver only.
static int newton_step_host(const double* g, const double* H, const EvalState& st_in, NewtonCarry& carry, const NewtonParams& prm, EvalState* st_out) {
  if (!carry.first) {
    if (carry.nr_iterations > prm.max_iterations || (carry.nr_iterations && (fabs(carry.a_t) < prm.trans_eps))) return NEWTON_CONVERGED;
    carry.nr_iterations += 1;
  }
  carry.first = 0;
  double neg_g[6], dp[6];
  for (int i = 0; i < 6; ++i) neg_g[i] = -g[i];
  if (!gecp6_solve(H, neg_g, dp, 1e-6)) return NEWTON_HOST;
  double nrm2 = 0;
  for (int i = 0; i < 6; ++i) nrm2 += dp[i] * dp[i];
  const double nrm = sqrt(nrm2);
  if (nrm == 0 || nrm != nrm) return NEWTON_HOST;
  for (int i = 0; i < 6; ++i) dp[i] = dp[i] / nrm;
  double d_phi_0 = 0;
  for (int i = 0; i < 6; ++i) d_phi_0 += g[i] * dp[i];
  d_phi_0 = -d_phi_0;
  if (d_phi_0 >= 0 && d_phi_0 == 0) return NEWTON_HOST;
  if (d_phi_0 >= 0) { for (int i = 0; i < 6; ++i) dp[i] *= -1; }
  double a_t = nrm;
  a_t = std::min(a_t, prm.step_max);
  a_t = std::max(a_t, prm.step_min);
  double x_t[6];
  for (int i = 0; i < 6; ++i) x_t[i] = st_in.p[i] + dp[i] * a_t;
  carry.a_t = a_t;
  std::memset(st_out, 0, sizeof(*st_out));
  float T[16];
  pose_to_T(x_t, T);
  std::memcpy(st_out->T, T, 12 * sizeof(float));
  st_out->one = 1.0f;
  angle_tables(x_t, 10e-5, st_out->jang, st_out->hang);
  st_out->d1 = st_in.d1; st_out->d2 = st_in.d2; st_out->d3 = st_in.d3;
  std::memcpy(st_out->p, x_t, sizeof(x_t));
  return NEWTON_CONTINUE;
}

}  // namespace ndtb

// which = 0: device, 1: host.  out_state: T[12], jang[24], hang[45] as floats (81), then p[6] as doubles in out_p.
extern "C" int ndtb_test_newton_step(int which, int cuda_device, const double g[6], const double H[36], const double p[6], const double params[4],
                                     int max_iterations, int first, int nr_iterations, double a_t_prev, int* status, int* nr_out, double* a_t_out,
                                     float* out_state81, double* out_p6) {
  using namespace ndtb;
  EvalState st_in;
  std::memset(&st_in, 0, sizeof(st_in));
  std::memcpy(st_in.p, p, 48);
  st_in.d1 = 1.0; st_in.d2 = 2.0; st_in.d3 = 3.0; st_in.one = 1.0f;
  NewtonParams prm;
  prm.step_max = params[0]; prm.step_min = params[1]; prm.trans_eps = params[2]; prm.max_iterations = max_iterations; prm.pad = 0;
  NewtonCarry carry;
  carry.a_t = a_t_prev; carry.nr_iterations = nr_iterations; carry.first = first;
  NewtonStepOut o;
  std::memset(&o, 0, sizeof(o));
  if (which == 1) {
    o.status = newton_step_host(g, H, st_in, carry, prm, &o.st);
    o.nr_iterations = carry.nr_iterations; o.a_t = carry.a_t;
  } else {
    if (cudaSetDevice(cuda_device) != cudaSuccess) return NDTB_ERR_CUDA;
    double *d_g = nullptr, *d_H = nullptr;
    NewtonStepOut* d_o = nullptr;
    if (cudaMalloc(&d_g, 48) || cudaMalloc(&d_H, 288) || cudaMalloc(&d_o, sizeof(o))) return NDTB_ERR_CUDA;
    cudaMemcpy(d_g, g, 48, cudaMemcpyHostToDevice); cudaMemcpy(d_H, H, 288, cudaMemcpyHostToDevice);
    cudaMemset(d_o, 0, sizeof(o));
    newton_step_test_kernel<<<1, 32>>>(d_g, d_H, st_in, carry, prm, d_o);
    const cudaError_t e = cudaMemcpy(&o, d_o, sizeof(o), cudaMemcpyDeviceToHost);
    cudaFree(d_g); cudaFree(d_H); cudaFree(d_o);
    if (e != cudaSuccess) return NDTB_ERR_CUDA;
  }
  *status = o.status; *nr_out = o.nr_iterations; *a_t_out = o.a_t;
  std::memcpy(out_state81, o.st.T, 12 * 4); std::memcpy(out_state81 + 12, o.st.jang, 24 * 4); std::memcpy(out_state81 + 36, o.st.hang, 45 * 4);
  std::memcpy(out_p6, o.st.p, 48);
  return NDTB_OK;
}
