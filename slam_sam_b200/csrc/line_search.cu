// More-Thuente step length support (ndt_omp_impl.hpp:805-1079), used only with ndtb_params.use_line_search:
//   * computeHessian / updateHessian (:701-800): the all-FP64 Hessian-only pass the reference runs after a line
//     search that iterated -- serial there, one kernel here (thread per point, FP64 pair maths in the reference's
//     operation order, block partials summed in block order by the host, so the result is deterministic);
//   * the interval update and trial value selection (:805-911) live in newton_math.cuh.
#include "handle.cuh"
#include "newton_math.cuh"
#include <cstring>
#include <string>
#include <vector>

using namespace ndtb;

#define CKH(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { h->last_error = std::string(#x) + ": " + cudaGetErrorString(e_); return NDTB_ERR_CUDA; } } while (0)

namespace ndtb {

struct HessF64State {
  float T[12];
  double jd[24];  // 8 x 3 FP64 angle vectors j_ang_a_..h_ (ndt_omp_impl.hpp:433-440)
  double hd[45];  // 15 x 3 h_ang_a2_..f3_ (:469-488)
  double d1, d2;
};

__device__ __constant__ int c_off_ls[7][3] = {{0, 0, 0}, {1, 0, 0}, {-1, 0, 0}, {0, 1, 0}, {0, -1, 0}, {0, 0, 1}, {0, 0, -1}};

__device__ __forceinline__ double dot3d(const double* a, const double* b) { return (a[0] * b[0] + a[1] * b[1]) + a[2] * b[2]; }
__device__ __forceinline__ void mat3v(const double* C, const double* v, double* o) {
#pragma unroll
  for (int r = 0; r < 3; ++r) o[r] = (C[3 * r] * v[0] + C[3 * r + 1] * v[1]) + C[3 * r + 2] * v[2];
}

constexpr int HESS_THREADS = 128;

__global__ void __launch_bounds__(HESS_THREADS) hessian_f64_kernel(const float4* __restrict__ src, int n, GridMeta m, const HashSlot* __restrict__ table,
                                                                   const double* __restrict__ c_mean, const double* __restrict__ c_icov,
                                                                   const __grid_constant__ HessF64State st, int probes, double* __restrict__ block_out) {
  __shared__ double s_red[HESS_THREADS / 32][36];
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  double H[36];
#pragma unroll
  for (int i = 0; i < 36; ++i) H[i] = 0.0;
  if (idx < n) {
    const float4 p = __ldg(src + idx);
    float xt[3];
#pragma unroll
    for (int r = 0; r < 3; ++r) xt[r] = st.T[4 * r] * p.x + (st.T[4 * r + 1] * p.y + (st.T[4 * r + 2] * p.z + st.T[4 * r + 3]));
    const int i0 = static_cast<int>(floorf(__fdiv_rn(xt[0], m.leaf)));
    const int i1 = static_cast<int>(floorf(__fdiv_rn(xt[1], m.leaf)));
    const int i2 = static_cast<int>(floorf(__fdiv_rn(xt[2], m.leaf)));
    // computePointDerivatives, FP64 overload (:590-636): the same for every neighbour of the point
    const double x[3] = {static_cast<double>(p.x), static_cast<double>(p.y), static_cast<double>(p.z)};
    double Jp[3][6];
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
      for (int c = 0; c < 6; ++c) Jp[r][c] = (r == c) ? 1.0 : 0.0;
    Jp[1][3] = dot3d(x, st.jd + 0); Jp[2][3] = dot3d(x, st.jd + 3); Jp[0][4] = dot3d(x, st.jd + 6); Jp[1][4] = dot3d(x, st.jd + 9);
    Jp[2][4] = dot3d(x, st.jd + 12); Jp[0][5] = dot3d(x, st.jd + 15); Jp[1][5] = dot3d(x, st.jd + 18); Jp[2][5] = dot3d(x, st.jd + 21);
    double hv[6][3];  // a, b, c, d, e, f of Equation 6.21
    hv[0][0] = 0.0; hv[0][1] = dot3d(x, st.hd + 0); hv[0][2] = dot3d(x, st.hd + 3);
    hv[1][0] = 0.0; hv[1][1] = dot3d(x, st.hd + 6); hv[1][2] = dot3d(x, st.hd + 9);
    hv[2][0] = 0.0; hv[2][1] = dot3d(x, st.hd + 12); hv[2][2] = dot3d(x, st.hd + 15);
#pragma unroll
    for (int k = 0; k < 3; ++k) { hv[3][k] = dot3d(x, st.hd + 18 + 3 * k); hv[4][k] = dot3d(x, st.hd + 27 + 3 * k); hv[5][k] = dot3d(x, st.hd + 36 + 3 * k); }
    // block (3i, j) of point_hessian_ for i, j in 3..5: [[a b c] [b d e] [c e f]]; zero elsewhere
    const int hsel[3][3] = {{0, 1, 2}, {1, 3, 4}, {2, 4, 5}};
    for (int k = 0; k < probes; ++k) {
      const uint32_t key = [&]() -> uint32_t {
        const int ox = c_off_ls[k][0], oy = c_off_ls[k][1], oz = c_off_ls[k][2];
        if (m.min_b[0] - i0 > ox || m.max_b[0] - i0 < ox || m.min_b[1] - i1 > oy || m.max_b[1] - i1 < oy || m.min_b[2] - i2 > oz || m.max_b[2] - i2 < oz) return HASH_EMPTY;
        return static_cast<uint32_t>((i0 + ox - m.min_b[0]) * m.divb_mul[0] + (i1 + oy - m.min_b[1]) * m.divb_mul[1] + (i2 + oz - m.min_b[2]) * m.divb_mul[2]);
      }();
      if (key == HASH_EMPTY) continue;
      const int row = hash_find(table, m.hash_mask, key);
      if (row < 0) continue;
      double xr[3], C[9];
#pragma unroll
      for (int a = 0; a < 3; ++a) xr[a] = static_cast<double>(xt[a]) - __ldg(c_mean + 3 * row + a);
#pragma unroll
      for (int i = 0; i < 9; ++i) C[i] = __ldg(c_icov + 9 * row + i);
      // updateHessian (:772-800)
      double Cx[3];
      mat3v(C, xr, Cx);
      double e = st.d2 * exp(-st.d2 * dot3d(xr, Cx) / 2);
      if (e > 1 || e < 0 || e != e) continue;
      e *= st.d1;
      double CJ[6][3], xCJ[6];
#pragma unroll
      for (int i = 0; i < 6; ++i) {
        const double ci[3] = {Jp[0][i], Jp[1][i], Jp[2][i]};
        mat3v(C, ci, CJ[i]);
        xCJ[i] = dot3d(xr, CJ[i]);
      }
#pragma unroll
      for (int i = 0; i < 6; ++i) {
#pragma unroll
        for (int j = 0; j < 6; ++j) {
          const double cj[3] = {Jp[0][j], Jp[1][j], Jp[2][j]};
          double xCh = 0.0;  // x . (C * 0) = +0 for the zero blocks
          if (i >= 3 && j >= 3) {
            double Ch[3];
            mat3v(C, hv[hsel[i - 3][j - 3]], Ch);
            xCh = dot3d(xr, Ch);
          } else {
            const double z[3] = {0.0, 0.0, 0.0};
            double Ch[3];
            mat3v(C, z, Ch);
            xCh = dot3d(xr, Ch);  // keeps the sign-of-zero / NaN behaviour of the reference expression
          }
          H[6 * i + j] += e * (-st.d2 * xCJ[i] * xCJ[j] + xCh + dot3d(cj, CJ[i]));
        }
      }
    }
  }
#pragma unroll
  for (int i = 0; i < 36; ++i) {
    double v = H[i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5][i] = v;
  }
  __syncthreads();
  if (threadIdx.x < 36) block_out[36 * static_cast<size_t>(blockIdx.x) + threadIdx.x] =
      ((s_red[0][threadIdx.x] + s_red[1][threadIdx.x]) + s_red[2][threadIdx.x]) + s_red[3][threadIdx.x];
}

// FP64 angle vectors: the double rows behind angle_tables() (same expressions, not rounded to float)
static void angle_vectors_f64(const double* p, double* jd, double* hd) {
  const double snap = 10e-5;
  double cx = 1.0, sx = 0.0, cy = 1.0, sy = 0.0, cz = 1.0, sz = 0.0;
  if (!(fabs(p[3]) < snap)) { cx = cos(p[3]); sx = sin(p[3]); }
  if (!(fabs(p[4]) < snap)) { cy = cos(p[4]); sy = sin(p[4]); }
  if (!(fabs(p[5]) < snap)) { cz = cos(p[5]); sz = sin(p[5]); }
  const double j[24] = {(-sx * sz + cx * sy * cz), (-sx * cz - cx * sy * sz), (-cx * cy), (cx * sz + sx * sy * cz), (cx * cz - sx * sy * sz), (-sx * cy),
                        (-sy * cz), sy * sz, cy, sx * cy * cz, (-sx * cy * sz), sx * sy, (-cx * cy * cz), cx * cy * sz, (-cx * sy),
                        (-cy * sz), (-cy * cz), 0, (cx * cz - sx * sy * sz), (-cx * sz - sx * sy * cz), 0, (sx * cz + cx * sy * sz), (cx * sy * cz - sx * sz), 0};
  const double hh[45] = {(-cx * sz - sx * sy * cz), (-cx * cz + sx * sy * sz), sx * cy, (-sx * sz + cx * sy * cz), (-cx * sy * sz - sx * cz), (-cx * cy),
                         (cx * cy * cz), (-cx * cy * sz), (cx * sy), (sx * cy * cz), (-sx * cy * sz), (sx * sy),
                         (-sx * cz - cx * sy * sz), (sx * sz - cx * sy * cz), 0, (cx * cz - sx * sy * sz), (-sx * sy * cz - cx * sz), 0,
                         (-cy * cz), (cy * sz), (sy), (-sx * sy * cz), (sx * sy * sz), (sx * cy), (cx * sy * cz), (-cx * sy * sz), (-cx * cy),
                         (sy * sz), (sy * cz), 0, (-sx * cy * sz), (-sx * cy * cz), 0, (cx * cy * sz), (cx * cy * cz), 0,
                         (-cy * cz), (cy * sz), 0, (-cx * sz - sx * sy * cz), (-cx * cz + sx * sy * sz), 0, (-sx * sz + cx * sy * cz), (-cx * sy * sz - sx * cz), 0};
  std::memcpy(jd, j, sizeof(j));
  std::memcpy(hd, hh, sizeof(hh));
}

// computeHessian(hessian, trans_cloud, x_t): Hessian only, all FP64, at pose p / transform T16.
int compute_hessian_f64(ndtb_handle_s* h, const double* p, const float* T16, double* H) {
  const int n = static_cast<int>(h->n_source);
  const int blocks = (n + HESS_THREADS - 1) / HESS_THREADS;
  HessF64State st;
  std::memcpy(st.T, T16, 12 * sizeof(float));
  angle_vectors_f64(p, st.jd, st.hd);
  double d3;
  gauss_pclomp(h->params.outlier_ratio, h->params.resolution, &st.d1, &st.d2, &d3);
  double* d_out = nullptr;
  CKH(cudaMallocAsync(&d_out, sizeof(double) * 36 * static_cast<size_t>(blocks), h->stream));
  const int probes = (h->params.search_method == NDTB_DIRECT1) ? 1 : 7;
  hessian_f64_kernel<<<blocks, HESS_THREADS, 0, h->stream>>>(h->d_source, n, h->grid.meta, h->grid.table, h->grid.c_mean, h->grid.c_icov, st, probes, d_out);
  h->launches += 1;
  std::vector<double> out(36 * static_cast<size_t>(blocks));
  CKH(cudaMemcpyAsync(out.data(), d_out, sizeof(double) * out.size(), cudaMemcpyDeviceToHost, h->stream));
  CKH(cudaFreeAsync(d_out, h->stream));
  CKH(cudaStreamSynchronize(h->stream));
  CKH(cudaGetLastError());
  for (int i = 0; i < 36; ++i) H[i] = 0.0;
  for (int b = 0; b < blocks; ++b)
    for (int i = 0; i < 36; ++i) H[i] += out[36 * static_cast<size_t>(b) + i];
  return NDTB_OK;
}

}  // namespace ndtb
