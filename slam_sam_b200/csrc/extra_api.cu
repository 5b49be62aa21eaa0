// C ABI, remaining entry points: FP64 score-only passes and the batched replay.
//   calculateTransformationProbability / calculateNearestVoxelTransformationLikelihood
//   (ndt_omp_impl.hpp:1169-1281): serial FP64 passes in the reference, one kernel here.
#include "handle.cuh"
#include <cstring>
#include <string>
#include <vector>
#include <cmath>
#include <algorithm>
#include <array>
#include <fstream>
#include <iomanip>
#include <map>
#include <set>

using namespace ndtb;

#define CKH(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { h->last_error = std::string(#x) + ": " + cudaGetErrorString(e_); return NDTB_ERR_CUDA; } } while (0)

namespace ndtb {

__device__ __constant__ int c_off7[7][3] = {{0, 0, 0}, {1, 0, 0}, {-1, 0, 0}, {0, 1, 0}, {0, -1, 0}, {0, 0, 1}, {0, 0, -1}};
// DIRECT26: pcl::getAllNeighborCellIndices (13 half displacements, then their negatives; no centre cell), see derivatives.cu
__device__ __constant__ int c_off26s[26][3] = {
    {-1, -1, -1}, {-1, 0, -1}, {-1, 1, -1}, {0, -1, -1}, {0, 0, -1}, {0, 1, -1}, {1, -1, -1}, {1, 0, -1}, {1, 1, -1},
    {-1, -1, 0}, {0, -1, 0}, {1, -1, 0}, {-1, 0, 0},
    {1, 1, 1}, {1, 0, 1}, {1, -1, 1}, {0, 1, 1}, {0, 0, 1}, {0, -1, 1}, {-1, 1, 1}, {-1, 0, 1}, {-1, -1, 1},
    {1, 1, 0}, {0, 1, 0}, {-1, 1, 0}, {1, 0, 0}};
// KDTREE: the 27 cells a radius search with radius = leaf size can reach (+ the float distance test below), see derivatives.cu
__device__ __constant__ int c_off27s[27][3] = {
    {0, 0, 0}, {1, 0, 0}, {-1, 0, 0}, {0, 1, 0}, {0, -1, 0}, {0, 0, 1}, {0, 0, -1},
    {1, 1, 0}, {1, -1, 0}, {-1, 1, 0}, {-1, -1, 0}, {1, 0, 1}, {1, 0, -1}, {-1, 0, 1}, {-1, 0, -1}, {0, 1, 1}, {0, 1, -1}, {0, -1, 1}, {0, -1, -1},
    {1, 1, 1}, {1, 1, -1}, {1, -1, 1}, {1, -1, -1}, {-1, 1, 1}, {-1, 1, -1}, {-1, -1, 1}, {-1, -1, -1}};

// out[0] = sum of scores, out[1] = sum of per-point max score, out[2] = points with neighbours.
// Block partials are combined in block order by the host (tiny), so the result is deterministic.
__global__ void __launch_bounds__(128) score_kernel(const float4* __restrict__ src, int n, GridMeta m, const HashSlot* __restrict__ table,
                                                    const double* __restrict__ c_mean, const double* __restrict__ c_icov, const float* __restrict__ T,
                                                    double d1, double d2, int probes, double* __restrict__ block_out) {
  __shared__ double s_red[4][3];
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  double a0 = 0, a1 = 0, a2 = 0;
  if (idx < n) {
    const float4 p = __ldg(src + idx);
    float xt[3];
#pragma unroll
    for (int r = 0; r < 3; ++r) xt[r] = T[4 * r] * p.x + (T[4 * r + 1] * p.y + (T[4 * r + 2] * p.z + T[4 * r + 3]));
    const int i0 = static_cast<int>(floorf(__fdiv_rn(xt[0], m.leaf)));
    const int i1 = static_cast<int>(floorf(__fdiv_rn(xt[1], m.leaf)));
    const int i2 = static_cast<int>(floorf(__fdiv_rn(xt[2], m.leaf)));
    double best = 0;
    int nn = 0;
    for (int k = 0; k < probes; ++k) {
      const int* off = (probes == 27) ? c_off27s[k] : ((probes == 26) ? c_off26s[k] : c_off7[k]);
      const int ox = off[0], oy = off[1], oz = off[2];
      if (m.min_b[0] - i0 > ox || m.max_b[0] - i0 < ox || m.min_b[1] - i1 > oy || m.max_b[1] - i1 < oy || m.min_b[2] - i2 > oz || m.max_b[2] - i2 < oz) continue;
      const int lin = (i0 + ox - m.min_b[0]) * m.divb_mul[0] + (i1 + oy - m.min_b[1]) * m.divb_mul[1] + (i2 + oz - m.min_b[2]) * m.divb_mul[2];
      int row = hash_find(table, m.hash_mask, static_cast<uint32_t>(lin));
      if (probes == 27 && row < -1) row = -2 - row;   // flagged leaf: still in the reference's kd-tree (voxel_finalize.cu)
      if (row < 0) continue;
      if (probes == 27) {  // radiusSearch over float(leaf.mean_): L2_Simple<float>, dist < float(r * r)
        const float r2 = static_cast<float>(static_cast<double>(m.leaf) * static_cast<double>(m.leaf));
        const float dx = xt[0] - static_cast<float>(__ldg(c_mean + 3 * row)), dy = xt[1] - static_cast<float>(__ldg(c_mean + 3 * row + 1)),
                    dz = xt[2] - static_cast<float>(__ldg(c_mean + 3 * row + 2));
        float d = dx * dx; d = d + dy * dy; d = d + dz * dz;
        if (!(d < r2)) continue;
      }
      ++nn;
      double x[3], C[9];
      for (int a = 0; a < 3; ++a) x[a] = static_cast<double>(xt[a]) - __ldg(c_mean + 3 * row + a);
      for (int i = 0; i < 9; ++i) C[i] = __ldg(c_icov + 9 * row + i);
      double Cx[3];
      for (int r = 0; r < 3; ++r) Cx[r] = (C[3 * r] * x[0] + C[3 * r + 1] * x[1]) + C[3 * r + 2] * x[2];
      const double q = (x[0] * Cx[0] + x[1] * Cx[1]) + x[2] * Cx[2];
      const double s = -d1 * exp(-d2 * q / 2);
      a0 += s;
      if (s > best) best = s;
    }
    if (nn > 0) { a1 = best; a2 = 1.0; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    a0 += __shfl_xor_sync(0xffffffffu, a0, o);
    a1 += __shfl_xor_sync(0xffffffffu, a1, o);
    a2 += __shfl_xor_sync(0xffffffffu, a2, o);
  }
  if ((threadIdx.x & 31) == 0) { s_red[threadIdx.x >> 5][0] = a0; s_red[threadIdx.x >> 5][1] = a1; s_red[threadIdx.x >> 5][2] = a2; }
  __syncthreads();
  if (threadIdx.x < 3) block_out[3 * blockIdx.x + threadIdx.x] = ((s_red[0][threadIdx.x] + s_red[1][threadIdx.x]) + s_red[2][threadIdx.x]) + s_red[3][threadIdx.x];
}

// calculateScore (ndt_omp_impl.hpp:1083-1167): the score-only pass that also keeps a per-voxel score sum / pair count
// (voxel_score_map_, voxel_points_num) and the voxel ids of the points that found no neighbour (empty_voxels_).
// block_out[blockIdx] = block sum of scores (combined in block order by the host); per-voxel sums by FP64 atomics (their order is
// not fixed: the per-voxel means agree with a serial sum to rounding, not to the bit -- an off-path diagnostic).
__global__ void __launch_bounds__(128) score_map_kernel(const float4* __restrict__ src, int n, GridMeta m, const HashSlot* __restrict__ table,
                                                        const double* __restrict__ c_mean, const double* __restrict__ c_icov, const float* __restrict__ T,
                                                        double d1, double d2, int probes, double* __restrict__ block_out, double* __restrict__ vox_sum,
                                                        unsigned int* __restrict__ vox_cnt, int* __restrict__ empty_ijk, unsigned int* __restrict__ n_empty) {
  __shared__ double s_red[4];
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  double a0 = 0;
  if (idx < n) {
    const float4 p = __ldg(src + idx);
    float xt[3];
#pragma unroll
    for (int r = 0; r < 3; ++r) xt[r] = T[4 * r] * p.x + (T[4 * r + 1] * p.y + (T[4 * r + 2] * p.z + T[4 * r + 3]));
    const int i0 = static_cast<int>(floorf(__fdiv_rn(xt[0], m.leaf)));
    const int i1 = static_cast<int>(floorf(__fdiv_rn(xt[1], m.leaf)));
    const int i2 = static_cast<int>(floorf(__fdiv_rn(xt[2], m.leaf)));
    int nn = 0;
    for (int k = 0; k < probes; ++k) {
      const int* off = (probes == 27) ? c_off27s[k] : ((probes == 26) ? c_off26s[k] : c_off7[k]);
      const int ox = off[0], oy = off[1], oz = off[2];
      if (m.min_b[0] - i0 > ox || m.max_b[0] - i0 < ox || m.min_b[1] - i1 > oy || m.max_b[1] - i1 < oy || m.min_b[2] - i2 > oz || m.max_b[2] - i2 < oz) continue;
      const int lin = (i0 + ox - m.min_b[0]) * m.divb_mul[0] + (i1 + oy - m.min_b[1]) * m.divb_mul[1] + (i2 + oz - m.min_b[2]) * m.divb_mul[2];
      int row = hash_find(table, m.hash_mask, static_cast<uint32_t>(lin));
      if (probes == 27 && row < -1) row = -2 - row;   // flagged leaf: still in the reference's kd-tree (voxel_finalize.cu)
      if (row < 0) continue;
      if (probes == 27) {
        const float r2 = static_cast<float>(static_cast<double>(m.leaf) * static_cast<double>(m.leaf));
        const float dx = xt[0] - static_cast<float>(__ldg(c_mean + 3 * row)), dy = xt[1] - static_cast<float>(__ldg(c_mean + 3 * row + 1)),
                    dz = xt[2] - static_cast<float>(__ldg(c_mean + 3 * row + 2));
        float d = dx * dx; d = d + dy * dy; d = d + dz * dz;
        if (!(d < r2)) continue;
      }
      ++nn;
      double x[3], C[9];
      for (int a = 0; a < 3; ++a) x[a] = static_cast<double>(xt[a]) - __ldg(c_mean + 3 * row + a);
      for (int i = 0; i < 9; ++i) C[i] = __ldg(c_icov + 9 * row + i);
      double Cx[3];
      for (int r = 0; r < 3; ++r) Cx[r] = (C[3 * r] * x[0] + C[3 * r + 1] * x[1]) + C[3 * r + 2] * x[2];
      const double q = (x[0] * Cx[0] + x[1] * Cx[1]) + x[2] * Cx[2];
      const double s = -d1 * exp(-d2 * q / 2);
      a0 += s;
      atomicAdd(vox_sum + row, s);
      atomicAdd(vox_cnt + row, 1u);
    }
    if (nn == 0) {  // :1113-1122 (floor(x / resolution) on floats)
      const unsigned int e = atomicAdd(n_empty, 1u);
      empty_ijk[3 * e] = i0; empty_ijk[3 * e + 1] = i1; empty_ijk[3 * e + 2] = i2;
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) a0 += __shfl_xor_sync(0xffffffffu, a0, o);
  if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = a0;
  __syncthreads();
  if (threadIdx.x == 0) block_out[blockIdx.x] = ((s_red[0] + s_red[1]) + s_red[2]) + s_red[3];
}

}  // namespace ndtb

extern "C" {

int ndtb_calculate_scores(ndtb_handle h, const float T16[16], double* transformation_probability, double* nvtl) {
  if (!h || !T16) return NDTB_ERR_INVALID;
  cudaSetDevice(h->device);
  if (!h->grid.built || h->n_source == 0) return set_error(h, NDTB_ERR_NOT_READY, "target and source must be set");
  if (h->params.grid_flavour != NDTB_FLAVOUR_PCLOMP) return set_error(h, NDTB_ERR_INVALID, "pclomp entry point on an svn-flavour handle");
  const int n = static_cast<int>(h->n_source);
  const int blocks = (n + 127) / 128;
  double* d_out = nullptr; float* d_T = nullptr;
  CKH(cudaMalloc(&d_out, sizeof(double) * 3 * blocks));
  CKH(cudaMalloc(&d_T, 16 * sizeof(float)));
  CKH(cudaMemcpyAsync(d_T, T16, 16 * sizeof(float), cudaMemcpyHostToDevice, h->stream));
  double d1, d2, d3;
  {
    const double c1 = 10 * (1 - h->params.outlier_ratio);
    const double c2 = h->params.outlier_ratio / pow(static_cast<double>(h->params.resolution), 3);
    d3 = -log(c2); d1 = -log(c1 + c2) - d3; d2 = -2 * log((-log(c1 * exp(-0.5) + c2) - d3) / d1);
  }
  const int probes = (h->params.search_method == NDTB_DIRECT1) ? 1 : (h->params.search_method == NDTB_DIRECT26 ? 26 : (h->params.search_method == NDTB_KDTREE ? 27 : 7));
  score_kernel<<<blocks, 128, 0, h->stream>>>(h->d_source, n, h->grid.meta, h->grid.table, h->grid.c_mean, h->grid.c_icov, d_T, d1, d2, probes, d_out);
  h->launches += 1;
  std::vector<double> out(3 * static_cast<size_t>(blocks));
  CKH(cudaMemcpyAsync(out.data(), d_out, sizeof(double) * 3 * blocks, cudaMemcpyDeviceToHost, h->stream));
  CKH(cudaStreamSynchronize(h->stream));
  CKH(cudaGetLastError());
  cudaFree(d_out); cudaFree(d_T);
  double s = 0, nv = 0, found = 0;
  for (int b = 0; b < blocks; ++b) { s += out[3 * b]; nv += out[3 * b + 1]; found += out[3 * b + 2]; }
  if (transformation_probability) *transformation_probability = n ? s / static_cast<double>(n) : 0.0;
  if (nvtl) *nvtl = found != 0 ? nv / found : 0.0;
  return NDTB_OK;
}

int ndtb_calculate_score(ndtb_handle h, const float T16[16], double* score, size_t map_capacity, int64_t* voxel_idx, double* voxel_mean_score,
                         size_t* n_voxels, size_t empty_capacity, int32_t* empty_ijk, size_t* n_empty) {
  if (!h || !T16) return NDTB_ERR_INVALID;
  cudaSetDevice(h->device);
  if (!h->grid.built || h->n_source == 0) return set_error(h, NDTB_ERR_NOT_READY, "target and source must be set");
  if (h->params.grid_flavour != NDTB_FLAVOUR_PCLOMP) return set_error(h, NDTB_ERR_INVALID, "pclomp entry point on an svn-flavour handle");
  const int n = static_cast<int>(h->n_source);
  const size_t V = h->grid.V;
  const int blocks = (n + 127) / 128;
  double *d_out = nullptr, *d_sum = nullptr; float* d_T = nullptr; unsigned int *d_cnt = nullptr, *d_ne = nullptr; int* d_empty = nullptr;
  CKH(cudaMalloc(&d_out, sizeof(double) * blocks));
  CKH(cudaMalloc(&d_sum, sizeof(double) * std::max<size_t>(V, 1)));
  CKH(cudaMalloc(&d_cnt, sizeof(unsigned int) * std::max<size_t>(V, 1)));
  CKH(cudaMalloc(&d_empty, sizeof(int) * 3 * static_cast<size_t>(n)));
  CKH(cudaMalloc(&d_ne, sizeof(unsigned int)));
  CKH(cudaMalloc(&d_T, 16 * sizeof(float)));
  CKH(cudaMemsetAsync(d_sum, 0, sizeof(double) * std::max<size_t>(V, 1), h->stream));
  CKH(cudaMemsetAsync(d_cnt, 0, sizeof(unsigned int) * std::max<size_t>(V, 1), h->stream));
  CKH(cudaMemsetAsync(d_ne, 0, sizeof(unsigned int), h->stream));
  CKH(cudaMemcpyAsync(d_T, T16, 16 * sizeof(float), cudaMemcpyHostToDevice, h->stream));
  double d1, d2, d3;
  {
    const double c1 = 10 * (1 - h->params.outlier_ratio);
    const double c2 = h->params.outlier_ratio / pow(static_cast<double>(h->params.resolution), 3);
    d3 = -log(c2); d1 = -log(c1 + c2) - d3; d2 = -2 * log((-log(c1 * exp(-0.5) + c2) - d3) / d1);
  }
  const int probes = (h->params.search_method == NDTB_DIRECT1) ? 1 : (h->params.search_method == NDTB_DIRECT26 ? 26 : (h->params.search_method == NDTB_KDTREE ? 27 : 7));
  score_map_kernel<<<blocks, 128, 0, h->stream>>>(h->d_source, n, h->grid.meta, h->grid.table, h->grid.c_mean, h->grid.c_icov, d_T, d1, d2, probes, d_out,
                                                  d_sum, d_cnt, d_empty, d_ne);
  h->launches += 1;
  std::vector<double> out(blocks), sum(V), mean(3 * V);
  std::vector<unsigned int> cnt(V);
  unsigned int ne = 0;
  CKH(cudaMemcpyAsync(out.data(), d_out, sizeof(double) * blocks, cudaMemcpyDeviceToHost, h->stream));
  if (V) {
    CKH(cudaMemcpyAsync(sum.data(), d_sum, sizeof(double) * V, cudaMemcpyDeviceToHost, h->stream));
    CKH(cudaMemcpyAsync(cnt.data(), d_cnt, sizeof(unsigned int) * V, cudaMemcpyDeviceToHost, h->stream));
    CKH(cudaMemcpyAsync(mean.data(), h->grid.c_mean, sizeof(double) * 3 * V, cudaMemcpyDeviceToHost, h->stream));
  }
  CKH(cudaMemcpyAsync(&ne, d_ne, sizeof(unsigned int), cudaMemcpyDeviceToHost, h->stream));
  CKH(cudaStreamSynchronize(h->stream));
  std::vector<int> empt(3 * static_cast<size_t>(ne));
  if (ne) CKH(cudaMemcpy(empt.data(), d_empty, sizeof(int) * 3 * ne, cudaMemcpyDeviceToHost));
  CKH(cudaGetLastError());
  cudaFree(d_out); cudaFree(d_sum); cudaFree(d_cnt); cudaFree(d_empty); cudaFree(d_ne); cudaFree(d_T);
  double s = 0;
  for (int b = 0; b < blocks; ++b) s += out[b];
  if (score) *score = n ? s / static_cast<double>(n) : 0.0;
  // getScoreMap(): keyed by getLeafIndex(cell mean) (voxel_grid_covariance_omp.h:340-349), value = sum / pairs
  const GridMeta& m = h->grid.meta;
  std::map<size_t, std::pair<double, size_t>> smap;
  for (size_t v = 0; v < V; ++v) {
    if (!cnt[v]) continue;
    int ijk[3];
    for (int a = 0; a < 3; ++a) ijk[a] = static_cast<int>(floor(mean[3 * v + a] * static_cast<double>(m.inv_leaf)) - m.min_b[a]);
    const size_t key = static_cast<size_t>(ijk[0] * m.divb_mul[0] + ijk[1] * m.divb_mul[1] + ijk[2] * m.divb_mul[2]);
    auto& e = smap[key];
    e.first += sum[v]; e.second += cnt[v];
  }
  size_t c = 0;
  for (auto& kv : smap) {
    if (c < map_capacity) { if (voxel_idx) voxel_idx[c] = static_cast<int64_t>(kv.first); if (voxel_mean_score) voxel_mean_score[c] = kv.second.first / static_cast<double>(kv.second.second); }
    ++c;
  }
  if (n_voxels) *n_voxels = c;
  // getEmptyVoxels(): std::set ordered by EigenCmp (ndt_omp.h:57-65) = lexicographic (i, j, k)
  std::set<std::array<int, 3>> es;
  for (unsigned int e = 0; e < ne; ++e) es.insert({empt[3 * e], empt[3 * e + 1], empt[3 * e + 2]});
  c = 0;
  for (auto& v : es) { if (c < empty_capacity && empty_ijk) { empty_ijk[3 * c] = v[0]; empty_ijk[3 * c + 1] = v[1]; empty_ijk[3 * c + 2] = v[2]; } ++c; }
  if (n_empty) *n_empty = c;
  return NDTB_OK;
}

int ndtb_write_ndt_files(ndtb_handle h, const char* ellipsoid_filename, const char* voxel_filename, const char* points_filename,
                         const float* points, size_t n_points, size_t stride_bytes) {
  if (!h || !ellipsoid_filename || !voxel_filename) return NDTB_ERR_INVALID;
  if (!h->grid.built) return set_error(h, NDTB_ERR_NOT_READY, "no voxel map");
  if (h->params.grid_flavour != NDTB_FLAVOUR_PCLOMP) return set_error(h, NDTB_ERR_INVALID, "pclomp entry point on an svn-flavour handle");
  size_t V = 0;
  int rc = ndtb_get_voxel_count(h, &V, nullptr, nullptr, nullptr);
  if (rc != NDTB_OK) return rc;
  std::vector<int64_t> idx(V); std::vector<int32_t> n(V);
  std::vector<double> mean(3 * V), evecs(9 * V), evals(3 * V);
  if (V) { rc = ndtb_get_voxels(h, idx.data(), n.data(), mean.data(), nullptr, nullptr, evecs.data(), evals.data()); if (rc != NDTB_OK) return rc; }
  std::ofstream ef(ellipsoid_filename), vf(voxel_filename);
  if (!ef || !vf) return set_error(h, NDTB_ERR_INVALID, "cannot open the export files");
  ef << std::fixed << std::setprecision(8);
  vf << std::fixed << std::setprecision(8);
  const GridMeta& m = h->grid.meta;
  const size_t min_points = static_cast<size_t>(h->params.min_points_per_voxel);
  const float resolution = h->params.resolution;
  for (size_t v = 0; v < V; ++v) {
    // `leaf.getPointCount() >= min_points` compares an int with a size_t (include/pipeline.hpp:194): a leaf flagged -1 passes
    if (!(static_cast<size_t>(n[v]) >= min_points)) continue;
    ef << mean[3 * v] << " " << mean[3 * v + 1] << " " << mean[3 * v + 2] << " ";
    for (int i = 0; i < 9; ++i) ef << evecs[9 * v + i] << " ";
    ef << evals[3 * v] << " " << evals[3 * v + 1] << " " << evals[3 * v + 2] << " " << static_cast<size_t>(n[v]) << "\n";
    // getLeafCenter(index), voxel_grid_covariance_omp.h:352-367
    const size_t index = static_cast<size_t>(idx[v]);
    const int ijk2 = static_cast<int>(index / m.divb_mul[2]);
    const int ijk1 = static_cast<int>((index % m.divb_mul[2]) / m.divb_mul[1]);
    const int ijk0 = static_cast<int>(((index % m.divb_mul[2]) % m.divb_mul[1]) / m.divb_mul[0]);
    const double inv = static_cast<double>(m.inv_leaf);
    vf << (static_cast<double>(ijk0 + m.min_b[0]) + 0.5) / inv << " " << (static_cast<double>(ijk1 + m.min_b[1]) + 0.5) / inv << " "
       << (static_cast<double>(ijk2 + m.min_b[2]) + 0.5) / inv << " " << resolution << "\n";
  }
  if (points_filename && points && n_points) {   // the copy of the map cloud, x y z intensity (PointXYZI: intensity is the fifth float)
    std::ofstream pf(points_filename);
    if (!pf) return set_error(h, NDTB_ERR_INVALID, "cannot open the points file");
    pf << std::fixed << std::setprecision(8);
    for (size_t i = 0; i < n_points; ++i) {
      const float* q = reinterpret_cast<const float*>(reinterpret_cast<const unsigned char*>(points) + i * stride_bytes);
      const float intensity = stride_bytes >= 20 ? q[4] : 0.0f;
      pf << q[0] << " " << q[1] << " " << q[2] << " " << intensity << "\n";
    }
  }
  return NDTB_OK;
}

}  // extern "C"
