// Target voxel-map build on the GPU:
//   k0 bbox_kernel        : float min/max over the target (pcl::getMinMax3D)
//   k1 voxel_key_kernel   : per-point linear voxel index, float32 -> int32 exactly as the reference
//   k2 radix_*            : stable LSD radix sort of (key, point index)   [hand-written, 8-bit digits]
//   k3 seg_*              : segment heads -> voxel numbering in ascending linear index
//   k4 voxel_finalize     : ONE THREAD PER VOXEL sums its points IN INPUT ORDER in FP64 (this is
//                           the order-preserving segmented reduction: the reference accumulates
//                           Sum x, Sum x x^T in point order, voxel_grid_covariance_omp_impl.hpp:235-239),
//                           then mean / covariance / eigen-inflate / inverse and the hash insert.
// Roofline: HBM streaming. Algorithmic bytes B_build = 2*16*N_t + 64*V.
#include "ndtb_internal.cuh"
#include "build_api.cuh"
#pragma once

namespace ndtb {

// ------------------------------------------------------------------------------------------ k0
__global__ void bbox_kernel(const float4* __restrict__ pts, int n, int is_dense, int* __restrict__ out /*6 ordered ints*/) {
  float mn[3] = {3.402823466e+38f, 3.402823466e+38f, 3.402823466e+38f};
  float mx[3] = {-3.402823466e+38f, -3.402823466e+38f, -3.402823466e+38f};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const float4 p = __ldg(pts + i);
    if (!is_dense && !(isfinite(p.x) && isfinite(p.y) && isfinite(p.z))) continue;
    mn[0] = fminf(mn[0], p.x); mn[1] = fminf(mn[1], p.y); mn[2] = fminf(mn[2], p.z);
    mx[0] = fmaxf(mx[0], p.x); mx[1] = fmaxf(mx[1], p.y); mx[2] = fmaxf(mx[2], p.z);
  }
#pragma unroll
  for (int a = 0; a < 3; ++a) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      mn[a] = fminf(mn[a], __shfl_xor_sync(0xffffffffu, mn[a], o));
      mx[a] = fmaxf(mx[a], __shfl_xor_sync(0xffffffffu, mx[a], o));
    }
  }
  if ((threadIdx.x & 31) == 0) {
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      atomicMin(out + a, float_to_ordered(mn[a]));
      atomicMax(out + 3 + a, float_to_ordered(mx[a]));
    }
  }
}

// ------------------------------------------------------------------------------------------ k1
// pclomp: int(floorf(x * inv) - float(min_b))   voxel_grid_covariance_omp_impl.hpp:219-227
// svn   : int(floorf(x * inv) - min_b)          voxel_grid_covariance_impl.hpp:222-225 (float - int -> float)
// Both subtractions are exact float operations for |index| < 2^24, so one code path serves both.
__global__ void voxel_key_kernel(const float4* __restrict__ pts, int n, GridMeta m, int is_dense, uint32_t* __restrict__ keys,
                                 uint32_t* __restrict__ vals, unsigned int* __restrict__ n_skipped) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float4 p = __ldg(pts + i);
  const bool finite = isfinite(p.x) && isfinite(p.y) && isfinite(p.z);
  const bool skip = (m.flavour == FLAVOUR_SVN) ? !finite : (!is_dense && !finite);
  uint32_t key;
  if (skip) {
    key = m.total_cells;
    atomicAdd(n_skipped, 1u);
  } else {
    const int i0 = static_cast<int>(floorf(p.x * m.inv_leaf) - static_cast<float>(m.min_b[0]));
    const int i1 = static_cast<int>(floorf(p.y * m.inv_leaf) - static_cast<float>(m.min_b[1]));
    const int i2 = static_cast<int>(floorf(p.z * m.inv_leaf) - static_cast<float>(m.min_b[2]));
    key = static_cast<uint32_t>(i0 * m.divb_mul[0] + i1 * m.divb_mul[1] + i2 * m.divb_mul[2]);
  }
  keys[i] = key;
  vals[i] = static_cast<uint32_t>(i);
}

// ------------------------------------------------------------------------------------------ k2
constexpr int RS_THREADS = 256;
constexpr int RS_ITEMS = 8;
constexpr int RS_TILE = RS_THREADS * RS_ITEMS;

__global__ void __launch_bounds__(RS_THREADS) radix_hist_kernel(const uint32_t* __restrict__ keys, int n, int shift,
                                                                uint32_t* __restrict__ hist, int nb) {
  __shared__ uint32_t h[256];
  h[threadIdx.x] = 0;
  __syncthreads();
  const int base = blockIdx.x * RS_TILE;
#pragma unroll
  for (int i = 0; i < RS_ITEMS; ++i) {
    const int idx = base + i * RS_THREADS + threadIdx.x;
    if (idx < n) atomicAdd(&h[(__ldg(keys + idx) >> shift) & 255u], 1u);
  }
  __syncthreads();
  hist[threadIdx.x * nb + blockIdx.x] = h[threadIdx.x];
}

// Single-block exclusive scan of `n` uint32 (contiguous chunk per thread).  Optional: writes total.
__global__ void __launch_bounds__(1024) scan_u32_kernel(uint32_t* __restrict__ data, int n, uint32_t* __restrict__ total) {
  __shared__ uint32_t warp_sums[32];
  const int tid = threadIdx.x;
  const int per = (n + 1023) / 1024;
  const int lo = min(tid * per, n), hi = min(lo + per, n);
  uint32_t s = 0;
  for (int i = lo; i < hi; ++i) s += data[i];
  uint32_t incl = s;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
    if ((tid & 31) >= o) incl += t;
  }
  if ((tid & 31) == 31) warp_sums[tid >> 5] = incl;
  __syncthreads();
  if (tid < 32) {
    uint32_t w = warp_sums[tid], wi = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, wi, o);
      if (tid >= o) wi += t;
    }
    warp_sums[tid] = wi - w;
    if (tid == 31 && total) *total = wi;
  }
  __syncthreads();
  uint32_t run = warp_sums[tid >> 5] + (incl - s);
  for (int i = lo; i < hi; ++i) { const uint32_t v = data[i]; data[i] = run; run += v; }
}

__global__ void __launch_bounds__(RS_THREADS) radix_scatter_kernel(const uint32_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                                                                   uint32_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out, int n,
                                                                   int shift, const uint32_t* __restrict__ hist, int nb) {
  __shared__ uint32_t cnt[RS_THREADS / 32][256];
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
#pragma unroll
  for (int i = 0; i < RS_THREADS / 32; ++i) cnt[i][tid] = 0;
  __syncthreads();
  const int base = blockIdx.x * RS_TILE + w * (32 * RS_ITEMS);
  uint32_t key[RS_ITEMS], val[RS_ITEMS], rank[RS_ITEMS];
  const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
  for (int i = 0; i < RS_ITEMS; ++i) {
    const int idx = base + i * 32 + lane;
    const bool valid = idx < n;
    key[i] = valid ? __ldg(keys_in + idx) : 0u;
    val[i] = valid ? __ldg(vals_in + idx) : 0u;
    const uint32_t d = valid ? ((key[i] >> shift) & 255u) : 256u;
    const uint32_t peers = __match_any_sync(0xffffffffu, d);
    uint32_t prefix = 0;
    if (valid) prefix = cnt[w][d];
    __syncwarp();
    rank[i] = prefix + __popc(peers & lt);
    if (valid && (peers & lt) == 0) cnt[w][d] = prefix + __popc(peers);
    __syncwarp();
  }
  __syncthreads();
  {
    uint32_t run = __ldg(hist + tid * nb + blockIdx.x);
#pragma unroll
    for (int i = 0; i < RS_THREADS / 32; ++i) { const uint32_t c = cnt[i][tid]; cnt[i][tid] = run; run += c; }
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < RS_ITEMS; ++i) {
    const int idx = base + i * 32 + lane;
    if (idx < n) {
      const uint32_t pos = cnt[w][(key[i] >> shift) & 255u] + rank[i];
      keys_out[pos] = key[i];
      vals_out[pos] = val[i];
    }
  }
}

// ------------------------------------------------------------------------------------------ k3
constexpr int SEG_THREADS = 256;
constexpr int SEG_ITEMS = 8;
constexpr int SEG_TILE = SEG_THREADS * SEG_ITEMS;

__device__ __forceinline__ bool is_head(const uint32_t* __restrict__ keys, int i, uint32_t sentinel) {
  const uint32_t k = __ldg(keys + i);
  if (k >= sentinel) return false;
  return i == 0 || __ldg(keys + i - 1) != k;
}

__global__ void __launch_bounds__(SEG_THREADS) seg_count_kernel(const uint32_t* __restrict__ keys, int n, uint32_t sentinel,
                                                                uint32_t* __restrict__ block_counts) {
  __shared__ uint32_t s;
  if (threadIdx.x == 0) s = 0;
  __syncthreads();
  const int base = blockIdx.x * SEG_TILE + threadIdx.x * SEG_ITEMS;
  uint32_t c = 0;
#pragma unroll
  for (int i = 0; i < SEG_ITEMS; ++i) if (base + i < n && is_head(keys, base + i, sentinel)) ++c;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
  if ((threadIdx.x & 31) == 0) atomicAdd(&s, c);
  __syncthreads();
  if (threadIdx.x == 0) block_counts[blockIdx.x] = s;
}

__global__ void __launch_bounds__(SEG_THREADS) seg_write_kernel(const uint32_t* __restrict__ keys, int n, uint32_t sentinel,
                                                                const uint32_t* __restrict__ block_offsets, uint32_t* __restrict__ seg_start,
                                                                uint32_t* __restrict__ vox_key) {
  __shared__ uint32_t warp_sums[SEG_THREADS / 32];
  const int tid = threadIdx.x;
  const int base = blockIdx.x * SEG_TILE + tid * SEG_ITEMS;
  bool head[SEG_ITEMS];
  uint32_t c = 0;
#pragma unroll
  for (int i = 0; i < SEG_ITEMS; ++i) { head[i] = (base + i < n) && is_head(keys, base + i, sentinel); c += head[i]; }
  uint32_t incl = c;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o); if ((tid & 31) >= o) incl += t; }
  if ((tid & 31) == 31) warp_sums[tid >> 5] = incl;
  __syncthreads();
  uint32_t woff = 0;
  for (int k = 0; k < (tid >> 5); ++k) woff += warp_sums[k];
  uint32_t v = block_offsets[blockIdx.x] + woff + (incl - c);
#pragma unroll
  for (int i = 0; i < SEG_ITEMS; ++i) if (head[i]) { seg_start[v] = base + i; vox_key[v] = __ldg(keys + base + i); ++v; }
}

}  // namespace ndtb
