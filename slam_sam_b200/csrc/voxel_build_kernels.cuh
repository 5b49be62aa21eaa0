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
// pcl::getMinMax3D.  Block reduction, 6 atomics per block; the last block publishes the box to the host.
constexpr int BBOX_THREADS = 256;
__global__ void __launch_bounds__(BBOX_THREADS) bbox_kernel(const float4* __restrict__ pts, int n, int is_dense, int* __restrict__ out /*6 ordered ints*/,
                                                            unsigned int* __restrict__ ticket, volatile unsigned long long* host_slots,
                                                            unsigned long long key) {
  __shared__ float s_mn[BBOX_THREADS / 32][3], s_mx[BBOX_THREADS / 32][3];
  __shared__ bool s_last;
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
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (lane == 0) {
#pragma unroll
    for (int a = 0; a < 3; ++a) { s_mn[w][a] = mn[a]; s_mx[w][a] = mx[a]; }
  }
  __syncthreads();
  if (threadIdx.x < 3) {
    float v = s_mn[0][threadIdx.x];
    for (int k = 1; k < BBOX_THREADS / 32; ++k) v = fminf(v, s_mn[k][threadIdx.x]);
    atomicMin(out + threadIdx.x, float_to_ordered(v));
  } else if (threadIdx.x < 6) {
    float v = s_mx[0][threadIdx.x - 3];
    for (int k = 1; k < BBOX_THREADS / 32; ++k) v = fmaxf(v, s_mx[k][threadIdx.x - 3]);
    atomicMax(out + threadIdx.x, float_to_ordered(v));
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned int tk;
    asm volatile("atom.acq_rel.gpu.global.add.u32 %0, [%1], 1;" : "=r"(tk) : "l"(ticket) : "memory");
    s_last = (tk == gridDim.x - 1);
  }
  __syncthreads();
  if (s_last && threadIdx.x < 6) {
    const int v = atomicAdd(out + threadIdx.x, 0);  // read at L2
    publish_slot(host_slots, threadIdx.x, static_cast<unsigned long long>(static_cast<unsigned int>(v)), key);
    if (threadIdx.x == 0) *ticket = 0;
  }
}

// ------------------------------------------------------------------------------------------ k1
// pclomp: int(floorf(x * inv) - float(min_b))   voxel_grid_covariance_omp_impl.hpp:219-227
// svn   : int(floorf(x * inv) - min_b)          voxel_grid_covariance_impl.hpp:222-225 (float - int -> float)
// Both subtractions are exact float operations for |index| < 2^24, so one code path serves both.
__device__ __forceinline__ uint32_t voxel_key(const float4 p, const GridMeta& m, int is_dense, bool& skipped) {
  const bool finite = isfinite(p.x) && isfinite(p.y) && isfinite(p.z);
  skipped = (m.flavour == FLAVOUR_SVN) ? !finite : (!is_dense && !finite);
  if (skipped) return m.total_cells;
  const int i0 = static_cast<int>(floorf(p.x * m.inv_leaf) - static_cast<float>(m.min_b[0]));
  const int i1 = static_cast<int>(floorf(p.y * m.inv_leaf) - static_cast<float>(m.min_b[1]));
  const int i2 = static_cast<int>(floorf(p.z * m.inv_leaf) - static_cast<float>(m.min_b[2]));
  return static_cast<uint32_t>(i0 * m.divb_mul[0] + i1 * m.divb_mul[1] + i2 * m.divb_mul[2]);
}

// ------------------------------------------------------------------------------------------ k2
// Stable LSD radix sort, 8-bit digits.  Pass 0 computes the keys from the points on the fly (k1 fused)
// and its scatter generates the point index instead of reading it.
constexpr int RS_THREADS = 256;
constexpr int RS_ITEMS = 8;
constexpr int RS_TILE = RS_THREADS * RS_ITEMS;

template <bool FIRST>
__global__ void __launch_bounds__(RS_THREADS) radix_hist_kernel(const float4* __restrict__ pts, GridMeta m, int is_dense, uint32_t* __restrict__ keys,
                                                                int n, int shift, uint32_t* __restrict__ hist, int nb,
                                                                unsigned int* __restrict__ n_skipped) {
  __shared__ uint32_t h[256];
  h[threadIdx.x] = 0;
  __syncthreads();
  const int base = blockIdx.x * RS_TILE;
  uint32_t skipped_here = 0;
#pragma unroll
  for (int i = 0; i < RS_ITEMS; ++i) {
    const int idx = base + i * RS_THREADS + threadIdx.x;
    if (idx < n) {
      uint32_t key;
      if (FIRST) {
        bool sk;
        key = voxel_key(__ldg(pts + idx), m, is_dense, sk);
        keys[idx] = key;
        skipped_here += sk ? 1u : 0u;
      } else {
        key = __ldg(keys + idx);
      }
      atomicAdd(&h[(key >> shift) & 255u], 1u);
    }
  }
  if (FIRST && skipped_here) atomicAdd(n_skipped, skipped_here);
  __syncthreads();
  hist[threadIdx.x * nb + blockIdx.x] = h[threadIdx.x];
}

// One block per digit: exclusive scan of that digit's per-tile counts (in place) + the digit total.
__global__ void __launch_bounds__(256) radix_scan_kernel(uint32_t* __restrict__ hist, int nb, uint32_t* __restrict__ digit_totals) {
  __shared__ uint32_t warp_sums[8];
  __shared__ uint32_t carry_s;
  uint32_t* row = hist + static_cast<size_t>(blockIdx.x) * nb;
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  if (tid == 0) carry_s = 0;
  __syncthreads();
  for (int base = 0; base < nb; base += 256) {
    const int i = base + tid;
    const uint32_t v = (i < nb) ? row[i] : 0u;
    uint32_t incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
    if (lane == 31) warp_sums[w] = incl;
    __syncthreads();
    uint32_t woff = 0, total = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) { const uint32_t c = warp_sums[k]; if (k < w) woff += c; total += c; }
    const uint32_t carry = carry_s;
    if (i < nb) row[i] = carry + woff + incl - v;
    __syncthreads();
    if (tid == 0) carry_s = carry + total;
    __syncthreads();
  }
  if (tid == 0) digit_totals[blockIdx.x] = carry_s;
}

template <bool FIRST>
__global__ void __launch_bounds__(RS_THREADS) radix_scatter_kernel(const uint32_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                                                                   uint32_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out, int n,
                                                                   int shift, const uint32_t* __restrict__ hist, int nb,
                                                                   const uint32_t* __restrict__ digit_totals) {
  __shared__ uint32_t cnt[RS_THREADS / 32][256];
  __shared__ uint32_t digit_base[256];
  __shared__ uint32_t ws[8];
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  {  // exclusive scan of the 256 digit totals
    const uint32_t v = __ldg(digit_totals + tid);
    uint32_t incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
    if (lane == 31) ws[w] = incl;
#pragma unroll
    for (int i = 0; i < RS_THREADS / 32; ++i) cnt[i][tid] = 0;
    __syncthreads();
    uint32_t woff = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) if (k < w) woff += ws[k];
    digit_base[tid] = woff + incl - v;
  }
  __syncthreads();
  const int base = blockIdx.x * RS_TILE + w * (32 * RS_ITEMS);
  uint32_t key[RS_ITEMS], val[RS_ITEMS], rank[RS_ITEMS], peers[RS_ITEMS];
  const uint32_t lt = (1u << lane) - 1u;
  // all loads and all eight peer masks first (independent: they pipeline), then the serial shared-memory counter chain
#pragma unroll
  for (int i = 0; i < RS_ITEMS; ++i) {
    const int idx = base + i * 32 + lane;
    const bool valid = idx < n;
    key[i] = valid ? __ldg(keys_in + idx) : 0u;
    val[i] = valid ? (FIRST ? static_cast<uint32_t>(idx) : __ldg(vals_in + idx)) : 0u;
  }
#pragma unroll
  for (int i = 0; i < RS_ITEMS; ++i) {
    const bool valid = base + i * 32 + lane < n;
    const uint32_t d = valid ? ((key[i] >> shift) & 255u) : 256u;
    peers[i] = __match_any_sync(0xffffffffu, d);
  }
#pragma unroll
  for (int i = 0; i < RS_ITEMS; ++i) {
    const bool valid = base + i * 32 + lane < n;
    const uint32_t d = (key[i] >> shift) & 255u;
    uint32_t prefix = 0;
    if (valid) prefix = cnt[w][d];
    __syncwarp();
    rank[i] = prefix + __popc(peers[i] & lt);
    if (valid && (peers[i] & lt) == 0) cnt[w][d] = prefix + __popc(peers[i]);
    __syncwarp();
  }
  __syncthreads();
  {
    uint32_t run = digit_base[tid] + __ldg(hist + tid * nb + blockIdx.x);
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

// sorted_xyz[i] = double(pts[vals[i]]): the finalize kernel then streams each voxel's points sequentially,
// already widened to FP64 (the float -> double conversions leave the per-voxel serial chains).
__global__ void __launch_bounds__(256) gather_points_kernel(const float4* __restrict__ pts, const uint32_t* __restrict__ vals, int n,
                                                            double* __restrict__ sorted_xyz) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    const float4 p = __ldg(pts + __ldg(vals + i));
    double* o = sorted_xyz + 3 * static_cast<size_t>(i);
    o[0] = static_cast<double>(p.x); o[1] = static_cast<double>(p.y); o[2] = static_cast<double>(p.z);
  }
}

// Single-block exclusive scan (coalesced, 1024 elements per step with a running carry).  Optional total.
__global__ void __launch_bounds__(1024) scan_small_kernel(uint32_t* __restrict__ data, int n, uint32_t* __restrict__ total_out) {
  __shared__ uint32_t warp_sums[32];
  __shared__ uint32_t carry_s;
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  if (tid == 0) carry_s = 0;
  __syncthreads();
  for (int base = 0; base < n; base += 1024) {
    const int i = base + tid;
    const uint32_t v = (i < n) ? data[i] : 0u;
    uint32_t incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
    if (lane == 31) warp_sums[w] = incl;
    __syncthreads();
    if (w == 0) {
      const uint32_t x = warp_sums[lane];
      uint32_t xi = x;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, xi, o); if (lane >= o) xi += t; }
      warp_sums[lane] = xi - x;  // exclusive offsets of the 32 warps
    }
    __syncthreads();
    const uint32_t carry = carry_s;
    if (i < n) data[i] = carry + warp_sums[w] + incl - v;
    __syncthreads();
    // step total = exclusive offset of the last warp + that warp's inclusive sum (held by thread 1023)
    if (tid == 1023) carry_s = carry + warp_sums[31] + incl;
    __syncthreads();
  }
  if (tid == 0 && total_out) *total_out = carry_s;
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

// pcl::VoxelGrid centroid pass (PCL >= 1.14 filters/impl/voxel_grid.hpp, CentroidPoint<PointT>/AccumulatorXYZ): one output
// point per occupied leaf, in ascending leaf index = float32 sum of the leaf's points / float(count).  One thread per
// leaf walks its segment in sorted order (stable sort: input order inside a leaf).  The last block publishes the count.
__global__ void __launch_bounds__(256) centroid_kernel(const float4* __restrict__ pts, const uint32_t* __restrict__ vals,
                                                       const uint32_t* __restrict__ seg_start, const unsigned int* __restrict__ counters,
                                                       uint32_t n_points, float4* __restrict__ out, unsigned int* __restrict__ ticket,
                                                       volatile unsigned long long* host_slots, unsigned long long key) {
  const uint32_t V = counters[1];
  const uint32_t n_valid_pts = n_points - counters[0];
  for (uint32_t v = blockIdx.x * blockDim.x + threadIdx.x; v < V; v += gridDim.x * blockDim.x) {
    const uint32_t lo = seg_start[v], hi = (v + 1 < V) ? seg_start[v + 1] : n_valid_pts;
    float sx = 0.0f, sy = 0.0f, sz = 0.0f;
    for (uint32_t i = lo; i < hi; ++i) {
      const float4 p = __ldg(pts + __ldg(vals + i));
      sx = sx + p.x; sy = sy + p.y; sz = sz + p.z;
    }
    const float cnt = static_cast<float>(hi - lo);
    out[v] = make_float4(__fdiv_rn(sx, cnt), __fdiv_rn(sy, cnt), __fdiv_rn(sz, cnt), 1.0f);
  }
  __shared__ bool last;
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    last = atomicAdd(ticket, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (last && threadIdx.x == 0) {
    *ticket = 0;
    publish_slot(host_slots, 12, V, key);
  }
}

}  // namespace ndtb
