// Target voxel-map build on the GPU: pieces shared by the kernels of a build.
//   make_build_meta : grid geometry from the bounding box + the radix plan, derived ON THE DEVICE (BuildMeta)
//   voxel_key       : per-point linear voxel index, float32 -> int32 exactly as the reference
//   centroid_kernel : pcl::VoxelGrid centroids (the down-sampling filter)
// The sort half of a build (bbox -> keys -> stable radix sort -> segments) is ONE cooperative kernel, sort_kernel.cuh; the
// statistics kernels (sums in input order, eigen-inflation, inverse, hash insert) are in voxel_finalize.cu.
// Roofline: HBM streaming. Algorithmic bytes B_build = 2*16*N_t + 64*V.
#pragma once
#include "ndtb_internal.cuh"
#include "build_api.cuh"
#include "../../include/ndtb.h"

namespace ndtb {

// Grid geometry from the bounding box, float maths as in applyFilter (voxel_grid_covariance_omp_impl.hpp:83-113 /
// voxel_grid_covariance_impl.hpp:103-140), and the sort plan.  One thread.
__device__ inline void make_build_meta(BuildMeta& B, const int* ordered_bbox, const BuildParams& prm) {
  float min_p[3], max_p[3];
#pragma unroll
  for (int a = 0; a < 3; ++a) { min_p[a] = ordered_to_float_bits(ordered_bbox[a]); max_p[a] = ordered_to_float_bits(ordered_bbox[3 + a]); }
  GridMeta& m = B.m;
  m.leaf = prm.resolution;
  m.inv_leaf = prm.inv_leaf;  // pcl::VoxelGrid::setLeafSize: 1.0f / leaf, computed by the host
  m.min_pts = prm.min_pts; m.flavour = prm.flavour; m.eig_mult = prm.eig_mult;
  m.hash_mask = 0; m.total_cells = 0;
  B.error = 0; B.bits = 0; B.digit_bits = 0; B.passes = 0; B.final_buf = 0;
  B.n_skipped = 0; B.V = 0; B.n_valid = 0; B.chunk_counter = 0; B.n_heavy = 0;
  B.tickets[0] = 0; B.tickets[1] = 0; B.tickets[2] = 0; B.n_flagged = 0;
  const long long dx = static_cast<long long>((max_p[0] - min_p[0]) * m.inv_leaf) + 1;
  const long long dy = static_cast<long long>((max_p[1] - min_p[1]) * m.inv_leaf) + 1;
  const long long dz = static_cast<long long>((max_p[2] - min_p[2]) * m.inv_leaf) + 1;
  if (prm.flavour == FLAVOUR_SVN && (dx < 0 || dy < 0 || dz < 0)) { B.error = NDTB_ERR_OVERFLOW; return; }
  if ((dx * dy * dz) > 2147483647ll) { B.error = NDTB_ERR_OVERFLOW; return; }
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    m.min_b[a] = static_cast<int>(floorf(min_p[a] * m.inv_leaf));
    m.max_b[a] = static_cast<int>(floorf(max_p[a] * m.inv_leaf));
    m.div_b[a] = m.max_b[a] - m.min_b[a] + 1;
    m.lo_bound[a] = static_cast<float>(m.min_b[a]) * m.leaf;
    m.hi_bound[a] = static_cast<float>(m.max_b[a] + 1) * m.leaf;
  }
  m.divb_mul[0] = 1; m.divb_mul[1] = m.div_b[0]; m.divb_mul[2] = m.div_b[0] * m.div_b[1];
  const long long cells = static_cast<long long>(m.div_b[0]) * m.div_b[1] * m.div_b[2];
  if (cells > 2147483647ll || cells <= 0) { B.error = NDTB_ERR_OVERFLOW; return; }
  m.total_cells = static_cast<uint32_t>(cells);
  {  // hash capacity 2^m >= 2 * min(N, cells): load factor <= 0.5, and whenever the cell space itself fits (cells <= 2^m) the
     // identity hash `key & mask` is collision-free -- every probe of the derivative kernels is a single load
    const unsigned long long need = 2ull * (static_cast<unsigned long long>(prm.n) < static_cast<unsigned long long>(cells) ? static_cast<unsigned long long>(prm.n) : static_cast<unsigned long long>(cells));
    unsigned long long cap = 16;
    while (cap < need) cap <<= 1;
    m.hash_mask = static_cast<uint32_t>(cap - 1);
  }
  int bits = 1;
  while (bits < 32 && (static_cast<unsigned long long>(m.total_cells) >> bits) != 0) ++bits;  // key <= total_cells (the "skipped" key)
  B.bits = bits;
  B.passes = (bits + RADIX_MAX_BITS - 1) / RADIX_MAX_BITS;
  B.digit_bits = (bits + B.passes - 1) / B.passes;
  B.final_buf = (B.passes - 1) & 1;   // pass p writes pairs[p & 1] (sort_kernel.cuh)
}

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

// Tiles of the radix sort: 8192 elements, one 512-thread CTA; digits of BuildMeta::digit_bits (<= 10) bits.
constexpr int RS_THREADS = 512;
constexpr int RS_WARPS = RS_THREADS / 32;
constexpr int RS_ITEMS = 16;
constexpr int RS_TILE = RS_THREADS * RS_ITEMS;
constexpr int RS_MAX_BINS = 1 << RADIX_MAX_BITS;
// Dynamic shared memory of the sort kernel (<= 136 KB): elems[RS_TILE] u64 | wcnt[RS_WARPS][bins] | goff[bins] | lbase[bins] | ws[2 * RS_WARPS]
constexpr size_t RS_SCATTER_SMEM = static_cast<size_t>(RS_TILE) * 8 + (static_cast<size_t>(RS_WARPS + 2) * RS_MAX_BINS + 2 * RS_WARPS) * sizeof(uint32_t);

// pcl::VoxelGrid centroid pass (PCL >= 1.14 filters/impl/voxel_grid.hpp, CentroidPoint<PointT>/AccumulatorXYZ): one output
// point per occupied leaf, in ascending leaf index = float32 sum of the leaf's points / float(count).  One thread per
// leaf walks its segment in sorted order (stable sort: input order inside a leaf).  The last block publishes BuildMeta.
__global__ void __launch_bounds__(256) centroid_kernel(const float4* __restrict__ pts, BuildMeta* __restrict__ meta, const unsigned long long* __restrict__ pairs0,
                                                       const unsigned long long* __restrict__ pairs1, const uint32_t* __restrict__ seg_start, uint32_t n_points,
                                                       float4* __restrict__ out, volatile unsigned long long* host_slots, unsigned long long key) {
  __shared__ bool last;
  pdl_enter();
  if (!meta->error) {
    const uint32_t V = meta->V;
    const uint32_t n_valid_pts = n_points - meta->n_skipped;
    const unsigned long long* pairs = meta->final_buf ? pairs1 : pairs0;
    for (uint32_t v = blockIdx.x * blockDim.x + threadIdx.x; v < V; v += gridDim.x * blockDim.x) {
      const uint32_t lo = seg_start[v], hi = (v + 1 < V) ? seg_start[v + 1] : n_valid_pts;
      float sx = 0.0f, sy = 0.0f, sz = 0.0f;
      for (uint32_t i = lo; i < hi; ++i) {
        const float4 p = __ldg(pts + static_cast<uint32_t>(__ldg(pairs + i)));
        sx = sx + p.x; sy = sy + p.y; sz = sz + p.z;
      }
      const float cnt = static_cast<float>(hi - lo);
      out[v] = make_float4(__fdiv_rn(sx, cnt), __fdiv_rn(sy, cnt), __fdiv_rn(sz, cnt), 1.0f);
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    last = atomicAdd(&meta->tickets[2], 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (last) {
    if (threadIdx.x == 0) meta->tickets[2] = 0;
    __syncthreads();
    const unsigned long long* words = reinterpret_cast<const unsigned long long*>(meta);
    for (int i = threadIdx.x; i < BUILD_META_WORDS; i += blockDim.x) publish_slot(host_slots, i, __ldcg(words + i), key);
  }
}

}  // namespace ndtb
