// Host driver of the target voxel-map build: setInputTarget -> VoxelGridCovariance::filter()
// (ndt_omp.h:324-330, voxel_grid_covariance_omp_impl.hpp:55-364; svn_ndt_impl.hpp:137-159,
// voxel_grid_covariance_impl.hpp:77-379).  See voxel_build_kernels.cuh for the kernels.
#include "voxel_build_kernels.cuh"
#include "../../include/ndtb.h"
#include <cmath>
#include <cstring>
#include <limits>
#include <algorithm>

namespace ndtb {

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) return NDTB_ERR_CUDA; } while (0)

static int ensure_capacity(DeviceGrid& g, size_t n, bool voxel_arrays) {
  if (!g.d_bbox) {
    CK(cudaMalloc(&g.d_bbox, 6 * sizeof(int)));
    CK(cudaMalloc(&g.d_counters, 4 * sizeof(unsigned int)));
    CK(cudaMallocHost(&g.h_pinned, 16 * sizeof(int)));
    CK(cudaMalloc(&g.digit_totals, 256 * sizeof(uint32_t)));
    CK(cudaMalloc(&g.d_ticket, 2 * sizeof(unsigned int)));
    CK(cudaMemset(g.d_ticket, 0, 2 * sizeof(unsigned int)));
    CK(cudaMallocHost(&g.h_slots, 32 * sizeof(unsigned long long)));
    std::memset(const_cast<unsigned long long*>(g.h_slots), 0, 32 * sizeof(unsigned long long));
  }
  if (n > g.pts_alloc) {
    for (int i = 0; i < 2; ++i) { cudaFree(g.keys[i]); cudaFree(g.vals[i]); g.keys[i] = nullptr; g.vals[i] = nullptr; }
    cudaFree(g.seg_start); cudaFree(g.block_counts); cudaFree(g.hist); cudaFree(g.sorted_xyz); cudaFree(g.c_key);
    // nothing dangles if an allocation below fails: the capacity is only set once all of them succeeded
    g.seg_start = nullptr; g.block_counts = nullptr; g.hist = nullptr; g.sorted_xyz = nullptr; g.c_key = nullptr;
    g.pts_alloc = 0;
    const size_t cap = n + n / 8 + 1024;
    CK(cudaMalloc(&g.sorted_xyz, cap * 3 * sizeof(double)));
    CK(cudaMalloc(&g.c_key, cap * 4));  // leaf keys, written with the segments
    for (int i = 0; i < 2; ++i) { CK(cudaMalloc(&g.keys[i], cap * 4)); CK(cudaMalloc(&g.vals[i], cap * 4)); }
    CK(cudaMalloc(&g.seg_start, (cap + 1) * 4));
    const size_t nb = (cap + RS_TILE - 1) / RS_TILE + 1;
    CK(cudaMalloc(&g.block_counts, nb * 4));
    CK(cudaMalloc(&g.hist, nb * 256 * 4));
    g.pts_alloc = cap;
  }
  if (voxel_arrays && n > g.vox_alloc) {  // voxel arrays: V <= n (not needed by the down-sampling scratch)
    const size_t cap = n + n / 8 + 1024;
    cudaFree(g.recs); cudaFree(g.c_n); cudaFree(g.c_mean); cudaFree(g.c_cov); cudaFree(g.c_icov);
    cudaFree(g.c_evecs); cudaFree(g.c_evals); cudaFree(g.vox_sums);
    g.recs = nullptr; g.c_n = nullptr; g.c_mean = g.c_cov = g.c_icov = g.c_evecs = g.c_evals = nullptr; g.vox_sums = nullptr;
    g.vox_alloc = 0;
    CK(cudaMalloc(&g.vox_sums, cap * 72));
    CK(cudaMalloc(&g.recs, cap * sizeof(VoxRec)));
    CK(cudaMalloc(&g.c_n, cap * 4));
    CK(cudaMalloc(&g.c_mean, cap * 24));
    CK(cudaMalloc(&g.c_cov, cap * 72));
    CK(cudaMalloc(&g.c_icov, cap * 72));
    CK(cudaMalloc(&g.c_evecs, cap * 72));
    CK(cudaMalloc(&g.c_evals, cap * 24));
    g.vox_alloc = cap;
  }
  return NDTB_OK;
}

void free_device_grid(DeviceGrid& g) {
  for (int i = 0; i < 2; ++i) { cudaFree(g.keys[i]); cudaFree(g.vals[i]); }
  cudaFree(g.seg_start); cudaFree(g.block_counts); cudaFree(g.hist); cudaFree(g.sorted_xyz); cudaFree(g.vox_sums); cudaFree(g.digit_totals); cudaFree(g.d_ticket);
  if (g.h_slots) cudaFreeHost(const_cast<unsigned long long*>(g.h_slots));
  cudaFree(g.recs); cudaFree(g.table); cudaFree(g.c_key); cudaFree(g.c_n); cudaFree(g.c_mean); cudaFree(g.c_cov);
  cudaFree(g.c_icov); cudaFree(g.c_evecs); cudaFree(g.c_evals); cudaFree(g.d_bbox); cudaFree(g.d_counters);
  if (g.h_pinned) cudaFreeHost(g.h_pinned);
  const uint64_t keep = g.launches;
  g = DeviceGrid();
  g.launches = keep;
}

// Poll `count` self-validating host slots starting at `first` (see publish_slot); a stream query
// every 4096 spins catches a failed kernel.
static int wait_slots(DeviceGrid& g, int first, int count, unsigned long long key, cudaStream_t stream) {
  unsigned int spins = 0;
  for (int i = first + count - 1; i >= first; --i) {
    for (;;) {
      const unsigned long long w0 = g.h_slots[2 * i], w1 = g.h_slots[2 * i + 1];
      if ((w0 ^ w1) == key) break;
      if ((++spins & 4095u) == 0) {
        const cudaError_t q = cudaStreamQuery(stream);
        if (q != cudaSuccess && q != cudaErrorNotReady) return NDTB_ERR_CUDA;
        if (q == cudaSuccess) {
          const unsigned long long x0 = g.h_slots[2 * i], x1 = g.h_slots[2 * i + 1];
          if ((x0 ^ x1) != key) return NDTB_ERR_CUDA;
        }
      }
    }
  }
  __sync_synchronize();
  return NDTB_OK;
}

// bbox -> grid geometry -> voxel keys -> stable radix sort -> segments (one per occupied leaf, ascending linear index).
// Leaves g.meta, g.keys/vals[*cur], g.seg_start, g.c_key and d_counters[0] (skipped points) / [1] (leaf count) ready on the stream.
static int sort_into_segments(DeviceGrid& g, const float4* d_pts, int ni, float resolution, int min_pts, double eig_mult, int flavour,
                              int is_dense, cudaStream_t stream, int* cur_out, int* passes_out) {
  const size_t n = static_cast<size_t>(ni);

  // ---- k0: bounding box (pcl::getMinMax3D); the last block publishes it to pinned host slots -----------
  {
    int init[6];
    init[0] = init[1] = init[2] = 0x7f7fffff;                                     // ordered encoding of +FLT_MAX
    init[3] = init[4] = init[5] = static_cast<int>(0xff7fffffu) ^ 0x7fffffff;     // ordered encoding of -FLT_MAX
    std::memcpy(g.h_pinned, init, sizeof(init));
    CK(cudaMemcpyAsync(g.d_bbox, g.h_pinned, sizeof(init), cudaMemcpyHostToDevice, stream));
    const int blocks = std::min((ni + BBOX_THREADS - 1) / BBOX_THREADS, 148 * 4);
    const unsigned long long key = 0x9E3779B97F4A7C15ull * (++g.slot_seq) | 1ull;
    bbox_kernel<<<blocks, BBOX_THREADS, 0, stream>>>(d_pts, ni, is_dense, g.d_bbox, g.d_ticket, g.h_slots, key);
    CK(cudaGetLastError());
    int rc2 = wait_slots(g, 0, 6, key, stream);
    if (rc2 != NDTB_OK) return rc2;
    for (int a = 0; a < 6; ++a) g.h_pinned[a] = static_cast<int>(static_cast<unsigned int>(g.h_slots[2 * a]));
  }
  float min_p[3], max_p[3];
  for (int a = 0; a < 3; ++a) { min_p[a] = ordered_to_float_bits(g.h_pinned[a]); max_p[a] = ordered_to_float_bits(g.h_pinned[3 + a]); }

  // ---- grid geometry on the host, float maths as in applyFilter :83-113 ---------------------
  GridMeta& m = g.meta;
  m.leaf = resolution;
  m.inv_leaf = 1.0f / resolution;  // pcl::VoxelGrid::setLeafSize
  m.min_pts = min_pts; m.flavour = flavour; m.eig_mult = eig_mult;
  const int64_t dx = static_cast<int64_t>((max_p[0] - min_p[0]) * m.inv_leaf) + 1;
  const int64_t dy = static_cast<int64_t>((max_p[1] - min_p[1]) * m.inv_leaf) + 1;
  const int64_t dz = static_cast<int64_t>((max_p[2] - min_p[2]) * m.inv_leaf) + 1;
  if (flavour == FLAVOUR_SVN && (dx < 0 || dy < 0 || dz < 0)) return NDTB_ERR_OVERFLOW;
  if ((dx * dy * dz) > static_cast<int64_t>(std::numeric_limits<int32_t>::max())) return NDTB_ERR_OVERFLOW;
  for (int a = 0; a < 3; ++a) {
    m.min_b[a] = static_cast<int>(std::floor(min_p[a] * m.inv_leaf));
    m.max_b[a] = static_cast<int>(std::floor(max_p[a] * m.inv_leaf));
    m.div_b[a] = m.max_b[a] - m.min_b[a] + 1;
    m.lo_bound[a] = static_cast<float>(m.min_b[a]) * m.leaf;
    m.hi_bound[a] = static_cast<float>(m.max_b[a] + 1) * m.leaf;
  }
  m.divb_mul[0] = 1; m.divb_mul[1] = m.div_b[0]; m.divb_mul[2] = m.div_b[0] * m.div_b[1];
  const int64_t cells = static_cast<int64_t>(m.div_b[0]) * m.div_b[1] * m.div_b[2];
  if (cells > static_cast<int64_t>(std::numeric_limits<int32_t>::max())) return NDTB_ERR_OVERFLOW;
  m.total_cells = static_cast<uint32_t>(cells);

  (void)n;
  CK(cudaMemsetAsync(g.d_counters, 0, 4 * sizeof(unsigned int), stream));

  // ---- k1 + k2: keys (fused into the first histogram) and the stable LSD radix sort over the significant key bits
  int bits = 1;
  while (bits < 32 && (static_cast<uint64_t>(m.total_cells) >> bits) != 0) ++bits;  // key <= total_cells
  const int passes = (bits + 7) / 8;
  const int nb = (ni + RS_TILE - 1) / RS_TILE;
  int cur = 0;
  for (int p = 0; p < passes; ++p) {
    if (p == 0) radix_hist_kernel<true><<<nb, RS_THREADS, 0, stream>>>(d_pts, m, is_dense, g.keys[0], ni, 0, g.hist, nb, g.d_counters);
    else radix_hist_kernel<false><<<nb, RS_THREADS, 0, stream>>>(nullptr, m, is_dense, g.keys[cur], ni, 8 * p, g.hist, nb, g.d_counters);
    radix_scan_kernel<<<256, 256, 0, stream>>>(g.hist, nb, g.digit_totals);
    if (p == 0) radix_scatter_kernel<true><<<nb, RS_THREADS, 0, stream>>>(g.keys[0], nullptr, g.keys[1], g.vals[1], ni, 0, g.hist, nb, g.digit_totals);
    else radix_scatter_kernel<false><<<nb, RS_THREADS, 0, stream>>>(g.keys[cur], g.vals[cur], g.keys[cur ^ 1], g.vals[cur ^ 1], ni, 8 * p, g.hist, nb, g.digit_totals);
    cur ^= 1;
  }

  // ---- k3: segments -> voxel numbering in ascending linear index; points gathered into sorted order -----
  const int sb = (ni + SEG_TILE - 1) / SEG_TILE;
  seg_count_kernel<<<sb, SEG_THREADS, 0, stream>>>(g.keys[cur], ni, m.total_cells, g.block_counts);
  scan_small_kernel<<<1, 1024, 0, stream>>>(g.block_counts, sb, g.d_counters + 1);
  seg_write_kernel<<<sb, SEG_THREADS, 0, stream>>>(g.keys[cur], ni, m.total_cells, g.block_counts, g.seg_start, g.c_key);
  CK(cudaGetLastError());
  *cur_out = cur;
  *passes_out = passes;
  return NDTB_OK;
}

int build_voxel_map(DeviceGrid& g, const float4* d_pts, size_t n, float resolution, int min_pts, double eig_mult, int flavour,
                    int is_dense, cudaStream_t stream) {
  g.built = false; g.V = 0; g.n_valid = 0;
  if (n == 0) return NDTB_ERR_EMPTY;
  if (n > 0x7fffffffu) return NDTB_ERR_INVALID;
  int rc = ensure_capacity(g, n, true);
  if (rc != NDTB_OK) return rc;
  const int ni = static_cast<int>(n);
  int cur = 0, passes = 0;
  rc = sort_into_segments(g, d_pts, ni, resolution, min_pts, eig_mult, flavour, is_dense, stream, &cur, &passes);
  if (rc != NDTB_OK) return rc;
  GridMeta& m = g.meta;
  // ---- hash capacity: 2^m >= 2 * min(N, cells)  (load factor <= 0.5); the table is first touched by the finalize kernel
  {
    const uint64_t need = 2ull * std::min<uint64_t>(n, static_cast<uint64_t>(m.total_cells));
    size_t cap = 16;
    while (cap < need) cap <<= 1;
    if (cap > g.table_alloc) {
      cudaFree(g.table);
      CK(cudaMalloc(&g.table, cap * sizeof(HashSlot)));
      g.table_alloc = cap;
    }
    g.table_cap = cap;
    m.hash_mask = static_cast<uint32_t>(cap - 1);
    CK(cudaMemsetAsync(g.table, 0xFF, cap * sizeof(HashSlot), stream));
  }
  gather_points_kernel<<<(ni + 255) / 256, 256, 0, stream>>>(d_pts, g.vals[cur], ni, g.sorted_xyz);

  // ---- k4: per-voxel statistics + hash insert; its last block publishes the counters to the host ---------
  const unsigned long long key = 0x9E3779B97F4A7C15ull * (++g.slot_seq) | 1ull;
  launch_voxel_finalize(g, static_cast<uint32_t>(ni), key, stream);
  g.launches += 1 + 3 * static_cast<uint64_t>(passes) + 4 + 2;  // bbox, radix passes, segments + gather, sums + finalize
  CK(cudaGetLastError());
  rc = wait_slots(g, 8, 3, key, stream);
  if (rc != NDTB_OK) return rc;
  for (int a = 0; a < 3; ++a) g.h_pinned[a] = static_cast<int>(static_cast<unsigned int>(g.h_slots[2 * (8 + a)]));
  g.V = static_cast<uint32_t>(g.h_pinned[1]);
  g.n_valid = static_cast<uint32_t>(g.h_pinned[2]);
  g.built = true;
  return NDTB_OK;
}

// pcl::VoxelGrid::filter on the device: same bbox / index / sort / segment machinery as the map build, then centroids.
// `g` is scratch (its NDT voxel arrays are not allocated); d_out must hold n points; *n_out = occupied leaves.
int voxelgrid_downsample(DeviceGrid& g, const float4* d_pts, size_t n, float leaf, int is_dense, float4* d_out, size_t* n_out, cudaStream_t stream) {
  if (n == 0) return NDTB_ERR_EMPTY;
  if (n > 0x7fffffffu || !(leaf > 0.0f)) return NDTB_ERR_INVALID;
  int rc = ensure_capacity(g, n, false);
  if (rc != NDTB_OK) return rc;
  const int ni = static_cast<int>(n);
  int cur = 0, passes = 0;
  rc = sort_into_segments(g, d_pts, ni, leaf, 0, 0.0, FLAVOUR_PCLOMP, is_dense, stream, &cur, &passes);
  if (rc != NDTB_OK) return rc;
  const unsigned long long key = 0x9E3779B97F4A7C15ull * (++g.slot_seq) | 1ull;
  const int blocks = std::min((ni + 255) / 256, 148 * 8);
  centroid_kernel<<<blocks, 256, 0, stream>>>(d_pts, g.vals[cur], g.seg_start, g.d_counters, static_cast<uint32_t>(ni), d_out, g.d_ticket + 1, g.h_slots, key);
  CK(cudaGetLastError());
  g.launches += 1 + 3 * static_cast<uint64_t>(passes) + 3 + 1;
  rc = wait_slots(g, 12, 1, key, stream);
  if (rc != NDTB_OK) return rc;
  *n_out = static_cast<size_t>(static_cast<unsigned int>(g.h_slots[2 * 12]));
  return NDTB_OK;
}

}  // namespace ndtb
