// Host driver of the target voxel-map build: setInputTarget -> VoxelGridCovariance::filter()
// (ndt_omp.h:324-330, voxel_grid_covariance_omp_impl.hpp:55-364; svn_ndt_impl.hpp:137-159,
// voxel_grid_covariance_impl.hpp:77-379).  See voxel_build_kernels.cuh / voxel_finalize.cu for the kernels.
//
// A build is two launches queued back to back that never wait for the host: the cooperative sort kernel (bbox -> geometry ->
// keys -> radix sort -> segments -> sums of the heavy leaves, sort_kernel.cuh) and the statistics kernel (voxel_finalize.cu).  The grid
// geometry, the radix plan, the leaf count and the hash capacity are derived on the device (BuildMeta).  The host waits ONCE,
// for the BuildMeta the last kernel publishes into pinned self-validating slots.
#include "voxel_build_kernels.cuh"
#include "sort_kernel.cuh"
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <algorithm>

namespace ndtb {

// Development aid: NDTB_BUILD_TIMING=1 records a CUDA event after every launch of a build and prints the per-kernel times.
static const bool g_build_timing = getenv("NDTB_BUILD_TIMING") != nullptr;
static cudaEvent_t g_tev[32];
static const char* g_tname[32];
static int g_tn = 0;
static void tmark(const char* name, cudaStream_t s) {
  if (!g_build_timing || g_tn >= 32) return;
  if (!g_tev[g_tn]) cudaEventCreate(&g_tev[g_tn]);
  cudaEventRecord(g_tev[g_tn], s); g_tname[g_tn] = name; ++g_tn;
}
static void tprint() {
  if (!g_build_timing || g_tn < 2) { g_tn = 0; return; }
  cudaEventSynchronize(g_tev[g_tn - 1]);
  float tot = 0; cudaEventElapsedTime(&tot, g_tev[0], g_tev[g_tn - 1]);
  fprintf(stderr, "[build timing] total %.1f us:", tot * 1e3f);
  for (int i = 1; i < g_tn; ++i) { float ms = 0; cudaEventElapsedTime(&ms, g_tev[i - 1], g_tev[i]); fprintf(stderr, " %s %.1f", g_tname[i], ms * 1e3f); }
  fprintf(stderr, "\n"); g_tn = 0;
}

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) return NDTB_ERR_CUDA; } while (0)

static int ensure_capacity(DeviceGrid& g, size_t n, bool voxel_arrays) {
  if (!g.d_bbox) {
    CK(cudaMalloc(&g.d_bbox, 16 * sizeof(int)));                                  // [0..5] box, [8..9] barrier / exit counters of the sort kernel
    int init[16] = {0};
    init[0] = init[1] = init[2] = 0x7f7fffff;                                     // ordered encoding of +FLT_MAX
    init[3] = init[4] = init[5] = static_cast<int>(0xff7fffffu) ^ 0x7fffffff;     // ordered encoding of -FLT_MAX
    CK(cudaMemcpy(g.d_bbox, init, sizeof(init), cudaMemcpyHostToDevice));         // the sort kernel re-arms both after every build
    CK(cudaMalloc(&g.d_meta, sizeof(BuildMeta)));
    CK(cudaMemset(g.d_meta, 0, sizeof(BuildMeta)));
    CK(cudaMalloc(&g.digit_base, RS_MAX_BINS * sizeof(uint32_t)));
    CK(cudaMallocHost(&g.h_slots, 2 * BUILD_META_WORDS * sizeof(unsigned long long)));
    std::memset(const_cast<unsigned long long*>(g.h_slots), 0, 2 * BUILD_META_WORDS * sizeof(unsigned long long));
  }
  if (n > g.pts_alloc) {
    for (int i = 0; i < 2; ++i) { cudaFree(g.pairs[i]); g.pairs[i] = nullptr; }
    cudaFree(g.seg_start); cudaFree(g.tile_heads); cudaFree(g.hist); cudaFree(g.c_key);
    // nothing dangles if an allocation below fails: the capacity is only set once all of them succeeded
    g.seg_start = nullptr; g.tile_heads = nullptr; g.hist = nullptr; g.c_key = nullptr;
    g.pts_alloc = 0;
    const size_t cap = n + n / 8 + 1024;
    CK(cudaMalloc(&g.c_key, cap * 4));  // leaf keys, written with the segments
    for (int i = 0; i < 2; ++i) CK(cudaMalloc(&g.pairs[i], cap * 8));
    CK(cudaMalloc(&g.seg_start, (cap + 1) * 4));
    const size_t nb = (cap + RS_TILE - 1) / RS_TILE + 1;
    CK(cudaMalloc(&g.tile_heads, nb * sizeof(uint32_t)));
    CK(cudaMalloc(&g.hist, nb * RS_MAX_BINS * 4));
    g.pts_alloc = cap;
  }
  if (voxel_arrays && n > g.vox_alloc) {  // voxel arrays: V <= n (not needed by the down-sampling scratch)
    const size_t cap = n + n / 8 + 1024;
    cudaFree(g.recs); cudaFree(g.c_n); cudaFree(g.c_mean); cudaFree(g.c_cov); cudaFree(g.c_icov);
    cudaFree(g.c_evecs); cudaFree(g.c_evals); cudaFree(g.table); cudaFree(g.vox_sums); cudaFree(g.sorted_pts);
    g.vox_sums = nullptr; g.sorted_pts = nullptr; g.recs = nullptr; g.c_n = nullptr; g.c_mean = g.c_cov = g.c_icov = g.c_evecs = g.c_evals = nullptr; g.table = nullptr;
    g.vox_alloc = 0; g.table_alloc = 0;
    CK(cudaMalloc(&g.vox_sums, cap * 72));
    CK(cudaMalloc(&g.sorted_pts, cap * sizeof(float4)));
    CK(cudaMalloc(&g.recs, cap * sizeof(VoxRec)));
    CK(cudaMalloc(&g.c_n, cap * 4));
    CK(cudaMalloc(&g.c_mean, cap * 24));
    CK(cudaMalloc(&g.c_cov, cap * 72));
    CK(cudaMalloc(&g.c_icov, cap * 72));
    CK(cudaMalloc(&g.c_evecs, cap * 72));
    CK(cudaMalloc(&g.c_evals, cap * 24));
    // hash table for the worst case: 2^m >= 2 n slots (the build uses, and clears, 2^m' >= 2 min(n, cells) of them)
    size_t tcap = 16;
    while (tcap < 2 * cap) tcap <<= 1;
    CK(cudaMalloc(&g.table, tcap * sizeof(HashSlot)));
    g.table_alloc = tcap;
    g.vox_alloc = cap;
  }
  return NDTB_OK;
}

void free_device_grid(DeviceGrid& g) {
  for (int i = 0; i < 2; ++i) cudaFree(g.pairs[i]);
  cudaFree(g.seg_start); cudaFree(g.tile_heads); cudaFree(g.hist); cudaFree(g.digit_base); cudaFree(g.d_meta);
  if (g.h_slots) cudaFreeHost(const_cast<unsigned long long*>(g.h_slots));
  cudaFree(g.vox_sums); cudaFree(g.sorted_pts); cudaFree(g.recs); cudaFree(g.table); cudaFree(g.c_key); cudaFree(g.c_n); cudaFree(g.c_mean); cudaFree(g.c_cov);
  cudaFree(g.c_icov); cudaFree(g.c_evecs); cudaFree(g.c_evals); cudaFree(g.d_bbox);
  const uint64_t keep = g.launches;
  g = DeviceGrid();
  g.launches = keep;
}

// Poll the BuildMeta slots (see publish_slot); a stream query every 4096 spins catches a failed kernel.
static int wait_meta(DeviceGrid& g, unsigned long long key, cudaStream_t stream, BuildMeta* out) {
  unsigned int spins = 0;
  unsigned long long words[BUILD_META_WORDS];
  for (int i = BUILD_META_WORDS - 1; i >= 0; --i) {
    for (;;) {
      const unsigned long long w0 = g.h_slots[2 * i], w1 = g.h_slots[2 * i + 1];
      if ((w0 ^ w1) == key) { words[i] = w0; break; }
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
  std::memcpy(out, words, sizeof(BuildMeta));
  return NDTB_OK;
}

// Co-resident CTAs of the sort kernel on the current device (one per SM: 137 KB of shared memory each).
static int sort_grid_limit() {
  static int limit[MAX_DEVICES] = {};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= MAX_DEVICES) return 0;
  if (!limit[dev]) {
    int occ = 0, sms = 0;
    if (cudaFuncSetAttribute(sort_segments_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(SORT_SMEM)) != cudaSuccess) return 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, sort_segments_kernel, RS_THREADS, SORT_SMEM) != cudaSuccess) return 0;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return 0;
    limit[dev] = occ * sms;
  }
  return limit[dev];
}

// Queues the sort half of a build: bbox -> geometry -> voxel keys -> stable radix sort -> segments (one per occupied leaf,
// ascending linear index), ONE cooperative launch (sort_kernel.cuh).  Leaves BuildMeta, pairs[final_buf], seg_start, c_key
// (and sorted_pts / the emptied hash table for a map build) ready on the stream; nothing is waited for.
static int queue_sort_into_segments(DeviceGrid& g, const float4* d_pts, int ni, float resolution, int min_pts, double eig_mult, int flavour,
                                    int is_dense, bool map_build, cudaStream_t stream, int* launches) {
  SortArgs a;
  a.pts = d_pts;
  a.prm.resolution = resolution;
  a.prm.inv_leaf = 1.0f / resolution;  // pcl::VoxelGrid::setLeafSize
  a.prm.min_pts = min_pts; a.prm.flavour = flavour; a.prm.is_dense = is_dense; a.prm.eig_mult = eig_mult;
  a.prm.n = ni; a.prm.reserved = 0;
  a.bbox = g.d_bbox;
  a.sync = reinterpret_cast<unsigned int*>(g.d_bbox + 8);
  a.meta = g.d_meta;
  a.pairs[0] = g.pairs[0]; a.pairs[1] = g.pairs[1];
  a.sorted_pts = map_build ? g.sorted_pts : nullptr;
  a.hist = g.hist;
  a.digit_totals = g.digit_base;
  a.tile_heads = g.tile_heads;
  a.seg_start = g.seg_start;
  a.vox_key = g.c_key;
  a.table = map_build ? g.table : nullptr;
  a.vox_sums = map_build ? g.vox_sums : nullptr;
  a.nb = (ni + RS_TILE - 1) / RS_TILE;
  static const bool dbg_on = getenv("NDTB_SORT_DEBUG") != nullptr;
  static unsigned long long* d_dbg = nullptr;
  if (dbg_on && !d_dbg) cudaMalloc(&d_dbg, 64 * sizeof(unsigned long long));
  if (dbg_on) cudaMemsetAsync(d_dbg, 0, 64 * sizeof(unsigned long long), stream);
  a.dbg = dbg_on ? d_dbg : nullptr;
  const int limit = sort_grid_limit();
  if (limit <= 0) return NDTB_ERR_CUDA;
  tmark("start", stream);
  void* args[] = {&a};
  CK(cudaLaunchCooperativeKernel(reinterpret_cast<const void*>(sort_segments_kernel), dim3(std::min(a.nb, limit)), dim3(RS_THREADS), args, SORT_SMEM, stream));
  tmark("sort", stream);
  if (dbg_on) {
    unsigned long long st[64];
    cudaStreamSynchronize(stream);
    cudaMemcpy(st, d_dbg, sizeof(st), cudaMemcpyDeviceToHost);
    fprintf(stderr, "[sort dbg n=%d grid=%d] ns between stamps of CTA 0:", ni, std::min(a.nb, limit));
    for (int i = 1; i < 64 && st[i]; ++i) fprintf(stderr, " %llu", st[i] - st[i - 1]);
    fprintf(stderr, "\n");
  }
  *launches = 1;
  return NDTB_OK;
}

// Takes the published BuildMeta over: error -> return code.  A barrier time-out leaves the counters of the sort kernel in an
// unknown state: they are re-armed here, off the hot path.
static int adopt_meta(DeviceGrid& g, const BuildMeta& B, cudaStream_t stream) {
  if (B.error == NDTB_ERR_CUDA) {
    int init[16] = {0};
    init[0] = init[1] = init[2] = 0x7f7fffff;
    init[3] = init[4] = init[5] = static_cast<int>(0xff7fffffu) ^ 0x7fffffff;
    cudaStreamSynchronize(stream);
    cudaMemcpy(g.d_bbox, init, sizeof(init), cudaMemcpyHostToDevice);
  }
  if (B.error) return B.error;
  g.meta = B.m;
  g.V = B.V; g.n_valid = B.n_valid; g.n_skipped = B.n_skipped; g.n_flagged = B.n_flagged; g.final_buf = B.final_buf;
  g.table_cap = static_cast<size_t>(B.m.hash_mask) + 1;
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
  int launches = 0;
  rc = queue_sort_into_segments(g, d_pts, ni, resolution, min_pts, eig_mult, flavour, is_dense, true, stream, &launches);
  if (rc != NDTB_OK) return rc;
  const unsigned long long key = 0x9E3779B97F4A7C15ull * (++g.slot_seq) | 1ull;
  launch_voxel_finalize(g, static_cast<uint32_t>(ni), key, stream);
  tmark("stats", stream);
  CK(cudaGetLastError());
  g.launches += static_cast<uint64_t>(launches) + 1;
  g.last_build_launches = launches + 1;
  BuildMeta B;
  rc = wait_meta(g, key, stream, &B);
  tprint();
  if (rc != NDTB_OK) return rc;
  rc = adopt_meta(g, B, stream);
  if (rc != NDTB_OK) return rc;
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
  int launches = 0;
  rc = queue_sort_into_segments(g, d_pts, ni, leaf, 0, 0.0, FLAVOUR_PCLOMP, is_dense, false, stream, &launches);
  if (rc != NDTB_OK) return rc;
  const unsigned long long key = 0x9E3779B97F4A7C15ull * (++g.slot_seq) | 1ull;
  const int blocks = std::min((ni + 255) / 256, 148 * 8);
  launch_pdl(centroid_kernel, dim3(blocks), dim3(256), 0, stream, d_pts, g.d_meta, g.pairs[0], g.pairs[1], g.seg_start, static_cast<uint32_t>(ni), d_out, g.h_slots, key);
  CK(cudaGetLastError());
  g.launches += static_cast<uint64_t>(launches) + 1;
  BuildMeta B;
  rc = wait_meta(g, key, stream, &B);
  if (rc != NDTB_OK) return rc;
  rc = adopt_meta(g, B, stream);
  if (rc != NDTB_OK) return rc;
  *n_out = static_cast<size_t>(B.V);
  return NDTB_OK;
}

}  // namespace ndtb
