// Host driver of the target voxel-map build: setInputTarget -> VoxelGridCovariance::filter()
// (ndt_omp.h:324-330, voxel_grid_covariance_omp_impl.hpp:55-364; svn_ndt_impl.hpp:137-159,
// voxel_grid_covariance_impl.hpp:77-379).  See voxel_build_kernels.cuh / voxel_finalize.cu for the kernels.
//
// The kernels of one build are queued back to back and never wait for the host: the grid geometry, the radix plan, the leaf
// count and the hash capacity are derived on the device (BuildMeta).  The host waits ONCE, for the BuildMeta the last kernel
// publishes into pinned self-validating slots.  The only thing the host has to guess is how many radix passes to queue
// (DeviceGrid::pass_hint: digits of <= 10 bits, 2 passes up to 1 M cells); a wrong guess is reported by the device and the
// build is queued again with the right count.
#include "voxel_build_kernels.cuh"
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
    CK(cudaMalloc(&g.d_bbox, 6 * sizeof(int)));
    int init[6];
    init[0] = init[1] = init[2] = 0x7f7fffff;                                     // ordered encoding of +FLT_MAX
    init[3] = init[4] = init[5] = static_cast<int>(0xff7fffffu) ^ 0x7fffffff;     // ordered encoding of -FLT_MAX
    CK(cudaMemcpy(g.d_bbox, init, sizeof(init), cudaMemcpyHostToDevice));         // the bbox kernel re-arms it after every build
    CK(cudaMalloc(&g.d_meta, sizeof(BuildMeta)));
    CK(cudaMemset(g.d_meta, 0, sizeof(BuildMeta)));
    CK(cudaMalloc(&g.digit_base, RS_MAX_BINS * sizeof(uint32_t)));
    CK(cudaMallocHost(&g.h_slots, 2 * BUILD_META_WORDS * sizeof(unsigned long long)));
    std::memset(const_cast<unsigned long long*>(g.h_slots), 0, 2 * BUILD_META_WORDS * sizeof(unsigned long long));
    static bool configured = false;
    if (!configured) {
      CK(cudaFuncSetAttribute(radix_scatter_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(RS_SCATTER_SMEM)));
      configured = true;
    }
  }
  if (n > g.pts_alloc) {
    for (int i = 0; i < 2; ++i) { cudaFree(g.pairs[i]); g.pairs[i] = nullptr; }
    cudaFree(g.seg_start); cudaFree(g.seg_state); cudaFree(g.hist); cudaFree(g.c_key);
    // nothing dangles if an allocation below fails: the capacity is only set once all of them succeeded
    g.seg_start = nullptr; g.seg_state = nullptr; g.hist = nullptr; g.c_key = nullptr;
    g.pts_alloc = 0;
    const size_t cap = n + n / 8 + 1024;
    CK(cudaMalloc(&g.c_key, cap * 4));  // leaf keys, written with the segments
    for (int i = 0; i < 2; ++i) CK(cudaMalloc(&g.pairs[i], cap * 8));
    CK(cudaMalloc(&g.seg_start, (cap + 1) * 4));
    const size_t nb = (cap + RS_TILE - 1) / RS_TILE + 1;
    const size_t nseg = (cap + SEG_TILE - 1) / SEG_TILE + 1;
    CK(cudaMalloc(&g.seg_state, nseg * sizeof(unsigned long long)));
    CK(cudaMemset(g.seg_state, 0, nseg * sizeof(unsigned long long)));
    CK(cudaMalloc(&g.hist, nb * RS_MAX_BINS * 4));
    g.pts_alloc = cap;
  }
  if (voxel_arrays && n > g.vox_alloc) {  // voxel arrays: V <= n (not needed by the down-sampling scratch)
    const size_t cap = n + n / 8 + 1024;
    cudaFree(g.recs); cudaFree(g.c_n); cudaFree(g.c_mean); cudaFree(g.c_cov); cudaFree(g.c_icov);
    cudaFree(g.c_evecs); cudaFree(g.c_evals); cudaFree(g.table); cudaFree(g.vox_sums);
    g.vox_sums = nullptr; g.recs = nullptr; g.c_n = nullptr; g.c_mean = g.c_cov = g.c_icov = g.c_evecs = g.c_evals = nullptr; g.table = nullptr;
    g.vox_alloc = 0; g.table_alloc = 0;
    CK(cudaMalloc(&g.vox_sums, cap * 72));
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
  cudaFree(g.seg_start); cudaFree(g.seg_state); cudaFree(g.hist); cudaFree(g.digit_base); cudaFree(g.d_meta);
  if (g.h_slots) cudaFreeHost(const_cast<unsigned long long*>(g.h_slots));
  cudaFree(g.vox_sums); cudaFree(g.recs); cudaFree(g.table); cudaFree(g.c_key); cudaFree(g.c_n); cudaFree(g.c_mean); cudaFree(g.c_cov);
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

// Queues bbox -> geometry -> voxel keys -> stable radix sort -> segments (one per occupied leaf, ascending linear index).
// Leaves BuildMeta, keys/vals[final_buf], seg_start, c_key ready on the stream; nothing is waited for.
static int queue_sort_into_segments(DeviceGrid& g, const float4* d_pts, int ni, float resolution, int min_pts, double eig_mult, int flavour,
                                    int is_dense, int passes, cudaStream_t stream, int* launches) {
  BuildParams prm;
  prm.resolution = resolution;
  prm.inv_leaf = 1.0f / resolution;  // pcl::VoxelGrid::setLeafSize
  prm.min_pts = min_pts; prm.flavour = flavour; prm.is_dense = is_dense; prm.eig_mult = eig_mult;
  prm.n = ni; prm.passes_queued = passes;
  const int bb_blocks = std::min((ni + BBOX_THREADS - 1) / BBOX_THREADS, 148 * 4);
  tmark("start", stream);
  launch_pdl(bbox_kernel, dim3(bb_blocks), dim3(BBOX_THREADS), 0, stream, d_pts, g.d_bbox, g.d_meta, &g.d_meta->tickets[0], prm);
  tmark("bbox", stream);
  const int nb = (ni + RS_TILE - 1) / RS_TILE;
  for (int p = 0; p < passes; ++p) {
    if (p == 0) launch_pdl(radix_hist_kernel<true>, dim3(nb), dim3(RS_THREADS), 0, stream, d_pts, g.d_meta, is_dense, g.pairs[0], g.pairs[1], ni, 0, g.hist, nb);
    else launch_pdl(radix_hist_kernel<false>, dim3(nb), dim3(RS_THREADS), 0, stream, static_cast<const float4*>(nullptr), g.d_meta, is_dense, g.pairs[0], g.pairs[1], ni, p, g.hist, nb);
    tmark("hist", stream);
    launch_pdl(radix_scan_kernel, dim3(RS_MAX_BINS / 8), dim3(256), 0, stream, g.d_meta, p, g.hist, nb, g.digit_base);
    tmark("scan", stream);
    launch_pdl(radix_scatter_kernel, dim3(nb), dim3(RS_THREADS), RS_SCATTER_SMEM, stream, g.d_meta, p, g.pairs[0], g.pairs[1], ni, g.hist, nb, g.digit_base);
    tmark("scatter", stream);
  }
  const int sb = (ni + SEG_TILE - 1) / SEG_TILE;
  launch_pdl(seg_kernel, dim3(sb), dim3(SEG_THREADS), 0, stream, g.d_meta, g.pairs[0], g.pairs[1], ni, g.seg_state, ++g.epoch, g.seg_start, g.c_key);
  tmark("seg", stream);
  CK(cudaGetLastError());
  *launches = 1 + 3 * passes + 1;
  return NDTB_OK;
}

// Takes the published BuildMeta over: error -> return code (a wrong pass guess asks for a retry through *retry_passes).
static int adopt_meta(DeviceGrid& g, const BuildMeta& B, int* retry_passes) {
  *retry_passes = 0;
  if (B.error >= BUILD_NEED_PASSES) { *retry_passes = B.error - BUILD_NEED_PASSES; return NDTB_OK; }
  if (B.error) return B.error;
  g.meta = B.m;
  g.V = B.V; g.n_valid = B.n_valid; g.n_skipped = B.n_skipped; g.n_flagged = B.n_flagged; g.final_buf = B.final_buf;
  g.table_cap = static_cast<size_t>(B.m.hash_mask) + 1;
  g.pass_hint = B.passes;   // the next build of a map like this one queues exactly the passes it needs
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
  for (int attempt = 0; attempt < 2; ++attempt) {
    int launches = 0;
    rc = queue_sort_into_segments(g, d_pts, ni, resolution, min_pts, eig_mult, flavour, is_dense, g.pass_hint, stream, &launches);
    if (rc != NDTB_OK) return rc;
    launch_pdl(hash_clear_kernel, dim3(148 * 2), dim3(256), 0, stream, g.d_meta, g.table);
    tmark("clear", stream);
    const unsigned long long key = 0x9E3779B97F4A7C15ull * (++g.slot_seq) | 1ull;
    launch_voxel_stats(g, d_pts, static_cast<uint32_t>(ni), key, stream);
    tmark("stats", stream);
    CK(cudaGetLastError());
    g.launches += static_cast<uint64_t>(launches) + 3;
    g.last_build_launches = launches + 3;
    BuildMeta B;
    rc = wait_meta(g, key, stream, &B);
    tprint();
    if (rc != NDTB_OK) return rc;
    int retry = 0;
    rc = adopt_meta(g, B, &retry);
    if (rc != NDTB_OK) return rc;
    if (!retry) { g.built = true; return NDTB_OK; }
    g.pass_hint = retry;
  }
  return NDTB_ERR_CUDA;
}

// pcl::VoxelGrid::filter on the device: same bbox / index / sort / segment machinery as the map build, then centroids.
// `g` is scratch (its NDT voxel arrays are not allocated); d_out must hold n points; *n_out = occupied leaves.
int voxelgrid_downsample(DeviceGrid& g, const float4* d_pts, size_t n, float leaf, int is_dense, float4* d_out, size_t* n_out, cudaStream_t stream) {
  if (n == 0) return NDTB_ERR_EMPTY;
  if (n > 0x7fffffffu || !(leaf > 0.0f)) return NDTB_ERR_INVALID;
  int rc = ensure_capacity(g, n, false);
  if (rc != NDTB_OK) return rc;
  const int ni = static_cast<int>(n);
  for (int attempt = 0; attempt < 2; ++attempt) {
    int launches = 0;
    rc = queue_sort_into_segments(g, d_pts, ni, leaf, 0, 0.0, FLAVOUR_PCLOMP, is_dense, g.pass_hint, stream, &launches);
    if (rc != NDTB_OK) return rc;
    const unsigned long long key = 0x9E3779B97F4A7C15ull * (++g.slot_seq) | 1ull;
    const int blocks = std::min((ni + 255) / 256, 148 * 8);
    launch_pdl(centroid_kernel, dim3(blocks), dim3(256), 0, stream, d_pts, g.d_meta, g.pairs[0], g.pairs[1], g.seg_start, static_cast<uint32_t>(ni), d_out, g.h_slots, key);
    CK(cudaGetLastError());
    g.launches += static_cast<uint64_t>(launches) + 1;
    BuildMeta B;
    rc = wait_meta(g, key, stream, &B);
    if (rc != NDTB_OK) return rc;
    int retry = 0;
    rc = adopt_meta(g, B, &retry);
    if (rc != NDTB_OK) return rc;
    if (!retry) { *n_out = static_cast<size_t>(B.V); return NDTB_OK; }
    g.pass_hint = retry;
  }
  return NDTB_ERR_CUDA;
}

}  // namespace ndtb
