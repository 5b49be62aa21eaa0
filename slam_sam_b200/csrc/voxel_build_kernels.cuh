// Target voxel-map build on the GPU, kernels of the sort half:
//   k0 bbox_kernel     : float min/max over the target (pcl::getMinMax3D); its last block derives the whole grid geometry and
//                        the radix plan ON THE DEVICE (BuildMeta) -- no host round trip inside a build
//   k1 radix_hist      : per-tile digit histograms; pass 0 computes the per-point linear voxel index, float32 -> int32 exactly
//                        as the reference (voxel_key)
//   k2 radix_scan      : one block per digit: exclusive scan of that digit's per-tile counts
//   k3 radix_scatter   : stable LSD scatter of packed (key << 32 | point index) elements, reordered by digit in shared
//                        memory so that the global stores are coalesced; 8192-element tiles, digits of <= 10 bits
//   k4 seg_kernel      : segment heads -> leaf numbering in ascending linear index (the std::map order of the reference), ONE
//                        kernel with a decoupled look-back over the tiles; its last tile fixes the leaf count
//   k5 hash_clear      : empties exactly the part of the table the build uses
// The statistics kernel (sums in input order, eigen-inflation, inverse, hash insert) is in voxel_finalize.cu.
// Roofline: HBM streaming. Algorithmic bytes B_build = 2*16*N_t + 64*V.
#pragma once
#include "ndtb_internal.cuh"
#include "build_api.cuh"
#include "../../include/ndtb.h"

namespace ndtb {

// ------------------------------------------------------------------------------------------ k0
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
  B.n_skipped = 0; B.V = 0; B.n_valid = 0; B.chunk_counter = 0; B.tile_ticket = 0;
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
  B.final_buf = B.passes & 1;
  if (B.passes > prm.passes_queued) B.error = BUILD_NEED_PASSES + B.passes;
}

// pcl::getMinMax3D.  Block reduction, 6 atomics per block; the last block writes BuildMeta.
constexpr int BBOX_THREADS = 256;
__global__ void __launch_bounds__(BBOX_THREADS) bbox_kernel(const float4* __restrict__ pts, int* __restrict__ bbox /*6 ordered ints*/,
                                                            BuildMeta* __restrict__ meta, unsigned int* __restrict__ ticket, BuildParams prm) {
  __shared__ float s_mn[BBOX_THREADS / 32][3], s_mx[BBOX_THREADS / 32][3];
  __shared__ bool s_last;
  pdl_enter();
  const int n = prm.n;
  float mn[3] = {3.402823466e+38f, 3.402823466e+38f, 3.402823466e+38f};
  float mx[3] = {-3.402823466e+38f, -3.402823466e+38f, -3.402823466e+38f};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const float4 p = __ldg(pts + i);
    if (!prm.is_dense && !(isfinite(p.x) && isfinite(p.y) && isfinite(p.z))) continue;
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
    atomicMin(bbox + threadIdx.x, float_to_ordered(v));
  } else if (threadIdx.x < 6) {
    float v = s_mx[0][threadIdx.x - 3];
    for (int k = 1; k < BBOX_THREADS / 32; ++k) v = fmaxf(v, s_mx[k][threadIdx.x - 3]);
    atomicMax(bbox + threadIdx.x, float_to_ordered(v));
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned int tk;
    asm volatile("atom.acq_rel.gpu.global.add.u32 %0, [%1], 1;" : "=r"(tk) : "l"(ticket) : "memory");
    s_last = (tk == gridDim.x - 1);
  }
  __syncthreads();
  if (s_last && threadIdx.x == 0) {
    int bb[6];
#pragma unroll
    for (int a = 0; a < 6; ++a) bb[a] = atomicAdd(bbox + a, 0);  // read at L2
    BuildMeta B;
    make_build_meta(B, bb, prm);
    *meta = B;
    // re-arm for the next build: box back to (+FLT_MAX, -FLT_MAX), ticket to zero
#pragma unroll
    for (int a = 0; a < 3; ++a) { bbox[a] = 0x7f7fffff; bbox[3 + a] = static_cast<int>(0xff7fffffu) ^ 0x7fffffff; }
    *ticket = 0;
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
// Stable LSD radix sort of packed 64-bit elements (voxel key << 32 | point index) on digits of BuildMeta::digit_bits (<= 10)
// bits: the 20 key bits of a 1 M-point / 1 m map take two passes of 10 bits.  Tiles of 8192 elements (one 512-thread CTA,
// 123 tiles for 1 M points: a single wave).  Per pass: per-tile digit histograms, a scan over the tiles, and the scatter.
// A pass the plan does not need (pass >= passes), or any pass of a build that failed on the device (overflow), returns at once.
constexpr int RS_THREADS = 512;
constexpr int RS_WARPS = RS_THREADS / 32;
constexpr int RS_ITEMS = 16;
constexpr int RS_TILE = RS_THREADS * RS_ITEMS;
constexpr int RS_MAX_BINS = 1 << RADIX_MAX_BITS;

// hist[d * nb + tile].  Pass 0 computes the voxel keys from the points on the fly and writes the packed elements.
template <bool FIRST>
__global__ void __launch_bounds__(RS_THREADS) radix_hist_kernel(const float4* __restrict__ pts, BuildMeta* __restrict__ meta, int is_dense,
                                                                unsigned long long* __restrict__ pairs0, const unsigned long long* __restrict__ pairs1,
                                                                int n, int pass, uint32_t* __restrict__ hist, int nb) {
  __shared__ uint32_t h[RS_MAX_BINS];
  pdl_enter();
  if (meta->error || pass >= meta->passes) return;
  const int dbits = meta->digit_bits, bins = 1 << dbits, shift = 32 + pass * dbits;
  const unsigned long long mask = static_cast<unsigned long long>(bins - 1);
  for (int b = threadIdx.x; b < bins; b += RS_THREADS) h[b] = 0;
  __syncthreads();
  const int base = blockIdx.x * RS_TILE;
  if (FIRST) {
    const GridMeta m = meta->m;
    uint32_t skipped_here = 0;
#pragma unroll 4
    for (int i = 0; i < RS_ITEMS; ++i) {
      const int idx = base + i * RS_THREADS + threadIdx.x;
      if (idx < n) {
        bool sk;
        const uint32_t key = voxel_key(__ldg(pts + idx), m, is_dense, sk);
        const unsigned long long e = (static_cast<unsigned long long>(key) << 32) | static_cast<unsigned int>(idx);
        pairs0[idx] = e;
        skipped_here += sk ? 1u : 0u;
        atomicAdd(&h[(e >> shift) & mask], 1u);
      }
    }
    if (skipped_here) atomicAdd(&meta->n_skipped, skipped_here);
  } else {
    const unsigned long long* in = (pass & 1) ? pairs1 : pairs0;
#pragma unroll 4
    for (int i = 0; i < RS_ITEMS; ++i) {
      const int idx = base + i * RS_THREADS + threadIdx.x;
      if (idx < n) atomicAdd(&h[(__ldg(in + idx) >> shift) & mask], 1u);
    }
  }
  __syncthreads();
  for (int b = threadIdx.x; b < bins; b += RS_THREADS) hist[static_cast<size_t>(b) * nb + blockIdx.x] = h[b];
}

// One WARP per digit: exclusive scan of that digit's per-tile counts (in place, 32 tiles per step) + the digit total.
__global__ void __launch_bounds__(256) radix_scan_kernel(const BuildMeta* __restrict__ meta, int pass, uint32_t* __restrict__ hist, int nb,
                                                         uint32_t* __restrict__ digit_totals) {
  pdl_enter();
  if (meta->error || pass >= meta->passes) return;
  const int lane = threadIdx.x & 31;
  const int d = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (d >= (1 << meta->digit_bits)) return;
  uint32_t* row = hist + static_cast<size_t>(d) * nb;
  uint32_t carry = 0;
  for (int base = 0; base < nb; base += 32) {
    const int i = base + lane;
    const uint32_t v = (i < nb) ? row[i] : 0u;
    uint32_t incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
    if (i < nb) row[i] = carry + incl - v;
    carry += __shfl_sync(0xffffffffu, incl, 31);
  }
  if (lane == 0) digit_totals[d] = carry;
}

// Dynamic shared memory of the scatter (<= 136 KB): elems[RS_TILE] u64 | wcnt[RS_WARPS][bins] | goff[bins] | lbase[bins] | ws[2 * RS_WARPS]
constexpr size_t RS_SCATTER_SMEM = static_cast<size_t>(RS_TILE) * 8 + (static_cast<size_t>(RS_WARPS + 2) * RS_MAX_BINS + 2 * RS_WARPS) * sizeof(uint32_t);
// The tile is ranked (stable: warp w owns 512 consecutive elements, round i its i-th 32; __match_any_sync gives the rank among
// the 32, a per-warp shared-memory counter the rank among the warp's earlier rounds), reordered by digit IN SHARED MEMORY and
// written out from there, so that consecutive threads store consecutive elements of one digit: 8-byte coalesced stores.
__global__ void __launch_bounds__(RS_THREADS) radix_scatter_kernel(const BuildMeta* __restrict__ meta, int pass, unsigned long long* __restrict__ pairs0,
                                                                   unsigned long long* __restrict__ pairs1, int n, const uint32_t* __restrict__ hist,
                                                                   int nb, const uint32_t* __restrict__ digit_totals) {
  extern __shared__ __align__(16) unsigned char rs_raw[];
  pdl_enter();
  if (meta->error || pass >= meta->passes) return;
  unsigned long long* elems = reinterpret_cast<unsigned long long*>(rs_raw);
  uint32_t* wcnt = reinterpret_cast<uint32_t*>(rs_raw + static_cast<size_t>(RS_TILE) * 8);   // [RS_WARPS][bins]
  const int dbits = meta->digit_bits, bins = 1 << dbits, shift = 32 + pass * dbits;
  uint32_t* goff = wcnt + RS_WARPS * bins;
  uint32_t* lbase = goff + bins;
  uint32_t* ws = lbase + bins;
  const unsigned long long mask = static_cast<unsigned long long>(bins - 1);
  const unsigned long long* in = (pass & 1) ? pairs1 : pairs0;
  unsigned long long* out = (pass & 1) ? pairs0 : pairs1;
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  for (int i = tid; i < RS_WARPS * bins; i += RS_THREADS) wcnt[i] = 0;
  const int tile_base = blockIdx.x * RS_TILE;
  const int base = tile_base + w * (32 * RS_ITEMS);
  const int tile_cnt = min(RS_TILE, n - tile_base);
  unsigned long long e[RS_ITEMS];
  uint32_t rank[RS_ITEMS];
#pragma unroll
  for (int i = 0; i < RS_ITEMS; ++i) {
    const int idx = base + i * 32 + lane;
    e[i] = (idx < n) ? __ldg(in + idx) : ~0ull;
  }
  __syncthreads();
  const uint32_t lt = (1u << lane) - 1u;
  uint32_t* my = wcnt + w * bins;
#pragma unroll
  for (int i = 0; i < RS_ITEMS; ++i) {
    const bool valid = base + i * 32 + lane < n;
    const uint32_t d = valid ? static_cast<uint32_t>((e[i] >> shift) & mask) : static_cast<uint32_t>(bins);
    const uint32_t peers = __match_any_sync(0xffffffffu, d);
    uint32_t prefix = 0;
    if (valid) prefix = my[d];
    __syncwarp();
    rank[i] = prefix + __popc(peers & lt);
    if (valid && (peers & lt) == 0) my[d] = prefix + __popc(peers);
    __syncwarp();
  }
  __syncthreads();
  // per digit: exclusive offsets of the warps inside the tile, the digit's count in the tile, then the exclusive scan over the
  // digits (position of the digit's run in the reordered tile) and the global offset of that run.  Thread t owns digits 2t, 2t+1.
  {
    uint32_t cnt2[2] = {0, 0}, tot2[2] = {0, 0};
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int d = 2 * tid + j;
      if (d < bins) {
        uint32_t run = 0;
#pragma unroll
        for (int k = 0; k < RS_WARPS; ++k) { const uint32_t c = wcnt[k * bins + d]; wcnt[k * bins + d] = run; run += c; }
        cnt2[j] = run;
        tot2[j] = __ldg(digit_totals + d);
      }
    }
    const uint32_t sc = cnt2[0] + cnt2[1], st = tot2[0] + tot2[1];
    uint32_t ic = sc, it = st;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t a2 = __shfl_up_sync(0xffffffffu, ic, o), b2 = __shfl_up_sync(0xffffffffu, it, o);
      if (lane >= o) { ic += a2; it += b2; }
    }
    if (lane == 31) { ws[w] = ic; ws[RS_WARPS + w] = it; }
    __syncthreads();
    uint32_t oc = ic - sc, ot = it - st;   // exclusive inside the warp
#pragma unroll
    for (int k = 0; k < RS_WARPS; ++k) if (k < w) { oc += ws[k]; ot += ws[RS_WARPS + k]; }
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int d = 2 * tid + j;
      if (d < bins) {
        lbase[d] = oc;                                                                              // start of digit d in the reordered tile
        goff[d] = ot + __ldg(hist + static_cast<size_t>(d) * nb + blockIdx.x) - oc;               // global position = goff[d] + local position
      }
      oc += cnt2[j]; ot += tot2[j];
    }
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < RS_ITEMS; ++i) {
    if (base + i * 32 + lane < n) {
      const uint32_t d = static_cast<uint32_t>((e[i] >> shift) & mask);
      elems[lbase[d] + my[d] + rank[i]] = e[i];
    }
  }
  __syncthreads();
  for (int j = tid; j < tile_cnt; j += RS_THREADS) {
    const unsigned long long v = elems[j];
    out[goff[(v >> shift) & mask] + j] = v;
  }
}

// ------------------------------------------------------------------------------------------ k4
// Segment heads of the sorted list -> seg_start[v], vox_key[v] with v ascending = ascending linear index.  One kernel:
// tiles take their index from a ticket (so a tile only ever waits for tiles that are already running), count their heads,
// and get the number of heads before them by a decoupled look-back over one 64-bit word per tile, 256 predecessors at a time:
//   word = epoch (30 bits) << 34 | flag (2 bits: 1 = tile aggregate, 2 = inclusive prefix) << 32 | value (32 bits)
// A word of another build's epoch reads as "not yet published", so the array is never cleared.  The tile holding the end of
// the list publishes the leaf count.
constexpr int SEG_THREADS = 256;
constexpr int SEG_ITEMS = 16;
constexpr int SEG_TILE = SEG_THREADS * SEG_ITEMS;

__global__ void __launch_bounds__(SEG_THREADS) seg_kernel(BuildMeta* __restrict__ meta, const unsigned long long* __restrict__ pairs0,
                                                          const unsigned long long* __restrict__ pairs1, int n, unsigned long long* __restrict__ state,
                                                          unsigned int epoch, uint32_t* __restrict__ seg_start, uint32_t* __restrict__ vox_key) {
  __shared__ uint32_t warp_sums[SEG_THREADS / 32];
  __shared__ uint32_t s_tile;
  __shared__ uint32_t s_wsum[SEG_THREADS / 32];
  __shared__ int s_wfirst[SEG_THREADS / 32];
  pdl_enter();
  if (meta->error) return;
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  if (tid == 0) s_tile = atomicAdd(&meta->tile_ticket, 1u);
  __syncthreads();
  const uint32_t tile = s_tile;
  const unsigned long long* pairs = meta->final_buf ? pairs1 : pairs0;
  const uint32_t sentinel = meta->m.total_cells;
  const int base = static_cast<int>(tile) * SEG_TILE + tid * SEG_ITEMS;
  uint32_t prev = (base > 0 && base - 1 < n) ? static_cast<uint32_t>(__ldg(pairs + base - 1) >> 32) : 0xFFFFFFFFu;
  uint32_t heads = 0;   // bit i: element base + i starts a leaf
  uint32_t c = 0;
  uint32_t k[SEG_ITEMS];
#pragma unroll
  for (int i = 0; i < SEG_ITEMS; ++i) {
    k[i] = (base + i < n) ? static_cast<uint32_t>(__ldg(pairs + base + i) >> 32) : sentinel;
    const bool hd = (base + i < n) && k[i] < sentinel && (base + i == 0 || k[i] != prev);
    prev = k[i];
    heads |= hd ? (1u << i) : 0u;
    c += hd ? 1u : 0u;
  }
  uint32_t incl = c;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
  if (lane == 31) warp_sums[w] = incl;
  __syncthreads();
  uint32_t woff = 0, total = 0;
#pragma unroll
  for (int j = 0; j < SEG_THREADS / 32; ++j) { const uint32_t t = warp_sums[j]; if (j < w) woff += t; total += t; }
  // publish, then look back with the WHOLE block: thread i examines predecessor tile - 1 - i (one L2 round trip covers 256 tiles)
  const unsigned long long tag = static_cast<unsigned long long>(epoch & 0x3fffffffu) << 34;
  uint32_t excl = 0;
  if (tile == 0) {
    if (tid == 0) asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(state), "l"(tag | (2ull << 32) | total) : "memory");
  } else {
    if (tid == 0) asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(state + tile), "l"(tag | (1ull << 32) | total) : "memory");
    int hi_t = static_cast<int>(tile) - 1;   // newest predecessor not yet accounted for
    for (;;) {
      const int t = hi_t - tid;              // thread 0 looks at the nearest predecessor
      unsigned long long wv = 0;
      unsigned int flag = 2;                 // positions before tile 0 count as "inclusive, value 0"
      if (t >= 0) {
        do {                                 // the tile holds an earlier ticket: it is running and will publish
          asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(wv) : "l"(state + t) : "memory");
          flag = ((wv >> 34) == (tag >> 34)) ? (static_cast<unsigned int>(wv >> 32) & 3u) : 0u;
        } while (flag == 0);
      }
      const unsigned int incl_mask = __ballot_sync(0xffffffffu, flag == 2);
      const int first_incl = __ffs(incl_mask) - 1;
      uint32_t contrib = (t >= 0 && (first_incl < 0 || lane <= first_incl)) ? static_cast<uint32_t>(wv) : 0u;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) contrib += __shfl_xor_sync(0xffffffffu, contrib, o);
      if (lane == 0) { s_wsum[w] = contrib; s_wfirst[w] = first_incl; }
      __syncthreads();
      bool found = false;
#pragma unroll
      for (int j = 0; j < SEG_THREADS / 32; ++j) {
        if (!found) { excl += s_wsum[j]; found = s_wfirst[j] >= 0; }
      }
      __syncthreads();
      if (found) break;
      hi_t -= SEG_THREADS;
    }
    if (tid == 0) asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(state + tile), "l"(tag | (2ull << 32) | (excl + total)) : "memory");
  }
  if (tid == 0) {
    const uint32_t n_tiles = static_cast<uint32_t>((n + SEG_TILE - 1) / SEG_TILE);
    if (tile == n_tiles - 1) {
      meta->V = excl + total;
      meta->tile_ticket = 0;   // every tile has taken its ticket: re-arm (a later launch on this grid starts from zero)
    }
  }
  uint32_t v = excl + woff + (incl - c);
#pragma unroll
  for (int i = 0; i < SEG_ITEMS; ++i)
    if (heads & (1u << i)) { seg_start[v] = static_cast<uint32_t>(base + i); vox_key[v] = k[i]; ++v; }
}

// ------------------------------------------------------------------------------------------ k5
__global__ void __launch_bounds__(256) hash_clear_kernel(const BuildMeta* __restrict__ meta, HashSlot* __restrict__ table) {
  pdl_enter();
  if (meta->error) return;
  const size_t cap = static_cast<size_t>(meta->m.hash_mask) + 1;   // slots of 8 bytes, cap >= 16: a multiple of 2
  uint4* t = reinterpret_cast<uint4*>(table);
  const uint4 e = make_uint4(HASH_EMPTY, 0xFFFFFFFFu, HASH_EMPTY, 0xFFFFFFFFu);
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < cap / 2; i += static_cast<size_t>(gridDim.x) * blockDim.x) t[i] = e;
}

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
