// The sort half of a voxel-map build as ONE cooperative kernel: bounding box -> grid geometry and radix
// plan -> stable LSD radix sort of (voxel key << 32 | point index) -> segment heads, with grid-wide barriers between the steps
// instead of kernel boundaries.  One 512-thread CTA per SM; a CTA owns the tiles t = blockIdx, blockIdx + grid, ... (8192
// elements each; the 123 tiles of a 1 M-point map are one tile per CTA, whose elements and ranks then stay in registers across
// the barriers).  Per radix pass: rank the tile by digit and publish its digit counts | barrier | one warp per digit scans the
// counts over the tiles | barrier | reorder the tile by digit in shared memory and store it coalesced (the final pass of a map
// build also stores the POINTS in sorted order) | barrier.  Nothing returns to the host in between; the number of passes is
// whatever the key range asks for.  A map build ends with the sums of the heavy leaves (two more barriers).
//   pcl::getMinMax3D + applyFilter geometry: voxel_grid_covariance_omp_impl.hpp:83-113 / voxel_grid_covariance_impl.hpp:103-140
//   leaf index of a point:                   voxel_grid_covariance_omp_impl.hpp:219-227 / voxel_grid_covariance_impl.hpp:222-225
//   leaves in ascending index (std::map):    voxel_grid_covariance_omp.h leaves_
// Launched with cudaLaunchCooperativeKernel (all CTAs co-resident or the launch fails -- no deadlock on the barriers); a barrier
// that does not complete within 2 s marks the build as failed and lets every CTA leave.
#pragma once
#include "voxel_build_kernels.cuh"

namespace ndtb {

struct SortArgs {
  const float4* pts;
  BuildParams prm;                 // prm.n points
  int* bbox;                       // 6 ordered ints, (+FLT_MAX, -FLT_MAX) between builds
  unsigned int* sync;              // [0] barrier arrivals, [1] exits; both zero between builds
  BuildMeta* meta;
  unsigned long long* pairs[2];    // pass p writes pairs[p & 1]
  float4* sorted_pts;              // map build: points in sorted order (w = point index); nullptr: not wanted
  uint32_t* hist;                  // [bins][nb]
  uint32_t* digit_totals;          // [bins]
  uint32_t* tile_heads;            // [nb]
  uint32_t* seg_start;
  uint32_t* vox_key;
  HashSlot* table;                 // cleared over the capacity the build uses; nullptr: no table (down-sampling)
  double* vox_sums;                // map build: [V][9] sums of the heavy leaves; nullptr: not wanted
  int nb;                          // tiles
  unsigned long long* dbg;         // development aid (NDTB_SORT_DEBUG): globaltimer stamps of CTA 0 at the phase boundaries
};

constexpr size_t SORT_SMEM = RS_SCATTER_SMEM;
constexpr unsigned long long SORT_BARRIER_TIMEOUT_NS = 2000000000ull;

__device__ __forceinline__ unsigned long long sort_globaltimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// All CTAs of the (cooperative) grid.  `target` counts the arrivals this CTA expects in total; *dead is set when the wait timed out.
__device__ __forceinline__ void sort_grid_barrier(unsigned int* counter, unsigned int& target, int* dead) {
  __syncthreads();
  if (threadIdx.x == 0) {
    target += gridDim.x;
    __threadfence();
    atomicAdd(counter, 1u);
    unsigned int seen, spins = 0;
    unsigned long long t0 = 0;
    for (;;) {
      asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(seen) : "l"(counter) : "memory");
      if (seen >= target || *dead) break;
      if ((++spins & 4095u) == 0) {
        const unsigned long long now = sort_globaltimer();
        if (!t0) t0 = now;
        if (now - t0 > SORT_BARRIER_TIMEOUT_NS) { *dead = 1; break; }
      }
    }
    __threadfence();
  }
  __syncthreads();
}

__global__ void __launch_bounds__(RS_THREADS, 1) sort_segments_kernel(const SortArgs a) {
  extern __shared__ __align__(16) unsigned char rs_raw[];
  __shared__ BuildMeta sB;
  __shared__ float s_mn[RS_WARPS][3], s_mx[RS_WARPS][3];
  __shared__ uint32_t s_warp[RS_WARPS];
  __shared__ int s_dead;
  unsigned long long* elems = reinterpret_cast<unsigned long long*>(rs_raw);
  uint32_t* wcnt = reinterpret_cast<uint32_t*>(rs_raw + static_cast<size_t>(RS_TILE) * 8);   // [RS_WARPS][bins]
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const int n = a.prm.n, nb = a.nb, G = gridDim.x;
  const bool keep = nb <= G;   // one tile per CTA: its elements and ranks stay in registers across the barriers
  unsigned int target = 0;
  if (tid == 0) s_dead = 0;
  int dbg_k = 0;
  auto stamp = [&]() { if (a.dbg && blockIdx.x == 0 && tid == 0 && dbg_k < 64) a.dbg[dbg_k] = sort_globaltimer(); ++dbg_k; };
  stamp();

  // ---------------------------------------------------------------------------------------------- bounding box
  {
    float mn[3] = {3.402823466e+38f, 3.402823466e+38f, 3.402823466e+38f};
    float mx[3] = {-3.402823466e+38f, -3.402823466e+38f, -3.402823466e+38f};
#pragma unroll 4
    for (int i = blockIdx.x * RS_THREADS + tid; i < n; i += G * RS_THREADS) {
      const float4 p = __ldg(a.pts + i);
      if (!a.prm.is_dense && !(isfinite(p.x) && isfinite(p.y) && isfinite(p.z))) continue;
      mn[0] = fminf(mn[0], p.x); mn[1] = fminf(mn[1], p.y); mn[2] = fminf(mn[2], p.z);
      mx[0] = fmaxf(mx[0], p.x); mx[1] = fmaxf(mx[1], p.y); mx[2] = fmaxf(mx[2], p.z);
    }
#pragma unroll
    for (int c = 0; c < 3; ++c) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        mn[c] = fminf(mn[c], __shfl_xor_sync(0xffffffffu, mn[c], o));
        mx[c] = fmaxf(mx[c], __shfl_xor_sync(0xffffffffu, mx[c], o));
      }
    }
    if (lane == 0) {
#pragma unroll
      for (int c = 0; c < 3; ++c) { s_mn[w][c] = mn[c]; s_mx[w][c] = mx[c]; }
    }
    __syncthreads();
    if (tid < 3) {
      float v = s_mn[0][tid];
      for (int k = 1; k < RS_WARPS; ++k) v = fminf(v, s_mn[k][tid]);
      atomicMin(a.bbox + tid, float_to_ordered(v));
    } else if (tid < 6) {
      float v = s_mx[0][tid - 3];
      for (int k = 1; k < RS_WARPS; ++k) v = fmaxf(v, s_mx[k][tid - 3]);
      atomicMax(a.bbox + tid, float_to_ordered(v));
    }
  }
  stamp();
  sort_grid_barrier(a.sync, target, &s_dead);
  stamp();

  // ---------------------------------------------------------------------------------------------- geometry + radix plan
  if (tid == 0) {
    int bb[6];
#pragma unroll
    for (int c = 0; c < 6; ++c) bb[c] = __ldcg(a.bbox + c);
    make_build_meta(sB, bb, a.prm);
    if (s_dead) sB.error = NDTB_ERR_CUDA;
    if (blockIdx.x == 0) *a.meta = sB;
  }
  __syncthreads();
  if (sB.error) {
    // every CTA has derived the same error; the box is re-armed once all of them have read it
    sort_grid_barrier(a.sync, target, &s_dead);
    if (blockIdx.x == 0 && tid == 0) {
#pragma unroll
      for (int c = 0; c < 3; ++c) { a.bbox[c] = 0x7f7fffff; a.bbox[3 + c] = static_cast<int>(0xff7fffffu) ^ 0x7fffffff; }
    }
    if (tid == 0) {
      const unsigned int t = atomicAdd(a.sync + 1, 1u);
      if (t == static_cast<unsigned int>(G) - 1u) { a.sync[0] = 0; a.sync[1] = 0; }
    }
    return;
  }
  const GridMeta m = sB.m;
  const int passes = sB.passes, dbits = sB.digit_bits, bins = 1 << dbits;
  const unsigned long long mask = static_cast<unsigned long long>(bins - 1);
  uint32_t* goff = wcnt + RS_WARPS * bins;
  uint32_t* lbase = goff + bins;
  uint32_t* ws = lbase + bins;
  const uint32_t lt = (1u << lane) - 1u;
  uint32_t* my = wcnt + w * bins;

  unsigned long long e[RS_ITEMS];
  uint32_t rank[RS_ITEMS];
  uint32_t cnt2[2] = {0, 0};
  uint32_t skipped_cta = 0;   // non-finite points this THREAD has seen in pass 0

  for (int p = 0; p < passes; ++p) {
    const int shift = 32 + p * dbits;
    const unsigned long long* in = a.pairs[(p + 1) & 1];
    unsigned long long* out = a.pairs[p & 1];
    // Loads tile T (pass 0: the voxel keys are computed from the points), ranks it by digit -- stable: warp w owns 512 consecutive
    // elements, round i its i-th 32; __match_any_sync gives the rank among the 32, a per-warp shared-memory counter the rank among
    // the warp's earlier rounds -- and turns the per-warp counters into exclusive offsets.  Thread t owns digits 2t, 2t + 1.
    auto load_rank = [&](int T, bool count_skipped) {
      for (int i = tid; i < RS_WARPS * bins; i += RS_THREADS) wcnt[i] = 0;
      const int base = T * RS_TILE + w * (32 * RS_ITEMS);
      if (p == 0) {
        uint32_t sk_n = 0;
#pragma unroll
        for (int i = 0; i < RS_ITEMS; ++i) {
          const int idx = base + i * 32 + lane;
          e[i] = ~0ull;
          if (idx < n) {
            bool sk;
            const uint32_t key = voxel_key(__ldg(a.pts + idx), m, a.prm.is_dense, sk);
            e[i] = (static_cast<unsigned long long>(key) << 32) | static_cast<unsigned int>(idx);
            sk_n += sk ? 1u : 0u;
          }
        }
        if (count_skipped) skipped_cta += sk_n;
      } else {
#pragma unroll
        for (int i = 0; i < RS_ITEMS; ++i) {
          const int idx = base + i * 32 + lane;
          e[i] = (idx < n) ? __ldcg(in + idx) : ~0ull;
        }
      }
      __syncthreads();
      stamp();
#pragma unroll
      for (int i = 0; i < RS_ITEMS; ++i) {
        const bool valid = base + i * 32 + lane < n;
        const uint32_t d = static_cast<uint32_t>((e[i] >> shift) & mask);
        // lanes holding the same digit: one ballot per digit bit (as fast as MATCH.ANY here, and independent of its throughput)
        uint32_t peers = __ballot_sync(0xffffffffu, valid);
#pragma unroll
        for (int kb = 0; kb < RADIX_MAX_BITS; ++kb) {
          if (kb < dbits) {
            const bool bit = (d >> kb) & 1u;
            const uint32_t mb = __ballot_sync(0xffffffffu, bit);
            peers &= bit ? mb : ~mb;
          }
        }
        uint32_t prefix = 0;
        if (valid) prefix = my[d];
        __syncwarp();
        rank[i] = prefix + __popc(peers & lt);
        if (valid && (peers & lt) == 0) my[d] = prefix + __popc(peers);
        __syncwarp();
      }
      __syncthreads();
      stamp();
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const int d = 2 * tid + j;
        cnt2[j] = 0;
        if (d < bins) {
          uint32_t run = 0;
#pragma unroll
          for (int k = 0; k < RS_WARPS; ++k) { const uint32_t c = wcnt[k * bins + d]; wcnt[k * bins + d] = run; run += c; }
          cnt2[j] = run;
        }
      }
      stamp();
    };

    // ---- count: digit counts of every tile ----------------------------------------------------------------------------------
    for (int T = blockIdx.x; T < nb; T += G) {
      load_rank(T, p == 0);
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const int d = 2 * tid + j;
        if (d < bins) a.hist[static_cast<size_t>(d) * nb + T] = cnt2[j];
      }
      if (!keep) __syncthreads();
    }
    stamp();
    sort_grid_barrier(a.sync, target, &s_dead);
    stamp();
    if (p == 0 && skipped_cta) atomicAdd(&a.meta->n_skipped, skipped_cta);   // per thread; non-finite points are rare
    if (p == 0 && tid == 0) {
      if (blockIdx.x == 0) {   // every CTA has read the box: re-arm it for the next build
#pragma unroll
        for (int c = 0; c < 3; ++c) { a.bbox[c] = 0x7f7fffff; a.bbox[3 + c] = static_cast<int>(0xff7fffffu) ^ 0x7fffffff; }
      }
    }
    // ---- scan: one warp per digit, exclusive scan of the digit's counts over the tiles (in place) + the digit total ------------
    for (int d = blockIdx.x * RS_WARPS + w; d < bins; d += G * RS_WARPS) {
      uint32_t* row = a.hist + static_cast<size_t>(d) * nb;
      uint32_t carry = 0;
      for (int b0 = 0; b0 < nb; b0 += 32) {
        const int i = b0 + lane;
        const uint32_t v = (i < nb) ? __ldcg(row + i) : 0u;
        uint32_t incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
        if (i < nb) row[i] = carry + incl - v;
        carry += __shfl_sync(0xffffffffu, incl, 31);
      }
      if (lane == 0) a.digit_totals[d] = carry;
    }
    stamp();
    sort_grid_barrier(a.sync, target, &s_dead);
    stamp();
    // ---- write: reorder the tile by digit in shared memory, store it coalesced ------------------------------------------------
    const bool carry_points = a.sorted_pts != nullptr && p == passes - 1;
    for (int T = blockIdx.x; T < nb; T += G) {
      if (!keep) load_rank(T, false);
      uint32_t tot2[2] = {0, 0};
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const int d = 2 * tid + j;
        if (d < bins) tot2[j] = __ldcg(a.digit_totals + d);
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
          lbase[d] = oc;                                                                         // start of digit d in the reordered tile
          goff[d] = ot + __ldcg(a.hist + static_cast<size_t>(d) * nb + T) - oc;                // global position = goff[d] + local position
        }
        oc += cnt2[j]; ot += tot2[j];
      }
      __syncthreads();
      stamp();
      const int tile_base = T * RS_TILE;
      const int base = tile_base + w * (32 * RS_ITEMS);
      const int tile_cnt = min(RS_TILE, n - tile_base);
#pragma unroll
      for (int i = 0; i < RS_ITEMS; ++i) {
        if (base + i * 32 + lane < n) {
          const uint32_t d = static_cast<uint32_t>((e[i] >> shift) & mask);
          elems[lbase[d] + my[d] + rank[i]] = e[i];
        }
      }
      __syncthreads();
      stamp();
      if (!carry_points) {
        for (int j = tid; j < tile_cnt; j += RS_THREADS) {
          const unsigned long long v = elems[j];
          out[goff[(v >> shift) & mask] + j] = v;
        }
      } else {
        // final pass of a map build: the points follow their elements (w <- point index), so that the statistics kernels read
        // every leaf as a contiguous run instead of chasing indices.  Sixteen independent gathers in flight per thread.
        float4 q[RS_ITEMS];
#pragma unroll
        for (int u = 0; u < RS_ITEMS; ++u) {
          const int j = tid + u * RS_THREADS;
          if (j < tile_cnt) q[u] = __ldg(a.pts + static_cast<uint32_t>(elems[j]));
        }
#pragma unroll
        for (int u = 0; u < RS_ITEMS; ++u) {
          const int j = tid + u * RS_THREADS;
          if (j < tile_cnt) {
            const unsigned long long v = elems[j];
            const uint32_t dst = goff[(v >> shift) & mask] + j;
            out[dst] = v;
            q[u].w = __uint_as_float(static_cast<uint32_t>(v));
            a.sorted_pts[dst] = q[u];
          }
        }
      }
      __syncthreads();
    }
    stamp();
    sort_grid_barrier(a.sync, target, &s_dead);
    stamp();
  }

  // ---------------------------------------------------------------------------------------------- segment heads
  // leaf v = the v-th distinct key of the sorted list (ascending linear index = the std::map order of the reference):
  // seg_start[v] = position of its first element, vox_key[v] = its key.  Thread t of a tile looks at 16 consecutive elements.
  {
    const unsigned long long* fin = a.pairs[(passes - 1) & 1];
    const uint32_t sentinel = m.total_cells;
    uint32_t k[RS_ITEMS];
    uint32_t heads = 0, c = 0, toff = 0, total = 0;
    uint32_t* skeys = reinterpret_cast<uint32_t*>(rs_raw);   // the tile's keys, key j at word j + j / 16 (conflict-free blocked reads)
    auto scan_tile = [&](int T) {
      const int tile_base = T * RS_TILE;
      __syncthreads();
#pragma unroll
      for (int i = 0; i < RS_ITEMS; ++i) {
        const int j = tid + i * RS_THREADS;
        const uint32_t key = (tile_base + j < n) ? static_cast<uint32_t>(__ldcg(fin + tile_base + j) >> 32) : sentinel;
        skeys[j + (j >> 4)] = key;
      }
      __syncthreads();
      const int base = tile_base + tid * RS_ITEMS;
      uint32_t prev = 0xFFFFFFFFu;
      if (tid > 0) prev = skeys[(tid * RS_ITEMS - 1) + ((tid * RS_ITEMS - 1) >> 4)];
      else if (base > 0 && base - 1 < n) prev = static_cast<uint32_t>(__ldcg(fin + base - 1) >> 32);
      heads = 0; c = 0;
#pragma unroll
      for (int i = 0; i < RS_ITEMS; ++i) {
        k[i] = skeys[tid * (RS_ITEMS + 1) + i];
        const bool hd = (base + i < n) && k[i] < sentinel && (base + i == 0 || k[i] != prev);
        prev = k[i];
        heads |= hd ? (1u << i) : 0u;
        c += hd ? 1u : 0u;
      }
      uint32_t incl = c;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
      __syncthreads();
      if (lane == 31) s_warp[w] = incl;
      __syncthreads();
      uint32_t woff = 0;
      total = 0;
#pragma unroll
      for (int j = 0; j < RS_WARPS; ++j) { const uint32_t t = s_warp[j]; if (j < w) woff += t; total += t; }
      toff = woff + (incl - c);
    };
    for (int T = blockIdx.x; T < nb; T += G) {
      scan_tile(T);
      if (tid == 0) a.tile_heads[T] = total;
    }
    stamp();
    sort_grid_barrier(a.sync, target, &s_dead);
    stamp();
    for (int T = blockIdx.x; T < nb; T += G) {
      if (!keep) scan_tile(T);
      uint32_t part = 0;
      for (int t = tid; t < T; t += RS_THREADS) part += __ldcg(a.tile_heads + t);
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
      __syncthreads();
      if (lane == 0) s_warp[w] = part;
      __syncthreads();
      uint32_t excl = 0;
#pragma unroll
      for (int j = 0; j < RS_WARPS; ++j) excl += s_warp[j];
      if (T == nb - 1 && tid == 0) a.meta->V = excl + total;
      uint32_t v = excl + toff;
      const int base = T * RS_TILE + tid * RS_ITEMS;
#pragma unroll
      for (int i = 0; i < RS_ITEMS; ++i)
        if (heads & (1u << i)) { a.seg_start[v] = static_cast<uint32_t>(base + i); a.vox_key[v] = k[i]; ++v; }
    }
  }
  // ---------------------------------------------------------------------------------------------- heavy leaves
  // Sum x and Sum x x^T of the leaves with more than HEAVY_LEAF_POINTS points, IN INPUT ORDER (FP64, one serial chain per component
  // exactly like the reference's loop, voxel_grid_covariance_omp_impl.hpp:235-239 / voxel_grid_covariance_impl.hpp:237-239): a
  // lane of the statistics kernel would walk such a leaf for microseconds while its 31 neighbours wait.  The leaves are listed
  // first (any order: it only decides which warp takes which leaf), then one warp per leaf: 32 lanes load 32 consecutive points
  // of sorted_pts and form the nine products; lane c < 9 runs chain c over them, so the nine chains advance in the same
  // instruction.  Two 32-point tiles are in flight ahead of the chain.
  if (a.vox_sums != nullptr) {
    stamp();
    sort_grid_barrier(a.sync, target, &s_dead);   // seg_start and the leaf count are complete
    stamp();
    const uint32_t V = __ldcg(&a.meta->V);
    const uint32_t n_valid_pts = static_cast<uint32_t>(n) - __ldcg(&a.meta->n_skipped);
    uint32_t* heavy_list = a.hist;                  // free since the last radix pass; V_heavy <= n / 49 < its n / 8 words
    for (uint32_t v = blockIdx.x * RS_THREADS + tid; v < V; v += static_cast<uint32_t>(G) * RS_THREADS) {
      const uint32_t lo = __ldcg(a.seg_start + v), hi = (v + 1 < V) ? __ldcg(a.seg_start + v + 1) : n_valid_pts;
      if (hi - lo > static_cast<uint32_t>(HEAVY_LEAF_POINTS)) heavy_list[atomicAdd(&a.meta->n_heavy, 1u)] = v;
    }
    sort_grid_barrier(a.sync, target, &s_dead);
    stamp();
    const uint32_t n_heavy = __ldcg(&a.meta->n_heavy);
    double (*s_prod)[9] = reinterpret_cast<double (*)[9]>(rs_raw + static_cast<size_t>(w) * 32 * 9 * sizeof(double));
    for (uint32_t hi_i = blockIdx.x * RS_WARPS + w; hi_i < n_heavy; hi_i += static_cast<uint32_t>(G) * RS_WARPS) {
      const uint32_t v = __ldcg(heavy_list + hi_i);
      const uint32_t wlo = __ldcg(a.seg_start + v);
      const int wn = static_cast<int>(((v + 1 < V) ? __ldcg(a.seg_start + v + 1) : n_valid_pts) - wlo);
      // the reference's Leaf starts cov_ at identity and adds x x^T onto it: the diagonal chains start at 1.0
      double acc = (lane == 3 || lane == 6 || lane == 8) ? 1.0 : 0.0;
      const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
      const float4* src = a.sorted_pts + wlo + lane;
      auto tile = [&](const float4 pt, int base) {
        const double x = static_cast<double>(pt.x), y = static_cast<double>(pt.y), z = static_cast<double>(pt.z);
        double* pr = s_prod[lane];
        pr[0] = x; pr[1] = y; pr[2] = z; pr[3] = x * x; pr[4] = x * y; pr[5] = x * z; pr[6] = y * y; pr[7] = y * z; pr[8] = z * z;
        __syncwarp();
        const int cnt = min(32, wn - base);
        if (lane < 9) {
          int kk = 0;
          for (; kk + 8 <= cnt; kk += 8) {
            double t8[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) t8[u] = s_prod[kk + u][lane];
#pragma unroll
            for (int u = 0; u < 8; ++u) acc += t8[u];
          }
          for (; kk < cnt; ++kk) acc += s_prod[kk][lane];
        }
        __syncwarp();
      };
      float4 pa = (lane < wn) ? __ldcg(src) : zero;
      float4 pb = (32 + lane < wn) ? __ldcg(src + 32) : zero;
      for (int base = 0; base < wn; base += 64) {
        const float4 na = (base + 64 + lane < wn) ? __ldcg(src + base + 64) : zero;
        const float4 nb2 = (base + 96 + lane < wn) ? __ldcg(src + base + 96) : zero;
        tile(pa, base);
        if (base + 32 < wn) tile(pb, base + 32);
        pa = na; pb = nb2;
      }
      if (lane < 9) a.vox_sums[9 * static_cast<size_t>(v) + lane] = acc;
    }
  }
  // ---------------------------------------------------------------------------------------------- empty hash table
  if (a.table) {   // exactly the part of the table the build uses; the stores drain while the statistics kernels start
    const size_t cap2 = (static_cast<size_t>(m.hash_mask) + 1) / 2;   // slots of 8 bytes, cap >= 16
    uint4* t = reinterpret_cast<uint4*>(a.table);
    const uint4 ev = make_uint4(HASH_EMPTY, 0xFFFFFFFFu, HASH_EMPTY, 0xFFFFFFFFu);
    for (size_t i = static_cast<size_t>(blockIdx.x) * RS_THREADS + tid; i < cap2; i += static_cast<size_t>(G) * RS_THREADS) t[i] = ev;
  }
  // ---------------------------------------------------------------------------------------------- leave
  stamp();
  if (tid == 0) {
    if (s_dead) a.meta->error = NDTB_ERR_CUDA;
    __threadfence();
    const unsigned int t = atomicAdd(a.sync + 1, 1u);
    if (t == static_cast<unsigned int>(G) - 1u) { a.sync[0] = 0; a.sync[1] = 0; }
  }
}

}  // namespace ndtb
