// pclomp::MultiGridNormalDistributionsTransform on the device (TIER IV's dynamic-map NDT, the variant autoware-style
// localisers drive with addTarget / removeTarget as map tiles stream in):
//   extern/ndt_omp/include/multigrid_pclomp/multigrid_ndt_omp.h:143-165 (setInputTarget / addTarget / removeTarget /
//     createVoxelKdtree) and multigrid_ndt_omp_impl.hpp:338-494 (computeDerivatives, FP64), :590-697 (point and pair
//     maths, FP64), :701-801 (computeHessian);
//   extern/ndt_omp/include/multigrid_pclomp/multi_voxel_grid_covariance_omp_impl.hpp:157-258, :291-439 (the set of
//     voxel grids, one per target id, and the radius search over their leaf centroids).
// Device layout: every target id owns a full voxel map (the same build as ndtb_set_target: sort, per-voxel FP64 sums,
// eigen-inflation, hash table) plus one float4 centroid per leaf (float running sum in input order / float n, the
// point the reference's kd-tree is built over).  ndtb_mg_create_kdtree freezes the current set into a device table of
// grid descriptors -- the "kd-tree": a radius search with radius = resolution = leaf size can only reach the 3x3x3
// cells around the query in each grid, so it is 27 hash probes per overlapping grid plus the float distance test
// FLANN applies (L2_Simple<float>, dist < float(r*r)).
// Kernel: one thread per source point (grid-stride), FP64 pair maths with a fixed operation order, 46 FP64 accumulators per thread, shuffle + shared-memory block reduction, block partials summed in
// block order by a second one-block kernel -> deterministic, ~1e-13 from the CPU restatement.
#include "handle.cuh"
#include "newton_math.cuh"
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

using namespace ndtb;

#define CKH(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { h->last_error = std::string(#x) + ": " + cudaGetErrorString(e_); return NDTB_ERR_CUDA; } } while (0)

namespace ndtb {

struct MgGridDev {
  GridMeta meta;
  const HashSlot* table;
  const double* c_mean;
  const double* c_icov;
  const float4* centroid;
};

struct MgGrid {
  DeviceGrid grid;
  float4* d_centroid = nullptr;
  size_t centroid_alloc = 0;
};

struct MgState {
  std::map<std::string, MgGrid*> grids;  // sid_to_iid_ + grid_list_: createKdtree walks the ids in map order
  MgGridDev* d_table = nullptr;          // descriptors of the grids the current "kd-tree" covers
  int table_alloc = 0, n_tree = 0;
  size_t tree_leaves = 0;
  std::vector<std::string> tree_ids;
  bool dirty = false;
  float tree_resolution = 0.f;
  float4* d_in = nullptr; size_t in_alloc = 0; bool in_owned = true;  // upload scratch of add_target
  double* d_partials = nullptr; size_t partials_alloc = 0;
  double* h_out = nullptr;               // pinned, RED_N self-validating slots {bits, bits ^ key} (2 RED_N words)
};

struct MgEvalState {
  float T[12];
  double jd[24];  // 8 x 3 FP64 angle rows (multigrid_ndt_omp_impl.hpp:535-544)
  double hd[45];  // 15 x 3 (:548-571)
  double d1, d2;
  float r2;       // float(radius * radius), pcl::KdTreeFLANN::radiusSearch
};

// leaf.centroid_ (multi_voxel_grid_covariance_omp_impl.hpp:404-405, :358): float running sum in input order / float(n).
// The three float chains of a leaf are sequential by definition (input order).  Light leaves: one lane each.  Heavy
// leaves (a dense patch of ground holds thousands of points): the whole warp takes them one at a time -- 32 lanes
// load 32 points coalesced into shared memory, lanes 0..2 run the x / y / z chains over the tile (8 values in
// flight), so a heavy leaf costs one dependent FADD per point instead of one dependent global load.
constexpr int CEN_THREADS = 128;
constexpr int CEN_HEAVY = 64;
__global__ void __launch_bounds__(CEN_THREADS) mg_centroid_kernel(const float4* __restrict__ pts, const unsigned long long* __restrict__ vals,
                                                                  const uint32_t* __restrict__ seg_start, uint32_t V,
                                                                  uint32_t n_valid_pts, float4* __restrict__ out) {
  __shared__ float s_tile[CEN_THREADS / 32][3][32];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const uint32_t Vr = (V + 31u) & ~31u;
  for (uint32_t v = blockIdx.x * blockDim.x + threadIdx.x; v < Vr; v += gridDim.x * blockDim.x) {  // warp-uniform trip count
    const bool has = v < V;
    const uint32_t lo = has ? seg_start[v] : 0u;
    const uint32_t hi = has ? ((v + 1 < V) ? seg_start[v + 1] : n_valid_pts) : 0u;
    const int n = static_cast<int>(hi - lo);
    const bool heavy = n > CEN_HEAVY;
    float s0 = 0.f, s1 = 0.f, s2 = 0.f;
    if (has && !heavy) {
      const unsigned long long* q = vals + lo;   // the leaf's (key << 32 | point index) elements, input order (stable sort)
      int i = 0;
      for (; i + 4 <= n; i += 4) {
        float4 t[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) t[u] = __ldg(pts + static_cast<uint32_t>(__ldg(q + i + u)));
#pragma unroll
        for (int u = 0; u < 4; ++u) { s0 += t[u].x; s1 += t[u].y; s2 += t[u].z; }
      }
      for (; i < n; ++i) { const float4 t = __ldg(pts + static_cast<uint32_t>(__ldg(q + i))); s0 += t.x; s1 += t.y; s2 += t.z; }
    }
    unsigned int hm = __ballot_sync(0xffffffffu, heavy);
    while (hm) {
      const int o = __ffs(hm) - 1;
      hm &= hm - 1;
      const uint32_t wlo = __shfl_sync(0xffffffffu, lo, o);
      const int wn = __shfl_sync(0xffffffffu, n, o);
      const unsigned long long* q = vals + wlo;
      float acc = 0.f;
      float x = 0.f, y = 0.f, z = 0.f;
      if (lane < wn) { const float4 t = __ldg(pts + static_cast<uint32_t>(__ldg(q + lane))); x = t.x; y = t.y; z = t.z; }
      for (int base = 0; base < wn; base += 32) {
        s_tile[w][0][lane] = x; s_tile[w][1][lane] = y; s_tile[w][2][lane] = z;
        const int nxt = base + 32 + lane;
        if (nxt < wn) { const float4 t = __ldg(pts + static_cast<uint32_t>(__ldg(q + nxt))); x = t.x; y = t.y; z = t.z; }
        __syncwarp();
        const int cnt = min(32, wn - base);
        if (lane < 3) {
          int k = 0;
          for (; k + 8 <= cnt; k += 8) {
            float t8[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) t8[u] = s_tile[w][lane][k + u];
#pragma unroll
            for (int u = 0; u < 8; ++u) acc += t8[u];
          }
          for (; k < cnt; ++k) acc += s_tile[w][lane][k];
        }
        __syncwarp();
      }
      const float a0 = __shfl_sync(0xffffffffu, acc, 0), a1 = __shfl_sync(0xffffffffu, acc, 1), a2 = __shfl_sync(0xffffffffu, acc, 2);
      if (lane == o) { s0 = a0; s1 = a1; s2 = a2; }
    }
    if (has) {
      const float fn = static_cast<float>(n);
      out[v] = make_float4(s0 / fn, s1 / fn, s2 / fn, 0.f);
    }
  }
}

constexpr int MG_THREADS = 128;
#ifndef MG_CTAS_PER_SM
#define MG_CTAS_PER_SM 2
#endif
constexpr int MG_WARPS = MG_THREADS / 32;

__device__ __forceinline__ double dot3s(double a0, double a1, double a2, double b0, double b1, double b2) {
  double s = a0 * b0; s = s + a1 * b1; s = s + a2 * b2; return s;
}

// updateDerivatives, FP64 (multigrid_ndt_omp_impl.hpp:640-697) with every dot product k ascending, left to right (the order the
// CPU restatement used by the tests pins).  The point gradient is [I | J] with J = columns 3..5 (jc[c][r]); the products
// with the identity block are not executed: C*1 + C*0 + C*0 == C and x.(C e_c) is the expression of xC[c] itself, so
// the skipped work is bit-neutral for finite inverse covariances (and no other kind passes the voxel build).
// hv: vectors a..f of eq 6.21 (3 doubles each).  acc layout = RED_*.
template <bool HESS>
__device__ __forceinline__ double mg_pair(const double (&jc)[3][3], const double (&hv)[6][3], const double* x, const double* C, double d1, double d2,
                                          double* __restrict__ acc) {
  double xC[3];
#pragma unroll
  for (int c = 0; c < 3; ++c) xC[c] = dot3s(x[0], x[1], x[2], C[c], C[3 + c], C[6 + c]);
  const double q = dot3s(x[0], x[1], x[2], xC[0], xC[1], xC[2]);
  double e = exp(-d2 * q * 0.5);
  const double score_inc = -d1 * e;
  e = d2 * e;
  if (e > 1 || e < 0 || e != e) return 0.0;
  e *= d1;
  double CJ[3][6], xCJ[6];
#pragma unroll
  for (int r = 0; r < 3; ++r) {
#pragma unroll
    for (int c = 0; c < 3; ++c) CJ[r][c] = C[3 * r + c];
#pragma unroll
    for (int c = 0; c < 3; ++c) CJ[r][3 + c] = dot3s(C[3 * r], C[3 * r + 1], C[3 * r + 2], jc[c][0], jc[c][1], jc[c][2]);
  }
#pragma unroll
  for (int c = 0; c < 3; ++c) xCJ[c] = xC[c];
#pragma unroll
  for (int c = 3; c < 6; ++c) xCJ[c] = dot3s(x[0], x[1], x[2], CJ[0][c], CJ[1][c], CJ[2][c]);
#pragma unroll
  for (int c = 0; c < 6; ++c) acc[RED_G + c] += e * xCJ[c];
  if (HESS) {
    const int hsel[3][3] = {{0, 1, 2}, {1, 3, 4}, {2, 4, 5}};
#pragma unroll
    for (int i = 0; i < 6; ++i) {
#pragma unroll
      for (int j = 0; j < 6; ++j) {
        // (Jp^T C Jp)(j, i): column j of Jp dotted with column i of C Jp
        const double jcj = (j < 3) ? CJ[j][i] : dot3s(jc[j - 3][0], jc[j - 3][1], jc[j - 3][2], CJ[0][i], CJ[1][i], CJ[2][i]);
        double xch = 0.0;  // x C (zero block of point_hessian) = 0
        if (i >= 3 && j >= 3) {
          const int s = hsel[i - 3][j - 3];
          xch = dot3s(xC[0], xC[1], xC[2], hv[s][0], hv[s][1], hv[s][2]);
        }
        acc[RED_H + 6 * i + j] += e * (-d2 * xCJ[i] * xCJ[j] + xch + jcj);
      }
    }
  }
  return score_inc;
}

constexpr int MG_QCAP = 512;  // pair queue of one warp: < 32 left over + at most 32 x 9 appended per probe batch

struct __align__(16) MgSlot {  // one source point of the chunk a warp is working on: 208 B
  float xt[3];
  int pad;
  double jc[8];                      // the eight non-trivial entries of the point gradient (eq 6.18/6.19)
  double hv[15];                     // the non-trivial entries of a..f (eq 6.21)
  unsigned long long nearest_bits;   // max pair score of the point (scores are >= 0: the bit pattern orders like the value)
};

struct MgWarpScratch {
  MgSlot pt[32];
  int qrow[MG_QCAP];
  unsigned short qmeta[MG_QCAP];     // slot (5 bits) | grid << 5
};

template <bool HESS>
__device__ __forceinline__ void mg_eval_pair(MgWarpScratch& ws, int e, const MgGridDev* __restrict__ grids, const MgEvalState& st, double* __restrict__ acc) {
  const int meta = ws.qmeta[e & (MG_QCAP - 1)];
  const size_t row = static_cast<size_t>(ws.qrow[e & (MG_QCAP - 1)]);
  MgSlot& sl = ws.pt[meta & 31];
  const MgGridDev& G = grids[meta >> 5];
  double jc[3][3], hv[6][3];
  jc[0][0] = 0.0; jc[0][1] = sl.jc[0]; jc[0][2] = sl.jc[1];
  jc[1][0] = sl.jc[2]; jc[1][1] = sl.jc[3]; jc[1][2] = sl.jc[4];
  jc[2][0] = sl.jc[5]; jc[2][1] = sl.jc[6]; jc[2][2] = sl.jc[7];
  if (HESS) {
    hv[0][0] = 0.0; hv[0][1] = sl.hv[0]; hv[0][2] = sl.hv[1];
    hv[1][0] = 0.0; hv[1][1] = sl.hv[2]; hv[1][2] = sl.hv[3];
    hv[2][0] = 0.0; hv[2][1] = sl.hv[4]; hv[2][2] = sl.hv[5];
#pragma unroll
    for (int k = 0; k < 3; ++k) { hv[3][k] = sl.hv[6 + k]; hv[4][k] = sl.hv[9 + k]; hv[5][k] = sl.hv[12 + k]; }
  }
  double xr[3], C[9];
#pragma unroll
  for (int a = 0; a < 3; ++a) xr[a] = static_cast<double>(sl.xt[a]) - __ldg(G.c_mean + 3 * row + a);
#pragma unroll
  for (int i = 0; i < 9; ++i) C[i] = __ldg(G.c_icov + 9 * row + i);
  const double s = mg_pair<HESS>(jc, hv, xr, C, st.d1, st.d2, acc);
  acc[RED_SCORE] += s;
  if (s > 0.0) atomicMax(&sl.nearest_bits, static_cast<unsigned long long>(__double_as_longlong(s)));
}

// Every warp works on chunks of 32 consecutive points and alternates two phases through a pair queue of its own:
//   phase A (one lane per POINT): for every target grid whose bounding box some lane touches, three batches of nine
//            hash probes (one z layer of the 3x3x3 block; the nine loads are issued together), the centroid distance
//            test of the radius search, and the surviving (point, grid, leaf) pairs appended to the queue by a warp scan;
//   phase B (one lane per PAIR): the FP64 pair maths on 32 real pairs at a time -- no lane idles on a probe that missed
//            and no lane waits for a neighbour with a longer neighbourhood.
// The queue is drained at the end of a chunk (its slots are reused by the next one).  The point -> warp -> queue
// position -> lane mapping is static, so every accumulator sees its terms in a fixed order (run-to-run bit stable).
template <bool HESS>
__global__ void __launch_bounds__(MG_THREADS, MG_CTAS_PER_SM) mg_eval_kernel(const float4* __restrict__ src, int n, const MgGridDev* __restrict__ grids, int n_grids,
                                                             const __grid_constant__ MgEvalState st, double* __restrict__ block_out) {
  __shared__ MgWarpScratch s_warp[MG_WARPS];
  __shared__ double s_red[MG_WARPS][RED_N];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  MgWarpScratch& ws = s_warp[w];
  double acc[RED_N];
#pragma unroll
  for (int i = 0; i < RED_N; ++i) acc[i] = 0.0;
  const int n_chunks = (n + 31) / 32;
  const int n_warps = gridDim.x * MG_WARPS;
  int head = 0, tail = 0;
  for (int chunk = blockIdx.x * MG_WARPS + w; chunk < n_chunks; chunk += n_warps) {
    const int idx = chunk * 32 + lane;
    const bool have = idx < n;
    float xt[3] = {0.f, 0.f, 0.f};
    if (have) {
      const float4 p = __ldg(src + idx);
#pragma unroll
      for (int r = 0; r < 3; ++r) xt[r] = st.T[4 * r] * p.x + (st.T[4 * r + 1] * p.y + (st.T[4 * r + 2] * p.z + st.T[4 * r + 3]));
      // computePointDerivatives, FP64 (multigrid_ndt_omp_impl.hpp:590-636)
      const double x0 = static_cast<double>(p.x), x1 = static_cast<double>(p.y), x2 = static_cast<double>(p.z);
      MgSlot& sl = ws.pt[lane];
      sl.xt[0] = xt[0]; sl.xt[1] = xt[1]; sl.xt[2] = xt[2];
#pragma unroll
      for (int r = 0; r < 8; ++r) sl.jc[r] = dot3s(st.jd[3 * r], st.jd[3 * r + 1], st.jd[3 * r + 2], x0, x1, x2);
      if (HESS) {
#pragma unroll
        for (int r = 0; r < 15; ++r) sl.hv[r] = dot3s(st.hd[3 * r], st.hd[3 * r + 1], st.hd[3 * r + 2], x0, x1, x2);
      }
    }
    ws.pt[lane].nearest_bits = 0ull;
    int nn = 0;
    __syncwarp();
    for (int gi = 0; gi < n_grids; ++gi) {
      const MgGridDev& G = grids[gi];
      const float leaf = G.meta.leaf;
      const int i0 = static_cast<int>(floorf(__fdiv_rn(xt[0], leaf)));
      const int i1 = static_cast<int>(floorf(__fdiv_rn(xt[1], leaf)));
      const int i2 = static_cast<int>(floorf(__fdiv_rn(xt[2], leaf)));
      const bool touches = have && !(i0 + 1 < G.meta.min_b[0] || i0 - 1 > G.meta.max_b[0] || i1 + 1 < G.meta.min_b[1] || i1 - 1 > G.meta.max_b[1] ||
                                     i2 + 1 < G.meta.min_b[2] || i2 - 1 > G.meta.max_b[2]);
      if (!__any_sync(0xffffffffu, touches)) continue;
      const uint32_t mask = G.meta.hash_mask;
      for (int oz = -1; oz <= 1; ++oz) {
        const int c2 = i2 + oz;
        const bool layer = touches && c2 >= G.meta.min_b[2] && c2 <= G.meta.max_b[2];
        if (!__any_sync(0xffffffffu, layer)) continue;
        int rows[9];
        int found = 0;
        if (layer) {
          uint32_t key[9];
          uint2 slot[9];
#pragma unroll
          for (int q = 0; q < 9; ++q) {
            const int c0 = i0 + (q % 3) - 1, c1 = i1 + (q / 3) - 1;
            const bool in = c0 >= G.meta.min_b[0] && c0 <= G.meta.max_b[0] && c1 >= G.meta.min_b[1] && c1 <= G.meta.max_b[1];
            key[q] = in ? static_cast<uint32_t>((c0 - G.meta.min_b[0]) * G.meta.divb_mul[0] + (c1 - G.meta.min_b[1]) * G.meta.divb_mul[1] +
                                                (c2 - G.meta.min_b[2]) * G.meta.divb_mul[2])
                        : HASH_EMPTY;
          }
#pragma unroll
          for (int q = 0; q < 9; ++q)
            slot[q] = (key[q] != HASH_EMPTY) ? __ldg(reinterpret_cast<const uint2*>(G.table) + (key[q] & mask)) : make_uint2(HASH_EMPTY, 0u);
#pragma unroll
          for (int q = 0; q < 9; ++q) {
            uint2 v = slot[q];
            uint32_t b = key[q] & mask;
            while (v.x != key[q] && v.x != HASH_EMPTY) {  // displaced key (load <= 0.5): the usual linear probe
              b = (b + 1) & mask;
              v = __ldg(reinterpret_cast<const uint2*>(G.table) + b);
            }
            rows[q] = (v.x == HASH_EMPTY) ? -1 : static_cast<int>(v.y);
            if (rows[q] < -1) rows[q] = -2 - rows[q];  // flagged leaf (nr_points_ = -1): computeLeafParams keeps it in the node and the kd-tree (multi_voxel_grid_covariance_omp_impl.hpp:369-389)
          }
          float4 cen[9];
#pragma unroll
          for (int q = 0; q < 9; ++q) cen[q] = (rows[q] >= 0) ? __ldg(G.centroid + rows[q]) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
          for (int q = 0; q < 9; ++q) {
            if (rows[q] < 0) continue;
            // FLANN L2_Simple<float> against the leaf centroid, strict '<' (RadiusResultSet::addPoint)
            const float dx = xt[0] - cen[q].x, dy = xt[1] - cen[q].y, dz = xt[2] - cen[q].z;
            float d = dx * dx; d = d + dy * dy; d = d + dz * dz;
            if (d < st.r2) ++found; else rows[q] = -1;
          }
        } else {
#pragma unroll
          for (int q = 0; q < 9; ++q) rows[q] = -1;
        }
        // warp scan of the hit counts -> queue positions
        int incl = found;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const int t = __shfl_up_sync(0xffffffffu, incl, o);
          if (lane >= o) incl += t;
        }
        const int total = __shfl_sync(0xffffffffu, incl, 31);
        if (found > 0) {
          int pos = tail + incl - found;
#pragma unroll
          for (int q = 0; q < 9; ++q) {
            if (rows[q] >= 0) {
              ws.qrow[pos & (MG_QCAP - 1)] = rows[q];
              ws.qmeta[pos & (MG_QCAP - 1)] = static_cast<unsigned short>(lane | (gi << 5));
              ++pos;
            }
          }
        }
        nn += found;
        tail += total;
        __syncwarp();
        while (tail - head >= 32) {  // phase B on full rounds
          mg_eval_pair<HESS>(ws, head + lane, grids, st, acc);
          head += 32;
        }
        __syncwarp();
      }
    }
    if (head < tail) {  // drain: the slots are reused by the next chunk
      if (head + lane < tail) mg_eval_pair<HESS>(ws, head + lane, grids, st, acc);
      head = tail;
    }
    __syncwarp();
    if (nn > 0) {
      acc[RED_NEAREST] += __longlong_as_double(static_cast<long long>(ws.pt[lane].nearest_bits));
      acc[RED_FOUND] += 1.0;
      acc[RED_COUNT] += static_cast<double>(nn);
    }
    __syncwarp();
  }
#pragma unroll
  for (int i = 0; i < RED_N; ++i) {
    double v = acc[i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) s_red[w][i] = v;
  }
  __syncthreads();
  if (threadIdx.x < RED_N) {
    double v = s_red[0][threadIdx.x];
#pragma unroll
    for (int k = 1; k < MG_WARPS; ++k) v += s_red[k][threadIdx.x];
    block_out[RED_N * static_cast<size_t>(blockIdx.x) + threadIdx.x] = v;
  }
}

// Block partials -> RED_N sums: one warp per component, lane l adds blocks l, l+32, ... in order, then a fixed
// shuffle tree -> deterministic.  Published to pinned host memory as self-validating slots {bits, bits ^ key} (publish_slot:
// one coalesced run of 16-byte stores per CTA), which the host polls instead of synchronising the stream.
constexpr int RED_WARPS = RED_N / 2;  // two CTAs of 23 warps
__global__ void __launch_bounds__(32 * RED_WARPS) mg_reduce_kernel(const double* __restrict__ block_out, int blocks, volatile unsigned long long* out_slots,
                                                                   unsigned long long key) {
  __shared__ double s_v[RED_WARPS];
  const int c = blockIdx.x * RED_WARPS + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  double v = 0.0;
  for (int b = lane; b < blocks; b += 32) v += block_out[RED_N * static_cast<size_t>(b) + c];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  if (lane == 0) s_v[threadIdx.x >> 5] = v;
  __syncthreads();
  if (threadIdx.x < RED_WARPS)
    publish_slot(out_slots, blockIdx.x * RED_WARPS + threadIdx.x, static_cast<unsigned long long>(__double_as_longlong(s_v[threadIdx.x])), key);
}

static void mg_angle_vectors(const double* p, double* jd, double* hd) {  // computeAngleDerivatives :498-586 (FP64 rows, snap 10e-5)
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

static MgState* mg_state(ndtb_handle_s* h) {
  if (!h->mg) h->mg = new MgState();
  return h->mg;
}

void mg_free(ndtb_handle_s* h) {
  MgState* s = h->mg;
  if (!s) return;
  for (auto& kv : s->grids) { free_device_grid(kv.second->grid); cudaFree(kv.second->d_centroid); delete kv.second; }
  cudaFree(s->d_table);
  if (s->in_owned) cudaFree(s->d_in);
  cudaFree(s->d_partials);
  cudaFreeHost(s->h_out);
  delete s;
  h->mg = nullptr;
}

bool mg_ready(ndtb_handle_s* h) { return h->mg && h->mg->n_tree > 0 && !h->mg->dirty; }

// One multigrid evaluation (computeDerivatives :338-494); compute_hessian = false leaves H zero.
int mg_eval(ndtb_handle_s* h, const double* p, const float* T16, bool compute_hessian, double* score, double* g, double* H, double* nvtl) {
  MgState* s = h->mg;
  if (!s || s->n_tree == 0) return set_error(h, NDTB_ERR_NOT_READY, "multigrid: no target in the kd-tree (ndtb_mg_add_target + ndtb_mg_create_kdtree)");
  if (s->dirty) return set_error(h, NDTB_ERR_NOT_READY, "multigrid: targets changed since ndtb_mg_create_kdtree");
  const int n = static_cast<int>(h->n_source);
  if (n == 0) return set_error(h, NDTB_ERR_NOT_READY, "source must be set");
  static const int blocks_cap = std::getenv("NDTB_MG_BLOCKS") ? std::atoi(std::getenv("NDTB_MG_BLOCKS")) : 148 * MG_CTAS_PER_SM;
  const int blocks = std::max(1, std::min((n + MG_THREADS - 1) / MG_THREADS, std::min(blocks_cap, 148 * 8)));
  if (static_cast<size_t>(blocks) * RED_N > s->partials_alloc) {
    cudaFree(s->d_partials);
    CKH(cudaMalloc(&s->d_partials, sizeof(double) * RED_N * 148 * 8));
    s->partials_alloc = static_cast<size_t>(RED_N) * 148 * 8;
  }
  if (!s->h_out) {
    CKH(cudaHostAlloc(&s->h_out, sizeof(double) * 2 * RED_N, cudaHostAllocDefault));
    std::memset(s->h_out, 0, sizeof(double) * 2 * RED_N);
  }
  const unsigned long long key = (0x9E3779B97F4A7C15ull * static_cast<unsigned long long>(++h->flag_seq)) | 1ull;
  MgEvalState st;
  std::memcpy(st.T, T16, 12 * sizeof(float));
  mg_angle_vectors(p, st.jd, st.hd);
  double d3;
  gauss_pclomp(h->params.outlier_ratio, s->tree_resolution, &st.d1, &st.d2, &d3);
  const double radius = static_cast<double>(s->tree_resolution);
  st.r2 = static_cast<float>(radius * radius);
  if (compute_hessian) mg_eval_kernel<true><<<blocks, MG_THREADS, 0, h->stream>>>(h->d_source, n, s->d_table, s->n_tree, st, s->d_partials);
  else mg_eval_kernel<false><<<blocks, MG_THREADS, 0, h->stream>>>(h->d_source, n, s->d_table, s->n_tree, st, s->d_partials);
  volatile unsigned long long* slots = reinterpret_cast<volatile unsigned long long*>(s->h_out);
  mg_reduce_kernel<<<2, 32 * RED_WARPS, 0, h->stream>>>(s->d_partials, blocks, slots, key);
  h->launches += 2;
  h->eval_launches += 1;
  CKH(cudaGetLastError());
  // poll the slots (as eval_pclomp does) instead of synchronising the stream; a stream query every 4096 probes catches a failed launch
  double r[RED_N];
  unsigned int spins = 0;
  for (int c = RED_N - 1; c >= 0; --c) {
    for (;;) {
      const unsigned long long w0 = slots[2 * c], w1 = slots[2 * c + 1];
      if ((w0 ^ w1) == key) { std::memcpy(&r[c], &w0, 8); break; }
#if defined(__x86_64__) || defined(__i386__)
      __builtin_ia32_pause();
#endif
      if ((++spins & 4095u) == 0) {
        const cudaError_t q = cudaStreamQuery(h->stream);
        if (q != cudaSuccess && q != cudaErrorNotReady) { h->last_error = std::string("multigrid eval kernel: ") + cudaGetErrorString(q); return NDTB_ERR_CUDA; }
        if (q == cudaSuccess) {
          const unsigned long long x0 = slots[2 * c], x1 = slots[2 * c + 1];
          if ((x0 ^ x1) == key) continue;
          h->last_error = "multigrid eval kernel finished without publishing its result";
          return NDTB_ERR_CUDA;
        }
      }
    }
  }
  __sync_synchronize();
  finish_eval(h, n, r, p, score, g, H, nvtl);  // regulariser (:446-480) and NVTL (:482-488): same host epilogue as pclomp
  return NDTB_OK;
}

}  // namespace ndtb

extern "C" {

int ndtb_mg_add_target(ndtb_handle h, const char* grid_id, const float* xyz, size_t n, size_t stride_bytes) {
  if (!h || !grid_id) return NDTB_ERR_INVALID;
  cudaSetDevice(h->device);
  if (!xyz || n == 0) return set_error(h, NDTB_ERR_EMPTY, "empty target cloud");
  MgState* s = mg_state(h);
  int rc = upload_cloud(h, xyz, n, stride_bytes, &s->d_in, &s->in_alloc, &s->in_owned);
  if (rc != NDTB_OK) return rc;
  // an existing id is replaced: the reference orphans the old node (sid_to_iid_[id] = new index) and createKdtree drops it
  MgGrid* g = nullptr;
  auto it = s->grids.find(grid_id);
  if (it != s->grids.end()) g = it->second; else { g = new MgGrid(); s->grids[grid_id] = g; }
  s->dirty = true;
  const ndtb_params& P = h->params;
  const uint64_t launches_before = g->grid.launches;
  rc = build_voxel_map(g->grid, s->d_in, n, P.resolution, P.min_points_per_voxel, P.min_covar_eigvalue_mult, FLAVOUR_PCLOMP, P.is_dense, h->stream);
  if (rc != NDTB_OK) {  // applyFilter :307-313 warns and leaves the node empty; here the id is not registered at all
    free_device_grid(g->grid); cudaFree(g->d_centroid); delete g; s->grids.erase(grid_id);
    return set_error(h, rc, "multigrid: voxel map build failed");
  }
  const uint32_t V = g->grid.V;
  if (V > g->centroid_alloc) {
    cudaFree(g->d_centroid);
    CKH(cudaMalloc(&g->d_centroid, sizeof(float4) * V));
    g->centroid_alloc = V;
  }
  const uint32_t n_valid_pts = static_cast<uint32_t>(n) - g->grid.n_skipped;
  mg_centroid_kernel<<<std::max(1, std::min<int>((V + CEN_THREADS - 1) / CEN_THREADS, 148 * 8)), CEN_THREADS, 0, h->stream>>>(s->d_in, g->grid.pairs[g->grid.final_buf], g->grid.seg_start, V, n_valid_pts, g->d_centroid);
  g->grid.launches += 1;
  h->launches += g->grid.launches - launches_before;  // ndtb_get_launch_count: the per-tile builds are this handle's kernels too
  CKH(cudaGetLastError());
  CKH(cudaStreamSynchronize(h->stream));
  return NDTB_OK;
}

int ndtb_mg_remove_target(ndtb_handle h, const char* grid_id) {
  if (!h || !grid_id) return NDTB_ERR_INVALID;
  cudaSetDevice(h->device);
  MgState* s = mg_state(h);
  auto it = s->grids.find(grid_id);
  if (it == s->grids.end()) return NDTB_OK;  // removeCloud :183-185: unknown ids are ignored
  CKH(cudaStreamSynchronize(h->stream));
  free_device_grid(it->second->grid); cudaFree(it->second->d_centroid); delete it->second;
  s->grids.erase(it);
  s->dirty = true;
  return NDTB_OK;
}

int ndtb_mg_create_kdtree(ndtb_handle h) {
  if (!h) return NDTB_ERR_INVALID;
  cudaSetDevice(h->device);
  MgState* s = mg_state(h);
  std::vector<MgGridDev> tab;
  s->tree_ids.clear();
  s->tree_leaves = 0;
  float res = 0.f;
  for (auto& kv : s->grids) {
    MgGrid* g = kv.second;
    if (!g->grid.built) continue;
    if (res != 0.f && g->grid.meta.leaf != res) return set_error(h, NDTB_ERR_INVALID, "multigrid: targets were added with different resolutions");
    res = g->grid.meta.leaf;
    MgGridDev d;
    d.meta = g->grid.meta; d.table = g->grid.table; d.c_mean = g->grid.c_mean; d.c_icov = g->grid.c_icov; d.centroid = g->d_centroid;
    tab.push_back(d);
    s->tree_ids.push_back(kv.first);
    s->tree_leaves += g->grid.n_valid + g->grid.n_flagged;
  }
  if (tab.size() > 2047) return set_error(h, NDTB_ERR_INVALID, "multigrid: at most 2047 targets in one kd-tree");
  if (static_cast<int>(tab.size()) > s->table_alloc) {
    cudaFree(s->d_table);
    const int cap = std::max<int>(16, 2 * static_cast<int>(tab.size()));
    CKH(cudaMalloc(&s->d_table, sizeof(MgGridDev) * cap));
    s->table_alloc = cap;
  }
  if (!tab.empty()) CKH(cudaMemcpyAsync(s->d_table, tab.data(), sizeof(MgGridDev) * tab.size(), cudaMemcpyHostToDevice, h->stream));
  CKH(cudaStreamSynchronize(h->stream));
  s->n_tree = static_cast<int>(tab.size());
  s->tree_resolution = res;
  s->dirty = false;
  return NDTB_OK;
}

int ndtb_mg_set_input_target(ndtb_handle h, const float* xyz, size_t n, size_t stride_bytes) {  // multigrid_ndt_omp.h:143-148
  const int rc = ndtb_mg_add_target(h, "default", xyz, n, stride_bytes);
  if (rc != NDTB_OK) return rc;
  return ndtb_mg_create_kdtree(h);
}

int ndtb_mg_get_info(ndtb_handle h, int* n_grids_total, int* n_grids_in_tree, size_t* leaves_in_tree, int* tree_is_current) {
  if (!h) return NDTB_ERR_INVALID;
  MgState* s = mg_state(h);
  if (n_grids_total) *n_grids_total = static_cast<int>(s->grids.size());
  if (n_grids_in_tree) *n_grids_in_tree = s->n_tree;
  if (leaves_in_tree) *leaves_in_tree = s->tree_leaves;
  if (tree_is_current) *tree_is_current = s->dirty ? 0 : 1;
  return NDTB_OK;
}

// getCurrentMapIDs (multi_voxel_grid_covariance_omp_impl.hpp:277-288): ids joined by '\n' into buf (NUL-terminated);
// *needed = bytes required including the terminator.
int ndtb_mg_get_grid_ids(ndtb_handle h, char* buf, size_t capacity, size_t* needed) {
  if (!h) return NDTB_ERR_INVALID;
  MgState* s = mg_state(h);
  std::string all;
  for (auto& kv : s->grids) { if (!all.empty()) all += '\n'; all += kv.first; }
  if (needed) *needed = all.size() + 1;
  if (buf && capacity > 0) { const size_t c = std::min(capacity - 1, all.size()); std::memcpy(buf, all.data(), c); buf[c] = 0; }
  return NDTB_OK;
}

// Leaves of the current kd-tree, in tree order (grid id order, ascending linear index inside a grid; searchable
// leaves only): getVoxelPCD-style export for parity tests.  Any output pointer may be NULL.
int ndtb_mg_get_leaves(ndtb_handle h, size_t capacity, int32_t* nr_points, double* mean, double* icov, float* centroid_xyz, size_t* count) {
  if (!h) return NDTB_ERR_INVALID;
  cudaSetDevice(h->device);
  MgState* s = mg_state(h);
  if (s->dirty) return set_error(h, NDTB_ERR_NOT_READY, "multigrid: targets changed since ndtb_mg_create_kdtree");
  size_t k = 0;
  for (const std::string& id : s->tree_ids) {
    MgGrid* g = s->grids[id];
    const size_t V = g->grid.V;
    std::vector<int32_t> n(V); std::vector<double> m(3 * V), ic(9 * V); std::vector<float4> cen(V);
    CKH(cudaMemcpy(n.data(), g->grid.c_n, sizeof(int32_t) * V, cudaMemcpyDeviceToHost));
    CKH(cudaMemcpy(m.data(), g->grid.c_mean, sizeof(double) * 3 * V, cudaMemcpyDeviceToHost));
    CKH(cudaMemcpy(ic.data(), g->grid.c_icov, sizeof(double) * 9 * V, cudaMemcpyDeviceToHost));
    CKH(cudaMemcpy(cen.data(), g->d_centroid, sizeof(float4) * V, cudaMemcpyDeviceToHost));
    for (size_t v = 0; v < V; ++v) {
      if (n[v] < h->params.min_points_per_voxel && n[v] != -1) continue;  // dropped by applyFilter :369-372; leaves flagged -1 by computeLeafParams stay in the node
      if (k < capacity) {
        if (nr_points) nr_points[k] = n[v];
        if (mean) std::memcpy(mean + 3 * k, &m[3 * v], 24);
        if (icov) std::memcpy(icov + 9 * k, &ic[9 * v], 72);
        if (centroid_xyz) { centroid_xyz[3 * k] = cen[v].x; centroid_xyz[3 * k + 1] = cen[v].y; centroid_xyz[3 * k + 2] = cen[v].z; }
      }
      ++k;
    }
  }
  if (count) *count = k;
  return NDTB_OK;
}

// calculateTransformationProbability / calculateNearestVoxelTransformationLikelihood (multigrid_ndt_omp_impl.hpp:1054-1145)
// for the source cloud transformed by T16: sum of the pair scores / N and mean of the per-point best pair score over the
// points with neighbours.  Same pair score as computeDerivatives (no regulariser, no gradient work kept).
int ndtb_mg_calculate_scores(ndtb_handle h, const float T16[16], double* transform_probability, double* nvtl) {
  if (!h || !T16) return NDTB_ERR_INVALID;
  cudaSetDevice(h->device);
  const bool reg = h->has_reg_pose;
  h->has_reg_pose = false;
  const double p0[6] = {0, 0, 0, 0, 0, 0};
  double score = 0, g[6], H[36], nv = 0;
  const int rc = mg_eval(h, p0, T16, false, &score, g, H, &nv);
  h->has_reg_pose = reg;
  if (rc != NDTB_OK) return rc;
  if (transform_probability) *transform_probability = h->n_source ? score / static_cast<double>(h->n_source) : 0.0;
  if (nvtl) *nvtl = nv;
  return NDTB_OK;
}

int ndtb_mg_compute_derivatives(ndtb_handle h, const double p[6], const float* T16, int compute_hessian, double* score, double gradient[6],
                                double hessian[36], double* nvtl) {
  if (!h || !p || !score || !gradient || !hessian) return NDTB_ERR_INVALID;
  cudaSetDevice(h->device);
  float T[16];
  if (T16) std::memcpy(T, T16, sizeof(T)); else pose_to_T(p, T);
  return mg_eval(h, p, T, compute_hessian != 0, score, gradient, hessian, nvtl);
}

}  // extern "C"
