// k6: per-voxel statistics (SURVEY 8a rows a2 pass 1, a3, a11), see voxel_stats_kernel.
// One lane per voxel.  The lane walks its segment of the (stably) sorted point list, i.e. its
// points IN INPUT ORDER, and accumulates Sum x and Sum x x^T in FP64 exactly like the reference's
// serial loop (voxel_grid_covariance_omp_impl.hpp:235-239 / voxel_grid_covariance_impl.hpp:237-239),
// so means and covariances are reproduced bit for bit, not just to rounding.
// Then: pclomp flavour  voxel_grid_covariance_omp_impl.hpp:286-360
//       svn flavour     voxel_grid_covariance_impl.hpp:270-343
// and the voxel is inserted into the open-addressing hash (atomicCAS linear probing).
#include "build_api.cuh"
#include <algorithm>

namespace ndtb {

__device__ __forceinline__ void d_mat3_mul(const double* a, const double* b, double* out) {
  double r[9];
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      double s = a[3 * i + 0] * b[0 + j];
      s = s + a[3 * i + 1] * b[3 + j];
      s = s + a[3 * i + 2] * b[6 + j];
      r[3 * i + j] = s;
    }
#pragma unroll
  for (int i = 0; i < 9; ++i) out[i] = r[i];
}

// cofactor / determinant inverse (Eigen's 3x3 path)
__device__ __forceinline__ void d_mat3_inverse(const double* m, double* r) {
  const double c00 = m[4] * m[8] - m[5] * m[7];
  const double c10 = m[5] * m[6] - m[3] * m[8];
  const double c20 = m[3] * m[7] - m[4] * m[6];
  const double det = (m[0] * c00 + m[1] * c10) + m[2] * c20;
  const double invdet = 1.0 / det;
  double t[9];
  t[0] = c00 * invdet;
  t[3] = c10 * invdet;
  t[6] = c20 * invdet;
  t[1] = (m[2] * m[7] - m[1] * m[8]) * invdet;
  t[4] = (m[0] * m[8] - m[2] * m[6]) * invdet;
  t[7] = (m[1] * m[6] - m[0] * m[7]) * invdet;
  t[2] = (m[1] * m[5] - m[2] * m[4]) * invdet;
  t[5] = (m[2] * m[3] - m[0] * m[5]) * invdet;
  t[8] = (m[0] * m[4] - m[1] * m[3]) * invdet;
#pragma unroll
  for (int i = 0; i < 9; ++i) r[i] = t[i];
}

// One Jacobi rotation on the (p,q) plane of a symmetric 3x3 held in a[9] (full storage), v = vectors.
template <int P, int Q>
__device__ __forceinline__ void jacobi_rotate(double* a, double* v, int sweep) {
  constexpr int R = 3 - P - Q;
  const double apq = a[3 * P + Q];
  if (apq == 0.0) return;
  const double g = 100.0 * fabs(apq);
  const double app = a[4 * P], aqq = a[4 * Q];
  if (sweep > 3 && fabs(app) + g == fabs(app) && fabs(aqq) + g == fabs(aqq)) {
    a[3 * P + Q] = 0.0;
    a[3 * Q + P] = 0.0;
    return;
  }
  const double h = aqq - app;
  double t;
  if (fabs(h) + g == fabs(h)) {
    t = apq / h;
  } else {
    const double theta = 0.5 * h / apq;
    t = 1.0 / (fabs(theta) + sqrt(1.0 + theta * theta));
    if (theta < 0.0) t = -t;
  }
  const double c = 1.0 / sqrt(1.0 + t * t);
  const double s = t * c;
  const double tau = s / (1.0 + c);
  const double hh = t * apq;
  a[4 * P] = app - hh;
  a[4 * Q] = aqq + hh;
  a[3 * P + Q] = 0.0;
  a[3 * Q + P] = 0.0;
  {
    const double arp = a[3 * R + P], arq = a[3 * R + Q];
    const double nrp = arp - s * (arq + arp * tau);
    const double nrq = arq + s * (arp - arq * tau);
    a[3 * R + P] = nrp; a[3 * P + R] = nrp;
    a[3 * R + Q] = nrq; a[3 * Q + R] = nrq;
  }
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const double vkp = v[3 * k + P], vkq = v[3 * k + Q];
    v[3 * k + P] = vkp - s * (vkq + vkp * tau);
    v[3 * k + Q] = vkq + s * (vkp - vkq * tau);
  }
}

// Cyclic Jacobi, lower triangle read, ascending eigenvalues, eigenvectors in columns (row-major).
__device__ void d_eig3_sym(const double* A, double* evals, double* evecs) {
  double a[9], v[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) a[3 * i + j] = (i >= j) ? A[3 * i + j] : A[3 * j + i];
  for (int sweep = 0; sweep < 60; ++sweep) {
    const double off = (fabs(a[1]) + fabs(a[2])) + fabs(a[5]);
    if (off == 0.0) break;
    jacobi_rotate<0, 1>(a, v, sweep);
    jacobi_rotate<0, 2>(a, v, sweep);
    jacobi_rotate<1, 2>(a, v, sweep);
  }
  const double d0 = a[0], d1 = a[4], d2 = a[8];
  // stable insertion sort of 3 (same comparisons as the oracle)
  int i0 = 0, i1 = 1, i2 = 2;
  double e0 = d0, e1 = d1, e2 = d2;
  if (e1 < e0) { double t = e0; e0 = e1; e1 = t; int ti = i0; i0 = i1; i1 = ti; }
  if (e2 < e1) {
    double t = e1; e1 = e2; e2 = t; int ti = i1; i1 = i2; i2 = ti;
    if (e1 < e0) { t = e0; e0 = e1; e1 = t; ti = i0; i0 = i1; i1 = ti; }
  }
  evals[0] = e0; evals[1] = e1; evals[2] = e2;
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    evecs[3 * r + 0] = (i0 == 0) ? v[3 * r] : (i0 == 1 ? v[3 * r + 1] : v[3 * r + 2]);
    evecs[3 * r + 1] = (i1 == 0) ? v[3 * r] : (i1 == 1 ? v[3 * r + 1] : v[3 * r + 2]);
    evecs[3 * r + 2] = (i2 == 0) ? v[3 * r] : (i2 == 1 ? v[3 * r + 1] : v[3 * r + 2]);
  }
}

__device__ __forceinline__ void hash_insert(HashSlot* table, uint32_t mask, uint32_t key, int row) {
  uint32_t b = key & mask;
  for (;;) {
    const uint32_t prev = atomicCAS(&table[b].key, HASH_EMPTY, key);
    if (prev == HASH_EMPTY) { table[b].row = row; return; }
    b = (b + 1) & mask;
  }
}

constexpr int ST_THREADS = 128;   // 4 warps per CTA
constexpr int ST_CHUNK = 32;      // leaves per chunk: one per lane
constexpr int ST_HEAVY = HEAVY_LEAF_POINTS;   // leaves with more points arrive summed (end of sort_segments_kernel)
constexpr int FIN_THREADS = 64;            // finalize: 2 warps per CTA, 8 CTAs per SM (128 registers, no spill)
constexpr int FIN_MIN_BLOCKS = 8;

// Everything per voxel after the sums.  `n`, sums in; writes the cold arrays, the hot record and the hash slot.
// Returns 1 for a leaf the searches can return, 2 for a pclomp leaf flagged -1 (radius searches only), 0 otherwise.
__device__ int finalize_voxel(uint32_t v, int n, double sx, double sy, double sz, double sxx, double sxy, double sxz, double syy, double syz, double szz,
                              const uint32_t* __restrict__ vox_key, const GridMeta& m, VoxRec* __restrict__ recs, HashSlot* __restrict__ table,
                              int32_t* __restrict__ c_n, double* __restrict__ c_mean, double* __restrict__ c_cov, double* __restrict__ c_icov,
                              double* __restrict__ c_evecs, double* __restrict__ c_evals);

// Per-leaf statistics, one thread per leaf, chunks of 32 consecutive leaves per warp (handed out dynamically: the leaf count is
// only known on the device).  A light leaf (<= ST_HEAVY points) is summed here: the lane walks ITS run of sorted_pts -- its
// points in input order -- widening to FP64 and accumulating Sum x, Sum x x^T exactly like the reference's serial loop
// (bit-exact means and covariances); the 32 runs of a chunk are adjacent, every 128-byte line is fetched once and then served
// from L1.  Heavy leaves arrive summed (heavy_leaf_sums at the end of sort_segments_kernel).  Then mean, covariance, Jacobi eigen-decomposition, inflation,
// inverse, hot record, hash insert (finalize_voxel).  The last block publishes BuildMeta (geometry, leaf counts, error) to the host.
__global__ void __launch_bounds__(FIN_THREADS, FIN_MIN_BLOCKS) voxel_stats_kernel(const float4* __restrict__ sorted_pts, BuildMeta* __restrict__ meta,
                                                                    const uint32_t* __restrict__ seg_start, const uint32_t* __restrict__ vox_key,
                                                                    uint32_t n_points, const double* __restrict__ vox_sums, VoxRec* __restrict__ recs,
                                                                    HashSlot* __restrict__ table, int32_t* __restrict__ c_n,
                                                                    double* __restrict__ c_mean, double* __restrict__ c_cov,
                                                                    double* __restrict__ c_icov, double* __restrict__ c_evecs,
                                                                    double* __restrict__ c_evals, volatile unsigned long long* host_slots,
                                                                    unsigned long long publish_key) {
  __shared__ unsigned int s_cnt[2];
  __shared__ bool s_last;
  pdl_enter();
  if (threadIdx.x < 2) s_cnt[threadIdx.x] = 0;
  __syncthreads();
  if (!meta->error) {
    const GridMeta m = meta->m;
    const uint32_t V = meta->V;
    const uint32_t n_valid_pts = n_points - meta->n_skipped;
    const uint32_t n_chunks = (V + ST_CHUNK - 1) / ST_CHUNK;
    const int lane = threadIdx.x & 31;
    for (;;) {
      uint32_t chunk = 0;
      if (lane == 0) chunk = atomicAdd(&meta->chunk_counter, 1u);
      chunk = __shfl_sync(0xffffffffu, chunk, 0);
      if (chunk >= n_chunks) break;
      const uint32_t v = chunk * ST_CHUNK + lane;
      int kind = 0;
      if (v < V) {
        const uint32_t lo = seg_start[v], hi = (v + 1 < V) ? seg_start[v + 1] : n_valid_pts;
        const int n = static_cast<int>(hi - lo);
        // sx sy sz sxx sxy sxz syy syz szz.  The reference's Leaf starts cov_ at identity and adds x x^T onto it
        // (voxel_grid_covariance_omp.h:113 + _impl.hpp:239; svn voxel_grid_covariance.h:109 + _impl.hpp:240), so the
        // diagonal chains start at 1.0, not 0.
        double sum[9] = {0, 0, 0, 1.0, 0, 0, 1.0, 0, 1.0};
        if (n > ST_HEAVY) {
          const double* q = vox_sums + 9 * static_cast<size_t>(v);
#pragma unroll
          for (int c = 0; c < 9; ++c) sum[c] = __ldcg(q + c);
        } else {
          const float4* src = sorted_pts + lo;
          int k = 0;
          for (; k + 4 <= n; k += 4) {
            float4 p[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) p[u] = __ldg(src + k + u);
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const double x = static_cast<double>(p[u].x), y = static_cast<double>(p[u].y), z = static_cast<double>(p[u].z);
              sum[0] += x; sum[1] += y; sum[2] += z;
              sum[3] += x * x; sum[4] += x * y; sum[5] += x * z; sum[6] += y * y; sum[7] += y * z; sum[8] += z * z;
            }
          }
          for (; k < n; ++k) {
            const float4 p = __ldg(src + k);
            const double x = static_cast<double>(p.x), y = static_cast<double>(p.y), z = static_cast<double>(p.z);
            sum[0] += x; sum[1] += y; sum[2] += z;
            sum[3] += x * x; sum[4] += x * y; sum[5] += x * z; sum[6] += y * y; sum[7] += y * z; sum[8] += z * z;
          }
        }
        kind = finalize_voxel(v, n, sum[0], sum[1], sum[2], sum[3], sum[4], sum[5], sum[6], sum[7], sum[8], vox_key, m, recs, table, c_n, c_mean,
                              c_cov, c_icov, c_evecs, c_evals);
      }
      const unsigned int mv = __ballot_sync(0xffffffffu, kind == 1), mf = __ballot_sync(0xffffffffu, kind == 2);
      if (lane == 0) {
        if (mv) atomicAdd(&s_cnt[0], static_cast<unsigned int>(__popc(mv)));
        if (mf) atomicAdd(&s_cnt[1], static_cast<unsigned int>(__popc(mf)));
      }
    }
  }
  // ---- last block done: publish BuildMeta to the host slots -------------------------------------------------------------
  __syncthreads();
  if (threadIdx.x == 0) {
    if (s_cnt[0]) atomicAdd(&meta->n_valid, s_cnt[0]);
    if (s_cnt[1]) atomicAdd(&meta->n_flagged, s_cnt[1]);
    __threadfence();
    unsigned int tk;
    asm volatile("atom.acq_rel.gpu.global.add.u32 %0, [%1], 1;" : "=r"(tk) : "l"(&meta->tickets[1]) : "memory");
    s_last = (tk == gridDim.x - 1);
  }
  __syncthreads();
  if (s_last) {
    if (threadIdx.x == 0) { meta->tickets[1] = 0; meta->chunk_counter = 0; __threadfence(); }
    __syncthreads();
    const unsigned long long* words = reinterpret_cast<const unsigned long long*>(meta);
    for (int i = threadIdx.x; i < BUILD_META_WORDS; i += blockDim.x) publish_slot(host_slots, i, __ldcg(words + i), publish_key);
  }
}

__device__ int finalize_voxel(uint32_t v, int n, double sx, double sy, double sz, double sxx, double sxy, double sxz, double syy, double syz, double szz,
                              const uint32_t* __restrict__ vox_key, const GridMeta& m, VoxRec* __restrict__ recs, HashSlot* __restrict__ table,
                              int32_t* __restrict__ c_n, double* __restrict__ c_mean, double* __restrict__ c_cov, double* __restrict__ c_icov,
                              double* __restrict__ c_evecs, double* __restrict__ c_evals) {
  const double dn = static_cast<double>(n);
  double mean[3] = {sx, sy, sz};
  double cov[9] = {sxx, sxy, sxz, sxy, syy, syz, sxz, syz, szz};
  double icov[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}, evecs[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}, evals[3] = {0, 0, 0};
  bool valid = false;
  if (m.flavour == FLAVOUR_PCLOMP) {
    const double pt_sum[3] = {sx, sy, sz};
#pragma unroll
    for (int a = 0; a < 3; ++a) mean[a] = mean[a] / dn;
    if (n >= m.min_pts) {
#pragma unroll
      for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) cov[3 * i + j] = (cov[3 * i + j] - 2.0 * (pt_sum[i] * mean[j])) / dn + mean[i] * mean[j];
      const double f = (dn - 1.0) / dn;
#pragma unroll
      for (int i = 0; i < 9; ++i) cov[i] = cov[i] * f;
      double ev[3];
      d_eig3_sym(cov, ev, evecs);
      if (ev[0] < 0 || ev[1] < 0 || ev[2] <= 0) {
        n = -1;
      } else {
        const double min_ev = m.eig_mult * ev[2];
        if (ev[0] < min_ev) {
          ev[0] = min_ev;
          if (ev[1] < min_ev) ev[1] = min_ev;
          double vd[9], vinv[9];
#pragma unroll
          for (int i = 0; i < 3; ++i)
#pragma unroll
            for (int j = 0; j < 3; ++j) vd[3 * i + j] = evecs[3 * i + j] * ev[j];
          d_mat3_inverse(evecs, vinv);
          d_mat3_mul(vd, vinv, cov);
        }
        evals[0] = ev[0]; evals[1] = ev[1]; evals[2] = ev[2];
        d_mat3_inverse(cov, icov);
        double mx = icov[0], mn = icov[0];
#pragma unroll
        for (int i = 1; i < 9; ++i) { mx = fmax(mx, icov[i]); mn = fmin(mn, icov[i]); }
        const double inf = __longlong_as_double(0x7ff0000000000000LL);
        if (mx == inf || mn == -inf) n = -1; else valid = true;
      }
    }
  } else {
    if (n >= m.min_pts) {
#pragma unroll
      for (int a = 0; a < 3; ++a) mean[a] = mean[a] / dn;
#pragma unroll
      for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) cov[3 * i + j] = (cov[3 * i + j] / dn) - (mean[i] * mean[j]);
      if (dn > 1) {
        const double f = dn / (dn - 1.0);
#pragma unroll
        for (int i = 0; i < 9; ++i) cov[i] = cov[i] * f;
      } else {
#pragma unroll
        for (int i = 0; i < 9; ++i) cov[i] = (i % 4 == 0) ? 1e-9 : 0.0;
      }
      double ev[3];
      d_eig3_sym(cov, ev, evecs);
      const double thr = 1e-12;
      if (!(ev[0] < 0 || ev[1] < 0 || ev[2] < thr)) {
        const double min_ok = fmax(thr, ev[2] * m.eig_mult);
        bool recompose = false;
        if (ev[0] < min_ok) { ev[0] = min_ok; recompose = true; }
        if (ev[1] < min_ok) { ev[1] = min_ok; recompose = true; }
        evals[0] = ev[0]; evals[1] = ev[1]; evals[2] = ev[2];
        if (recompose) {
          double vd[9], vt[9];
#pragma unroll
          for (int i = 0; i < 3; ++i)
#pragma unroll
            for (int j = 0; j < 3; ++j) { vd[3 * i + j] = evecs[3 * i + j] * ev[j]; vt[3 * i + j] = evecs[3 * j + i]; }
          d_mat3_mul(vd, vt, cov);
        }
        d_mat3_inverse(cov, icov);
        double mx = 0;
        bool fin = true;
#pragma unroll
        for (int i = 0; i < 9; ++i) { if (!isfinite(icov[i])) fin = false; mx = fmax(mx, fabs(icov[i])); }
        if (fin && !(mx > 1e12)) valid = true;
      }
    }
    if (!valid) n = -2;  // erased leaf (svn): filtered out of the export
  }
  c_n[v] = n;
#pragma unroll
  for (int a = 0; a < 3; ++a) { c_mean[3 * v + a] = mean[a]; c_evals[3 * v + a] = evals[a]; }
#pragma unroll
  for (int i = 0; i < 9; ++i) { c_cov[9 * v + i] = cov[i]; c_icov[9 * v + i] = icov[i]; c_evecs[9 * v + i] = evecs[i]; }
  // pclomp: a leaf flagged nr_points_ = -1 (negative eigenvalue / infinite inverse, voxel_grid_covariance_omp_impl.hpp:335-338,
  // 355-359) had already entered the centroid cloud (:298-323), so the reference's kd-tree still returns it -- with icov_ as the
  // leaf holds it (zero, or the infinite inverse).  It goes into the table under the row code -2 - v: the DIRECT searches skip it
  // (they test nr_points_ >= min_points, :390-394), the radius searches decode it.
  const bool flagged = (m.flavour == FLAVOUR_PCLOMP) && n == -1;
  if (valid || flagged) {
    VoxRec r;
    r.mean[0] = mean[0]; r.mean[1] = mean[1]; r.mean[2] = mean[2];
#pragma unroll
    for (int i = 0; i < 9; ++i) r.icov[i] = static_cast<float>(icov[i]);
    r.n = n;
    // 64-byte record written as four 16-byte stores
    const uint4* src = reinterpret_cast<const uint4*>(&r);
    uint4* dst = reinterpret_cast<uint4*>(recs + v);
    dst[0] = src[0]; dst[1] = src[1]; dst[2] = src[2]; dst[3] = src[3];
    hash_insert(table, m.hash_mask, vox_key[v], valid ? static_cast<int>(v) : -2 - static_cast<int>(v));
  }
  return valid ? 1 : (flagged ? 2 : 0);
}

void launch_voxel_finalize(const DeviceGrid& g, uint32_t n_points, unsigned long long publish_key, cudaStream_t stream) {
  const int fin = static_cast<int>((n_points + FIN_THREADS - 1) / FIN_THREADS);
  launch_pdl(voxel_stats_kernel, dim3(std::max(1, std::min(fin, 148 * FIN_MIN_BLOCKS))), dim3(FIN_THREADS), 0, stream,
             static_cast<const float4*>(g.sorted_pts), g.d_meta, static_cast<const uint32_t*>(g.seg_start), static_cast<const uint32_t*>(g.c_key),
             n_points, static_cast<const double*>(g.vox_sums), g.recs, g.table, g.c_n, g.c_mean, g.c_cov, g.c_icov, g.c_evecs, g.c_evals, g.h_slots,
             publish_key);
}

}  // namespace ndtb
