// k4: per-voxel statistics (SURVEY 8a rows a2 pass 1, a3, a11).
// One thread per voxel.  The thread walks its segment of the (stably) sorted point list, i.e. its
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

constexpr int FIN_BATCH = 4;   // points per software-pipeline stage (light voxels)
constexpr int FIN_HEAVY = 48;  // voxels with more points are summed by the whole warp

// Everything per voxel after the sums.  `n`, sums in; writes the cold arrays, the hot record and the hash slot.
__device__ void finalize_voxel(uint32_t v, int n, double sx, double sy, double sz, double sxx, double sxy, double sxz, double syy, double syz, double szz,
                               const uint32_t* __restrict__ vox_key, const GridMeta& m, VoxRec* __restrict__ recs, HashSlot* __restrict__ table,
                               int32_t* __restrict__ c_n, double* __restrict__ c_mean, double* __restrict__ c_cov, double* __restrict__ c_icov,
                               double* __restrict__ c_evecs, double* __restrict__ c_evals, unsigned int* __restrict__ n_valid_out);

// k4a: per-voxel sums (Sum x, Sum x x^T in input order) -> vox_sums[v][0..8], one resident wave.
__global__ void __launch_bounds__(128) voxel_sum_kernel(const double* __restrict__ sorted_xyz, const uint32_t* __restrict__ seg_start,
                                                        const unsigned int* __restrict__ counters, uint32_t n_points,
                                                        double* __restrict__ vox_sums) {
  __shared__ double s_prod[4][32][9];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const uint32_t V = counters[1];
  const uint32_t n_valid_pts = n_points - counters[0];
  // Voxels are dealt to threads through a multiplicative permutation of [0, Vr), Vr = V rounded up to a warp:
  // neighbouring voxels (a wall is a run of equally heavy voxels) land in different warps, so no warp collects
  // all the long segments.  2654435761 is prime, hence coprime with Vr: t -> t * P mod Vr is a bijection.
  const uint32_t Vr = (V + 31u) & ~31u;
  for (uint32_t t = blockIdx.x * blockDim.x + threadIdx.x; t < Vr; t += gridDim.x * blockDim.x) {  // warp-uniform trip count
    const uint32_t v = static_cast<uint32_t>((static_cast<unsigned long long>(t) * 2654435761ull) % Vr);
    const bool has = v < V;
    const uint32_t lo = has ? seg_start[v] : 0u;
    const uint32_t hi = has ? ((v + 1 < V) ? seg_start[v + 1] : n_valid_pts) : 0u;
    const int n = static_cast<int>(hi - lo);
    const bool heavy = n > FIN_HEAVY;
    // sx sy sz sxx sxy sxz syy syz szz.  The reference's Leaf starts cov_ at identity and adds x x^T onto it
    // (voxel_grid_covariance_omp.h:113 + _impl.hpp:239; svn voxel_grid_covariance.h:109 + _impl.hpp:240), so the
    // diagonal chains start at 1.0, not 0.
    double sum[9] = {0, 0, 0, 1.0, 0, 0, 1.0, 0, 1.0};
    if (has && !heavy) {
      // Light voxel: this lane adds its points IN INPUT ORDER (stable sort; already widened to FP64 by the
      // gather kernel), loads software-pipelined one batch ahead of the sequential FP64 chains.
      const double* q = sorted_xyz + 3 * static_cast<size_t>(lo);
      double cur[3 * FIN_BATCH], nxt[3 * FIN_BATCH];
      const int full = n / FIN_BATCH;
      if (full > 0) {
#pragma unroll
        for (int u = 0; u < 3 * FIN_BATCH; ++u) cur[u] = __ldg(q + u);
      }
      for (int bt = 0; bt < full; ++bt) {
        if (bt + 1 < full) {
#pragma unroll
          for (int u = 0; u < 3 * FIN_BATCH; ++u) nxt[u] = __ldg(q + 3 * FIN_BATCH * (bt + 1) + u);
        }
#pragma unroll
        for (int u = 0; u < FIN_BATCH; ++u) {
          const double x = cur[3 * u], y = cur[3 * u + 1], z = cur[3 * u + 2];
          sum[0] += x; sum[1] += y; sum[2] += z;
          sum[3] += x * x; sum[4] += x * y; sum[5] += x * z; sum[6] += y * y; sum[7] += y * z; sum[8] += z * z;
        }
#pragma unroll
        for (int u = 0; u < 3 * FIN_BATCH; ++u) cur[u] = nxt[u];
      }
      for (int i = full * FIN_BATCH; i < n; ++i) {
        const double x = __ldg(q + 3 * i), y = __ldg(q + 3 * i + 1), z = __ldg(q + 3 * i + 2);
        sum[0] += x; sum[1] += y; sum[2] += z;
        sum[3] += x * x; sum[4] += x * y; sum[5] += x * z; sum[6] += y * y; sum[7] += y * z; sum[8] += z * z;
      }
    }
    // Heavy voxels: the whole warp takes them one at a time.  32 lanes load 32 points (coalesced) and form the
    // nine products; lane c < 9 then runs chain c over those 32 points in order, so the nine serial FP64 chains
    // advance in the same instruction and a point costs one dependent DADD instead of nine.
    unsigned int hm = __ballot_sync(0xffffffffu, heavy);
    while (hm) {
      const int o = __ffs(hm) - 1;
      hm &= hm - 1;
      const uint32_t wlo = __shfl_sync(0xffffffffu, lo, o);
      const int wn = __shfl_sync(0xffffffffu, n, o);
      const double* q = sorted_xyz + 3 * static_cast<size_t>(wlo);
      double acc = (lane == 3 || lane == 6 || lane == 8) ? 1.0 : 0.0;  // identity-initialised second moment (see above)
      double x = 0, y = 0, z = 0;
      if (lane < wn) { x = __ldg(q + 3 * lane); y = __ldg(q + 3 * lane + 1); z = __ldg(q + 3 * lane + 2); }
      for (int base = 0; base < wn; base += 32) {
        double* pr = s_prod[w][lane];
        pr[0] = x; pr[1] = y; pr[2] = z; pr[3] = x * x; pr[4] = x * y; pr[5] = x * z; pr[6] = y * y; pr[7] = y * z; pr[8] = z * z;
        const int nxt_i = base + 32 + lane;
        if (nxt_i < wn) { x = __ldg(q + 3 * nxt_i); y = __ldg(q + 3 * nxt_i + 1); z = __ldg(q + 3 * nxt_i + 2); }  // next tile in flight
        __syncwarp();
        const int cnt = min(32, wn - base);
        if (lane < 9) {
          int k = 0;
          for (; k + 8 <= cnt; k += 8) {
            double t8[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) t8[u] = s_prod[w][k + u][lane];
#pragma unroll
            for (int u = 0; u < 8; ++u) acc += t8[u];
          }
          for (; k < cnt; ++k) acc += s_prod[w][k][lane];
        }
        __syncwarp();
      }
#pragma unroll
      for (int c = 0; c < 9; ++c) {
        const double tv = __shfl_sync(0xffffffffu, acc, c);
        if (lane == o) sum[c] = tv;
      }
    }
    if (has) {
      double* o = vox_sums + 9 * static_cast<size_t>(v);
#pragma unroll
      for (int c = 0; c < 9; ++c) o[c] = sum[c];
    }
  }
}

// k4b: one thread per voxel: mean / covariance / eigen-inflation / inverse, the hot record and the hash insert.
__global__ void __launch_bounds__(128, 5) voxel_finalize_kernel(const double* __restrict__ vox_sums, const uint32_t* __restrict__ seg_start,
                                                                const uint32_t* __restrict__ vox_key, unsigned int* __restrict__ counters,
                                                                uint32_t n_points, GridMeta m, VoxRec* __restrict__ recs, HashSlot* __restrict__ table,
                                                                int32_t* __restrict__ c_n, double* __restrict__ c_mean, double* __restrict__ c_cov,
                                                                double* __restrict__ c_icov, double* __restrict__ c_evecs, double* __restrict__ c_evals,
                                                                unsigned int* __restrict__ ticket, volatile unsigned long long* host_slots,
                                                                unsigned long long publish_key) {
  __shared__ bool s_last;
  const uint32_t V = counters[1];
  const uint32_t n_valid_pts = n_points - counters[0];
  for (uint32_t v = blockIdx.x * blockDim.x + threadIdx.x; v < V; v += gridDim.x * blockDim.x) {
    const uint32_t lo = seg_start[v];
    const uint32_t hi = (v + 1 < V) ? seg_start[v + 1] : n_valid_pts;
    const double* q = vox_sums + 9 * static_cast<size_t>(v);
    finalize_voxel(v, static_cast<int>(hi - lo), q[0], q[1], q[2], q[3], q[4], q[5], q[6], q[7], q[8], vox_key, m, recs, table, c_n, c_mean, c_cov,
                   c_icov, c_evecs, c_evals, counters + 2);
  }
  // ---- last block done: publish {skipped, V, n_valid} to the host slots 8..10 ----------------------------
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned int tk;
    asm volatile("atom.acq_rel.gpu.global.add.u32 %0, [%1], 1;" : "=r"(tk) : "l"(ticket) : "memory");
    s_last = (tk == gridDim.x - 1);
  }
  __syncthreads();
  if (s_last && threadIdx.x < 3) {
    const unsigned int c = atomicAdd(counters + threadIdx.x, 0u);
    publish_slot(host_slots, 8 + threadIdx.x, static_cast<unsigned long long>(c), publish_key);
    if (threadIdx.x == 0) *ticket = 0;
  }
}

__device__ void finalize_voxel(uint32_t v, int n, double sx, double sy, double sz, double sxx, double sxy, double sxz, double syy, double syz, double szz,
                               const uint32_t* __restrict__ vox_key, const GridMeta& m, VoxRec* __restrict__ recs, HashSlot* __restrict__ table,
                               int32_t* __restrict__ c_n, double* __restrict__ c_mean, double* __restrict__ c_cov, double* __restrict__ c_icov,
                               double* __restrict__ c_evecs, double* __restrict__ c_evals, unsigned int* __restrict__ n_valid_out) {
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
  if (valid) {
    VoxRec r;
    r.mean[0] = mean[0]; r.mean[1] = mean[1]; r.mean[2] = mean[2];
#pragma unroll
    for (int i = 0; i < 9; ++i) r.icov[i] = static_cast<float>(icov[i]);
    r.n = n;
    // 64-byte record written as four 16-byte stores
    const uint4* src = reinterpret_cast<const uint4*>(&r);
    uint4* dst = reinterpret_cast<uint4*>(recs + v);
    dst[0] = src[0]; dst[1] = src[1]; dst[2] = src[2]; dst[3] = src[3];
    hash_insert(table, m.hash_mask, vox_key[v], static_cast<int>(v));
    atomicAdd(n_valid_out, 1u);
  }
}

void launch_voxel_finalize(const DeviceGrid& g, uint32_t n_points_upper, unsigned long long publish_key, cudaStream_t stream) {
  const int threads = 128;
  const int need = static_cast<int>((n_points_upper + threads - 1) / threads);
  voxel_sum_kernel<<<std::min(need, 148 * 4), threads, 0, stream>>>(g.sorted_xyz, g.seg_start, g.d_counters, n_points_upper, g.vox_sums);
  voxel_finalize_kernel<<<std::min(need, 148 * 5), threads, 0, stream>>>(g.vox_sums, g.seg_start, g.c_key, g.d_counters, n_points_upper, g.meta, g.recs,
                                                                      g.table, g.c_n, g.c_mean, g.c_cov, g.c_icov, g.c_evecs, g.c_evals,
                                                                      g.d_ticket + 1, g.h_slots, publish_key);
}

}  // namespace ndtb
