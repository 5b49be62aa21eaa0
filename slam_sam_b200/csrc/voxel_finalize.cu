// k4: per-voxel statistics (SURVEY 8a rows a2 pass 1, a3, a11).
// One thread per voxel.  The thread walks its segment of the (stably) sorted point list, i.e. its
// points IN INPUT ORDER, and accumulates Sum x and Sum x x^T in FP64 exactly like the reference's
// serial loop (voxel_grid_covariance_omp_impl.hpp:235-239 / voxel_grid_covariance_impl.hpp:237-239),
// so means and covariances are reproduced bit for bit, not just to rounding.
// Then: pclomp flavour  voxel_grid_covariance_omp_impl.hpp:286-360
//       svn flavour     voxel_grid_covariance_impl.hpp:270-343
// and the voxel is inserted into the open-addressing hash (atomicCAS linear probing).
#include "build_api.cuh"

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

__global__ void __launch_bounds__(128) voxel_finalize_kernel(const float4* __restrict__ pts, const uint32_t* __restrict__ sorted_vals,
                                                             const uint32_t* __restrict__ seg_start, const uint32_t* __restrict__ vox_key,
                                                             const unsigned int* __restrict__ counters, uint32_t n_points, GridMeta m,
                                                             VoxRec* __restrict__ recs, HashSlot* __restrict__ table, int32_t* __restrict__ c_n,
                                                             double* __restrict__ c_mean, double* __restrict__ c_cov, double* __restrict__ c_icov,
                                                             double* __restrict__ c_evecs, double* __restrict__ c_evals,
                                                             unsigned int* __restrict__ n_valid_out) {
  const uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t V = counters[1];
  if (v >= V) return;
  const uint32_t n_valid_pts = n_points - counters[0];
  const uint32_t lo = seg_start[v];
  const uint32_t hi = (v + 1 < V) ? seg_start[v + 1] : n_valid_pts;
  double sx = 0, sy = 0, sz = 0, sxx = 0, sxy = 0, sxz = 0, syy = 0, syz = 0, szz = 0;
  for (uint32_t i = lo; i < hi; ++i) {
    const float4 p = __ldg(pts + __ldg(sorted_vals + i));
    const double x = p.x, y = p.y, z = p.z;
    sx += x; sy += y; sz += z;
    sxx += x * x; sxy += x * y; sxz += x * z; syy += y * y; syz += y * z; szz += z * z;
  }
  int n = static_cast<int>(hi - lo);
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

void launch_voxel_finalize(const DeviceGrid& g, const float4* d_pts, const uint32_t* sorted_vals, uint32_t n_points_upper,
                           cudaStream_t stream) {
  const int threads = 128;
  const int blocks = static_cast<int>((n_points_upper + threads - 1) / threads);
  voxel_finalize_kernel<<<blocks, threads, 0, stream>>>(d_pts, sorted_vals, g.seg_start, g.c_key, g.d_counters, n_points_upper, g.meta,
                                                        g.recs, g.table, g.c_n, g.c_mean, g.c_cov, g.c_icov, g.c_evecs, g.c_evals,
                                                        g.d_counters + 2);
}

}  // namespace ndtb
