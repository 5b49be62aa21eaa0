// The hot kernel: fused rigid transform + DIRECT1/DIRECT7 voxel gather + per-pair NDT
// score/gradient/Hessian + deterministic FP64 reduction  (SURVEY 8a rows a4-a8, a12-a15).
//
//   pclomp flavour: computeDerivatives   ndt_omp_impl.hpp:217-347
//                   getNeighborhoodAtPoint voxel_grid_covariance_omp_impl.hpp:368-399
//                   computePointDerivatives :542-587, updateDerivatives :640-697
//   svn flavour   : computeParticleDerivatives svn_ndt_impl.hpp:518-647, updateDerivatives :401-513,
//                   getLeaf / getNeighborhoodAtPoint7 voxel_grid_covariance.h:280-304, _impl.hpp:560-615
//
// One thread per source point (grid.y = particle).  The transformed cloud is never materialised.
// Per-pair maths is float32 in the reference's evaluation order with structural zeros/ones of the
// 4x6 / 24x6 point Jacobian/Hessian skipped (x*1, x*0 and +0 are exact, so the result is bit-identical
// to the dense Eigen expressions for finite inputs); accumulation is FP64 per thread in neighbour
// order, then a fixed-shape tree (warp shuffles -> shared memory -> last-block pass), so the result
// is run-to-run bit stable and differs from the reference's serial in-index-order sum only by FP64
// summation order.  Bound: HBM/L2 gather, B_eval = N*(16 + 8*P + 64*hbar) bytes.
#include "eval_api.cuh"
#include "handle.cuh"       // ndtb_compute_derivatives_static (end of this file)
#include "newton_math.cuh"
#include "device_newton.cuh"
#include <algorithm>
#include <cstring>
#include <string>

namespace ndtb {

__device__ __forceinline__ float dot3(float a0, float a1, float a2, float b0, float b1, float b2) {
  float s = a0 * b0;
  s = s + a1 * b1;
  s = s + a2 * b2;
  return s;
}

struct PointTables {  // per-point angular parts of Jp (8) and Hp (15)
  float xj[8];
  float xh[15];
};

// ------------------------------------------------------------------------------ pclomp pair maths
template <bool HESS>
__device__ __forceinline__ float pair_pclomp(const PointTables& pt, const float x4[3], const float C[9], float d2f, double d1,
                                             double* __restrict__ acc /* RED_N */, bool& counted) {
  float xC[3];
#pragma unroll
  for (int c = 0; c < 3; ++c) xC[c] = dot3(x4[0], x4[1], x4[2], C[c], C[3 + c], C[6 + c]);
  const float q = dot3(x4[0], x4[1], x4[2], xC[0], xC[1], xC[2]);
  float e = static_cast<float>(exp(static_cast<double>(-d2f * q * 0.5f)));
  const float score_inc = static_cast<float>(-d1 * static_cast<double>(e));
  e = d2f * e;
  counted = true;
  if (e > 1 || e < 0 || e != e) return 0.0f;
  e = static_cast<float>(static_cast<double>(e) * d1);
  const float* xj = pt.xj;
  float CJ[3][6];
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    CJ[r][0] = C[3 * r + 0]; CJ[r][1] = C[3 * r + 1]; CJ[r][2] = C[3 * r + 2];
    CJ[r][3] = C[3 * r + 1] * xj[0] + C[3 * r + 2] * xj[1];
    CJ[r][4] = dot3(C[3 * r], C[3 * r + 1], C[3 * r + 2], xj[2], xj[3], xj[4]);
    CJ[r][5] = dot3(C[3 * r], C[3 * r + 1], C[3 * r + 2], xj[5], xj[6], xj[7]);
  }
  float xCJ[6];
#pragma unroll
  for (int c = 0; c < 6; ++c) xCJ[c] = dot3(x4[0], x4[1], x4[2], CJ[0][c], CJ[1][c], CJ[2][c]);
#pragma unroll
  for (int c = 0; c < 6; ++c) acc[RED_G + c] += static_cast<double>(e * xCJ[c]);
  if (HESS) {
    const float* xh = pt.xh;
    // xCH for the 3x3 angular block: rows/cols 3..5 -> vectors a b c / b d e / c e f
    const float h_a = xC[1] * xh[0] + xC[2] * xh[1];
    const float h_b = xC[1] * xh[2] + xC[2] * xh[3];
    const float h_c = xC[1] * xh[4] + xC[2] * xh[5];
    const float h_d = dot3(xC[0], xC[1], xC[2], xh[6], xh[7], xh[8]);
    const float h_e = dot3(xC[0], xC[1], xC[2], xh[9], xh[10], xh[11]);
    const float h_f = dot3(xC[0], xC[1], xC[2], xh[12], xh[13], xh[14]);
    const float xCH[3][3] = {{h_a, h_b, h_c}, {h_b, h_d, h_e}, {h_c, h_e, h_f}};
#pragma unroll
    for (int i = 0; i < 6; ++i) {
#pragma unroll
      for (int j = 0; j < 6; ++j) {
        // JCJ(j,i) = Jp(:,j)^T * CJ(:,i)
        float jcj;
        if (j < 3) jcj = CJ[j][i];
        else if (j == 3) jcj = xj[0] * CJ[1][i] + xj[1] * CJ[2][i];
        else if (j == 4) jcj = dot3(xj[2], xj[3], xj[4], CJ[0][i], CJ[1][i], CJ[2][i]);
        else jcj = dot3(xj[5], xj[6], xj[7], CJ[0][i], CJ[1][i], CJ[2][i]);
        float t = -d2f * xCJ[i] * xCJ[j];
        if (i >= 3 && j >= 3) t = t + xCH[i - 3][j - 3];
        t = t + jcj;
        acc[RED_H + 6 * i + j] += static_cast<double>(e * t);
      }
    }
  }
  return score_inc;
}

// Same arithmetic as pair_pclomp, but the 43 float terms (score, 6 gradient, 36 Hessian entries) are
// RETURNED instead of accumulated: t[0] = score_inc, t[1..6] = e*xCJ, t[7..42] = e*(...), exactly the
// floats the reference widens to double and adds (ndt_omp_impl.hpp:657,671,691).  A rejected pair
// (:662) contributes nothing: all terms are +0.
template <bool HESS>
__device__ __forceinline__ void pclomp_terms(const PointTables& pt, const float x4[3], const float C[9], float d2f, double d1, float* __restrict__ t) {
  float xC[3];
#pragma unroll
  for (int c = 0; c < 3; ++c) xC[c] = dot3(x4[0], x4[1], x4[2], C[c], C[3 + c], C[6 + c]);
  const float q = dot3(x4[0], x4[1], x4[2], xC[0], xC[1], xC[2]);
  float e = static_cast<float>(exp(static_cast<double>(-d2f * q * 0.5f)));
  const float score_inc = static_cast<float>(-d1 * static_cast<double>(e));
  e = d2f * e;
  const bool reject = (e > 1 || e < 0 || e != e);
  e = static_cast<float>(static_cast<double>(e) * d1);
  // a rejected pair adds nothing (:662).  Its terms are e * (finite) with e forced to 0, i.e. +-0: adding a
  // signed zero to an FP64 sum leaves it unchanged, so no per-term select is needed.
  if (reject) e = 0.0f;
  const float* xj = pt.xj;
  float CJ[3][6];
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    CJ[r][0] = C[3 * r + 0]; CJ[r][1] = C[3 * r + 1]; CJ[r][2] = C[3 * r + 2];
    CJ[r][3] = C[3 * r + 1] * xj[0] + C[3 * r + 2] * xj[1];
    CJ[r][4] = dot3(C[3 * r], C[3 * r + 1], C[3 * r + 2], xj[2], xj[3], xj[4]);
    CJ[r][5] = dot3(C[3 * r], C[3 * r + 1], C[3 * r + 2], xj[5], xj[6], xj[7]);
  }
  float xCJ[6];
#pragma unroll
  for (int c = 0; c < 6; ++c) xCJ[c] = dot3(x4[0], x4[1], x4[2], CJ[0][c], CJ[1][c], CJ[2][c]);
  t[0] = reject ? 0.0f : score_inc;
#pragma unroll
  for (int c = 0; c < 6; ++c) t[1 + c] = e * xCJ[c];
  if (HESS) {
    const float* xh = pt.xh;
    const float h_a = xC[1] * xh[0] + xC[2] * xh[1];
    const float h_b = xC[1] * xh[2] + xC[2] * xh[3];
    const float h_c = xC[1] * xh[4] + xC[2] * xh[5];
    const float h_d = dot3(xC[0], xC[1], xC[2], xh[6], xh[7], xh[8]);
    const float h_e = dot3(xC[0], xC[1], xC[2], xh[9], xh[10], xh[11]);
    const float h_f = dot3(xC[0], xC[1], xC[2], xh[12], xh[13], xh[14]);
    const float xCH[3][3] = {{h_a, h_b, h_c}, {h_b, h_d, h_e}, {h_c, h_e, h_f}};
#pragma unroll
    for (int i = 0; i < 6; ++i) {
#pragma unroll
      for (int j = 0; j < 6; ++j) {
        float jcj;  // JCJ(j,i) = Jp(:,j)^T * CJ(:,i)
        if (j < 3) jcj = CJ[j][i];
        else if (j == 3) jcj = xj[0] * CJ[1][i] + xj[1] * CJ[2][i];
        else if (j == 4) jcj = dot3(xj[2], xj[3], xj[4], CJ[0][i], CJ[1][i], CJ[2][i]);
        else jcj = dot3(xj[5], xj[6], xj[7], CJ[0][i], CJ[1][i], CJ[2][i]);
        float v = -d2f * xCJ[i] * xCJ[j];
        if (i >= 3 && j >= 3) v = v + xCH[i - 3][j - 3];
        v = v + jcj;
        t[7 + 6 * i + j] = e * v;
      }
    }
  }
}

// Packed-FP32 version of pclomp_terms for sm_100a: the same IEEE round-to-nearest multiplies and adds in the
// same association, with the multiplies issued two at a time as FMUL2 (__fmul2_rn), which removes a quarter of
// the issue slots of the per-pair maths.  Element results are bit-identical to the scalar code.  The point tables come from a
// PtSlot written in the PAIRED layout (see phase_a<..., PAIRED = true>):
//   words 4,5: xj0 xj1 | 6..11: (xj2,xj5) (xj3,xj6) (xj4,xj7) | 12..15: (xh0,xh2) (xh1,xh3)
//   16..21: (xh6,xh9) (xh7,xh10) (xh8,xh11) | 22,23: xh4 xh5 | 24..26: xh12 xh13 xh14
__device__ __forceinline__ float2 bc2(float x) { return make_float2(x, x); }
__device__ __forceinline__ float2 mul2(float2 a, float2 b) { return __fmul2_rn(a, b); }
// Packed additions: ptxas 12.9 contracts mul.rn.f32x2 + add.rn.f32x2 into FFMA2 even with explicit .rn and
// --fmad=false (checked with cuobjdump), which would change the reference's two-rounding arithmetic.  An addition
// written as fma(a, 1, b) cannot be contracted with the multiply that produced `a` (a*1+b rounds once and is
// exactly a+b, signed zeros and NaNs included), provided ptxas cannot see that the multiplier is 1: it comes from
// the evaluation state (EvalState::one), i.e. from the constant bank or shared memory.
__device__ __forceinline__ float2 add2(float2 a, float2 b, float2 one2) { return __ffma2_rn(a, one2, b); }
// (a0*b0 + a1*b1) + a2*b2 on pairs: dot3's association
__device__ __forceinline__ float2 dot3_2(float a0, float a1, float a2, float2 b0, float2 b1, float2 b2, float2 one2) {
  return add2(add2(mul2(bc2(a0), b0), mul2(bc2(a1), b1), one2), mul2(bc2(a2), b2), one2);
}

template <bool HESS>
// Output: 22 PAIRS in the order the packed multiplies produce them (no register shuffling before the stores):
//   tp[0] = (score, 0) | tp[1..3] = (g0,g1) (g2,g3) (g4,g5) | tp[4 + 3j + p] = (H(2p, j), H(2p+1, j)), j = 0..5, p = 0..2
// pair_word_component() maps a word of this layout back to the RED_* index.
__device__ __forceinline__ void pclomp_terms2(const float4* __restrict__ sp, const float x4[3], const float C[9], float d2f, double d1,
                                              float one, float2* __restrict__ tp) {
  const float2 one2 = bc2(one);
  const float4 s1 = sp[1], s2 = sp[2];
  const float xj0 = s1.x, xj1 = s1.y;
  const float2 J25 = make_float2(s1.z, s1.w), J36 = make_float2(s2.x, s2.y), J47 = make_float2(s2.z, s2.w);
  float xC[3];
#pragma unroll
  for (int c = 0; c < 3; ++c) xC[c] = dot3(x4[0], x4[1], x4[2], C[c], C[3 + c], C[6 + c]);
  const float q = dot3(x4[0], x4[1], x4[2], xC[0], xC[1], xC[2]);
  float e = static_cast<float>(exp(static_cast<double>(-d2f * q * 0.5f)));
  const float score_inc = static_cast<float>(-d1 * static_cast<double>(e));
  e = d2f * e;
  const bool reject = (e > 1 || e < 0 || e != e);
  e = static_cast<float>(static_cast<double>(e) * d1);
  if (reject) e = 0.0f;  // see pclomp_terms
  // CJ = C4 * Jp, rows r, columns as pairs (0,1) (2,3) (4,5)
  float2 cj[3][3];
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    const float c0 = C[3 * r], c1 = C[3 * r + 1], c2 = C[3 * r + 2];
    cj[r][0] = make_float2(c0, c1);
    cj[r][1] = make_float2(c2, c1 * xj0 + c2 * xj1);
    cj[r][2] = dot3_2(c0, c1, c2, J25, J36, J47, one2);
  }
  float2 xcj[3];
#pragma unroll
  for (int p = 0; p < 3; ++p) xcj[p] = dot3_2(x4[0], x4[1], x4[2], cj[0][p], cj[1][p], cj[2][p], one2);
  tp[0] = make_float2(reject ? 0.0f : score_inc, 0.0f);
  const float2 e2 = bc2(e);
#pragma unroll
  for (int p = 0; p < 3; ++p) tp[1 + p] = mul2(e2, xcj[p]);
  if (HESS) {
    const float4 s3 = sp[3], s4 = sp[4], s5 = sp[5], s6 = sp[6];
    const float2 hab = add2(mul2(bc2(xC[1]), make_float2(s3.x, s3.y)), mul2(bc2(xC[2]), make_float2(s3.z, s3.w)), one2);  // (h_a, h_b)
    const float2 hde = dot3_2(xC[0], xC[1], xC[2], make_float2(s4.x, s4.y), make_float2(s4.z, s4.w), make_float2(s5.x, s5.y), one2);  // (h_d, h_e)
    const float h_c = xC[1] * s5.z + xC[2] * s5.w;
    const float h_f = dot3(xC[0], xC[1], xC[2], s6.x, s6.y, s6.z);
    // xCH rows for i = 3,4,5 and columns j = 3,4,5: [a b c; b d e; c e f]
    const float xch3[3] = {hab.x, hab.y, h_c};                                                          // row i = 3, j = 3..5
    const float2 xch45[3] = {make_float2(hab.y, h_c), make_float2(hde.x, hde.y), make_float2(hde.y, h_f)};  // (row 4, row 5) for j = 3..5
    const float2 nd2 = bc2(-d2f);
    float2 ap[3];
#pragma unroll
    for (int p = 0; p < 3; ++p) ap[p] = mul2(nd2, xcj[p]);  // (-d2 * xCJ_i) for i = 2p, 2p+1
    const float xjs[8] = {xj0, xj1, J25.x, J36.x, J47.x, J25.y, J36.y, J47.y};
#pragma unroll
    for (int j = 0; j < 6; ++j) {
      const float xcjj = (j & 1) ? xcj[j >> 1].y : xcj[j >> 1].x;
      const float2 xcjj2 = bc2(xcjj);
#pragma unroll
      for (int p = 0; p < 3; ++p) {
        float2 jc;  // JCJ(j, i) for i = 2p, 2p+1
        if (j < 3) jc = cj[j][p];
        else if (j == 3) jc = add2(mul2(bc2(xjs[0]), cj[1][p]), mul2(bc2(xjs[1]), cj[2][p]), one2);
        else if (j == 4) jc = dot3_2(xjs[2], xjs[3], xjs[4], cj[0][p], cj[1][p], cj[2][p], one2);
        else jc = dot3_2(xjs[5], xjs[6], xjs[7], cj[0][p], cj[1][p], cj[2][p], one2);
        float2 v = mul2(ap[p], xcjj2);
        if (j >= 3 && p == 1) v.y = v.y + xch3[j - 3];
        if (j >= 3 && p == 2) v = add2(v, xch45[j - 3], one2);
        v = add2(v, jc, one2);
        tp[4 + 3 * j + p] = mul2(e2, v);
      }
    }
  }
}

// ------------------------------------------------------------------------------ svn pair maths
template <bool GN>
__device__ __forceinline__ double pair_svn(const PointTables& pt, const double xr[3], const double Cd[9], double d1, double d2,
                                           double* __restrict__ acc) {
  float x4[3], C[9];
#pragma unroll
  for (int a = 0; a < 3; ++a) x4[a] = static_cast<float>(xr[a]);
#pragma unroll
  for (int i = 0; i < 9; ++i) C[i] = static_cast<float>(Cd[i]);
  double Cx[3];
#pragma unroll
  for (int r = 0; r < 3; ++r) Cx[r] = (Cd[3 * r] * xr[0] + Cd[3 * r + 1] * xr[1]) + Cd[3 * r + 2] * xr[2];
  double mahal = (xr[0] * Cx[0] + xr[1] * Cx[1]) + xr[2] * Cx[2];
  if (!isfinite(mahal) || mahal < -1e-9) return 0.0;
  if (mahal < 0.0) mahal = 0.0;
  const double earg = d2 * mahal * 0.5;
  if (earg > 50.0) return 0.0;
  const double exp_term = exp(-earg);
  const double score_inc = -d1 * exp_term;
  const double factor = d1 * d2 * exp_term;
  if (!isfinite(factor) || fabs(factor) < 1e-15) return score_inc;
  const float* xj = pt.xj;
  float CJ[3][6];
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    CJ[r][0] = C[3 * r + 0]; CJ[r][1] = C[3 * r + 1]; CJ[r][2] = C[3 * r + 2];
    CJ[r][3] = C[3 * r + 1] * xj[0] + C[3 * r + 2] * xj[1];
    CJ[r][4] = dot3(C[3 * r], C[3 * r + 1], C[3 * r + 2], xj[2], xj[3], xj[4]);
    CJ[r][5] = dot3(C[3 * r], C[3 * r + 1], C[3 * r + 2], xj[5], xj[6], xj[7]);
  }
  float gc[6];
  double ginc[6];
  bool gfin = true;
#pragma unroll
  for (int c = 0; c < 6; ++c) {
    gc[c] = dot3(x4[0], x4[1], x4[2], CJ[0][c], CJ[1][c], CJ[2][c]);
    ginc[c] = factor * static_cast<double>(gc[c]);
    if (!isfinite(ginc[c])) gfin = false;
  }
  if (gfin) {
#pragma unroll
    for (int c = 0; c < 6; ++c) acc[RED_G + c] += ginc[c];
  }
  double hc[36];
  bool hfin = true;
  float xC[3];
  float xCH[3][3];
  if (!GN) {
#pragma unroll
    for (int c = 0; c < 3; ++c) xC[c] = dot3(x4[0], x4[1], x4[2], C[c], C[3 + c], C[6 + c]);
    const float* xh = pt.xh;
    const float h_a = xC[1] * xh[0] + xC[2] * xh[1];
    const float h_b = xC[1] * xh[2] + xC[2] * xh[3];
    const float h_c = xC[1] * xh[4] + xC[2] * xh[5];
    const float h_d = dot3(xC[0], xC[1], xC[2], xh[6], xh[7], xh[8]);
    const float h_e = dot3(xC[0], xC[1], xC[2], xh[9], xh[10], xh[11]);
    const float h_f = dot3(xC[0], xC[1], xC[2], xh[12], xh[13], xh[14]);
    xCH[0][0] = h_a; xCH[0][1] = h_b; xCH[0][2] = h_c;
    xCH[1][0] = h_b; xCH[1][1] = h_d; xCH[1][2] = h_e;
    xCH[2][0] = h_c; xCH[2][1] = h_e; xCH[2][2] = h_f;
  }
#pragma unroll
  for (int i = 0; i < 6; ++i) {
#pragma unroll
    for (int j = 0; j < 6; ++j) {
      // term2(i,j) = Jp(:,i)^T * CJ(:,j)
      float jcj;
      if (i < 3) jcj = CJ[i][j];
      else if (i == 3) jcj = xj[0] * CJ[1][j] + xj[1] * CJ[2][j];
      else if (i == 4) jcj = dot3(xj[2], xj[3], xj[4], CJ[0][j], CJ[1][j], CJ[2][j]);
      else jcj = dot3(xj[5], xj[6], xj[7], CJ[0][j], CJ[1][j], CJ[2][j]);
      double h = static_cast<double>(jcj);
      if (!GN) {
        const double term1 = -d2 * (static_cast<double>(gc[i]) * static_cast<double>(gc[j]));
        const double term3 = (i >= 3 && j >= 3) ? static_cast<double>(xCH[i - 3][j - 3]) : 0.0;
        h = (term1 + h) + term3;
      }
      h = h * factor;
      if (!isfinite(h)) hfin = false;
      hc[6 * i + j] = h;
    }
  }
  if (hfin) {
#pragma unroll
    for (int i = 0; i < 36; ++i) acc[RED_H + i] += hc[i];
  }
  return score_inc;
}

// ------------------------------------------------------------------------------ neighbour lookup
// Linear index of the neighbour cell, or HASH_EMPTY when the displacement leaves the bounding box
// (voxel_grid_covariance_omp_impl.hpp:379-394: displacement must lie in [min_b - ijk, max_b - ijk]).
__device__ __forceinline__ uint32_t cell_key_pclomp(const GridMeta& m, int i0, int i1, int i2, int ox, int oy, int oz) {
  if (m.min_b[0] - i0 > ox || m.max_b[0] - i0 < ox) return HASH_EMPTY;
  if (m.min_b[1] - i1 > oy || m.max_b[1] - i1 < oy) return HASH_EMPTY;
  if (m.min_b[2] - i2 > oz || m.max_b[2] - i2 < oz) return HASH_EMPTY;
  return static_cast<uint32_t>((i0 + ox - m.min_b[0]) * m.divb_mul[0] + (i1 + oy - m.min_b[1]) * m.divb_mul[1] + (i2 + oz - m.min_b[2]) * m.divb_mul[2]);
}

// Cell key of svn's getLeaf (voxel_grid_covariance_impl.hpp:59-70 + voxel_grid_covariance.h:289-303)
__device__ __forceinline__ uint32_t cell_key_svn(const GridMeta& m, float px, float py, float pz) {
  if (!(px >= m.lo_bound[0] && px < m.hi_bound[0] && py >= m.lo_bound[1] && py < m.hi_bound[1] && pz >= m.lo_bound[2] && pz < m.hi_bound[2]))
    return HASH_EMPTY;
  const int i0 = static_cast<int>(floorf(px * m.inv_leaf) - static_cast<float>(m.min_b[0]));
  const int i1 = static_cast<int>(floorf(py * m.inv_leaf) - static_cast<float>(m.min_b[1]));
  const int i2 = static_cast<int>(floorf(pz * m.inv_leaf) - static_cast<float>(m.min_b[2]));
  return static_cast<uint32_t>(i0 * m.divb_mul[0] + i1 * m.divb_mul[1] + i2 * m.divb_mul[2]);
}

// All P home slots are fetched before any is examined (one L2 round trip instead of P); the rare
// displaced key (load factor <= 0.5) continues with the usual linear probe.
template <int PROBES>
__device__ __forceinline__ void hash_find_batch(const HashSlot* __restrict__ table, uint32_t mask, const uint32_t* key, int* rows) {
  uint2 s[PROBES];
#pragma unroll
  for (int q = 0; q < PROBES; ++q) s[q] = (key[q] != HASH_EMPTY) ? __ldg(reinterpret_cast<const uint2*>(table) + (key[q] & mask)) : make_uint2(HASH_EMPTY, 0u);
#pragma unroll
  for (int q = 0; q < PROBES; ++q) {
    uint2 v = s[q];
    uint32_t b = key[q] & mask;
    while (v.x != key[q] && v.x != HASH_EMPTY) {
      b = (b + 1) & mask;
      v = __ldg(reinterpret_cast<const uint2*>(table) + b);
    }
    rows[q] = (v.x == HASH_EMPTY) ? -1 : static_cast<int>(v.y);
  }
}

__device__ __constant__ int c_off[7][3] = {{0, 0, 0}, {1, 0, 0}, {-1, 0, 0}, {0, 1, 0}, {0, -1, 0}, {0, 0, 1}, {0, 0, -1}};
// KDTREE (radius search, radius = resolution = leaf size): only the 27 cells around the query can hold a leaf whose mean is closer
// than one leaf size, so the kd-tree is 27 hash probes + FLANN's float distance test (phase_a).
__device__ __constant__ int c_off27[27][3] = {
    {0, 0, 0}, {1, 0, 0}, {-1, 0, 0}, {0, 1, 0}, {0, -1, 0}, {0, 0, 1}, {0, 0, -1},
    {1, 1, 0}, {1, -1, 0}, {-1, 1, 0}, {-1, -1, 0}, {1, 0, 1}, {1, 0, -1}, {-1, 0, 1}, {-1, 0, -1}, {0, 1, 1}, {0, 1, -1}, {0, -1, 1}, {0, -1, -1},
    {1, 1, 1}, {1, 1, -1}, {1, -1, 1}, {1, -1, -1}, {-1, 1, 1}, {-1, 1, -1}, {-1, -1, 1}, {-1, -1, -1}};
// DIRECT26 (voxel_grid_covariance_omp_impl.hpp:366-411): pcl::getAllNeighborCellIndices() = the 13 "half" displacements of
// pcl::getHalfNeighborCellIndices() followed by their negatives; the centre cell is NOT among them.
__device__ __constant__ int c_off26[26][3] = {
    {-1, -1, -1}, {-1, 0, -1}, {-1, 1, -1}, {0, -1, -1}, {0, 0, -1}, {0, 1, -1}, {1, -1, -1}, {1, 0, -1}, {1, 1, -1},
    {-1, -1, 0}, {0, -1, 0}, {1, -1, 0}, {-1, 0, 0},
    {1, 1, 1}, {1, 0, 1}, {1, -1, 1}, {0, 1, 1}, {0, 0, 1}, {0, -1, 1}, {-1, 1, 1}, {-1, 0, 1}, {-1, -1, 1},
    {1, 1, 0}, {0, 1, 0}, {-1, 1, 0}, {1, 0, 0}};

// ------------------------------------------------------------------------------ reductions
constexpr int RED_PAD = 48;  // RED_N padded so the vector can be halved 4 times (48,24,12,6,3)

// Transpose-reduce across a warp: at each butterfly step a lane keeps one half of its values and
// sends the other half, so 48 values cost 24+12+6+3+3 = 48 64-bit shuffles instead of 48*5.
// Afterwards even lanes hold 3 finished sums each; component index = b4*24 + b3*12 + b2*6 + b1*3 + j.
// The summation tree is fixed by lane ids -> run-to-run bit stable.
template <int N, int OFF>
__device__ __forceinline__ void halve_step(double* v, int lane) {
  const bool up = (lane & OFF) != 0;
#pragma unroll
  for (int i = 0; i < N / 2; ++i) {
    const double keep = up ? v[N / 2 + i] : v[i];
    const double send = up ? v[i] : v[N / 2 + i];
    v[i] = keep + __shfl_xor_sync(0xffffffffu, send, OFF);
  }
}

// Reduce acc[RED_PAD] over the block and store RED_N sums to `out` (fixed order: warp tree, then
// warps 0..W-1 in order).
__device__ __forceinline__ void block_reduce_store(double* acc, double (*s_red)[RED_PAD], double* __restrict__ out, size_t stride) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  halve_step<48, 16>(acc, lane);
  halve_step<24, 8>(acc, lane);
  halve_step<12, 4>(acc, lane);
  halve_step<6, 2>(acc, lane);
#pragma unroll
  for (int j = 0; j < 3; ++j) acc[j] += __shfl_xor_sync(0xffffffffu, acc[j], 1);
  if ((lane & 1) == 0) {
    const int base = ((lane >> 4) & 1) * 24 + ((lane >> 3) & 1) * 12 + ((lane >> 2) & 1) * 6 + ((lane >> 1) & 1) * 3;
#pragma unroll
    for (int j = 0; j < 3; ++j) s_red[w][base + j] = acc[j];
  }
  __syncthreads();
  if (threadIdx.x < RED_N) {
    double v = s_red[0][threadIdx.x];
#pragma unroll
    for (int k = 1; k < EVAL_THREADS / 32; ++k) v += s_red[k][threadIdx.x];
    out[threadIdx.x * stride] = v;
  }
}

// ------------------------------------------------------------------------------ the kernel
// Every WARP is autonomous (no block barrier in the main loop) and alternates two phases that are
// decoupled through a small shared-memory pair queue of its own:
//   phase A (one lane per POINT, 32 points): transform, the P hash probes (independent loads -> one
//            or two L2 round trips), an L1 prefetch of every hit voxel record, the per-point
//            Jacobian/Hessian tables (computePointDerivatives) into a shared-memory slot, and the
//            point's (slot,row) pairs appended to the queue by a warp scan;
//   phase B (one lane per PAIR): 32 real (point,voxel) pairs evaluate updateDerivatives, so no lane
//            idles on a missed probe (h-bar/P of the lanes would otherwise be masked off).
// Pairs that do not fill a round stay queued for the next chunk.  Points are dealt to warps in
// groups of PT_GROUP consecutive points, round-robin, so every warp sees the same hit density.
// The point -> warp -> queue position -> lane mapping is static: every accumulator sees its terms
// in a fixed order (run-to-run bit stable).
constexpr int PT_GROUP = 4;                   // consecutive points per deal (64 B of the cloud)
constexpr int GROUPS_PER_CHUNK = 32 / PT_GROUP;
constexpr int Q_CAP = 256;                    // >= 31 + 7 * 32, power of two
constexpr int SLOT_RING = 64;                 // point slots of the current and the previous chunk
constexpr int EVAL_WARPS = EVAL_THREADS / 32;

struct __align__(16) PtSlot {  // 28 words
  float xt[3];        // transformed point
  int nearest_bits;   // max pair score of the point as float bits (>= 0): NVTL, ndt_omp_impl.hpp:384-389
  float xj[8];
  float xh[16];
};

constexpr int Q_CAP26 = 512;                  // >= 31 + 26 * 16: DIRECT26 through the generic kernel, 16 points per chunk (static smem <= 48 KB)
template <int QC>
struct WarpScratchT {
  PtSlot pts[SLOT_RING];
  int qrow[QC];
  unsigned char qslot[QC];
};
using WarpScratch = WarpScratchT<Q_CAP>;

template <int FLAVOUR, bool HESS, bool GN>
__device__ __forceinline__ void eval_pair(PtSlot* __restrict__ s_pts, int slot, int row, const EvalState& st, const VoxRec* __restrict__ recs,
                                          const double* __restrict__ c_mean, const double* __restrict__ c_icov, float d2f,
                                          double* __restrict__ acc) {
  PointTables pt;
  float xt[3];
  {
    const float4* sp = reinterpret_cast<const float4*>(s_pts + slot);
    const float4 a = sp[0], b = sp[1], c = sp[2];
    xt[0] = a.x; xt[1] = a.y; xt[2] = a.z;
    pt.xj[0] = b.x; pt.xj[1] = b.y; pt.xj[2] = b.z; pt.xj[3] = b.w; pt.xj[4] = c.x; pt.xj[5] = c.y; pt.xj[6] = c.z; pt.xj[7] = c.w;
    if (HESS && !(FLAVOUR == FLAVOUR_SVN && GN)) {
      const float4 d = sp[3], e = sp[4], f = sp[5], g = sp[6];
      pt.xh[0] = d.x; pt.xh[1] = d.y; pt.xh[2] = d.z; pt.xh[3] = d.w; pt.xh[4] = e.x; pt.xh[5] = e.y; pt.xh[6] = e.z; pt.xh[7] = e.w;
      pt.xh[8] = f.x; pt.xh[9] = f.y; pt.xh[10] = f.z; pt.xh[11] = f.w; pt.xh[12] = g.x; pt.xh[13] = g.y; pt.xh[14] = g.z;
    }
  }
  if (FLAVOUR == FLAVOUR_PCLOMP) {
    const uint4* rp = reinterpret_cast<const uint4*>(recs + row);
    const uint4 q0 = __ldg(rp), q1 = __ldg(rp + 1), q2 = __ldg(rp + 2), q3 = __ldg(rp + 3);
    const double m0 = __hiloint2double(q0.y, q0.x), m1 = __hiloint2double(q0.w, q0.z), m2 = __hiloint2double(q1.y, q1.x);
    float C[9];
    C[0] = __uint_as_float(q1.z); C[1] = __uint_as_float(q1.w);
    C[2] = __uint_as_float(q2.x); C[3] = __uint_as_float(q2.y); C[4] = __uint_as_float(q2.z); C[5] = __uint_as_float(q2.w);
    C[6] = __uint_as_float(q3.x); C[7] = __uint_as_float(q3.y); C[8] = __uint_as_float(q3.z);
    float x4[3];
    x4[0] = static_cast<float>(static_cast<double>(xt[0]) - m0);
    x4[1] = static_cast<float>(static_cast<double>(xt[1]) - m1);
    x4[2] = static_cast<float>(static_cast<double>(xt[2]) - m2);
    bool counted;
    const float sf = pair_pclomp<HESS>(pt, x4, C, d2f, st.d1, acc, counted);
    acc[RED_SCORE] += static_cast<double>(sf);
    if (sf > 0.0f) atomicMax(&s_pts[slot].nearest_bits, __float_as_int(sf));
  } else {
    double xr[3], Cd[9];
#pragma unroll
    for (int a = 0; a < 3; ++a) xr[a] = static_cast<double>(xt[a]) - __ldg(c_mean + 3 * row + a);
#pragma unroll
    for (int i = 0; i < 9; ++i) Cd[i] = __ldg(c_icov + 9 * row + i);
    acc[RED_SCORE] += pair_svn<GN>(pt, xr, Cd, st.d1, st.d2, acc);
  }
}

__device__ __forceinline__ void prefetch_l1(const void* p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }

// Phase A for one chunk: lane `lane` handles point `idx` (if `have`), fills its slot, appends its pairs
// to the warp's queue at `tail`; returns the number of pairs the warp appended.
template <int FLAVOUR, int PROBES, bool HESS, bool GN, bool PAIRED = false, int QC = Q_CAP>
__device__ __forceinline__ int phase_a(const float4* __restrict__ src, int idx, bool have, int slot, int lane, int tail, const GridMeta& m,
                                       const VoxRec* __restrict__ recs, const HashSlot* __restrict__ table, const EvalState& st,
                                       PtSlot* __restrict__ s_pts, int* __restrict__ s_qrow, unsigned char* __restrict__ s_qslot,
                                       double& found, double& count) {
    int rows[PROBES];
    int nn = 0;
    bool filled = false;
    if (have) {
      const float4 p = __ldg(src + idx);
      // pcl::transformPointCloud (SSE2 Transformer association): m0*x + (m1*y + (m2*z + m3))
      float xt[3];
#pragma unroll
      for (int r = 0; r < 3; ++r) xt[r] = st.T[4 * r] * p.x + (st.T[4 * r + 1] * p.y + (st.T[4 * r + 2] * p.z + st.T[4 * r + 3]));
      bool point_ok = true;
      if (FLAVOUR == FLAVOUR_SVN) point_ok = isfinite(xt[0]) && isfinite(xt[1]) && isfinite(xt[2]);  // svn_ndt_impl.hpp:573
      if (point_ok) {
        uint32_t key[PROBES];
        if (FLAVOUR == FLAVOUR_PCLOMP || PROBES > 7) {  // (svn KDTREE: the 27 cells a radius-`leaf` ball around the query can reach)
          const int i0 = static_cast<int>(floorf(__fdiv_rn(xt[0], m.leaf)));
          const int i1 = static_cast<int>(floorf(__fdiv_rn(xt[1], m.leaf)));
          const int i2 = static_cast<int>(floorf(__fdiv_rn(xt[2], m.leaf)));
#pragma unroll
          for (int q = 0; q < PROBES; ++q) {
            if (PROBES == 27) key[q] = cell_key_pclomp(m, i0, i1, i2, c_off27[q][0], c_off27[q][1], c_off27[q][2]);
            else if (PROBES == 26) key[q] = cell_key_pclomp(m, i0, i1, i2, c_off26[q][0], c_off26[q][1], c_off26[q][2]);
            else key[q] = cell_key_pclomp(m, i0, i1, i2, c_off[q < 7 ? q : 0][0], c_off[q < 7 ? q : 0][1], c_off[q < 7 ? q : 0][2]);
          }
        } else {
#pragma unroll
          for (int q = 0; q < PROBES; ++q) {
            const float lw = m.leaf;
            const float qx = (q == 1) ? xt[0] + lw : ((q == 2) ? xt[0] - lw : xt[0]);
            const float qy = (q == 3) ? xt[1] + lw : ((q == 4) ? xt[1] - lw : xt[1]);
            const float qz = (q == 5) ? xt[2] + lw : ((q == 6) ? xt[2] - lw : xt[2]);
            key[q] = cell_key_svn(m, qx, qy, qz);
          }
        }
        hash_find_batch<PROBES>(table, m.hash_mask, key, rows);
        if (PROBES == 27) {
          // KDTREE (pclomp, and svn_ndt_impl.hpp:576-579 / voxel_grid_covariance_impl.hpp:383-436, 505-555): radiusSearch(x_trans, resolution_) over the cloud of float(leaf.mean_) (voxel_grid_covariance_omp.h:458-485,
          // _omp_impl.hpp:301-303); pcl::KdTreeFLANN hands float(radius * radius) to FLANN, L2_Simple<float> accumulates
          // dx*dx + dy*dy + dz*dz in float, RadiusResultSet keeps dist < r2 (strict)
          const float r2 = static_cast<float>(static_cast<double>(m.leaf) * static_cast<double>(m.leaf));
#pragma unroll
          for (int q = 0; q < PROBES; ++q) {
            if (FLAVOUR == FLAVOUR_PCLOMP && rows[q] < -1) rows[q] = -2 - rows[q];  // flagged leaf: still in the reference's kd-tree (voxel_finalize.cu)
            if (rows[q] >= 0) {
              const double* mu = recs[rows[q]].mean;
              const float dx = xt[0] - static_cast<float>(__ldg(mu)), dy = xt[1] - static_cast<float>(__ldg(mu + 1)), dz = xt[2] - static_cast<float>(__ldg(mu + 2));
              float d = dx * dx; d = d + dy * dy; d = d + dz * dz;
              if (!(d < r2)) rows[q] = -1;
            }
          }
        }
#pragma unroll
        for (int q = 0; q < PROBES; ++q) {
          if (rows[q] >= 0) {
            ++nn;
            if (FLAVOUR == FLAVOUR_PCLOMP) prefetch_l1(recs + rows[q]);
          }
        }
        // computePointDerivatives (ndt_omp_impl.hpp:542-587) once per point
        float4* sp = reinterpret_cast<float4*>(s_pts + slot);
        float xj[8];
#pragma unroll
        for (int r = 0; r < 8; ++r) xj[r] = dot3(st.jang[3 * r], st.jang[3 * r + 1], st.jang[3 * r + 2], p.x, p.y, p.z);
        sp[0] = make_float4(xt[0], xt[1], xt[2], 0.0f);  // .w = nearest_bits = 0
        if (PAIRED) {  // layout of pclomp_terms2
          sp[1] = make_float4(xj[0], xj[1], xj[2], xj[5]);
          sp[2] = make_float4(xj[3], xj[6], xj[4], xj[7]);
        } else {
          sp[1] = make_float4(xj[0], xj[1], xj[2], xj[3]);
          sp[2] = make_float4(xj[4], xj[5], xj[6], xj[7]);
        }
        if (HESS && !(FLAVOUR == FLAVOUR_SVN && GN)) {
          float xh[16];
#pragma unroll
          for (int r = 0; r < 15; ++r) xh[r] = dot3(st.hang[3 * r], st.hang[3 * r + 1], st.hang[3 * r + 2], p.x, p.y, p.z);
          xh[15] = 0.0f;
          if (PAIRED) {
            sp[3] = make_float4(xh[0], xh[2], xh[1], xh[3]);
            sp[4] = make_float4(xh[6], xh[9], xh[7], xh[10]);
            sp[5] = make_float4(xh[8], xh[11], xh[4], xh[5]);
            sp[6] = make_float4(xh[12], xh[13], xh[14], xh[15]);
          } else {
            sp[3] = make_float4(xh[0], xh[1], xh[2], xh[3]);
            sp[4] = make_float4(xh[4], xh[5], xh[6], xh[7]);
            sp[5] = make_float4(xh[8], xh[9], xh[10], xh[11]);
            sp[6] = make_float4(xh[12], xh[13], xh[14], xh[15]);
          }
        }
        filled = true;
        found += nn > 0 ? 1.0 : 0.0;
        count += static_cast<double>(nn);
      }
    }
    if (!filled) s_pts[slot].nearest_bits = 0;
    // warp exclusive scan of nn -> queue positions
    int incl = nn;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += t;
    }
    const int total = __shfl_sync(0xffffffffu, incl, 31);
    if (nn > 0) {
      int pos = tail + incl - nn;
#pragma unroll
      for (int q = 0; q < PROBES; ++q) {
        if (rows[q] >= 0) {
          s_qrow[pos & (QC - 1)] = rows[q];
          s_qslot[pos & (QC - 1)] = static_cast<unsigned char>(slot);
          ++pos;
        }
      }
    }
    return total;
}

template <int FLAVOUR, int PROBES, bool HESS, bool GN>
__global__ void __launch_bounds__(EVAL_THREADS, EVAL_MIN_BLOCKS) eval_kernel(const float4* __restrict__ src, int n, GridMeta m,
                                                                          const VoxRec* __restrict__ recs, const HashSlot* __restrict__ table,
                                                                          const double* __restrict__ c_mean, const double* __restrict__ c_icov,
                                                                          const EvalState* __restrict__ states, double* __restrict__ partials,
                                                                          double* __restrict__ results, unsigned int* __restrict__ tickets,
                                                                          int particle_offset, const SvnExchange xch) {
  // programmatic dependent launch: wait for the predecessor (the state kernel), then let the successor be scheduled while this grid
  // drains -- the Stein kernel of the device-resident SVN loop does not wait for this kernel to END, it acquires the record flags
  pdl_enter();
  if (FLAVOUR == FLAVOUR_SVN && xch.done && *xch.done) return;  // speculatively queued iteration after the align converged
  __shared__ EvalState st;
  constexpr int QC = (PROBES > 7) ? Q_CAP26 : Q_CAP;
  constexpr int GPC = (PROBES > 7) ? GROUPS_PER_CHUNK / 2 : GROUPS_PER_CHUNK;  // point groups per chunk
  __shared__ WarpScratchT<QC> s_warp[EVAL_WARPS];
  __shared__ double s_red[EVAL_WARPS][RED_PAD];
  __shared__ bool s_last;
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const int particle = blockIdx.y + particle_offset;   // row of partials / tickets / results
  // svn units (SvnExchange::vshards > 1): the row is a (particle, point shard) pair; states are indexed by particle
  const int vsh = (FLAVOUR == FLAVOUR_SVN) ? xch.vshards : 1;
  const int unit = (vsh > 1) ? xch.rank + particle * xch.world : 0;
  const int shard = (vsh > 1) ? unit % vsh : 0;
  const int state_idx = (vsh > 1) ? unit / vsh : particle;
  {
    const int words = sizeof(EvalState) / 4;
    const uint32_t* g = reinterpret_cast<const uint32_t*>(states + state_idx);
    uint32_t* s = reinterpret_cast<uint32_t*>(&st);
    for (int i = tid; i < words; i += blockDim.x) s[i] = g[i];
  }
  PtSlot* const s_pts = s_warp[w].pts;
  int* const s_qrow = s_warp[w].qrow;
  unsigned char* const s_qslot = s_warp[w].qslot;
  s_pts[lane].nearest_bits = 0;
  s_pts[32 + lane].nearest_bits = 0;
  __syncthreads();
  double acc[RED_PAD];
#pragma unroll
  for (int i = 0; i < RED_PAD; ++i) acc[i] = 0.0;
  const float d2f = static_cast<float>(st.d2);

  const int n_groups = (n + PT_GROUP - 1) / PT_GROUP;
  const int n_warps = gridDim.x * EVAL_WARPS * vsh;   // the shards of a particle deal the point groups out between them
  const int warp_global = (shard * gridDim.x + blockIdx.x) * EVAL_WARPS + w;
  // this warp's groups: warp_global + k * n_warps, k = 0 .. my_groups-1
  const int my_groups = (n_groups > warp_global) ? (n_groups - warp_global + n_warps - 1) / n_warps : 0;
  int head = 0, tail = 0;  // queue positions (monotonic; entry = pos & (QC - 1))
  int chunk = 0;
  for (int k0 = 0; k0 < my_groups; k0 += GPC, ++chunk) {
    // ---------------------------------------------------------------- phase A: one lane per point
    const int k = k0 + lane / PT_GROUP;
    const int idx = (warp_global + k * n_warps) * PT_GROUP + (lane % PT_GROUP);
    const bool have = (lane < GPC * PT_GROUP) && (k < my_groups) && (idx < n);
    const int slot = (chunk & 1) * 32 + lane;
    // the slot's previous occupant (two chunks ago) has had all its pairs evaluated: retire its NVTL term
    if (FLAVOUR == FLAVOUR_PCLOMP) acc[RED_NEAREST] += static_cast<double>(__int_as_float(s_pts[slot].nearest_bits));
    const int total = phase_a<FLAVOUR, PROBES, HESS, GN, false, QC>(src, idx, have, slot, lane, tail, m, recs, table, st, s_pts, s_qrow, s_qslot,
                                                         acc[RED_FOUND], acc[RED_COUNT]);
    const int must = tail;  // everything queued before this chunk must be consumed before its slots are reused
    tail += total;
    __syncwarp();
    // ---------------------------------------------------------------- phase B: one lane per pair
    while (tail - head >= 32 || head < must) {
      const int e = head + lane;
      if (e < tail) eval_pair<FLAVOUR, HESS, GN>(s_pts, s_qslot[e & (QC - 1)], s_qrow[e & (QC - 1)], st, recs, c_mean, c_icov, d2f, acc);
      head = min(head + 32, tail);
    }
    __syncwarp();
  }
  while (head < tail) {  // flush
    const int e = head + lane;
    if (e < tail) eval_pair<FLAVOUR, HESS, GN>(s_pts, s_qslot[e & (QC - 1)], s_qrow[e & (QC - 1)], st, recs, c_mean, c_icov, d2f, acc);
    head = min(head + 32, tail);
  }
  __syncwarp();
  if (FLAVOUR == FLAVOUR_PCLOMP) {
    acc[RED_NEAREST] += static_cast<double>(__int_as_float(s_pts[(chunk & 1) * 32 + lane].nearest_bits));
    acc[RED_NEAREST] += static_cast<double>(__int_as_float(s_pts[((chunk + 1) & 1) * 32 + lane].nearest_bits));
  }

  // ---- deterministic block reduction ----------------------------------------------------------
  // partials are component-major [particle][RED_N][gridDim.x] so that the final pass reads them coalesced
  block_reduce_store(acc, s_red, partials + static_cast<size_t>(particle) * gridDim.x * RED_N + blockIdx.x, gridDim.x);
  // ---- last block done: fixed-order reduction over all block partials --------------------------
  const int nblk = gridDim.x;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int t = atomicAdd(tickets + particle, 1u);
    s_last = (t == static_cast<unsigned int>(nblk) - 1u);
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  const double* pp = partials + static_cast<size_t>(particle) * nblk * RED_N;
#pragma unroll
  for (int i = 0; i < RED_PAD; ++i) acc[i] = 0.0;
  for (int b = threadIdx.x; b < nblk; b += EVAL_THREADS) {
#pragma unroll
    for (int i = 0; i < RED_N; ++i) acc[i] += __ldcg(pp + static_cast<size_t>(i) * nblk + b);
  }
  __syncthreads();
  block_reduce_store(acc, s_red, results + static_cast<size_t>(particle) * RED_N, 1);
  if (threadIdx.x == 0) tickets[particle] = 0;  // self-reset for the next launch
  if (FLAVOUR == FLAVOUR_SVN && xch.n_dest > 0) {
    // ---- fused exchange: this particle's record goes to every rank's window over NVLink (see SvnExchange) ----
    __shared__ int s_bad[2];
    if (threadIdx.x == 0) { s_bad[0] = 0; s_bad[1] = 0; }
    __syncthreads();  // also orders the stores of block_reduce_store before the reads below
    const int t = threadIdx.x;
    double v = 0.0;
    if (t < SVN_REC) {
      v = __ldcg(results + static_cast<size_t>(particle) * RED_N + t);   // RED_SCORE, RED_G, RED_H are laid out like the record
      if (vsh == 1) {
        if (t >= 7 && (t - 7) % 7 == 0) v += 1e-6;                       // H += 1e-6 I  (svn_ndt_impl.hpp:650-653)
        if (t >= 1 && t < 7 && !isfinite(v)) s_bad[0] = 1;
        if (t >= 7 && !isfinite(v)) s_bad[1] = 1;
      }
    }
    __syncthreads();
    const int k = xch.rank + (blockIdx.y + particle_offset) * xch.world;   // particle, or unit when the particles are sharded
    const size_t row = static_cast<size_t>(xch.parity) * xch.capacity + k;
    if (t < SVN_REC) {
      if (t >= 1 && t < 7 && s_bad[0]) v = 0.0;                          // non-finite gradient -> zero (:656-659)
      if (t >= 7 && s_bad[1]) v = ((t - 7) % 7 == 0) ? 1.0 : 0.0;        // non-finite Hessian -> identity (:660-663)
      for (int r = 0; r < xch.n_dest; ++r) xch.rec[r][row * SVN_REC + t] = v;
      __threadfence_system();
    }
    __syncthreads();
    if (t < xch.n_dest) asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(xch.flag[t] + row), "l"(xch.stamp) : "memory");
  }
}

// ------------------------------------------------------------------------------ the pclomp kernel
// Same two phases, plus a register diet: a lane does not keep 43 FP64 accumulators.  Each round the
// 32 lanes write their 43 float terms to a shared-memory tile and read it back TRANSPOSED: lane
// (q = lane >> 3, g = lane & 7) owns 6 of every 24 terms and adds the terms of the four lanes
// g, g+8, g+16, g+24 (widened to FP64, fixed order).  12 + 3 accumulators per lane instead of 46
// -> <= 128 registers -> 16 resident warps per SM, one 512-thread CTA per SM, 148 partial rows.
constexpr int PCL_THREADS = 512;
constexpr int PCL_WARPS = PCL_THREADS / 32;
constexpr int TR_STRIDE = 28;  // 24 terms + 4 pad words: conflict-free 16-byte stores and 8-byte transposed loads

struct WarpScratchP {
  PtSlot pts[SLOT_RING];
  int qrow[Q_CAP];
  unsigned char qslot[Q_CAP];
  float tr[32 * TR_STRIDE];
};

struct PclShared {
  WarpScratchP warp[PCL_WARPS];
  double red[PCL_WARPS][RED_PAD];
  bool last;
};

// Owner q of a pass holds terms {4q..4q+3} and {16+2q, 16+2q+1}: one 16-byte and one 8-byte load per source lane.
__device__ __forceinline__ void transposed_add(const float* __restrict__ s_tr, int q, int g, double* __restrict__ a) {
#pragma unroll
  for (int s = 0; s < 4; ++s) {
    const float* r = s_tr + (g + 8 * s) * TR_STRIDE;
    const float4 v0 = *reinterpret_cast<const float4*>(r + 4 * q);
    const float2 v1 = *reinterpret_cast<const float2*>(r + 16 + 2 * q);
    a[0] += static_cast<double>(v0.x); a[1] += static_cast<double>(v0.y);
    a[2] += static_cast<double>(v0.z); a[3] += static_cast<double>(v0.w);
    a[4] += static_cast<double>(v1.x); a[5] += static_cast<double>(v1.y);
  }
}

// term index (0..47) of accumulator i (0..5) of pass p for owner q
__device__ __forceinline__ int owner_term(int p, int q, int i) { return 24 * p + (i < 4 ? 4 * q + i : 16 + 2 * q + (i - 4)); }

// One round of phase B: every lane takes part (inactive lanes contribute +0 terms).
template <bool HESS>
__device__ __forceinline__ void pclomp_round(bool active, int slot, int row, PtSlot* __restrict__ s_pts, float* __restrict__ s_tr, int lane,
                                             const EvalState& st, const VoxRec* __restrict__ recs, float d2f, double* __restrict__ a) {
  float2 tp[22];
  if (!active) {
#pragma unroll
    for (int i = 0; i < 22; ++i) tp[i] = make_float2(0.0f, 0.0f);
  } else {
    const float4* sp = reinterpret_cast<const float4*>(s_pts + slot);
    const float4 pa = sp[0];
    const uint4* rp = reinterpret_cast<const uint4*>(recs + row);
    const uint4 q0 = __ldg(rp), q1 = __ldg(rp + 1), q2 = __ldg(rp + 2), q3 = __ldg(rp + 3);
    const double m0 = __hiloint2double(q0.y, q0.x), m1 = __hiloint2double(q0.w, q0.z), m2 = __hiloint2double(q1.y, q1.x);
    float C[9];
    C[0] = __uint_as_float(q1.z); C[1] = __uint_as_float(q1.w);
    C[2] = __uint_as_float(q2.x); C[3] = __uint_as_float(q2.y); C[4] = __uint_as_float(q2.z); C[5] = __uint_as_float(q2.w);
    C[6] = __uint_as_float(q3.x); C[7] = __uint_as_float(q3.y); C[8] = __uint_as_float(q3.z);
    float x4[3];
    x4[0] = static_cast<float>(static_cast<double>(pa.x) - m0);
    x4[1] = static_cast<float>(static_cast<double>(pa.y) - m1);
    x4[2] = static_cast<float>(static_cast<double>(pa.z) - m2);
    pclomp_terms2<HESS>(sp, x4, C, d2f, st.d1, st.one, tp);
    if (!HESS) {
#pragma unroll
      for (int i = 4; i < 22; ++i) tp[i] = make_float2(0.0f, 0.0f);
    }
    if (tp[0].x > 0.0f) atomicMax(&s_pts[slot].nearest_bits, __float_as_int(tp[0].x));
  }
  const int q = lane >> 3, g = lane & 7;
  float2* my = reinterpret_cast<float2*>(s_tr + lane * TR_STRIDE);
#pragma unroll
  for (int i = 0; i < 12; ++i) my[i] = tp[i];
  __syncwarp();
  transposed_add(s_tr, q, g, a);
  __syncwarp();
  if (HESS) {
#pragma unroll
    for (int i = 0; i < 10; ++i) my[i] = tp[12 + i];
    my[10] = make_float2(0.0f, 0.0f);
    my[11] = make_float2(0.0f, 0.0f);
    __syncwarp();
    transposed_add(s_tr, q, g, a + 6);
    __syncwarp();
  }
}

// RED_* index of word `w` (0..47) of the pair layout of pclomp_terms2, or -1 for padding.
__device__ __forceinline__ int pair_word_component(int w) {
  const int P = w >> 1, half = w & 1;
  if (P == 0) return half ? -1 : RED_SCORE;
  if (P < 4) return RED_G + 2 * (P - 1) + half;
  if (P >= 22) return -1;
  const int hp = P - 4, j = hp / 3, pp = hp % 3;
  return RED_H + 6 * (2 * pp + half) + j;
}

// The evaluation proper, shared by the one-shot kernel (state in the constant bank) and the alignment-session
// kernel (state in shared memory).  Every thread of the CTA calls it; it ends with a CTA-wide barrier-free tail
// (only the last CTA to take a ticket runs the final reduction), so callers that loop must __syncthreads() after it.
template <int PROBES, bool HESS>
__device__ __forceinline__ void pclomp_eval_body(PclShared& sh, const float4* __restrict__ src, int n, const GridMeta& m,
                                                 const VoxRec* __restrict__ recs, const HashSlot* __restrict__ table, const EvalState& st,
                                                 double* __restrict__ partials, double* __restrict__ results, unsigned int* __restrict__ tickets,
                                                 unsigned long long result_key, unsigned long long* __restrict__ dbg /* optional: 8 stamps per block */,
                                                 unsigned long long t_seen /* session: globaltimer when the command was seen, else 0 */,
                                                 double* __restrict__ s_final = nullptr /* session: shared-memory copy of the RED_N sums (last CTA) */) {
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  auto stamp = [&](int k) {
    if (dbg && tid == 0) { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); dbg[8 * blockIdx.x + k] = t; }
  };
  stamp(0);
  PtSlot* const s_pts = sh.warp[w].pts;
  int* const s_qrow = sh.warp[w].qrow;
  unsigned char* const s_qslot = sh.warp[w].qslot;
  float* const s_tr = sh.warp[w].tr;
  s_pts[lane].nearest_bits = 0;
  s_pts[32 + lane].nearest_bits = 0;
  __syncwarp();
  double a[12];
#pragma unroll
  for (int i = 0; i < 12; ++i) a[i] = 0.0;
  double nearest = 0.0, found = 0.0, count = 0.0;
  const float d2f = static_cast<float>(st.d2);

  const int n_groups = (n + PT_GROUP - 1) / PT_GROUP;
  const int n_warps = gridDim.x * PCL_WARPS;
  const int warp_global = blockIdx.x * PCL_WARPS + w;
  const int my_groups = (n_groups > warp_global) ? (n_groups - warp_global + n_warps - 1) / n_warps : 0;
  int head = 0, tail = 0, chunk = 0;
  for (int k0 = 0; k0 < my_groups; k0 += GROUPS_PER_CHUNK, ++chunk) {
    const int k = k0 + lane / PT_GROUP;
    const int idx = (warp_global + k * n_warps) * PT_GROUP + (lane % PT_GROUP);
    const bool have = (k < my_groups) && (idx < n);
    const int slot = (chunk & 1) * 32 + lane;
    nearest += static_cast<double>(__int_as_float(s_pts[slot].nearest_bits));  // retire the slot's previous occupant
    const int total = phase_a<FLAVOUR_PCLOMP, PROBES, HESS, false, true>(src, idx, have, slot, lane, tail, m, recs, table, st, s_pts, s_qrow, s_qslot,
                                                                   found, count);
    const int must = tail;
    tail += total;
    __syncwarp();
    while (tail - head >= 32 || head < must) {
      const int e = head + lane;
      const bool active = e < tail;
      pclomp_round<HESS>(active, active ? s_qslot[e & (Q_CAP - 1)] : 0, active ? s_qrow[e & (Q_CAP - 1)] : 0, s_pts, s_tr, lane, st, recs, d2f, a);
      head = min(head + 32, tail);
    }
    __syncwarp();
  }
  while (head < tail) {  // flush
    const int e = head + lane;
    const bool active = e < tail;
    pclomp_round<HESS>(active, active ? s_qslot[e & (Q_CAP - 1)] : 0, active ? s_qrow[e & (Q_CAP - 1)] : 0, s_pts, s_tr, lane, st, recs, d2f, a);
    head = min(head + 32, tail);
  }
  __syncwarp();
  nearest += static_cast<double>(__int_as_float(s_pts[(chunk & 1) * 32 + lane].nearest_bits));
  nearest += static_cast<double>(__int_as_float(s_pts[((chunk + 1) & 1) * 32 + lane].nearest_bits));

  stamp(1);
  // ---- warp reduction (fixed tree): the 8 lanes that share q hold partial sums of the same 12 terms
#pragma unroll
  for (int o = 1; o < 8; o <<= 1) {
#pragma unroll
    for (int i = 0; i < 12; ++i) a[i] += __shfl_xor_sync(0xffffffffu, a[i], o);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    nearest += __shfl_xor_sync(0xffffffffu, nearest, o);
    found += __shfl_xor_sync(0xffffffffu, found, o);
    count += __shfl_xor_sync(0xffffffffu, count, o);
  }
  if ((lane & 7) == 0) {
    const int q = lane >> 3;
#pragma unroll
    for (int i = 0; i < 6; ++i) {
      const int c0 = pair_word_component(owner_term(0, q, i)), c1 = pair_word_component(owner_term(1, q, i));
      if (c0 >= 0) sh.red[w][c0] = a[i];
      if (c1 >= 0) sh.red[w][c1] = a[6 + i];
    }
  }
  if (lane == 0) { sh.red[w][RED_NEAREST] = nearest; sh.red[w][RED_FOUND] = found; sh.red[w][RED_COUNT] = count; }
  __syncthreads();
  stamp(2);
  const int nblk = gridDim.x;
  if (tid < RED_N) {
    double v = sh.red[0][tid];
#pragma unroll
    for (int kk = 1; kk < PCL_WARPS; ++kk) v += sh.red[kk][tid];
    partials[static_cast<size_t>(tid) * nblk + blockIdx.x] = v;  // component-major
  }
  // ---- last block done: fixed-order reduction over the block partials --------------------------
  // Release/acquire on the ticket (ordered with the CTA barriers) instead of a membar in every thread.
  __syncthreads();
  if (tid == 0) {
    unsigned int tk;
    asm volatile("atom.acq_rel.gpu.global.add.u32 %0, [%1], 1;" : "=r"(tk) : "l"(tickets) : "memory");
    sh.last = (tk == static_cast<unsigned int>(nblk) - 1u);
  }
  __syncthreads();
  stamp(3);
  if (sh.last) {
    // reset the ticket BEFORE the results are published: the next evaluation (next launch, or the next command of a
    // session, which the host only sends after it has seen these results) finds it at zero
    if (tid == 0) { tickets[0] = 0; __threadfence(); }
    __syncthreads();
    if (tid < ((RED_N * 8 + 31) & ~31)) {  // whole warps: 8 lanes per component
      const int c = tid >> 3, l = tid & 7;
      const bool valid = c < RED_N;
      constexpr int MAXI = (148 + 7) / 8;
      double v[MAXI];
#pragma unroll
      for (int i = 0; i < MAXI; ++i) {
        const int b = l + 8 * i;
        v[i] = (valid && b < nblk) ? __ldcg(partials + static_cast<size_t>(c) * nblk + b) : 0.0;
      }
      double sum = v[0];
#pragma unroll
      for (int i = 1; i < MAXI; ++i) sum += v[i];
      if (valid) for (int b = l + 8 * MAXI; b < nblk; b += 8) sum += __ldcg(partials + static_cast<size_t>(c) * nblk + b);
#pragma unroll
      for (int o = 1; o < 8; o <<= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
      if (l == 0 && valid) {
#ifdef NDTB_RESULTS_PER_LANE   // A/B aid: the 46 owners store their slot themselves (92 scattered 8-byte writes over PCIe)
        const unsigned long long bits = static_cast<unsigned long long>(__double_as_longlong(sum));
        volatile unsigned long long* slot = reinterpret_cast<volatile unsigned long long*>(results) + 2 * c;
        slot[0] = bits;
        slot[1] = bits ^ result_key;
#else
        sh.red[0][c] = sum;   // the block partials of sh.red were consumed two barriers ago
#endif
        if (s_final) s_final[c] = sum;
      }
    }
#ifndef NDTB_RESULTS_PER_LANE
    // Self-validating result slots {value, value ^ key}: no system fence and no separate flag -- the host accepts slot c once
    // w0 ^ w1 == key (a stale half cannot satisfy it), see eval_pclomp().  ONE warp publishes them as two coalesced runs of
    // 16-byte stores (a handful of PCIe write packets instead of 92).
    __syncthreads();
    if (tid < 32) {
      for (int c = tid; c < RED_N; c += 32) {
        const unsigned long long bits = static_cast<unsigned long long>(__double_as_longlong(sh.red[0][c]));
        asm volatile("st.volatile.global.v2.u64 [%0], {%1, %2};" ::"l"(reinterpret_cast<unsigned long long*>(results) + 2 * c), "l"(bits), "l"(bits ^ result_key) : "memory");
      }
    }
#endif
    stamp(4);
    if (t_seen && tid == 0) {  // session: device-side duration of this evaluation (command seen -> results stored), slot RED_N
      unsigned long long t;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
      publish_slot(reinterpret_cast<volatile unsigned long long*>(results), RED_N, t - t_seen, result_key);
    }
  }
  stamp(5);
}

template <int PROBES, bool HESS>
__global__ void __launch_bounds__(PCL_THREADS, 1) eval_kernel_pclomp(const float4* __restrict__ src, int n, GridMeta m,
                                                                     const VoxRec* __restrict__ recs, const HashSlot* __restrict__ table,
                                                                     const __grid_constant__ EvalState st, double* __restrict__ partials,
                                                                     double* __restrict__ results, unsigned int* __restrict__ tickets,
                                                                     unsigned long long result_key,
                                                                     unsigned long long* __restrict__ dbg /* optional: 8 stamps per block */) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  PclShared& sh = *reinterpret_cast<PclShared*>(smem_raw);
  pclomp_eval_body<PROBES, HESS>(sh, src, n, m, recs, table, st, partials, results, tickets, result_key, dbg, 0ull);
}

// ---- alignment session: ONE resident launch serves every evaluation of a Newton loop -----------------------
// ndtb_align launches this kernel once (one CTA per SM, cooperative launch so that all CTAs are co-resident) and
// then sends one command per evaluation through a 1 KB mailbox in pinned host memory: 64 self-validating 16-byte
// units {word, word ^ session_key(seq)}.  Warp 0 of CTA 0 polls the mailbox over PCIe, copies the validated command
// to device memory and releases `seq` to the other CTAs, which spin on it in L2.  Results leave through the same
// host slots as the one-shot kernel.  This removes the per-iteration launch path (API call, front-end, CTA
// start-up) from the host <-> device ping-pong.  The kernel leaves by itself on a quit command or after
// `idle_timeout_ns` without a command (the host then falls back to one-shot launches), so it cannot hang the GPU.
struct PclSessionShared {
  PclShared sh;
  SessionCmd cmd;
  unsigned long long t_seen;
};
// 196 KB is a shared-memory carve-out step of the SM: one byte more (with the 1 KB the system reserves per CTA) and the L1 that
// caches the voxel-record gathers shrinks from 60 KB to 28 KB -- measured as +4 % on every evaluation.
static_assert(sizeof(PclSessionShared) + 1024 <= 196 * 1024, "the session kernel must stay inside the 196 KB carve-out");
// Device Newton step (SESSION_SPECULATE): scratch of warp 0 of the last CTA of an evaluation.  It lives in the transposition tile
// of warp 1, which is idle then: every warp of the CTA has left its main loop before the final reduction writes final_sums, and
// none re-enters it before warp 0 is back at the loop's barrier; the tile is rewritten before it is read in every round.
struct NewtonArea {
  double final_sums[RED_N];
  SessionCmd next;
  NewtonScratch newton;
};
static_assert(sizeof(NewtonArea) <= sizeof(WarpScratchP::tr) && offsetof(WarpScratchP, tr) % 16 == 0 && sizeof(WarpScratchP) % 16 == 0, "NewtonArea aliases warp 1's tile");

__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ unsigned long long ld_acquire_u64(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_u64(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// Speculation (SESSION_SPECULATE in a command): warp 0 of the LAST CTA of the evaluation takes the Newton step on the device
// (newton_step_warp: direction, normalisation, step clamp, pose -> T and angle tables), writes the next command into d_cmd and
// releases the next sequence number itself -- the other CTAs start the next evaluation ~3 us after the previous one ended
// instead of after a PCIe round trip through the host (~7 us).  The host still receives every result, repeats the step with
// its own code (glibc sin / cos) and compares the command the device built, word for word, with its own: equal -> nothing to
// do, the evaluation is already running; different (an ulp of sin / cos that survives the rounding to float: ~1e-8 of the
// steps) -> the host lets the kernel run out and repeats the evaluation with its own state, so results never depend on the
// device's libm.  When the step needs the host (ill-conditioned H -> SVD, zero slope, convergence) the warp declines and CTA 0
// polls the mailbox as usual.
template <int PROBES, bool SPEC>
__global__ void __launch_bounds__(PCL_THREADS, 1) align_session_kernel(const float4* __restrict__ src, int n, GridMeta m,
                                                                       const VoxRec* __restrict__ recs, const HashSlot* __restrict__ table,
                                                                       const ulonglong2* mailbox, unsigned long long* d_cmd,
                                                                       double* __restrict__ partials, double* __restrict__ results,
                                                                       ulonglong2* spec_record, unsigned int* __restrict__ tickets,
                                                                       unsigned long long seq0, unsigned long long idle_timeout_ns) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  PclSessionShared& ss = *reinterpret_cast<PclSessionShared*>(smem_raw);
  NewtonArea& na = *reinterpret_cast<NewtonArea*>(ss.sh.warp[1].tr);
  const int tid = threadIdx.x, lane = tid & 31;
  unsigned long long seq = seq0;
  bool spec_prev = false;  // the evaluation that has just finished speculates: its last CTA releases `seq` or declines it
  for (;;) {
    ++seq;
    if (blockIdx.x == 0 && tid < 32) {
      int released = 0;
      if (SPEC && spec_prev) {
        if (lane == 0) {  // no time-out: the last CTA does one or the other after a bounded amount of work of its own
          for (;;) {
            if (ld_acquire_u64(d_cmd + SESSION_SEQ_WORD) == seq) { released = 1; break; }
            if (ld_acquire_u64(d_cmd + SESSION_DECLINE_WORD) == seq) break;
          }
        }
        released = __shfl_sync(0xffffffffu, released, 0);
      }
      if (!released) {
        const unsigned long long key = session_key(seq);
        const unsigned long long t_wait = globaltimer_ns();
        unsigned long long w0 = 0, w1 = 0;
        bool quit = false;
        for (;;) {
#ifdef NDTB_RESULTS_PER_LANE
          const volatile ulonglong2* mb = mailbox;
          const unsigned long long a0 = mb[lane].x, a1 = mb[lane].y, b0 = mb[32 + lane].x, b1 = mb[32 + lane].y;
#else
          unsigned long long a0, a1, b0, b1;   // two 16-byte loads per lane: half the PCIe read requests of four 8-byte ones
          asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(a0), "=l"(a1) : "l"(mailbox + lane) : "memory");
          asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(b0), "=l"(b1) : "l"(mailbox + 32 + lane) : "memory");
#endif
          w0 = a0; w1 = b0;
          if (__all_sync(0xffffffffu, ((a0 ^ a1) == key) && ((b0 ^ b1) == key))) break;
          const bool late = globaltimer_ns() - t_wait > idle_timeout_ns;
          if (__shfl_sync(0xffffffffu, late ? 1 : 0, 0)) { quit = true; break; }
        }
        d_cmd[lane] = w0;
        d_cmd[32 + lane] = w1;
        __syncwarp();
        if (lane == 0) {
          if (quit) d_cmd[SESSION_FLAGS_WORD] = SESSION_QUIT;
          d_cmd[SESSION_TSEEN_WORD] = globaltimer_ns();
          __threadfence();
          st_release_u64(d_cmd + SESSION_SEQ_WORD, seq);
        }
      }
    }
    if (tid == 0) {
      while (ld_acquire_u64(d_cmd + SESSION_SEQ_WORD) != seq) {}
    }
    __syncthreads();
    if (tid < SESSION_CMD_WORDS) reinterpret_cast<unsigned long long*>(&ss.cmd)[tid] = __ldcg(d_cmd + tid);
    if (tid == 64) ss.t_seen = __ldcg(d_cmd + SESSION_TSEEN_WORD);
    __syncthreads();
    if (ss.cmd.flags & SESSION_QUIT) return;
    const bool spec = SPEC && (ss.cmd.flags & SESSION_SPECULATE) != 0;   // SPEC = false: the kernel without the step's code (the default)
    // with the step, results and records are double-buffered by the parity of key_seq (eval_api.cuh): the host may be reading
    // the previous evaluation's while this one is published
    double* const res = SPEC ? reinterpret_cast<double*>(spec_record + SESSION_SPEC_RESULT_UNIT0 + (ss.cmd.key_seq & 1ull) * SESSION_SPEC_RESULT_UNITS) : results;
    pclomp_eval_body<PROBES, true>(ss.sh, src, n, m, recs, table, ss.cmd.st, partials, res, tickets, ss.cmd.result_key, nullptr, ss.t_seen,
                                   spec ? na.final_sums : nullptr);
    __syncthreads();
#ifdef NDTB_NEWTON_TIMING
    const unsigned long long t_reduced = globaltimer_ns();
#endif
    if constexpr (SPEC) if (spec && ss.sh.last && tid < 32) {
      // flow control: the next evaluation's results land where those of evaluation key_seq - 1 are -- the host must have read
      // them.  The acknowledgement is fetched over PCIe now and looked at after the step (~1.5 us, hidden behind it).
      const volatile unsigned long long* ack = reinterpret_cast<const volatile unsigned long long*>(spec_record + SESSION_SPEC_ACK_UNIT);
      const unsigned long long ack0 = ack[0];   // one aligned 8-byte word: no tearing, and an old value only errs on the safe side
      NewtonParams prm;
      prm.step_max = ss.cmd.step_max; prm.step_min = ss.cmd.step_min; prm.trans_eps = ss.cmd.trans_eps; prm.max_iterations = ss.cmd.max_iterations; prm.pad = 0;
      NewtonCarry carry;
      carry.a_t = ss.cmd.a_t; carry.nr_iterations = static_cast<int>(ss.cmd.nr_iterations); carry.first = ss.cmd.first;
      int status = newton_step_warp(na.final_sums + RED_G, na.final_sums + RED_H, ss.cmd.st, carry, prm, &na.next.st, na.newton);
      if (status == NEWTON_CONTINUE && !(ack0 + 1 >= ss.cmd.key_seq)) status = NEWTON_FLOW;
      const unsigned long long next_key = result_key_of(ss.cmd.key_seq + 1);
      const unsigned long long* nw = reinterpret_cast<const unsigned long long*>(&na.next);
      if (status == NEWTON_CONTINUE) {
        if (lane == 0) {
          na.next.result_key = next_key; na.next.flags = SESSION_SPECULATE;
          na.next.step_max = prm.step_max; na.next.step_min = prm.step_min; na.next.trans_eps = prm.trans_eps; na.next.max_iterations = prm.max_iterations;
          na.next.first = 0; na.next.a_t = carry.a_t; na.next.nr_iterations = static_cast<unsigned long long>(carry.nr_iterations);
          na.next.key_seq = ss.cmd.key_seq + 1;
        }
        __syncwarp();
        for (int i = lane; i < SESSION_CMD_WORDS; i += 32) d_cmd[i] = nw[i];
        __syncwarp();
        if (lane == 0) {
          d_cmd[SESSION_TSEEN_WORD] = globaltimer_ns();
          __threadfence();
          st_release_u64(d_cmd + SESSION_SEQ_WORD, seq + 1);
        }
      } else if (lane == 0) {
        st_release_u64(d_cmd + SESSION_DECLINE_WORD, seq + 1);
      }
#ifdef NDTB_NEWTON_TIMING
      if (lane == 0 && status == NEWTON_CONTINUE) {   // d_cmd[67..]: summed phase durations + count, printed by session_end
        const unsigned long long t_end = globaltimer_ns();
        for (int k = 0; k < 6; ++k) d_cmd[67 + k] += na.newton.stamp[k + 1] - na.newton.stamp[k];
        d_cmd[73] += t_end - na.newton.stamp[6];
        d_cmd[74] += na.newton.stamp[0] - t_reduced;
        d_cmd[75] += 1;
      }
#endif
      // the record the host compares with its own step (self-validating units, off the critical path)
      volatile ulonglong2* rec = spec_record + ((ss.cmd.key_seq + 1) & 1ull) * SESSION_UNITS;
      if (status == NEWTON_CONTINUE)
        for (int i = lane; i < SESSION_CMD_WORDS; i += 32) { rec[i].x = nw[i]; rec[i].y = nw[i] ^ next_key; }
      if (lane == 0) {
        const unsigned long long sw = ((seq + 1) << 8) | static_cast<unsigned long long>(status);
        rec[SESSION_SPEC_STATUS_UNIT].x = sw;
        rec[SESSION_SPEC_STATUS_UNIT].y = sw ^ next_key;
      }
    }
    spec_prev = spec;
    __syncthreads();
  }
}

template <int PROBES, bool SPEC>
static cudaError_t launch_session_t(const float4* d_src, int n, const GridMeta& meta, const VoxRec* recs, const HashSlot* table,
                                    const ulonglong2* mailbox, unsigned long long* d_cmd, const EvalBuffers& b, double* results,
                                    ulonglong2* spec_record, unsigned long long seq0, unsigned long long idle_timeout_ns, cudaStream_t stream) {
  static bool configured[MAX_DEVICES] = {};
  if (needs_func_setup(configured)) {
    cudaError_t e = cudaFuncSetAttribute(align_session_kernel<PROBES, SPEC>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(sizeof(PclSessionShared)));
    if (e != cudaSuccess) return e;
  }
  GridMeta m = meta;
  double* partials = b.d_partials;
  unsigned int* tickets = b.d_tickets;
  void* args[] = {&d_src, &n, &m, &recs, &table, &mailbox, &d_cmd, &partials, &results, &spec_record, &tickets, &seq0, &idle_timeout_ns};
  // cooperative launch: fails (instead of deadlocking on the inter-CTA spin) if the CTAs cannot all be resident at once
  return cudaLaunchCooperativeKernel(reinterpret_cast<const void*>(align_session_kernel<PROBES, SPEC>), dim3(eval_blocks_pclomp(n, b.free_sms)), dim3(PCL_THREADS), args,
                                     sizeof(PclSessionShared), stream);
}

cudaError_t launch_align_session(int n_probe, const float4* d_src, int n, const GridMeta& meta, const VoxRec* recs, const HashSlot* table,
                                 const ulonglong2* mailbox, unsigned long long* d_cmd, const EvalBuffers& b, double* results,
                                 ulonglong2* spec_record, unsigned long long seq0, unsigned long long idle_timeout_ns, cudaStream_t stream) {
  // spec_record == nullptr: the kernel without the device Newton step (it ignores SESSION_SPECULATE)
  if (spec_record)
    return n_probe == 7 ? launch_session_t<7, true>(d_src, n, meta, recs, table, mailbox, d_cmd, b, results, spec_record, seq0, idle_timeout_ns, stream)
                        : launch_session_t<1, true>(d_src, n, meta, recs, table, mailbox, d_cmd, b, results, spec_record, seq0, idle_timeout_ns, stream);
  return n_probe == 7 ? launch_session_t<7, false>(d_src, n, meta, recs, table, mailbox, d_cmd, b, results, spec_record, seq0, idle_timeout_ns, stream)
                      : launch_session_t<1, false>(d_src, n, meta, recs, table, mailbox, d_cmd, b, results, spec_record, seq0, idle_timeout_ns, stream);
}

// Generic (svn) kernel: one resident wave of 148 SMs x EVAL_MIN_BLOCKS CTAs (register-limited); the cloud is dealt to the warps round-robin.
int eval_blocks_for(int n) {
  const int b = (n + EVAL_THREADS - 1) / EVAL_THREADS;
  return b < 148 * EVAL_MIN_BLOCKS ? b : 148 * EVAL_MIN_BLOCKS;
}

int eval_blocks_pclomp(int n, int free_sms) {
  const int per_block = PCL_WARPS * 32;  // at least one full chunk per warp
  const int b = (n + per_block - 1) / per_block;
  const int cap = std::max(1, 148 - std::max(0, free_sms));
  return b < 1 ? 1 : (b < cap ? b : cap);
}

template <int PROBES, bool HESS>
static cudaError_t launch_pclomp_t(const float4* d_src, int n, const GridMeta& meta, const VoxRec* recs, const HashSlot* table, const EvalState& st,
                                   const EvalBuffers& b, double* results, unsigned long long result_key,
                                   cudaStream_t stream, unsigned long long* dbg) {
  static bool configured[MAX_DEVICES] = {};  // one handle per caller thread, but the attribute is per function and device: benign if set twice
  if (needs_func_setup(configured)) {
    cudaError_t e = cudaFuncSetAttribute(eval_kernel_pclomp<PROBES, HESS>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(sizeof(PclShared)));
    if (e != cudaSuccess) return e;
  }
  eval_kernel_pclomp<PROBES, HESS><<<eval_blocks_pclomp(n, b.free_sms), PCL_THREADS, sizeof(PclShared), stream>>>(d_src, n, meta, recs, table, st, b.d_partials,
                                                                                                  results, b.d_tickets, result_key, dbg);
  return cudaGetLastError();
}

cudaError_t launch_eval_pclomp(int n_probe, bool compute_hessian, const float4* d_src, int n, const GridMeta& meta, const VoxRec* recs,
                               const HashSlot* table, const EvalState& st, const EvalBuffers& b, double* results, unsigned long long result_key,
                               cudaStream_t stream, unsigned long long* dbg) {
  if (n_probe == 7) {
    return compute_hessian ? launch_pclomp_t<7, true>(d_src, n, meta, recs, table, st, b, results, result_key, stream, dbg)
                           : launch_pclomp_t<7, false>(d_src, n, meta, recs, table, st, b, results, result_key, stream, dbg);
  }
  return compute_hessian ? launch_pclomp_t<1, true>(d_src, n, meta, recs, table, st, b, results, result_key, stream, dbg)
                         : launch_pclomp_t<1, false>(d_src, n, meta, recs, table, st, b, results, result_key, stream, dbg);
}

template <int FLAVOUR, int PROBES, bool HESS, bool GN>
static void launch_t(const float4* d_src, int n, const GridMeta& meta, const VoxRec* recs, const HashSlot* table, const double* c_mean,
                     const double* c_icov, const EvalBuffers& b, int n_particles, int particle_offset, cudaStream_t stream, const SvnExchange& xch) {
  const int vsh = (FLAVOUR == FLAVOUR_SVN && xch.vshards > 1) ? xch.vshards : 1;
  dim3 grid((vsh > 1 && xch.blocks_per_unit > 0) ? xch.blocks_per_unit : std::max(1, eval_blocks_for(n) / vsh), n_particles);
  launch_pdl(eval_kernel<FLAVOUR, PROBES, HESS, GN>, grid, dim3(EVAL_THREADS), 0, stream, d_src, n, meta, recs, table, c_mean, c_icov,
             static_cast<const EvalState*>(b.d_states), b.d_partials, b.d_results, b.d_tickets, particle_offset, xch);
}

void launch_eval(int flavour, int n_probe, bool compute_hessian, bool gn, const float4* d_src, int n, const GridMeta& meta,
                 const VoxRec* recs, const HashSlot* table, const double* c_mean, const double* c_icov, const EvalBuffers& b,
                 int n_particles, int particle_offset, cudaStream_t stream, const SvnExchange* xch_in) {
  const SvnExchange xch = xch_in ? *xch_in : SvnExchange();
#define ARGS d_src, n, meta, recs, table, c_mean, c_icov, b, n_particles, particle_offset, stream, xch
  if (flavour == FLAVOUR_PCLOMP) {
    // DIRECT1 / DIRECT7 evaluations go through launch_eval_pclomp; DIRECT26 and KDTREE (27 cells + radius test), off the hot path, use this generic kernel
    if (n_probe == 26) { if (compute_hessian) launch_t<FLAVOUR_PCLOMP, 26, true, false>(ARGS); else launch_t<FLAVOUR_PCLOMP, 26, false, false>(ARGS); }
    if (n_probe == 27) { if (compute_hessian) launch_t<FLAVOUR_PCLOMP, 27, true, false>(ARGS); else launch_t<FLAVOUR_PCLOMP, 27, false, false>(ARGS); }
    return;
  } else {
    if (n_probe == 7) { if (gn) launch_t<FLAVOUR_SVN, 7, true, true>(ARGS); else launch_t<FLAVOUR_SVN, 7, true, false>(ARGS); }
    else if (n_probe == 27) { if (gn) launch_t<FLAVOUR_SVN, 27, true, true>(ARGS); else launch_t<FLAVOUR_SVN, 27, true, false>(ARGS); }
    else { if (gn) launch_t<FLAVOUR_SVN, 1, true, true>(ARGS); else launch_t<FLAVOUR_SVN, 1, true, false>(ARGS); }
  }
#undef ARGS
}

}  // namespace ndtb

// ---- allocation-free evaluation (ndtb_compute_derivatives_static, include/ndtb.h) ------------------------------------
// The one-shot DIRECT1 / DIRECT7 kernel above needs scratch memory: the per-block partial sums, the last-block-done
// ticket and -- when the result is not written to pinned host memory -- the 46 result slots.  ndtb_compute_derivatives
// takes them from buffers that ensure_eval_buffers() allocates; callers that may not allocate (code under CUDA graph
// capture, tight serving loops) get the same kernel over a scratch that is part of the library image: a file-scope
// __device__ array of STATIC_SCRATCH_BYTES, held by every GPU that loads the library until the process exits.
// Every call on a device shares the array: calls that overlap on that GPU race, and the array keeps the previous call's
// contents -- the ticket and the result slots are cleared at the start of every call, the partial sums are all
// rewritten before they are read.

namespace ndtb {

constexpr int STATIC_MAX_BLOCKS = 148;  // eval_blocks_pclomp() never launches more
struct StaticEvalScratch {
  double partials[RED_N * STATIC_MAX_BLOCKS];   // component-major block partials
  unsigned long long slots[2 * (RED_N + 1)];    // self-validating result slots {bits, bits ^ key}
  unsigned int ticket[4];
};
constexpr size_t STATIC_SCRATCH_BYTES = (sizeof(StaticEvalScratch) + 255) & ~static_cast<size_t>(255);
static __device__ __align__(256) unsigned char g_static_eval_scratch[STATIC_SCRATCH_BYTES];

// In host code `g_static_eval_scratch` is only a stand-in for the symbol: the device address comes from
// cudaGetSymbolAddress, asked once per device with that device current.
constexpr int STATIC_MAX_DEVICES = 64;
static StaticEvalScratch* g_static_scratch_ptr[STATIC_MAX_DEVICES] = {};

static cudaError_t static_scratch_for_current_device(int device, StaticEvalScratch** out) {
  if (device < 0 || device >= STATIC_MAX_DEVICES) return cudaErrorInvalidDevice;
  if (!g_static_scratch_ptr[device]) {
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_static_eval_scratch);
    if (e != cudaSuccess) return e;
    g_static_scratch_ptr[device] = static_cast<StaticEvalScratch*>(p);
  }
  *out = g_static_scratch_ptr[device];
  return cudaSuccess;
}

}  // namespace ndtb

extern "C" int ndtb_compute_derivatives_static(ndtb_handle h, const void* d_source_xyzw, size_t n_source, const double p[6], const float* T16,
                                               int compute_hessian, double* score, double gradient[6], double hessian[36], double* nvtl) {
  using namespace ndtb;
  if (!h || !p || !score || !gradient || !hessian) return NDTB_ERR_INVALID;
  if (cudaSetDevice(h->device) != cudaSuccess) return set_error(h, NDTB_ERR_CUDA, "cudaSetDevice failed");
  if (!h->grid.built) return set_error(h, NDTB_ERR_NOT_READY, "the target must be set before ndtb_compute_derivatives_static");
  if (h->params.grid_flavour != NDTB_FLAVOUR_PCLOMP) return set_error(h, NDTB_ERR_INVALID, "pclomp entry point on an svn-flavour handle");
  if (h->params.search_method != NDTB_DIRECT1 && h->params.search_method != NDTB_DIRECT7)
    return set_error(h, NDTB_ERR_INVALID, "ndtb_compute_derivatives_static serves DIRECT1 and DIRECT7 only");
  const float4* d_src = static_cast<const float4*>(d_source_xyzw);
  if (!d_src) { d_src = h->d_source; n_source = h->n_source; }
  if (!d_src || n_source == 0) return set_error(h, NDTB_ERR_NOT_READY, "no source cloud: pass a device pointer or call ndtb_set_source first");
  if (n_source > static_cast<size_t>(NDTB_STATIC_MAX_POINTS))
    return set_error(h, NDTB_ERR_INVALID, "ndtb_compute_derivatives_static: more than NDTB_STATIC_MAX_POINTS (4194304) source points");
  if (h->session_open) return set_error(h, NDTB_ERR_NOT_READY, "an alignment session is running on this handle");
  const int n = static_cast<int>(n_source);
  if (eval_blocks_pclomp(n) > STATIC_MAX_BLOCKS) return set_error(h, NDTB_ERR_INVALID, "ndtb_compute_derivatives_static: grid exceeds the static scratch");
  StaticEvalScratch* sc = nullptr;
  cudaError_t e = static_scratch_for_current_device(h->device, &sc);
  if (e != cudaSuccess) { h->last_error = std::string("cudaGetSymbolAddress(static scratch): ") + cudaGetErrorString(e); return NDTB_ERR_CUDA; }
  float T[16];
  if (T16) std::memcpy(T, T16, sizeof(T)); else pose_to_T(p, T);
  EvalState st;
  fill_eval_state(h, st, p, T, 10e-5, false);
  // what must start at zero: the ticket (a call that failed half-way may have left it counted up) and the result slots
  e = cudaMemsetAsync(sc->slots, 0, sizeof(sc->slots) + sizeof(sc->ticket), h->stream);
  if (e != cudaSuccess) { h->last_error = std::string("static scratch clear: ") + cudaGetErrorString(e); return NDTB_ERR_CUDA; }
  EvalBuffers b;  // views into the static scratch, nothing owned
  b.d_partials = sc->partials;
  b.d_tickets = sc->ticket;
  b.max_particles = 1; b.max_blocks = STATIC_MAX_BLOCKS;
  const unsigned long long key = 0x9E3779B97F4A7C15ull * static_cast<unsigned long long>(++h->flag_seq) | 1ull;
  const int n_probe = (h->params.search_method == NDTB_DIRECT1) ? 1 : 7;
  e = launch_eval_pclomp(n_probe, compute_hessian != 0, d_src, n, h->grid.meta, h->grid.recs, h->grid.table, st, b,
                         reinterpret_cast<double*>(sc->slots), key, h->stream, nullptr);
  if (e != cudaSuccess) { h->last_error = std::string("static eval kernel launch: ") + cudaGetErrorString(e); return NDTB_ERR_CUDA; }
  h->launches += 1;
  unsigned long long words[2 * RED_N];
  e = cudaMemcpyAsync(words, sc->slots, sizeof(words), cudaMemcpyDeviceToHost, h->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
  if (e != cudaSuccess) { h->last_error = std::string("static eval kernel: ") + cudaGetErrorString(e); return NDTB_ERR_CUDA; }
  double r[RED_N];
  for (int c = 0; c < RED_N; ++c) {
    if ((words[2 * c] ^ words[2 * c + 1]) != key) return set_error(h, NDTB_ERR_CUDA, "static eval kernel finished without publishing its result");
    std::memcpy(&r[c], &words[2 * c], 8);
  }
  finish_eval(h, n, r, p, score, gradient, hessian, nvtl);
  return NDTB_OK;
}
