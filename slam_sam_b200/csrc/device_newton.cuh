// The Newton step of ndtb_align (ndt_omp_impl.hpp:128-190, no line search) as a WARP-cooperative device function: direction
// by elimination with complete pivoting + one refinement step (gecp6_solve of newton_math.cuh, the SAME operations in the same
// order, spread over the lanes), normalisation, descent-direction flip, step clamp (computeStepLengthMT's first trial, :915-990),
// pose -> float 4x4 and the angle tables of the next evaluation.  One thread would need ~10 us for this chain (as long as the PCIe
// round trip to the host it replaces); a warp needs ~3 us.  Every arithmetic result is bit-identical to the host code: each
// matrix entry sees the same operations, sums keep their order, only independent entries are computed side by side.  libm: the
// float sin / cos of T restate glibc's sincosf operation by operation (nm_sincosf_glibc, newton_math.cuh: bit-identical); the
// FP64 sin / cos of the angle tables come from CUDA's routines, an ulp from glibc's for some arguments, which survives the
// rounding to float of a table entry in ~1e-8 of the cases -- the host compares every step (eval_pclomp) and takes over then.
#pragma once
#include "ndtb_internal.cuh"
#include "newton_math.cuh"

namespace ndtb {

struct NewtonParams {
  double step_max, step_min, trans_eps;
  int max_iterations;
  int pad;
};
struct NewtonCarry {       // travels from one evaluation of an align to the next
  double a_t;              // step length of the iteration whose evaluation has just finished
  int nr_iterations;
  int first;               // 1: the evaluation that has just finished is the one at the initial guess
};
constexpr int NEWTON_CONTINUE = 0;   // st_out holds the state of the next evaluation
constexpr int NEWTON_CONVERGED = 1;  // the loop's convergence test fired: no further evaluation
constexpr int NEWTON_FLOW = 3;       // (session kernel) the host has not yet read the results two evaluations back: wait for it
constexpr int NEWTON_HOST = 2;       // the host decides: ill-conditioned H (SVD), zero or NaN step, zero slope

#ifdef NDTB_NEWTON_TIMING   // development aid: globaltimer stamps of the phases of one step (lane 0), read by the session kernel
#define NEWTON_STAMP(k) do { if (lane == 0) { unsigned long long t_; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_)); s.stamp[k] = t_; } } while (0)
#else
#define NEWTON_STAMP(k) do { } while (0)
#endif

struct NewtonScratch {     // shared memory of the warp
  double M[36];
  double rhs[6], y[6], sol[6], dp[6], x_t[6];
  double sc[12];           // sin / cos of the three angles: double argument (tables) and float argument (T)
  int rowperm[6], colperm[6];
  int status;
#ifdef NDTB_NEWTON_TIMING
  unsigned long long stamp[8];
#endif
};

// All 32 lanes of one warp call this with the same arguments.  g / H: gradient and Hessian of the evaluation at st_in.p.
__device__ inline int newton_step_warp(const double* __restrict__ g, const double* __restrict__ H, const EvalState& st_in, NewtonCarry& carry,
                                       const NewtonParams& prm, EvalState* __restrict__ st_out, NewtonScratch& s) {
  const int lane = threadIdx.x & 31;
  NEWTON_STAMP(0);
  // ---- finish the iteration whose evaluation this was (ndt_omp_impl.hpp:176-186)
  if (!carry.first) {
    if (carry.nr_iterations > prm.max_iterations || (carry.nr_iterations && (fabs(carry.a_t) < prm.trans_eps))) return NEWTON_CONVERGED;
    carry.nr_iterations += 1;
  }
  carry.first = 0;
  // ---- gecp6_solve(H, -g, dp, 1e-6), entry (i, j) of the 6x6 on lane 6 i + j (lanes 0..3 also own entries 32..35)
  for (int e = lane; e < 36; e += 32) s.M[e] = H[e];
  if (lane < 6) { s.rowperm[lane] = lane; s.colperm[lane] = lane; s.rhs[lane] = -g[lane]; s.sol[lane] = 0.0; }
  if (lane == 0) s.status = NEWTON_CONTINUE;
  __syncwarp();
  double first_pivot = 0.0;
  for (int k = 0; k < 6; ++k) {
    // complete pivoting: the largest |M(i, j)|, i, j >= k; ties -> the first in row-major order (the serial loop's strict '>')
    double best = -1.0;
    int bidx = 64;
    for (int e = lane; e < 36; e += 32) {
      const int i = e / 6, j = e - 6 * i;
      if (i >= k && j >= k) {
        const double a = fabs(s.M[e]);
        if (a > best) { best = a; bidx = e; }
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const double ob = __shfl_xor_sync(0xffffffffu, best, o);
      const int oi = __shfl_xor_sync(0xffffffffu, bidx, o);
      if (ob > best || (ob == best && oi < bidx)) { best = ob; bidx = oi; }
    }
    if (k == 0) first_pivot = best;
    if (!(best >= first_pivot * 1e-6) || !(best > 0.0) || best > DBL_MAX) return NEWTON_HOST;  // also catches NaN (best stays -1)
    const int pr = bidx / 6, pc = bidx - 6 * pr;
    if (pr != k) {
      if (lane < 6) { const double t = s.M[6 * k + lane]; s.M[6 * k + lane] = s.M[6 * pr + lane]; s.M[6 * pr + lane] = t; }
      if (lane == 0) { const int t = s.rowperm[k]; s.rowperm[k] = s.rowperm[pr]; s.rowperm[pr] = t; }
      __syncwarp();
    }
    if (pc != k) {
      if (lane < 6) { const double t = s.M[6 * lane + k]; s.M[6 * lane + k] = s.M[6 * lane + pc]; s.M[6 * lane + pc] = t; }
      if (lane == 0) { const int t = s.colperm[k]; s.colperm[k] = s.colperm[pc]; s.colperm[pc] = t; }
      __syncwarp();
    }
    const double inv = 1.0 / s.M[7 * k];
    double upd[2] = {0.0, 0.0};
    int n_upd = 0;
    for (int e = lane; e < 36; e += 32) {
      const int i = e / 6, j = e - 6 * i;
      if (i > k && j >= k) {
        const double f = s.M[6 * i + k] * inv;
        upd[n_upd] = (j == k) ? f : (s.M[e] - f * s.M[6 * k + j]);
      }
      ++n_upd;
    }
    __syncwarp();
    n_upd = 0;
    for (int e = lane; e < 36; e += 32) {
      const int i = e / 6, j = e - 6 * i;
      if (i > k && j >= k) s.M[e] = upd[n_upd];
      ++n_upd;
    }
    __syncwarp();
  }
  NEWTON_STAMP(1);
  // the two triangular solves + one refinement step: short dependent chains, lane 0; the residual on six lanes
  for (int pass = 0; pass < 2; ++pass) {
    if (lane == 0) {
      double y[6];
      for (int i = 0; i < 6; ++i) {
        double sum = s.rhs[s.rowperm[i]];
        for (int j = 0; j < i; ++j) sum = sum - s.M[6 * i + j] * y[j];
        y[i] = sum;
      }
      for (int i = 5; i >= 0; --i) {
        double sum = y[i];
        for (int j = i + 1; j < 6; ++j) sum = sum - s.M[6 * i + j] * y[j];
        y[i] = sum / s.M[7 * i];
      }
      for (int i = 0; i < 6; ++i) s.sol[s.colperm[i]] = s.sol[s.colperm[i]] + y[i];
    }
    __syncwarp();
    if (pass == 0) {
      double r = 0.0;
      if (lane < 6) {
        r = -g[lane];
        for (int j = 0; j < 6; ++j) r = r - H[6 * lane + j] * s.sol[j];
      }
      __syncwarp();
      if (lane < 6) s.rhs[lane] = r;
      __syncwarp();
    }
  }
  NEWTON_STAMP(2);
  // ---- the step (lane 0: a handful of dependent operations)
  if (lane == 0) {
    int status = NEWTON_CONTINUE;
    double dp[6];
    bool finite = true;
    for (int i = 0; i < 6; ++i) { dp[i] = s.sol[i]; if (!(fabs(dp[i]) <= DBL_MAX)) finite = false; }
    if (!finite) status = NEWTON_HOST;   // gecp6_solve returns false -> the host takes the SVD
    double nrm2 = 0;
    for (int i = 0; i < 6; ++i) nrm2 += dp[i] * dp[i];
    const double nrm = sqrt(nrm2);
    if (nrm == 0 || nrm != nrm) status = NEWTON_HOST;
    if (status == NEWTON_CONTINUE) {
      for (int i = 0; i < 6; ++i) dp[i] = dp[i] / nrm;
      double d_phi_0 = 0;
      for (int i = 0; i < 6; ++i) d_phi_0 += g[i] * dp[i];
      d_phi_0 = -d_phi_0;
      if (d_phi_0 >= 0 && d_phi_0 == 0) status = NEWTON_HOST;   // zero slope: a_t = 0, no evaluation (:944-950)
      else {
        if (d_phi_0 >= 0) { for (int i = 0; i < 6; ++i) dp[i] *= -1; }
        double a_t = nrm;
        a_t = (a_t < prm.step_max) ? a_t : prm.step_max;   // std::min(a_t, step_max)
        a_t = (a_t < prm.step_min) ? prm.step_min : a_t;   // std::max(a_t, step_min)
        for (int i = 0; i < 6; ++i) s.x_t[i] = st_in.p[i] + dp[i] * a_t;
        carry.a_t = a_t;
      }
    }
    s.status = status;
  }
  __syncwarp();
  const int status = s.status;
  carry.a_t = __shfl_sync(0xffffffffu, carry.a_t, 0);
  if (status != NEWTON_CONTINUE) return status;
  NEWTON_STAMP(3);
  // ---- state of the next evaluation: sin / cos side by side -- lanes 0..5 the double arguments (angle tables), lanes 6..8 the
  // float arguments (T) through the restatement of the host's sincosf (newton_math.cuh); an angle it does not cover -> host
  bool covered = true;
  if (lane < 6) {
    const int a = lane % 3;
    const double x = s.x_t[3 + a];
    s.sc[lane] = (lane >= 3) ? cos(x) : sin(x);
  } else if (lane < 9) {
    const int a = lane - 6;
    float sv = 0.0f, cv = 1.0f;
    covered = nm_sincosf_glibc(static_cast<float>(s.x_t[3 + a]), &sv, &cv);
    s.sc[6 + a] = static_cast<double>(sv);
    s.sc[9 + a] = static_cast<double>(cv);
  }
  if (!__all_sync(0xffffffffu, covered)) return NEWTON_HOST;
  __syncwarp();
  NEWTON_STAMP(4);
  if (lane == 0) {
    // angle_tables(x_t, 10e-5): the snap of ndt_omp_impl.hpp:401-437
    const double snap = 10e-5;
    double sx = 0.0, cx = 1.0, sy = 0.0, cy = 1.0, sz = 0.0, cz = 1.0;
    if (!(fabs(s.x_t[3]) < snap)) { sx = s.sc[0]; cx = s.sc[3]; }
    if (!(fabs(s.x_t[4]) < snap)) { sy = s.sc[1]; cy = s.sc[4]; }
    if (!(fabs(s.x_t[5]) < snap)) { sz = s.sc[2]; cz = s.sc[5]; }
    s.sc[0] = sx; s.sc[1] = sy; s.sc[2] = sz; s.sc[3] = cx; s.sc[4] = cy; s.sc[5] = cz;
  }
  __syncwarp();
  {
    const double sx = s.sc[0], sy = s.sc[1], sz = s.sc[2], cx = s.sc[3], cy = s.sc[4], cz = s.sc[5];
    // the 24 + 45 entries of angle_tables (newton_math.cuh), entry e on lane e % 32
    for (int e = lane; e < 69; e += 32) {
      double v = 0.0;
      switch (e) {
        case 0: v = (-sx * sz + cx * sy * cz); break;  case 1: v = (-sx * cz - cx * sy * sz); break;  case 2: v = (-cx * cy); break;
        case 3: v = (cx * sz + sx * sy * cz); break;   case 4: v = (cx * cz - sx * sy * sz); break;   case 5: v = (-sx * cy); break;
        case 6: v = (-sy * cz); break;                 case 7: v = sy * sz; break;                    case 8: v = cy; break;
        case 9: v = sx * cy * cz; break;               case 10: v = (-sx * cy * sz); break;           case 11: v = sx * sy; break;
        case 12: v = (-cx * cy * cz); break;           case 13: v = cx * cy * sz; break;              case 14: v = (-cx * sy); break;
        case 15: v = (-cy * sz); break;                case 16: v = (-cy * cz); break;                case 17: v = 0; break;
        case 18: v = (cx * cz - sx * sy * sz); break;  case 19: v = (-cx * sz - sx * sy * cz); break; case 20: v = 0; break;
        case 21: v = (sx * cz + cx * sy * sz); break;  case 22: v = (cx * sy * cz - sx * sz); break;  case 23: v = 0; break;
        case 24: v = (-cx * sz - sx * sy * cz); break; case 25: v = (-cx * cz + sx * sy * sz); break; case 26: v = sx * cy; break;
        case 27: v = (-sx * sz + cx * sy * cz); break; case 28: v = (-cx * sy * sz - sx * cz); break; case 29: v = (-cx * cy); break;
        case 30: v = (cx * cy * cz); break;            case 31: v = (-cx * cy * sz); break;           case 32: v = (cx * sy); break;
        case 33: v = (sx * cy * cz); break;            case 34: v = (-sx * cy * sz); break;           case 35: v = (sx * sy); break;
        case 36: v = (-sx * cz - cx * sy * sz); break; case 37: v = (sx * sz - cx * sy * cz); break;  case 38: v = 0; break;
        case 39: v = (cx * cz - sx * sy * sz); break;  case 40: v = (-sx * sy * cz - cx * sz); break; case 41: v = 0; break;
        case 42: v = (-cy * cz); break;                case 43: v = (cy * sz); break;                 case 44: v = (sy); break;
        case 45: v = (-sx * sy * cz); break;           case 46: v = (sx * sy * sz); break;            case 47: v = (sx * cy); break;
        case 48: v = (cx * sy * cz); break;            case 49: v = (-cx * sy * sz); break;           case 50: v = (-cx * cy); break;
        case 51: v = (sy * sz); break;                 case 52: v = (sy * cz); break;                 case 53: v = 0; break;
        case 54: v = (-sx * cy * sz); break;           case 55: v = (-sx * cy * cz); break;           case 56: v = 0; break;
        case 57: v = (cx * cy * sz); break;            case 58: v = (cx * cy * cz); break;            case 59: v = 0; break;
        case 60: v = (-cy * cz); break;                case 61: v = (cy * sz); break;                 case 62: v = 0; break;
        case 63: v = (-cx * sz - sx * sy * cz); break; case 64: v = (-cx * cz + sx * sy * sz); break; case 65: v = 0; break;
        case 66: v = (-sx * sz + cx * sy * cz); break; case 67: v = (-cx * sy * sz - sx * cz); break; default: v = 0; break;
      }
      if (e < 24) st_out->jang[e] = static_cast<float>(v); else st_out->hang[e - 24] = static_cast<float>(v);
    }
  }
  NEWTON_STAMP(5);
  if (lane == 0) {
    // pose_to_T(x_t): Translation * Rx * Ry * Rz in float (Eigen::AngleAxis<float>::toRotationMatrix for a unit axis)
    float Rm[3][9];
    for (int axis = 0; axis < 3; ++axis) {
      float u[3] = {0.0f, 0.0f, 0.0f};
      u[axis] = 1.0f;
      const float sn = static_cast<float>(s.sc[6 + axis]), c = static_cast<float>(s.sc[9 + axis]);
      const float su[3] = {sn * u[0], sn * u[1], sn * u[2]};
      const float cu[3] = {(1.0f - c) * u[0], (1.0f - c) * u[1], (1.0f - c) * u[2]};
      float* R = Rm[axis];
      float t = cu[0] * u[1];
      R[1] = t - su[2]; R[3] = t + su[2];
      t = cu[0] * u[2];
      R[2] = t + su[1]; R[6] = t - su[1];
      t = cu[1] * u[2];
      R[5] = t - su[0]; R[7] = t + su[0];
      R[0] = cu[0] * u[0] + c; R[4] = cu[1] * u[1] + c; R[8] = cu[2] * u[2] + c;
    }
    float R[9];
    f33_mul(Rm[0], Rm[1], R);
    f33_mul(R, Rm[2], R);
    for (int r = 0; r < 3; ++r) {
      st_out->T[4 * r] = R[3 * r]; st_out->T[4 * r + 1] = R[3 * r + 1]; st_out->T[4 * r + 2] = R[3 * r + 2];
      st_out->T[4 * r + 3] = static_cast<float>(s.x_t[r]);
    }
    st_out->one = 1.0f;
    st_out->d1 = st_in.d1; st_out->d2 = st_in.d2; st_out->d3 = st_in.d3;
    for (int i = 0; i < 6; ++i) st_out->p[i] = s.x_t[i];
  }
  __syncwarp();
  NEWTON_STAMP(6);
  return NEWTON_CONTINUE;
}

}  // namespace ndtb
