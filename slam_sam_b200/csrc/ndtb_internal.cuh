// Internal device-side data layout and helpers of libndtb (B200 / sm_100a).
//
// HBM layout (one handle):
//   points      : float4 (x,y,z,pad) packed, 16 B/point, source and target
//   VoxRec[V]   : the HOT voxel record, 64 B, 64-B aligned = 2 sectors: mean 3 x f64 | icov 9 x f32 | n
//                 -- exactly what updateDerivatives reads (ndt_omp_impl.hpp:306-311, 646-650)
//   HashSlot[C] : open-addressing table, 8 B/slot {key = linear voxel index, row}, C = 2^m >= 2 V,
//                 bucket = key & (C-1)  (identity hash + power-of-two mask + load <= 0.5: the
//                 keying svn's grid inherits from tsl::robin_map, robin_growth_policy.h:124-126)
//   cold arrays : idx / n / mean / cov / icov / evecs / evals in FP64 for getLeaves()-style export
// All arithmetic that decides bits (voxel indices, per-pair float maths, FP64 sums) is written
// with explicit operation order and compiled with -fmad=false: no FMA contraction anywhere.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace ndtb {

constexpr int FLAVOUR_PCLOMP = 0;
constexpr int FLAVOUR_SVN = 1;
constexpr uint32_t HASH_EMPTY = 0xFFFFFFFFu;

struct __align__(16) VoxRec {
  double mean[3];
  float icov[9];
  int n;
};
static_assert(sizeof(VoxRec) == 64, "VoxRec must be 64 bytes");

struct HashSlot {
  uint32_t key;
  int32_t row;
};

// Grid geometry, filled on the host exactly as applyFilter does
// (voxel_grid_covariance_omp_impl.hpp:80-113 / voxel_grid_covariance_impl.hpp:103-140).
struct GridMeta {
  int min_b[3], max_b[3], div_b[3], divb_mul[3];
  float leaf, inv_leaf;
  int min_pts;
  int flavour;
  uint32_t hash_mask;
  uint32_t total_cells;  // dx*dy*dz of the bbox in cells (<= INT32_MAX), also the "skipped point" sort key
  float lo_bound[3], hi_bound[3];  // svn isPointWithinBounds (voxel_grid_covariance_impl.hpp:59-64)
  double eig_mult;
};

// Per-evaluation state read by the derivative kernel (global memory, written by the host or by the
// previous evaluation's epilogue).
struct EvalState {
  float T[12];       // rows 0..2 of the float 4x4 that transforms the source cloud
  float jang[24];    // j_ang  8 x 3 (ndt_omp_impl.hpp:442-464)
  float hang[45];    // h_ang 15 x 3 (ndt_omp_impl.hpp:490-536)
  float one;         // 1.0f, opaque to the compiler: multiplier of the packed fma(a, 1, b) additions (derivatives.cu add2)
  double d1, d2, d3; // gauss constants (ndt_omp_impl.hpp:95-99)
  double p[6];       // pose vector of this evaluation
};

// Reduction vector per evaluation: score, g[6], H[36], nearest-voxel score sum, #points with
// neighbours, total neighbourhood count  (ndt_omp_impl.hpp:340-347).
constexpr int RED_N = 46;
constexpr int RED_SCORE = 0, RED_G = 1, RED_H = 7, RED_NEAREST = 43, RED_FOUND = 44, RED_COUNT = 45;

__device__ __forceinline__ int float_to_ordered(float f) {
  int i = __float_as_int(f);
  return i >= 0 ? i : i ^ 0x7fffffff;
}
__host__ __device__ __forceinline__ float ordered_to_float_bits(int i) {
  int j = i >= 0 ? i : i ^ 0x7fffffff;
#ifdef __CUDA_ARCH__
  return __int_as_float(j);
#else
  union { int i; float f; } u; u.i = j; return u.f;
#endif
}

// Self-validating host slot (same scheme as the eval kernel's results): {bits, bits ^ key} as ONE 16-byte store into
// pinned host memory (slot i is 16-byte aligned); the host polls w0 ^ w1 == key instead of synchronising the stream.
// Consecutive lanes -> consecutive slots -> full sectors, a few PCIe write packets (two 8-byte stores per slot had measured
// 2.7 us more per evaluation on the eval kernel's 46 results).
__device__ __forceinline__ void publish_slot(volatile unsigned long long* slots, int i, unsigned long long bits, unsigned long long key) {
  asm volatile("st.volatile.global.v2.u64 [%0], {%1, %2};" ::"l"(const_cast<unsigned long long*>(slots) + 2 * i), "l"(bits), "l"(bits ^ key) : "memory");
}

// Hash lookup: linear probing from key & mask; returns row or -1.
__device__ __forceinline__ int hash_find(const HashSlot* __restrict__ table, uint32_t mask, uint32_t key) {
  uint32_t b = key & mask;
  for (;;) {
    const uint2 s = __ldg(reinterpret_cast<const uint2*>(table) + b);
    if (s.x == key) return static_cast<int>(s.y);
    if (s.x == HASH_EMPTY) return -1;
    b = (b + 1) & mask;
  }
}

// Programmatic dependent launch (sm_90+): the kernels of a build are a chain, each consuming what the previous one wrote.  Every
// kernel first waits for its predecessor to have completed and flushed (griddepcontrol.wait -- a no-op when the launch carries
// no programmatic dependency), then lets its successor be launched: the successor's CTAs are scheduled while this kernel's
// last CTAs drain and sit at their own wait, so launch latency and CTA ramp-up leave the critical path.
__device__ __forceinline__ void pdl_enter() {
  asm volatile("griddepcontrol.wait;" ::: "memory");
  asm volatile("griddepcontrol.launch_dependents;");
}
// Function attributes (dynamic shared memory opt-in) are per DEVICE: a process that drives several GPUs sets them once on each.
// `flags` is the call site's static array; returns true when the current device still needs the attribute.
constexpr int MAX_DEVICES = 64;
inline bool needs_func_setup(bool (&flags)[MAX_DEVICES]) {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= MAX_DEVICES) return true;
  if (flags[dev]) return false;
  flags[dev] = true;
  return true;
}
template <class... KArgs, class... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

}  // namespace ndtb
