// Development aid: throughput of F2F.F64.F32, DADD, integer float->double widening, per SM.
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ double widen_int(float f) {
  const unsigned int b = __float_as_uint(f);
  const unsigned int a = b & 0x7fffffffu;
  unsigned int hi = (a >> 3) + 0x38000000u;
  hi = (a < 0x00800000u) ? 0u : hi;          // zero / denormal -> 0
  hi |= b & 0x80000000u;
  return __hiloint2double(hi, b << 29);
}
template <int MODE>
__global__ void k(const float* in, double* out, int iters) {
  float f[8];
  for (int i = 0; i < 8; ++i) f[i] = in[threadIdx.x + 32 * i];
  double acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (MODE == 0) acc[i] += static_cast<double>(f[i]);          // F2F + DADD
      if (MODE == 1) acc[i] += widen_int(f[i]);                    // int widen + DADD
      if (MODE == 2) acc[i] += __longlong_as_double(__float_as_uint(f[i]));  // DADD only (+ int->reg moves)
      if (MODE == 3) acc[i] += static_cast<double>(f[i]) + static_cast<double>(f[i] + 1.5f);  // 2 F2F + 2 DADD
      f[i] = f[i] * 1.0001f + 0.25f;                               // keep the conversion inside the loop
    }
  }
  double s = 0;
  for (int i = 0; i < 8; ++i) s += acc[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int MODE> void run(const char* name, float* in, double* out) {
  const int iters = 4096, blocks = 148 * 4, threads = 512;
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  k<MODE><<<blocks, threads>>>(in, out, 16);
  cudaEventRecord(a); k<MODE><<<blocks, threads>>>(in, out, iters); cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b);
  const double ops = double(blocks) * threads * iters * 8;
  printf("%-22s %.3f ms  -> %.1f element-ops / clk / SM (at 1.9 GHz)\n", name, ms, ops / (ms * 1e-3) / 148 / 1.9e9);
}
int main() {
  float* in; double* out; cudaMalloc(&in, 4096 * 4); cudaMemset(in, 0x3f, 4096 * 4); cudaMalloc(&out, 148 * 4 * 512 * 8);
  run<0>("F2F.F64.F32 + DADD", in, out);
  run<1>("int widen + DADD", in, out);
  run<2>("DADD only", in, out);
  run<3>("F2F f32->f64->f32", in, out);
  return 0;
}
