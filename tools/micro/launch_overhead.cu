// Development aid: what does a launch cost between two CUDA events, as a function of CTA size / dynamic smem / params?
#include <cstdio>
#include <cuda_runtime.h>
struct Big { float f[100]; };
__global__ void k_empty(unsigned long long* out) { if (out && threadIdx.x == 0 && blockIdx.x == 0) { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); out[0] = t; } }
__global__ void k_smem(unsigned long long* out) { extern __shared__ float s[]; if (threadIdx.x == 0) s[0] = 1.f; if (out && threadIdx.x == 0 && blockIdx.x == 0) { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); out[0] = t + (s[0] > 2.f); } }
__global__ void k_param(const __grid_constant__ Big b, unsigned long long* out) { if (out && threadIdx.x == 0 && blockIdx.x == 0) out[0] = (unsigned long long)b.f[3]; }
template <class F> float timeit(F f, cudaStream_t s, int reps = 200) {
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  for (int i = 0; i < 10; ++i) f();
  cudaStreamSynchronize(s);
  float tot = 0;
  for (int i = 0; i < reps; ++i) { cudaEventRecord(a, s); f(); cudaEventRecord(b, s); cudaEventSynchronize(b); float ms; cudaEventElapsedTime(&ms, a, b); tot += ms; }
  return tot / reps * 1e3f;
}
int main() {
  cudaStream_t s; cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking);
  unsigned long long* d; cudaMalloc(&d, 64);
  cudaFuncSetAttribute(k_smem, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  Big big{};
  printf("empty 1x32           : %.2f us\n", timeit([&] { k_empty<<<1, 32, 0, s>>>(d); }, s));
  printf("empty 148x512        : %.2f us\n", timeit([&] { k_empty<<<148, 512, 0, s>>>(d); }, s));
  printf("empty 444x128        : %.2f us\n", timeit([&] { k_empty<<<444, 128, 0, s>>>(d); }, s));
  printf("smem 32KB 148x512    : %.2f us\n", timeit([&] { k_smem<<<148, 512, 32 * 1024, s>>>(d); }, s));
  printf("smem 100KB 148x512   : %.2f us\n", timeit([&] { k_smem<<<148, 512, 100 * 1024, s>>>(d); }, s));
  printf("smem 199KB 148x512   : %.2f us\n", timeit([&] { k_smem<<<148, 512, 199 * 1024, s>>>(d); }, s));
  printf("param 400B 148x512   : %.2f us\n", timeit([&] { k_param<<<148, 512, 0, s>>>(big, d); }, s));
  printf("alternate 199KB/0KB  : %.2f us (pair)\n", timeit([&] { k_smem<<<148, 512, 199 * 1024, s>>>(d); k_empty<<<148, 512, 0, s>>>(d); }, s));
  return 0;
}
