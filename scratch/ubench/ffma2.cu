// microbenchmark: scalar FFMA vs packed FFMA2 throughput on sm_100a (scratch, not part of the library).
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ffma2_bench ffma2.cu   result on B200 (r01k): both forms reach
// ~125 FMA/clk/SM: FFMA2 halves the issue slots per FMA, not the pipe time.
#include <cuda_runtime.h>
#include <stdio.h>
__device__ __forceinline__ unsigned long long fma2(unsigned long long a, unsigned long long b, unsigned long long c) {
  unsigned long long d;
  asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__global__ void k_scalar(float* out, float a, float b, int iters) {
  float x[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) x[i] = threadIdx.x + i;
  for (int t = 0; t < iters; ++t) {
#pragma unroll
    for (int i = 0; i < 16; ++i) x[i] = __fmaf_rn(x[i], a, b);
  }
  float s = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += x[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_packed(float* out, float a, float b, int iters) {
  unsigned long long x[8];
  float2 av = make_float2(a, a), bv = make_float2(b, b);
  unsigned long long a2 = *reinterpret_cast<unsigned long long*>(&av), b2 = *reinterpret_cast<unsigned long long*>(&bv);
#pragma unroll
  for (int i = 0; i < 8; ++i) { float2 v = make_float2(threadIdx.x + 2 * i, threadIdx.x + 2 * i + 1); x[i] = *reinterpret_cast<unsigned long long*>(&v); }
  for (int t = 0; t < iters; ++t) {
#pragma unroll
    for (int i = 0; i < 8; ++i) x[i] = fma2(x[i], a2, b2);
  }
  float s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) { float2 v = *reinterpret_cast<float2*>(&x[i]); s += v.x + v.y; }
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
int main() {
  float* out;
  cudaMalloc(&out, 148 * 8 * 256 * 4);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int iters = 20000;
  for (int rep = 0; rep < 2; ++rep) {
    float ms;
    cudaEventRecord(e0); k_scalar<<<148 * 8, 256>>>(out, 0.999f, 0.001f, iters); cudaEventRecord(e1); cudaEventSynchronize(e1);
    cudaEventElapsedTime(&ms, e0, e1);
    double fmas = 148.0 * 8 * 256 * 16 * iters;
    printf("scalar FFMA : %.3f ms  %.1f TFMA/s  (%.1f FMA/clk/SM at 1.965 GHz)\n", ms, fmas / ms / 1e9, fmas / (ms * 1e-3) / 148 / 1.965e9);
    cudaEventRecord(e0); k_packed<<<148 * 8, 256>>>(out, 0.999f, 0.001f, iters); cudaEventRecord(e1); cudaEventSynchronize(e1);
    cudaEventElapsedTime(&ms, e0, e1);
    printf("packed FFMA2: %.3f ms  %.1f TFMA/s  (%.1f FMA/clk/SM)\n", ms, fmas / ms / 1e9, fmas / (ms * 1e-3) / 148 / 1.965e9);
  }
  printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
