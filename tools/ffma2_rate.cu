// Issue rate of packed fp32 (fma.rn.f32x2 -> FFMA2) against scalar FFMA on one SM: 8 warps x 8 independent chains.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/ffma2_rate tools/ffma2_rate.cu && /tmp/ffma2_rate
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ float2 fma2(float2 a, float2 b, float2 c) {
  unsigned long long d;
  asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(*(unsigned long long*)&a), "l"(*(unsigned long long*)&b), "l"(*(unsigned long long*)&c));
  return *(float2*)&d;
}
template <int MODE>
__global__ void k(float* out, long long* cyc, int iters) {
  float2 a[8];
  for (int i = 0; i < 8; ++i) a[i] = make_float2(threadIdx.x * 0.001f + i, i * 0.5f);
  const float2 b = make_float2(1.0001f, 0.9999f), c = make_float2(0.001f, -0.001f);
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (MODE == 0) {
        asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a[i].x) : "f"(b.x), "f"(c.x));
        asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a[i].y) : "f"(b.y), "f"(c.y));
      } else {
        a[i] = fma2(a[i], b, c);
      }
    }
  }
  const long long t1 = clock64();
  float s = 0;
  for (int i = 0; i < 8; ++i) s += a[i].x + a[i].y;
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
int main() {
  float* out; long long* cyc;
  cudaMalloc(&out, 1 << 20); cudaMalloc(&cyc, 1024);
  const int iters = 4096;
  for (int warps : {4, 8, 16}) {
    long long h0, h1;
    k<0><<<1, warps * 32>>>(out, cyc, iters); cudaMemcpy(&h0, cyc, 8, cudaMemcpyDeviceToHost);
    k<1><<<1, warps * 32>>>(out, cyc, iters); cudaMemcpy(&h1, cyc, 8, cudaMemcpyDeviceToHost);
    const double fmas = (double)iters * 16 * warps * 32;
    printf("warps %2d: scalar FFMA %lld cyc (%.1f FMA/clk/SM)   FFMA2 %lld cyc (%.1f FMA/clk/SM)\n", warps, h0, fmas / h0, h1, fmas / h1);
  }
  return 0;
}
