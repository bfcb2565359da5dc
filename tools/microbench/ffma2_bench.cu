// FFMA vs packed FFMA2 (fma.rn.f32x2) issue throughput on sm_100a: is the packed form full rate?
#include <cstdio>
#include <cuda_runtime.h>
template <int MODE>
__global__ void k(float* out, int iters, long long* clk) {
  float a[16];
  for (int i = 0; i < 16; ++i) a[i] = threadIdx.x * 0.001f + i;
  const float b = 1.0001f, c = 0.5f;
  unsigned long long pa[8];
  for (int i = 0; i < 8; ++i) pa[i] = ((unsigned long long)__float_as_uint(a[2 * i + 1]) << 32) | __float_as_uint(a[2 * i]);
  const unsigned long long pb = ((unsigned long long)__float_as_uint(b) << 32) | __float_as_uint(b);
  const unsigned long long pc = ((unsigned long long)__float_as_uint(c) << 32) | __float_as_uint(c);
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    if (MODE == 0) {
#pragma unroll
      for (int i = 0; i < 16; ++i) a[i] = fmaf(a[i], b, c);
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(pa[i]) : "l"(pb), "l"(pc));
    }
  }
  long long t1 = clock64();
  float s = 0;
  for (int i = 0; i < 16; ++i) s += a[i];
  for (int i = 0; i < 8; ++i) s += __uint_as_float((unsigned)pa[i]) + __uint_as_float((unsigned)(pa[i] >> 32));
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) clk[blockIdx.x] = t1 - t0;
}
int main() {
  float* out; long long* clk; cudaMalloc(&out, 4 * 1024 * 148); cudaMalloc(&clk, 8 * 148);
  const int iters = 4096;
  for (int warps : {4, 8, 16}) {
    for (int mode = 0; mode < 2; ++mode) {
      if (mode == 0) k<0><<<148, warps * 32>>>(out, iters, clk); else k<1><<<148, warps * 32>>>(out, iters, clk);
      cudaDeviceSynchronize();
      long long h; cudaMemcpy(&h, clk, 8, cudaMemcpyDeviceToHost);
      const double flops_per_clk = (double)iters * 16 * warps * 32 / (double)h;   // scalar FMAs per clk per SM
      printf("warps/SM=%2d %s: %lld clk, %.1f FMA lanes/clk/SM (%s)\n", warps, mode ? "FFMA2" : "FFMA ", h, flops_per_clk, cudaGetErrorString(cudaGetLastError()));
    }
  }
  return 0;
}
