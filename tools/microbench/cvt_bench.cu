// Microbenchmark: cost of narrowing FP64 -> FP32 on B200 (F2F.F32.F64 vs an integer bit trick)
// and whether it contends with the DMMA pipe.  Feeds the FP32 screen of pair_tile's epilogue.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o cvt_bench cvt_bench.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#define CK(x) do{cudaError_t e=(x); if(e!=cudaSuccess){printf("CUDA error %s at %d\n",cudaGetErrorString(e),__LINE__); exit(1);} }while(0)

__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1},{%2},{%3},{%0,%1};"
               : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
__device__ __forceinline__ float trick(double x) {
  const int hi = __double2hiint(x), lo = __double2loint(x);
  const int t = (hi & 0x7fffffff) - 0x38000000;
  unsigned f = __funnelshift_l((unsigned)lo, (unsigned)t, 3);
  f = (t < 0x00100000) ? 0u : f;           // below the float normal range (incl. 0) -> 0
  return __uint_as_float(f | ((unsigned)hi & 0x80000000u));
}

// mode 0: all warps F2F; 1: all warps trick; 2: all DMMA; 3: even DMMA / odd F2F; 4: even DMMA / odd trick
__global__ void __launch_bounds__(1024) k(float* out, const double* in, int iters, int mode, long long* clk) {
  const int warp = threadIdx.x >> 5;
  double v[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = in[(threadIdx.x + i) & 63];
  float acc[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) acc[i] = 0.f;
  double c[16][2];
#pragma unroll
  for (int i = 0; i < 16; ++i) { c[i][0] = 0; c[i][1] = 0; }
  __syncthreads();
  const long long t0 = clock64();
  const bool do_mma = (mode == 2) || ((mode == 3 || mode == 4) && !(warp & 1));
  const bool use_trick = (mode == 1 || mode == 4);
  if (do_mma) {
    for (int it = 0; it < iters; ++it)
#pragma unroll
      for (int i = 0; i < 16; ++i) dmma(c[i][0], c[i][1], v[0], v[1]);
  } else if (use_trick) {
    for (int it = 0; it < iters; ++it)
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        const float f = trick(v[i]);
        acc[i] += f;
        v[i] = __hiloint2double(__double2hiint(v[i]) ^ (it & 1), __double2loint(v[i]));  // defeat hoisting
      }
  } else {
    for (int it = 0; it < iters; ++it)
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        const float f = __double2float_rn(v[i]);
        acc[i] += f;
        v[i] = __hiloint2double(__double2hiint(v[i]) ^ (it & 1), __double2loint(v[i]));
      }
  }
  const long long t1 = clock64();
  float s = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += acc[i] + (float)(c[i][0] + c[i][1]);
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) clk[blockIdx.x] = t1 - t0;
}

int main() {
  double h[64];
  for (int i = 0; i < 64; ++i) h[i] = 1.0 + i * 0.37;
  double* din; float* dout; long long* dclk;
  CK(cudaMalloc(&din, sizeof(h))); CK(cudaMemcpy(din, h, sizeof(h), cudaMemcpyHostToDevice));
  CK(cudaMalloc(&dout, 148 * 1024 * 4)); CK(cudaMalloc(&dclk, 148 * 8));
  const char* names[] = {"all warps F2F.F32.F64", "all warps int-trick", "all warps DMMA",
                         "even DMMA / odd F2F", "even DMMA / odd int-trick"};
  for (int warps : {8, 16}) {
    for (int mode = 0; mode < 5; ++mode) {
      const int iters = 2000;
      k<<<148, warps * 32, 0>>>(dout, din, iters, mode, dclk);
      CK(cudaDeviceSynchronize());
      k<<<148, warps * 32, 0>>>(dout, din, iters, mode, dclk);
      CK(cudaDeviceSynchronize());
      long long c[148]; CK(cudaMemcpy(c, dclk, sizeof(c), cudaMemcpyDeviceToHost));
      double avg = 0; for (int i = 0; i < 148; ++i) avg += c[i]; avg /= 148;
      const double ops_per_warp = (double)iters * 16;
      printf("warps/SM=%2d %-28s clk=%9.0f  clk per warp-op (whole SM, one warp's stream) = %.2f ; per-SMSP issue interval = %.2f clk\n",
             warps, names[mode], avg, avg / ops_per_warp, avg / ops_per_warp / (warps / 4.0) * ((mode >= 3) ? 2.0 : 1.0));
    }
  }
  return 0;
}
