// Microbenchmark: FP64 pipes on B200 (sm_100a).  Answers the design questions of DESIGN.md §3:
//   (1) DMMA.8x8x4 issue rate per SM (FP64 tensor peak), (2) DFMA rate, (3) do DMMA and DFMA
//   overlap (separate pipes) or serialise (shared pipe), (4) DMMA dependent-issue latency,
//   (5) FP32 FFMA alongside DMMA, (6) MUFU.RCP64H / DP division cost.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o fp64_pipes fp64_pipes.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#define CK(x) do{cudaError_t e=(x); if(e!=cudaSuccess){printf("CUDA error %s at %d\n",cudaGetErrorString(e),__LINE__); exit(1);} }while(0)

__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1},{%2},{%3},{%0,%1};"
               : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

// mode bit0: warps with (warp&1)==0 run DMMA ; bit1: the other warps run DFMA; bit2: others FFMA
// mode 8: every warp interleaves DMMA and DFMA in one instruction stream
template <int NACC>
__global__ void __launch_bounds__(1024) pipes(double* out, const double* in, int iters, int mode,
                                              long long* clk) {
  const int warp = threadIdx.x >> 5;
  double a = in[threadIdx.x & 31], b = in[32 + (threadIdx.x & 31)];
  double c[NACC][2];
#pragma unroll
  for (int i = 0; i < NACC; ++i) { c[i][0] = 0.0; c[i][1] = 0.0; }
  float fa = (float)a, fb = (float)b, fc[2 * NACC];
#pragma unroll
  for (int i = 0; i < 2 * NACC; ++i) fc[i] = 0.f;
  __syncthreads();
  long long t0 = clock64();
  const bool even = (warp & 1) == 0;
  if (mode == 8) {
    for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int i = 0; i < NACC / 2; ++i) dmma(c[i][0], c[i][1], a, b);
#pragma unroll
      for (int i = NACC / 2; i < NACC; ++i) {
        c[i][0] = fma(a, b, c[i][0]); c[i][1] = fma(a, b, c[i][1]);
        c[i][0] = fma(a, b, c[i][0]); c[i][1] = fma(a, b, c[i][1]);
        c[i][0] = fma(a, b, c[i][0]); c[i][1] = fma(a, b, c[i][1]);
        c[i][0] = fma(a, b, c[i][0]); c[i][1] = fma(a, b, c[i][1]);
      }
    }
  } else if ((mode & 1) && (even || mode == 1)) {
    for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int i = 0; i < NACC; ++i) dmma(c[i][0], c[i][1], a, b);
    }
  } else if ((mode & 2) && (!even || mode == 2)) {
    for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int i = 0; i < NACC; ++i) {
        // 8 DFMA = the FMA count of one DMMA.8x8x4 per warp
        c[i][0] = fma(a, b, c[i][0]); c[i][1] = fma(a, b, c[i][1]);
        c[i][0] = fma(a, b, c[i][0]); c[i][1] = fma(a, b, c[i][1]);
        c[i][0] = fma(a, b, c[i][0]); c[i][1] = fma(a, b, c[i][1]);
        c[i][0] = fma(a, b, c[i][0]); c[i][1] = fma(a, b, c[i][1]);
      }
    }
  } else if ((mode & 4) && (!even || mode == 4)) {
    for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int r = 0; r < 8; ++r)
#pragma unroll
        for (int i = 0; i < 2 * NACC; ++i) fc[i] = fmaf(fa, fb, fc[i]);
    }
  }
  long long t1 = clock64();
  double s = 0; float fs = 0;
#pragma unroll
  for (int i = 0; i < NACC; ++i) s += c[i][0] + c[i][1];
#pragma unroll
  for (int i = 0; i < 2 * NACC; ++i) fs += fc[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s + fs;
  if (threadIdx.x == 0) clk[blockIdx.x] = t1 - t0;
}

__global__ void dmma_latency(double* out, const double* in, int iters, long long* clk) {
  double a = in[threadIdx.x], b = in[32 + threadIdx.x], c0 = 0, c1 = 0;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) dmma(c0, c1, a, b);
  long long t1 = clock64();
  out[threadIdx.x] = c0 + c1;
  if (threadIdx.x == 0) clk[0] = t1 - t0;
}
__global__ void dfma_latency(double* out, const double* in, int iters, long long* clk) {
  double a = in[threadIdx.x], b = in[32 + threadIdx.x], c0 = 0;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) c0 = fma(a, c0, b);
  long long t1 = clock64();
  out[threadIdx.x] = c0;
  if (threadIdx.x == 0) clk[0] = t1 - t0;
}
__global__ void ddiv_rate(double* out, const double* in, int iters, long long* clk) {
  double a = in[threadIdx.x & 31] + 1.5, c[4] = {1.1, 1.2, 1.3, 1.4};
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 4; ++i) c[i] = a / (c[i] + 1.0);
  }
  long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = c[0] + c[1] + c[2] + c[3];
  if (threadIdx.x == 0) clk[blockIdx.x] = t1 - t0;
}
__global__ void dsqrt_rate(double* out, const double* in, int iters, long long* clk) {
  double a = in[threadIdx.x & 31] + 1.5, c[4] = {1.1, 1.2, 1.3, 1.4};
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 4; ++i) c[i] = sqrt(c[i] + a);
  }
  long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = c[0] + c[1] + c[2] + c[3];
  if (threadIdx.x == 0) clk[blockIdx.x] = t1 - t0;
}

int main() {
  cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
  const int nsm = p.multiProcessorCount;
  printf("device %s, %d SMs, clockRate %d kHz\n", p.name, nsm, p.clockRate);
  double *in, *out; long long* clk;
  CK(cudaMalloc(&in, 4096)); CK(cudaMalloc(&out, sizeof(double) * 1024 * nsm * 4));
  CK(cudaMalloc(&clk, sizeof(long long) * nsm * 4));
  double h[512]; for (int i = 0; i < 512; ++i) h[i] = 1e-3 * (i % 7) ;
  CK(cudaMemcpy(in, h, 4096, cudaMemcpyHostToDevice));
  cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  const int iters = 4000;
  constexpr int NACC = 8;
  struct { int mode; const char* name; } modes[] = {
    {1, "DMMA only (all warps)"}, {2, "DFMA only (all warps)"}, {4, "FFMA only (all warps)"},
    {3, "even warps DMMA, odd warps DFMA"}, {5, "even warps DMMA, odd warps FFMA"},
    {8, "each warp interleaves DMMA+DFMA"}};
  for (int warps : {4, 8, 16, 32}) {
    for (auto& m : modes) {
      pipes<NACC><<<nsm, warps * 32>>>(out, in, 10, m.mode, clk);  // warm
      CK(cudaDeviceSynchronize());
      CK(cudaEventRecord(e0));
      pipes<NACC><<<nsm, warps * 32>>>(out, in, iters, m.mode, clk);
      CK(cudaEventRecord(e1)); CK(cudaDeviceSynchronize());
      float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
      long long hc[1024]; CK(cudaMemcpy(hc, clk, sizeof(long long) * nsm, cudaMemcpyDeviceToHost));
      double avgclk = 0; for (int i = 0; i < nsm; ++i) avgclk += hc[i]; avgclk /= nsm;
      // op counts per SM
      double n_dmma = 0, n_dfma = 0, n_ffma = 0;
      const int wd = (m.mode == 1 || m.mode == 8) ? warps : ((m.mode & 1) ? warps / 2 : 0);
      const int wf = (m.mode == 2 || m.mode == 8) ? warps : ((m.mode & 2) ? warps / 2 : 0);
      const int ws = (m.mode == 4) ? warps : ((m.mode & 4) ? warps / 2 : 0);
      const double accd = (m.mode == 8) ? NACC / 2 : NACC;
      n_dmma = (double)wd * iters * accd;              // warp-level DMMA instr per SM
      n_dfma = (double)wf * iters * accd * 8;          // warp-level DFMA instr per SM
      n_ffma = (double)ws * iters * NACC * 2 * 8;
      const double flops = (n_dmma * 512 + n_dfma * 64 + n_ffma * 64) * nsm;
      printf("warps/SM=%2d %-36s ms=%8.3f clk/SM=%10.0f  DMMA/clk/SM=%6.4f DFMA/clk/SM=%6.4f FFMA/clk/SM=%6.4f  total %.2f TFLOP/s (eff clk %.0f MHz)\n",
             warps, m.name, ms, avgclk, n_dmma / avgclk, n_dfma / avgclk, n_ffma / avgclk,
             flops / (ms * 1e-3) / 1e12, avgclk / (ms * 1e-3) / 1e6);
    }
  }
  dmma_latency<<<1, 32>>>(out, in, 10000, clk); CK(cudaDeviceSynchronize());
  long long c; CK(cudaMemcpy(&c, clk, 8, cudaMemcpyDeviceToHost));
  printf("DMMA dependent latency: %.2f clk\n", c / 10000.0);
  dfma_latency<<<1, 32>>>(out, in, 10000, clk); CK(cudaDeviceSynchronize());
  CK(cudaMemcpy(&c, clk, 8, cudaMemcpyDeviceToHost));
  printf("DFMA dependent latency: %.2f clk\n", c / 10000.0);
  for (int warps : {4, 16}) {
    ddiv_rate<<<nsm, warps * 32>>>(out, in, 2000, clk); CK(cudaDeviceSynchronize());
    long long hc[1024]; CK(cudaMemcpy(hc, clk, sizeof(long long) * nsm, cudaMemcpyDeviceToHost));
    printf("DP div: warps=%d  clk per warp-div per SM = %.2f\n", warps, hc[0] / (2000.0 * 4 * warps));
    dsqrt_rate<<<nsm, warps * 32>>>(out, in, 2000, clk); CK(cudaDeviceSynchronize());
    CK(cudaMemcpy(hc, clk, sizeof(long long) * nsm, cudaMemcpyDeviceToHost));
    printf("DP sqrt: warps=%d  clk per warp-sqrt per SM = %.2f\n", warps, hc[0] / (2000.0 * 4 * warps));
  }
  return 0;
}
