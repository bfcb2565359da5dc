// TMEM -> register bandwidth (tcgen05.ld.32x32b.xN) on sm_100a, alone and against packed FP32 work
// in the same warps / in other warps of the same sub-partition.  Answers: is the pair_tf32 epilogue
// bound by TMEM bandwidth (then loads and FFMA2 must add up) or by latency / handshakes?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tmem_ld_bench tmem_ld_bench.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ void ld4(uint32_t t, uint32_t* v) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]) : "r"(t));
}
__device__ __forceinline__ void ld8(uint32_t t, uint32_t* v) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
               : "r"(t));
}
__device__ __forceinline__ void ld16(uint32_t t, uint32_t* v) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                 "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
               : "r"(t));
}
__device__ __forceinline__ void wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// SHAPE: 4 / 8 / 16 registers per load; each "batch" moves 32 (SHAPE 8,16) or 36 (SHAPE 4) registers per
// lane, like one j-group share of an epilogue warp.  FMA = packed FFMA2 instructions per batch.
// MODE 0: loads only   1: FMA only   2: load batch, wait, FMA on it (same warp, like the kernel)
// MODE 3: even warps load, odd warps do FMA (different warps of one sub-partition)
// MODE 4: software pipelined: issue loads of batch n+1, FMA on batch n, then wait
template <int SHAPE, int MODE>
__global__ void __launch_bounds__(512) k(float* out, int iters, int nfma, long long* clk) {
  __shared__ uint32_t tbase;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;"
                 :: "r"((uint32_t)__cvta_generic_to_shared(&tbase)));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");
  const uint32_t t0a = tbase + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)((warp >> 2) * 48) % 320;
  constexpr int NREG = (SHAPE == 4) ? 36 : 32;
  uint32_t v[NREG], w[NREG];
  for (int i = 0; i < NREG; ++i) v[i] = w[i] = lane + i;
  unsigned long long acc[8];
  for (int i = 0; i < 8; ++i) acc[i] = 0x3f8000003f800000ull + i;
  const unsigned long long pb = 0x3f8000103f800010ull;
  __syncthreads();
  const long long c0 = clock64();
  const bool do_ld = (MODE == 0 || MODE == 2 || MODE == 4 || (MODE == 3 && (warp & 4) == 0));
  const bool do_fma = (MODE == 1 || MODE == 2 || MODE == 4 || (MODE == 3 && (warp & 4) != 0));
  auto loads = [&](uint32_t* r) {
    if (SHAPE == 4) { for (int b = 0; b < 9; ++b) ld4(t0a + b * 4, r + 4 * b); }
    if (SHAPE == 8) { for (int b = 0; b < 4; ++b) ld8(t0a + b * 8, r + 8 * b); }
    if (SHAPE == 16) { for (int b = 0; b < 2; ++b) ld16(t0a + b * 16, r + 16 * b); }
  };
  auto fmas = [&](const uint32_t* r) {
    // nfma packed instructions, operands taken from the loaded registers so they are live
    for (int f = 0; f < nfma; f += 8) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const unsigned long long pc = ((unsigned long long)r[(2 * i + 1) % NREG] << 32) | r[(2 * i) % NREG];
        asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(acc[i]) : "l"(pb), "l"(pc));
      }
    }
  };
  if (MODE == 4) {
    loads(v); wait_ld();
    for (int it = 0; it < iters; it += 2) {
      loads(w); fmas(v); wait_ld();
      loads(v); fmas(w); wait_ld();
    }
  } else {
    for (int it = 0; it < iters; ++it) {
      if (do_ld) { loads(v); wait_ld(); }
      if (do_fma) fmas(v);
    }
  }
  const long long c1 = clock64();
  float s = 0;
  for (int i = 0; i < NREG; ++i) s += __uint_as_float(v[i] ^ w[i]);
  for (int i = 0; i < 8; ++i) s += __uint_as_float((unsigned)acc[i]) + __uint_as_float((unsigned)(acc[i] >> 32));
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  __syncthreads();
  if (threadIdx.x == 0) clk[blockIdx.x] = c1 - c0;
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" :: "r"(tbase));
}

template <int SHAPE, int MODE>
static void run(int warps, int iters, int nfma, float* out, long long* clk) {
  k<SHAPE, MODE><<<148, warps * 32>>>(out, iters, nfma, clk);
  cudaError_t e = cudaDeviceSynchronize();
  long long h = 0; cudaMemcpy(&h, clk, 8, cudaMemcpyDeviceToHost);
  const int nreg = SHAPE == 4 ? 36 : 32;
  const int ldw = (MODE == 3) ? warps / 2 : warps, fw = (MODE == 3) ? warps / 2 : warps;
  const double bytes = (MODE == 1) ? 0.0 : (double)iters * ldw * 32 * nreg * 4;
  const double lanes = (MODE == 0) ? 0.0 : (double)iters * fw * 32 * nfma * 2;
  static const char* names[] = {"ld only", "fma only", "ld;wait;fma", "split warps", "sw pipelined"};
  printf("x%-2d warps=%2d %-12s: %9lld clk  %6.1f clk/batch-round  TMEM %6.1f B/clk/SM  FMA %6.1f lanes/clk/SM  (%s)\n",
         SHAPE, warps, names[MODE], h, (double)h / iters, bytes / (double)h, lanes / (double)h, cudaGetErrorString(e));
}

int main() {
  float* out; long long* clk; cudaMalloc(&out, 4 * 512 * 148); cudaMalloc(&clk, 8 * 148);
  const int iters = 2048, nfma = 112;   // 112 packed instructions per batch = the screen of 4 pairs
  for (int warps : {4, 8, 16}) {
    run<4, 0>(warps, iters, nfma, out, clk);
    run<8, 0>(warps, iters, nfma, out, clk);
    run<16, 0>(warps, iters, nfma, out, clk);
    run<4, 1>(warps, iters, nfma, out, clk);
    run<4, 2>(warps, iters, nfma, out, clk);
    run<16, 2>(warps, iters, nfma, out, clk);
    if (warps >= 8) run<4, 3>(warps, iters, nfma, out, clk);
    run<4, 4>(warps, iters, nfma, out, clk);
    run<16, 4>(warps, iters, nfma, out, clk);
  }
  return 0;
}
