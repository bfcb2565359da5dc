// Kernels (2)+(3): all-pairs 3x3 covariance blocks on the FP64 tensor path (DMMA m8n8k4) with the
// per-pair epilogue fused in registers.
//
// Replaces the reference's pair loops K1/K2/K3 (src/cregen.f90:1266-1349: serial i, OpenMP j,
// two rmsd() calls per pair = src/ls_rmsd.f90:72-174) and the duplicate test of K4
// (src/cregen.f90:1353-1373).  No float32 rmat band is materialised: the thresholds are applied
// where the covariance lives.
//
// Tiling: a CTA owns a 32(i) x 64(j) pair tile; 8 consumer warps (2 x 4) each own 16 x 16 pairs =
// 2 x 2 DMMA tiles x 9 covariance components (a,b in xyz) = 72 FP64 accumulators per lane, so a
// lane ends the K loop holding the complete 3x3 R of its 8 pairs - no shuffle before the epilogue.
// One producer warp streams (panel, slab) chunks of the packed operand X with 1-D bulk async
// copies (TMA engine, cp.async.bulk -> UBLKCP) into an mbarrier ring; the ring runs across tile
// boundaries so the next tile's operands land while the current epilogue executes.
// tcgen05/UMMA has no FP64 kind, so the FP64 contraction is warp-level mma.sync (DMMA); the
// microbenchmark (profiles/r01_fp64_pipes_microbench.txt) shows DMMA and DFMA share one pipe on
// B200, which is why the epilogue is written to minimise FP64 instruction count.
#include "common.cuh"
#include "qcp.cuh"
#include "pipe.cuh"

// make PROF=1 builds the phase-instrumented kernel (clock64 counters per warp, CREST_CUDA_DEBUG=2)
#ifndef CREST_PAIR_PROF
#define CREST_PAIR_PROF 0
#endif

namespace crest {

constexpr int kMaxStages = 8;

// Tile index -> (i-panel I, j-tile J).  A CTA walks t = t0, t0+grid, ... : after one binary search
// the row is advanced incrementally (a dependent-load search per tile costs ~5k clk of latency).
struct TileCursor {
  int I = -1;
  long long row_beg = 0, row_end = 0;   // rowstart[I], rowstart[I+1]
  int jf = 0;
  __device__ __forceinline__ void seek(const PairArgs& a, long long t, int& Iout, int& Jout) {
    if (a.tiles) {                      // explicit device-generated tile list
      const int4 tl = a.tiles[t];
      Iout = tl.x; Jout = tl.y;
      return;
    }
    if (I < 0) {
      int lo = 0, hi = a.nI - 1;        // largest I with rowstart[I] <= t
      while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (a.rowstart[mid] <= t) lo = mid; else hi = mid - 1;
      }
      I = lo;
      row_beg = a.rowstart[I]; row_end = a.rowstart[I + 1]; jf = a.jfirst[I];
    }
    while (t >= row_end) {              // also skips empty rows
      ++I;
      row_beg = row_end; row_end = a.rowstart[I + 1]; jf = a.jfirst[I];
    }
    Iout = I;
    Jout = jf + (int)(t - row_beg);
  }
};

// SINK (compare_blocks / intraconf, blockpair.cu) is a template parameter for the reason given in pair_col.cu:
// compiled into the CREGEN instantiation, the out-of-line sink call put the argument struct into local memory
// (stack frame 72 -> 1040 bytes).
template <bool SINK>
__global__ void __launch_bounds__(kPairThreads, 1)
pair_tile_kernel(const PairArgs a, const int nstage) {
  const long long tile_end = a.work_count ? a.tile_begin + (long long)*a.work_count : a.tile_end;
  if (a.tile_begin + blockIdx.x >= tile_end) return;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[kMaxStages];
  __shared__ __align__(8) uint64_t empty_bar[kMaxStages];

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const PackLayout L = a.L;
  const size_t chunk = L.chunk_doubles();
  const uint32_t chunk_bytes = (uint32_t)(chunk * sizeof(double));
  const size_t stage_doubles = chunk * (1 + kJPanels);
  double* const sbase = reinterpret_cast<double*>(smem_raw);

  if (threadIdx.x == 0) {
    for (int s = 0; s < nstage; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], kConsumerWarps);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  // -------------------------------- consumer warps --------------------------------
  const int wi = warp & 1;        // 16-row half of the i panel
  const int wj = warp >> 1;       // 16-column quarter of the 64 j columns
  const int c = lane >> 2, tq = lane & 3;
  // fragment base offsets (doubles) inside a stage: row (conformer) * KS + 2*tq
  const int KS = L.KS;
  const int a_off = (wi * 16 + c) * KS + 2 * tq;                       // + mi*8*KS + comp*32*KS
  const int jrow = wj * 16 + c;                                        // 0..63 (+ nj*8)
  const int b_off0 = (int)chunk * (1 + (jrow >> 5)) + (jrow & 31) * KS + 2 * tq;
  const int b_off1 = (int)chunk * (1 + ((jrow + 8) >> 5)) + ((jrow + 8) & 31) * KS + 2 * tq;
  const int comp_stride = kPanel * KS;
  const int nblk = L.KU >> 3;

  // Producer duty is folded into lane 0 of warp 0 (a 9th warp would be accounted as 12 warps by
  // the register allocator and cap the kernel at 168 registers).  An item = one K slab of one
  // tile (3 bulk copies); the ring is kept `dist` items ahead of warp 0.  Measured with the
  // CREST_PAIR_PROF build: issuing one item costs warp 0 ~1000 clk, which is why few large items
  // beat a deep ring of small ones.
  const bool is_producer = (threadIdx.x == 0);
  long long p_tile = a.tile_begin + blockIdx.x;
  int p_slab = 0, p_I = 0, p_J = 0;
  uint32_t p_it = 0;
  TileCursor p_cur;
  auto produce_one = [&]() {
    if (p_tile >= tile_end) return;
    if (p_slab == 0) p_cur.seek(a, p_tile, p_I, p_J);
    const int st = p_it % nstage;
    const uint32_t ph = (p_it / nstage) & 1;
    mbar_wait(&empty_bar[st], ph ^ 1);
    double* dst = sbase + (size_t)st * stage_doubles;
    mbar_expect_tx(&full_bar[st], chunk_bytes * (1 + kJPanels));
    bulk_g2s(dst, a.X + ((size_t)p_I * L.nslab + p_slab) * chunk, chunk_bytes, &full_bar[st]);
#pragma unroll
    for (int p = 0; p < kJPanels; ++p)
      bulk_g2s(dst + chunk * (1 + p),
               ((SINK && a.sink.mode == 1) ? a.sink.Xj : a.X) + ((size_t)(p_J * kJPanels + p) * L.nslab + p_slab) * chunk,
               chunk_bytes, &full_bar[st]);
    ++p_it;
    if (++p_slab == L.nslab) { p_slab = 0; p_tile += gridDim.x; }
  };
  if (is_producer)
    for (int s = 0; s < a.dist; ++s) produce_one();

  // Anti-phase start: the two warps of an SM sub-partition (w, w+4) share the DMMA pipe.  Warps
  // 4-7 begin half a tile late, so one warp's epilogue (FFMA/ALU pipes) runs under the other's
  // DMMA stream instead of both idling the tensor pipe together; the offset is conserved by the
  // pipe sharing itself and bounded by the ring depth (nstage - dist - 1 items).
#if CREST_PAIR_PROF
  if (warp >= 4 && a.skew_clk > 0) {
    const long long t0 = clock64();
    while (clock64() - t0 < a.skew_clk) {}
  }
#endif
  unsigned long long my_pairs = 0;
  uint32_t it = 0;
  long long tw = 0, tm = 0, te = 0, tp = 0, tc = 0;   // debug cycle counters
#if CREST_PAIR_PROF
  const bool prof = a.dbg != nullptr;
#else
  constexpr bool prof = false;
#endif
  TileCursor cur;
  for (long long t = a.tile_begin + blockIdx.x; t < tile_end; t += gridDim.x) {
    int I, J;
    cur.seek(a, t, I, J);
    const int i0 = I * kTileI + wi * 16;       // first i of this warp
    const int j0 = J * kTileJ + wj * 16;       // first j of this warp
    // warp-level activity: any pair with j < i, i < nall
    const bool active = (SINK && a.sink.mode == 1) ? (i0 < a.nall) && (j0 < a.sink.nj)
                                         : (j0 < i0 + 16) && (i0 < a.nall) && (j0 < a.nall);

    double acc[2][2][9][2];
#pragma unroll
    for (int x = 0; x < 2; ++x)
#pragma unroll
      for (int y = 0; y < 2; ++y)
#pragma unroll
        for (int z = 0; z < 9; ++z) { acc[x][y][z][0] = 0.0; acc[x][y][z][1] = 0.0; }

    for (int s = 0; s < L.nslab; ++s, ++it) {
      const int st = it % nstage;
      const uint32_t ph = (it / nstage) & 1;
      if (prof) tc = clock64();
      if (is_producer) produce_one();
      __syncwarp();
      if (prof) { const long long t1 = clock64(); tp += t1 - tc; tc = t1; }
      mbar_wait(&full_bar[st], ph);
      if (prof) { const long long t1 = clock64(); tw += t1 - tc; tc = t1; }
      if (active) {
        const double* sb = sbase + (size_t)st * stage_doubles;
#pragma unroll 1
        for (int kb = 0; kb < nblk; ++kb) {
          double2 fa[2][3];
#pragma unroll
          for (int mi = 0; mi < 2; ++mi)
#pragma unroll
            for (int x = 0; x < 3; ++x)
              fa[mi][x] = *reinterpret_cast<const double2*>(
                  sb + a_off + mi * 8 * KS + x * comp_stride + kb * 8);
#pragma unroll
          for (int nj = 0; nj < 2; ++nj) {
            double2 fb[3];
            const int bo = nj ? b_off1 : b_off0;
#pragma unroll
            for (int y = 0; y < 3; ++y)
              fb[y] = *reinterpret_cast<const double2*>(sb + bo + y * comp_stride + kb * 8);
#pragma unroll
            for (int mi = 0; mi < 2; ++mi)
#pragma unroll
              for (int x = 0; x < 3; ++x)
#pragma unroll
                for (int y = 0; y < 3; ++y) {
                  dmma(acc[mi][nj][3 * x + y][0], acc[mi][nj][3 * x + y][1], fa[mi][x].x, fb[y].x);
                  dmma(acc[mi][nj][3 * x + y][0], acc[mi][nj][3 * x + y][1], fa[mi][x].y, fb[y].y);
                }
          }
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty_bar[st]);
      if (prof) { const long long t1 = clock64(); tm += t1 - tc; tc = t1; }
    }
    if (!active || a.g.debug == 1) continue;
    if (prof) tc = clock64();
    if constexpr (SINK) {
      // rmsd() callers beside cregen_CRE (blockpair.cu): every valid pair to the sink
#pragma unroll 1
      for (int q = 0; q < 8; ++q) {
        const int mi = q >> 2, nj = (q >> 1) & 1, p = q & 1;
        const int i = i0 + mi * 8 + c, j = j0 + nj * 8 + 2 * tq + p;
        const bool ok = i < a.nall && (a.sink.mode == 1 ? j < a.sink.nj : j < i);
        if (!ok) continue;
        const double xni = a.xn[i];
        const double xnjv = a.sink.mode == 1 ? a.sink.xnj[j] : a.xn[j];
        sink_pair(a.sink, i, j, acc[mi][nj][0][p], acc[mi][nj][1][p], acc[mi][nj][2][p], acc[mi][nj][3][p],
                  acc[mi][nj][4][p], acc[mi][nj][5][p], acc[mi][nj][6][p], acc[mi][nj][7][p], acc[mi][nj][8][p],
                  xni + xnjv);
      }
      continue;
    }

    // ------------------------------ fused epilogue ------------------------------
    // Pass 1 (branch-free, 8 independent pairs per lane): quartic invariants of each 3x3 R and
    // the sign screen at lambda* = (s - E2)/2.  Pass 2 (rare): exact Newton/Jacobi solve only for
    // pairs the screen could not prove to lie beyond 2*rthr.
    // energy gate can only fail in tiles on the band edge: max de sits at (largest i, smallest j)
    const int ti_last = min(I * kTileI + kTileI - 1, a.nall - 1);
    const bool edge =
        !(__dmul_rn(__dsub_rn(a.er[ti_last], a.er[J * kTileJ]), kAutokcal) < a.g.ethr);
    double xnj[2][2];
    float lam_j[2][2];          // 0.5*(xn_j*(1-2e-6) - E2) rounded down; + 0.5*xn_i*(1-2e-6) gives lam
    const float e2_up = a.g.E2_up;
#pragma unroll
    for (int nj = 0; nj < 2; ++nj)
#pragma unroll
      for (int p = 0; p < 2; ++p) {
        const int j = j0 + nj * 8 + 2 * tq + p;
        xnj[nj][p] = (SINK && a.sink.mode == 1) ? ((j < a.sink.nj) ? a.sink.xnj[j] : 0.0) : ((j < a.nall) ? a.xn[j] : 0.0);
        lam_j[nj][p] = 0.5f * __fsub_rd(f64_to_f32_trunc(xnj[nj][p]) * 0.999998f, e2_up);
      }
    unsigned cand = 0;
#pragma unroll
    for (int mi = 0; mi < 2; ++mi) {
      const int i = i0 + mi * 8 + c;
      const bool iok = i < a.nall;
      const double xni = iok ? a.xn[i] : 0.0;
      const double eri = (edge && iok) ? a.er[i] : 0.0;
      const float lam_i = 0.5f * (f64_to_f32_trunc(xni) * 0.999998f);
#pragma unroll
      for (int nj = 0; nj < 2; ++nj)
#pragma unroll
        for (int p = 0; p < 2; ++p) {
          const int j = j0 + nj * 8 + 2 * tq + p;
          bool ok = iok && (j < i);
          if (edge && ok) ok = __dmul_rn(__dsub_rn(eri, a.er[j]), kAutokcal) < a.g.ethr;
          float rf[9];
#pragma unroll
          for (int z = 0; z < 9; ++z) rf[z] = f64_to_f32_trunc(acc[mi][nj][z][p]);
          // lam <= (s - E2)/2 - 1e-6*s: below the exact decision point by far more than the
          // exact path's own rounding (relative 2e-6 covers the three float roundings)
          const float lam_f = __fadd_rd(lam_i, lam_j[nj][p]);
          // lam_f < 2^99 bounds every |R_ab| (<= s/2) inside the float range of the narrowing
          const bool far = (lam_f < 6.0e29f) & screen_far_f32(rf, lam_f, a.g.enantio);
          my_pairs += ok;
          if (ok && (!far || a.rmat)) cand |= 1u << (mi * 4 + nj * 2 + p);
        }
    }
    if (prof) te += clock64() - tc;
    if (!a.mask && !__any_sync(0xffffffffu, cand != 0)) continue;
#pragma unroll
    for (int mi = 0; mi < 2; ++mi) {
      const int i = i0 + mi * 8 + c;
      const double xni = (i < a.nall) ? a.xn[i] : 0.0;
#pragma unroll
      for (int nj = 0; nj < 2; ++nj) {
        unsigned hit[2] = {0u, 0u};
#pragma unroll
        for (int p = 0; p < 2; ++p) {
          const int j = j0 + nj * 8 + 2 * tq + p;
          bool dupl = false;
          if (cand & (1u << (mi * 4 + nj * 2 + p))) {
            const double s = xni + xnj[nj][p];
            bool far = false;
            if (!a.rmat) {   // second level: the same screen in FP64 (rigorous much closer to E2)
              double r[9];
#pragma unroll
              for (int z = 0; z < 9; ++z) r[z] = acc[mi][nj][z][p];
              const Quartic q = quartic_of(r);
              const double c1 = a.g.enantio ? -q.c1abs : -8.0 * q.det;
              far = screen_far(q, c1, s, fma(1e-8, s, a.g.E2));
            }
            if (!far) {
              const double e = pair_e_exact(acc[mi][nj][0][p], acc[mi][nj][1][p],
                                            acc[mi][nj][2][p], acc[mi][nj][3][p],
                                            acc[mi][nj][4][p], acc[mi][nj][5][p],
                                            acc[mi][nj][6][p], acc[mi][nj][7][p],
                                            acc[mi][nj][8][p], s, a.g.enantio);
              if (a.rmat) {
                const double rr = sqrt(fmax(0.0, e) / (double)L.A);
                a.rmat[(size_t)i * (i - 1) / 2 + j] = (float)rr;
              }
              if (e < a.g.E2) {
                dupl = e < a.g.E1;
                if (!dupl)
                  dupl = equalrotaniso_dev(a.rot + 3 * (size_t)i, a.rot + 3 * (size_t)j,
                                           a.aniso[i], a.aniso[j], a.g.bthr_half, a.g.bthrmax,
                                           a.g.bthrshift);
                if (dupl) atomicMax_system(a.dup + i, j + 1);
              }
            }
          }
          if (a.mask) hit[p] = __ballot_sync(0xffffffffu, dupl);
        }
        if (a.mask && lane == 0 && (hit[0] | hit[1])) {
          // ballot bit (row*4 + tq) of parity p  ->  tile bit row*8 + 2*tq + p
          unsigned long long w = 0;
#pragma unroll
          for (int p = 0; p < 2; ++p) {
            unsigned long long x = hit[p];
            x = (x | (x << 16)) & 0x0000FFFF0000FFFFull;
            x = (x | (x << 8)) & 0x00FF00FF00FF00FFull;
            x = (x | (x << 4)) & 0x0F0F0F0F0F0F0F0Full;
            x = (x | (x << 2)) & 0x3333333333333333ull;
            x = (x | (x << 1)) & 0x5555555555555555ull;
            w |= x << p;
          }
          const int ti8 = (i0 + mi * 8) >> 3, tj8 = (j0 + nj * 8) >> 3;
          a.mask[(size_t)ti8 * a.n8 + tj8] = w;
        }
      }
    }
  }
  if (prof && lane == 0) {
    atomicAdd((unsigned long long*)a.dbg + warp * 4 + 0, (unsigned long long)tp);
    atomicAdd((unsigned long long*)a.dbg + warp * 4 + 1, (unsigned long long)tw);
    atomicAdd((unsigned long long*)a.dbg + warp * 4 + 2, (unsigned long long)tm);
    atomicAdd((unsigned long long*)a.dbg + warp * 4 + 3, (unsigned long long)te);
  }
  // one atomic per warp for the pair counter
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) my_pairs += __shfl_xor_sync(0xffffffffu, my_pairs, o);
  if (lane == 0 && my_pairs) atomicAdd(a.npairs, my_pairs);
}

int pair_smem_bytes(const PackLayout& L, int* nstage_out) {
  const size_t stage = L.chunk_doubles() * (1 + kJPanels) * sizeof(double);
  int ns = (int)((size_t)(227 * 1024 - 1024) / stage);
  if (ns > kMaxStages) ns = kMaxStages;
  if (tune().stages > 0 && tune().stages < ns) ns = tune().stages;
  if (ns < 1) ns = 1;
  if (nstage_out) *nstage_out = ns;
  return (int)(stage * ns);
}

int launch_pair_tiles(const PairArgs& a_in, int nstage, cudaStream_t stream, int* grid_out,
                      int* smem_out) {
  PairArgs a = a_in;
  const long long ntiles = a.work_count ? (1ll << 40) : a.tile_end - a.tile_begin;
  if (grid_out) *grid_out = 0;
  if (ntiles <= 0) return 0;
  int dev = 0, nsm = 148;
  CREST_CUDA_TRY(cudaGetDevice(&dev));
  CREST_CUDA_TRY(cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev));
  int ns = nstage;
  if (ns <= 0) pair_smem_bytes(a.L, &ns);
  const size_t stage = a.L.chunk_doubles() * (1 + kJPanels) * sizeof(double);
  const int smem_used = (int)(stage * ns);
  // ring shape: the producer keeps up to `dist` items in flight beyond the one being consumed.
  // Skew = half of a tile's DMMA time on one sub-partition (2 warps x 72 DMMA per 8-atom block,
  // 16 clk of pipe each); it needs ring slots to live in, so only with >= 3 stages.
  a.dist = tune().dist > 0 ? tune().dist : ns - 1;
  if (a.dist > ns - 1) a.dist = ns - 1;
  if (a.dist < 0) a.dist = 0;
  a.skew_clk = 0;          // measured: no gain while all 8 warps share one ring
  if (tune().skew >= 0) a.skew_clk = tune().skew;
  const int grid = (int)(ntiles < nsm ? ntiles : nsm);   // persistent: one CTA per SM
  if (a.sink.mode != 0) {
    CREST_CUDA_TRY(cudaFuncSetAttribute(pair_tile_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_used));
    pair_tile_kernel<true><<<grid, kPairThreads, smem_used, stream>>>(a, ns);
  } else {
    CREST_CUDA_TRY(cudaFuncSetAttribute(pair_tile_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_used));
    pair_tile_kernel<false><<<grid, kPairThreads, smem_used, stream>>>(a, ns);
  }
  CREST_CUDA_TRY(cudaGetLastError());
  if (grid_out) *grid_out = grid;
  if (smem_out) *smem_out = smem_used;
  return 0;
}

}  // namespace crest
