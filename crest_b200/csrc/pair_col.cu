// Kernels (2)+(3), column-sweep variant for A <= 88 atoms: all-pairs 3x3 covariance blocks on the
// FP64 tensor path (DMMA m8n8k4) with the per-pair screen SOFTWARE-PIPELINED under the next
// unit's DMMA stream.
//
// Replaces the reference's pair loops K1/K2/K3 (src/cregen.f90:1266-1349, two rmsd() calls per
// pair = src/ls_rmsd.f90:72-174) and the duplicate test of K4 (src/cregen.f90:1353-1373).
//
// Why a second kernel (measured on B200): in the row-tile kernel (pair_tile.cu) a
// warp's 144 accumulators are all live until its epilogue, both warps of an SM sub-partition reach
// the epilogue together, and the DMMA pipe idles for ~16 % of the time; each ring item also costs
// warp 0 ~1000 clk of bulk-copy issue.  Here
//   * a CTA keeps one j-tile (64 conformers x all K) resident in shared memory and streams 16-row
//     i-units past it: ONE bulk copy per unit, 3.4x less L2->SM traffic;
//   * a warp owns 8(i) x 16(j) pairs of a unit = 72 accumulators, and holds TWO such sets: while
//     the DMMAs of unit u fill one set, the FP32 screen of unit u-1 (quartic invariants + sign
//     test, qcp.cuh) is issued from the other set in the same straight-line block, i.e. in the 15
//     of 16 issue slots the DMMA stream leaves free.  Only candidate pairs (rare) reach the exact
//     FP64 solve after the block.
// tcgen05/UMMA has no FP64 kind, so the contraction is warp-level mma.sync (DMMA).
#include "common.cuh"
#include "qcp.cuh"
#include "pipe.cuh"

#ifndef CREST_PAIR_PROF
#define CREST_PAIR_PROF 0
#endif

namespace crest {

// non-volatile: lets the compiler interleave independent screen arithmetic between DMMAs
__device__ __forceinline__ void dmma_nv(double& c0, double& c1, double a, double b) {
  asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1},{%2},{%3},{%0,%1};"
      : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

constexpr int kColStagesMax = 4;

// K block after which pair q (0..3) of the previous unit is screened.  The two warps of an SM
// sub-partition (w, w+4) get PHASE 0 / 1 so that their screen blocks fall into different K blocks
// whenever NKB allows (NKB >= 8: disjoint): while one runs ~200 ALU/FFMA instructions the other
// keeps the DMMA pipe fed.
template <int NKB, int PHASE>
__host__ __device__ constexpr int screen_slot(int q) {
  return NKB >= 8 ? (q * (NKB / 4)) + PHASE * (NKB / 8)
                  : (NKB >= 4 ? (PHASE ? NKB - 1 - q : q) : (q * NKB) / 4);
}

// what a warp remembers about a unit between its MMA block and its screen / exact pass
struct UnitMeta {
  int i;            // this lane's conformer row (i0w + lane/4); may be >= nall
  int j0w;          // first j column of this warp (J*64 + wj*16)
  unsigned ok;      // bit (nj*2+p): pair passes j<i, bounds and the energy gate
  double xni;       // |x_i|^2
  double xnj[2][2]; // |x_j|^2 of the lane's 4 columns
};

// One straight-line block: DMMA stream of the current unit into `cur` + FP32 screen of the
// previous unit from `prev`.  Returns the candidate bits of the previous unit.
template <int NKB, bool HALF, bool SCREEN>
__device__ __forceinline__ unsigned unit_block(double (&cur)[2][9][2], const double (&prev)[2][9][2],
                                               const double* __restrict__ su,   // unit stage + lane offset
                                               const double* __restrict__ sj0,  // j-tile + lane offset, nj=0
                                               const double* __restrict__ sj1,  // nj=1
                                               int comp_u, int comp_j, const UnitMeta& pm,
                                               float e2_up, int enantio, bool all_cand) {
#pragma unroll
  for (int y = 0; y < 2; ++y)
#pragma unroll
    for (int z = 0; z < 9; ++z) { cur[y][z][0] = 0.0; cur[y][z][1] = 0.0; }
  unsigned cand = 0;
  const float lam_i = 0.5f * (f64_to_f32_trunc(pm.xni) * 0.999998f);
#pragma unroll
  for (int kb = 0; kb < NKB; ++kb) {
    double2 fa[3], fb[2][3];
#pragma unroll
    for (int x = 0; x < 3; ++x) fa[x] = *reinterpret_cast<const double2*>(su + x * comp_u + kb * 8);
#pragma unroll
    for (int y = 0; y < 3; ++y) {
      fb[0][y] = *reinterpret_cast<const double2*>(sj0 + y * comp_j + kb * 8);
      fb[1][y] = *reinterpret_cast<const double2*>(sj1 + y * comp_j + kb * 8);
    }
#pragma unroll
    for (int nj = 0; nj < 2; ++nj)
#pragma unroll
      for (int x = 0; x < 3; ++x)
#pragma unroll
        for (int y = 0; y < 3; ++y) {
          dmma_nv(cur[nj][3 * x + y][0], cur[nj][3 * x + y][1], fa[x].x, fb[nj][y].x);
          if (!(HALF && kb == NKB - 1))     // half tail: the odd slots of the last block are padding
            dmma_nv(cur[nj][3 * x + y][0], cur[nj][3 * x + y][1], fa[x].y, fb[nj][y].y);
        }
    // screen slice: the 4 pairs of the previous unit are spread over the K blocks
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      if (!SCREEN) continue;
      if (screen_slot<NKB, 0>(q) != kb) continue;       // compile-time after unrolling
      const int nj = q >> 1, p = q & 1;
      float rf[9];
#pragma unroll
      for (int z = 0; z < 9; ++z) rf[z] = f64_to_f32_trunc(prev[nj][z][p]);
      const float lam_j = 0.5f * __fsub_rd(f64_to_f32_trunc(pm.xnj[nj][p]) * 0.999998f, e2_up);
      const float lam_f = __fadd_rd(lam_i, lam_j);
      // lam_f <= (s - E2)/2 - 1e-6*s; lam_f < 2^99 bounds every |R_ab| (<= s/2) for the narrowing
      const bool far = (lam_f < 6.0e29f) & screen_far_f32(rf, lam_f, enantio);
      if (((pm.ok >> q) & 1u) && (!far || all_cand)) cand |= 1u << q;
    }
  }
  if (!SCREEN) {   // timing-only variant: keep the accumulators live without screening
    double t = 0.0;
#pragma unroll
    for (int y = 0; y < 2; ++y)
#pragma unroll
      for (int z = 0; z < 9; ++z) t += prev[y][z][0] + prev[y][z][1];
    if (t == 1.2345e-300) cand = 1;
  }
  return cand;
}

// j-tile, i-units [u_begin, u_end); edge != 0: some pair of the item may fail the energy gate
struct ColItem { int J; int u_begin; int u_end; int edge; };

struct ColArgs {
  PairSink sink;         // mode != 0: every valid pair goes to sink_pair (blockpair.cu)
  const double* X;       // column layout: [unit][comp][16][KS]
  const double* xn;
  const double* er;
  const double* rot;
  const double* aniso;
  const ColItem* items;
  int item_begin, item_end;   // this shard's slice of the item list
  int nall;
  int* dup;
  unsigned long long* npairs;
  unsigned long long* mask;
  int n8;
  float* rmat;
  Gates g;
  int A, KS, nstage;
  long long* dbg;        // make PROF=1 + CREST_CUDA_DEBUG=2: per-warp cycle counters
  const unsigned* work_count; // non-NULL: items [item_begin, item_begin + *work_count) written by the
                              // device (dense items behind the tensor-core screen)
};

// Per-warp constants of the kernel, bundled so the two ping-pong call sites stay short.
struct ColWarp {
  int lane, c, tq, wi, wj;
  int a_off, b_off0, b_off1, comp;   // fragment offsets (doubles)
  unsigned long long my_pairs;
};

// Exact pass for the (rare) candidate pairs of a finished unit: second-level FP64 screen, then
// Newton/Jacobi, float32-threshold emulation, rot gate, atomicMax — identical decisions to
// pair_tile.cu's pass 2.  Executed by all lanes (ballots for the optional bitmask).
// SINK is a template parameter on purpose: with the sink call compiled into the CREGEN instantiation the
// kernel argument struct had to live in local memory (sink_pair takes it by reference, out of line) and the
// stack frame of pair_col_kernel<8,1,1> grew from 128 to 1248 bytes, spills from 0.1 to 1.5 KB per thread -
// 178 -> 245 ms on the C3 triangle.  The CREGEN kernels contain no trace of the sink now.
template <bool SINK>
__device__ __forceinline__ void unit_exact(const ColArgs& a, const ColWarp& w,
                                           const double (&acc)[2][9][2], const UnitMeta& m,
                                           unsigned cand) {
  const int i = m.i;
  if constexpr (SINK) {
#pragma unroll 1
    for (int q = 0; q < 4; ++q) {
      if (!((cand >> q) & 1u)) continue;
      const int nj = q >> 1, p = q & 1;
      const int j = m.j0w + nj * 8 + 2 * w.tq + p;
      sink_pair(a.sink, i, j, acc[nj][0][p], acc[nj][1][p], acc[nj][2][p], acc[nj][3][p], acc[nj][4][p],
                acc[nj][5][p], acc[nj][6][p], acc[nj][7][p], acc[nj][8][p], m.xni + m.xnj[nj][p]);
    }
    return;
  }
#pragma unroll
  for (int nj = 0; nj < 2; ++nj) {
    unsigned hit[2] = {0u, 0u};
#pragma unroll
    for (int p = 0; p < 2; ++p) {
      const int j = m.j0w + nj * 8 + 2 * w.tq + p;
      bool dupl = false;
      if (cand & (1u << (nj * 2 + p))) {
        const double s = m.xni + m.xnj[nj][p];
        bool far = false;
        if (!a.rmat) {
          double r[9];
#pragma unroll
          for (int z = 0; z < 9; ++z) r[z] = acc[nj][z][p];
          const Quartic q = quartic_of(r);
          const double c1 = a.g.enantio ? -q.c1abs : -8.0 * q.det;
          far = screen_far(q, c1, s, fma(1e-8, s, a.g.E2));
        }
        if (!far) {
          const double e = pair_e_exact(acc[nj][0][p], acc[nj][1][p], acc[nj][2][p], acc[nj][3][p],
                                        acc[nj][4][p], acc[nj][5][p], acc[nj][6][p], acc[nj][7][p],
                                        acc[nj][8][p], s, a.g.enantio);
          if (a.rmat) {
            const double rr = sqrt(fmax(0.0, e) / (double)a.A);
            a.rmat[(size_t)i * (i - 1) / 2 + j] = (float)rr;
          }
          if (e < a.g.E2) {
            dupl = e < a.g.E1;
            if (!dupl)
              dupl = equalrotaniso_dev(a.rot + 3 * (size_t)i, a.rot + 3 * (size_t)j, a.aniso[i],
                                       a.aniso[j], a.g.bthr_half, a.g.bthrmax, a.g.bthrshift);
            if (dupl) atomicMax_system(a.dup + i, j + 1);
          }
        }
      }
      if (a.mask) hit[p] = __ballot_sync(0xffffffffu, dupl);
    }
    if (a.mask && w.lane == 0 && (hit[0] | hit[1])) {
      // ballot bit (row*4 + tq) of parity p  ->  tile bit row*8 + 2*tq + p
      unsigned long long word = 0;
#pragma unroll
      for (int p = 0; p < 2; ++p) {
        unsigned long long x = hit[p];
        x = (x | (x << 16)) & 0x0000FFFF0000FFFFull;
        x = (x | (x << 8)) & 0x00FF00FF00FF00FFull;
        x = (x | (x << 4)) & 0x0F0F0F0F0F0F0F0Full;
        x = (x | (x << 2)) & 0x3333333333333333ull;
        x = (x | (x << 1)) & 0x5555555555555555ull;
        word |= x << p;
      }
      const int ti8 = (m.i - w.c) >> 3, tj8 = (m.j0w + nj * 8) >> 3;
      a.mask[(size_t)ti8 * a.n8 + tj8] = word;
    }
  }
}

// Fill the meta of the unit that is about to be multiplied.  Nothing loaded here is consumed before
// the next step (the xn_i load hides under the DMMA block); only items on the band edge take the
// slow path that reads energies and gates pair by pair (reference cregen.f90:1275).
template <bool SINK>
__device__ __forceinline__ void unit_meta(const ColArgs& a, const ColWarp& w, int unit, int J,
                                          int item_edge, const double (&xnj_item)[2][2],
                                          UnitMeta& m) {
  const int i0 = unit * kUnit;
  const int i = i0 + w.wi * 8 + w.c;
  m.i = i;
  m.j0w = J * kColJ + w.wj * 16;
  const bool iok = i < a.nall;
  m.xni = a.xn[min(i, a.nall - 1)];
  unsigned ok = 0;
#pragma unroll
  for (int nj = 0; nj < 2; ++nj)
#pragma unroll
    for (int p = 0; p < 2; ++p) {
      const int j = m.j0w + nj * 8 + 2 * w.tq + p;
      m.xnj[nj][p] = xnj_item[nj][p];
      // CREGEN and intraconf: the triangle j < i; compare (sink mode 1): the whole rectangle
      const bool jok = (SINK && a.sink.mode == 1) ? j < a.sink.nj : j < i;
      ok |= ((iok && jok) ? 1u : 0u) << (nj * 2 + p);
    }
  if (item_edge) {
    // max de of the unit sits at (largest i, smallest j)
    const int i_last = min(i0 + kUnit - 1, a.nall - 1);
    if (!(__dmul_rn(__dsub_rn(a.er[i_last], a.er[J * kColJ]), kAutokcal) < a.g.ethr)) {
      const double eri = a.er[min(i, a.nall - 1)];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int j = m.j0w + (q >> 1) * 8 + 2 * w.tq + (q & 1);
        if (((ok >> q) & 1u) && !(__dmul_rn(__dsub_rn(eri, a.er[j]), kAutokcal) < a.g.ethr))
          ok &= ~(1u << q);
      }
    }
  }
  m.ok = ok;
}

template <int NKB, bool HALF, bool SCREEN, bool SINK>
__device__ __forceinline__ void pair_col_body(const ColArgs& a, unsigned char* smem_raw,
                                              uint64_t* full_bar, uint64_t* empty_bar,
                                              uint64_t& jfull_bar, uint64_t& jempty_bar,
                                              const int item_end) {
  const int warp = threadIdx.x >> 5;
  const int nstage = a.nstage;
  const int KS = a.KS;
  const size_t unit_doubles = (size_t)3 * kUnit * KS;
  const uint32_t unit_bytes = (uint32_t)(unit_doubles * sizeof(double));
  double* const sJ = reinterpret_cast<double*>(smem_raw);             // 4 units = the j-tile
  double* const sU = sJ + 4 * unit_doubles;                           // ring of i-units

  ColWarp w;
  w.lane = threadIdx.x & 31;
  w.c = w.lane >> 2; w.tq = w.lane & 3;
  w.wi = warp & 1; w.wj = warp >> 1;
  w.comp = kUnit * KS;
  w.a_off = (w.wi * 8 + w.c) * KS + 2 * w.tq;
  w.b_off0 = w.wj * (int)unit_doubles + w.c * KS + 2 * w.tq;
  w.b_off1 = w.b_off0 + 8 * KS;
  w.my_pairs = 0;

  // ---- producer state (lane 0 of warp 0): flat stream of units over this CTA's items
  const bool is_producer = (threadIdx.x == 0);
  int pk = a.item_begin + blockIdx.x, pu = 0, pu_end = 0;
  uint32_t p_it = 0;
  if (is_producer && pk < item_end) { pu = a.items[pk].u_begin; pu_end = a.items[pk].u_end; }
  auto produce_one = [&]() {
    if (pk >= item_end) return;
    const int st = p_it % nstage;
    const uint32_t ph = (p_it / nstage) & 1;
    mbar_wait(&empty_bar[st], ph ^ 1);
    mbar_expect_tx(&full_bar[st], unit_bytes);
    bulk_g2s(sU + (size_t)st * unit_doubles, a.X + (size_t)pu * unit_doubles, unit_bytes,
             &full_bar[st]);
    ++p_it;
    if (++pu == pu_end) {
      pk += gridDim.x;
      if (pk < item_end) { pu = a.items[pk].u_begin; pu_end = a.items[pk].u_end; }
    }
  };
  if (is_producer)
    for (int s = 0; s < nstage - 1; ++s) produce_one();

  // ---- consumer state: same flat stream
  int k = a.item_begin + blockIdx.x - gridDim.x, u = 0, u_end = 0, J = 0, item_edge = 0;
  double xnj_item[2][2] = {{0.0, 0.0}, {0.0, 0.0}};
  uint32_t it = 0, jn = 0;
  bool started = false;
  // advance to the next unit; at item boundaries hand the j-tile over (all warps done -> reload)
  auto next_unit = [&]() -> bool {
    if (started && u + 1 < u_end) { ++u; return true; }
    if (started) {
      __syncwarp();
      if (w.lane == 0) mbar_arrive(&jempty_bar);
    }
    k += gridDim.x;
    if (k >= item_end) return false;
    const ColItem item = a.items[k];
    J = item.J; u = item.u_begin; u_end = item.u_end; item_edge = item.edge;
#pragma unroll
    for (int nj = 0; nj < 2; ++nj)
#pragma unroll
      for (int p = 0; p < 2; ++p) {
        const int j = J * kColJ + w.wj * 16 + nj * 8 + 2 * w.tq + p;
        xnj_item[nj][p] = (SINK && a.sink.mode == 1) ? a.sink.xnj[min(j, a.sink.nj - 1)] : a.xn[min(j, a.nall - 1)];
      }
    if (is_producer) {
      mbar_wait(&jempty_bar, (jn & 1) ^ 1);      // previous j-tile released by all 8 warps
      mbar_expect_tx(&jfull_bar, 4 * unit_bytes);
      bulk_g2s(sJ, ((SINK && a.sink.mode == 1) ? a.sink.Xj : a.X) + (size_t)(4 * J) * unit_doubles, 4 * unit_bytes, &jfull_bar);
    }
    __syncwarp();
    mbar_wait(&jfull_bar, jn & 1);
    ++jn;
    started = true;
    return true;
  };

  double accA[2][9][2], accB[2][9][2];
  UnitMeta mA, mB;
  mA.i = 0; mA.j0w = 0; mA.ok = 0; mA.xni = 0.0;
  mB = mA;
#pragma unroll
  for (int y = 0; y < 2; ++y)
#pragma unroll
    for (int z = 0; z < 9; ++z) { accB[y][z][0] = accB[y][z][1] = 0.0; }
#pragma unroll
  for (int y = 0; y < 2; ++y)
#pragma unroll
    for (int p = 0; p < 2; ++p) { mA.xnj[y][p] = 0.0; mB.xnj[y][p] = 0.0; }
  const bool all_cand = a.rmat != nullptr || SINK;
  const float e2_up = a.g.E2_up;

  // one pipeline step: multiply unit (cur) while screening the previous one (prev), then finish prev
#if CREST_PAIR_PROF
  const bool prof = a.dbg != nullptr;
#else
  constexpr bool prof = false;
#endif
  long long tp = 0, tw = 0, tm = 0, te = 0, tc = 0;
  auto step = [&](double (&cur)[2][9][2], UnitMeta& mc, double (&prev)[2][9][2], UnitMeta& mp) {
    const int st = it % nstage;
    const uint32_t ph = (it / nstage) & 1;
    if (prof) tc = clock64();
    if (is_producer) produce_one();
    unit_meta<SINK>(a, w, u, J, item_edge, xnj_item, mc);
    __syncwarp();
    if (prof) { const long long t1 = clock64(); tp += t1 - tc; tc = t1; }
    mbar_wait(&full_bar[st], ph);
    if (prof) { const long long t1 = clock64(); tw += t1 - tc; tc = t1; }
    const double* su = sU + (size_t)st * unit_doubles + w.a_off;
    const unsigned cand = unit_block<NKB, HALF, SCREEN>(cur, prev, su, sJ + w.b_off0, sJ + w.b_off1, w.comp,
                                          w.comp, mp, e2_up, a.g.enantio, all_cand);
    __syncwarp();
    if (w.lane == 0) mbar_arrive(&empty_bar[st]);
    ++it;
    if (prof) { const long long t1 = clock64(); tm += t1 - tc; tc = t1; }
    w.my_pairs += __popc(mp.ok);
    if (__any_sync(0xffffffffu, cand != 0)) unit_exact<SINK>(a, w, prev, mp, cand);
    if (prof) { const long long t1 = clock64(); te += t1 - tc; tc = t1; }
  };

  for (;;) {
    if (!next_unit()) break;
    step(accA, mA, accB, mB);
    mB.ok = 0;
    if (!next_unit()) {
      // flush A: screen it without new data (the DMMAs re-read a resident stage; result unused)
      const unsigned cand = unit_block<NKB, HALF, SCREEN>(accB, accA, sU + w.a_off, sJ + w.b_off0, sJ + w.b_off1,
                                            w.comp, w.comp, mA, e2_up, a.g.enantio, all_cand);
      w.my_pairs += __popc(mA.ok);
      if (__any_sync(0xffffffffu, cand != 0)) unit_exact<SINK>(a, w, accA, mA, cand);
      mA.ok = 0;
      break;
    }
    step(accB, mB, accA, mA);
    mA.ok = 0;
  }
  if (__any_sync(0xffffffffu, mB.ok != 0)) {
    const unsigned cand = unit_block<NKB, HALF, SCREEN>(accA, accB, sU + w.a_off, sJ + w.b_off0, sJ + w.b_off1,
                                          w.comp, w.comp, mB, e2_up, a.g.enantio, all_cand);
    w.my_pairs += __popc(mB.ok);
    if (__any_sync(0xffffffffu, cand != 0)) unit_exact<SINK>(a, w, accB, mB, cand);
  }
  if (prof && w.lane == 0) {
    atomicAdd((unsigned long long*)a.dbg + warp * 4 + 0, (unsigned long long)tp);
    atomicAdd((unsigned long long*)a.dbg + warp * 4 + 1, (unsigned long long)tw);
    atomicAdd((unsigned long long*)a.dbg + warp * 4 + 2, (unsigned long long)tm);
    atomicAdd((unsigned long long*)a.dbg + warp * 4 + 3, (unsigned long long)te);
  }
  unsigned long long my_pairs = w.my_pairs;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) my_pairs += __shfl_xor_sync(0xffffffffu, my_pairs, o);
  if (w.lane == 0 && my_pairs) atomicAdd(a.npairs, my_pairs);
}

template <int NKB, bool HALF, bool SCREEN, bool SINK>
__global__ void __launch_bounds__(kPairThreads, 1)
pair_col_kernel(const ColArgs a) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[kColStagesMax];
  __shared__ __align__(8) uint64_t empty_bar[kColStagesMax];
  __shared__ __align__(8) uint64_t jfull_bar, jempty_bar;
  const int item_end = a.work_count ? a.item_begin + (int)*a.work_count : a.item_end;
  if (a.item_begin + (int)blockIdx.x >= item_end) return;
  if (threadIdx.x == 0) {
    for (int s = 0; s < a.nstage; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], kConsumerWarps);
    }
    mbar_init(&jfull_bar, 1);
    mbar_init(&jempty_bar, kConsumerWarps);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  pair_col_body<NKB, HALF, SCREEN, SINK>(a, smem_raw, full_bar, empty_bar, jfull_bar, jempty_bar, item_end);
}

int col_smem_bytes(int KS, int* nstage_out) {
  const size_t unit = (size_t)3 * kUnit * KS * sizeof(double);
  int ns = (int)(((size_t)(227 * 1024 - 1024) - 4 * unit) / unit);
  if (ns > kColStagesMax) ns = kColStagesMax;
  if (tune().stages > 0 && tune().stages < ns) ns = tune().stages;
  if (ns < 2) ns = 2;
  if (nstage_out) *nstage_out = ns;
  return (int)(unit * (4 + ns));
}

template <int NKB, bool HALF, bool SCREEN, bool SINK>
static int launch_col_inst(const ColArgs& a, int grid, int smem, cudaStream_t stream) {
  CREST_CUDA_TRY(cudaFuncSetAttribute(pair_col_kernel<NKB, HALF, SCREEN, SINK>,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  pair_col_kernel<NKB, HALF, SCREEN, SINK><<<grid, kPairThreads, smem, stream>>>(a);
  CREST_CUDA_TRY(cudaGetLastError());
  return 0;
}

template <int NKB>
static int launch_col_nkb(const ColArgs& a, int half, int grid, int smem, cudaStream_t stream) {
  if (a.sink.mode != 0)    // compare_blocks / intraconf (blockpair.cu): every valid pair to the sink
    return half ? launch_col_inst<NKB, true, true, true>(a, grid, smem, stream)
                : launch_col_inst<NKB, false, true, true>(a, grid, smem, stream);
  if (a.g.debug == 1)      // timing-only variant without the screen (CREST_CUDA_DEBUG=1)
    return half ? launch_col_inst<NKB, true, false, false>(a, grid, smem, stream)
                : launch_col_inst<NKB, false, false, false>(a, grid, smem, stream);
  return half ? launch_col_inst<NKB, true, true, false>(a, grid, smem, stream)
              : launch_col_inst<NKB, false, true, false>(a, grid, smem, stream);
}

int launch_pair_cols(const PairArgs& pa, const void* d_items, int item_begin, int item_end,
                     cudaStream_t stream, int* grid_out, int* smem_out) {
  if (grid_out) *grid_out = 0;
  const int nitems = pa.work_count ? (1 << 30) : item_end - item_begin;
  if (nitems <= 0) return 0;
  ColArgs a{};
  a.sink = pa.sink;
  a.X = pa.X; a.xn = pa.xn; a.er = pa.er; a.rot = pa.rot; a.aniso = pa.aniso;
  a.items = static_cast<const ColItem*>(d_items);
  a.item_begin = item_begin; a.item_end = item_end;
  a.nall = pa.nall; a.dup = pa.dup; a.npairs = pa.npairs; a.mask = pa.mask; a.n8 = pa.n8;
  a.rmat = pa.rmat; a.g = pa.g; a.A = pa.L.A; a.KS = pa.L.KS; a.dbg = pa.dbg; a.work_count = pa.work_count;
  int ns = 0;
  const int smem = col_smem_bytes(a.KS, &ns);
  a.nstage = ns;
  int dev = 0, nsm = 148;
  CREST_CUDA_TRY(cudaGetDevice(&dev));
  CREST_CUDA_TRY(cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev));
  const int grid = nitems < nsm ? nitems : nsm;     // persistent: one CTA per SM
  int rc = 1;
  switch (pa.L.KU / 8) {
    case 1: rc = launch_col_nkb<1>(a, pa.L.half_tail, grid, smem, stream); break;
    case 2: rc = launch_col_nkb<2>(a, pa.L.half_tail, grid, smem, stream); break;
    case 3: rc = launch_col_nkb<3>(a, pa.L.half_tail, grid, smem, stream); break;
    case 4: rc = launch_col_nkb<4>(a, pa.L.half_tail, grid, smem, stream); break;
    case 5: rc = launch_col_nkb<5>(a, pa.L.half_tail, grid, smem, stream); break;
    case 6: rc = launch_col_nkb<6>(a, pa.L.half_tail, grid, smem, stream); break;
    case 7: rc = launch_col_nkb<7>(a, pa.L.half_tail, grid, smem, stream); break;
    case 8: rc = launch_col_nkb<8>(a, pa.L.half_tail, grid, smem, stream); break;
    case 9: rc = launch_col_nkb<9>(a, pa.L.half_tail, grid, smem, stream); break;
    case 10: rc = launch_col_nkb<10>(a, pa.L.half_tail, grid, smem, stream); break;
    case 11: rc = launch_col_nkb<11>(a, pa.L.half_tail, grid, smem, stream); break;
    default: set_error("column kernel: unsupported K extent %d", pa.L.KU); return 1;
  }
  if (rc) return rc;
  if (grid_out) *grid_out = grid;
  if (smem_out) *smem_out = smem;
  return 0;
}

}  // namespace crest
