// Kernels (2)+(3), tensor-core SCREEN variant for A <= 88 atoms: the all-pairs 3x3 covariance
// blocks are contracted on the 5th-generation tensor cores (tcgen05.mma kind::tf32, accumulators
// in TMEM), every pair is screened in FP32 with a RIGOROUS error bound, and only the pairs the
// screen cannot prove "far" (duplicates and near-threshold pairs: ~1e-4 of an ensemble) are
// re-contracted in FP64 and decided exactly.
//
// Replaces the reference's pair loops K1/K2/K3 (src/cregen.f90:1266-1349, two rmsd() calls per
// pair = src/ls_rmsd.f90:72-174) and the duplicate test of K4 (src/cregen.f90:1353-1373); the
// decisions are the ones of the FP64 kernels (pair_col.cu / pair_tile.cu), because a pair leaves
// the FP32 path only with a proof.
//
// Why this is allowed: CREGEN never needs the RMSD of a far pair, only the fact
// rmsd >= 2*rthr, i.e. lambda_max(S(R)) < lam* = (s - E2)/2 with s = |x|^2+|y|^2.  With the operands
// rounded to TF32 (relative error u <= 2^-11 each) and FP32 accumulation,
//     |dR_ab| <= rho * |x_a||y_b|,  rho = 2u + u^2 + gamma  (gamma: accumulation, budget 6e-5)
//     ||dR||_F <= rho * |x||y| <= rho * s/2
//     |lambda_max(R) - lambda_max(R~)| <= ||dR||_* <= sqrt(3) ||dR||_F <= 1.78e-3 * s/2
// (S is linear in R and every eigenvalue of S(dR) is a signed sum of the singular values of dR), so
// the two-sign test of qcp.cuh (lam^2 > I1 and P(lam) > 0; screen_margin_f32x2) evaluated on R~ at
// lam = (1 - 1.8e-3) s/2 - E2/2 proves lambda_max(R) < lam*.  B200's FP64 tensor path tops out at ~36 TFLOP/s; the TF32 path is >20x
// wider, which moves the bound of this stage from the DMMA pipe to the FP32 screen arithmetic.
//
// Shape: a CTA owns one i-tile (128 conformers, all K, 3 components = 3 A operands [128 x K]) and
// streams j-groups of 16 conformers past it; a j-group is ONE B operand [48 x K] (row n = 12*s + 4*b + p
// for conformer j0 + 4*s + p), so three MMAs per K step (M = 128, N = 48, K = 8) leave, in TMEM lane i,
// the full 3x3 covariance of the 16 pairs (i, j0..j0+15): D_a[i][12 s + 4 b + p].  Operands are written by tf32_convert_kernel in the
// canonical no-swizzle K-major core-matrix layout, so one 1-D bulk copy (TMA engine) brings a
// tile/group in exactly the byte order the UMMA descriptor describes.  For K <= 64 the i-tile is moved
// once per work item from shared memory into TMEM (tcgen05.st) and the MMAs take A from TMEM: an MMA
// then reads only its 1.5 KB B slice from shared memory instead of 5.5 KB (measured: MMA path alone
// 12.8 ms -> 8.2 ms on C3).
// Warp roles: warps 0-15 epilogue (TMEM -> registers -> screen), warp 16 producer (bulk copies),
// warp 17 MMA issuer (one elected lane).  Pipelines: B ring (producer <-> MMA, mbarrier + tcgen05.commit),
// TMEM accumulator stages (MMA <-> epilogue), i-tile (producer -> epilogue warps 0-3 -> MMA).
// Measured bound: ~770 clk of packed FP32 arithmetic per j-group and sub-partition;
// the kernel runs at ~1120 clk per 2048 pairs (FMA pipe 71 % busy), limited by the instruction count of
// the epilogue loop at 4 warps per sub-partition; TMEM loads hide completely (tools/microbench).
#include "common.cuh"
#include "qcp.cuh"
#include "pipe.cuh"
#include <algorithm>
#include <cmath>

namespace crest {

constexpr int kTfEpiWarps = 16;        // (TMEM lane quarter q) x (quarter s of the 16 j's of a group)
constexpr int kTfThreads = (kTfEpiWarps + 2) * 32;
constexpr int kTfBStagesMax = 8;
constexpr int kTfAccStride = 48;       // TMEM columns between the three accumulators of a stage
constexpr int kTfStageStride = 160;    // TMEM columns between accumulator stages (3 x 48 + pad)
constexpr int kTfDStagesMax = 3;
// A operand placement.  K <= 64 ("TS"): the i-tile lives in TMEM (lane = conformer, column = atom
// slot, 3 x K columns) next to TWO accumulator stages, and an MMA reads only its 1.5 KB B slice from
// shared memory.  K > 64 ("SS"): A stays in shared memory (4 KB re-read per MMA, smem-bound) with
// THREE accumulator stages.  512 columns either way.
__host__ __device__ constexpr bool tf_ts_mode(int ksteps) { return ksteps <= 8; }
__host__ __device__ constexpr int tf_d_stages(int ksteps) { return tf_ts_mode(ksteps) ? 2 : 3; }
__host__ __device__ constexpr int tf_d_base(int ksteps) { return tf_ts_mode(ksteps) ? 192 : 0; }
constexpr int kTfCols = 512;

// ---- tcgen05 primitives (sm_100a PTX) ------------------------------------------------------
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];"
               ::"r"(smem_u32(bar)) : "memory");
}
// D[tmem] (+)= A[smem] * B[smem]^T, kind::tf32, one CTA
__device__ __forceinline__ void tc_mma_tf32(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc,
                                            uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
// same, A operand in TMEM (lane = row, 8 columns = the K extent of one TF32 MMA)
__device__ __forceinline__ void tc_mma_tf32_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t bdesc,
                                               uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n"
      "}\n" ::"r"(d_tmem), "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void tc_st8(uint32_t taddr, const uint32_t (&v)[8]) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
               ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]),
                 "r"(v[6]), "r"(v[7]) : "memory");
}
__device__ __forceinline__ void tc_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tc_ld8(uint32_t taddr, uint32_t (&v)[8]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]),
                 "=r"(v[7]) : "r"(taddr));
}
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n.reg .pred p;\nelect.sync _|p, 0xffffffff;\nselp.u32 %0, 1, 0, p;\n}" : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ void tc_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// K-major, no-swizzle shared-memory matrix descriptor (cute::UMMA::SmemDescriptor bit layout):
// core matrix = 8 rows x 16 B, contiguous (128 B); SBO = byte distance between 8-row groups,
// LBO = byte distance between the two 16-byte K chunks of one MMA; version 1 (Blackwell).
__device__ __forceinline__ uint64_t umma_desc_kmajor(uint32_t smem_addr, uint32_t lbo_bytes,
                                                     uint32_t sbo_bytes) {
  return (uint64_t)((smem_addr >> 4) & 0x3FFFu) | ((uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16) |
         ((uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32) | (1ull << 46);
}
// instruction descriptor (cute::UMMA::InstrDescriptor): D=F32, A=B=TF32, both K-major, M=128, N=48
constexpr uint32_t kTfIdesc = (1u << 4) | (2u << 7) | (2u << 10) | ((48u >> 3) << 17) | ((128u >> 4) << 24);

__device__ __forceinline__ void tc_ld4(uint32_t taddr, uint32_t (&v)[4]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]) : "r"(taddr));
}
// The 48 B rows of a j-group are ordered (j-quarter s, component b, j within the quarter), so TMEM
// column 48a + 12s + 4b + p holds R_ab(lane, j0 + 4s + p): the 36 values of a warp's 4 pairs are three
// runs of 12 columns, read with one x8 and one x4 load each.  v[3a+b][p] = R_ab of pair (lane, p).
__device__ __forceinline__ void ld_batch(uint32_t taddr, uint32_t (&v)[9][4]) {
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const uint32_t t = taddr + (uint32_t)(c * kTfAccStride);
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(v[3 * c][0]), "=r"(v[3 * c][1]), "=r"(v[3 * c][2]), "=r"(v[3 * c][3]),
                   "=r"(v[3 * c + 1][0]), "=r"(v[3 * c + 1][1]), "=r"(v[3 * c + 1][2]), "=r"(v[3 * c + 1][3])
                 : "r"(t));
    tc_ld4(t + 8u, v[3 * c + 2]);
  }
}

struct TfItem { int I; int g_begin; int g_end; int g_gate_end; };

// Row of conformer i in the candidate bitmask of a rank that owns the i-tiles I % nranks == rank: its
// owned tiles are stored back to back (tile ordinal I / nranks), so the mask holds no rows of other ranks.
__host__ __device__ __forceinline__ long long mask_row(int i, int nranks) {
  return (long long)((i >> 7) / nranks) * 128 + (i & 127);
}

struct TfArgs {
  const float* Xa;        // [tile][comp][chunk][128][4]
  const float* Xb;        // [group][chunk][48][4]
  const float* hn;        // rd(0.5*|x|^2*(1-kappa)) per conformer, zero padded to the tile grid
  const double* er;
  const TfItem* items;
  int item_begin, item_end;
  int nall;
  int KU;                 // K extent (multiple of 8); per slab in the slab kernel
  int nslab;              // K slabs (slab kernel only)
  int nbstage;
  float cE;               // -0.5 * E2_up
  double ethr;
  int enantio;
  int dbg;                // timing experiments (CREST_CUDA_DEBUG=4: no screen arithmetic, 5: no MMAs, 6: neither)
  int all_cand;           // debug: every pair is a candidate, fmask bit = screen said far
  // candidate bitmask of this pass: bit (j - cm_col0) of row (i - cm_row0), row stride
  // cm_stride 32-bit words; written with red.or, read top-down and cleared by cand_scan_kernel
  unsigned* cmask;
  unsigned* fmask;        // debug (all_cand): same geometry, bit set = the screen called the pair far
  int cm_row0, cm_col0;   // cm_row0 = mask_row of the first row the pass can touch
  int cm_nranks;
  long long cm_stride;
  float* dump;            // accumulate-error probe (debug 7): raw TMEM covariances, [i][j][9], row stride dump_n
  int dump_n;
  const int* dup;         // current links (1-based parent or 0): candidates below them are never needed
  unsigned* item_cnt;     // candidates emitted per work item (classify_items_kernel reads + clears)
  unsigned* cand_count;   // [0] exact evaluations, [1] converter "unsafe magnitudes" seen by the launch, [2] screen violations (debug), [3] candidates emitted, [4] converter flag, [5] dense items
  unsigned long long* npairs;
};

// FP32 screen of the 4 pairs (i, j0..j0+3) of this lane + candidate emission.
// DBG = false is the production instantiation: no timing switches, no self-check tagging.
// rowfast (per lane): the row exists and its half test point rd(hn_i + cE) is positive, so every lam of
// the row is (hn_j >= 0, rounding down is monotone).  The screen returns margin words whose sign bit is
// clear iff the pair is proven far (screen_margin_f32x2; finiteness is guaranteed by the converter's
// magnitude check, see tf32_convert_kernel).  Fast path (groups off the band edge and the diagonal,
// live rows): all four pairs count, and one OR + sign test decides that nothing has to be emitted.
template <bool DBG>
__device__ __forceinline__ void screen_batch(const TfArgs& a, const uint32_t (&v)[9][4], const float* hnp,
                                             int j0, int i, bool iok, f2_t hic2, bool rowfast,
                                             bool gate, bool slow, uint32_t* slot, unsigned& my_pairs) {
  if (DBG && a.dump) {     // probe of the tensor core's FP32 accumulation (tf32_accum_check_kernel)
    if (iok)
#pragma unroll
      for (int p = 0; p < 4; ++p)
        if (j0 + p < i)
#pragma unroll
          for (int z = 0; z < 9; ++z)
            a.dump[((size_t)i * a.dump_n + (j0 + p)) * 9 + z] = __uint_as_float(v[z][p]);
  }
  const float4 h0 = __ldg(reinterpret_cast<const float4*>(hnp));
  const float hj[4] = {h0.x, h0.y, h0.z, h0.w};
  uint32_t sg[4];                        // sign bit set = not proven far
  if (DBG && (a.dbg == 4 || a.dbg >= 6)) {   // timing experiment: keep the loads live, skip the screen
    unsigned x = 0;
#pragma unroll
    for (int z = 0; z < 9; ++z)
#pragma unroll
      for (int p = 0; p < 4; ++p) x ^= v[z][p];
    sg[0] = (x == 0x12345678u) ? 0x80000000u : 0u;
    sg[1] = sg[2] = sg[3] = 0u;
  } else {
#pragma unroll
    for (int p = 0; p < 4; p += 2) {     // two pairs per packed instruction (qcp.cuh)
      f2_t rp[9];
#pragma unroll
      for (int z = 0; z < 9; ++z) rp[z] = f2_pack_u(v[z][p], v[z][p + 1]);
      const f2_t lam2 = f2_add_rd(hic2, f2_pack(hj[p], hj[p + 1]));   // rd(rd(hn_i + cE) + hn_j) <= the exact sum
      const f2_t m = screen_margin_f32x2(rp, lam2, a.enantio);
      sg[p] = (uint32_t)m;
      sg[p + 1] = (uint32_t)(m >> 32);
    }
  }
  unsigned cand, farbits = 0;
  if (!DBG && !slow && rowfast) {
    // pairs of this path are counted per item in epilogue_groups
    if ((int)(sg[0] | sg[1] | sg[2] | sg[3]) >= 0) return;           // all four proven far
    cand = (sg[0] >> 31) | ((sg[1] >> 31) << 1) | ((sg[2] >> 31) << 2) | ((sg[3] >> 31) << 3);
  } else {
    // which pairs count: j < i, inside the ensemble, inside the energy band (cregen.f90:1275)
    unsigned ok = iok ? 0xFu : 0u;
    if (slow) {
      ok = 0;
      const double eri = gate && iok ? a.er[i] : 0.0;
#pragma unroll
      for (int p = 0; p < 4; ++p) {
        const int j = j0 + p;
        bool o = iok && (j < i);
        if (o && gate) o = __dmul_rn(__dsub_rn(eri, a.er[j]), kAutokcal) < a.ethr;
        ok |= (o ? 1u : 0u) << p;
      }
    }
    if (DBG || slow || !rowfast) my_pairs += __popc(ok);
    farbits = ~((sg[0] >> 31) | ((sg[1] >> 31) << 1) | ((sg[2] >> 31) << 2) | ((sg[3] >> 31) << 3)) & 0xFu;
    if (!rowfast) farbits = 0;
    cand = (DBG && a.all_cand) ? 0xFu : (~farbits & 0xFu);
    cand &= ok;
  }
  if (cand) {
    // the reference keeps the LARGEST matching j (cregen.f90:1353-1373): nothing at or below the link
    // this row already has can change it
    // (the link of the row and the item's candidate count live in this thread's two shared-memory words,
    // not in registers of the issue-bound loop and not in global memory: with a global load of dup[i] and a
    // same-address atomicAdd per emitting lane, an ensemble where most groups emit - C3adv - spent 62 of its
    // 73 ms screen time here)
    if (!(DBG && a.all_cand) && j0 + 3 < (int)slot[0]) return;
    const int rel = j0 - a.cm_col0;
    const size_t w = (size_t)(mask_row(i, a.cm_nranks) - a.cm_row0) * (size_t)a.cm_stride + (size_t)(rel >> 5);
    atomicOr(a.cmask + w, cand << (rel & 31));
    if (DBG && a.all_cand && farbits) atomicOr(a.fmask + w, (farbits & cand) << (rel & 31));
    slot[kTfEpiWarps * 32] += (unsigned)__popc(cand);
  }
}

// Epilogue of one work item, shared by both kernels: per j-group wait for the accumulator stage, pull
// this warp's 4 pairs per lane out of TMEM, hand the stage back, screen.  Barrier addresses and the
// ring position live in registers (pinned through an empty asm so that they are not rematerialised
// from special registers every group).
template <int NST, bool DBG>
__device__ __forceinline__ void epilogue_groups(const TfArgs& a, const TfItem& item, int i, bool iok, int sq, int lane,
                                                uint32_t lane_base, uint32_t dfull0, uint32_t dempty0,
                                                uint32_t& t, uint32_t& ph, int item_index, uint32_t* slot) {
  // slot[0] = the row's current link (links only change between kernels), slot[512] = candidates of this lane
  slot[0] = iok ? (uint32_t)a.dup[i] : 0u;
  slot[kTfEpiWarps * 32] = 0u;
  const float hic = __fadd_rd(a.hn[i], a.cE);
  unsigned my_pairs = 0;                                  // <= 4 * 256 per item and lane
  const f2_t hic2 = f2_pack(hic, hic);
  const bool rowfast = iok && hic > 0.0f;
  const int g_diag = item.I * 8;                          // first group with g*16 + 15 >= I*128
  // the loop runs on ONE induction variable, j0 (first j of this warp's four pairs): the group's norms sit
  // at hn + j0, and a group is "slow" (per-pair gate and/or diagonal) unless j0 lies in [f0, f0 + flen)
  const int fast_b = max(item.g_begin, item.g_gate_end), fast_e = max(fast_b, min(item.g_end, g_diag));
  if (!DBG && rowfast) my_pairs += 4u * (unsigned)(fast_e - fast_b);   // the fast path's pairs: 4 per lane and group
  const int f0 = fast_b * 16 + sq * 4;
  const unsigned flen = (unsigned)(fast_e - fast_b) * 16u;
  const int j_end = item.g_end * 16 + sq * 4;
  for (int j0 = item.g_begin * 16 + sq * 4; j0 < j_end; j0 += 16) {
    const bool slow = (unsigned)(j0 - f0) >= flen;
    const bool gate = j0 < f0;                             // below the fast range: pairs may fail the energy gate
    uint32_t v[9][4];
    mbar_wait_a(dfull0 + 8u * t, ph);
    tc_fence_after();
    ld_batch(lane_base + t * (uint32_t)kTfStageStride, v);
    tc_wait_ld();
    tc_fence_before();
    __syncwarp();
    if (elect_one()) mbar_arrive_a(dempty0 + 8u * t);    // this warp's 4 pairs are in registers
    screen_batch<DBG>(a, v, a.hn + j0, j0, i, iok, hic2, rowfast, gate, slow, slot, my_pairs);
    if (NST == 2) { ph ^= t; t ^= 1u; }
    else if (++t == (uint32_t)NST) { t = 0; ph ^= 1u; }
  }
  // one counter update per warp and item: nothing of it stays live across items
  my_pairs = __reduce_add_sync(0xffffffffu, my_pairs);
  if (lane == 0 && my_pairs) atomicAdd(a.npairs, (unsigned long long)my_pairs);
  const unsigned ncand = __reduce_add_sync(0xffffffffu, slot[kTfEpiWarps * 32]);
  if (lane == 0 && ncand) atomicAdd(a.item_cnt + item_index, ncand);
}

template <int KSTEPS, bool DBG>
__global__ void __launch_bounds__(kTfThreads, 1)
pair_tf32_kernel(const TfArgs a) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ __align__(8) uint64_t a_full, a_empty, at_full, at_empty;
  constexpr bool TS = tf_ts_mode(KSTEPS);
  constexpr int kTfDStages = tf_d_stages(KSTEPS);
  constexpr uint32_t kDBase = (uint32_t)tf_d_base(KSTEPS);
  __shared__ __align__(8) uint64_t b_full[kTfBStagesMax], b_empty[kTfBStagesMax];
  __shared__ __align__(8) uint64_t d_full[kTfDStagesMax], d_empty[kTfDStagesMax];
  __shared__ uint32_t tmem_base_smem;
  __shared__ uint32_t epi_slot[2 * kTfEpiWarps * 32];   // per epilogue thread: row link, candidate count (screen_batch)

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr int nchunk = 2 * KSTEPS;
  constexpr uint32_t a_comp_bytes = (uint32_t)nchunk * 2048u;       // one component of the i-tile
  constexpr uint32_t a_bytes = 3u * a_comp_bytes;
  constexpr uint32_t b_bytes = (uint32_t)nchunk * 768u;         // one j-group
  unsigned char* const sA = smem_raw;
  unsigned char* const sB = smem_raw + a_bytes;
  const int nbs = a.nbstage;

  if (threadIdx.x == 0) {
    mbar_init(&a_full, 1);
    mbar_init(&a_empty, TS ? 4 : 1);
    mbar_init(&at_full, 4);
    mbar_init(&at_empty, 1);
    for (int s = 0; s < nbs; ++s) { mbar_init(&b_full[s], 1); mbar_init(&b_empty[s], 1); }
    for (int t = 0; t < kTfDStages; ++t) { mbar_init(&d_full[t], 1); mbar_init(&d_empty[t], kTfEpiWarps); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == kTfEpiWarps) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                 ::"r"(smem_u32(&tmem_base_smem)), "r"((uint32_t)kTfCols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_smem;
  if (blockIdx.x == 0 && threadIdx.x == 0 && a.cand_count[4]) a.cand_count[1] = 1u;   // converter's magnitude check

  if (warp == kTfEpiWarps) {
    // ===================== producer: bulk copies of the i-tile and the j-group ring ==========
    if (lane == 0) {
      uint32_t sb = 0, phb = 0, pha = 0;
      for (int k = a.item_begin + blockIdx.x; k < a.item_end; k += gridDim.x) {
        const TfItem item = a.items[k];
        mbar_wait(&a_empty, pha ^ 1);                   // every MMA on the previous tile is done
        pha ^= 1;
        mbar_expect_tx(&a_full, a_bytes);
        bulk_g2s(sA, a.Xa + (size_t)item.I * (a_bytes >> 2), a_bytes, &a_full);
        for (int g = item.g_begin; g < item.g_end; ++g) {
          mbar_wait(&b_empty[sb], phb ^ 1);
          mbar_expect_tx(&b_full[sb], b_bytes);
          bulk_g2s(sB + (size_t)sb * b_bytes, a.Xb + (size_t)g * (b_bytes >> 2), b_bytes, &b_full[sb]);
          if (++sb == (uint32_t)nbs) { sb = 0; phb ^= 1; }
        }
      }
    }
  } else if (warp == kTfEpiWarps + 1) {
    // ===================== MMA issuer: the warp runs the control flow, ONE elected lane issues
    // 3 x KSTEPS tcgen05.mma per j-group.  Descriptors differ only in the 14-bit start-address
    // field, so every MMA of the unrolled block is base + compile-time constant.
    uint32_t sb = 0, phb = 0, td = 0, phd = 0, pha = 0;
    const uint32_t a_lo0 = ((smem_u32(sA) >> 4) & 0x3FFFu) | ((2048u >> 4) << 16);
    const uint32_t b_lo00 = ((smem_u32(sB) >> 4) & 0x3FFFu) | ((768u >> 4) << 16);
    constexpr uint32_t desc_hi = (128u >> 4) | (1u << 14);      // SBO = 128 B, version 1
    for (int k = a.item_begin + blockIdx.x; k < a.item_end; k += gridDim.x) {
      const int g_begin = a.items[k].g_begin, g_end = a.items[k].g_end;
      mbar_wait(TS ? &at_full : &a_full, pha);         // i-tile in TMEM (TS) / in shared memory (SS)
      pha ^= 1;
      for (int g = g_begin; g < g_end; ++g) {
        mbar_wait(&d_empty[td], phd ^ 1);              // epilogue drained this TMEM stage
        mbar_wait(&b_full[sb], phb);
        tc_fence_after();
        if (!(DBG && a.dbg >= 5) && elect_one()) {
          const uint32_t b_lo0 = b_lo00 + sb * (b_bytes >> 4);
          const uint32_t d0 = tmem_base + kDBase + td * kTfStageStride;
#pragma unroll
          for (int ks = 0; ks < KSTEPS; ++ks) {
            const uint64_t bdesc = ((uint64_t)desc_hi << 32) | (b_lo0 + (uint32_t)ks * (2u * 768u >> 4));
#pragma unroll
            for (int c = 0; c < 3; ++c) {
              if (TS) {
                tc_mma_tf32_ts(d0 + (uint32_t)(c * kTfAccStride),
                               tmem_base + (uint32_t)(c * 8 * KSTEPS + ks * 8), bdesc, kTfIdesc,
                               ks > 0 ? 1u : 0u);
              } else {
                const uint64_t adesc = ((uint64_t)desc_hi << 32) |
                                       (a_lo0 + (((uint32_t)c * a_comp_bytes + (uint32_t)ks * 4096u) >> 4));
                tc_mma_tf32(d0 + (uint32_t)(c * kTfAccStride), adesc, bdesc, kTfIdesc, ks > 0 ? 1u : 0u);
              }
            }
          }
          tc_commit(&b_empty[sb]);     // smem slot reusable once these MMAs have read it
          tc_commit(&d_full[td]);      // accumulators of this group complete
        } else if (DBG && a.dbg >= 5 && elect_one()) {
          tc_commit(&b_empty[sb]);
          tc_commit(&d_full[td]);
        }
        __syncwarp();
        if (++sb == (uint32_t)nbs) { sb = 0; phb ^= 1; }
        if (++td == (uint32_t)kTfDStages) { td = 0; phd ^= 1; }
      }
      if (elect_one()) tc_commit(TS ? &at_empty : &a_empty);   // i-tile reusable once every MMA of the item is done
      __syncwarp();
    }
  } else {
    // ===================== epilogue: TMEM -> registers -> FP32 screen -> candidate list ======
    // Four epilogue warps per TMEM lane quarter q, one per quarter s of a group's 16 j's: a warp
    // pulls its 4 pairs per lane (36 registers of covariance) out of TMEM, hands the accumulator
    // stage back at once and screens them.  Four warps per SM sub-partition at different points of
    // wait / load / screen keep the FMA pipe fed.
    int q = warp & 3, sq = warp >> 2, lane_p = lane;
    asm volatile("" : "+r"(q), "+r"(sq), "+r"(lane_p));   // keep them in registers (no S2R per group)
    uint32_t t = 0, ph = 0, pht = 0;
    uint32_t dfull0 = smem_u32(&d_full[0]), dempty0 = smem_u32(&d_empty[0]);
    asm volatile("" : "+r"(dfull0), "+r"(dempty0));
    const uint32_t lane_base = tmem_base + ((uint32_t)(q * 32) << 16) + kDBase + (uint32_t)(sq * 12);
    for (int k = a.item_begin + blockIdx.x; k < a.item_end; k += gridDim.x) {
      const TfItem item = a.items[k];
      const int i = item.I * 128 + q * 32 + lane;
      const bool iok = i < a.nall;
      if (TS && warp < 4) {
        // move the i-tile shared memory -> TMEM: this lane's row, 3 x K columns.  The previous
        // tile's MMAs must have completed (at_empty), the new tile must have landed (a_full).
        mbar_wait(&at_empty, pht ^ 1);
        mbar_wait(&a_full, pht);
        pht ^= 1;
        tc_fence_after();
        const int row = q * 32 + lane;
#pragma unroll
        for (int c = 0; c < 3; ++c)
#pragma unroll
          for (int ks = 0; ks < KSTEPS; ++ks) {
            const uint4 lo = *reinterpret_cast<const uint4*>(sA + ((size_t)(c * nchunk + 2 * ks) * 128 + row) * 16);
            const uint4 up = *reinterpret_cast<const uint4*>(sA + ((size_t)(c * nchunk + 2 * ks + 1) * 128 + row) * 16);
            const uint32_t w[8] = {lo.x, lo.y, lo.z, lo.w, up.x, up.y, up.z, up.w};
            tc_st8(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(c * 8 * KSTEPS + ks * 8), w);
          }
        tc_wait_st();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) { mbar_arrive(&at_full); mbar_arrive(&a_empty); }
      }
      epilogue_groups<kTfDStages, DBG>(a, item, i, iok, sq, lane_p, lane_base, dfull0, dempty0, t, ph, k,
                                       epi_slot + threadIdx.x);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kTfEpiWarps) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;"
                 ::"r"(tmem_base), "r"((uint32_t)kTfCols) : "memory");
  }
}

// ---- operand conversion: FP64 column layout X -> TF32 operands + screen norms ------------------
// X: [unit][comp][16][KS] doubles (pack_axis.cu).  One thread per (conformer, comp, 4-atom chunk).
__device__ __forceinline__ float to_tf32_rn(double x) {
  unsigned u = __float_as_uint(__double2float_rn(x));
  u = (u + 0x1000u) & 0xFFFFE000u;       // round to 10 mantissa bits (half-ulp added to the magnitude)
  return __uint_as_float(u);
}

// ---- K-slab variant for A > 88 atoms (BASELINE config C4: 300 atoms) -----------------------------
// Same decomposition, but a conformer's K extent no longer fits on chip, so K is cut into slabs of
// 48 atom slots and the accumulators of a BLOCK of three consecutive j-groups (= the three TMEM
// stages) stay live across the slab loop:
//     for slab s:  A slab (i-tile, 72 KB) -> smem ring of 2
//        for g in block:  B slab (9 KB) -> ring;  18 MMAs accumulate into stage(g)   (overwrite at s = 0)
// after the last slab each stage is committed to the epilogue, which is the same code as above.
// Operand traffic from L2 per 3 groups: nslab x (72 + 27) KB; the MMAs read A from shared memory
// (SS mode).  With 5400 FLOP per pair this shape is tensor-bound, not screen-bound.
constexpr int kSlabKsteps = 6;                       // 48 atom slots per slab
constexpr int kSlabBlock = 3;                        // j-groups per block = TMEM stages

template <bool DBG>
__global__ void __launch_bounds__(kTfThreads, 1)
pair_tf32_slab_kernel(const TfArgs a) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ __align__(8) uint64_t a_full[2], a_empty[2];
  __shared__ __align__(8) uint64_t b_full[kTfBStagesMax], b_empty[kTfBStagesMax];
  __shared__ __align__(8) uint64_t d_full[kSlabBlock], d_empty[kSlabBlock];
  __shared__ uint32_t tmem_base_smem;
  __shared__ uint32_t epi_slot[2 * kTfEpiWarps * 32];   // per epilogue thread: row link, candidate count (screen_batch)
  constexpr int nchunk = 2 * kSlabKsteps;
  constexpr uint32_t a_comp_bytes = (uint32_t)nchunk * 2048u;
  constexpr uint32_t a_bytes = 3u * a_comp_bytes;            // one slab of the i-tile
  constexpr uint32_t b_bytes = (uint32_t)nchunk * 768u;      // one slab of a j-group
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  unsigned char* const sA = smem_raw;
  unsigned char* const sB = smem_raw + 2 * a_bytes;
  const int nbs = a.nbstage, nslab = a.nslab;

  if (threadIdx.x == 0) {
    for (int s = 0; s < 2; ++s) { mbar_init(&a_full[s], 1); mbar_init(&a_empty[s], 1); }
    for (int s = 0; s < nbs; ++s) { mbar_init(&b_full[s], 1); mbar_init(&b_empty[s], 1); }
    for (int t = 0; t < kSlabBlock; ++t) { mbar_init(&d_full[t], 1); mbar_init(&d_empty[t], kTfEpiWarps); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == kTfEpiWarps) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                 ::"r"(smem_u32(&tmem_base_smem)), "r"((uint32_t)kTfCols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_smem;
  if (blockIdx.x == 0 && threadIdx.x == 0 && a.cand_count[4]) a.cand_count[1] = 1u;   // converter's magnitude check

  if (warp == kTfEpiWarps) {
    // ===================== producer =====================
    if (lane == 0) {
      uint32_t ua = 0, sb = 0, phb = 0;
      for (int k = a.item_begin + blockIdx.x; k < a.item_end; k += gridDim.x) {
        const TfItem item = a.items[k];
        for (int gb = item.g_begin; gb < item.g_end; gb += kSlabBlock) {
          const int ge = min(gb + kSlabBlock, item.g_end);
          for (int s = 0; s < nslab; ++s) {
            const uint32_t slot = ua & 1u;
            mbar_wait(&a_empty[slot], ((ua >> 1) & 1u) ^ 1u);
            mbar_expect_tx(&a_full[slot], a_bytes);
            bulk_g2s(sA + (size_t)slot * a_bytes,
                     a.Xa + ((size_t)item.I * nslab + s) * (a_bytes >> 2), a_bytes, &a_full[slot]);
            ++ua;
            for (int g = gb; g < ge; ++g) {
              mbar_wait(&b_empty[sb], phb ^ 1u);
              mbar_expect_tx(&b_full[sb], b_bytes);
              bulk_g2s(sB + (size_t)sb * b_bytes, a.Xb + ((size_t)g * nslab + s) * (b_bytes >> 2),
                       b_bytes, &b_full[sb]);
              if (++sb == (uint32_t)nbs) { sb = 0; phb ^= 1u; }
            }
          }
        }
      }
    }
  } else if (warp == kTfEpiWarps + 1) {
    // ===================== MMA issuer =====================
    uint32_t ua = 0, sb = 0, phb = 0, nd = 0;
    const uint32_t a_lo00 = ((smem_u32(sA) >> 4) & 0x3FFFu) | ((2048u >> 4) << 16);
    const uint32_t b_lo00 = ((smem_u32(sB) >> 4) & 0x3FFFu) | ((768u >> 4) << 16);
    constexpr uint32_t desc_hi = (128u >> 4) | (1u << 14);
    for (int k = a.item_begin + blockIdx.x; k < a.item_end; k += gridDim.x) {
      const int g_begin = a.items[k].g_begin, g_end = a.items[k].g_end;
      for (int gb = g_begin; gb < g_end; gb += kSlabBlock) {
        const int cnt = min(kSlabBlock, g_end - gb);
        for (int s = 0; s < nslab; ++s) {
          const uint32_t slot = ua & 1u;
          mbar_wait(&a_full[slot], (ua >> 1) & 1u);
          ++ua;
          const uint32_t a_lo0 = a_lo00 + slot * (a_bytes >> 4);
          for (int gi = 0; gi < cnt; ++gi) {
            const uint32_t my = nd + (uint32_t)gi, t = my % kSlabBlock;
            if (s == 0) mbar_wait(&d_empty[t], ((my / kSlabBlock) & 1u) ^ 1u);
            mbar_wait(&b_full[sb], phb);
            tc_fence_after();
            if (elect_one()) {
              const uint32_t b_lo0 = b_lo00 + sb * (b_bytes >> 4);
              const uint32_t d0 = tmem_base + t * kTfStageStride;
#pragma unroll
              for (int ks = 0; ks < kSlabKsteps; ++ks) {
                const uint64_t bdesc = ((uint64_t)desc_hi << 32) | (b_lo0 + (uint32_t)ks * (2u * 768u >> 4));
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                  const uint64_t adesc = ((uint64_t)desc_hi << 32) |
                                         (a_lo0 + (((uint32_t)c * a_comp_bytes + (uint32_t)ks * 4096u) >> 4));
                  tc_mma_tf32(d0 + (uint32_t)(c * kTfAccStride), adesc, bdesc, kTfIdesc,
                              (s > 0 || ks > 0) ? 1u : 0u);
                }
              }
              tc_commit(&b_empty[sb]);
              if (s == nslab - 1) tc_commit(&d_full[t]);
            }
            __syncwarp();
            if (++sb == (uint32_t)nbs) { sb = 0; phb ^= 1u; }
          }
          if (elect_one()) tc_commit(&a_empty[slot]);      // slab slot reusable once its MMAs are done
          __syncwarp();
        }
        nd += (uint32_t)cnt;
      }
    }
  } else {
    // ===================== epilogue (as in pair_tf32_kernel, three stages, D at column 0) =====
    int q = warp & 3, sq = warp >> 2, lane_p = lane;
    asm volatile("" : "+r"(q), "+r"(sq), "+r"(lane_p));   // keep them in registers (no S2R per group)
    uint32_t t = 0, ph = 0;
    uint32_t dfull0 = smem_u32(&d_full[0]), dempty0 = smem_u32(&d_empty[0]);
    asm volatile("" : "+r"(dfull0), "+r"(dempty0));
    const uint32_t lane_base = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(sq * 12);
    for (int k = a.item_begin + blockIdx.x; k < a.item_end; k += gridDim.x) {
      const TfItem item = a.items[k];
      const int i = item.I * 128 + q * 32 + lane;
      const bool iok = i < a.nall;
      epilogue_groups<kSlabBlock, DBG>(a, item, i, iok, sq, lane_p, lane_base, dfull0, dempty0, t, ph, k,
                                       epi_slot + threadIdx.x);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kTfEpiWarps) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;"
                 ::"r"(tmem_base), "r"((uint32_t)kTfCols) : "memory");
  }
}

// FP64 operand as pack_axis wrote it: chunks [panel][slab][comp][rows][KS]; column layout (A <= 88):
// rows = 16, one slab; row-tile layout (A > 88): rows = 32, slabs of KU atom slots.
struct XView {
  const double* X;
  int rows, nslab, KU, KS, npanel;
  __device__ __forceinline__ const double* row(int i, int comp, int slab) const {
    return X + ((((size_t)(i / rows) * nslab + slab) * 3 + comp) * rows + (i % rows)) * KS;
  }
};

// One thread per (conformer, TF32 slab, comp, 4-slot chunk).  TF32 slot t = FP64 slot t (slabs of the
// FP64 layout concatenated); kslab slots per TF32 slab, zero beyond the FP64 extent.
__global__ void __launch_bounds__(256)
tf32_convert_kernel(const XView xv, const double* __restrict__ xn, int nall, int tile0, int ntile, int nslab_t,
                    int kslab, float* __restrict__ Xa, float* __restrict__ Xb,
                    float* __restrict__ hn, double kappa, unsigned* __restrict__ unsafe) {
  const int nchunk = kslab >> 2;
  const long long total = (long long)ntile * 128 * 3 * nchunk * nslab_t;
  const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (tid < (long long)ntile * 128) {
    const int i = tile0 * 128 + (int)tid;
    const float h = i < nall ? __double2float_rd(0.5 * xn[i] * (1.0 - kappa)) : 0.0f;
    hn[i] = h;
    // the screen's sign tests assume finite FP32 arithmetic: lam = hn_i + hn_j + cE < 1e9 keeps every
    // intermediate (<= lam^4) below 1e36.  A structure beyond that (or a NaN) sends the whole launch
    // to the FP64 kernel through the overflow flag.
    if (!(h < 5.0e8f)) *unsafe = 1u;
  }
  if (tid >= total) return;
  const int r = (int)(tid % 128);
  long long rest = tid / 128;
  const int c = (int)(rest % nchunk); rest /= nchunk;
  const int comp = (int)(rest % 3); rest /= 3;
  const int sl = (int)(rest % nslab_t);
  const int I = tile0 + (int)(rest / nslab_t);
  const int i = I * 128 + r;
  float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
  const int t = sl * kslab + 4 * c;                   // first of 4 consecutive slots (same FP64 slab)
  if (i < nall && i / xv.rows < xv.npanel && t < xv.nslab * xv.KU) {
    const double* src = xv.row(i, comp, t / xv.KU) + (t % xv.KU);
    const double2 p0 = *reinterpret_cast<const double2*>(src);
    const double2 p1 = *reinterpret_cast<const double2*>(src + 2);
    o = make_float4(to_tf32_rn(p0.x), to_tf32_rn(p0.y), to_tf32_rn(p1.x), to_tf32_rn(p1.y));
  }
  reinterpret_cast<float4*>(Xa)[((((size_t)I * nslab_t + sl) * 3 + comp) * nchunk + c) * 128 + r] = o;
  const int g = i >> 4, jj = i & 15;
  reinterpret_cast<float4*>(Xb)[(((size_t)g * nslab_t + sl) * nchunk + c) * 48 + (jj >> 2) * 12 + comp * 4 + (jj & 3)] = o;
}

// ---- exact pass over the candidate bitmask: top-down per row, FP64 covariance, Newton/Jacobi, gates ----
// The reference's K4 (src/cregen.f90:1353-1373) keeps, per row i, the LARGEST in-band j that matches
// (`double(i)=j` without leaving the loop), so a row is decided by its candidates in DESCENDING j: the
// first exact match ends the row, and nothing at or below the link the row already has (earlier pass,
// other kernel) is looked at.  One warp per row: the row's words of this pass are read 128 at a time
// (512 B per warp load), non-zero words are cleared on the way (the mask is all-zero again when the
// kernel ends - no memset per call) and their set bits evaluated four at a time, 8 lanes per candidate.
// Decisions are those of pair_col.cu's unit_exact (same second-level screen, same pair_e_exact, same
// float32-threshold emulation and rot gate).  Words of items the classifier handed to the DMMA kernel
// (dense != 0) are cleared without evaluation.
struct ScanArgs {
  XView xv; const double* xn; const double* rot; const double* aniso;
  unsigned* cmask; unsigned* fmask;
  int row0, col0, col_end, nall;      // rows [row0, nall) x columns [col0, col_end) of this pass
  int mrow0, nranks, rank;            // mask_row of row0; rows of i-tiles I % nranks == rank only
  long long stride;                   // words per row (multiple of 128)
  const unsigned char* dense;         // [segment][i-tile] != 0: candidate-dense item (classify_items_kernel)
  unsigned char* unfinished;          // [row] set when a row gives up inside a dense item (-> DMMA re-run)
  int dense_budget;                   // exact evaluations a row may spend inside dense items: > 0 fixed,
                                      // 0 none (every dense item re-done), < 0 adaptive (see the kernel)
  int segcols, ntile;                 // columns per segment (multiple of 128), i-tiles
  Gates g;
  int* dup;
  unsigned* counters;                 // [0] += exact evaluations, [2] += screen violations (debug)
  int debug_all;                      // self-check: evaluate every candidate, no pruning, count violations
};

// exact decision for candidate (i, j) by the 8 lanes of group `grp` (sub = lane & 7); returns on the
// sub == 0 lane: bit 0 = duplicate link, bit 1 = exact e < E2 (for the self-check)
__device__ __forceinline__ unsigned exact_pair8(const ScanArgs& a, int i, int j, bool live, int sub,
                                                unsigned gmask) {
  const size_t cs = (size_t)a.xv.rows * a.xv.KS;         // distance between components
  double r[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
  if (live)
    for (int sd = 0; sd < a.xv.nslab; ++sd) {
      const double* xi = a.xv.row(i, 0, sd);
      const double* xj = a.xv.row(j, 0, sd);
      for (int k = sub; k < a.xv.KU; k += 8) {
        const double x0 = xi[k], x1 = xi[cs + k], x2 = xi[2 * cs + k];
        const double y0 = xj[k], y1 = xj[cs + k], y2 = xj[2 * cs + k];
        r[0] = fma(x0, y0, r[0]); r[1] = fma(x0, y1, r[1]); r[2] = fma(x0, y2, r[2]);
        r[3] = fma(x1, y0, r[3]); r[4] = fma(x1, y1, r[4]); r[5] = fma(x1, y2, r[5]);
        r[6] = fma(x2, y0, r[6]); r[7] = fma(x2, y1, r[7]); r[8] = fma(x2, y2, r[8]);
      }
    }
#pragma unroll
  for (int z = 0; z < 9; ++z) {
    r[z] += __shfl_xor_sync(gmask, r[z], 1);
    r[z] += __shfl_xor_sync(gmask, r[z], 2);
    r[z] += __shfl_xor_sync(gmask, r[z], 4);
  }
  unsigned res = 0;
  if (live && sub == 0) {
    const double s = a.xn[i] + a.xn[j];
    const Quartic q = quartic_of(r);
    const double c1 = a.g.enantio ? -q.c1abs : -8.0 * q.det;
    if (!screen_far(q, c1, s, fma(1e-8, s, a.g.E2))) {
      const double e = pair_e_exact(r[0], r[1], r[2], r[3], r[4], r[5], r[6], r[7], r[8], s, a.g.enantio);
      if (e < a.g.E2) {
        res = 2u;
        bool dupl = e < a.g.E1;
        if (!dupl)
          dupl = equalrotaniso_dev(a.rot + 3 * (size_t)i, a.rot + 3 * (size_t)j, a.aniso[i], a.aniso[j],
                                   a.g.bthr_half, a.g.bthrmax, a.g.bthrshift);
        if (dupl) res |= 1u;
      }
    }
  }
  return res;
}

// The same decision for EVERY set bit of one 32-column word at once, one lane per candidate (bit b -> lane b):
// each lane contracts its own pair serially (the row of i is a broadcast load).  One such step costs about
// three of the 8-lanes-per-candidate steps above and decides up to 32 candidates instead of 4 - the path of
// rows that keep missing inside a block of near-identical structures, where the one-warp-per-row scan used to
// end in a tail of rows with hundreds of sequential steps (C3dense: ~1 ms per launch on any number of GPUs).
// Returns the ballot of duplicate links.
__device__ __noinline__ unsigned exact_word32(const ScanArgs& a, int i, int cbase, unsigned livebits, int lane) {
  const bool live = (livebits >> lane) & 1u;
  const int j = cbase + lane;
  bool dupl = false;
  if (live) {
    const size_t cs = (size_t)a.xv.rows * a.xv.KS;
    double r[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    for (int sd = 0; sd < a.xv.nslab; ++sd) {
      const double* xi = a.xv.row(i, 0, sd);
      const double* xj = a.xv.row(j, 0, sd);
#pragma unroll 2
      for (int k = 0; k < a.xv.KU; ++k) {
        const double x0 = xi[k], x1 = xi[cs + k], x2 = xi[2 * cs + k];
        const double y0 = xj[k], y1 = xj[cs + k], y2 = xj[2 * cs + k];
        r[0] = fma(x0, y0, r[0]); r[1] = fma(x0, y1, r[1]); r[2] = fma(x0, y2, r[2]);
        r[3] = fma(x1, y0, r[3]); r[4] = fma(x1, y1, r[4]); r[5] = fma(x1, y2, r[5]);
        r[6] = fma(x2, y0, r[6]); r[7] = fma(x2, y1, r[7]); r[8] = fma(x2, y2, r[8]);
      }
    }
    const double s = a.xn[i] + a.xn[j];
    const Quartic q = quartic_of(r);
    const double c1 = a.g.enantio ? -q.c1abs : -8.0 * q.det;
    if (!screen_far(q, c1, s, fma(1e-8, s, a.g.E2))) {
      const double e = pair_e_exact(r[0], r[1], r[2], r[3], r[4], r[5], r[6], r[7], r[8], s, a.g.enantio);
      if (e < a.g.E2) {
        dupl = e < a.g.E1;
        if (!dupl)
          dupl = equalrotaniso_dev(a.rot + 3 * (size_t)i, a.rot + 3 * (size_t)j, a.aniso[i], a.aniso[j],
                                   a.g.bthr_half, a.g.bthrmax, a.g.bthrshift);
      }
    }
  }
  return __ballot_sync(0xffffffffu, dupl);
}

template <int kMinBlocks>
__global__ void __launch_bounds__(256, kMinBlocks)
cand_scan_kernel(const ScanArgs a) {
  const int lane = threadIdx.x & 31, sub = lane & 7, grp = lane >> 3;
  const unsigned gmask = 0xFFu << (lane & 24);
  // rows are handed out through a queue (counters[10], zeroed per launch), longest first: a row that keeps
  // missing inside a dense block is hundreds of sequential steps, and with rows dealt round-robin the slowest
  // SM of a C3dense launch was busy 5.0 M cycles against 2.9 M on average (profiles/r02_h_scan_dense_ncu.txt)
  unsigned n_eval = 0, n_viol = 0;
  // owned rows only, longest first (descending i), round-robin over the warps of the grid
  const long long mtop = mask_row(a.nall - 1, a.nranks) + 128;       // upper bound of the compact row index
  for (;;) {
    unsigned ticket = 0;
    if (lane == 0) ticket = atomicAdd(a.counters + 10, 1u);
    ticket = __shfl_sync(0xffffffffu, ticket, 0);
    const long long mr = mtop - 1 - (long long)ticket;
    if (mr < a.mrow0) break;
    const int i = (int)(((mr >> 7) * a.nranks + a.rank) << 7) + (int)(mr & 127);
    if (i >= a.nall || i <= a.col0 || i < a.row0) continue;
    const int jend = min(i, a.col_end);                       // columns [col0, jend) hold pairs of row i
    if (jend <= a.col0) continue;
    unsigned* const rowp = a.cmask + (size_t)(mr - a.mrow0) * (size_t)a.stride;
    const unsigned* const frow = a.fmask ? a.fmask + (size_t)(mr - a.mrow0) * (size_t)a.stride : nullptr;
    const int m0 = a.dup[i];
    int m = m0;                                               // best link so far (j + 1)
    bool done = false;                                        // row decided: only clearing remains
    // Evaluations this row may spend inside candidate-dense items before its share goes to the FP64 tensor
    // kernel.  Measured: where rows DO end in a match (collapsed ensembles, extended chains
    // with many copies) scanning on is 2x cheaper than the re-run even after thousands of misses; where
    // nothing matches (near-miss ensemble) every evaluation is wasted.  The rows of a launch tell which
    // case it is: counters[8] rows that met a dense item, counters[9] those that still ended in a match.
    // Start with a probe of 512; once 256 rows have reported, rows get 4096 while >= 1/8 of them succeed
    // and 64 otherwise.  Budgets only steer work between two exact paths - the links do not depend on them.
    int budget = a.dense_budget;
    if (budget < 0) {
      const unsigned tried = a.counters[8], won = a.counters[9];
      budget = tried < 256u ? 512 : (won * 8u >= tried ? 4096 : 64);
    }
    bool gave_up = false, met_dense = false;
    int misses = 0;                                           // candidates of this row evaluated without a match
    const unsigned char* const dense_i = a.dense ? a.dense + (i >> 7) : nullptr;
    // The row's words are read 8 blocks (8 x 512 B per warp) at a time: the loads of a batch are independent,
    // and nearly every block is empty - read one by one, the dependent DRAM round trips (12 per row on C3,
    // 3552 resident warps) were most of this kernel's time.
    constexpr int kScanBatch = 8;
    constexpr int kScanWideAfter = 8, kScanWideBits = 6;     // wide steps after 8 misses, for words with >= 6 candidates
    for (int btop = (jend - 1 - a.col0) >> 12; btop >= 0; btop -= kScanBatch) {     // block = 128 words = 4096 columns
      unsigned nonempty = 0u;
      {
        unsigned orw[kScanBatch];
#pragma unroll
        for (int t = 0; t < kScanBatch; ++t) {
          orw[t] = 0u;
          if (btop - t >= 0) {
            const uint4 v = *(reinterpret_cast<const uint4*>(rowp + (size_t)(btop - t) * 128) + lane);
            orw[t] = v.x | v.y | v.z | v.w;
          }
        }
#pragma unroll
        for (int t = 0; t < kScanBatch; ++t)
          if (__any_sync(0xffffffffu, orw[t] != 0u)) nonempty |= 1u << t;
      }
      while (nonempty) {
      const int tsel = __ffs(nonempty) - 1;                  // highest block of the batch first
      nonempty &= nonempty - 1u;
      const int blk = btop - tsel;
      uint4* const wp = reinterpret_cast<uint4*>(rowp + (size_t)blk * 128) + lane;
      uint4 w = *wp;                                         // a rare re-read (L2): the batch kept one OR word per block
      const bool nz = (w.x | w.y | w.z | w.w) != 0u;
      unsigned any = __ballot_sync(0xffffffffu, nz);
      if (nz) *wp = make_uint4(0u, 0u, 0u, 0u);
      if (done) continue;
      while (any && !done) {
        const int src = 31 - __clz(any);
        any &= ~(1u << src);
        const unsigned w4x = __shfl_sync(0xffffffffu, w.x, src), w4y = __shfl_sync(0xffffffffu, w.y, src);
        const unsigned w4z = __shfl_sync(0xffffffffu, w.z, src), w4w = __shfl_sync(0xffffffffu, w.w, src);
        // ONE copy of the evaluation code: unrolled over the four words the kernel was 129 KB of SASS and a third
        // of its stall samples were instruction fetches (`no_instruction`, profiles/r02_d_scan_ncu.txt)
#pragma unroll 1
        for (int q = 3; q >= 0 && !done; --q) {
          unsigned word = q == 3 ? w4w : (q == 2 ? w4z : (q == 1 ? w4y : w4x));
          if (!word) continue;
          const int widx = blk * 128 + src * 4 + q;
          const int cbase = a.col0 + widx * 32;
          if (!a.debug_all && cbase + 32 <= m) { done = true; break; }               // whole word below the link
          const bool in_dense = dense_i && dense_i[(size_t)(cbase / a.segcols) * a.ntile];
          met_dense |= in_dense;
          if (in_dense && budget <= 0) {       // this row's share of the item goes to the FP64 tensor kernel
            gave_up = true;
            continue;
          }
          const unsigned fword = frow ? frow[widx] : 0u;
          while (word && !done) {
            if (!a.debug_all && misses >= kScanWideAfter && __popc(word) >= kScanWideBits) {
              // wide step: all remaining candidates of the word, one lane each
              const int below = m - cbase;                    // bits < below are at or under the current link
              const unsigned livebits = below <= 0 ? word : (below >= 32 ? 0u : word & (0xFFFFFFFFu << below));
              const unsigned hits = livebits ? exact_word32(a, i, cbase, livebits, lane) : 0u;
              if ((livebits >> lane) & 1u) ++n_eval;
              if (hits) m = cbase + (31 - __clz(hits)) + 1;
              if (in_dense) budget -= __popc(livebits);
              misses += __popc(livebits);
              if (hits || livebits != word) done = true;      // a match, or candidates under the link: the row is decided
              word = 0u;
              continue;
            }
            // the four highest candidates of the word, one per 8-lane group
            int bsel = -1;
#pragma unroll
            for (int t = 0; t < 4; ++t) {
              const int bit = word ? 31 - __clz(word) : -1;
              if (bit >= 0) word &= ~(1u << bit);
              if (t == grp) bsel = bit;
            }
            const int j = cbase + bsel;
            // descending order: once a candidate is at or below the current link, so is everything after it
            const bool valid = bsel >= 0;
            const bool pruned = valid && !a.debug_all && j + 1 <= m;
            const bool live = valid && !pruned;
            const unsigned res = exact_pair8(a, i, live ? j : 0, live, sub, gmask);
            if (live && sub == 0) {
              ++n_eval;
              if (a.debug_all && (res & 2u) && ((fword >> bsel) & 1u)) ++n_viol;
            }
            int hit = (live && sub == 0 && (res & 1u)) ? j + 1 : 0;
            hit = __reduce_max_sync(0xffffffffu, hit);
            if (hit > m) m = hit;
            if (in_dense) budget -= 4;
            misses += 4;
            // a match ends the row (the reference keeps the largest j), and so does a pruned candidate
            if (!a.debug_all && (hit > 0 || __any_sync(0xffffffffu, pruned))) done = true;
          }
        }
      }
      }   // while (nonempty)
    }
    if (lane == 0 && m > m0) atomicMax_system(a.dup + i, m);
    // a row that skipped candidates inside a dense item needs the FP64 tensor kernel for them even if it
    // found a match further down: the skipped ones sit at HIGHER j
    if (lane == 0 && gave_up) a.unfinished[i] = 1;
    if (lane == 0 && met_dense && a.dense_budget < 0) {
      atomicAdd(a.counters + 8, 1u);
      if (!gave_up) atomicAdd(a.counters + 9, 1u);
    }
  }
  n_eval = __reduce_add_sync(0xffffffffu, n_eval);
  n_viol = __reduce_add_sync(0xffffffffu, n_viol);
  if (lane == 0) {
    if (n_eval) atomicAdd(a.counters + 0, n_eval);
    if (n_viol) atomicAdd(a.counters + 2, n_viol);
  }
}

// ---- per-item verdict after the screen, dense leftovers on the FP64 tensor kernel --------------------------
// An item whose candidates exceed 1/16 of its pairs (a block of near-identical structures, or a molecule so
// large that the TF32 bound decides little) is flagged DENSE.  Measured (C3dense, C4ext):
// re-doing every dense item as a DMMA contraction costs more than the top-down scan, because nearly every
// row of such an item ends at its first exact match.  So the scan gets a small budget of evaluations per
// row inside dense items; only rows that spend it without a match ("unfinished": many near, non-matching
// neighbours) send their 16-row unit of the item to pair_col_kernel ({J, u_begin, u_end, edge} items) or
// their 32-row panel to pair_tile_kernel ({panel, j-tile} tiles) - the worst case is bounded by the FP64
// tensor kernel over the dense items.  A launch whose converter saw unsafe magnitudes (FP32 screen
// arithmetic not guaranteed finite) sends every unit of every item there.  Item counters are cleared for
// the next launch.
struct ClassifyArgs {
  const TfItem* items; int item_begin, item_end;
  unsigned* item_cnt;
  unsigned char* dense; int chunk, ntile;
  const unsigned char* unfinished;
  int col_layout;          // 1: column items for pair_col, 0: (panel, j-tile) tiles for pair_tile
  int4* out_items; unsigned out_cap;
  unsigned* counters;      // [3] += candidates, [4] converter flag (read), [5] += dense items, [6] list length, [7] list overflow
  int nunits;              // 16-row units of the column layout / 32-row panels of the row-tile layout
  int nall;
  int force_dense;         // developer knob: 1 every item dense and re-done, -1 none (except unsafe magnitudes)
};

__global__ void __launch_bounds__(256)
classify_items_kernel(const ClassifyArgs a) {
  const int k = a.item_begin + blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= a.item_end) return;
  const TfItem it = a.items[k];
  const unsigned c = a.item_cnt[k];
  if (c) { a.item_cnt[k] = 0u; atomicAdd(a.counters + 3, c); }
  const unsigned pairs = 128u * 16u * (unsigned)(it.g_end - it.g_begin);
  const bool all = a.force_dense > 0 || a.counters[4] != 0u;
  const bool dense = all || (a.force_dense == 0 && c > pairs / 16u);
  a.dense[(size_t)(it.g_begin / a.chunk) * a.ntile + it.I] = all ? 2 : (dense ? 1 : 0);
  if (dense) atomicAdd(a.counters + 5, 1u);
}

// after the scan: one thread per dense item; runs of consecutive units that still need work become one
// column item per j-tile (the column kernel streams the i-units of an item past a resident j-tile)
__global__ void __launch_bounds__(256)
expand_dense_kernel(const ClassifyArgs a) {
  const int per = a.col_layout ? 8 : 4, rows = a.col_layout ? 16 : 32;
  const int k = a.item_begin + blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= a.item_end) return;
  const TfItem it = a.items[k];
  const unsigned char d = a.dense[(size_t)(it.g_begin / a.chunk) * a.ntile + it.I];
  if (!d) return;
  unsigned need = 0;
  for (int sub = 0; sub < per; ++sub) {
    const int r = it.I * per + sub;                    // unit / panel index
    if (r >= a.nunits) break;
    bool nd = d == 2;
    if (!nd) {
      const int i0 = r * rows, i1 = min(i0 + rows, a.nall);
      for (int i = i0; i < i1 && !nd; ++i) nd = a.unfinished[i] != 0;
    }
    if (nd) need |= 1u << sub;
  }
  if (!need) return;
  const int J0 = (it.g_begin * 16) / 64, J1 = (it.g_end * 16 - 1) / 64;
  const unsigned nJ = (unsigned)(J1 - J0 + 1);
  unsigned nrun = 0;
  if (a.col_layout) { for (int sub = 0; sub < per; ++sub) nrun += ((need >> sub) & 1u) && !(sub && ((need >> (sub - 1)) & 1u)); }
  else nrun = __popc(need);
  const unsigned n = nrun * nJ;
  unsigned at = atomicAdd(a.counters + 6, n);
  if (at + n > a.out_cap) { atomicSub(a.counters + 6, n); a.counters[7] = 1u; return; }
  for (int sub = 0; sub < per; ++sub) {
    if (!((need >> sub) & 1u)) continue;
    const int r0 = it.I * per + sub;
    if (a.col_layout) {
      int e = sub;
      while (e + 1 < per && ((need >> (e + 1)) & 1u)) ++e;
      for (int J = J0; J <= J1; ++J) a.out_items[at++] = make_int4(J, r0, it.I * per + e + 1, 1);
      sub = e;
    } else {
      for (int J = J0; J <= J1; ++J) a.out_items[at++] = make_int4(r0, J, 0, 0);
    }
  }
}

// ---- probe of the tensor core's FP32 accumulation (tools/gpu_tf32_accum.py) ----------
// The screen's bound budgets gamma for what tcgen05.mma kind::tf32 loses while it accumulates K products
// in FP32 - a property of the hardware that is not architecturally specified.  This kernel measures it:
// for every pair (i > j) and every one of the nine components it recomputes the dot product of the SAME
// TF32-rounded operands (read back from the A operand copy) exactly (products of two TF32 numbers fit
// FP64 with room to spare) and reports the largest |R~_ab - R_ab| / (|x_a| |y_b|), the quantity gamma
// bounds.  One thread per pair.
struct AccumArgs {
  const float* Xa; int nslab, kslab; int n;
  const float* dump; int dump_n;
  unsigned long long* out;   // [0] max rel error (bits of a double), [1] dot products checked
};
__device__ __forceinline__ float xa_at(const AccumArgs& a, int i, int comp, int t) {
  const int I = i >> 7, r = i & 127, sl = t / a.kslab, tt = t % a.kslab, nchunk = a.kslab >> 2;
  return a.Xa[(((((size_t)I * a.nslab + sl) * 3 + comp) * nchunk + (tt >> 2)) * 128 + r) * 4 + (tt & 3)];
}
__global__ void __launch_bounds__(256)
tf32_accum_check_kernel(const AccumArgs a) {
  const long long total = (long long)a.n * a.n;
  double worst = 0.0;
  unsigned long long cnt = 0;
  for (long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x; p < total; p += (long long)gridDim.x * blockDim.x) {
    const int i = (int)(p / a.n), j = (int)(p % a.n);
    if (j >= i) continue;
    const int K = a.nslab * a.kslab;
    double r[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}, nx[3] = {0, 0, 0}, ny[3] = {0, 0, 0};
    for (int t = 0; t < K; ++t) {
      double x[3], y[3];
#pragma unroll
      for (int c = 0; c < 3; ++c) { x[c] = (double)xa_at(a, i, c, t); y[c] = (double)xa_at(a, j, c, t); }
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        nx[c] = fma(x[c], x[c], nx[c]); ny[c] = fma(y[c], y[c], ny[c]);
#pragma unroll
        for (int d = 0; d < 3; ++d) r[3 * c + d] = fma(x[c], y[d], r[3 * c + d]);
      }
    }
#pragma unroll
    for (int c = 0; c < 3; ++c)
#pragma unroll
      for (int d = 0; d < 3; ++d) {
        const double got = (double)a.dump[((size_t)i * a.dump_n + j) * 9 + 3 * c + d];
        const double den = sqrt(nx[c] * ny[d]);
        if (den > 0.0) worst = fmax(worst, fabs(got - r[3 * c + d]) / den);
        ++cnt;
      }
  }
  for (int o = 16; o > 0; o >>= 1) {
    worst = fmax(worst, __shfl_xor_sync(0xffffffffu, worst, o));
    cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
  }
  if ((threadIdx.x & 31) == 0) {
    atomicMax(a.out, (unsigned long long)__double_as_longlong(worst));     // non-negative doubles order as integers
    atomicAdd(a.out + 1, cnt);
  }
}

int launch_tf32_accum_check(const Tf32Buffers& tb, int nall, const float* d_dump, unsigned long long* d_out,
                            cudaStream_t stream) {
  AccumArgs a{};
  a.Xa = tb.Xa; a.nslab = tb.nslab; a.kslab = tb.kslab; a.n = nall; a.dump = d_dump; a.dump_n = nall; a.out = d_out;
  tf32_accum_check_kernel<<<148 * 16, 256, 0, stream>>>(a);
  CREST_CUDA_TRY(cudaGetLastError());
  return 0;
}

// ---- host side -------------------------------------------------------------------------------
int tf32_smem_bytes(int KU, int* nbstage_out) {
  const size_t a_bytes = (size_t)3 * (KU / 4) * 2048, b_bytes = (size_t)(KU / 4) * 768;
  int ns = (int)(((size_t)(227 * 1024 - 2048) - a_bytes) / b_bytes);
  if (ns > kTfBStagesMax) ns = kTfBStagesMax;
  if (tune().stages > 0 && tune().stages < ns) ns = tune().stages;
  if (nbstage_out) *nbstage_out = ns;
  return (int)(a_bytes + (size_t)ns * b_bytes);
}

static XView make_xview(const double* d_X, const PackLayout& L) {
  XView v;
  v.X = d_X; v.rows = L.rows; v.nslab = L.nslab; v.KU = L.KU; v.KS = L.KS; v.npanel = L.npanel;
  return v;
}

// Error budget of the screen for a K extent of `kslots` atom slots: operands rounded to
// TF32 by us (u), products exact in FP32, FP32 accumulation over the chained MMAs budgeted at
// gamma = max(6e-5, 2 * kslots * 2^-23) (one truncated add per product, twice the truncation ulp);
// kappa = sqrt(3) * (2u + u^2 + gamma) with 0.2 % slack, never below the 1.8e-3 of the A <= 88 analysis.
double tf32_kappa(int kslots) {
  const double u = std::ldexp(1.0, -11) * (1.0 + std::ldexp(1.0, -12));
  const double gamma = std::max(6.0e-5, 2.0 * (double)kslots * std::ldexp(1.0, -23));
  const double k = 1.002 * std::sqrt(3.0) * (2.0 * u + u * u + gamma);
  return k > 1.8e-3 ? k : 1.8e-3;
}

int launch_tf32_convert(const double* d_X, const double* d_xn, int nall, const PackLayout& L,
                        int tile0, int tile1, int nslab_t, int kslab, double kappa, float* d_Xa,
                        float* d_Xb, float* d_hn, unsigned* d_unsafe, cudaStream_t stream) {
  const int ntile = tile1 - tile0;
  const long long total = (long long)ntile * 128 * 3 * (kslab / 4) * nslab_t;
  if (total <= 0) return 0;
  const int blocks = (int)((total + 255) / 256);
  tf32_convert_kernel<<<blocks, 256, 0, stream>>>(make_xview(d_X, L), d_xn, nall, tile0, ntile, nslab_t, kslab,
                                                  d_Xa, d_Xb, d_hn, kappa, d_unsafe);
  CREST_CUDA_TRY(cudaGetLastError());
  return 0;
}

template <int KSTEPS>
static int launch_tf32_inst(const TfArgs& a, int grid, int smem, cudaStream_t stream) {
  if (a.dbg != 0 || a.all_cand) {
    CREST_CUDA_TRY(cudaFuncSetAttribute(pair_tf32_kernel<KSTEPS, true>,
                                        cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    pair_tf32_kernel<KSTEPS, true><<<grid, kTfThreads, smem, stream>>>(a);
    CREST_CUDA_TRY(cudaGetLastError());
    return 0;
  }
  CREST_CUDA_TRY(cudaFuncSetAttribute(pair_tf32_kernel<KSTEPS, false>,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  pair_tf32_kernel<KSTEPS, false><<<grid, kTfThreads, smem, stream>>>(a);
  CREST_CUDA_TRY(cudaGetLastError());
  return 0;
}

// words of the candidate bitmask of a pass over columns [col0, col_end): rows [col0, nall), row stride
// rounded up to 128 words (one warp-wide uint4 load of cand_scan_kernel)
static long long mask_stride(int col0, int col_end) {
  const long long w = ((long long)(col_end - col0) + 31) / 32;
  return (w + 127) / 128 * 128;
}
size_t tf32_mask_words(int row_lo, int row_hi, int col0, int col_end, int nranks) {
  const int r0 = row_lo > col0 ? row_lo : col0;
  if (row_hi <= r0) return 0;
  const long long rows = mask_row(row_hi - 1, nranks) + 128 - ((long long)((r0 >> 7) / nranks) * 128);
  return (size_t)rows * (size_t)mask_stride(col0, col_end);
}

static int launch_tf32_screen(const TfArgs& a, bool slab, int grid, int smem, cudaStream_t stream) {
  if (slab) {
    if (a.dbg != 0 || a.all_cand) {
      CREST_CUDA_TRY(cudaFuncSetAttribute(pair_tf32_slab_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
      pair_tf32_slab_kernel<true><<<grid, kTfThreads, smem, stream>>>(a);
    } else {
      CREST_CUDA_TRY(cudaFuncSetAttribute(pair_tf32_slab_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
      pair_tf32_slab_kernel<false><<<grid, kTfThreads, smem, stream>>>(a);
    }
    CREST_CUDA_TRY(cudaGetLastError());
    return 0;
  }
  switch (a.KU / 8) {
    case 1: return launch_tf32_inst<1>(a, grid, smem, stream);
    case 2: return launch_tf32_inst<2>(a, grid, smem, stream);
    case 3: return launch_tf32_inst<3>(a, grid, smem, stream);
    case 4: return launch_tf32_inst<4>(a, grid, smem, stream);
    case 5: return launch_tf32_inst<5>(a, grid, smem, stream);
    case 6: return launch_tf32_inst<6>(a, grid, smem, stream);
    case 7: return launch_tf32_inst<7>(a, grid, smem, stream);
    case 8: return launch_tf32_inst<8>(a, grid, smem, stream);
    case 9: return launch_tf32_inst<9>(a, grid, smem, stream);
    case 10: return launch_tf32_inst<10>(a, grid, smem, stream);
    case 11: return launch_tf32_inst<11>(a, grid, smem, stream);
    default: set_error("tf32 kernel: unsupported K extent %d", a.KU); return 1;
  }
}

// Screen + exact decisions of the items [item_begin, item_end) of a segment-major work list.  The list
// is cut into PASSES of whole segments (all items of a pass share the column range [col0, col_end)),
// run from the highest columns down so that links found in one pass prune the candidates of the next:
//   pair_tf32 screen -> classify_items -> FP64 tensor kernel on dense items -> cand_scan (top-down)
int launch_pair_tf32(const PairArgs& pa, const PairArgs& fb_in, const Tf32Buffers& tb, const void* d_items,
                     const int* h_items, int item_begin, int item_end, int row_lo, int row_hi,
                     int rank, int nranks, int all_cand, int nstage_tile, cudaStream_t stream, int* grid_out,
                     int* smem_out, long long* launches_out) {
  if (grid_out) *grid_out = 0;
  if (item_end - item_begin <= 0) return 0;
  TfArgs a{};
  a.Xa = tb.Xa; a.Xb = tb.Xb; a.hn = tb.hn; a.er = pa.er;
  a.items = static_cast<const TfItem*>(d_items);
  a.nall = pa.nall; a.KU = tb.kslab; a.nslab = tb.nslab;
  const bool slab = tb.nslab > 1 || tb.kslab != pa.L.KU;
  int ns = 0;
  int smem = 0;
  if (slab) {
    if (tb.kslab != 8 * kSlabKsteps) { set_error("tf32 slab kernel: slab of %d slots", tb.kslab); return 1; }
    ns = kTfBStagesMax;
    smem = 2 * 3 * (tb.kslab / 4) * 2048 + ns * (tb.kslab / 4) * 768;
  } else {
    smem = tf32_smem_bytes(a.KU, &ns);
  }
  if (ns < 2) { set_error("tf32 kernel: K extent %d does not fit shared memory", a.KU); return 1; }
  a.nbstage = ns;
  a.cE = -0.5f * pa.g.E2_up;
  a.ethr = pa.g.ethr; a.enantio = pa.g.enantio; a.all_cand = all_cand; a.dbg = pa.g.debug;
  a.cmask = tb.cmask; a.fmask = all_cand ? tb.fmask : nullptr;
  a.dup = pa.dup; a.item_cnt = tb.item_cnt; a.cand_count = tb.cand_count;
  a.npairs = pa.npairs;
  a.dump = tb.dump; a.dump_n = pa.nall;
  if (tb.dump && a.dbg == 0) a.dbg = 3;       // the probe lives in the debug instantiation (3: no timing switch)
  if (all_cand && !tb.fmask) { set_error("self-check needs the far-verdict mask"); return 1; }
  int dev = 0, nsm = 148;
  CREST_CUDA_TRY(cudaGetDevice(&dev));
  CREST_CUDA_TRY(cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev));
  const int chunk = tb.chunk, segcols = chunk * 16;
  const int ntile = (pa.nall + 127) / 128;
  const bool col_layout = pa.L.rows == kUnit;
  int grid_max = 0;
  long long launches = 0;
  auto seg_of = [&](int k) { return h_items[4 * k + 1] / chunk; };
  int pe = item_end;
  while (pe > item_begin) {
    // grow the pass downwards by whole segments while its bitmask fits
    const int s_hi = seg_of(pe - 1);
    const int col_end = std::min((s_hi + 1) * segcols, row_hi);
    int pb = pe, s_lo = s_hi;
    for (;;) {
      int q = pb;
      const int s = seg_of(q - 1);
      while (q > item_begin && seg_of(q - 1) == s) --q;
      if (tf32_mask_words(row_lo, row_hi, s * segcols, col_end, nranks) > tb.cm_words) {
        if (pb == pe) { set_error("candidate bitmask too small for one segment (%zu words)", tb.cm_words); return 1; }
        break;
      }
      pb = q; s_lo = s;
      if (pb == item_begin) break;
    }
    const int col0 = s_lo * segcols;
    a.item_begin = pb; a.item_end = pe;
    const int row0 = std::max(row_lo, col0);
    const long long mrow0 = (long long)((row0 >> 7) / nranks) * 128;
    a.cm_row0 = (int)mrow0; a.cm_nranks = nranks; a.cm_col0 = col0; a.cm_stride = mask_stride(col0, col_end);
    const int nitems = pe - pb;
    const int grid = nitems < nsm ? nitems : nsm;
    if (grid > grid_max) grid_max = grid;
    CREST_CUDA_TRY(cudaMemsetAsync(tb.cand_count + 6, 0, sizeof(unsigned), stream));
    CREST_CUDA_TRY(cudaMemsetAsync(tb.cand_count + 10, 0, sizeof(unsigned), stream));   // row queue of cand_scan_kernel
    if (all_cand)
      CREST_CUDA_TRY(cudaMemsetAsync(tb.fmask, 0, sizeof(unsigned) * tf32_mask_words(row_lo, row_hi, col0, col_end, nranks), stream));
    if (launch_tf32_screen(a, slab, grid, smem, stream)) return 1;
    ClassifyArgs c{};
    c.items = a.items; c.item_begin = pb; c.item_end = pe; c.item_cnt = tb.item_cnt;
    c.dense = tb.dense; c.chunk = chunk; c.ntile = ntile; c.col_layout = col_layout ? 1 : 0;
    c.unfinished = tb.unfinished; c.nall = pa.nall;
    c.out_items = tb.dense_items; c.out_cap = tb.dense_items_cap; c.counters = tb.cand_count;
    c.nunits = pa.L.npanel; c.force_dense = tune().force_dense ? 1 : (tune().no_dense ? -1 : 0);
    classify_items_kernel<<<(nitems + 255) / 256, 256, 0, stream>>>(c);
    CREST_CUDA_TRY(cudaGetLastError());
    CREST_CUDA_TRY(cudaMemsetAsync(tb.unfinished + row_lo, 0, (size_t)(row_hi - row_lo), stream));
    ScanArgs e{};
    e.xv = make_xview(pa.X, pa.L); e.xn = pa.xn; e.rot = pa.rot; e.aniso = pa.aniso;
    e.cmask = tb.cmask; e.fmask = a.fmask;
    e.row0 = row0; e.col0 = col0; e.col_end = col_end; e.nall = row_hi; e.stride = a.cm_stride;
    e.mrow0 = (int)mrow0; e.nranks = nranks; e.rank = nranks > 1 ? rank : 0;
    e.dense = all_cand ? nullptr : tb.dense; e.unfinished = tb.unfinished;
    e.dense_budget = (c.force_dense > 0) ? 0 : (tune().dense_budget > 0 ? tune().dense_budget : -1);
    e.segcols = segcols; e.ntile = ntile;
    e.g = pa.g; e.dup = pa.dup; e.counters = tb.cand_count; e.debug_all = all_cand ? 1 : 0;
    const int rows = (row_hi - row0) / nranks + 128;
    int sgrid = (rows + 7) / 8;
    if (sgrid > nsm * 12) sgrid = nsm * 12;          // whole waves at 2, 3 or 4 resident blocks per SM
    if (sgrid < 1) sgrid = 1;
    // resident warps per SM: the scan is a chain of latencies per row (mask words, operand gathers, the quartic
    // solve), so it is paid in warps - 2 blocks at 128 registers, 3 at 80 (some spills), 4 at 64 (`scanocc=`)
    // Measured (profiles/r02_d_scan_occupancy.txt): up to 88 atoms 4 blocks win (C3 422 us against 482, C5 500 us
    // per pass against 620); with K slabs the evaluation loop is long and the spills cost more than the warps
    // bring (C4ext: 24.6 ms pair stage at 4, 21.9 at 3, 20.2 at 2)
    switch (tune().scan_occ > 0 ? tune().scan_occ : (col_layout ? 4 : 2)) {
      case 2: cand_scan_kernel<2><<<sgrid, 256, 0, stream>>>(e); break;
      case 3: cand_scan_kernel<3><<<sgrid, 256, 0, stream>>>(e); break;
      default: cand_scan_kernel<4><<<sgrid, 256, 0, stream>>>(e); break;
    }
    CREST_CUDA_TRY(cudaGetLastError());
    // unfinished units of dense items on the FP64 tensor kernel (idle CTAs return at once when the list is empty)
    if (!all_cand) {
      expand_dense_kernel<<<(nitems + 255) / 256, 256, 0, stream>>>(c);
      CREST_CUDA_TRY(cudaGetLastError());
      PairArgs fb = fb_in;
      fb.work_count = tb.cand_count + 6;
      int g2 = 0, s2 = 0;
      if (col_layout) {
        if (launch_pair_cols(fb, tb.dense_items, 0, 0, stream, &g2, &s2)) return 1;
      } else {
        fb.tiles = tb.dense_items; fb.tile_begin = 0; fb.tile_end = 0;
        if (launch_pair_tiles(fb, nstage_tile, stream, &g2, &s2)) return 1;
      }
    }
    launches += all_cand ? 3 : 5;
    pe = pb;
  }
  if (grid_out) *grid_out = grid_max;
  if (smem_out) *smem_out = smem;
  if (launches_out) *launches_out += launches;
  return 0;
}

}  // namespace crest
