// Shared declarations of the CREGEN CUDA path (sm_100a).  See DESIGN.md §3 for the HBM layout.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>

namespace crest {

constexpr int kPanel = 32;        // conformers per panel (= i-tile height)
constexpr int kJPanels = 2;       // panels per j-tile (j-tile width 64)
constexpr int kTileI = kPanel;
constexpr int kTileJ = kPanel * kJPanels;
constexpr int kConsumerWarps = 8; // 2 (i) x 4 (j) warps, each 16x16 pairs
constexpr int kPairThreads = kConsumerWarps * 32;   // producer duty folded into warp 0
constexpr double kAutokcal = 627.50947428;               // reference src/crest_pars.f90:23

// Packed-coordinate layout in HBM (written by pack_axis, read by pair_tile through bulk copies):
//   X[panel][slab][comp(3)][conf(32)][KS] doubles; one (panel,slab) chunk is contiguous.
// KU atoms of each slab are used by the MMA loop (multiple of 8, zero padded), KS >= KU is the
// row stride with KS % 16 == 8 so that 128-bit fragment loads are bank-conflict free.
struct PackLayout {
  int A;        // atoms entering the RMSD
  int nslab;    // K slabs
  int KU;       // atoms used per slab
  int KS;       // row stride (doubles)
  int npanel;   // panels (even, >= ceil(nall/32)) / 16-row units (multiple of 4) in the column layout
  int rows;     // conformers per chunk: kPanel (row-tile kernel) or kUnit (column kernel)
  int half_tail; // column layout: the last 8-atom block holds <= 4 atoms, stored at its EVEN slots
                 // (the k-step that reads the odd slots is skipped: no MMA work on K padding)
  __host__ __device__ size_t chunk_doubles() const { return (size_t)3 * rows * KS; }
  __host__ __device__ size_t total_doubles() const { return chunk_doubles() * nslab * npanel; }
};

// Developer knobs (env CREST_CUDA_TUNE="slab=<8-atom blocks per slab>,stages=<n>,dist=<n>,skew=<clk>"),
// read once; 0 = automatic.  Used to A/B the pipeline shape on the GPU box without rebuilding.
struct Tune { int blocks = 0, even_blocks = 0, block_f1 = 0, block_gain = 0, no_stage = 0, force_dense = 0, no_dense = 0, mask_mb = 0, dense_budget = 0, slab_blocks = 0, stages = 0, dist = 0, skew = -1, force_tile = 0, no_half = 0, no_tf32 = 0, cand_cap = 0, chunk = 0, seq_group = 0, group_cap = 0, copy_threads = -1, pack_g = 0, pack_old = 0, scan_occ = 0; };
const Tune& tune();
void set_tune(const char* spec);   // re-parse the knobs (tests A/B kernel variants in one process)

// K slabs of at most 40 atoms (5 blocks of 8); KS >= KU is the smem row stride with KS % 16 == 8
// (bank rule for the 128-bit fragment loads).  Fewer, larger ring items measured faster than a
// deep ring of small ones: each item costs warp 0 ~1000 clk of bulk-copy issue.
inline PackLayout make_layout(int A, int nall) {
  PackLayout L;
  L.A = A;
  L.rows = kPanel;
  L.half_tail = 0;
  const int nb = (A + 7) / 8 > 0 ? (A + 7) / 8 : 1;   // 8-atom blocks
  int b = tune().slab_blocks;
  if (b <= 0) {
    const int nslab = (nb + 4) / 5;
    b = (nb + nslab - 1) / nslab;
  }
  L.nslab = (nb + b - 1) / b;
  L.KU = b * 8;
  L.KS = (L.KU % 16 == 8) ? L.KU : L.KU + 8;
  int np = (nall + kPanel - 1) / kPanel;
  if (np % kJPanels) np += kJPanels - (np % kJPanels);
  if (np < kJPanels) np = kJPanels;
  L.npanel = np;
  return L;
}

// Column-sweep kernel (pair_col.cu): the whole K extent of a conformer is one row, chunks are
// 16-row units, a j-tile is 4 consecutive units.  Usable while a j-tile plus >= 2 ring stages fit
// in shared memory (A <= 88).
constexpr int kUnit = 16;
constexpr int kColJ = 64;          // j-tile width = 4 units
constexpr int kColMaxNkb = 11;
inline bool col_layout_fits(int A) { return (A + 7) / 8 <= kColMaxNkb; }
inline PackLayout make_col_layout(int A, int nall) {
  PackLayout L;
  L.A = A;
  L.rows = kUnit;
  const int nb = (A + 7) / 8 > 0 ? (A + 7) / 8 : 1;
  L.half_tail = (A - 8 * (nb - 1) <= 4 && !tune().no_half) ? 1 : 0;
  L.nslab = 1;
  L.KU = nb * 8;
  L.KS = (L.KU % 16 == 8) ? L.KU : L.KU + 8;
  int nu = (nall + kUnit - 1) / kUnit;
  nu = (nu + 3) / 4 * 4;
  if (nu < 4) nu = 4;
  L.npanel = nu;
  return L;
}

// per-launch gate constants, precomputed on the host (capi.cu: make_gates)
struct Gates {
  double ethr;       // kcal/mol
  double E1;         // e < E1  <=>  real(rmsd,sp) < rthr        (e = xn+yn-2*lambda)
  double E2;         // e < E2  <=>  real(rmsd,sp) < 2*rthr
  double bthr_half;  // 0.5*bthr (cregen.f90:1368)
  double bthrmax, bthrshift;
  float E2_up;       // E2 rounded up to float (FP32 screen)
  int enantio;
  int debug;         // profiling switch (env CREST_CUDA_DEBUG): 1 = skip the epilogue (timing only)
};

}  // namespace crest

#define CREST_CUDA_TRY(expr)                                                          \
  do {                                                                                \
    cudaError_t _e = (expr);                                                          \
    if (_e != cudaSuccess) {                                                          \
      crest::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
      return 1;                                                                       \
    }                                                                                 \
  } while (0)

namespace crest {
void set_error(const char* fmt, ...);

// kernel launchers (implemented in the .cu files)
int launch_pack_axis(int nat, int nall, const int* d_at, const double* d_xyz, const int* d_sel,
                     int nsel, int rot_from_mask, const PackLayout& L, double* d_xyz_axis,
                     double* d_rot, double* d_aniso, double* d_X, double* d_xn,
                     cudaStream_t stream);
int launch_rotconst(int nat, int n, const int* d_at, const double* d_xyz, double* d_rot,
                    cudaStream_t stream);
// Pair SINK: the callers of rmsd() beside cregen_CRE (blockpair.cu: compare_ensembles, intraconfRMSD) run
// the SAME FP64 tensor mainloops (pair_col.cu / pair_tile.cu) over rectangular or block-diagonal item
// lists; instead of the CREGEN decisions every valid pair gets its exact RMSD and goes to the sink.
struct PairSink {
  int mode;               // 0 none (CREGEN), 1 compare_blocks (min per block pair), 2 intraconf
  const double* Xj;       // j-side operand and norms (mode 1: the second ensemble; else the i side)
  const double* xnj;
  int ni, nj;             // structures on either side
  const int* blk_i;       // block (conformer) id per structure, -1 = none
  const int* blk_j;
  int nconf1;
  unsigned long long* rmat_bits;   // mode 1: rmat(nconf1, nconf2) as ordered bit patterns, atomicMin
  int* has_earlier;       // mode 2: [ni] 1 = some j < i of the same conformer with rmsd < rthr
  int* enantio;           // mode 2: [nconf] 1 = enantiomer probe fired
  double rthr, rthr_enantio;
  int A;                  // atoms in the RMSD
};

struct PairArgs {
  PairSink sink;         // zero-initialised = CREGEN
  const double* X;       // packed coordinates
  const double* xn;      // centred squared norms
  const double* er;      // energies (Hartree, ascending)
  const double* rot;     // (3,nall) MHz
  const double* aniso;   // rotaniso per structure
  const long long* rowstart;  // [nI+1] prefix of tiles per i-panel
  const int* jfirst;     // [nI] first j-tile of each i-panel
  int nI;
  int nall;
  long long tile_begin, tile_end;   // this shard's slice of the tile list
  int* dup;              // out, atomicMax of (j+1)
  unsigned long long* npairs;       // out, atomicAdd
  unsigned long long* mask;         // optional 8x8-tile bitmask [N8][N8]
  int n8;
  float* rmat;           // optional all-pairs float32 band (diagnostics)
  Gates g;
  PackLayout L;
  int dist;              // producer runs `dist` ring items ahead of warp 0
  int skew_clk;          // start delay of warps 4-7 (half a tile), 0 = none
  long long* dbg;        // CREST_CUDA_DEBUG=2: per-warp cycle counters [8 warps][4 phases]
  // Device-generated work (dense items behind the tensor-core screen, pair_tf32.cu classify_items_kernel):
  // if non-NULL the kernel takes its item / tile COUNT from *work_count (items start at item_begin /
  // tile_begin = 0) and, in the row-tile kernel, the tiles themselves from `tiles` {panel, j-tile, -, -}.
  const unsigned* work_count;
  const int4* tiles;
};
int launch_pair_tiles(const PairArgs& a, int nstage, cudaStream_t stream, int* grid_out,
                      int* smem_out);
int pair_smem_bytes(const PackLayout& L, int* nstage_out);
// column-sweep kernel (pair_col.cu); items = device array of {J, u_begin, u_end, edge} int quadruples
int launch_pair_cols(const PairArgs& a, const void* d_items, int item_begin, int item_end,
                     cudaStream_t stream, int* grid_out, int* smem_out);
int col_smem_bytes(int KS, int* nstage_out);
// tensor-core screen (pair_tf32.cu): TF32 operand copies, screen norms and the candidate list
struct Tf32Buffers {
  float* Xa = nullptr;        // [tile][comp][chunk][128][4]  A operands (i-tiles of 128 conformers)
  float* Xb = nullptr;        // [group][chunk][48][4]        B operands (j-groups of 16 conformers)
  float* hn = nullptr;        // [ntile*128] rd(0.5*|x|^2*(1-kappa))
  unsigned* cmask = nullptr;  // candidate bitmask of one pass (rows x columns of the pass), all-zero between launches
  unsigned* fmask = nullptr;  // self-check only (CREST_CUDA_DEBUG=3): the screen's "far" verdicts
  size_t cm_words = 0;        // capacity of cmask (and fmask) in 32-bit words
  unsigned* item_cnt = nullptr;     // candidates per work item, zero between launches
  unsigned char* dense = nullptr;   // [segment][i-tile] verdict of classify_items_kernel
  unsigned char* unfinished = nullptr;  // [row] rows that gave up inside a dense item (cand_scan_kernel)
  size_t unfinished_cap = 0;
  size_t dense_cap = 0;
  int4* dense_items = nullptr;      // work list of the FP64 tensor kernel for dense items
  unsigned dense_items_cap = 0;
  unsigned* cand_count = nullptr;   // [0] exact evaluations, [1] unsafe magnitudes, [2] debug violations, [3] candidates, [4] converter flag, [5] dense items, [6] dense list length, [7] dense list overflow
  float* dump = nullptr;      // accumulate-error probe only: raw covariances [nall][nall][9] (crest_cuda_ctx_tf32_accum)
  int chunk = 0;              // j-groups per segment of the work list (multiple of 8)
  double kappa = 0.0;         // relative lowering of the screen's test point (error budget of this atom count)
  int nslab = 1;              // TF32 K slabs (1 for A <= 88; slabs of 48 slots for the row-tile layout)
  int kslab = 0;              // atom slots per slab (multiple of 8)
};
constexpr int kTfSlabSlots = 48;
int tf32_smem_bytes(int KU, int* nbstage_out);
// i-tiles [tile0, tile1); *d_unsafe is only ever set (the caller zeroes it once per call)
int launch_tf32_convert(const double* d_X, const double* d_xn, int nall, const PackLayout& L,
                        int tile0, int tile1, int nslab_t, int kslab, double kappa, float* d_Xa,
                        float* d_Xb, float* d_hn, unsigned* d_unsafe, cudaStream_t stream);
double tf32_kappa(int kslots);
int launch_tf32_accum_check(const Tf32Buffers& tb, int nall, const float* d_dump, unsigned long long* d_out,
                            cudaStream_t stream);
// items = device array of {I, g_begin, g_end, g_gate_end} int quadruples
// h_items = host copy of the list (the launcher cuts [item_begin, item_end) into passes of whole
// segments whose candidate bitmask fits tb.cm_words); fb = arguments of the FP64 tensor kernel that
// re-does dense items (its own pair counter); launches_out += kernels launched
// rows [row_lo, row_hi) = the i-tiles the items belong to (a stage; 0, nall for the whole ensemble)
int launch_pair_tf32(const PairArgs& a, const PairArgs& fb, const Tf32Buffers& tb, const void* d_items,
                     const int* h_items, int item_begin, int item_end, int row_lo, int row_hi,
                     int rank, int nranks, int all_cand, int nstage_tile, cudaStream_t stream, int* grid_out,
                     int* smem_out, long long* launches_out);
size_t tf32_mask_words(int row_lo, int row_hi, int col0, int col_end, int nranks);
int launch_backtrack(int nall, int* d_dup, cudaStream_t stream);
int launch_rmsd_pairs(int A, int n, const double* d_packed, long long npairs, const int* d_i,
                      const int* d_j, double* d_rmsd, double* d_rmsd_m, cudaStream_t stream);
int launch_rmsdalign(int nat, int nall, const int* d_at, double* d_xyz, cudaStream_t stream);
int launch_permute(int nat, int nall, const int* d_src, const double* d_xyz_in, const double* d_er_in,
                   const double* d_rot_in, double* d_xyz_out, double* d_er_out, double* d_rot_out,
                   cudaStream_t stream);
int launch_enuc(int nat, int n, const int* d_at, const double* d_xyz, double* d_enuc,
                cudaStream_t stream);
int launch_group(int n, const double* d_er, const double* d_rot, const double* d_enuc,
                 double ethr, double bthr, double bthrmax, double bthrshift, int* d_group,
                 cudaStream_t stream);
}  // namespace crest
