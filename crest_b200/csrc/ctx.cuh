// Internal: the device-resident ensemble context and the host-side helpers shared by capi.cu (C ABI,
// context API) and driver.cu (one-call entry points: row-block pipeline, single-process multi-GPU).
#pragma once
#include "../../include/crest_cuda.h"
#include "common.cuh"
#include <limits>
#include <vector>

struct crest_cuda_ctx {
  int device = 0;
  int nat = 0, nall = 0, nsel = 0, rot_from_mask = 0;
  crest::PackLayout L{};
  std::vector<double> h_er;
  // device buffers
  int* d_at = nullptr;
  int* d_sel = nullptr;
  double* d_xyz = nullptr;
  double* d_xyz_axis = nullptr;
  double* d_er = nullptr;
  double* d_rot = nullptr;
  double* d_aniso = nullptr;
  double* d_X = nullptr;
  double* d_xn = nullptr;
  int* d_dup = nullptr;
  unsigned long long* d_npairs = nullptr;
  long long* d_rowstart = nullptr;
  int* d_jfirst = nullptr;
  unsigned long long* d_mask = nullptr;
  float* d_rmat = nullptr;
  size_t cap_xyz = 0, cap_conf = 0, cap_X = 0, cap_at = 0, cap_rows = 0, cap_mask = 0;
  // tile list (host)
  std::vector<long long> rowstart;
  std::vector<int> jfirst;
  double tiles_ethr = std::numeric_limits<double>::quiet_NaN();
  // column-sweep kernel work list (pair_col.cu): {J, u_begin, u_end, edge} quadruples
  bool use_col = false;
  std::vector<int> items;
  int* d_items = nullptr;
  size_t cap_items = 0;
  double items_ethr = std::numeric_limits<double>::quiet_NaN();
  long long launches = 0;
  int last_grid = 0, last_smem = 0;
  // tensor-core screen (pair_tf32.cu): TF32 operand copies, candidate list, its own work list
  bool fp64_only = false;              // set before ctx_upload: no TF32 operands / candidate mask (blockpair.cu)
  bool use_tf32 = false;
  int ntile = 0;                       // i-tiles of 128 conformers
  crest::Tf32Buffers tf{};
  size_t cap_tf_x = 0, cap_tf_hn = 0;
  unsigned long long* d_npairs_scratch = nullptr;   // pair counter of a fallback re-run (discarded)
  // work list of the screen, one segment-major list per STAGE (= a block of i-tiles; one stage unless the
  // one-call driver pipelines the upload, driver.cu), concatenated
  std::vector<int> tf_items;
  std::vector<int> stage_tile;  // nstage+1 tile boundaries
  std::vector<int> stage_item;  // nstage+1 item offsets into tf_items (in quadruples)
  std::vector<int> stage_chunk; // j-groups per segment of each stage's list
  int* d_tf_items = nullptr;
  size_t cap_tf_items = 0;
  double tf_items_ethr = std::numeric_limits<double>::quiet_NaN();
  int tf_items_nranks = 0;     // the item size depends on the number of ranks sharing the list
  int tf_items_rank = -1;      // the rank whose rows the list holds
  // links of this context's pairs go here: its own d_dup, or the d_dup of device 0 over NVLink peer
  // memory in a single-process multi-GPU call (driver.cu)
  int* dup_target = nullptr;
  cudaStream_t stream = nullptr;   // the context's own non-blocking stream (one-call entry points)
  // scratch of the whole-CREGEN call (stage B), kept across calls
  int* d_src = nullptr; int* d_group = nullptr; double* d_er2 = nullptr; double* d_rot2 = nullptr; double* d_enuc = nullptr;
  size_t cap_stageb = 0;
  // d_X padding (atom slots beyond the selection, rows beyond nall) is zeroed once per layout: pack
  // rewrites every used slot of every live row on each call and nothing else writes d_X
  bool X_zeroed = false;
  crest::PackLayout X_zero_L{};
  int X_zero_nall = -1;
  int last_variant = 0;                // 0 row-tile DMMA, 1 column DMMA, 2 TF32 screen + exact list
};


namespace crest {
// stage = the i-tiles [tile0, tile1) of the ensemble (all of them: 0, ntile).  Launches the pair kernels of
// shard `rank` of `nranks` for the stage on `st`; links go to c->dup_target (atomic max, system scope).
int pairs_stage(crest_cuda_ctx* c, int enantio, const crest_cuda_thresholds* thr, int rank, int nranks,
                int stage, int want_mask, float* d_rmat, cudaStream_t st);
// (re)build the per-stage work lists of the screen for tile boundaries `bounds` (nstage+1 entries, first 0,
// last ntile) and upload them; cached on (ethr, nranks, bounds)
int plan_stages(crest_cuda_ctx* c, double ethr, int nranks, int rank, const std::vector<int>& bounds, cudaStream_t st);
bool tf32_path(const crest_cuda_ctx* c, int want_mask, const float* d_rmat);
// K0 on the rows [r0, r1) (r0 a multiple of 128): pack_axis + TF32 operand conversion
int pack_rows(crest_cuda_ctx* c, int r0, int r1, cudaStream_t st);
// zero the per-call accumulators (dup, pair counter, operand padding once per layout)
int begin_call(crest_cuda_ctx* c, cudaStream_t st);
int check_thr(const crest_cuda_thresholds* t);
int copy_threads();
struct CopySlots;
}  // namespace crest
