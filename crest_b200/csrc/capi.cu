// C ABI of the CREGEN CUDA path (include/crest_cuda.h): argument checks, context, host<->device
// copies, the host-side band/tile bookkeeping (reference src/cregen.f90:1240-1256) and the exact
// float32-threshold emulation (cregen.f90:1286,1359-1367).  No CPU fallback lives here: every
// compute entry point needs a CUDA device and fails otherwise.
#include "ctx.cuh"
#include "hostcopy.cuh"
#include <cmath>
#include <cstdarg>
#include <cstdlib>
#include <functional>
#include <cstring>
#include <limits>
#include <utility>
#include <vector>
#include <chrono>
#include <thread>

namespace crest {

static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

// Smallest e such that NOT ( real( sqrt(max(0,e)/n), sp ) < thr ).  All operations involved are
// monotone non-decreasing in e, so `e < E` reproduces the reference's comparison of the
// float32-rounded RMSD (rmat is real(sp), cregen.f90:1176,1286) with the double threshold bit
// for bit, without a sqrt or a division per pair.
static double e_threshold(double thr, int n) {
  if (!(thr > 0.0)) return -std::numeric_limits<double>::infinity();
  auto below = [&](double e) {
    const double r = std::sqrt(std::max(0.0, e) / (double)n);
    return (double)(float)r < thr;
  };
  uint64_t lo = 0, hi;                       // bit patterns of non-negative doubles are ordered
  double big = std::numeric_limits<double>::max();
  if (below(big)) return std::numeric_limits<double>::infinity();
  std::memcpy(&hi, &big, 8);
  // invariant: below(lo) true, below(hi) false
  while (hi - lo > 1) {
    const uint64_t mid = lo + (hi - lo) / 2;
    double m;
    std::memcpy(&m, &mid, 8);
    if (below(m)) lo = mid; else hi = mid;
  }
  double E;
  std::memcpy(&E, &hi, 8);
  return E;
}

static void parse_tune(const char* e, Tune& t) {
  t = Tune();
  if (!e) return;
  int v;
  if (const char* q = std::strstr(e, "slab=")) if (std::sscanf(q + 5, "%d", &v) == 1) t.slab_blocks = v;
  if (const char* q = std::strstr(e, "stages=")) if (std::sscanf(q + 7, "%d", &v) == 1) t.stages = v;
  if (const char* q = std::strstr(e, "dist=")) if (std::sscanf(q + 5, "%d", &v) == 1) t.dist = v;
  if (std::strstr(e, "tile")) t.force_tile = 1;
  if (std::strstr(e, "nohalf")) t.no_half = 1;
  if (std::strstr(e, "notf32")) t.no_tf32 = 1;
  if (std::strstr(e, "seqgroup")) t.seq_group = 1;
  if (std::strstr(e, "forcedense")) t.force_dense = 1;
  if (std::strstr(e, "nodense")) t.no_dense = 1;
  if (std::strstr(e, "nostage")) t.no_stage = 1;
  if (std::strstr(e, "evenblocks")) t.even_blocks = 1;
  if (const char* q = std::strstr(e, "blockf1=")) if (std::sscanf(q + 8, "%d", &v) == 1) t.block_f1 = v;
  if (const char* q = std::strstr(e, "blockgain=")) if (std::sscanf(q + 10, "%d", &v) == 1) t.block_gain = v;
  if (std::strstr(e, "packold")) t.pack_old = 1;
  if (const char* q = std::strstr(e, "blocks=")) if (std::sscanf(q + 7, "%d", &v) == 1) t.blocks = v;
  if (const char* q = std::strstr(e, "densebudget=")) if (std::sscanf(q + 12, "%d", &v) == 1) t.dense_budget = v;
  if (const char* q = std::strstr(e, "maskmb=")) if (std::sscanf(q + 7, "%d", &v) == 1) t.mask_mb = v;
  if (const char* q = std::strstr(e, "scanocc=")) if (std::sscanf(q + 8, "%d", &v) == 1) t.scan_occ = v;
  if (const char* q = std::strstr(e, "packg=")) if (std::sscanf(q + 6, "%d", &v) == 1) t.pack_g = v;
  if (const char* q = std::strstr(e, "copythr=")) if (std::sscanf(q + 8, "%d", &v) == 1) t.copy_threads = v;
  if (const char* q = std::strstr(e, "groupcap=")) if (std::sscanf(q + 9, "%d", &v) == 1) t.group_cap = v;
  if (const char* q = std::strstr(e, "chunk=")) if (std::sscanf(q + 6, "%d", &v) == 1) t.chunk = v;
  if (const char* q = std::strstr(e, "skew=")) if (std::sscanf(q + 5, "%d", &v) == 1) t.skew = v;
}
static Tune g_tune;
static bool g_tune_init = false;
const Tune& tune() {
  if (!g_tune_init) {
    g_tune_init = true;
    parse_tune(std::getenv("CREST_CUDA_TUNE"), g_tune);
  }
  return g_tune;
}
void set_tune(const char* spec) {
  g_tune_init = true;
  parse_tune(spec, g_tune);
}

static Gates make_gates(const crest_cuda_thresholds& t, int enantio, int A) {
  Gates g;
  g.ethr = t.ethr;
  g.E1 = e_threshold(t.rthr, A);
  g.E2 = e_threshold(2.0 * t.rthr, A);
  g.E2_up = std::nextafterf((float)g.E2, std::numeric_limits<float>::infinity());
  g.bthr_half = 0.5 * t.bthr;
  g.bthrmax = t.bthrmax;
  g.bthrshift = t.bthrshift;
  g.enantio = enantio ? 1 : 0;
  const char* dbg = std::getenv("CREST_CUDA_DEBUG");
  g.debug = dbg ? std::atoi(dbg) : 0;
  return g;
}

}  // namespace crest

using namespace crest;

namespace {

// In-band tile list: for i-panel I the j-tiles [jfirst, jlast]; energies ascend, so the first
// in-band j of the panel's first row bounds the whole panel (cregen.f90:1245-1255 rmap2).
void band_tiles_host(int n, const double* er, double ethr, std::vector<long long>& rowstart,
                     std::vector<int>& jfirst) {
  const int nI = (n + kTileI - 1) / kTileI;
  rowstart.assign(nI + 1, 0);
  jfirst.assign(nI > 0 ? nI : 1, 0);
  long long acc = 0;
  for (int I = 0; I < nI; ++I) {
    rowstart[I] = acc;
    const int ifirst = I * kTileI;
    const int ilast = std::min(ifirst + kTileI - 1, n - 1);
    const int jmax = ilast - 1;
    // first j with (er[ifirst]-er[j])*autokcal < ethr (monotone in j).  de grows with i, so the
    // panel's FIRST row admits the smallest j; the other rows are gated per pair in edge tiles.
    int lo = 0, hi = ifirst;   // answer in [0, ifirst]
    while (lo < hi) {
      const int mid = (lo + hi) / 2;
      const double de = (er[ifirst] - er[mid]) * kAutokcal;
      if (de < ethr) hi = mid; else lo = mid + 1;
    }
    const int jlo = lo;
    if (jmax >= 0 && jlo <= jmax) {
      jfirst[I] = jlo / kTileJ;
      acc += (long long)(jmax / kTileJ - jfirst[I] + 1);
    } else {
      jfirst[I] = 0;
    }
  }
  rowstart[nI] = acc;
}

void build_tiles(crest_cuda_ctx* c, double ethr) {
  band_tiles_host(c->nall, c->h_er.data(), ethr, c->rowstart, c->jfirst);
  c->tiles_ethr = ethr;
}

// Work list of the column-sweep kernel.  Column J = j-tile of 64 conformers; its in-band rows run
// from the diagonal to the last i with (er[i]-er[j_hi])*autokcal < ethr (the tile's largest j
// admits the largest i; the per-pair gate inside edge units does the rest, cregen.f90:1275).
// Rows are cut on a fixed grid of `seg` 16-row units and the list is ordered segment-major, so
// the CTAs that run at the same time stream the SAME i-units (L2-resident) past different j-tiles.
void col_items_host(int n, const double* er, double ethr, std::vector<int>& items) {
  items.clear();
  if (n < 2) return;
  const int nJ = (n + kColJ - 1) / kColJ;
  std::vector<int> uend(nJ, 0);
  long long total = 0;
  for (int J = 0; J < nJ; ++J) {
    const int jlo = J * kColJ, jhi = std::min(jlo + kColJ - 1, n - 1);
    // last i in [jlo, n-1] with de(i, jhi) < ethr; de grows with i.  i <= jhi always passes
    // (de <= 0 < ethr unless ethr <= 0, where nothing passes).
    int lo = jlo, hi = n - 1;
    if (!((er[lo] - er[jhi]) * kAutokcal < ethr)) { uend[J] = 0; continue; }
    while (lo < hi) {
      const int mid = (lo + hi + 1) / 2;
      if ((er[mid] - er[jhi]) * kAutokcal < ethr) lo = mid; else hi = mid - 1;
    }
    const int imax = lo;
    if (imax <= jlo) { uend[J] = 0; continue; }
    uend[J] = imax / kUnit + 1;
    total += uend[J] - 4 * J;
  }
  long long seg = total / (148LL * 8 * 24);
  if (seg < 8) seg = 8;
  if (seg > 256) seg = 256;
  const int nu = (n + kUnit - 1) / kUnit;
  const int nseg = (int)((nu + seg - 1) / seg);
  for (int sg = nseg - 1; sg >= 0; --sg) {          // bottom rows first: most columns, full items
    const int s0 = (int)(sg * seg), s1 = (int)std::min<long long>((sg + 1) * seg, nu);
    for (int J = 0; J < nJ; ++J) {
      if (uend[J] == 0) continue;
      const int b = std::max(4 * J, s0), e = std::min(uend[J], s1);
      if (b < e) {
        // edge: the item's largest i against the tile's smallest j may fail the gate
        const int ilast = std::min(e * kUnit - 1, n - 1);
        const int edge = ((er[ilast] - er[J * kColJ]) * kAutokcal < ethr) ? 0 : 1;
        items.push_back(J); items.push_back(b); items.push_back(e); items.push_back(edge);
      }
    }
  }
}

// Work list of the tensor-core screen kernel.  i-tile I = 128 conformers; its in-band j-groups (16
// conformers) run from the first j admitted by the tile's FIRST row (smallest de) to the diagonal.
// Groups below g_gate_end contain pairs that fail the energy gate for the tile's LAST row and are
// gated pair by pair (cregen.f90:1275); the rest of the band needs no energies at all.  The list is
// cut on a fixed grid of `chunk` groups and ordered segment-major so that concurrently running CTAs
// stream the same j-groups (L2-resident) past different i-tiles.
// tile0/tile1: the list of a STAGE, i.e. only the i-tiles [tile0, tile1) (appended to `items`).
// rank (0..nranks-1): only the tiles I with I % nranks == rank - ranks share the pair space by ROWS, so the
// per-row "largest j" needs no exchange beyond combining disjoint rows, and a rank sees all columns of its
// rows: links found in its high columns prune its low ones (pair_tf32.cu).  Rows interleave, so the
// triangle's growing row length is spread evenly.  rank < 0: all tiles.
int tf_items_host(int n, const double* er, double ethr, std::vector<int>& items, int nranks = 1,
                  int tile0 = 0, int tile1 = -1, int rank = -1) {
  if (n < 2) return 8;
  const int ntile = (n + 127) / 128;
  if (tile1 < 0 || tile1 > ntile) tile1 = ntile;
  std::vector<int> gb(ntile, 0), ge(ntile, 0), gg(ntile, 0);
  // first j in [0, i] with (er[i]-er[j])*autokcal < ethr; i+1 if none (monotone: er ascends)
  auto first_in_band = [&](int i) {
    int lo = 0, hi = i + 1;
    while (lo < hi) {
      const int mid = (lo + hi) / 2;
      if ((er[i] - er[mid]) * kAutokcal < ethr) hi = mid; else lo = mid + 1;
    }
    return lo;
  };
  const int nr = nranks > 0 ? nranks : 1;
  auto mine = [&](int I) { return rank < 0 || I % nr == rank; };
  long long total = 0;
  for (int I = tile0; I < tile1; ++I) {
    if (!mine(I)) continue;
    const int i0 = I * 128, ilast = std::min(i0 + 127, n - 1), jhi = ilast - 1;
    if (jhi < 0) continue;
    const int jlo = first_in_band(i0);
    if (jlo > jhi) continue;
    gb[I] = jlo / 16; ge[I] = jhi / 16 + 1;
    const int jint = first_in_band(ilast);
    gg[I] = (jint + 15) / 16;
    total += ge[I] - gb[I];
  }
  // >= 16 items per CTA and rank: the CTAs of a rank take its items round-robin, so the stage lasts
  // ceil(items / 148) items (measured on C3: 128-group items cost nothing against 256, 64 cost 1.3 %)
  // `total` already is this rank's share when a rank is given
  const long long per_cta = 148LL * (rank < 0 ? nr : 1);
  long long chunk = tune().chunk > 0 ? tune().chunk : total / (per_cta * 16);
  // small lists (the first row blocks of the upload pipeline, C2): an item re-loads its 96 KB i-tile, so
  // items below 64 groups trade balance for operand traffic - 4 items per CTA are enough there
  if (tune().chunk <= 0 && chunk < 64) chunk = std::min(64LL, std::max(chunk, total / (per_cta * 4)));
  chunk = chunk / 8 * 8;       // segments are whole 128-column blocks of the candidate bitmask
  if (chunk < 8) chunk = 8;
  if (chunk > 256) chunk = 256;
  const int ngroup = tile1 * 8;
  const int nseg = (int)((ngroup + chunk - 1) / chunk);
  for (int sg = 0; sg < nseg; ++sg) {
    const int s0 = (int)(sg * chunk), s1 = (int)std::min<long long>((sg + 1) * chunk, ngroup);
    for (int I = tile1 - 1; I >= tile0; --I) {
      if (!mine(I)) continue;
      const int b = std::max(gb[I], s0), e = std::min(ge[I], s1);
      if (b < e) { items.push_back(I); items.push_back(b); items.push_back(e); items.push_back(gg[I]); }
    }
  }
  return (int)chunk;
}

}  // namespace

namespace crest {
int plan_stages(crest_cuda_ctx* c, double ethr, int nranks, int rank, const std::vector<int>& bounds, cudaStream_t st) {
  if (c->tf_items_ethr == ethr && c->tf_items_nranks == nranks && c->tf_items_rank == rank && c->stage_tile == bounds) return 0;
  const int nstage = (int)bounds.size() - 1;
  c->tf_items.clear();
  c->stage_item.assign(1, 0);
  c->stage_chunk.clear();
  size_t nd = 1, per_item_max = 1;
  for (int s = 0; s < nstage; ++s) {
    const int chunk = tf_items_host(c->nall, c->h_er.data(), ethr, c->tf_items, nranks, bounds[s], bounds[s + 1],
                                    nranks > 1 ? rank : -1);
    c->stage_chunk.push_back(chunk);
    c->stage_item.push_back((int)(c->tf_items.size() / 4));
    const size_t nseg = ((size_t)bounds[s + 1] * 8 + chunk - 1) / chunk;
    nd = std::max(nd, nseg * (size_t)c->ntile);
    per_item_max = std::max(per_item_max, (size_t)(chunk / 4 + 1) * (c->use_col ? 8 : 4));
  }
  c->stage_tile = bounds;
  c->tf_items_ethr = ethr;
  c->tf_items_nranks = nranks;
  c->tf_items_rank = rank;
  const size_t n = c->tf_items.size();
  if (n > c->cap_tf_items) {
    cudaFree(c->d_tf_items); c->d_tf_items = nullptr; c->cap_tf_items = 0;
    cudaFree(c->tf.item_cnt); c->tf.item_cnt = nullptr;
    CREST_CUDA_TRY(cudaMalloc(&c->d_tf_items, sizeof(int) * n));
    CREST_CUDA_TRY(cudaMalloc(&c->tf.item_cnt, sizeof(unsigned) * (n / 4)));
    CREST_CUDA_TRY(cudaMemsetAsync(c->tf.item_cnt, 0, sizeof(unsigned) * (n / 4), st));
    c->cap_tf_items = n;
  }
  if (n) CREST_CUDA_TRY(cudaMemcpyAsync(c->d_tf_items, c->tf_items.data(), sizeof(int) * n,
                                        cudaMemcpyHostToDevice, st));
  // verdict per (segment, i-tile) and the work list of the FP64 tensor kernel for dense items
  if (nd > c->tf.dense_cap) {
    cudaFree(c->tf.dense); c->tf.dense = nullptr; c->tf.dense_cap = 0;
    CREST_CUDA_TRY(cudaMalloc(&c->tf.dense, nd));
    c->tf.dense_cap = nd;
  }
  size_t want = std::min<size_t>((n / 4) * per_item_max, (size_t)32 << 20);
  if (want < 1024) want = 1024;
  if (want > c->tf.dense_items_cap) {
    cudaFree(c->tf.dense_items); c->tf.dense_items = nullptr; c->tf.dense_items_cap = 0;
    CREST_CUDA_TRY(cudaMalloc(&c->tf.dense_items, sizeof(int4) * want));
    c->tf.dense_items_cap = (unsigned)want;
  }
  return 0;
}
}  // namespace crest

namespace {

int upload_items(crest_cuda_ctx* c, double ethr, cudaStream_t st) {
  col_items_host(c->nall, c->h_er.data(), ethr, c->items);
  c->items_ethr = ethr;
  const size_t n = c->items.size();
  if (n > c->cap_items) {
    cudaFree(c->d_items); c->d_items = nullptr; c->cap_items = 0;
    CREST_CUDA_TRY(cudaMalloc(&c->d_items, sizeof(int) * n));
    c->cap_items = n;
  }
  if (n) CREST_CUDA_TRY(cudaMemcpyAsync(c->d_items, c->items.data(), sizeof(int) * n,
                                        cudaMemcpyHostToDevice, st));
  return 0;
}

int upload_tiles(crest_cuda_ctx* c, cudaStream_t st) {
  const size_t nI = c->jfirst.size();
  if (nI + 1 > c->cap_rows) {
    if (c->d_rowstart) cudaFree(c->d_rowstart);
    if (c->d_jfirst) cudaFree(c->d_jfirst);
    c->d_rowstart = nullptr; c->d_jfirst = nullptr;
    CREST_CUDA_TRY(cudaMalloc(&c->d_rowstart, sizeof(long long) * (nI + 1)));
    CREST_CUDA_TRY(cudaMalloc(&c->d_jfirst, sizeof(int) * nI));
    c->cap_rows = nI + 1;
  }
  CREST_CUDA_TRY(cudaMemcpyAsync(c->d_rowstart, c->rowstart.data(),
                                 sizeof(long long) * c->rowstart.size(), cudaMemcpyHostToDevice, st));
  CREST_CUDA_TRY(cudaMemcpyAsync(c->d_jfirst, c->jfirst.data(), sizeof(int) * nI,
                                 cudaMemcpyHostToDevice, st));
  return 0;
}

}  // namespace

namespace crest {
int check_thr(const crest_cuda_thresholds* t) {
  if (!t) { set_error("thresholds pointer is NULL"); return 1; }
  if (!(t->rthr == t->rthr) || !(t->ethr == t->ethr)) { set_error("NaN threshold"); return 1; }
  return 0;
}
}  // namespace crest

extern "C" {

const char* crest_cuda_last_error(void) { return g_err; }

void crest_cuda_set_tune(const char* spec) { crest::set_tune(spec); }

int crest_cuda_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}

// ---- host-only bookkeeping (no device needed; also what the CPU tests exercise) ----
double crest_cuda_e_threshold(double thr, int n) { return e_threshold(thr, n > 0 ? n : 1); }

void crest_cuda_tile_shape(int* tile_i, int* tile_j) {
  if (tile_i) *tile_i = kTileI;
  if (tile_j) *tile_j = kTileJ;
}

int crest_cuda_band_tiles(int nall, const double* er, double ethr, long long* rowstart,
                          int* jfirst, long long* ntiles) {
  if (nall < 0 || (!er && nall > 0)) { set_error("bad argument"); return 1; }
  std::vector<long long> rs;
  std::vector<int> jf;
  band_tiles_host(nall, er, ethr, rs, jf);
  const int nI = (nall + kTileI - 1) / kTileI;
  if (rowstart) for (int i = 0; i <= nI; ++i) rowstart[i] = rs[i];
  if (jfirst) for (int i = 0; i < nI; ++i) jfirst[i] = jf[i];
  if (ntiles) *ntiles = rs[nI];
  return 0;
}

int crest_cuda_tf32_items_n(int nall, const double* er, double ethr, int nranks, int* items,
                            long long cap, long long* nitems) {
  if (nall < 0 || (!er && nall > 0) || !nitems || nranks < 1) { set_error("bad argument"); return 1; }
  std::vector<int> v;
  if (nranks == 1) tf_items_host(nall, er, ethr, v, 1);
  else for (int r = 0; r < nranks; ++r) tf_items_host(nall, er, ethr, v, nranks, 0, -1, r);   // rank lists, concatenated
  *nitems = (long long)(v.size() / 4);
  if (items) {
    if ((long long)v.size() > cap * 4) { set_error("items buffer too small"); return 1; }
    std::memcpy(items, v.data(), sizeof(int) * v.size());
  }
  return 0;
}

int crest_cuda_tf32_items_rank(int nall, const double* er, double ethr, int nranks, int rank, int* items,
                               long long cap, long long* nitems) {
  if (nall < 0 || (!er && nall > 0) || !nitems || nranks < 1 || rank < 0 || rank >= nranks) {
    set_error("bad argument");
    return 1;
  }
  std::vector<int> v;
  tf_items_host(nall, er, ethr, v, nranks, 0, -1, nranks > 1 ? rank : -1);
  *nitems = (long long)(v.size() / 4);
  if (items) {
    if ((long long)v.size() > cap * 4) { set_error("items buffer too small"); return 1; }
    std::memcpy(items, v.data(), sizeof(int) * v.size());
  }
  return 0;
}

int crest_cuda_tf32_items(int nall, const double* er, double ethr, int* items, long long cap,
                          long long* nitems) {
  return crest_cuda_tf32_items_n(nall, er, ethr, 1, items, cap, nitems);
}

int crest_cuda_shard_range(long long ntiles, int rank, int nranks, long long* begin,
                           long long* end) {
  if (nranks < 1 || rank < 0 || rank >= nranks || ntiles < 0) { set_error("bad shard"); return 1; }
  if (begin) *begin = ntiles * rank / nranks;
  if (end) *end = ntiles * (rank + 1) / nranks;
  return 0;
}

int crest_cuda_ctx_create(crest_cuda_ctx** out, int device) {
  if (!out) { set_error("ctx out pointer is NULL"); return 1; }
  *out = nullptr;
  int ndev = 0;
  CREST_CUDA_TRY(cudaGetDeviceCount(&ndev));
  if (ndev <= 0) { set_error("no CUDA device available (there is no CPU fallback)"); return 1; }
  if (device < 0) CREST_CUDA_TRY(cudaGetDevice(&device));
  if (device >= ndev) { set_error("device %d out of range (%d devices)", device, ndev); return 1; }
  CREST_CUDA_TRY(cudaSetDevice(device));
  {
    // the small per-call scratch buffers come from cudaMallocAsync: keep the pool's memory across
    // synchronisations instead of returning it to the driver every time (the default threshold is 0)
    cudaMemPool_t pool;
    if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess) {
      unsigned long long keep = ~0ull;
      cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
    }
    cudaGetLastError();
  }
  crest_cuda_ctx* c = new crest_cuda_ctx();
  c->device = device;
  if (cudaMalloc(&c->d_npairs, sizeof(unsigned long long)) != cudaSuccess) {
    set_error("cudaMalloc failed: %s", cudaGetErrorString(cudaGetLastError()));
    delete c;
    return 1;
  }
  *out = c;
  return 0;
}

void crest_cuda_ctx_destroy(crest_cuda_ctx* c) {
  if (!c) return;
  cudaSetDevice(c->device);
  cudaFree(c->d_at); cudaFree(c->d_sel); cudaFree(c->d_xyz); cudaFree(c->d_xyz_axis);
  cudaFree(c->d_er); cudaFree(c->d_rot); cudaFree(c->d_aniso); cudaFree(c->d_X); cudaFree(c->d_xn);
  cudaFree(c->d_dup); cudaFree(c->d_npairs); cudaFree(c->d_rowstart); cudaFree(c->d_jfirst); cudaFree(c->d_items);
  cudaFree(c->d_mask); cudaFree(c->d_rmat);
  cudaFree(c->tf.Xa); cudaFree(c->tf.Xb); cudaFree(c->tf.hn); cudaFree(c->tf.cmask); cudaFree(c->tf.fmask);
  cudaFree(c->tf.item_cnt); cudaFree(c->tf.dense); cudaFree(c->tf.dense_items); cudaFree(c->tf.unfinished);
  cudaFree(c->tf.cand_count); cudaFree(c->d_npairs_scratch); cudaFree(c->d_tf_items);
  delete c;
}

int crest_cuda_ctx_upload(crest_cuda_ctx* c, int nat, int nall, const int* at, const double* xyz,
                          const double* er, const int* rmsd_mask, int rot_from_mask) {
  if (!c) { set_error("ctx is NULL"); return 1; }
  if (nat <= 0 || nall < 0) { set_error("bad sizes nat=%d nall=%d", nat, nall); return 1; }
  // xyz == NULL: reserve only; the geometry arrives through crest_cuda_ctx_upload_xyz_rows and/or a
  // device-side all-gather into crest_cuda_ctx_xyz_devptr (multi-GPU: each rank uploads a slice)
  if (!at || (!er && nall > 0)) { set_error("NULL input array"); return 1; }
  for (int i = 1; i < nall; ++i)
    if (er[i] < er[i - 1]) {
      set_error("energies must ascend (cregen_esort post-condition): er[%d] < er[%d]", i, i - 1);
      return 1;
    }
  CREST_CUDA_TRY(cudaSetDevice(c->device));
  std::vector<int> sel;
  sel.reserve(nat);
  for (int k = 0; k < nat; ++k)
    if (!rmsd_mask || rmsd_mask[k] > 0) sel.push_back(k);
  if (sel.empty()) { set_error("RMSD atom mask selects no atom"); return 1; }
  c->nat = nat; c->nall = nall; c->nsel = (int)sel.size(); c->rot_from_mask = rot_from_mask ? 1 : 0;
  c->use_col = col_layout_fits(c->nsel) && !tune().force_tile;
  c->L = c->use_col ? make_col_layout(c->nsel, nall) : make_layout(c->nsel, nall);
  c->items_ethr = std::numeric_limits<double>::quiet_NaN();
  c->tf_items_ethr = std::numeric_limits<double>::quiet_NaN();
  c->stage_tile.clear();
  c->dup_target = nullptr;
  c->use_tf32 = !tune().no_tf32 && !c->fp64_only && nall >= 2;
  if (c->use_col) { c->tf.nslab = 1; c->tf.kslab = c->L.KU; }
  else { c->tf.kslab = kTfSlabSlots; c->tf.nslab = (c->L.nslab * c->L.KU + kTfSlabSlots - 1) / kTfSlabSlots; }
  c->ntile = (nall + 127) / 128;
  c->h_er.assign(er, er + nall);
  c->tiles_ethr = std::numeric_limits<double>::quiet_NaN();

  const size_t nxyz = (size_t)3 * nat * ((size_t)nall + 64);   // 64 rows of slack: equal slices for <= 64 ranks
  if ((size_t)nat > c->cap_at) {
    cudaFree(c->d_at); cudaFree(c->d_sel);
    c->d_at = nullptr; c->d_sel = nullptr; c->cap_at = 0;
    CREST_CUDA_TRY(cudaMalloc(&c->d_at, sizeof(int) * nat));
    CREST_CUDA_TRY(cudaMalloc(&c->d_sel, sizeof(int) * nat));
    c->cap_at = nat;
  }
  if (nxyz > c->cap_xyz) {
    cudaFree(c->d_xyz); cudaFree(c->d_xyz_axis);
    c->d_xyz = nullptr; c->d_xyz_axis = nullptr; c->cap_xyz = 0;
    CREST_CUDA_TRY(cudaMalloc(&c->d_xyz, sizeof(double) * (nxyz ? nxyz : 1)));
    CREST_CUDA_TRY(cudaMalloc(&c->d_xyz_axis, sizeof(double) * (nxyz ? nxyz : 1)));
    c->cap_xyz = nxyz;
  }
  if ((size_t)nall > c->cap_conf) {
    cudaFree(c->d_er); cudaFree(c->d_rot); cudaFree(c->d_aniso); cudaFree(c->d_xn); cudaFree(c->d_dup);
    c->d_er = c->d_rot = c->d_aniso = c->d_xn = nullptr; c->d_dup = nullptr; c->cap_conf = 0;
    const size_t m = nall ? nall : 1;
    CREST_CUDA_TRY(cudaMalloc(&c->d_er, sizeof(double) * m));
    CREST_CUDA_TRY(cudaMalloc(&c->d_rot, sizeof(double) * 3 * m));
    CREST_CUDA_TRY(cudaMalloc(&c->d_aniso, sizeof(double) * m));
    CREST_CUDA_TRY(cudaMalloc(&c->d_xn, sizeof(double) * m));
    CREST_CUDA_TRY(cudaMalloc(&c->d_dup, sizeof(int) * m));
    c->cap_conf = nall;
  }
  const size_t nX = c->L.total_doubles();
  if (nX > c->cap_X) {
    cudaFree(c->d_X); c->d_X = nullptr; c->cap_X = 0;
    CREST_CUDA_TRY(cudaMalloc(&c->d_X, sizeof(double) * nX));
    c->cap_X = nX;
    c->X_zeroed = false;
  }
  if (c->use_tf32) {
    const size_t nx = (size_t)c->ntile * 128 * 3 * c->tf.kslab * c->tf.nslab;    // floats, same for Xa and Xb
    if (nx > c->cap_tf_x) {
      cudaFree(c->tf.Xa); cudaFree(c->tf.Xb); c->tf.Xa = c->tf.Xb = nullptr; c->cap_tf_x = 0;
      CREST_CUDA_TRY(cudaMalloc(&c->tf.Xa, sizeof(float) * nx));
      CREST_CUDA_TRY(cudaMalloc(&c->tf.Xb, sizeof(float) * nx));
      c->cap_tf_x = nx;
    }
    const size_t nh = (size_t)c->ntile * 128;
    if (nh > c->cap_tf_hn) {
      cudaFree(c->tf.hn); c->tf.hn = nullptr; c->cap_tf_hn = 0;
      CREST_CUDA_TRY(cudaMalloc(&c->tf.hn, sizeof(float) * nh));
      c->cap_tf_hn = nh;
    }
    if (!c->tf.cand_count) {
      CREST_CUDA_TRY(cudaMalloc(&c->tf.cand_count, sizeof(unsigned) * 16));   // counters, see Tf32Buffers
      CREST_CUDA_TRY(cudaMemset(c->tf.cand_count, 0, sizeof(unsigned) * 16));
      CREST_CUDA_TRY(cudaMalloc(&c->d_npairs_scratch, sizeof(unsigned long long)));
    }
    // candidate bitmask (pair_tf32.cu): one pass over all columns if that fits the budget (default
    // 2 GB, CREST_CUDA_TUNE maskmb=<MB>), else passes of whole segments; never less than one segment
    // of 4096 columns.  Zeroed once: cand_scan_kernel clears what the screen sets.
    {
      const size_t full = tf32_mask_words(0, nall, 0, nall, 1);
      const size_t budget = ((size_t)(tune().mask_mb > 0 ? tune().mask_mb : 2048) << 20) / sizeof(unsigned);
      const size_t one_seg = ((size_t)nall + 256) * 128;
      const size_t want = std::max<size_t>(std::min(full, std::max(budget, one_seg)), 128);
      if (want > c->tf.cm_words || (tune().mask_mb > 0 && want != c->tf.cm_words)) {
        cudaFree(c->tf.cmask); cudaFree(c->tf.fmask);
        c->tf.cmask = c->tf.fmask = nullptr; c->tf.cm_words = 0;
        CREST_CUDA_TRY(cudaMalloc(&c->tf.cmask, sizeof(unsigned) * want));
        CREST_CUDA_TRY(cudaMemset(c->tf.cmask, 0, sizeof(unsigned) * want));
        c->tf.cm_words = want;
      }
    }
    if ((size_t)nall > c->tf.unfinished_cap) {
      cudaFree(c->tf.unfinished); c->tf.unfinished = nullptr; c->tf.unfinished_cap = 0;
      CREST_CUDA_TRY(cudaMalloc(&c->tf.unfinished, (size_t)nall));
      c->tf.unfinished_cap = nall;
    }
    c->tf.kappa = tf32_kappa(c->tf.kslab * c->tf.nslab);
  }
  CREST_CUDA_TRY(cudaMemcpy(c->d_at, at, sizeof(int) * nat, cudaMemcpyHostToDevice));
  CREST_CUDA_TRY(cudaMemcpy(c->d_sel, sel.data(), sizeof(int) * sel.size(), cudaMemcpyHostToDevice));
  if (nall > 0) {
    if (xyz)
      if (big_copy(c->d_xyz, xyz, sizeof(double) * 3 * nat * (size_t)nall, 0, c->device)) return 1;
    CREST_CUDA_TRY(cudaMemcpy(c->d_er, er, sizeof(double) * nall, cudaMemcpyHostToDevice));
  }
  return 0;
}

int crest_cuda_ctx_upload_xyz_rows(crest_cuda_ctx* c, int first, int count, const double* xyz_rows) {
  if (!c) { set_error("ctx is NULL"); return 1; }
  if (first < 0 || count < 0 || first + count > c->nall || (!xyz_rows && count > 0)) {
    set_error("bad row range [%d, %d) of %d", first, first + count, c->nall);
    return 1;
  }
  CREST_CUDA_TRY(cudaSetDevice(c->device));
  const size_t rec = (size_t)3 * c->nat;
  if (count > 0)
    if (big_copy(c->d_xyz + rec * first, xyz_rows, sizeof(double) * rec * count, 0, c->device)) return 1;
  return 0;
}

void* crest_cuda_ctx_xyz_devptr(crest_cuda_ctx* c) { return c ? (void*)c->d_xyz : nullptr; }

}  // extern "C"

namespace crest {
int begin_call(crest_cuda_ctx* c, cudaStream_t st) {
  CREST_CUDA_TRY(cudaSetDevice(c->device));
  // padding rows / atoms of X must be zero: they enter the MMA (atoms) or are masked (rows)
  if (!c->X_zeroed || std::memcmp(&c->L, &c->X_zero_L, sizeof(PackLayout)) != 0 ||
      c->nall != c->X_zero_nall) {
    CREST_CUDA_TRY(cudaMemsetAsync(c->d_X, 0, sizeof(double) * c->L.total_doubles(), st));
    c->X_zeroed = true; c->X_zero_L = c->L; c->X_zero_nall = c->nall;
  }
  CREST_CUDA_TRY(cudaMemsetAsync(c->d_dup, 0, sizeof(int) * (size_t)(c->nall ? c->nall : 1), st));
  CREST_CUDA_TRY(cudaMemsetAsync(c->d_npairs, 0, sizeof(unsigned long long), st));
  if (c->use_tf32) {
    CREST_CUDA_TRY(cudaMemsetAsync(c->tf.cand_count, 0, sizeof(unsigned) * 16, st));
  }
  c->launches = 0;
  return 0;
}

int pack_rows(crest_cuda_ctx* c, int r0, int r1, cudaStream_t st) {
  if (r1 <= r0) return 0;
  if (r0 % 128 != 0) { set_error("pack_rows: row block must start on a 128-row tile"); return 1; }
  const size_t rec = (size_t)3 * c->nat;
  const size_t xoff = (size_t)(r0 / c->L.rows) * c->L.nslab * c->L.chunk_doubles();
  if (launch_pack_axis(c->nat, r1 - r0, c->d_at, c->d_xyz + rec * r0, c->d_sel, c->nsel, c->rot_from_mask,
                       c->L, c->d_xyz_axis + rec * r0, c->d_rot + (size_t)3 * r0, c->d_aniso + r0,
                       c->d_X + xoff, c->d_xn + r0, st))
    return 1;
  c->launches += 1;
  if (c->use_tf32) {
    const int t0 = r0 / 128, t1 = (r1 + 127) / 128;
    if (launch_tf32_convert(c->d_X, c->d_xn, c->nall, c->L, t0, t1, c->tf.nslab, c->tf.kslab, c->tf.kappa,
                            c->tf.Xa, c->tf.Xb, c->tf.hn, c->tf.cand_count + 4, st))
      return 1;
    c->launches += 1;
  }
  return 0;
}
}  // namespace crest

extern "C" {

int crest_cuda_ctx_pack(crest_cuda_ctx* c, void* stream) {
  if (!c) { set_error("ctx is NULL"); return 1; }
  cudaStream_t st = (cudaStream_t)stream;
  if (begin_call(c, st)) return 1;
  return pack_rows(c, 0, c->nall, st);
}

}  // extern "C"

namespace crest {
bool tf32_path(const crest_cuda_ctx* c, int want_mask, const float* d_rmat) {
  const char* dbg = std::getenv("CREST_CUDA_DEBUG");
  const int debug = dbg ? std::atoi(dbg) : 0;
  return c->use_tf32 && !want_mask && !d_rmat && debug != 1 && debug != 2;
}

int pairs_stage(crest_cuda_ctx* c, int enantio, const crest_cuda_thresholds* thr, int rank, int nranks,
                int stage, int want_mask, float* d_rmat, cudaStream_t st) {
  if (check_thr(thr)) return 1;
  if (nranks < 1 || rank < 0 || rank >= nranks) { set_error("bad shard %d/%d", rank, nranks); return 1; }
  CREST_CUDA_TRY(cudaSetDevice(c->device));
  if (c->nall < 2) return 0;
  PairArgs a{};
  a.X = c->d_X; a.xn = c->d_xn; a.er = c->d_er; a.rot = c->d_rot; a.aniso = c->d_aniso;
  a.nall = c->nall;
  a.dup = c->dup_target ? c->dup_target : c->d_dup; a.npairs = c->d_npairs;
  a.mask = nullptr; a.n8 = (c->nall + 7) / 8;
  a.rmat = d_rmat;
  a.g = make_gates(*thr, enantio, c->nsel);
  a.L = c->L;
  a.dbg = nullptr;
  a.work_count = nullptr; a.tiles = nullptr;
  if (tf32_path(c, want_mask, d_rmat)) {
    // tensor-core screen -> per-item verdict -> top-down exact scan of the candidate bitmask -> FP64
    // tensor kernel on what dense items left unfinished (pair_tf32.cu), pass by pass over this rank's
    // slice of the stage's list
    if (c->stage_tile.empty() || !(c->tf_items_ethr == thr->ethr && c->tf_items_nranks == nranks &&
                                   c->tf_items_rank == rank)) {
      const std::vector<int> one{0, c->ntile};
      if (plan_stages(c, thr->ethr, nranks, rank, one, st)) return 1;
    }
    if (stage < 0 || stage + 1 >= (int)c->stage_tile.size()) { set_error("bad stage %d", stage); return 1; }
    // the list already holds this rank's rows only
    const long long i0 = c->stage_item[stage], i1 = c->stage_item[stage + 1];
    const long long tb = 0, te = i1 - i0;
    if (a.g.debug == 3 && !c->tf.fmask) {
      CREST_CUDA_TRY(cudaMalloc(&c->tf.fmask, sizeof(unsigned) * c->tf.cm_words));
    }
    PairArgs fb = a;
    fb.npairs = c->d_npairs_scratch;      // the screen already counted these pairs
    int nsf = 0;
    if (!c->use_col) pair_smem_bytes(c->L, &nsf);
    Tf32Buffers tb_ = c->tf;
    tb_.chunk = c->stage_chunk[stage];
    const int row_lo = c->stage_tile[stage] * 128, row_hi = std::min(c->stage_tile[stage + 1] * 128, c->nall);
    if (launch_pair_tf32(a, fb, tb_, c->d_tf_items, c->tf_items.data(), (int)(i0 + tb), (int)(i0 + te),
                         row_lo, row_hi, rank, nranks, a.g.debug == 3, nsf, st, &c->last_grid, &c->last_smem, &c->launches))
      return 1;
    c->last_variant = 2;
    return 0;
  }
  if (stage != 0) { set_error("the FP64-only kernels run as one stage"); return 1; }
  if (c->use_col) {
    if (!(c->items_ethr == thr->ethr) && upload_items(c, thr->ethr, st)) return 1;
  } else if (!(c->tiles_ethr == thr->ethr)) {
    build_tiles(c, thr->ethr);
    if (upload_tiles(c, st)) return 1;
  }
  const long long T = c->use_col ? (long long)(c->items.size() / 4) : c->rowstart.back();
  a.rowstart = c->d_rowstart; a.jfirst = c->d_jfirst;
  a.nI = (int)c->jfirst.size();
  crest_cuda_shard_range(T, rank, nranks, &a.tile_begin, &a.tile_end);
  if (want_mask) {
    const size_t need = (size_t)a.n8 * a.n8;
    if (need > c->cap_mask) {
      cudaFree(c->d_mask); c->d_mask = nullptr; c->cap_mask = 0;
      CREST_CUDA_TRY(cudaMalloc(&c->d_mask, sizeof(unsigned long long) * need));
      c->cap_mask = need;
    }
    CREST_CUDA_TRY(cudaMemsetAsync(c->d_mask, 0, sizeof(unsigned long long) * need, st));
    a.mask = c->d_mask;
  }
  if (a.g.debug == 2) {
    CREST_CUDA_TRY(cudaMalloc(&a.dbg, sizeof(long long) * 32));
    CREST_CUDA_TRY(cudaMemsetAsync(a.dbg, 0, sizeof(long long) * 32, st));
  }
  int ns = 0;
  c->last_variant = c->use_col ? 1 : 0;
  if (c->use_col) {
    if (launch_pair_cols(a, c->d_items, (int)a.tile_begin, (int)a.tile_end, st, &c->last_grid,
                         &c->last_smem))
      return 1;
  } else {
    pair_smem_bytes(c->L, &ns);
    if (launch_pair_tiles(a, ns, st, &c->last_grid, &c->last_smem)) return 1;
  }
  if (a.dbg) {
    long long h[32];
    CREST_CUDA_TRY(cudaStreamSynchronize(st));
    CREST_CUDA_TRY(cudaMemcpy(h, a.dbg, sizeof(h), cudaMemcpyDeviceToHost));
    cudaFree(a.dbg);
    const double g = c->last_grid > 0 ? c->last_grid : 1;
    for (int w = 0; w < 8; ++w)
      std::fprintf(stderr, "[pair dbg] warp %d  avg clk/CTA: produce+meta %.0f  wait_full %.0f  mma(+screen) %.0f  epilogue/exact %.0f\n",
                   w, h[4 * w] / g, h[4 * w + 1] / g, h[4 * w + 2] / g, h[4 * w + 3] / g);
  }
  if (c->last_grid > 0) c->launches += 1;
  return 0;
}
}  // namespace crest

extern "C" {

int crest_cuda_ctx_pairs(crest_cuda_ctx* c, int enantio, const crest_cuda_thresholds* thr,
                         int rank, int nranks, int want_mask, void* stream) {
  if (!c) { set_error("ctx is NULL"); return 1; }
  return pairs_stage(c, enantio, thr, rank, nranks, 0, want_mask, nullptr, (cudaStream_t)stream);
}

void* crest_cuda_ctx_dup_devptr(crest_cuda_ctx* c) { return c ? (void*)c->d_dup : nullptr; }
void* crest_cuda_ctx_npairs_devptr(crest_cuda_ctx* c) { return c ? (void*)c->d_npairs : nullptr; }

int crest_cuda_ctx_sweep(crest_cuda_ctx* c, void* stream) {
  if (!c) { set_error("ctx is NULL"); return 1; }
  CREST_CUDA_TRY(cudaSetDevice(c->device));
  if (launch_backtrack(c->nall, c->d_dup, (cudaStream_t)stream)) return 1;
  if (c->nall > 0) c->launches += 1;
  return 0;
}

int crest_cuda_ctx_download(crest_cuda_ctx* c, double* rot, double* xyz_axis, int* dup,
                            long long* npairs, void* stream) {
  if (!c) { set_error("ctx is NULL"); return 1; }
  cudaStream_t st = (cudaStream_t)stream;
  CREST_CUDA_TRY(cudaSetDevice(c->device));
  const size_t n = c->nall;
  if (rot && n)
    CREST_CUDA_TRY(cudaMemcpyAsync(rot, c->d_rot, sizeof(double) * 3 * n, cudaMemcpyDeviceToHost, st));
  if (xyz_axis && n) {
    CREST_CUDA_TRY(cudaStreamSynchronize(st));
    if (big_copy(xyz_axis, c->d_xyz_axis, sizeof(double) * 3 * c->nat * n, 1, c->device)) return 1;
  }
  if (dup && n)
    CREST_CUDA_TRY(cudaMemcpyAsync(dup, c->d_dup, sizeof(int) * n, cudaMemcpyDeviceToHost, st));
  if (npairs)
    CREST_CUDA_TRY(cudaMemcpyAsync(npairs, c->d_npairs, sizeof(long long), cudaMemcpyDeviceToHost, st));
  CREST_CUDA_TRY(cudaStreamSynchronize(st));
  return 0;
}

int crest_cuda_ctx_info(crest_cuda_ctx* c, const crest_cuda_thresholds* thr, int rank, int nranks,
                        long long* info) {
  if (!c || !info) { set_error("NULL argument"); return 1; }
  if (check_thr(thr)) return 1;
  if (nranks < 1 || rank < 0 || rank >= nranks) { set_error("bad shard %d/%d", rank, nranks); return 1; }
  if (c->last_variant == 2) {
    // work list of the tensor-core screen: built by ctx_pairs
  } else if (c->nall >= 2 && c->use_col) {
    if (!(c->items_ethr == thr->ethr)) {
      CREST_CUDA_TRY(cudaSetDevice(c->device));
      if (upload_items(c, thr->ethr, 0)) return 1;
      CREST_CUDA_TRY(cudaStreamSynchronize(0));
    }
  } else if (c->nall >= 2 && !(c->tiles_ethr == thr->ethr)) {
    build_tiles(c, thr->ethr);
    CREST_CUDA_TRY(cudaSetDevice(c->device));
    if (upload_tiles(c, 0)) return 1;
    CREST_CUDA_TRY(cudaStreamSynchronize(0));
  }
  long long T = c->use_col ? (long long)(c->items.size() / 4)
                           : (c->rowstart.empty() ? 0 : c->rowstart.back());
  if (c->last_variant == 2) T = (long long)(c->tf_items.size() / 4);   // this rank's list, all stages
  long long tb = 0, te = 0;
  crest_cuda_shard_range(T, rank, nranks, &tb, &te);
  if (c->last_variant == 2) { tb = 0; te = T; }
  info[7] = c->last_variant;
  info[8] = info[9] = info[10] = info[11] = info[12] = info[13] = info[14] = 0;
  if (c->last_variant == 2 && c->tf.cand_count) {
    unsigned h[16] = {};
    CREST_CUDA_TRY(cudaSetDevice(c->device));
    CREST_CUDA_TRY(cudaDeviceSynchronize());
    CREST_CUDA_TRY(cudaMemcpy(h, c->tf.cand_count, sizeof(h), cudaMemcpyDeviceToHost));
    info[8] = h[3]; info[9] = h[5]; info[10] = h[2]; info[11] = h[0]; info[12] = h[1] | (h[7] << 1);
    info[13] = h[8]; info[14] = h[9];
  }
  info[0] = te - tb;
  info[1] = T;
  info[2] = c->nsel;
  info[3] = (long long)c->L.nslab * c->L.KU;
  info[4] = c->last_grid;
  info[5] = c->last_smem;
  info[6] = c->launches;
  return 0;
}

// Probe of the tensor core's FP32 accumulation: runs the screen over all pairs of the packed ensemble with
// the raw TMEM covariances dumped, recomputes every dot product of the same TF32-rounded operands in FP64
// and returns the largest |R~_ab - R_ab| / (|x_a||y_b|) - the quantity the screen's gamma budgets.
int crest_cuda_ctx_tf32_accum(crest_cuda_ctx* c, double* max_rel_err, long long* ndots, double* gamma_budget,
                              double* kappa) {
  if (!c || !max_rel_err || !ndots) { set_error("NULL argument"); return 1; }
  if (!c->use_tf32 || c->nall < 2) { set_error("the probe needs the tensor-core path and >= 2 structures"); return 1; }
  CREST_CUDA_TRY(cudaSetDevice(c->device));
  const size_t n = c->nall;
  float* d_dump = nullptr;
  unsigned long long* d_out = nullptr;
  CREST_CUDA_TRY(cudaMalloc(&d_dump, sizeof(float) * 9 * n * n));
  if (cudaMalloc(&d_out, 16) != cudaSuccess) { cudaFree(d_dump); set_error("cudaMalloc failed"); return 1; }
  cudaMemset(d_dump, 0, sizeof(float) * 9 * n * n);
  cudaMemset(d_out, 0, 16);
  crest_cuda_thresholds t{1e300, 0.125, 0.01, 0.025, 0.5};
  c->tf.dump = d_dump;
  int rc = pairs_stage(c, 1, &t, 0, 1, 0, 0, nullptr, nullptr);
  c->tf.dump = nullptr;
  if (!rc) rc = launch_tf32_accum_check(c->tf, c->nall, d_dump, d_out, nullptr);
  unsigned long long h[2] = {0, 0};
  if (!rc && cudaMemcpy(h, d_out, 16, cudaMemcpyDeviceToHost) != cudaSuccess) {
    set_error("probe failed: %s", cudaGetErrorString(cudaGetLastError()));
    rc = 1;
  }
  cudaFree(d_dump); cudaFree(d_out);
  if (rc) return rc;
  std::memcpy(max_rel_err, &h[0], 8);
  *ndots = (long long)h[1];
  const int K = c->tf.kslab * c->tf.nslab;
  if (gamma_budget) *gamma_budget = std::max(6.0e-5, 2.0 * K * std::ldexp(1.0, -23));
  if (kappa) *kappa = c->tf.kappa;
  return 0;
}

// Copy of the device bitmask (tests): mask[n8*n8] 64-bit words, bit r*8+col of word (ti,tj).
int crest_cuda_ctx_download_mask(crest_cuda_ctx* c, unsigned long long* mask, void* stream) {
  if (!c || !c->d_mask || !mask) { set_error("no bitmask was requested"); return 1; }
  const size_t n8 = (c->nall + 7) / 8;
  CREST_CUDA_TRY(cudaMemcpyAsync(mask, c->d_mask, sizeof(unsigned long long) * n8 * n8,
                                 cudaMemcpyDeviceToHost, (cudaStream_t)stream));
  CREST_CUDA_TRY(cudaStreamSynchronize((cudaStream_t)stream));
  return 0;
}

// Host-side: the reference's energy sort is an UNSTABLE Hoare quicksort with middle pivot
// (src/quicksort.f90:140-165), so the order of equal energies - which every later index inherits -
// is a property of that exact partition sequence.  Restated with an explicit stack (the two
// sub-ranges are disjoint, their order does not matter); then maskinvert (quicksort.f90:278-291)
// and the ewin cut of cregen_esort (src/cregen.f90:1096-1113).
int crest_cuda_cregen_esort(int nat, int nall, double* xyz, double* energies, double ewin,
                            int* order, int* nallout) {
  if (nall < 0 || (!energies && nall) || !order || !nallout || (xyz && nat <= 0)) {
    set_error("bad argument");
    return 1;
  }
  std::vector<int> was(nall);              // after sorting: element at position p WAS at was[p]
  for (int i = 0; i < nall; ++i) was[i] = i;
  double* a = energies;
  // The reference's quicksort (quicksort.f90:140-165) - Hoare partition, middle pivot, unstable: the order of
  // equal energies is part of the contract.  After a partition the two sub-ranges never interact, so they can
  // be sorted by different threads without changing a single swap: the first few levels hand their left
  // range to a new thread (large ranges only), the rest is the serial loop on an explicit stack.
  std::function<void(int, int, int)> sort_range = [&](int lo, int hi, int depth) {
    std::vector<std::pair<int, int>> stack;
    std::vector<std::thread> helpers;
    stack.emplace_back(lo, hi);
    while (!stack.empty()) {
      const int first = stack.back().first, last = stack.back().second;
      stack.pop_back();
      // 1-based (first+last)/2 of the reference == 0-based (first+last+2)/2 - 1
      const double x = a[(first + last + 2) / 2 - 1];
      int i = first, j = last;
      for (;;) {
        while (a[i] < x) ++i;
        while (x < a[j]) --j;
        if (i >= j) break;
        std::swap(a[i], a[j]);
        std::swap(was[i], was[j]);
        ++i; --j;
      }
      const bool left = first < i - 1, right = j + 1 < last;
      if (left && right && depth < 4 && (i - 1 - first) > (1 << 15) && (last - j - 1) > (1 << 15)) {
        const int l0 = first, l1 = i - 1;
        helpers.emplace_back([&sort_range, l0, l1, depth] { sort_range(l0, l1, depth + 1); });
        ++depth;                                           // this thread keeps the right range
        stack.emplace_back(j + 1, last);
        continue;
      }
      if (left) stack.emplace_back(first, i - 1);
      if (right) stack.emplace_back(j + 1, last);
    }
    for (auto& h : helpers) h.join();
  };
  if (nall > 0) sort_range(0, nall - 1, 0);
  for (int p = 0; p < nall; ++p) order[was[p]] = p + 1;     // structure i IS NOW at order[i] (1-based)
  if (xyz && nall > 0) {
    // the permutation of the geometries moves 24*nat*nall bytes twice (C5: 2 x 0.96 GB): up to 16 threads,
    // each a contiguous range of destination records, non-temporal stores (hostcopy.cu fast_copy), and a
    // scratch buffer that is NOT zero-filled first (a std::vector would add a third pass over it)
    const size_t rec = (size_t)3 * nat;
    // scratch up to 256 MB is kept between calls (CREST sorts 5-20 ensembles per search; a fresh mapping costs
    // a page fault and a zeroed page per 4 KB - as much as the copy itself), larger ones are returned at once
    static double* keep = nullptr;
    static size_t keep_bytes = 0;
    const size_t need = sizeof(double) * rec * nall;
    double* tmp = nullptr;
    if (need <= ((size_t)256 << 20)) {
      if (need > keep_bytes) {
        std::free(keep);
        keep = static_cast<double*>(std::malloc(need));
        keep_bytes = keep ? need : 0;
      }
      tmp = keep;
    } else {
      tmp = static_cast<double*>(std::malloc(need));
    }
    if (!tmp) { set_error("cregen_esort: out of host memory (%zu bytes)", need); return 1; }
    int T = (int)std::min<size_t>(16, std::max<size_t>(1, (rec * nall * sizeof(double)) >> 23));
    const int hw = (int)std::thread::hardware_concurrency();
    if (hw > 0 && T > hw) T = hw;
    auto part = [&](int t) {
      const size_t p0 = (size_t)nall * t / T, p1 = (size_t)nall * (t + 1) / T;
      for (size_t p = p0; p < p1; ++p)
        fast_copy(tmp + rec * p, xyz + rec * was[p], rec * sizeof(double));
      fast_copy_fence();
    };
    auto back = [&](int t) {
      const size_t p0 = (size_t)nall * t / T, p1 = (size_t)nall * (t + 1) / T;
      const size_t step = (size_t)1 << 17;                  // doubles per copy: 1 MB
      for (size_t o = rec * p0; o < rec * p1; o += step)
        fast_copy(xyz + o, tmp + o, std::min(step, rec * p1 - o) * sizeof(double));
      fast_copy_fence();
    };
    for (int phase = 0; phase < 2; ++phase) {
      std::vector<std::thread> th;
      for (int t = 1; t < T; ++t) th.emplace_back([&, t] { phase ? back(t) : part(t); });
      phase ? back(0) : part(0);
      for (auto& x : th) x.join();
    }
    if (tmp != keep) std::free(tmp);
  }
  int nout = nall;
  if (nall > 0) {
    double emax = a[0];
    for (int i = 1; i < nall; ++i) emax = std::max(emax, a[i]);
    double de = (emax - a[0]) * kAutokcal;
    if (de > ewin) {
      nout = 1;
      for (int i = 1; i < nall; ++i) {
        de = (a[i] - a[0]) * kAutokcal;
        if (de < ewin) ++nout; else break;
      }
    }
  }
  *nallout = nout;
  return 0;
}

}  // extern "C"
