// blockpair.cu — exact FP64 RMSD sweeps over rectangular and block-diagonal pair spaces: the callers
// of rmsd() either side of cregen_CRE.
//
//   f2  compare_ensembles main loop      /root/reference/src/ensemblecomp.f90:213-248
//       rmat(i,j) = min over rotamers k of conformer i (ensemble 1), l of conformer j (ensemble 2)
//       of rmsd(nat, c1, c2) (all atoms, no mirror) — the reference's seed `rmat2 = 100` is
//       overwritten for every (k,l), so it never survives into minval(); then the
//       "Correlation between Conformers" walk (ensemblecomp.f90:258-300).
//   f3  intraconfRMSD                    /root/reference/src/entropy/entropic.f90:938-1097
//       per conformer: rmat over its rotamers on the atoms inc(k) == 1, enantiomer probe
//       (rmat > rthr and mirror rmsd < 0.75 rthr), uniqueCore(rmat, nrot, rthr) = number of
//       "backbone" structures, enantiofac spread over equal symmetry symbols (symmap).
//
// Both need the RMSD itself as a double (no float32 rounding, no "far" screen).  They run on the SAME FP64
// tensor mainloops as CREGEN's FP64 path (pair_col.cu for <= 88 atoms, pair_tile.cu above; DMMA m8n8k4,
// bulk-copy operand rings): each ensemble is packed by pack_axis into the operand layout (the rigid motion
// it applies does not change an RMSD), the sweep is a rectangular (compare) or block-diagonal (intraconf)
// item list, and every valid pair leaves the kernel through sink_pair (qcp.cuh) instead of the CREGEN
// decisions.  Round 1 ran these on the CUDA cores at 0.17 of the DGEMM peak.
//
// uniqueCore (entropic.f90:1074-1095) is order dependent as written, but its count is not: i opens
// a new group iff mp(i) == 0 when the loop reaches it, and mp(i) is set by exactly those earlier
// i' < i with mat(i',i) < thr (every i' writes ALL its neighbours, whether or not it opened a
// group).  So  uniqueCore = #{ i : no j < i with rmat(i,j) < thr }  — a per-row OR like K4.
#include <algorithm>
#include <vector>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include "ctx.cuh"
#include "qcp.cuh"

namespace crest {

namespace {

struct DevBuf {   // frees on scope exit; every entry point below is a self-contained call
  std::vector<void*> ptrs;
  ~DevBuf() { for (void* q : ptrs) cudaFree(q); }
  template <class T> int alloc(T** out, size_t count) {
    void* q = nullptr;
    if (cudaMalloc(&q, sizeof(T) * (count ? count : 1)) != cudaSuccess) {
      set_error("cudaMalloc of %zu bytes failed: %s", sizeof(T) * count,
                cudaGetErrorString(cudaGetLastError()));
      return 1;
    }
    ptrs.push_back(q);
    *out = static_cast<T*>(q);
    return 0;
  }
};

struct CtxOwner {   // a context for the duration of one call
  crest_cuda_ctx* c = nullptr;
  ~CtxOwner() { crest_cuda_ctx_destroy(c); }
};

// ensemble -> FP64 operand layout (X, centred squared norms) through the CREGEN pack kernel.  The masses
// only steer the rigid frame pack_axis chooses, which no RMSD depends on: every atom counts as carbon.
int pack_operand(CtxOwner& o, int nat, int n, const double* xyz, const int* mask) {
  if (crest_cuda_ctx_create(&o.c, -1)) return 1;
  o.c->fp64_only = true;
  std::vector<int> at((size_t)nat, 6);
  std::vector<double> er((size_t)(n > 0 ? n : 1), 0.0);
  if (crest_cuda_ctx_upload(o.c, nat, n, at.data(), xyz, er.data(), mask, 0)) return 1;
  return crest_cuda_ctx_pack(o.c, nullptr);
}

// the sweep: work = column items {J, u_begin, u_end, 0} (A <= 88) or row tiles {panel, j-tile, 0, 0}
int run_sink(const crest_cuda_ctx* ci, const PairSink& sink, const std::vector<int4>& work) {
  if (work.empty()) return 0;
  DevBuf buf;
  int4* d_work;
  if (buf.alloc(&d_work, work.size())) return 1;
  CREST_CUDA_TRY(cudaMemcpy(d_work, work.data(), sizeof(int4) * work.size(), cudaMemcpyHostToDevice));
  PairArgs a{};
  a.sink = sink;
  a.X = ci->d_X; a.xn = ci->d_xn; a.er = ci->d_er; a.rot = ci->d_rot; a.aniso = ci->d_aniso;
  a.nall = ci->nall; a.dup = ci->d_dup; a.npairs = ci->d_npairs; a.n8 = (ci->nall + 7) / 8;
  a.L = ci->L;
  a.g.ethr = 1e300; a.g.E1 = a.g.E2 = 0.0; a.g.E2_up = 0.0f; a.g.enantio = 0; a.g.debug = 0;
  int grid = 0, smem = 0;
  const bool timing = getenv("CREST_CUDA_TIMING") && atoi(getenv("CREST_CUDA_TIMING")) > 0;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  if (timing) { cudaEventCreate(&e0); cudaEventCreate(&e1); cudaEventRecord(e0, nullptr); }
  if (ci->use_col) {
    if (launch_pair_cols(a, d_work, 0, (int)work.size(), nullptr, &grid, &smem)) return 1;
  } else {
    int ns = 0;
    pair_smem_bytes(ci->L, &ns);
    a.tiles = d_work; a.tile_begin = 0; a.tile_end = (long long)work.size();
    if (launch_pair_tiles(a, ns, nullptr, &grid, &smem)) return 1;
  }
  if (timing) cudaEventRecord(e1, nullptr);
  CREST_CUDA_TRY(cudaDeviceSynchronize());
  if (timing) {
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    const double pairs = sink.mode == 1 ? (double)sink.ni * sink.nj : 0.0;
    fprintf(stderr, "[blockpair] sweep kernel (%s, %zu items, grid %d): %.3f ms%s", ci->use_col ? "pair_col" : "pair_tile",
            work.size(), grid, ms, pairs > 0 ? "" : "\n");
    if (pairs > 0) fprintf(stderr, " = %.3g pairs/s, %.2f TFLOP/s (18*A per pair)\n", pairs / (ms * 1e-3),
                           18.0 * sink.A * pairs / (ms * 1e-3) / 1e12);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
  }
  return 0;
}

// 1-based inclusive [b(c), e(c)] blocks -> block id per structure (-1 = in no block)
int block_ids(const char* what, int nall, int nconf, const int* b, const int* e, std::vector<int>& blk) {
  blk.assign((size_t)nall, -1);
  for (int c = 0; c < nconf; ++c) {
    if (b[c] < 1 || e[c] > nall || e[c] < b[c]) {
      set_error("%s: block %d = [%d,%d] outside 1..%d or empty", what, c + 1, b[c], e[c], nall);
      return 1;
    }
    for (int s = b[c] - 1; s < e[c]; ++s) blk[s] = c;   // later blocks win, as the reference's copies would
  }
  return 0;
}

}  // namespace
}  // namespace crest

using namespace crest;

extern "C" {

int crest_cuda_compare_blocks(int nat, int nall1, const double* xyz1, int nconf1, const int* b1,
                              const int* e1, int nall2, const double* xyz2, int nconf2,
                              const int* b2, const int* e2, double* rmat) {
  if (nat <= 0 || nall1 < 0 || nall2 < 0 || nconf1 < 0 || nconf2 < 0 || !rmat ||
      (nconf1 && (!b1 || !e1 || !xyz1)) || (nconf2 && (!b2 || !e2 || !xyz2))) {
    set_error("bad argument"); return 1;
  }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    set_error("no CUDA device (there is no CPU fallback)"); return 1;
  }
  const size_t nout = (size_t)nconf1 * nconf2;
  if (nout == 0) return 0;
  std::vector<int> blk1, blk2;
  if (block_ids("ensemble 1", nall1, nconf1, b1, e1, blk1)) return 1;
  if (block_ids("ensemble 2", nall2, nconf2, b2, e2, blk2)) return 1;
  // only the structures up to the last block matter (maxcompare cut, ensemblecomp.f90:119,157)
  int n1 = 0, n2 = 0;
  for (int c = 0; c < nconf1; ++c) n1 = e1[c] > n1 ? e1[c] : n1;
  for (int c = 0; c < nconf2; ++c) n2 = e2[c] > n2 ? e2[c] : n2;
  CtxOwner c1, c2;
  if (pack_operand(c1, nat, n1, xyz1, nullptr) || pack_operand(c2, nat, n2, xyz2, nullptr)) return 1;
  DevBuf buf;
  int *d_b1, *d_b2; unsigned long long* d_r;
  if (buf.alloc(&d_b1, (size_t)n1) || buf.alloc(&d_b2, (size_t)n2) || buf.alloc(&d_r, nout)) return 1;
  CREST_CUDA_TRY(cudaMemcpy(d_b1, blk1.data(), sizeof(int) * n1, cudaMemcpyHostToDevice));
  CREST_CUDA_TRY(cudaMemcpy(d_b2, blk2.data(), sizeof(int) * n2, cudaMemcpyHostToDevice));
  // +inf seed: the reference's 100 (ensemblecomp.f90:228) is overwritten by every rotamer pair
  std::vector<double> seed(nout, __builtin_inf());
  CREST_CUDA_TRY(cudaMemcpy(d_r, seed.data(), sizeof(double) * nout, cudaMemcpyHostToDevice));
  PairSink k{};
  k.mode = 1; k.Xj = c2.c->d_X; k.xnj = c2.c->d_xn; k.ni = n1; k.nj = n2;
  k.blk_i = d_b1; k.blk_j = d_b2; k.nconf1 = nconf1; k.rmat_bits = d_r; k.A = nat;
  // the whole n1 x n2 rectangle: ensemble 2 in resident j-tiles of 64, ensemble 1 streaming past them
  std::vector<int4> work;
  const int nJ = (n2 + 63) / 64;
  if (c1.c->use_col) {
    const int nu = (n1 + kUnit - 1) / kUnit;
    int nseg = (148 * 8 + nJ - 1) / nJ;                  // >= 8 items per SM
    if (nseg > nu) nseg = nu;
    if (nseg < 1) nseg = 1;
    for (int sg = 0; sg < nseg; ++sg) {                  // segment-major: concurrent CTAs stream the same i-units
      const int u0 = (int)((long long)nu * sg / nseg), u1 = (int)((long long)nu * (sg + 1) / nseg);
      if (u1 > u0) for (int J = 0; J < nJ; ++J) work.push_back(make_int4(J, u0, u1, 0));
    }
  } else {
    const int np = (n1 + kPanel - 1) / kPanel;
    for (int P = 0; P < np; ++P)
      for (int J = 0; J < nJ; ++J) work.push_back(make_int4(P, J, 0, 0));
  }
  if (run_sink(c1.c, k, work)) return 1;
  CREST_CUDA_TRY(cudaMemcpy(rmat, d_r, sizeof(double) * nout, cudaMemcpyDeviceToHost));
  return 0;
}

int crest_cuda_compare_correlate(int nconf1, int nconf2, const double* en1, const double* en2,
                                 const double* rmat, double rthr, int* left, int* right) {
  if (nconf1 < 0 || nconf2 < 0 || ((nconf1 + nconf2) && (!left || !right)) ||
      (nconf1 && !en1) || (nconf2 && !en2) || ((size_t)nconf1 * nconf2 && !rmat)) {
    set_error("bad argument"); return -1;
  }
  // ensemblecomp.f90:262-300: walk both lists from the top; a line pairs kk with ll when
  // rmat(kk,ll) <= rthr, else it shows the one with the higher (>=) energy
  int kk = nconf1, ll = nconf2, nl = 0;
  while (kk > 0 || ll > 0) {
    bool matching = false;
    if (kk > 0 && ll > 0 && rmat[(kk - 1) + (size_t)nconf1 * (ll - 1)] <= rthr) {
      left[nl] = kk--; right[nl] = ll--; ++nl;
      matching = true;
    }
    if (!matching) {
      // the reference runs its three "if" blocks in sequence within one pass, so one pass can
      // emit two lines (e.g. the first block empties ll, the second then prints kk)
      if (kk > 0 && ll > 0) {
        if (en1[kk - 1] >= en2[ll - 1]) { left[nl] = kk--; right[nl] = 0; ++nl; }
        else { left[nl] = 0; right[nl] = ll--; ++nl; }
      }
      if (kk > 0 && ll == 0) { left[nl] = kk--; right[nl] = 0; ++nl; }
      if (kk == 0 && ll > 0) { left[nl] = 0; right[nl] = ll--; ++nl; }
    }
  }
  return nl;
}

int crest_cuda_intraconf(int nat, int nall, const double* xyz, const int* inc, int nconf,
                         const int* low, const int* nrot, const char* symsym, double* fac,
                         double* enantiofac) {
  if (nat <= 0 || nall < 0 || nconf < 0 || !inc || (nconf && (!low || !nrot || !symsym || !fac || !enantiofac))) {
    set_error("bad argument"); return 1;
  }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    set_error("no CUDA device (there is no CPU fallback)"); return 1;
  }
  std::vector<int> sel;
  int rednat = 0;
  for (int k = 0; k < nat; ++k) {
    rednat += inc[k];                                  // rednat = sum(inc) (entropic.f90:985)
    if (inc[k] == 1) sel.push_back(k);                 // cpincluded copies inc(k) == 1
  }
  if (rednat < 2) {                                    // entropic.f90:988-991
    for (int c = 0; c < nconf; ++c) fac[c] = 1.0;
    return 0;
  }
  if ((int)sel.size() != rednat) { set_error("inc(:) must hold 0/1 flags"); return 1; }
  const double rthr = 0.125;                           // entropic.f90:974
  int nmax = 0;
  std::vector<int> conf_of((size_t)(nall > 0 ? nall : 1), -1);
  for (int c = 0; c < nconf; ++c) {
    if (nrot[c] <= 1) continue;
    const int lo = low[c] - 1, hi = lo + nrot[c];
    if (lo < 0 || hi > nall) { set_error("conformer %d: rotamers [%d,%d] outside 1..%d", c + 1, low[c], hi, nall); return 1; }
    nmax = hi > nmax ? hi : nmax;
    for (int s = lo; s < hi; ++s) conf_of[s] = c;      // later conformers win, as in the tile list of round 1
  }
  std::vector<int> has((size_t)(nmax > 0 ? nmax : 1), 0), en((size_t)(nconf > 0 ? nconf : 1), 0);
  if (nmax > 0) {
    std::vector<int> mask((size_t)nat, 0);
    for (int k : sel) mask[k] = 1;
    CtxOwner cx;
    if (pack_operand(cx, nat, nmax, xyz, mask.data())) return 1;
    DevBuf buf;
    int *d_conf, *d_has, *d_en;
    if (buf.alloc(&d_conf, (size_t)nmax) || buf.alloc(&d_has, has.size()) || buf.alloc(&d_en, en.size())) return 1;
    CREST_CUDA_TRY(cudaMemcpy(d_conf, conf_of.data(), sizeof(int) * nmax, cudaMemcpyHostToDevice));
    CREST_CUDA_TRY(cudaMemset(d_has, 0, sizeof(int) * has.size()));
    CREST_CUDA_TRY(cudaMemset(d_en, 0, sizeof(int) * en.size()));
    PairSink k{};
    k.mode = 2; k.ni = nmax; k.nj = nmax; k.blk_i = d_conf; k.blk_j = d_conf;
    k.has_earlier = d_has; k.enantio = d_en; k.rthr = rthr; k.rthr_enantio = rthr * 0.75; k.A = rednat;
    // block diagonal: per conformer the j-tiles its rotamers touch and, per j-tile, the rows at or below it;
    // a (row block, j-tile) shared by neighbouring small conformers is listed once
    std::vector<int4> work;
    std::vector<long long> seen;
    const bool col = cx.c->use_col;
    const int rb = col ? kUnit : kPanel;               // rows per work unit
    for (int c = 0; c < nconf; ++c) {
      if (nrot[c] <= 1) continue;
      const int lo = low[c] - 1, hi = lo + nrot[c];
      for (int J = lo / 64; J <= (hi - 1) / 64; ++J) {
        const int r0 = std::max(lo, J * 64) / rb, r1 = (hi - 1) / rb + 1;
        if (col) {
          const long long key = ((long long)J << 40) | ((long long)r0 << 20) | r1;
          if (std::find(seen.begin(), seen.end(), key) == seen.end()) {
            seen.push_back(key);
            work.push_back(make_int4(J, r0, r1, 0));
          }
        } else {
          for (int P = r0; P < r1; ++P) {
            const long long key = ((long long)J << 32) | P;
            if (std::find(seen.begin(), seen.end(), key) == seen.end()) {
              seen.push_back(key);
              work.push_back(make_int4(P, J, 0, 0));
            }
          }
        }
        if (seen.size() > 64) seen.erase(seen.begin(), seen.begin() + 32);   // neighbours only can collide
      }
    }
    if (run_sink(cx.c, k, work)) return 1;
    CREST_CUDA_TRY(cudaMemcpy(has.data(), d_has, sizeof(int) * has.size(), cudaMemcpyDeviceToHost));
    CREST_CUDA_TRY(cudaMemcpy(en.data(), d_en, sizeof(int) * en.size(), cudaMemcpyDeviceToHost));
  }
  // symmap (entropic.f90:1096-1116): first slot with an equal trimmed symbol or the first empty
  // slot of a 100-entry list; a full list maps to slot 1
  std::vector<std::string> symref;
  std::vector<double> symenantio(100, 1.0);
  auto symmap = [&](int c) {
    std::string s(symsym + 3 * (size_t)c, 3);
    while (!s.empty() && s.back() == ' ') s.pop_back();
    for (size_t i = 0; i < symref.size(); ++i) if (symref[i] == s) return (int)i;
    if (s.empty()) return (int)symref.size() < 100 ? (int)symref.size() : 0;   // matches the empty slot, not stored
    if (symref.size() < 100) { symref.push_back(s); return (int)symref.size() - 1; }
    return 0;
  };
  for (int c = 0; c < nconf; ++c) {
    int groups = 1;
    if (nrot[c] > 1) {
      groups = 0;
      for (int s = low[c] - 1; s < low[c] - 1 + nrot[c]; ++s) groups += has[s] == 0;
      if (en[c]) enantiofac[c] = 2.0;
    }
    const int k = symmap(c);
    if (enantiofac[c] > symenantio[k]) symenantio[k] = enantiofac[c];
    fac[c] = (double)groups;
  }
  // the second pass re-runs symmap on the now populated list
  for (int c = 0; c < nconf; ++c) enantiofac[c] = symenantio[symmap(c)];
  return 0;
}

}  // extern "C"
