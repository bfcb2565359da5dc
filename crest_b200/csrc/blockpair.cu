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
// Both need the RMSD itself as a double (no float32 rounding, no "far" screen), so the covariance
// is accumulated in FP64 on the CUDA cores from a (3A x N) coordinate matrix (row = component*A +
// atom, conformer index fastest: a 32-conformer tile row is one 256 B run) and the quartic of
// qcp.cuh is solved per pair.  Sizes are small next to the CREGEN triangle (maxcompare x rotamers).
//
// uniqueCore (entropic.f90:1074-1095) is order dependent as written, but its count is not: i opens
// a new group iff mp(i) == 0 when the loop reaches it, and mp(i) is set by exactly those earlier
// i' < i with mat(i',i) < thr (every i' writes ALL its neighbours, whether or not it opened a
// group).  So  uniqueCore = #{ i : no j < i with rmat(i,j) < thr }  — a per-row OR like K4.
#include <vector>
#include <cstring>
#include <string>
#include "common.cuh"
#include "qcp.cuh"
#include "../../include/crest_cuda.h"

namespace crest {

constexpr int kBT = 32;   // conformers per tile edge
constexpr int kBK = 16;   // atoms per shared-memory chunk

// One warp per structure: gather the selected atoms, remove their unweighted centroid
// (ls_rmsd.f90:105-119), write the (3A x npad) matrix and the centred squared norm.
__global__ void __launch_bounds__(128)
bp_center_kernel(int nat, int n, int A, const int* __restrict__ sel, const double* __restrict__ xyz,
                 int npad, double* __restrict__ Xc, double* __restrict__ xn) {
  const int lane = threadIdx.x & 31;
  const int s = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (s >= n) return;
  const double* x = xyz + (size_t)3 * nat * s;
  double c[3] = {0, 0, 0};
  for (int k = lane; k < A; k += 32) {
    const int at = sel ? sel[k] : k;
    for (int a = 0; a < 3; ++a) c[a] += x[3 * at + a];
  }
  for (int a = 0; a < 3; ++a) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) c[a] += __shfl_xor_sync(0xffffffffu, c[a], o);
    c[a] /= (double)A;
  }
  double sq = 0.0;
  for (int k = lane; k < A; k += 32) {
    const int at = sel ? sel[k] : k;
    for (int a = 0; a < 3; ++a) {
      const double v = x[3 * at + a] - c[a];
      sq += v * v;
      Xc[((size_t)a * A + k) * npad + s] = v;
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, o);
  if (lane == 0) xn[s] = sq;
}

struct BpArgs {
  const double* X1; const double* xn1; int n1, npad1;   // i side
  const double* X2; const double* xn2; int n2, npad2;   // j side
  int A;
  // MODE 0 (compare): dense rectangle, block ids per structure, rmat(nconf1,nconf2) as ordered bits
  const int* blk1; const int* blk2; int nconf1;
  unsigned long long* rmat_bits;
  // MODE 1 (intra): tile list {i0, j0, lim, co}; pairs j < i < lim
  const int4* tiles;
  int* has_earlier;    // [n1] 1 = some j < i of the same conformer with rmsd < rthr
  int* enantio;        // [nconf] 1 = enantiomer probe fired
  double rthr, rthr_enantio;
};

template <int MODE>
__global__ void __launch_bounds__(256, 2)   // <= 128 registers: two CTAs per SM (ncu: 130 registers gave one)
bp_pair_kernel(const BpArgs p, int ntj) {
  __shared__ __align__(16) double xs[kBK][3][kBT];
  __shared__ __align__(16) double ys[kBK][3][kBT];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  int i0, j0, lim = 0, co = 0;
  if (MODE == 0) {
    i0 = (blockIdx.x / ntj) * kBT;
    j0 = (blockIdx.x % ntj) * kBT;
  } else {
    const int4 t = p.tiles[blockIdx.x];
    i0 = t.x; j0 = t.y; lim = t.z; co = t.w;
  }
  double r[2][2][9];
#pragma unroll
  for (int u = 0; u < 2; ++u)
#pragma unroll
    for (int v = 0; v < 2; ++v)
#pragma unroll
      for (int z = 0; z < 9; ++z) r[u][v][z] = 0.0;

  for (int k0 = 0; k0 < p.A; k0 += kBK) {
    const int kc = min(kBK, p.A - k0);
    for (int e = tid; e < kBK * 3 * kBT; e += 256) {
      const int cf = e & 31, a = (e >> 5) % 3, kk = e / 96;
      double xv = 0.0, yv = 0.0;
      if (kk < kc) {
        // the matrices are zero padded to a multiple of 32 conformers (+ one tile for ragged tile
        // origins of the block-diagonal mode), so no conformer bound is needed
        xv = p.X1[((size_t)a * p.A + k0 + kk) * p.npad1 + i0 + cf];
        yv = p.X2[((size_t)a * p.A + k0 + kk) * p.npad2 + j0 + cf];
      }
      xs[kk][a][cf] = xv;
      ys[kk][a][cf] = yv;
    }
    __syncthreads();
#pragma unroll 4
    for (int kk = 0; kk < kBK; ++kk) {
      double2 xa[3], ya[3];
#pragma unroll
      for (int a = 0; a < 3; ++a) {
        xa[a] = *reinterpret_cast<const double2*>(&xs[kk][a][2 * ty]);
        ya[a] = *reinterpret_cast<const double2*>(&ys[kk][a][2 * tx]);
      }
#pragma unroll
      for (int a = 0; a < 3; ++a)
#pragma unroll
        for (int b = 0; b < 3; ++b) {
          r[0][0][3 * a + b] = fma(xa[a].x, ya[b].x, r[0][0][3 * a + b]);
          r[0][1][3 * a + b] = fma(xa[a].x, ya[b].y, r[0][1][3 * a + b]);
          r[1][0][3 * a + b] = fma(xa[a].y, ya[b].x, r[1][0][3 * a + b]);
          r[1][1][3 * a + b] = fma(xa[a].y, ya[b].y, r[1][1][3 * a + b]);
        }
    }
    __syncthreads();
  }

  const double invA = 1.0 / (double)p.A;
#pragma unroll
  for (int u = 0; u < 2; ++u)
#pragma unroll
    for (int v = 0; v < 2; ++v) {
      const int i = i0 + 2 * ty + u, j = j0 + 2 * tx + v;
      if (MODE == 0) {
        if (i >= p.n1 || j >= p.n2) continue;
        const int ci = p.blk1[i], cj = p.blk2[j];
        if (ci < 0 || cj < 0) continue;
        const double s = p.xn1[i] + p.xn2[j];
        const double e = pair_e(r[u][v], s, 0, 0);
        const double val = sqrt(fmax(0.0, e) * invA);      // ls_rmsd.f90:163
        // non-negative doubles order like their bit patterns; the plain read only filters
        // (the slot decreases monotonically), the atomic decides
        unsigned long long* slot = p.rmat_bits + ci + (size_t)p.nconf1 * cj;
        const unsigned long long bits = (unsigned long long)__double_as_longlong(val);
        if (bits < *reinterpret_cast<volatile unsigned long long*>(slot)) atomicMin(slot, bits);
      } else {
        if (!(j < i && i < lim)) continue;
        const double s = p.xn1[i] + p.xn2[j];
        const double e = pair_e(r[u][v], s, 0, 0);
        const double val = sqrt(fmax(0.0, e) * invA);
        if (val < p.rthr) p.has_earlier[i] = 1;            // entropic.f90:1088 via uniqueCore
        if (val > p.rthr) {                                // entropic.f90:1024-1031
          const double em = pair_e(r[u][v], s, 0, 1);
          const double vm = sqrt(fmax(0.0, em) * invA);
          if (vm < p.rthr_enantio) p.enantio[co] = 1;
        }
      }
    }
}

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

// upload xyz(3,nat,n), centre on the selected atoms, leave Xc (3A x npad, zero padded) and xn
int center_ensemble(DevBuf& buf, int nat, int n, int A, const int* d_sel, const double* xyz,
                    int* npad_out, double** Xc_out, double** xn_out) {
  const int npad = (n + kBT - 1) / kBT * kBT + kBT;
  double *d_xyz, *d_X, *d_xn;
  if (buf.alloc(&d_xyz, (size_t)3 * nat * n) || buf.alloc(&d_X, (size_t)3 * A * npad) ||
      buf.alloc(&d_xn, (size_t)npad)) return 1;
  CREST_CUDA_TRY(cudaMemcpy(d_xyz, xyz, sizeof(double) * 3 * nat * (size_t)n, cudaMemcpyHostToDevice));
  CREST_CUDA_TRY(cudaMemset(d_X, 0, sizeof(double) * 3 * A * (size_t)npad));
  CREST_CUDA_TRY(cudaMemset(d_xn, 0, sizeof(double) * npad));
  bp_center_kernel<<<(n + 3) / 4, 128>>>(nat, n, A, d_sel, d_xyz, npad, d_X, d_xn);
  CREST_CUDA_TRY(cudaGetLastError());
  *npad_out = npad; *Xc_out = d_X; *xn_out = d_xn;
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
  DevBuf buf;
  int npad1, npad2; double *X1, *X2, *xn1, *xn2;
  if (center_ensemble(buf, nat, n1, nat, nullptr, xyz1, &npad1, &X1, &xn1)) return 1;
  if (center_ensemble(buf, nat, n2, nat, nullptr, xyz2, &npad2, &X2, &xn2)) return 1;
  int *d_b1, *d_b2; unsigned long long* d_r;
  if (buf.alloc(&d_b1, (size_t)n1) || buf.alloc(&d_b2, (size_t)n2) || buf.alloc(&d_r, nout)) return 1;
  CREST_CUDA_TRY(cudaMemcpy(d_b1, blk1.data(), sizeof(int) * n1, cudaMemcpyHostToDevice));
  CREST_CUDA_TRY(cudaMemcpy(d_b2, blk2.data(), sizeof(int) * n2, cudaMemcpyHostToDevice));
  // +inf seed: the reference's 100 (ensemblecomp.f90:228) is overwritten by every rotamer pair
  std::vector<double> seed(nout, __builtin_inf());
  CREST_CUDA_TRY(cudaMemcpy(d_r, seed.data(), sizeof(double) * nout, cudaMemcpyHostToDevice));
  BpArgs p{};
  p.X1 = X1; p.xn1 = xn1; p.n1 = n1; p.npad1 = npad1;
  p.X2 = X2; p.xn2 = xn2; p.n2 = n2; p.npad2 = npad2;
  p.A = nat; p.blk1 = d_b1; p.blk2 = d_b2; p.nconf1 = nconf1; p.rmat_bits = d_r;
  const long long nti = (n1 + kBT - 1) / kBT, ntj = (n2 + kBT - 1) / kBT;
  if (nti * ntj > 0x7fffffffLL) { set_error("compare: too many tiles"); return 1; }
  bp_pair_kernel<0><<<(unsigned)(nti * ntj), 256>>>(p, (int)ntj);
  CREST_CUDA_TRY(cudaGetLastError());
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
  std::vector<int4> tiles;
  int nmax = 0;
  for (int c = 0; c < nconf; ++c) {
    if (nrot[c] <= 1) continue;
    const int lo = low[c] - 1, hi = lo + nrot[c];
    if (lo < 0 || hi > nall) { set_error("conformer %d: rotamers [%d,%d] outside 1..%d", c + 1, low[c], hi, nall); return 1; }
    nmax = hi > nmax ? hi : nmax;
    for (int i0 = lo; i0 < hi; i0 += kBT)
      for (int j0 = lo; j0 <= i0; j0 += kBT) tiles.push_back(make_int4(i0, j0, hi, c));
  }
  std::vector<int> has((size_t)(nmax > 0 ? nmax : 1), 0), en((size_t)(nconf > 0 ? nconf : 1), 0);
  if (!tiles.empty()) {
    DevBuf buf;
    int npad; double *X, *xn; int *d_sel, *d_has, *d_en; int4* d_tiles;
    if (buf.alloc(&d_sel, sel.size())) return 1;
    CREST_CUDA_TRY(cudaMemcpy(d_sel, sel.data(), sizeof(int) * sel.size(), cudaMemcpyHostToDevice));
    if (center_ensemble(buf, nat, nmax, rednat, d_sel, xyz, &npad, &X, &xn)) return 1;
    if (buf.alloc(&d_has, has.size()) || buf.alloc(&d_en, en.size()) || buf.alloc(&d_tiles, tiles.size())) return 1;
    CREST_CUDA_TRY(cudaMemset(d_has, 0, sizeof(int) * has.size()));
    CREST_CUDA_TRY(cudaMemset(d_en, 0, sizeof(int) * en.size()));
    CREST_CUDA_TRY(cudaMemcpy(d_tiles, tiles.data(), sizeof(int4) * tiles.size(), cudaMemcpyHostToDevice));
    BpArgs p{};
    p.X1 = X; p.xn1 = xn; p.n1 = nmax; p.npad1 = npad;
    p.X2 = X; p.xn2 = xn; p.n2 = nmax; p.npad2 = npad;
    p.A = rednat; p.tiles = d_tiles; p.has_earlier = d_has; p.enantio = d_en;
    p.rthr = rthr; p.rthr_enantio = rthr * 0.75;
    bp_pair_kernel<1><<<(unsigned)tiles.size(), 256>>>(p, 0);
    CREST_CUDA_TRY(cudaGetLastError());
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
