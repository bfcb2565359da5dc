// crossing.cu — the de-duplication inside CREST's genetic structure crossing.
//
//   /root/reference/src/algos/crossing.f90:227-338
// For every structure generated from a parent pair (GMETRY + CN clash check: not on the RMSD path, they
// stay with the caller) the reference computes  rval = rmsd(new, lowest conformer)  (263), discards it if
// rval <= rthr, and then, inside an `!$omp critical` (275-321), compares it with the ACCEPTED list: only
// entries whose own rval lies within rthr/100 are looked at (284), an RMSD < rthr to any of them rejects
// the structure (287-291); otherwise it is appended (299-302), or - list full - ends the search
// (huge_number, 303-304) or replaces the entry that maximises the spread of rms() (sdselect, 306-310,
// 401-432).  With one OpenMP thread the loop order (i, then j < i) is fixed and so is the result; that
// order is the contract here (with several threads the reference's own result depends on thread arrival).
//
// The accepted list is a sequential object, but what is expensive is not: whether two structures MATCH
// (|rval_a - rval_b| < rthr/100 and rmsd(a, b) < rthr) does not depend on the sweep.  So, like K6
// (sweep.cu), the work is split in two: the GPU evaluates rval for every candidate (one warp each) and
// the RMSD of every GATED candidate pair (candidates sorted by rval: the gate is a window in that order;
// 8 lanes per pair, one result bit per pair), and the host replays the reference's critical section on
// those bits - integer logic plus sdselect's arithmetic in the reference's order.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <numeric>
#include <vector>
#include "common.cuh"
#include "qcp.cuh"
#include "../../include/crest_cuda.h"

namespace crest {

// one warp per structure: centred copy (ls_rmsd.f90:105-119) + squared norm
__global__ void __launch_bounds__(128)
cx_center_kernel(int nat, int n, const double* __restrict__ xyz, double* __restrict__ xc,
                 double* __restrict__ xn) {
  const int lane = threadIdx.x & 31;
  const int s = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (s >= n) return;
  const double* x = xyz + (size_t)3 * nat * s;
  double* o = xc + (size_t)3 * nat * s;
  double c[3] = {0, 0, 0};
  for (int k = lane; k < nat; k += 32)
    for (int a = 0; a < 3; ++a) c[a] += x[3 * k + a];
  for (int a = 0; a < 3; ++a) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) c[a] += __shfl_xor_sync(0xffffffffu, c[a], d);
    c[a] /= (double)nat;
  }
  double sq = 0.0;
  for (int k = lane; k < nat; k += 32)
    for (int a = 0; a < 3; ++a) {
      const double v = x[3 * k + a] - c[a];
      o[3 * k + a] = v;
      sq += v * v;
    }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, d);
  if (lane == 0) xn[s] = sq;
}

// RMSD (proper rotation only, ls_rmsd.f90 option 0) of two centred structures by G lanes
template <int G>
__device__ __forceinline__ double rmsd_lanes(int nat, const double* __restrict__ x, const double* __restrict__ y,
                                             double s, int sub, unsigned gmask) {
  double r[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
  for (int k = sub; k < nat; k += G) {
    const double x0 = x[3 * k], x1 = x[3 * k + 1], x2 = x[3 * k + 2];
    const double y0 = y[3 * k], y1 = y[3 * k + 1], y2 = y[3 * k + 2];
    r[0] = fma(x0, y0, r[0]); r[1] = fma(x0, y1, r[1]); r[2] = fma(x0, y2, r[2]);
    r[3] = fma(x1, y0, r[3]); r[4] = fma(x1, y1, r[4]); r[5] = fma(x1, y2, r[5]);
    r[6] = fma(x2, y0, r[6]); r[7] = fma(x2, y1, r[7]); r[8] = fma(x2, y2, r[8]);
  }
#pragma unroll
  for (int z = 0; z < 9; ++z)
#pragma unroll
    for (int d = G / 2; d > 0; d >>= 1) r[z] += __shfl_xor_sync(gmask, r[z], d);
  const double e = pair_e(r, s, 0, 0);
  return sqrt(fmax(0.0, e) / (double)nat);
}

// rval(c) = rmsd(candidate c, reference) - crossing.f90:263; the reference structure is entry `ref`
__global__ void __launch_bounds__(128)
cx_ref_kernel(int nat, int n, int ref, const double* __restrict__ xc, const double* __restrict__ xn,
              double* __restrict__ rval) {
  const int lane = threadIdx.x & 31;
  const int c = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (c >= n) return;
  const double v = rmsd_lanes<32>(nat, xc + (size_t)3 * nat * c, xc + (size_t)3 * nat * ref, xn[c] + xn[ref],
                                  lane, 0xffffffffu);
  if (lane == 0) rval[c] = v;
}

// gated pairs: sorted position p against q in [lo[p], p); bit off[p] + (q - lo[p]) = rmsd < rthr
__global__ void __launch_bounds__(256)
cx_pair_kernel(int nat, int nlive, const int* __restrict__ order, const int* __restrict__ lo,
               const long long* __restrict__ off, const double* __restrict__ xc,
               const double* __restrict__ xn, double rthr, unsigned* __restrict__ bits) {
  const int lane = threadIdx.x & 31, sub = lane & 7, grp = lane >> 3;
  const unsigned gmask = 0xFFu << (lane & 24);
  const int warps = (gridDim.x * blockDim.x) >> 5;
  for (int p = nlive - 1 - ((blockIdx.x * blockDim.x + threadIdx.x) >> 5); p >= 0; p -= warps) {
    const int l = lo[p];
    const int a = order[p];
    const double* xa = xc + (size_t)3 * nat * a;
    for (int q0 = l; q0 < p; q0 += 4) {
      const int q = q0 + grp;
      const bool live = q < p;
      const int b = order[live ? q : l];
      const double v = rmsd_lanes<8>(nat, xa, xc + (size_t)3 * nat * b, xn[a] + xn[b], sub, gmask);
      if (live && sub == 0 && v < rthr) {
        const long long bit = off[p] + (q - l);
        atomicOr(bits + (bit >> 5), 1u << (bit & 31));
      }
    }
  }
}

// sdselect (crossing.f90:401-432), operation order of the reference: which entry to replace by rval so
// that the standard deviation of rms() grows most; 0 = none.  O(n^2) like the reference.
static int sdselect_host(int n, const double* rms, double rval) {
  int pos = 0;
  double sum = 0.0;
  for (int i = 0; i < n; ++i) sum = sum + rms[i];
  double avg = sum / (double)(float)n;
  double sdref = 0.0;
  for (int i = 0; i < n; ++i) sdref = sdref + (rms[i] - avg) * (rms[i] - avg);
  sdref = std::sqrt(sdref / (double)(float)n);
  for (int i = 0; i < n; ++i) {
    avg = sum - rms[i] + rval;
    avg = avg / (double)(float)n;
    double sdtmp = 0.0;
    for (int k = 0; k < n; ++k) sdtmp = sdtmp + (rms[k] - avg) * (rms[k] - avg);
    sdtmp = sdtmp - (rms[i] - avg) * (rms[i] - avg);
    sdtmp = sdtmp + (rval - avg) * (rval - avg);
    sdtmp = std::sqrt(sdtmp / (double)(float)n);
    if (sdtmp > sdref) { sdref = sdtmp; pos = i + 1; }
  }
  return pos;
}

struct DevBufC {
  void* p = nullptr;
  ~DevBufC() { cudaFree(p); }
  int get(size_t bytes) { return cudaMalloc(&p, bytes ? bytes : 1) == cudaSuccess ? 0 : 1; }
};

}  // namespace crest

using namespace crest;

extern "C" int crest_cuda_crossing_dedup(int nat, int ncand, const double* cand, const double* xyzref,
                                         double rthr, int maxgen2, int huge_number, int* taken,
                                         double* rms, int* ntaken, long long* counters, double* rmsdsum) {
  if (nat <= 0 || ncand < 0 || maxgen2 < 1 || !xyzref || (!cand && ncand) || !taken || !rms || !ntaken ||
      !counters || !rmsdsum) {
    set_error("bad argument");
    return 1;
  }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) {
    cudaGetLastError();
    set_error("no CUDA device available (there is no CPU fallback)");
    return 1;
  }
  for (int k = 0; k < maxgen2; ++k) { taken[k] = 0; rms[k] = 0.0; }
  *ntaken = 0; *rmsdsum = 0.0;
  counters[0] = counters[1] = counters[2] = counters[3] = 0;
  if (ncand == 0) return 0;
  const size_t rec = (size_t)3 * nat, n = (size_t)ncand + 1;
  DevBufC d_xyz, d_xc, d_xn, d_rval;
  if (d_xyz.get(sizeof(double) * rec * n) || d_xc.get(sizeof(double) * rec * n) || d_xn.get(sizeof(double) * n) ||
      d_rval.get(sizeof(double) * n)) {
    set_error("cudaMalloc failed: %s", cudaGetErrorString(cudaGetLastError()));
    return 1;
  }
  CREST_CUDA_TRY(cudaMemcpy(d_xyz.p, cand, sizeof(double) * rec * ncand, cudaMemcpyHostToDevice));
  CREST_CUDA_TRY(cudaMemcpy((double*)d_xyz.p + rec * ncand, xyzref, sizeof(double) * rec, cudaMemcpyHostToDevice));
  cx_center_kernel<<<(unsigned)((n + 3) / 4), 128>>>(nat, (int)n, (const double*)d_xyz.p, (double*)d_xc.p, (double*)d_xn.p);
  cx_ref_kernel<<<(unsigned)((ncand + 3) / 4), 128>>>(nat, ncand, ncand, (const double*)d_xc.p, (const double*)d_xn.p,
                                                     (double*)d_rval.p);
  CREST_CUDA_TRY(cudaGetLastError());
  std::vector<double> rval((size_t)ncand);
  CREST_CUDA_TRY(cudaMemcpy(rval.data(), d_rval.p, sizeof(double) * ncand, cudaMemcpyDeviceToHost));

  // gate: only structures whose rval differ by less than rthr/100 are ever compared (crossing.f90:284);
  // sorted by rval that is a window.  The window is taken a hair wider than the gate; the replay below
  // applies the gate itself.
  const double thr100 = rthr / 100.0;
  std::vector<int> order;
  order.reserve((size_t)ncand);
  for (int c = 0; c < ncand; ++c)
    if (!(rval[c] <= rthr)) order.push_back(c);
  std::sort(order.begin(), order.end(), [&](int a, int b) { return rval[a] < rval[b] || (rval[a] == rval[b] && a < b); });
  const int nlive = (int)order.size();
  std::vector<int> lo((size_t)nlive);
  std::vector<long long> off((size_t)nlive + 1, 0);
  {
    int l = 0;
    const double w = thr100 * (1.0 + 1e-9) + 1e-300;
    for (int p = 0; p < nlive; ++p) {
      while (rval[order[l]] < rval[order[p]] - w) ++l;
      lo[p] = l;
      off[p + 1] = off[p] + (p - l);
    }
  }
  const long long npair = off[nlive];
  if (npair > (1ll << 36)) { set_error("crossing: %lld gated pairs exceed the bit table", npair); return 1; }
  std::vector<unsigned> bits((size_t)((npair + 31) / 32) + 1, 0u);
  if (npair > 0) {
    DevBufC d_order, d_lo, d_off, d_bits;
    if (d_order.get(sizeof(int) * nlive) || d_lo.get(sizeof(int) * nlive) || d_off.get(sizeof(long long) * (nlive + 1)) ||
        d_bits.get(sizeof(unsigned) * bits.size())) {
      set_error("cudaMalloc failed: %s", cudaGetErrorString(cudaGetLastError()));
      return 1;
    }
    CREST_CUDA_TRY(cudaMemcpy(d_order.p, order.data(), sizeof(int) * nlive, cudaMemcpyHostToDevice));
    CREST_CUDA_TRY(cudaMemcpy(d_lo.p, lo.data(), sizeof(int) * nlive, cudaMemcpyHostToDevice));
    CREST_CUDA_TRY(cudaMemcpy(d_off.p, off.data(), sizeof(long long) * (nlive + 1), cudaMemcpyHostToDevice));
    CREST_CUDA_TRY(cudaMemset(d_bits.p, 0, sizeof(unsigned) * bits.size()));
    int nsm = 148, dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev);
    cx_pair_kernel<<<nsm * 8, 256>>>(nat, nlive, (const int*)d_order.p, (const int*)d_lo.p, (const long long*)d_off.p,
                                     (const double*)d_xc.p, (const double*)d_xn.p, rthr, (unsigned*)d_bits.p);
    CREST_CUDA_TRY(cudaGetLastError());
    CREST_CUDA_TRY(cudaMemcpy(bits.data(), d_bits.p, sizeof(unsigned) * bits.size(), cudaMemcpyDeviceToHost));
  }
  // matches of every candidate with EARLIER candidates (original order), CSR
  std::vector<long long> deg((size_t)ncand + 1, 0);
  for (int p = 0; p < nlive; ++p)
    for (int q = lo[p]; q < p; ++q) {
      const long long b = off[p] + (q - lo[p]);
      if ((bits[(size_t)(b >> 5)] >> (b & 31)) & 1u) ++deg[(size_t)std::max(order[p], order[q]) + 1];
    }
  for (int c = 0; c < ncand; ++c) deg[(size_t)c + 1] += deg[c];
  std::vector<int> adj((size_t)deg[ncand]);
  {
    std::vector<long long> fill(deg.begin(), deg.end() - 1);
    for (int p = 0; p < nlive; ++p)
      for (int q = lo[p]; q < p; ++q) {
        const long long b = off[p] + (q - lo[p]);
        if ((bits[(size_t)(b >> 5)] >> (b & 31)) & 1u) {
          const int hi = std::max(order[p], order[q]), lw = std::min(order[p], order[q]);
          adj[(size_t)fill[hi]++] = lw;
        }
      }
  }
  // replay of the critical section (crossing.f90:275-321) in candidate order
  std::vector<int> slot_of((size_t)ncand, -1);
  int nt = 0;
  long long ierr = 0, ident = 0, rcount = 0, processed = 0;
  double rsum = 0.0;
  for (int c = 0; c < ncand; ++c) {
    ++processed;
    const double rv = rval[c];
    if (rv <= rthr) { ++ierr; continue; }
    ++rcount;
    rsum = rsum + rv;
    bool fail = false;
    if (nt >= 1)
      for (long long e = deg[c]; e < deg[(size_t)c + 1] && !fail; ++e) {
        const int o = adj[(size_t)e];
        if (slot_of[o] >= 0 && std::fabs(rms[slot_of[o]] - rv) < thr100) fail = true;
      }
    if (fail) { ++ident; continue; }
    if (nt < maxgen2) {
      taken[nt] = c + 1; rms[nt] = rv; slot_of[c] = nt; ++nt;
    } else if (huge_number) {
      break;                                     // stop_crossing: nothing after this pair is looked at
    } else {
      const int k = sdselect_host(nt, rms, rv);
      if (k > 0) {
        slot_of[taken[k - 1] - 1] = -1;
        taken[k - 1] = c + 1; rms[k - 1] = rv; slot_of[c] = k - 1;
      }
    }
  }
  *ntaken = nt;
  *rmsdsum = rsum;
  counters[0] = ierr; counters[1] = ident; counters[2] = rcount; counters[3] = processed;
  return 0;
}
