// Kernel (4) and the small serial loops around the pair kernel.
//   backtrack      reference src/legacy_algos/cregen_old.f90:1396-1410, as used in cregen.f90:1375
//   enuc           reference src/cregen.f90:1422-1435   (K5)
//   rotamer groups reference src/cregen.f90:1437-1469   (K6 + group ids)
//   rmsd_pairs     reference src/ls_rmsd.f90:72-174      (diagnostic, explicit pair list)
#include <vector>
#include "common.cuh"
#include "qcp.cuh"
#include "../../include/crest_cuda.h"

namespace crest {

// dup_in[i] = 1-based parent (< i+1) or 0.  The reference compresses row by row, so after row i
// double(i) is the root of its chain; chasing the unmodified links gives the same root.
__global__ void backtrack_kernel(int n, const int* __restrict__ dup_in, int* __restrict__ dup_out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int p = dup_in[i];
  if (p > 0) {
    int q;
    while ((q = dup_in[p - 1]) > 0) p = q;
  }
  dup_out[i] = p;
}

int launch_backtrack(int nall, int* d_dup, cudaStream_t stream) {
  if (nall <= 0) return 0;
  int* tmp = nullptr;
  CREST_CUDA_TRY(cudaMallocAsync(&tmp, sizeof(int) * (size_t)nall, stream));
  backtrack_kernel<<<(nall + 255) / 256, 256, 0, stream>>>(nall, d_dup, tmp);
  CREST_CUDA_TRY(cudaGetLastError());
  CREST_CUDA_TRY(cudaMemcpyAsync(d_dup, tmp, sizeof(int) * (size_t)nall, cudaMemcpyDeviceToDevice,
                                 stream));
  CREST_CUDA_TRY(cudaFreeAsync(tmp, stream));
  return 0;
}

// ---- K5: enuc(k) = sum_{i<j} Z_i Z_j / (r_ij^2 + 1e-12), one warp per structure
__global__ void __launch_bounds__(128)
enuc_kernel(int nat, int n, const int* __restrict__ at, const double* __restrict__ xyz,
            double* __restrict__ enuc) {
  const int lane = threadIdx.x & 31;
  const int conf = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (conf >= n) return;
  const double* c = xyz + (size_t)3 * nat * conf;
  double e = 0.0;
  for (int i = 0; i < nat - 1; ++i) {
    const double xi = c[3 * i], yi = c[3 * i + 1], zi = c[3 * i + 2];
    const int zi_ = at[i];
    for (int j = i + 1 + lane; j < nat; j += 32) {
      const double dx = xi - c[3 * j], dy = yi - c[3 * j + 1], dz = zi - c[3 * j + 2];
      const double r = dx * dx + dy * dy + dz * dz + 1.0e-12;
      e += (double)(zi_ * at[j]) / r;
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) e += __shfl_xor_sync(0xffffffffu, e, o);
  if (lane == 0) enuc[conf] = e;
}

int launch_enuc(int nat, int n, const int* d_at, const double* d_xyz, double* d_enuc,
                cudaStream_t stream) {
  if (n <= 0) return 0;
  enuc_kernel<<<(n + 3) / 4, 128, 0, stream>>>(nat, n, d_at, d_xyz, d_enuc);
  CREST_CUDA_TRY(cudaGetLastError());
  return 0;
}

// ---- K6: energy-ordered greedy sweep.  Structure i joins the FIRST earlier group head j with
// |dE| < ethr, equal rot. constants (bthr) and relative enuc difference < 1e-3.  The dependency
// (whether j is a head) is sequential in i; the search over j is parallel in one persistent CTA.
// group[0] = number of groups, group[1..n] = group id.
__global__ void __launch_bounds__(1024)
group_kernel(int n, const double* __restrict__ er, const double* __restrict__ rot,
             const double* __restrict__ enuc, double ethr, double bthr, double bthrmax,
             double bthrshift, int* __restrict__ group, unsigned char* __restrict__ head,
             double* __restrict__ aniso) {
  __shared__ int s_min[32];
  __shared__ int s_ng;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) s_ng = 0;
  // rotaniso of every structure once (rotcompare.f90:109-122), not once per candidate pair
  for (int k = tid; k < n; k += blockDim.x) aniso[k] = rotaniso_dev(rot[3 * k], rot[3 * k + 1], rot[3 * k + 2]);
  __syncthreads();
  int jlo = 0;   // energies ascend: the |dE| window start only moves forward
  for (int i = 0; i < n; ++i) {
    const double ei = er[i];
    // all threads advance the window start identically (uniform)
    while (jlo < i && !(fabs(__dmul_rn(__dsub_rn(ei, er[jlo]), kAutokcal)) < ethr)) ++jlo;
    int best = 0x7fffffff;
    if (jlo < i) {
      const double ri[3] = {rot[3 * i], rot[3 * i + 1], rot[3 * i + 2]};
      const double ai = aniso[i];
      const double eni = enuc[i];
      for (int j = jlo + tid; j < i; j += blockDim.x) {
        if (!head[j]) continue;
        const double de = __dmul_rn(__dsub_rn(ei, er[j]), kAutokcal);
        if (!(fabs(de) < ethr)) continue;
        // the three conditions of cregen.f90:1444-1448 are a conjunction: cheapest first
        const double enj = enuc[j];
        const double rel = __ddiv_rn(__dmul_rn(2.0, fabs(__dsub_rn(eni, enj))), __dadd_rn(eni, enj));
        if (!(rel < 1.0e-3)) continue;
        const double rj[3] = {rot[3 * j], rot[3 * j + 1], rot[3 * j + 2]};
        if (!equalrotaniso_dev(ri, rj, ai, aniso[j], bthr, bthrmax, bthrshift)) continue;
        best = j;
        break;
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, o));
    if (lane == 0) s_min[warp] = best;
    __syncthreads();
    if (warp == 0) {
      int b = (lane < (int)(blockDim.x >> 5)) ? s_min[lane] : 0x7fffffff;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) b = min(b, __shfl_xor_sync(0xffffffffu, b, o));
      if (lane == 0) {
        if (b == 0x7fffffff) { head[i] = 1; group[i + 1] = ++s_ng; }
        else { head[i] = 0; group[i + 1] = group[b + 1]; }
      }
    }
    __syncthreads();
  }
  if (tid == 0) group[0] = s_ng;
}

// ---- K6, two phases.  The sequential part of the greedy sweep is only "is j still a head";
// whether (i, j) satisfy the three conditions of cregen.f90:1444-1448 does not depend on the sweep.
// Phase 1 (whole GPU, one warp per i): the first kGroupCap matching j < i in ascending order.
// Phase 2 (host, O(n * kGroupCap)): i joins the first listed j that is a head, else opens a group.
// A row whose listed matches are all non-heads AND whose list was cut needs the full row: then the
// single-CTA kernel above redoes the sweep (never seen on the test ensembles: kept structures that
// agree in energy window, rot. constants and enuc are rotamers of one another, a handful per group).
constexpr int kGroupCap = 8;   // default list length; CREST_CUDA_TUNE `groupcap=<n>` (tests: 1 forces the fallback)

__global__ void group_aniso_kernel(int n, const double* __restrict__ rot, double* __restrict__ aniso) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k < n) aniso[k] = rotaniso_dev(rot[3 * k], rot[3 * k + 1], rot[3 * k + 2]);
}

__global__ void __launch_bounds__(256)
group_match_kernel(int n, const double* __restrict__ er, const double* __restrict__ rot,
                   const double* __restrict__ enuc, const double* __restrict__ aniso, double ethr,
                   double bthr, double bthrmax, double bthrshift, int cap, int* __restrict__ cnt,
                   int* __restrict__ list) {
  const int lane = threadIdx.x & 31;
  // heavy rows (large i) first: the tail of the grid is then the cheap rows
  const int i = n - 1 - (blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5));
  if (i < 0) return;
  const double ei = er[i];
  // energies ascend, so |dE| < ethr holds on a suffix of 0..i-1 (same rounding as the test below)
  int lo = 0, hi = i;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (fabs(__dmul_rn(__dsub_rn(ei, er[mid]), kAutokcal)) < ethr) hi = mid; else lo = mid + 1;
  }
  const double ri[3] = {rot[3 * i], rot[3 * i + 1], rot[3 * i + 2]};
  const double ai = aniso[i];
  const double eni = enuc[i];
  int c = 0;
  for (int j0 = lo; j0 < i && c <= cap; j0 += 32) {
    const int j = j0 + lane;
    bool m = false;
    if (j < i) {
      const double de = __dmul_rn(__dsub_rn(ei, er[j]), kAutokcal);
      if (fabs(de) < ethr) {
        const double enj = enuc[j];
        const double rel = __ddiv_rn(__dmul_rn(2.0, fabs(__dsub_rn(eni, enj))), __dadd_rn(eni, enj));
        if (rel < 1.0e-3) {
          const double rj[3] = {rot[3 * j], rot[3 * j + 1], rot[3 * j + 2]};
          m = equalrotaniso_dev(ri, rj, ai, aniso[j], bthr, bthrmax, bthrshift);
        }
      }
    }
    unsigned bal = __ballot_sync(0xffffffffu, m);
    while (bal && c <= cap) {
      const int b = __ffs(bal) - 1;
      if (c < cap && lane == 0) list[(size_t)i * cap + c] = j0 + b;
      ++c;
      bal &= bal - 1;
    }
  }
  if (lane == 0) cnt[i] = c;     // cap + 1 = list cut
}

int launch_group(int n, const double* d_er, const double* d_rot, const double* d_enuc,
                 double ethr, double bthr, double bthrmax, double bthrshift, int* d_group,
                 cudaStream_t stream) {
  const size_t m = (size_t)(n > 0 ? n : 1);
  const int cap = tune().group_cap > 0 ? tune().group_cap : kGroupCap;
  unsigned char* head = nullptr;
  double* aniso = nullptr;
  int *d_cnt = nullptr, *d_list = nullptr;
  CREST_CUDA_TRY(cudaMallocAsync(&head, m, stream));
  CREST_CUDA_TRY(cudaMallocAsync(&aniso, sizeof(double) * m, stream));
  CREST_CUDA_TRY(cudaMallocAsync(&d_cnt, sizeof(int) * m, stream));
  CREST_CUDA_TRY(cudaMallocAsync(&d_list, sizeof(int) * m * cap, stream));
  // CREST_CUDA_TUNE knob `seqgroup`: force the single-CTA sweep (A/B test of the two paths)
  bool need_full = tune().seq_group != 0;
  if (n > 0) {
    group_aniso_kernel<<<(n + 255) / 256, 256, 0, stream>>>(n, d_rot, aniso);
    group_match_kernel<<<(n + 7) / 8, 256, 0, stream>>>(n, d_er, d_rot, d_enuc, aniso, ethr, bthr,
                                                        bthrmax, bthrshift, cap, d_cnt, d_list);
    CREST_CUDA_TRY(cudaGetLastError());
    std::vector<int> cnt(m), list(m * cap), group(m + 1);
    CREST_CUDA_TRY(cudaMemcpyAsync(cnt.data(), d_cnt, sizeof(int) * m, cudaMemcpyDeviceToHost, stream));
    CREST_CUDA_TRY(cudaMemcpyAsync(list.data(), d_list, sizeof(int) * m * cap,
                                   cudaMemcpyDeviceToHost, stream));
    CREST_CUDA_TRY(cudaStreamSynchronize(stream));
    std::vector<unsigned char> h((size_t)n);
    int ng = 0;
    for (int i = 0; i < n && !need_full; ++i) {
      int parent = -1;
      const int k = cnt[i] < cap ? cnt[i] : cap;
      for (int t = 0; t < k; ++t) {
        const int j = list[(size_t)i * cap + t];
        if (h[j]) { parent = j; break; }
      }
      if (parent < 0 && cnt[i] > cap) { need_full = true; break; }
      if (parent < 0) { h[i] = 1; group[i + 1] = ++ng; }
      else { h[i] = 0; group[i + 1] = group[parent + 1]; }
    }
    if (!need_full) {
      group[0] = ng;
      CREST_CUDA_TRY(cudaMemcpyAsync(d_group, group.data(), sizeof(int) * ((size_t)n + 1),
                                     cudaMemcpyHostToDevice, stream));
      CREST_CUDA_TRY(cudaStreamSynchronize(stream));   // `group` leaves scope
    }
  } else {
    need_full = true;   // n == 0: the kernel writes group[0] = 0
  }
  if (need_full) {
    group_kernel<<<1, 1024, 0, stream>>>(n, d_er, d_rot, d_enuc, ethr, bthr, bthrmax, bthrshift,
                                         d_group, head, aniso);
    CREST_CUDA_TRY(cudaGetLastError());
  }
  CREST_CUDA_TRY(cudaFreeAsync(head, stream));
  CREST_CUDA_TRY(cudaFreeAsync(aniso, stream));
  CREST_CUDA_TRY(cudaFreeAsync(d_cnt, stream));
  CREST_CUDA_TRY(cudaFreeAsync(d_list, stream));
  return 0;
}

// ---- diagnostic: explicit pair list, one warp per pair, proper and mirror RMSD
__global__ void __launch_bounds__(128)
rmsd_pairs_kernel(int A, const double* __restrict__ packed, long long npairs,
                  const int* __restrict__ ii, const int* __restrict__ jj,
                  double* __restrict__ out, double* __restrict__ out_m) {
  const int lane = threadIdx.x & 31;
  const long long p = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (p >= npairs) return;
  const double* x = packed + (size_t)3 * A * ii[p];
  const double* y = packed + (size_t)3 * A * jj[p];
  double cx[3] = {0, 0, 0}, cy[3] = {0, 0, 0};
  for (int k = lane; k < A; k += 32)
    for (int a = 0; a < 3; ++a) { cx[a] += x[3 * k + a]; cy[a] += y[3 * k + a]; }
  for (int a = 0; a < 3; ++a) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      cx[a] += __shfl_xor_sync(0xffffffffu, cx[a], o);
      cy[a] += __shfl_xor_sync(0xffffffffu, cy[a], o);
    }
    cx[a] /= (double)A; cy[a] /= (double)A;
  }
  double r[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}, s = 0.0;
  for (int k = lane; k < A; k += 32) {
    double xv[3], yv[3];
    for (int a = 0; a < 3; ++a) { xv[a] = x[3 * k + a] - cx[a]; yv[a] = y[3 * k + a] - cy[a]; }
    for (int a = 0; a < 3; ++a) {
      s += xv[a] * xv[a] + yv[a] * yv[a];
      for (int b = 0; b < 3; ++b) r[3 * a + b] += xv[a] * yv[b];
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    s += __shfl_xor_sync(0xffffffffu, s, o);
    for (int z = 0; z < 9; ++z) r[z] += __shfl_xor_sync(0xffffffffu, r[z], o);
  }
  const double e0 = pair_e(r, s, 0, 0);
  const double e1 = pair_e(r, s, 0, 1);
  if (lane == 0) {
    out[p] = sqrt(fmax(0.0, e0) / (double)A);
    out_m[p] = sqrt(fmax(0.0, e1) / (double)A);
  }
}

int launch_rmsd_pairs(int A, int n, const double* d_packed, long long npairs, const int* d_i,
                      const int* d_j, double* d_rmsd, double* d_rmsd_m, cudaStream_t stream) {
  (void)n;
  if (npairs <= 0) return 0;
  const long long blocks = (npairs + 3) / 4;
  rmsd_pairs_kernel<<<(unsigned)blocks, 128, 0, stream>>>(A, d_packed, npairs, d_i, d_j, d_rmsd,
                                                          d_rmsd_m);
  CREST_CUDA_TRY(cudaGetLastError());
  return 0;
}

// ---- cregen_rmsdalign: rotate every structure onto the first one (heavy atoms at > 2) ---------
// reference src/cregen.f90:2229-2283: rmsd(n, c1 = structure k, c0 = structure 1, option 1) gives the
// rotation U (src/ls_rmsd.f90:122-160: R(a,b) = sum x_a y_b on centroid-free coordinates, key matrix
// S, eigenvector q of lambda_max, rotation_matrix 176-221), then xyz(:,:,k) <- U * xyz(:,:,k) for ALL
// atoms, without any translation.  One warp per structure; q from a cyclic Jacobi with accumulated
// rotations (U is quadratic in q, so the sign of q does not matter).
static __device__ void jacobi4_vec(double S[4][4], double V[4][4]) {
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j) V[i][j] = (i == j) ? 1.0 : 0.0;
  for (int sweep = 0; sweep < 60; ++sweep) {
    double off = 0.0, diag = 0.0;
    for (int i = 0; i < 4; ++i) {
      diag += S[i][i] * S[i][i];
      for (int j = i + 1; j < 4; ++j) off += S[i][j] * S[i][j];
    }
    if (off <= 1e-36 * diag || off == 0.0) break;
    for (int p = 0; p < 3; ++p)
      for (int q = p + 1; q < 4; ++q) {
        const double apq = S[p][q];
        if (apq == 0.0) continue;
        const double theta = (S[q][q] - S[p][p]) / (2.0 * apq);
        const double t = (theta >= 0.0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
        const double c = 1.0 / sqrt(t * t + 1.0), sn = t * c;
        for (int k = 0; k < 4; ++k) {
          const double akp = S[k][p], akq = S[k][q];
          S[k][p] = c * akp - sn * akq;
          S[k][q] = sn * akp + c * akq;
        }
        for (int k = 0; k < 4; ++k) {
          const double apk = S[p][k], aqk = S[q][k];
          S[p][k] = c * apk - sn * aqk;
          S[q][k] = sn * apk + c * aqk;
        }
        for (int k = 0; k < 4; ++k) {
          const double vkp = V[k][p], vkq = V[k][q];
          V[k][p] = c * vkp - sn * vkq;
          V[k][q] = sn * vkp + c * vkq;
        }
      }
  }
}

__global__ void __launch_bounds__(128)
rmsdalign_kernel(int nat, int nall, const int* __restrict__ at, double* __restrict__ xyz) {
  const int lane = threadIdx.x & 31;
  const int k = 1 + blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);   // structure 0 is the target
  if (k >= nall) return;
  const double* ref = xyz;
  double* cur = xyz + (size_t)3 * nat * k;
  // centroids of the heavy atoms (ls_rmsd.f90:105-119)
  double cx[3] = {0, 0, 0}, cy[3] = {0, 0, 0};
  int n = 0;
  for (int j = lane; j < nat; j += 32)
    if (at[j] > 2) {
      ++n;
      for (int a = 0; a < 3; ++a) { cx[a] += cur[3 * j + a]; cy[a] += ref[3 * j + a]; }
    }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    n += __shfl_xor_sync(0xffffffffu, n, o);
    for (int a = 0; a < 3; ++a) {
      cx[a] += __shfl_xor_sync(0xffffffffu, cx[a], o);
      cy[a] += __shfl_xor_sync(0xffffffffu, cy[a], o);
    }
  }
  if (n == 0) return;                       // no heavy atom: nothing to align on
  for (int a = 0; a < 3; ++a) { cx[a] /= (double)n; cy[a] /= (double)n; }
  double r[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
  for (int j = lane; j < nat; j += 32)
    if (at[j] > 2) {
      double xv[3], yv[3];
      for (int a = 0; a < 3; ++a) { xv[a] = cur[3 * j + a] - cx[a]; yv[a] = ref[3 * j + a] - cy[a]; }
      for (int a = 0; a < 3; ++a)
        for (int b = 0; b < 3; ++b) r[3 * a + b] = fma(xv[a], yv[b], r[3 * a + b]);
    }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1)
    for (int z = 0; z < 9; ++z) r[z] += __shfl_xor_sync(0xffffffffu, r[z], o);
  // key matrix (ls_rmsd.f90:129-147), R(a,b) = r[3a+b]
  double S[4][4], V[4][4];
  S[0][0] = r[0] + r[4] + r[8];
  S[1][0] = r[5] - r[7];
  S[2][0] = r[6] - r[2];
  S[3][0] = r[1] - r[3];
  S[1][1] = r[0] - r[4] - r[8];
  S[2][1] = r[1] + r[3];
  S[3][1] = r[2] + r[6];
  S[2][2] = -r[0] + r[4] - r[8];
  S[3][2] = r[5] + r[7];
  S[3][3] = -r[0] - r[4] + r[8];
  for (int i = 0; i < 4; ++i)
    for (int j = i + 1; j < 4; ++j) S[i][j] = S[j][i];
  jacobi4_vec(S, V);
  int m = 0;
  for (int i = 1; i < 4; ++i)
    if (S[i][i] > S[m][m]) m = i;
  const double q0 = V[0][m], q1 = V[1][m], q2 = V[2][m], q3 = V[3][m];
  // rotation_matrix (ls_rmsd.f90:176-221)
  const double b0 = 2.0 * q0, b1 = 2.0 * q1, b2 = 2.0 * q2, b3 = 2.0 * q3;
  const double q00 = b0 * q0 - 1.0, q01 = b0 * q1, q02 = b0 * q2, q03 = b0 * q3;
  const double q11 = b1 * q1, q12 = b1 * q2, q13 = b1 * q3, q22 = b2 * q2, q23 = b2 * q3, q33 = b3 * q3;
  const double U[3][3] = {{q00 + q11, q12 - q03, q13 + q02},
                          {q12 + q03, q00 + q22, q23 - q01},
                          {q13 - q02, q23 + q01, q00 + q33}};
  for (int j = lane; j < nat; j += 32) {
    const double x = cur[3 * j], y = cur[3 * j + 1], z = cur[3 * j + 2];
    cur[3 * j] = U[0][0] * x + U[0][1] * y + U[0][2] * z;
    cur[3 * j + 1] = U[1][0] * x + U[1][1] * y + U[1][2] * z;
    cur[3 * j + 2] = U[2][0] * x + U[2][1] * y + U[2][2] * z;
  }
}

int launch_rmsdalign(int nat, int nall, const int* d_at, double* d_xyz, cudaStream_t stream) {
  if (nall <= 1) return 0;
  const int wpb = 4;
  rmsdalign_kernel<<<(nall - 1 + wpb - 1) / wpb, wpb * 32, 0, stream>>>(nat, nall, d_at, d_xyz);
  CREST_CUDA_TRY(cudaGetLastError());
  return 0;
}

// ---- kept-first partition (cregen.f90:1397-1419) applied on the device: dst position d takes input
// structure src[d]; geometry, energy and rot. constants move together.  One warp per structure.
__global__ void __launch_bounds__(128)
permute_kernel(int nat, int nall, const int* __restrict__ src, const double* __restrict__ xyz_in,
               const double* __restrict__ er_in, const double* __restrict__ rot_in,
               double* __restrict__ xyz_out, double* __restrict__ er_out, double* __restrict__ rot_out) {
  const int lane = threadIdx.x & 31;
  const int d = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (d >= nall) return;
  const int sidx = src[d];
  const double* a = xyz_in + (size_t)3 * nat * sidx;
  double* b = xyz_out + (size_t)3 * nat * d;
  for (int k = lane; k < 3 * nat; k += 32) b[k] = a[k];
  if (lane == 0) er_out[d] = er_in[sidx];
  if (lane < 3) rot_out[3 * d + lane] = rot_in[3 * sidx + lane];
}

int launch_permute(int nat, int nall, const int* d_src, const double* d_xyz_in, const double* d_er_in,
                   const double* d_rot_in, double* d_xyz_out, double* d_er_out, double* d_rot_out,
                   cudaStream_t stream) {
  if (nall <= 0) return 0;
  permute_kernel<<<(nall + 3) / 4, 128, 0, stream>>>(nat, nall, d_src, d_xyz_in, d_er_in, d_rot_in,
                                                     d_xyz_out, d_er_out, d_rot_out);
  CREST_CUDA_TRY(cudaGetLastError());
  return 0;
}

}  // namespace crest

// ---- K6 without any allocation (crest_cuda_cregen_group_noalloc, include/crest_cuda.h) --------------------
// For callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).
// No counterpart in the reference.  Inputs and the output are
// device arrays of the caller; the scratch of group_kernel (head flag + rotaniso per structure) lives in ONE
// file-scope device array that every GPU loading the library holds until exit.  Every call shares it and it
// keeps the last call's contents: calls that overlap on the GPU race.
namespace crest {

constexpr int kNoAllocMaxKept = 32768;                                        // 8 + 1 bytes each = 288 KB
static __device__ __align__(256) unsigned char g_group_scratch[(size_t)kNoAllocMaxKept * 9];

namespace {
// In host code `g_group_scratch` is a host-side stand-in, not a device address: the address the kernel may use
// comes from cudaGetSymbolAddress, once per device, with that device current.
unsigned char* group_scratch_devptr() {
  static unsigned char* ptr[64] = {};
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) {
    set_error("cudaGetDevice failed or device index %d not in 0..63", dev);
    return nullptr;
  }
  if (!ptr[dev]) {
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_group_scratch);
    if (e != cudaSuccess || !p) {
      set_error("cudaGetSymbolAddress(g_group_scratch) failed on device %d: %s", dev, cudaGetErrorString(e));
      return nullptr;
    }
    ptr[dev] = static_cast<unsigned char*>(p);
  }
  return ptr[dev];
}
}  // namespace

}  // namespace crest

extern "C" {

int crest_cuda_noalloc_max_kept(void) { return crest::kNoAllocMaxKept; }

int crest_cuda_cregen_group_noalloc(int nkept, const double* d_er, const double* d_rot, const double* d_enuc,
                                    const crest_cuda_thresholds* thr, int* d_group, void* stream) {
  using namespace crest;
  if (nkept < 0 || !thr || !d_group || (nkept && (!d_er || !d_rot || !d_enuc))) {
    set_error("bad argument");
    return 1;
  }
  if (nkept > kNoAllocMaxKept) {
    set_error("crest_cuda_cregen_group_noalloc: nkept = %d exceeds the static scratch (%d structures); "
              "use crest_cuda_cregen_group", nkept, kNoAllocMaxKept);
    return 1;
  }
  unsigned char* s = group_scratch_devptr();
  if (!s) return 1;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  double* aniso = reinterpret_cast<double*>(s);                       // [0, 8 * max): rotaniso per structure
  unsigned char* head = s + (size_t)kNoAllocMaxKept * 8;              // [8 * max, 9 * max): "is a group head"
  // the kernel writes aniso[k] and head[i] before it reads them; the clear only keeps a stale flag of the
  // previous call from ever being looked at
  CREST_CUDA_TRY(cudaMemsetAsync(head, 0, (size_t)(nkept > 0 ? nkept : 1), st));
  group_kernel<<<1, 1024, 0, st>>>(nkept, d_er, d_rot, d_enuc, thr->ethr, thr->bthr, thr->bthrmax,
                                   thr->bthrshift, d_group, head, aniso);
  CREST_CUDA_TRY(cudaGetLastError());
  return 0;
}

}  // extern "C"
