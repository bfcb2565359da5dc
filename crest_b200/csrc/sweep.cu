// Kernel (4) and the small serial loops around the pair kernel.
//   backtrack      reference src/legacy_algos/cregen_old.f90:1396-1410, as used in cregen.f90:1375
//   enuc           reference src/cregen.f90:1422-1435   (K5)
//   rotamer groups reference src/cregen.f90:1437-1469   (K6 + group ids)
//   rmsd_pairs     reference src/ls_rmsd.f90:72-174      (diagnostic, explicit pair list)
#include "common.cuh"
#include "qcp.cuh"

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
             double bthrshift, int* __restrict__ group, unsigned char* __restrict__ head) {
  __shared__ int s_min[32];
  __shared__ int s_ng;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) s_ng = 0;
  __syncthreads();
  int jlo = 0;   // energies ascend: the |dE| window start only moves forward
  for (int i = 0; i < n; ++i) {
    const double ei = er[i];
    // all threads advance the window start identically (uniform)
    while (jlo < i && !(fabs(__dmul_rn(__dsub_rn(ei, er[jlo]), kAutokcal)) < ethr)) ++jlo;
    int best = 0x7fffffff;
    if (jlo < i) {
      const double ri[3] = {rot[3 * i], rot[3 * i + 1], rot[3 * i + 2]};
      const double ai = rotaniso_dev(ri[0], ri[1], ri[2]);
      const double eni = enuc[i];
      for (int j = jlo + tid; j < i; j += blockDim.x) {
        if (!head[j]) continue;
        const double de = __dmul_rn(__dsub_rn(ei, er[j]), kAutokcal);
        if (!(fabs(de) < ethr)) continue;
        const double rj[3] = {rot[3 * j], rot[3 * j + 1], rot[3 * j + 2]};
        const double aj = rotaniso_dev(rj[0], rj[1], rj[2]);
        if (!equalrotaniso_dev(ri, rj, ai, aj, bthr, bthrmax, bthrshift)) continue;
        const double enj = enuc[j];
        const double rel = __ddiv_rn(__dmul_rn(2.0, fabs(__dsub_rn(eni, enj))), __dadd_rn(eni, enj));
        if (rel < 1.0e-3) { best = j; break; }
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

int launch_group(int n, const double* d_er, const double* d_rot, const double* d_enuc,
                 double ethr, double bthr, double bthrmax, double bthrshift, int* d_group,
                 cudaStream_t stream) {
  unsigned char* head = nullptr;
  CREST_CUDA_TRY(cudaMallocAsync(&head, (size_t)(n > 0 ? n : 1), stream));
  group_kernel<<<1, 1024, 0, stream>>>(n, d_er, d_rot, d_enuc, ethr, bthr, bthrmax, bthrshift,
                                       d_group, head);
  CREST_CUDA_TRY(cudaGetLastError());
  CREST_CUDA_TRY(cudaFreeAsync(head, stream));
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

}  // namespace crest
