// Kernel (1): centring, principal-axis frame, rotational constants, RMSD-subset packing.
//
// Replaces the serial per-structure loop of the reference
//   src/cregen.f90:1223-1231    call axis(nat,at,xyz(:,:,i)); call axis(nat0,at0,c1,rot(1:3,i),bdum)
// i.e. axis_4 -> axis_3 (src/axis_module.f90:249-316): CMA shift (384-412), inertia tensor with
// the t(k)=k*1e-10 seed (130-133), analytic eigenvectors (1088-1266), chirality fix (273-278),
// rotation coordout = coord*evec; then axis_1 -> axis_0 (96-176) on the rotated (sub)structure for
// rot(1:3) in MHz.  It also hoists what the reference redoes per pair: the heavy/subset masking
// (cregen.f90:1293-1299, 2056-2101) and rmsd()'s unweighted centroid removal + norms
// (src/ls_rmsd.f90:105-119), writing the packed operand X of the covariance contraction.
//
// One warp per 32 conformers; in the memory passes groups of 8 lanes stride over the atoms of one
// conformer each and group shuffles reduce; the 3x3 eigenproblems run one conformer per lane.
// This file is compiled with -fmad=false: the eigen-solver follows the reference's operation
// order so rot. constants agree to ~1e-12 relative away from degenerate moments.
#include "common.cuh"
#include "atmasses.cuh"
#include "qcp.cuh"

namespace crest {

__constant__ double c_ams[CREST_NUM_ELEMENTS] = CREST_AMS_TABLE_INIT;

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// sum over aligned groups of G lanes (G = 32: the whole warp); every lane of a group gets the sum
template <int G>
__device__ __forceinline__ double group_sum(double v) {
#pragma unroll
  for (int o = G / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__device__ __forceinline__ double mass_of(int z) {
  return (z >= 1 && z <= CREST_NUM_ELEMENTS) ? c_ams[z - 1] : 0.0;
}

struct Sym3 { double a11, a12, a13, a22, a23, a33; };
struct Mat3 { double m[3][3]; };   // m[row][col], 0-based

// trigonometric eigenvalues of a symmetric 3x3 (reference axis_module.f90:1051-1083)
__device__ void eigval_sym3(const Sym3& a, double w[3]) {
  const double twothirdpi = 8.0 * atan(1.0) / 3.0;
  double r = a.a12 * a.a12 + a.a13 * a.a13 + a.a23 * a.a23;
  const double q = (a.a11 + a.a22 + a.a33) / 3.0;
  w[0] = a.a11 - q;
  w[1] = a.a22 - q;
  w[2] = a.a33 - q;
  const double p = sqrt((w[0] * w[0] + w[1] * w[1] + w[2] * w[2] + 2 * r) / 6.0);
  r = (w[0] * (w[1] * w[2] - a.a23 * a.a23) - a.a12 * (a.a12 * w[2] - a.a23 * a.a13) +
       a.a13 * (a.a12 * a.a23 - w[1] * a.a13)) / (p * p * p) * 0.5;
  if (r <= -1.0) r = 0.5 * twothirdpi;
  else if (r >= 1.0) r = 0.0;
  else r = acos(r) / 3.0;
  w[2] = q + 2 * p * cos(r);
  w[0] = q + 2 * p * cos(r + twothirdpi);
  w[1] = 3 * q - w[0] - w[2];
}

// analytic eigenvectors (reference axis_module.f90:1088-1266), columns of q
__device__ void eigvec_sym3(Sym3 a, double w[3], Mat3& q) {
  const double eps = 2.220446049250313e-16;
  double norm, n1, n2, n3;
  w[0] = fmax(fabs(a.a11), fabs(a.a12));
  w[1] = fmax(fabs(a.a13), fabs(a.a22));
  w[2] = fmax(fabs(a.a23), fabs(a.a33));
  const double precon = fmax(w[0], fmax(w[1], w[2]));
  if (precon < eps) {
    w[0] = w[1] = w[2] = 0.0;
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) q.m[i][j] = (i == j) ? 1.0 : 0.0;
    return;
  }
  norm = 1.0 / precon;
  a.a11 *= norm; a.a12 *= norm; a.a22 *= norm; a.a13 *= norm; a.a23 *= norm; a.a33 *= norm;
  eigval_sym3(a, w);

  a.a11 -= w[0]; a.a22 -= w[0]; a.a33 -= w[0];
  q.m[0][0] = a.a12 * a.a23 - a.a13 * a.a22;
  q.m[1][0] = a.a13 * a.a12 - a.a11 * a.a23;
  q.m[2][0] = a.a11 * a.a22 - a.a12 * a.a12;
  q.m[0][1] = a.a12 * a.a33 - a.a13 * a.a23;
  q.m[1][1] = a.a13 * a.a13 - a.a11 * a.a33;
  q.m[2][1] = a.a11 * a.a23 - a.a12 * a.a13;
  q.m[0][2] = a.a22 * a.a33 - a.a23 * a.a23;
  q.m[1][2] = a.a23 * a.a13 - a.a12 * a.a33;
  q.m[2][2] = a.a12 * a.a23 - a.a22 * a.a13;
  n1 = q.m[0][0] * q.m[0][0] + q.m[1][0] * q.m[1][0] + q.m[2][0] * q.m[2][0];
  n2 = q.m[0][1] * q.m[0][1] + q.m[1][1] * q.m[1][1] + q.m[2][1] * q.m[2][1];
  n3 = q.m[0][2] * q.m[0][2] + q.m[1][2] * q.m[1][2] + q.m[2][2] * q.m[2][2];
  // column choice exactly as written in the reference (1150-1158): the running maximum is NOT
  // updated when column 2 is taken, so column 3 wins whenever n3 > n1
  int pick = 0;
  if (n2 > n1) pick = 1;
  if (n3 > n1) pick = 2;
  {
    const double nn = (pick == 0) ? n1 : ((pick == 1) ? n2 : n3);
    norm = sqrt(1.0 / nn);
    const double v0 = q.m[0][pick] * norm, v1 = q.m[1][pick] * norm, v2 = q.m[2][pick] * norm;
    q.m[0][0] = v0; q.m[1][0] = v1; q.m[2][0] = v2;
  }
  if (fabs(q.m[0][0]) > fabs(q.m[1][0])) {
    norm = sqrt(1.0 / (q.m[0][0] * q.m[0][0] + q.m[2][0] * q.m[2][0]));
    q.m[0][1] = -q.m[2][0] * norm; q.m[1][1] = 0.0; q.m[2][1] = +q.m[0][0] * norm;
  } else {
    norm = sqrt(1.0 / (q.m[1][0] * q.m[1][0] + q.m[2][0] * q.m[2][0]));
    q.m[0][1] = 0.0; q.m[1][1] = +q.m[2][0] * norm; q.m[2][1] = -q.m[1][0] * norm;
  }
  q.m[0][2] = q.m[1][0] * q.m[2][1] - q.m[2][0] * q.m[1][1];
  q.m[1][2] = q.m[2][0] * q.m[0][1] - q.m[0][0] * q.m[2][1];
  q.m[2][2] = q.m[0][0] * q.m[1][1] - q.m[1][0] * q.m[0][1];

  a.a11 += w[0]; a.a22 += w[0]; a.a33 += w[0];
  // A*U
  n1 = a.a11 * q.m[0][1] + a.a12 * q.m[1][1] + a.a13 * q.m[2][1];
  n2 = a.a12 * q.m[0][1] + a.a22 * q.m[1][1] + a.a23 * q.m[2][1];
  n3 = a.a13 * q.m[0][1] + a.a23 * q.m[1][1] + a.a33 * q.m[2][1];
  // A*V, stored over a(:,3) in the reference's (deliberately) permuted order (1200-1203)
  a.a33 = a.a13 * q.m[0][2] + a.a23 * q.m[1][2] + a.a33 * q.m[2][2];
  a.a13 = a.a11 * q.m[0][2] + a.a12 * q.m[1][2] + a.a13 * q.m[2][2];
  a.a23 = a.a12 * q.m[0][2] + a.a22 * q.m[1][2] + a.a23 * q.m[2][2];
  n1 = q.m[0][1] * n1 + q.m[1][1] * n2 + q.m[2][1] * n3 - w[1];
  n2 = q.m[0][1] * a.a13 + q.m[1][1] * a.a23 + q.m[2][1] * a.a33;
  n3 = q.m[0][2] * a.a13 + q.m[1][2] * a.a23 + q.m[2][2] * a.a33 - w[1];

  if (fabs(n1) >= fabs(n3)) {
    norm = fmax(fabs(n1), fabs(n2));
    if (norm > eps) {
      if (fabs(n1) >= fabs(n2)) { n2 = n2 / n1; n1 = sqrt(1.0 / (1.0 + n2 * n2)); n2 = n2 * n1; }
      else { n1 = n1 / n2; n2 = sqrt(1.0 / (1.0 + n1 * n1)); n1 = n1 * n2; }
      q.m[0][1] = n2 * q.m[0][1] - n1 * q.m[0][2];
      q.m[1][1] = n2 * q.m[1][1] - n1 * q.m[1][2];
      q.m[2][1] = n2 * q.m[2][1] - n1 * q.m[2][2];
    }
  } else {
    norm = fmax(fabs(n3), fabs(n2));
    if (norm > eps) {
      if (fabs(n3) >= fabs(n2)) { n2 = n2 / n3; n3 = sqrt(1.0 / (1.0 + n2 * n2)); n2 = n2 * n3; }
      else { n3 = n3 / n2; n2 = sqrt(1.0 / (1.0 + n3 * n3)); n3 = n3 * n2; }
      q.m[0][1] = n3 * q.m[0][1] - n2 * q.m[0][2];
      q.m[1][1] = n3 * q.m[1][1] - n2 * q.m[1][2];
      q.m[2][1] = n3 * q.m[2][1] - n2 * q.m[2][2];
    }
  }
  q.m[0][2] = q.m[1][0] * q.m[2][1] - q.m[2][0] * q.m[1][1];
  q.m[1][2] = q.m[2][0] * q.m[0][1] - q.m[0][0] * q.m[2][1];
  q.m[2][2] = q.m[0][0] * q.m[1][1] - q.m[1][0] * q.m[0][1];
  w[0] *= precon; w[1] *= precon; w[2] *= precon;
}

// axis_0 core on coordinates supplied through a functor: returns rot (MHz), evec, CMA.
// `get(k, z, px,py,pz)` yields atom k of the (sub)structure.  Lanes stride over atoms.
template <int G, class Get>
__device__ __forceinline__ void axis0_tensor(int n, Get get, int sl, Sym3& a, double cma[3]) {
  // CMA (axis_module.f90:384-412; sumw seeded with 1e-20); sl = lane within the group of G lanes
  double sw = 0.0, sx = 0.0, sy = 0.0, sz = 0.0;
  for (int k = sl; k < n; k += G) {
    int z; double x, y, zc;
    get(k, z, x, y, zc);
    const double m = mass_of(z);
    sw += m; sx += m * x; sy += m * y; sz += m * zc;
  }
  sw = group_sum<G>(sw) + 1.0e-20;
  sx = group_sum<G>(sx) / sw; sy = group_sum<G>(sy) / sw; sz = group_sum<G>(sz) / sw;
  cma[0] = sx; cma[1] = sy; cma[2] = sz;
  // inertia tensor (axis_module.f90:130-141)
  double t1 = 0, t2 = 0, t3 = 0, t4 = 0, t5 = 0, t6 = 0;
  for (int k = sl; k < n; k += G) {
    int z; double x, y, zc;
    get(k, z, x, y, zc);
    const double m = mass_of(z);
    x -= sx; y -= sy; zc -= sz;
    t1 += m * (y * y + zc * zc);
    t2 -= m * x * y;
    t3 += m * (zc * zc + x * x);
    t4 -= m * zc * x;
    t5 -= m * y * zc;
    t6 += m * (x * x + y * y);
  }
  a.a11 = group_sum<G>(t1) + 1.0e-10;
  a.a12 = group_sum<G>(t2) + 2.0e-10;
  a.a22 = group_sum<G>(t3) + 3.0e-10;
  a.a13 = group_sum<G>(t4) + 4.0e-10;
  a.a23 = group_sum<G>(t5) + 5.0e-10;
  a.a33 = group_sum<G>(t6) + 6.0e-10;
}

// same, with the masses from a table indexed by what get() returns in z (shared memory: the lookup
// through c_ams diverges over the lanes of a warp)
template <int G, class Get>
__device__ __forceinline__ void axis0_tensor_m(int n, Get get, const double* __restrict__ mass, int sl, Sym3& a,
                                               double cma[3]) {
  double sw = 0.0, sx = 0.0, sy = 0.0, sz = 0.0;
  for (int k = sl; k < n; k += G) {
    int z; double x, y, zc;
    get(k, z, x, y, zc);
    const double m = mass[z];
    sw += m; sx += m * x; sy += m * y; sz += m * zc;
  }
  sw = group_sum<G>(sw) + 1.0e-20;
  sx = group_sum<G>(sx) / sw; sy = group_sum<G>(sy) / sw; sz = group_sum<G>(sz) / sw;
  cma[0] = sx; cma[1] = sy; cma[2] = sz;
  double t1 = 0, t2 = 0, t3 = 0, t4 = 0, t5 = 0, t6 = 0;
  for (int k = sl; k < n; k += G) {
    int z; double x, y, zc;
    get(k, z, x, y, zc);
    const double m = mass[z];
    x -= sx; y -= sy; zc -= sz;
    t1 += m * (y * y + zc * zc);
    t2 -= m * x * y;
    t3 += m * (zc * zc + x * x);
    t4 -= m * zc * x;
    t5 -= m * y * zc;
    t6 += m * (x * x + y * y);
  }
  a.a11 = group_sum<G>(t1) + 1.0e-10;
  a.a12 = group_sum<G>(t2) + 2.0e-10;
  a.a22 = group_sum<G>(t3) + 3.0e-10;
  a.a13 = group_sum<G>(t4) + 4.0e-10;
  a.a23 = group_sum<G>(t5) + 5.0e-10;
  a.a33 = group_sum<G>(t6) + 6.0e-10;
}

// the eigenproblem of one structure: moments -> rot (MHz), eigenvectors.  Called per LANE by the
// pack kernel (32 different structures per warp) and lane-uniformly by rotconst_kernel.
__device__ __noinline__ void axis0_solve(const Sym3& a, double rot[3], Mat3& evec) {
  double eig[3];
  eigvec_sym3(a, eig, evec);
  for (int i = 0; i < 3; ++i) {
    for (int j = 0; j < 3; ++j)
      if (fabs(evec.m[i][j]) < 1e-9) evec.m[i][j] = 0.0;
    if (eig[i] < 3.0e-4) rot[i] = 0.0;
    else rot[i] = 2.9979245e+4 * 16.8576522 / eig[i];   // icm2MHz*Aamu2icm (axis_module.f90:25,34)
  }
}

template <class Get>
__device__ void axis0_warp(int n, Get get, int lane, double rot[3], Mat3& evec, double cma[3]) {
  Sym3 a;
  axis0_tensor<32>(n, get, lane, a, cma);
  axis0_solve(a, rot, evec);
}

// One warp owns 32 consecutive conformers.  Memory passes (CMA, inertia tensor, rotation, packing):
// groups of 8 lanes work on one conformer each, four conformers in flight per warp and eight rounds
// per pass (lanes stride over atoms: 8 x 24 B = 192 B contiguous per step).  The two 3x3
// eigenproblems per conformer -- FP64 trigonometry, divisions and square roots, by far the longest
// instruction stream of the kernel -- run ONE CONFORMER PER LANE (they used to run 32 times
// redundantly per warp).  Per-conformer arithmetic is the reference's operation order; only the
// association of the atom sums differs from a serial loop, as it always did.
// kPG = lanes per conformer in the memory passes (template parameter; CREST_CUDA_TUNE packg=4|8|16),
// kPC = 32 / kPG conformers in flight per warp

__device__ __forceinline__ double take(double v, int src) { return __shfl_sync(0xffffffffu, v, src); }

template <int kPG>
__global__ void __launch_bounds__(64, 12)   // <= 85 registers: 24 warps per SM, C3's 3125 warps are one wave
pack_axis_kernel(int nat, int nall, const int* __restrict__ at, const double* __restrict__ xyz,
                 const int* __restrict__ sel, int nsel, int rot_from_mask, PackLayout L,
                 double* __restrict__ xyz_axis, double* __restrict__ rot_out,
                 double* __restrict__ aniso_out, double* __restrict__ X,
                 double* __restrict__ xn_out, int W) {
  constexpr int kPC = 32 / kPG;
  const int lane = threadIdx.x & 31, g = lane / kPG, sl = lane % kPG;
  // W = conformers per warp (multiple of kPC, <= 32): 32 fills every lane of the eigen-solves; small
  // ensembles take fewer per warp so that the grid still covers the SMs
  const int c0 = (blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * W;
  if (c0 >= nall) return;
  const int nc = min(W, nall - c0);
  const int nround = (nc + kPC - 1) / kPC;

  // ---- pass 1a: CMA + inertia tensor; afterwards lane c holds conformer c0+c
  Sym3 myA{1.0, 0.0, 0.0, 2.0, 0.0, 3.0};     // idle lanes solve a harmless diagonal problem
  double myCma[3] = {0.0, 0.0, 0.0};
  for (int r = 0; r < nround; ++r) {
    const int c = r * kPC + g;
    const int cc = c < nc ? c : 0;             // idle groups recompute conformer c0 (discarded)
    const double* src = xyz + (size_t)3 * nat * (c0 + cc);
    auto get_all = [&](int k, int& z, double& x, double& y, double& zc) {
      z = at[k]; x = src[3 * k]; y = src[3 * k + 1]; zc = src[3 * k + 2];
    };
    Sym3 a; double cma[3];
    axis0_tensor<kPG>(nat, get_all, sl, a, cma);
    // lane L in [4r, 4r+4) takes the result of group L-4r
    const int from = kPG * (lane % kPC);
    const bool mine = lane / kPC == r && lane < nc;
    const double v[9] = {take(a.a11, from), take(a.a12, from), take(a.a13, from), take(a.a22, from),
                         take(a.a23, from), take(a.a33, from), take(cma[0], from), take(cma[1], from),
                         take(cma[2], from)};
    if (mine) {
      myA.a11 = v[0]; myA.a12 = v[1]; myA.a13 = v[2]; myA.a22 = v[3]; myA.a23 = v[4]; myA.a33 = v[5];
      myCma[0] = v[6]; myCma[1] = v[7]; myCma[2] = v[8];
    }
  }
  // ---- pass 1b: axis_4's eigenproblem, one conformer per lane
  double rot[3];
  Mat3 ev;
  axis0_solve(myA, rot, ev);
  {
    // calcxsum (axis_module.f90:372-378): keep a right-handed frame
    const double xsum = ev.m[0][0] * (ev.m[1][1] * ev.m[2][2] - ev.m[2][1] * ev.m[1][2]) +
                        ev.m[0][1] * (ev.m[1][2] * ev.m[2][0] - ev.m[1][0] * ev.m[2][2]) +
                        ev.m[0][2] * (ev.m[1][0] * ev.m[2][1] - ev.m[1][1] * ev.m[2][0]);
    if (xsum < 0.0) { ev.m[0][0] = -ev.m[0][0]; ev.m[1][0] = -ev.m[1][0]; ev.m[2][0] = -ev.m[2][0]; }
  }
  // ---- pass 1c + 2a: rotate into the principal-axis frame, then the tensor of the rotated
  // structure (all atoms, or the subset if substruc) for axis_1
  myA = Sym3{1.0, 0.0, 0.0, 2.0, 0.0, 3.0};
  for (int r = 0; r < nround; ++r) {
    const int c = r * kPC + g;
    const bool live = c < nc;
    const int cc = live ? c : 0;
    const double* src = xyz + (size_t)3 * nat * (c0 + cc);
    double* dst = xyz_axis + (size_t)3 * nat * (c0 + cc);
    double e[3][3], cm[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      cm[i] = take(myCma[i], cc);
#pragma unroll
      for (int j = 0; j < 3; ++j) e[i][j] = take(ev.m[i][j], cc);
    }
    if (live)
      for (int k = sl; k < nat; k += kPG) {
        const double x = src[3 * k] - cm[0], y = src[3 * k + 1] - cm[1], z = src[3 * k + 2] - cm[2];
#pragma unroll
        for (int j = 0; j < 3; ++j) {
          double s = 0.0;
          s = s + x * e[0][j];
          s = s + y * e[1][j];
          s = s + z * e[2][j];
          dst[3 * k + j] = s;
        }
      }
    __syncwarp();
    Sym3 a; double cma2[3];
    if (rot_from_mask) {
      auto get_sub = [&](int k, int& z, double& x, double& y, double& zc) {
        const int q = sel[k];
        z = at[q]; x = dst[3 * q]; y = dst[3 * q + 1]; zc = dst[3 * q + 2];
      };
      axis0_tensor<kPG>(nsel, get_sub, sl, a, cma2);
    } else {
      auto get_rot = [&](int k, int& z, double& x, double& y, double& zc) {
        z = at[k]; x = dst[3 * k]; y = dst[3 * k + 1]; zc = dst[3 * k + 2];
      };
      axis0_tensor<kPG>(nat, get_rot, sl, a, cma2);
    }
    const int from = kPG * (lane % kPC);
    const bool mine = lane / kPC == r && lane < nc;
    const double v[6] = {take(a.a11, from), take(a.a12, from), take(a.a13, from), take(a.a22, from),
                         take(a.a23, from), take(a.a33, from)};
    if (mine) { myA.a11 = v[0]; myA.a12 = v[1]; myA.a13 = v[2]; myA.a22 = v[3]; myA.a23 = v[4]; myA.a33 = v[5]; }
  }
  // ---- pass 2b: axis_1's eigenproblem per lane -> rot. constants
  axis0_solve(myA, rot, ev);
  if (lane < nc) {
    rot_out[3 * (c0 + lane) + 0] = rot[0];
    rot_out[3 * (c0 + lane) + 1] = rot[1];
    rot_out[3 * (c0 + lane) + 2] = rot[2];
    aniso_out[c0 + lane] = rotaniso_dev(rot[0], rot[1], rot[2]);
  }

  // ---- pass 3: RMSD subset, unweighted centroid (ls_rmsd.f90:105-119), packed operand + norm
  const double inv_n = 1.0 / (double)nsel;
  const int kpad = L.nslab * L.KU;
  const int last0 = kpad - 8;          // first slot of the last 8-atom block
  for (int r = 0; r < nround; ++r) {
    const int c = r * kPC + g;
    const bool live = c < nc;
    const int conf = c0 + (live ? c : 0);
    const double* dst = xyz_axis + (size_t)3 * nat * conf;
    double cx = 0.0, cy = 0.0, cz = 0.0;
    for (int k = sl; k < nsel; k += kPG) {
      const int a = sel[k];
      cx += dst[3 * a]; cy += dst[3 * a + 1]; cz += dst[3 * a + 2];
    }
    cx = group_sum<kPG>(cx) * inv_n; cy = group_sum<kPG>(cy) * inv_n; cz = group_sum<kPG>(cz) * inv_n;
    const int panel = conf / L.rows, row = conf % L.rows;
    double nrm = 0.0;
    for (int k = sl; k < kpad; k += kPG) {
      // slot k holds selected atom ks (or nothing): identity, except in a half tail where the
      // <= 4 atoms of the last block sit at its even slots
      int ks = k;
      if (L.half_tail && k >= last0) ks = ((k - last0) & 1) ? -1 : last0 + ((k - last0) >> 1);
      double x = 0.0, y = 0.0, z = 0.0;
      if (ks >= 0 && ks < nsel) {
        const int a = sel[ks];
        x = dst[3 * a] - cx; y = dst[3 * a + 1] - cy; z = dst[3 * a + 2] - cz;
        nrm += x * x + y * y + z * z;
      }
      if (live) {
        const int slab = k / L.KU, kk = k % L.KU;
        double* chunk = X + ((size_t)panel * L.nslab + slab) * L.chunk_doubles();
        chunk[((size_t)0 * L.rows + row) * L.KS + kk] = x;
        chunk[((size_t)1 * L.rows + row) * L.KS + kk] = y;
        chunk[((size_t)2 * L.rows + row) * L.KS + kk] = z;
      }
    }
    nrm = group_sum<kPG>(nrm);
    if (live && sl == 0) xn_out[conf] = nrm;
  }
}


// ---- pack_axis, one pass over the ensemble (round 2) ---------------------------------------------------
// The kernel above reads every conformer from global memory in seven dependent sweeps; ncu's source page
// puts 70 % of its stall samples on the FP64 adds that consume those loads (latency, one wave of warps)
// and 431 MB of DRAM reads on a 144 MB input.  Here a warp stages its NC = 32 / G conformers in shared
// memory ONCE (coalesced 16-byte loads, all independent), every pass - CMA, inertia tensor, rotation in
// place, second tensor, centroid, norm - runs out of shared memory with G lanes per conformer, and the two
// outputs leave as coalesced 16-byte stores: xyz_axis straight from the staged copy, the FP64 operand X
// row by row in ITS layout (lanes along the atom slots).  The conformer stride in shared memory is odd (in
// doubles), so the G-lanes-per-conformer passes are bank-conflict free.  All G lanes of a conformer hold
// its reduced tensor, so every lane runs the 3x3 eigen-solve (no exchange; the solve is ~2 % of the time).
// Per-conformer arithmetic is the reference's operation order as before (-fmad=false); only the
// association of the atom sums depends on G.
template <int G>
__global__ void __launch_bounds__(64, 8)
pack_axis_smem_kernel(int nat, int nall, const int* __restrict__ at, const double* __restrict__ xyz,
                      const int* __restrict__ sel, int nsel, int rot_from_mask, PackLayout L,
                      double* __restrict__ xyz_axis, double* __restrict__ rot_out,
                      double* __restrict__ aniso_out, double* __restrict__ X, double* __restrict__ xn_out,
                      int stride) {
  constexpr int NC = 32 / G;
  extern __shared__ double sm[];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  const int g = lane / G, sl = lane % G;
  const int n3 = 3 * nat;
  double* const mass = sm;                                   // [nat]            (block)
  int* const sel_s = reinterpret_cast<int*>(sm + nat);       // [nsel] -> 3 * atom index   (block)
  double* const cs = sm + nat + (nsel + 1) / 2 + (size_t)wib * ((size_t)NC * stride + 4 * NC);   // [NC][stride]
  double* const cen = cs + (size_t)NC * stride;              // [NC][4] centroid of the RMSD subset
  for (int k = threadIdx.x; k < nat; k += blockDim.x) mass[k] = mass_of(at[k]);
  for (int k = threadIdx.x; k < nsel; k += blockDim.x) sel_s[k] = 3 * sel[k];
  __syncthreads();
  const int c0 = (blockIdx.x * nwarp + wib) * NC;
  if (c0 >= nall) return;
  const int nc = min(NC, nall - c0);
  const bool live = g < nc;
  const double* const mine = cs + (size_t)(live ? g : 0) * stride;   // idle groups recompute conformer c0

  // ---- stage: global -> shared, 16 bytes per lane and step when 3*nat is even
  {
    const double* src = xyz + (size_t)n3 * c0;
    constexpr int NB = NC < 4 ? NC : 4;                       // conformers per batch of independent loads
    if ((n3 & 1) == 0) {
      for (int cb = 0; cb < nc; cb += NB)
        for (int r = 2 * lane; r < n3; r += 64) {
          double2 v[NB];
#pragma unroll
          for (int c = 0; c < NB; ++c)
            if (cb + c < nc) v[c] = *reinterpret_cast<const double2*>(src + (size_t)(cb + c) * n3 + r);
#pragma unroll
          for (int c = 0; c < NB; ++c)
            if (cb + c < nc) {
              cs[(size_t)(cb + c) * stride + r] = v[c].x;
              cs[(size_t)(cb + c) * stride + r + 1] = v[c].y;
            }
        }
    } else {
      for (int cb = 0; cb < nc; cb += NB)
        for (int r = lane; r < n3; r += 32) {
          double v[NB];
#pragma unroll
          for (int c = 0; c < NB; ++c)
            if (cb + c < nc) v[c] = src[(size_t)(cb + c) * n3 + r];
#pragma unroll
          for (int c = 0; c < NB; ++c)
            if (cb + c < nc) cs[(size_t)(cb + c) * stride + r] = v[c];
        }
    }
  }
  __syncwarp();

  // ---- axis_4: CMA + inertia tensor (all atoms), eigenproblem, right-handed frame
  Sym3 a;
  double cma[3];
  {
    auto get_all = [&](int k, int& z, double& x, double& y, double& zc) {
      z = k; x = mine[3 * k]; y = mine[3 * k + 1]; zc = mine[3 * k + 2];
    };
    axis0_tensor_m<G>(nat, get_all, mass, sl, a, cma);
  }
  double rot[3];
  Mat3 ev;
  axis0_solve(a, rot, ev);
  {
    const double xsum = ev.m[0][0] * (ev.m[1][1] * ev.m[2][2] - ev.m[2][1] * ev.m[1][2]) +
                        ev.m[0][1] * (ev.m[1][2] * ev.m[2][0] - ev.m[1][0] * ev.m[2][2]) +
                        ev.m[0][2] * (ev.m[1][0] * ev.m[2][1] - ev.m[1][1] * ev.m[2][0]);
    if (xsum < 0.0) { ev.m[0][0] = -ev.m[0][0]; ev.m[1][0] = -ev.m[1][0]; ev.m[2][0] = -ev.m[2][0]; }
  }
  // ---- rotate in place (each atom is owned by one lane)
  if (live) {
    double* own = cs + (size_t)g * stride;
    for (int k = sl; k < nat; k += G) {
      const double x = own[3 * k] - cma[0], y = own[3 * k + 1] - cma[1], z = own[3 * k + 2] - cma[2];
#pragma unroll
      for (int j = 0; j < 3; ++j) {
        double s = 0.0;
        s = s + x * ev.m[0][j];
        s = s + y * ev.m[1][j];
        s = s + z * ev.m[2][j];
        own[3 * k + j] = s;
      }
    }
  }
  __syncwarp();
  // ---- axis_1 on the rotated (sub)structure -> rot. constants
  {
    double cma2[3];
    if (rot_from_mask) {
      auto get_sub = [&](int k, int& z, double& x, double& y, double& zc) {
        const int q3 = sel_s[k];
        z = q3 / 3; x = mine[q3]; y = mine[q3 + 1]; zc = mine[q3 + 2];
      };
      axis0_tensor_m<G>(nsel, get_sub, mass, sl, a, cma2);
    } else {
      auto get_rot = [&](int k, int& z, double& x, double& y, double& zc) {
        z = k; x = mine[3 * k]; y = mine[3 * k + 1]; zc = mine[3 * k + 2];
      };
      axis0_tensor_m<G>(nat, get_rot, mass, sl, a, cma2);
    }
  }
  axis0_solve(a, rot, ev);
  if (live && sl == 0) {
    const int conf = c0 + g;
    rot_out[3 * conf + 0] = rot[0];
    rot_out[3 * conf + 1] = rot[1];
    rot_out[3 * conf + 2] = rot[2];
    aniso_out[conf] = rotaniso_dev(rot[0], rot[1], rot[2]);
  }
  // ---- RMSD subset: unweighted centroid (ls_rmsd.f90:105-119) and centred squared norm
  {
    double cx = 0.0, cy = 0.0, cz = 0.0;
    for (int k = sl; k < nsel; k += G) {
      const int q3 = sel_s[k];
      cx += mine[q3]; cy += mine[q3 + 1]; cz += mine[q3 + 2];
    }
    const double inv_n = 1.0 / (double)nsel;
    cx = group_sum<G>(cx) * inv_n; cy = group_sum<G>(cy) * inv_n; cz = group_sum<G>(cz) * inv_n;
    double nrm = 0.0;
    for (int k = sl; k < nsel; k += G) {
      const int q3 = sel_s[k];
      const double x = mine[q3] - cx, y = mine[q3 + 1] - cy, z = mine[q3 + 2] - cz;
      nrm += x * x + y * y + z * z;
    }
    nrm = group_sum<G>(nrm);
    if (live && sl == 0) {
      cen[4 * g] = cx; cen[4 * g + 1] = cy; cen[4 * g + 2] = cz;
      xn_out[c0 + g] = nrm;
    }
  }
  __syncwarp();
  // ---- xyz_axis: the staged copy, coalesced
  {
    double* dst = xyz_axis + (size_t)n3 * c0;
    if ((n3 & 1) == 0) {
      for (int c = 0; c < nc; ++c)
#pragma unroll 2
        for (int r = 2 * lane; r < n3; r += 64)
          *reinterpret_cast<double2*>(dst + (size_t)c * n3 + r) =
              make_double2(cs[(size_t)c * stride + r], cs[(size_t)c * stride + r + 1]);
    } else {
      for (int c = 0; c < nc; ++c)
        for (int r = lane; r < n3; r += 32) dst[(size_t)c * n3 + r] = cs[(size_t)c * stride + r];
    }
  }
  // ---- packed operand X: chunk [panel][slab][comp][rows][KS]; per (slab, comp, conformer) one row of KU
  // slots, lanes along the slots (two per lane: KU and KS are even).  Slot k holds selected atom ks:
  // identity, except in a half tail where the <= 4 atoms of the last block sit at its even slots.
  {
    const int kpad = L.nslab * L.KU, last0 = kpad - 8;
    const int panel = c0 / L.rows, row0 = c0 % L.rows;            // NC divides L.rows: one panel per warp
    for (int slab = 0; slab < L.nslab; ++slab) {
      double* chunk = X + ((size_t)panel * L.nslab + slab) * L.chunk_doubles();
      for (int kk = 2 * lane; kk < L.KU; kk += 64) {
        int q3[2];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const int k = slab * L.KU + kk + h;
          int ks = k;
          if (L.half_tail && k >= last0) ks = ((k - last0) & 1) ? -1 : last0 + ((k - last0) >> 1);
          q3[h] = (ks >= 0 && ks < nsel) ? sel_s[ks] : -1;
        }
#pragma unroll
        for (int comp = 0; comp < 3; ++comp) {
#pragma unroll 4
          for (int c = 0; c < nc; ++c) {
            {
              const double* src = cs + (size_t)c * stride;
              const double ce = cen[4 * c + comp];
              double2 v;
              v.x = q3[0] >= 0 ? src[q3[0] + comp] - ce : 0.0;
              v.y = q3[1] >= 0 ? src[q3[1] + comp] - ce : 0.0;
              *reinterpret_cast<double2*>(chunk + ((size_t)comp * L.rows + row0 + c) * L.KS + kk) = v;
            }
          }
        }
      }
    }
  }
}

__global__ void __launch_bounds__(128)
rotconst_kernel(int nat, int n, const int* __restrict__ at, const double* __restrict__ xyz,
                double* __restrict__ rot_out) {
  const int lane = threadIdx.x & 31;
  const int conf = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (conf >= n) return;
  const double* src = xyz + (size_t)3 * nat * conf;
  double rot[3], cma[3];
  Mat3 ev;
  auto get_all = [&](int k, int& z, double& x, double& y, double& zc) {
    z = at[k]; x = src[3 * k]; y = src[3 * k + 1]; zc = src[3 * k + 2];
  };
  axis0_warp(nat, get_all, lane, rot, ev, cma);
  if (lane == 0) {
    rot_out[3 * conf + 0] = rot[0]; rot_out[3 * conf + 1] = rot[1]; rot_out[3 * conf + 2] = rot[2];
  }
}

int launch_pack_axis(int nat, int nall, const int* d_at, const double* d_xyz, const int* d_sel,
                     int nsel, int rot_from_mask, const PackLayout& L, double* d_xyz_axis,
                     double* d_rot, double* d_aniso, double* d_X, double* d_xn,
                     cudaStream_t stream) {
  if (nall <= 0) return 0;
  int dev = 0, nsm = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev);
  // One-pass kernel: G lanes per conformer, NC = 32 / G conformers staged per warp.  The passes are chains
  // of FP64 operations on shared-memory operands: they need ~16 resident warps per SM to keep the pipes busy
  // (measured with 8: 12 % warps active, issue 27 %, no faster than the multi-sweep kernel), so G = the
  // smallest power of two whose staged copy stays under 16 KB per warp (C3, 60 atoms: G = 4, 8 conformers =
  // 11.6 KB; C5, 40 atoms: G = 2; C4, 300 atoms: G = 16, two conformers).
  // nat > 1200 (one conformer > 29 KB) and CREST_CUDA_TUNE `packold` take the multi-sweep kernel.
  const int stride = (3 * nat) | 1;                             // doubles per staged conformer, odd
  int G2 = 2;
  while (G2 < 32 && (size_t)(32 / G2) * stride * 8 > (size_t)16 << 10) G2 <<= 1;
  // G depends on the atom count ONLY: the association of the atom sums follows G, and rot / xyz_axis must not
  // depend on how the upload pipeline cut the ensemble into row blocks (test_engine_upload_pipeline_blocks)
  if (tune().pack_g < 0 && (tune().pack_g == -2 || tune().pack_g == -4 || tune().pack_g == -8 ||
                            tune().pack_g == -16 || tune().pack_g == -32)) G2 = -tune().pack_g;   // packg=-<G>: force
  const bool one_pass = !tune().pack_old && (size_t)(32 / G2) * stride * 8 <= (size_t)29 << 10;
  if (one_pass) {
    const int NC = 32 / G2, wpb = 2;
    const int nwarp = (nall + NC - 1) / NC;
    const dim3 grid((nwarp + wpb - 1) / wpb), block(wpb * 32);
    const size_t smem = sizeof(double) * ((size_t)nat + (nsel + 1) / 2 + (size_t)wpb * ((size_t)NC * stride + 4 * NC));
#define CREST_PACK1_ARGS nat, nall, d_at, d_xyz, d_sel, nsel, rot_from_mask, L, d_xyz_axis, d_rot, d_aniso, d_X, d_xn, stride
#define CREST_PACK1(GG)                                                                                     \
    case GG:                                                                                                \
      CREST_CUDA_TRY(cudaFuncSetAttribute(pack_axis_smem_kernel<GG>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
      pack_axis_smem_kernel<GG><<<grid, block, smem, stream>>>(CREST_PACK1_ARGS);                          \
      break;
    switch (G2) {
      CREST_PACK1(2) CREST_PACK1(4) CREST_PACK1(8) CREST_PACK1(16) CREST_PACK1(32)
      default: set_error("pack_axis: bad lane group %d", G2); return 1;
    }
#undef CREST_PACK1
#undef CREST_PACK1_ARGS
    CREST_CUDA_TRY(cudaGetLastError());
    return 0;
  }
  // Conformers per warp: 32 when that still gives every SM its 24 resident warps' worth of work
  // (C3: 3125 warps, one wave); otherwise halve down to one round of 32/G conformers so that small
  // ensembles (C4: 20 000 x 300) are not serialised inside a few hundred warps.
  int G = tune().pack_g;
  if (G != 4 && G != 16) G = 8;
  int W = 32;
  while (W > 32 / G && (nall + W - 1) / W < nsm * 16) W >>= 1;
  const int wpb = 2;
  const int nwarp = (nall + W - 1) / W;
  const dim3 grid((nwarp + wpb - 1) / wpb), block(wpb * 32);
#define CREST_PACK_ARGS nat, nall, d_at, d_xyz, d_sel, nsel, rot_from_mask, L, d_xyz_axis, d_rot, d_aniso, d_X, d_xn, W
  switch (G) {
    case 4: pack_axis_kernel<4><<<grid, block, 0, stream>>>(CREST_PACK_ARGS); break;
    case 16: pack_axis_kernel<16><<<grid, block, 0, stream>>>(CREST_PACK_ARGS); break;
    default: pack_axis_kernel<8><<<grid, block, 0, stream>>>(CREST_PACK_ARGS); break;
  }
#undef CREST_PACK_ARGS
  CREST_CUDA_TRY(cudaGetLastError());
  return 0;
}

int launch_rotconst(int nat, int n, const int* d_at, const double* d_xyz, double* d_rot,
                    cudaStream_t stream) {
  if (n <= 0) return 0;
  const int wpb = 4;
  rotconst_kernel<<<(n + wpb - 1) / wpb, wpb * 32, 0, stream>>>(nat, n, d_at, d_xyz, d_rot);
  CREST_CUDA_TRY(cudaGetLastError());
  return 0;
}

}  // namespace crest
