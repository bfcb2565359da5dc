// ORACLE — TEST INFRASTRUCTURE ONLY (see ls_rmsd_ref.hpp header).  PARITY UNPINNED.
// Plain-C entry points so tests/ and bench.py's cpu_baseline leg can drive the restatement
// through ctypes.  Nothing in crest_b200/ links or loads this library.
#include <cstdio>
#include <cstdlib>
#include <omp.h>
#include "cregen_ref.hpp"

using namespace crest_oracle;

extern "C" {

// out[0]=rmsd out[1]=lambda out[2]=xnorm out[3]=ynorm out[4..12]=R row-major, out[13..16]=q
double oracle_rmsd(int n, const double* c1, const double* c2, double* out) {
  std::vector<double> x((size_t)3 * n), y((size_t)3 * n);
  RmsdOut o;
  rmsd_full(n, c1, c2, x.data(), y.data(), o);
  if (out) {
    out[0] = o.error; out[1] = o.lambda; out[2] = o.xnorm; out[3] = o.ynorm;
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) out[4 + 3 * i + j] = o.R[i][j];
    for (int i = 0; i < 4; ++i) out[13 + i] = o.q[i];
  }
  return o.error;
}

// leading eigenvalue of a symmetric 4x4 (row-major S[16]); returns io
int oracle_dstmev(const double* S, double* lambda, double* evec) {
  M4 A;
  for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) A(i + 1, j + 1) = S[4 * i + j];
  return dstmev(A, *lambda, evec);
}

// axis(nat,at,xyz) in place for n structures + rot from all atoms
void oracle_axis(int nat, int n, const int* at, double* xyz, double* rot) {
  for (int i = 0; i < n; ++i) axis_inplace(nat, at, xyz + (size_t)3 * nat * i, rot + 3 * i);
}

// rot. constants only (axis_1)
void oracle_rotconst(int nat, int n, const int* at, const double* xyz, double* rot) {
  for (int i = 0; i < n; ++i) {
    double av; M3 ev;
    axis_0(nat, at, xyz + (size_t)3 * nat * i, rot + 3 * i, av, ev);
  }
}

int oracle_equalrotaniso(const double* roti, const double* rotj, double bthr, double bthrmax,
                         double bthrshift) {
  double r[6] = {roti[0], roti[1], roti[2], rotj[0], rotj[1], rotj[2]};
  return equalrotaniso(1, 2, r, bthr, bthrmax, bthrshift) ? 1 : 0;
}

double oracle_bthrerf(double bthr, double aniso, double bthrmax, double bthrshift) {
  return bthrerf(bthr, aniso, bthrmax, bthrshift);
}
double oracle_rotaniso(const double* rot3) { return rotaniso(1, rot3); }

// cregen.f90:2229-2283 cregen_rmsdalign: structures 2..nall rotated onto structure 1 by the
// heavy-atom (at > 2) rotation of rmsd(..., option 1) (ls_rmsd.f90:72-221); xyz (nall,nat,3) in place.
// Returns the number of structures whose 4x4 eigen-solve failed (left unrotated).
int oracle_rmsdalign(int nat, int nall, const int* at, double* xyz) {
  int nath = 0;
  for (int j = 0; j < nat; ++j) nath += at[j] > 2;
  if (nath == 0 || nall < 2) return 0;
  std::vector<double> c0(3 * (size_t)nath), c1(3 * (size_t)nath), x(3 * (size_t)nath),
      y(3 * (size_t)nath), c2(3 * (size_t)nat);
  int i = 0;
  for (int j = 0; j < nat; ++j)
    if (at[j] > 2) { for (int a = 0; a < 3; ++a) c0[3 * i + a] = xyz[3 * j + a]; ++i; }
  int nfail = 0;
  for (int k = 1; k < nall; ++k) {
    double* cur = xyz + (size_t)3 * nat * k;
    i = 0;
    for (int j = 0; j < nat; ++j)
      if (at[j] > 2) { for (int a = 0; a < 3; ++a) c1[3 * i + a] = cur[3 * j + a]; ++i; }
    RmsdOut o;
    rmsd_full(nath, c1.data(), c0.data(), x.data(), y.data(), o);
    if (o.error < 0.0) { ++nfail; continue; }
    double U[3][3];
    rotation_matrix(o.q, U);
    for (int j = 0; j < nat; ++j)                      // c2 = matmul(U, xyz(:,:,k))
      for (int a = 0; a < 3; ++a)
        c2[3 * j + a] = U[a][0] * cur[3 * j] + U[a][1] * cur[3 * j + 1] + U[a][2] * cur[3 * j + 2];
    for (int j = 0; j < 3 * nat; ++j) cur[j] = c2[j];
  }
  return nfail;
}

int oracle_esort(int nall, double* energies, int* pos, double ewin) {
  return esort(nall, energies, pos, ewin);
}

void oracle_backtrack(int n, int* arr1 /*n+1, [0] unused*/) {
  for (int i = 1; i <= n; ++i) { int k; backtrack(arr1, i, k); }
}

// Full cregen_CRE.  Outputs (caller allocated): rot(3,nall), er_out(nall), dup(nall) 1-based
// parents, orderref(nall) 1-based, group(0:nall), enuc(nall), rmat (klong floats or NULL),
// info[0]=nallout info[1]=nkept info[2]=klong info[3]=nfail.
int oracle_cregen_cre(int nat, int nall, const int* at, double* xyz, const double* er,
                      const int* include_rmsd, int substruc, int heavy, int enantio, int nosort,
                      double ethr, double rthr, double bthr, double bthrmax, double bthrshift,
                      double* rot, double* er_out, int* dup, int* orderref, int* group,
                      double* enuc, float* rmat, long long rmat_cap, long long* info) {
  CreParams p;
  p.ethr = ethr; p.rthr = rthr; p.bthr = bthr; p.bthrmax = bthrmax; p.bthrshift = bthrshift;
  p.enantio = enantio; p.heavy = heavy; p.substruc = substruc; p.include_rmsd = include_rmsd;
  p.nosort = nosort; p.keep_rmat = rmat != nullptr;
  CreResult r;
  cregen_cre(nat, nall, at, xyz, er, p, r);
  for (int i = 0; i < nall; ++i) {
    if (rot) { rot[3 * i] = r.rot[3 * i]; rot[3 * i + 1] = r.rot[3 * i + 1]; rot[3 * i + 2] = r.rot[3 * i + 2]; }
    if (er_out) er_out[i] = r.er[i];
    if (dup) dup[i] = r.dup[i + 1];
    if (orderref) orderref[i] = r.orderref[i + 1];
  }
  if (group) for (int i = 0; i <= nall; ++i) group[i] = r.group[i];
  if (enuc) for (size_t i = 0; i < r.enuc.size(); ++i) enuc[i] = r.enuc[i];
  if (rmat) {
    if ((long long)r.rmat.size() > rmat_cap) return 2;
    for (size_t i = 0; i < r.rmat.size(); ++i) rmat[i] = r.rmat[i];
  }
  if (info) { info[0] = r.nallout; info[1] = r.nkept; info[2] = r.klong; info[3] = r.nfail; }
  return 0;
}

// Bounded CPU-baseline sample: the reference's K1 loop (cregen.f90:1269-1291 shape: serial i,
// `omp for` over j, two rmsd calls per pair, float32 store) for rows i0..i1-1 (0-based) of an
// ensemble already in the principal-axis frame and packed to the RMSD atoms (3,n,nall).
// rows_out: float32, row r holds j=0..i-1 contiguous at offset sum_{i'<i, i'>=i0} i'.
// Returns the number of pairs evaluated.
long long oracle_pair_rows(int n, int nall, const double* xyz, const double* er, double ethr,
                           int enantio, int i0, int i1, float* rows_out) {
  long long npairs = 0;
  const size_t stride = (size_t)3 * n;
  size_t off = 0;
  for (int i = i0; i < i1 && i < nall; ++i) {
    const double* c0 = xyz + stride * i;
    long long cnt = 0;
#pragma omp parallel reduction(+ : cnt)
    {
      std::vector<double> c1(stride), wx(stride), wy(stride);
#pragma omp for
      for (int j = 0; j < i; ++j) {
        const double de = (er[i] - er[j]) * kAutokcal;
        if (de < ethr) {
          std::memcpy(c1.data(), xyz + stride * j, stride * sizeof(double));
          const double rdum = rmsd(n, c0, c1.data(), wx.data(), wy.data());
          double rdum2 = rdum;
          if (enantio) {
            for (int k = 0; k < n; ++k) c1[3 * k] = -c1[3 * k];
            rdum2 = rmsd(n, c0, c1.data(), wx.data(), wy.data());
          }
          if (rows_out) rows_out[off + j] = (float)std::min(rdum, rdum2);
          cnt += 1;
        } else if (rows_out) {
          rows_out[off + j] = -2.0f;
        }
      }
    }
    npairs += cnt;
    off += (size_t)i;
  }
  return npairs;
}

int oracle_num_threads(void) { return omp_get_max_threads(); }

}  // extern "C"
