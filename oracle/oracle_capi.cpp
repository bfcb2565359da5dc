// ORACLE — TEST INFRASTRUCTURE ONLY (see ls_rmsd_ref.hpp header).  PARITY UNPINNED.
// Plain-C entry points so tests/ and bench.py's cpu_baseline leg can drive the restatement
// through ctypes.  Nothing in crest_b200/ links or loads this library.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <algorithm>
#include <vector>
#include <omp.h>
#include "cregen_ref.hpp"

using namespace crest_oracle;

static double g_boundary_tol = -1.0;
static std::vector<int> g_boundary_rows;

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
// Test listing: rows whose band holds an RMSD within `tol` of rthr / 2*rthr in the NEXT
// oracle_cregen_cre call (tol < 0 switches the listing off); read back with oracle_boundary_rows.
void oracle_set_boundary_tol(double tol) { g_boundary_tol = tol; }
int oracle_boundary_rows(int* rows, int cap) {
  const int n = (int)g_boundary_rows.size();
  if (rows) for (int i = 0; i < n && i < cap; ++i) rows[i] = g_boundary_rows[i];
  return n;
}

int oracle_cregen_cre(int nat, int nall, const int* at, double* xyz, const double* er,
                      const int* include_rmsd, int substruc, int heavy, int enantio, int nosort,
                      double ethr, double rthr, double bthr, double bthrmax, double bthrshift,
                      double* rot, double* er_out, int* dup, int* orderref, int* group,
                      double* enuc, float* rmat, long long rmat_cap, long long* info) {
  CreParams p;
  p.ethr = ethr; p.rthr = rthr; p.bthr = bthr; p.bthrmax = bthrmax; p.bthrshift = bthrshift;
  p.enantio = enantio; p.heavy = heavy; p.substruc = substruc; p.include_rmsd = include_rmsd;
  p.nosort = nosort; p.keep_rmat = rmat != nullptr;
  p.boundary_tol = g_boundary_tol;
  CreResult r;
  cregen_cre(nat, nall, at, xyz, er, p, r);
  g_boundary_rows = r.boundary_rows;
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

// ---- ensemblecomp.f90:213-248 (rmat = block-wise minimum RMSD, seeded with 100) ----
// xyz1(3,nat,nall1), xyz2(3,nat,nall2); b/e 1-based inclusive; rmat(nconf1,nconf2) column-major.
void oracle_compare_blocks(int nat, const double* xyz1, int nconf1, const int* b1, const int* e1,
                           const double* xyz2, int nconf2, const int* b2, const int* e2,
                           double* rmat) {
  const size_t stride = (size_t)3 * nat;
  for (int i = 0; i < nconf1; ++i) {
    for (int j = 0; j < nconf2; ++j) {
      const int nrot1 = e1[i] - b1[i] + 1, nrot2 = e2[j] - b2[j] + 1;
      std::vector<double> rmat2((size_t)nrot1 * nrot2, 100.0);
      for (int k = 0; k < nrot1; ++k) {
        const double* c1 = xyz1 + stride * (b1[i] - 1 + k);
#pragma omp parallel
        {
          std::vector<double> wx(stride), wy(stride);
#pragma omp for
          for (int l = 0; l < nrot2; ++l) {
            const double* c2 = xyz2 + stride * (b2[j] - 1 + l);
            rmat2[(size_t)k * nrot2 + l] = rmsd(nat, c1, c2, wx.data(), wy.data());
          }
        }
      }
      double m = rmat2[0];
      for (double v : rmat2) m = std::min(m, v);
      rmat[i + (size_t)nconf1 * j] = m;
    }
  }
}

// ensemblecomp.f90:258-300 as data: returns the number of lines; left/right 1-based, 0 = blank
int oracle_compare_correlate(int nconf1, int nconf2, const double* en1, const double* en2,
                             const double* rmat, double rthr, int* left, int* right) {
  int kk = nconf1, ll = nconf2, nl = 0;
  if (kk == 0 && ll == 0) return 0;
  for (;;) {
    bool matching = false;
    if (kk > 0 && ll > 0) {
      const double rval = rmat[(kk - 1) + (size_t)nconf1 * (ll - 1)];
      if (rval <= rthr) {
        left[nl] = kk; right[nl] = ll; ++nl;
        kk = kk - 1; ll = ll - 1;
        matching = true;
      }
    }
    if (!matching) {
      if (kk > 0 && ll > 0) {
        const double min_1 = en1[kk - 1], min_2 = en2[ll - 1];
        if (min_1 >= min_2) { left[nl] = kk; right[nl] = 0; ++nl; kk = kk - 1; }
        else { left[nl] = 0; right[nl] = ll; ++nl; ll = ll - 1; }
      }
      if (kk > 0 && ll == 0) { left[nl] = kk; right[nl] = 0; ++nl; kk = kk - 1; }
      if (kk == 0 && ll > 0) { left[nl] = 0; right[nl] = ll; ++nl; ll = ll - 1; }
    }
    if (kk == 0 && ll == 0) break;
  }
  return nl;
}

// ---- entropy/entropic.f90:938-1097 intraconfRMSD, statement by statement ----
static int unique_core(const std::vector<double>& mat, int n, double thr) {   // 1074-1095
  std::vector<int> mp((size_t)n, 0);
  int gr = 0;
  for (int i = 0; i < n; ++i) {
    if (mp[i] == 0) { gr = gr + 1; mp[i] = gr; }
    for (int j = 0; j < n; ++j)
      if (mat[(size_t)i + (size_t)n * j] < thr) mp[j] = gr;
  }
  return gr;
}

static int symmap_ref(char symlist[100][3], const char* symflag) {            // 1096-1116
  auto len_trim = [](const char* s) { int n = 3; while (n > 0 && s[n - 1] == ' ') --n; return n; };
  int res = 1;
  for (int i = 1; i <= 100; ++i) {
    if (std::memcmp(symlist[i - 1], symflag, 3) == 0) { res = i; break; }   // blank-padded compare
    if (len_trim(symlist[i - 1]) < 1) { std::memcpy(symlist[i - 1], symflag, 3); res = i; break; }
  }
  return res;
}

// xyz(3,nat,nall); inc(nat) 0/1; low/nrot per conformer (1-based low); symsym 3 chars each;
// fac(nconf) out, enantiofac(nconf) in/out.  rmat_out (optional) receives the rmat of conformer
// `rmat_conf` (0-based) for RMSD-level checks, nrot x nrot column-major.
void oracle_intraconf(int nat, int nall, const double* xyz, const int* inc, int nconf,
                      const int* low, const int* nrot, const char* symsym, double* fac,
                      double* enantiofac, int rmat_conf, double* rmat_out) {
  int rednat = 0;
  for (int k = 0; k < nat; ++k) rednat += inc[k];
  if (rednat < 2) { for (int c = 0; c < nconf; ++c) fac[c] = 1.0; return; }
  const double rthr = 0.125;
  char symref[100][3];
  std::memset(symref, ' ', sizeof(symref));
  double symenantio[100];
  for (double& v : symenantio) v = 1.0;
  const size_t stride = (size_t)3 * nat;
  auto cpincluded = [&](const double* c, bool flipx, double* cr) {
    int k = 0;
    for (int i = 0; i < nat; ++i)
      if (inc[i] == 1) {
        cr[3 * k] = flipx ? -c[3 * i] : c[3 * i];
        cr[3 * k + 1] = c[3 * i + 1]; cr[3 * k + 2] = c[3 * i + 2];
        ++k;
      }
  };
  std::vector<double> c1r((size_t)3 * rednat), c2r((size_t)3 * rednat), wx((size_t)3 * rednat),
      wy((size_t)3 * rednat);
  for (int co = 0; co < nconf; ++co) {
    const int nr = nrot[co];
    int groups = 1;
    if (nr > 1) {
      std::vector<double> rmat((size_t)nr * nr, 0.0);
      bool enantio = false;
      const double* base = xyz + stride * (low[co] - 1);
      for (int i = 0; i < nr; ++i) {
        cpincluded(base + stride * i, false, c1r.data());
        for (int j = 0; j < i; ++j) {
          cpincluded(base + stride * j, false, c2r.data());
          const double r = rmsd(rednat, c1r.data(), c2r.data(), wx.data(), wy.data());
          rmat[(size_t)i + (size_t)nr * j] = r;
          rmat[(size_t)j + (size_t)nr * i] = r;
          if (r > rthr) {
            cpincluded(base + stride * j, true, c2r.data());
            const double re = rmsd(rednat, c1r.data(), c2r.data(), wx.data(), wy.data());
            if (re < rthr * 0.75) enantio = true;
          }
        }
      }
      groups = unique_core(rmat, nr, rthr);
      if (enantio) enantiofac[co] = 2.0;
      if (rmat_out && co == rmat_conf) std::memcpy(rmat_out, rmat.data(), sizeof(double) * rmat.size());
    }
    const int k = symmap_ref(symref, symsym + 3 * (size_t)co);
    if (enantiofac[co] > symenantio[k - 1]) symenantio[k - 1] = enantiofac[co];
    fac[co] = (double)groups;
  }
  for (int co = 0; co < nconf; ++co) {
    const int k = symmap_ref(symref, symsym + 3 * (size_t)co);
    enantiofac[co] = symenantio[k - 1];
  }
}

// GC crossing de-duplication, /root/reference/src/algos/crossing.f90:227-338 with ONE thread (loop order
// i, then j < i = the order of `cand`): rval against the reference structure (263-273), the accepted-list
// check inside the critical section (275-321) and sdselect (401-432), statement by statement.  `cand`
// holds the structures that survived GMETRY and the CN clash check (not on the RMSD path).
// counters: [0] discarded by rval <= rthr, [1] ident, [2] rcount, [3] candidates looked at before
// stop_crossing.  taken[k] = 1-based candidate index in accepted slot k (0 = empty).
static int sdselect_ref(int n, const double* rms, double rval) {
  int pos = 0;
  double avg = 0.0;
  for (int i = 0; i < n; ++i) avg = avg + rms[i];
  const double sum = avg;
  avg = avg / (double)(float)n;
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

void oracle_crossing_dedup(int nat, int ncand, const double* cand, const double* xyzref, double rthr,
                           int maxgen2, int huge_number, int* taken, double* rms, int* ntaken_out,
                           long long* counters, double* rmsdsum, double* rval_out) {
  const size_t rec = (size_t)3 * nat;
  std::vector<double> wx(rec), wy(rec), xyzgen((size_t)maxgen2 * rec);
  const double rthr100 = rthr / 100.0;
  int ntaken = 0;
  long long ierr = 0, ident = 0, rcount = 0, processed = 0;
  double rmsdavg = 0.0;
  bool stop_crossing = false;
  for (int k = 0; k < maxgen2; ++k) { taken[k] = 0; rms[k] = 0.0; }
  for (int c = 0; c < ncand; ++c) {
    if (stop_crossing) continue;
    processed++;
    const double* cdum = cand + rec * c;
    const double rval = rmsd(nat, cdum, xyzref, wx.data(), wy.data());
    if (rval_out) rval_out[c] = rval;
    if (rval <= rthr) { ierr = ierr + 1; continue; }
    rcount = rcount + 1;
    rmsdavg = rmsdavg + rval;
    if (ntaken < 1) {
      ntaken = ntaken + 1;
      std::memcpy(&xyzgen[(size_t)(ntaken - 1) * rec], cdum, rec * sizeof(double));
      rms[ntaken - 1] = rval; taken[ntaken - 1] = c + 1;
    } else {
      bool fail = false;
      for (int m = 1; m <= ntaken; ++m) {
        if (std::fabs(rms[m - 1] - rval) < rthr100) {
          const double rval2 = rmsd(nat, cdum, &xyzgen[(size_t)(m - 1) * rec], wx.data(), wy.data());
          if (rval2 < rthr) { ident = ident + 1; fail = true; break; }
        }
      }
      if (!fail) {
        if (ntaken < maxgen2) {
          ntaken = ntaken + 1;
          std::memcpy(&xyzgen[(size_t)(ntaken - 1) * rec], cdum, rec * sizeof(double));
          rms[ntaken - 1] = rval; taken[ntaken - 1] = c + 1;
        } else if (huge_number) {
          stop_crossing = true;
        } else {
          const int k = sdselect_ref(ntaken, rms, rval);
          if (k > 0) {
            std::memcpy(&xyzgen[(size_t)(k - 1) * rec], cdum, rec * sizeof(double));
            rms[k - 1] = rval; taken[k - 1] = c + 1;
          }
        }
      }
    }
  }
  *ntaken_out = ntaken;
  *rmsdsum = rmsdavg;
  counters[0] = ierr; counters[1] = ident; counters[2] = rcount; counters[3] = processed;
}

// Under torch.distributed.run OMP_NUM_THREADS is 1 in the environment: the timed CPU arm sets its thread
// count explicitly (bench.py).
void oracle_set_num_threads(int n) { if (n > 0) omp_set_num_threads(n); }

int oracle_num_threads(void) { return omp_get_max_threads(); }

}  // extern "C"
