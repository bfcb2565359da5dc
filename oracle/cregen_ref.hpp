// ORACLE — TEST INFRASTRUCTURE ONLY (see ls_rmsd_ref.hpp header).  PARITY UNPINNED.
//
// Restatement of the CREGEN duplicate / rotamer sorting
//   /root/reference/src/cregen.f90:1125-1506           cregen_CRE
//   /root/reference/src/cregen.f90:1038-1121           cregen_esort
//   /root/reference/src/cregen.f90:2011-2101           xyzqsort, maskedxyz, maskedxyz2
//   /root/reference/src/rotcompare.f90:109-169         rotaniso, bthrerf, equalrotaniso
//   /root/reference/src/utilmod.f90:134-141            linr
//   /root/reference/src/legacy_algos/cregen_old.f90:1396-1410  backtrack
//   /root/reference/src/quicksort.f90:140-165,278-291  qsort, maskinvert
//   /root/reference/src/crest_pars.f90:23              autokcal
// Loop structure is kept (serial i, OpenMP `for` over j inside a per-i parallel region,
// two rmsd() calls per pair, float32 band storage) so it also serves as the timed CPU baseline.
#pragma once
#include <cstdint>
#include <cstring>
#include <vector>
#include <cmath>
#include "rmsd_ref.hpp"
#include "axis_ref.hpp"

namespace crest_oracle {

constexpr double kAutokcal = 627.50947428;  // crest_pars.f90:23

// rotcompare.f90:109-122 ; rot is (3,nall) column-major, i is 1-based
inline double rotaniso(int i, const double* rot) {
  const double a = rot[3 * (i - 1) + 0], b = rot[3 * (i - 1) + 1], c = rot[3 * (i - 1) + 2];
  const double av = (a + b + c) / 3.0;
  double r = std::sqrt((a - av) * (a - av) + (b - av) * (b - av) + (c - av) * (c - av));
  r = r / av;
  r = r / (3.0 * std::sqrt(2.0 / 3.0));
  return r;
}

// rotcompare.f90:127-142
inline double bthrerf(double bthr, double aniso, double bthrmax, double bthrshift) {
  const double c = ((bthrmax * 100.0) - (bthr * 100.0)) / 2.0;
  const double a = -std::erf(-2.5) * c + (bthr * 100.0);
  const double b = 4.0 / 0.8;
  const double d = bthrshift / 0.15;
  double thr = std::erf(aniso * b - d) * c + a;
  thr = thr / 100.0;
  return thr;
}

// rotcompare.f90:146-169 ; i,j 1-based
inline bool equalrotaniso(int i, int j, const double* rot, double bthr, double bthrmax,
                          double bthrshift) {
  const double anisoi = rotaniso(i, rot);
  const double anisoj = rotaniso(j, rot);
  const double av = (anisoi + anisoj) / 2.0;
  const double thr = bthrerf(bthr, av, bthrmax, bthrshift);
  const double* ri = rot + 3 * (i - 1);
  const double* rj = rot + 3 * (j - 1);
  const bool r1 = std::fabs((ri[0] / rj[0]) - 1.0) <= thr;
  const bool r2 = std::fabs((ri[1] / rj[1]) - 1.0) <= thr;
  const bool r3 = std::fabs((ri[2] / rj[2]) - 1.0) <= thr;
  return r1 && r2 && r3;
}

// utilmod.f90:134-141
inline int64_t linr(int64_t o1, int o2, int i) { return (o1 + 1) + i - o2; }

// cregen_old.f90:1396-1410 ; arr is 1-based (arr[0] unused)
inline void backtrack(int* arr, int i, int& val) {
  if (arr[i] > 0) {
    backtrack(arr, arr[i], val);
    arr[i] = val;
  } else {
    val = i;
  }
}

// quicksort.f90:140-165 ; 1-based a[], ind[]
inline void qsort_ref(double* a, int first, int last, int* ind) {
  const double x = a[(first + last) / 2];
  int i = first, j = last;
  for (;;) {
    while (a[i] < x) i = i + 1;
    while (x < a[j]) j = j - 1;
    if (i >= j) break;
    std::swap(a[i], a[j]);
    std::swap(ind[i], ind[j]);
    i = i + 1;
    j = j - 1;
  }
  if (first < i - 1) qsort_ref(a, first, i - 1, ind);
  if (j + 1 < last) qsort_ref(a, j + 1, last, ind);
}

// cregen.f90:1038-1121 (sorting part): returns `pos` (1-based): structure i moves to pos[i],
// energies sorted ascending in place, and nallout after the ewin cut.
inline int esort(int nall, double* energies /*0-based, sorted in place*/, int* pos /*0-based out*/,
                 double ewin) {
  std::vector<double> e(nall + 1);
  std::vector<int> orderref(nall + 1), inv(nall + 1);
  for (int i = 1; i <= nall; ++i) { e[i] = energies[i - 1]; orderref[i] = i; }
  if (nall > 0) qsort_ref(e.data(), 1, nall, orderref.data());
  for (int i = 1; i <= nall; ++i) inv[orderref[i]] = i;  // maskinvert, quicksort.f90:278-291
  for (int i = 1; i <= nall; ++i) { pos[i - 1] = inv[i]; energies[i - 1] = e[i]; }
  int nallout = nall;
  if (nall > 0) {
    double emax = e[1];
    for (int i = 2; i <= nall; ++i) emax = std::max(emax, e[i]);
    double de = (emax - e[1]) * kAutokcal;
    if (de > ewin) {
      nallout = 1;
      for (int i = 2; i <= nall; ++i) {
        de = (e[i] - e[1]) * kAutokcal;
        if (de < ewin) nallout = nallout + 1; else break;
      }
    }
  }
  return nallout;
}

struct CreParams {
  double ethr = 0.05, rthr = 0.125, bthr = 0.01, bthrmax = 0.025, bthrshift = 0.5;
  int enantio = 1;           // cregen.f90:1156,1196
  int heavy = 0;             // env%heavyrmsd
  int substruc = 0;          // nat /= rednat .and. subRMSD
  const int* include_rmsd = nullptr;  // mask(nat), >0 = atom enters RMSD (substruc only)
  int nosort = 0;
  int keep_rmat = 0;
  double boundary_tol = -1.0;  // >= 0: record rows with a band RMSD within tol of rthr / 2*rthr
};

struct CreResult {
  std::vector<double> rot;      // (3,nall) MHz, permuted like xyz unless nosort
  std::vector<double> er;       // energies, permuted
  std::vector<int> dup;         // stage-A double(1:nall) BEFORE the permutation (0 = kept)
  std::vector<int> orderref;    // new 1-based position of original structure i (identity if nosort)
  std::vector<int> group;       // group(0:nall)
  std::vector<double> enuc;     // enuc(1:nallout)
  std::vector<float> rmat;      // band storage (only if keep_rmat)
  std::vector<int64_t> rmap1;   // 1-based index arrays, [0] unused
  std::vector<int> rmap2;
  int64_t klong = 0;
  int nallout = 0;              // value returned to the caller
  int nkept = 0;                // nall - #duplicates
  int nfail = 0;                // number of rmsd() calls that returned -1
  std::vector<int> boundary_rows;  // 1-based rows i with |dr-rthr| or |dr-2rthr| <= boundary_tol (test listing)
};

// Apply "structure i goes to position pos[i]" to an array of records (what the reference's
// xyzqsort/maskqsort/matqsort/stringqsort achieve, whatever the swap order).
template <class T>
inline void permute_records(T* data, size_t rec, int nall, const int* pos1) {
  std::vector<T> tmp((size_t)nall * rec);
  for (int i = 0; i < nall; ++i)
    std::memcpy(&tmp[(size_t)(pos1[i + 1] - 1) * rec], &data[(size_t)i * rec], rec * sizeof(T));
  std::memcpy(data, tmp.data(), tmp.size() * sizeof(T));
}

// cregen.f90:1125-1506.  xyz (3,nat,nall) column-major, Angstrom, modified in place.
// er_in: energies (Hartree) per structure, ascending (output of esort).
void cregen_cre(int nat, int nall, const int* at, double* xyz, const double* er_in,
                const CreParams& p, CreResult& out);

}  // namespace crest_oracle
