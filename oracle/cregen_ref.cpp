// ORACLE — TEST INFRASTRUCTURE ONLY (see ls_rmsd_ref.hpp header).  PARITY UNPINNED.
// cregen_CRE restated from /root/reference/src/cregen.f90:1125-1506 (see cregen_ref.hpp).
#include "cregen_ref.hpp"
#include <omp.h>

namespace crest_oracle {

namespace {
// cregen.f90:2056-2101 maskedxyz/maskedxyz2: gather atoms with mask>0
inline int masked_xyz(int n, const double* c, double* cm, const int* mask) {
  int k = 0;
  for (int i = 0; i < n; ++i)
    if (mask[i] > 0) {
      cm[3 * k + 0] = c[3 * i + 0]; cm[3 * k + 1] = c[3 * i + 1]; cm[3 * k + 2] = c[3 * i + 2];
      ++k;
    }
  return k;
}
}  // namespace

void cregen_cre(int nat, int nall, const int* at, double* xyz, const double* er_in,
                const CreParams& p, CreResult& out) {
  const double ethr = p.ethr, rthr = p.rthr, bthr = p.bthr;
  const bool enantio = p.enantio != 0;
  const bool substruc = p.substruc != 0;
  const bool heavy = p.heavy != 0;
  const size_t stride = (size_t)3 * nat;

  std::vector<double> er(nall + 1);
  for (int i = 1; i <= nall; ++i) er[i] = er_in[i - 1];
  out.rot.assign((size_t)3 * nall, 0.0);
  double* rot = out.rot.data();

  // RMSD atom selection (cregen.f90:1199-1202, 1293-1299)
  std::vector<int> mask(nat, 1);
  int nat0 = nat;
  std::vector<int> at0(at, at + nat);
  if (substruc) {
    nat0 = 0;
    for (int i = 0; i < nat; ++i) { mask[i] = p.include_rmsd[i] > 0 ? 1 : 0; }
    at0.clear();
    for (int i = 0; i < nat; ++i) if (mask[i]) { at0.push_back(at[i]); ++nat0; }
  }
  int nrmsd = nat0;
  if (!substruc && heavy) {
    nrmsd = 0;
    for (int i = 0; i < nat; ++i) { mask[i] = (at[i] != 1) ? 1 : 0; nrmsd += mask[i]; }
  }

  // K0: CMA + principal axes in place, rot. constants (cregen.f90:1223-1231)
  {
    std::vector<double> c1((size_t)3 * nat);
    for (int i = 0; i < nall; ++i) {
      double rdum[3];
      axis_inplace(nat, at, xyz + stride * i, rdum);
      const double* src = xyz + stride * i;
      if (substruc) masked_xyz(nat, src, c1.data(), mask.data());
      else std::memcpy(c1.data(), src, stride * sizeof(double));
      double avmom; M3 dum;
      axis_0(nat0, at0.data(), c1.data(), rot + 3 * i, avmom, dum);
    }
  }

  // band bookkeeping (cregen.f90:1240-1256)
  out.dup.assign(nall + 1, 0);
  int* dbl = out.dup.data();
  out.rmap1.assign(nall + 1, 0);
  out.rmap2.assign(nall + 1, 0);
  int64_t klong = 0;
  for (int i = 1; i <= nall; ++i) {
    out.rmap1[i] = klong;
    bool l1 = true;
    for (int j = 1; j <= i - 1; ++j) {
      const double de = (er[i] - er[j]) * kAutokcal;
      if (de < ethr) {
        klong = klong + 1;
        if (l1) { out.rmap2[i] = j; l1 = false; }
      }
    }
  }
  out.klong = klong;
  std::vector<float> rmat((size_t)klong, 0.0f);
  int nfail = 0;

  // K1/K2/K3 (cregen.f90:1266-1349): serial i, parallel j
  const bool use_mask = substruc || heavy;
  std::vector<double> c0((size_t)3 * nrmsd);
  for (int i = 1; i <= nall; ++i) {
    if (use_mask) masked_xyz(nat, xyz + stride * (i - 1), c0.data(), mask.data());
    else std::memcpy(c0.data(), xyz + stride * (i - 1), stride * sizeof(double));
    const int64_t o1 = out.rmap1[i];
    const int o2 = out.rmap2[i];
#pragma omp parallel reduction(+ : nfail)
    {
      std::vector<double> c1((size_t)3 * nrmsd), wx((size_t)3 * nrmsd), wy((size_t)3 * nrmsd);
#pragma omp for
      for (int j = 1; j <= i - 1; ++j) {
        const double de = (er[i] - er[j]) * kAutokcal;
        if (de < ethr) {
          if (use_mask) masked_xyz(nat, xyz + stride * (j - 1), c1.data(), mask.data());
          else std::memcpy(c1.data(), xyz + stride * (j - 1), stride * sizeof(double));
          const double rdum = rmsd(nrmsd, c0.data(), c1.data(), wx.data(), wy.data());
          double rdum2;
          if (enantio) {
            for (int k = 0; k < nrmsd; ++k) c1[3 * k] = -c1[3 * k];
            rdum2 = rmsd(nrmsd, c0.data(), c1.data(), wx.data(), wy.data());
          } else {
            rdum2 = rdum;
          }
          if (rdum < 0.0 || rdum2 < 0.0) nfail += 1;
          const int64_t kl = linr(o1, o2, j);
          rmat[(size_t)(kl - 1)] = (float)std::min(rdum, rdum2);
        }
      }
    }
  }
  out.nfail = nfail;

  // K4 duplicate sweep (cregen.f90:1353-1376)
  for (int i = 1; i <= nall; ++i) {
    for (int j = 1; j <= i - 1; ++j) {
      const double de = (er[i] - er[j]) * kAutokcal;
      if (!(de < ethr)) continue;
      const int64_t kl = linr(out.rmap1[i], out.rmap2[i], j);
      const double dr = (double)rmat[(size_t)(kl - 1)];
      if (p.boundary_tol >= 0.0 &&
          (std::fabs(dr - rthr) <= p.boundary_tol || std::fabs(dr - 2.0 * rthr) <= p.boundary_tol) &&
          (out.boundary_rows.empty() || out.boundary_rows.back() != i))
        out.boundary_rows.push_back(i);
      if (dr < rthr) {
        dbl[i] = j;
      } else if (dr < 2.0 * rthr) {
        if (equalrotaniso(i, j, rot, 0.5 * bthr, p.bthrmax, p.bthrshift)) dbl[i] = j;
      }
    }
    int k;
    backtrack(dbl, i, k);
  }
  if (p.keep_rmat) out.rmat = rmat;
  std::vector<float>().swap(rmat);

  // count + partition (cregen.f90:1385-1419)
  int ndup = 0;
  for (int i = 1; i <= nall; ++i) if (dbl[i] != 0) ++ndup;
  out.nkept = nall - ndup;
  int nallout = nall - ndup;
  if (p.nosort) nallout = nall;
  out.orderref.assign(nall + 1, 0);
  for (int i = 1; i <= nall; ++i) out.orderref[i] = i;
  if (!p.nosort) {
    int j = 0, l = nall + 1;
    for (int i = 1; i <= nall; ++i) {
      if (dbl[i] == 0) { j = j + 1; out.orderref[i] = j; }
      else { l = l - 1; out.orderref[i] = l; }
    }
    permute_records(xyz, stride, nall, out.orderref.data());
    permute_records(er.data() + 1, 1, nall, out.orderref.data());
    permute_records(rot, 3, nall, out.orderref.data());
  }
  out.er.assign(er.begin() + 1, er.end());

  // K5 enuc (cregen.f90:1422-1435)
  out.enuc.assign(nallout, 0.0);
  for (int k = 0; k < nallout; ++k) {
    const double* c1 = xyz + stride * k;
    double e = 0.0;
    for (int i = 0; i < nat - 1; ++i)
      for (int j = i + 1; j < nat; ++j) {
        const double dx = c1[3 * i] - c1[3 * j], dy = c1[3 * i + 1] - c1[3 * j + 1],
                     dz = c1[3 * i + 2] - c1[3 * j + 2];
        const double r = dx * dx + dy * dy + dz * dz + 1.0e-12;
        e = e + (double)(at[i] * at[j]) / r;
      }
    out.enuc[k] = e;
  }

  // K6 rotamer grouping (cregen.f90:1437-1456); `double` re-used
  std::vector<int> dbl2(nall + 1, 0);
  const double* enuc = out.enuc.data() - 1;
  for (int i = 1; i <= nallout; ++i) {
    for (int j = 1; j <= i - 1; ++j) {
      const double de = (er[i] - er[j]) * kAutokcal;
      const bool l3 = dbl2[j] == 0;
      if (!l3) continue;
      if (std::fabs(de) < ethr) {
        const bool l1 = equalrotaniso(i, j, rot, bthr, p.bthrmax, p.bthrshift);
        const bool l2 = 2.0 * std::fabs(enuc[i] - enuc[j]) / (enuc[i] + enuc[j]) < 1.0e-3;
        if (l1 && l2 && l3) {
          dbl2[i] = j;
          int k;
          backtrack(dbl2.data(), i, k);
          break;
        }
      }
    }
  }
  // group ids (cregen.f90:1458-1474)
  out.group.assign(nall + 1, 0);
  int k = 0;
  for (int i = 1; i <= nallout; ++i) {
    if (dbl2[i] == 0) { k = k + 1; out.group[i] = k; }
    else out.group[i] = out.group[dbl2[i]];
  }
  out.group[0] = k;
  if (p.nosort) nallout = k;
  out.nallout = nallout;
}

}  // namespace crest_oracle
