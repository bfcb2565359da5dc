// ORACLE — TEST INFRASTRUCTURE ONLY (see ls_rmsd_ref.hpp header).  PARITY UNPINNED.
//
// Restatement of the principal-axis / rotational-constant routines
//   /root/reference/src/axis_module.f90:96-176    axis_0
//   /root/reference/src/axis_module.f90:249-293   axis_3 (axis_4 = axis_3 in place, 302-316)
//   /root/reference/src/axis_module.f90:372-378   calcxsum
//   /root/reference/src/axis_module.f90:384-412   CMAxyz
//   /root/reference/src/axis_module.f90:1051-1083 eigval3x3
//   /root/reference/src/axis_module.f90:1088-1266 eigvec3x3
//   /root/reference/src/atmasses.f90:27-56        ams (table in atmasses_ref.hpp)
#pragma once
#include <cmath>
#include <vector>
#include <limits>
#include <algorithm>
#include "atmasses_ref.hpp"

namespace crest_oracle {

constexpr double kIcm2MHz = 2.9979245e+4;   // axis_module.f90:25
constexpr double kAamu2icm = 16.8576522;    // axis_module.f90:34

// 1-based 3x3
struct M3 {
  double v[4][4];
  M3() { for (auto& r : v) for (auto& x : r) x = 0.0; }
  double& operator()(int i, int j) { return v[i][j]; }
  double operator()(int i, int j) const { return v[i][j]; }
};

// axis_module.f90:1051-1083 (trigonometric solution); twothirdpi = 8*atan(1)/3 (f90:37)
inline void eigval3x3(const M3& a, double w[4]) {
  const double twothirdpi = 8.0 * std::atan(1.0) / 3.0;
  double r = a(1, 2) * a(1, 2) + a(1, 3) * a(1, 3) + a(2, 3) * a(2, 3);
  const double q = (a(1, 1) + a(2, 2) + a(3, 3)) / 3.0;
  w[1] = a(1, 1) - q;
  w[2] = a(2, 2) - q;
  w[3] = a(3, 3) - q;
  const double p = std::sqrt((w[1] * w[1] + w[2] * w[2] + w[3] * w[3] + 2 * r) / 6.0);
  r = (w[1] * (w[2] * w[3] - a(2, 3) * a(2, 3))
       - a(1, 2) * (a(1, 2) * w[3] - a(2, 3) * a(1, 3))
       + a(1, 3) * (a(1, 2) * a(2, 3) - w[2] * a(1, 3))) / (p * p * p) * 0.5;
  if (r <= -1.0) r = 0.5 * twothirdpi;
  else if (r >= 1.0) r = 0.0;
  else r = std::acos(r) / 3.0;
  w[3] = q + 2 * p * std::cos(r);
  w[1] = q + 2 * p * std::cos(r + twothirdpi);
  w[2] = 3 * q - w[1] - w[3];
}

// axis_module.f90:1088-1266.  a is destroyed (as in the reference).
inline void eigvec3x3(M3& a, double w[4], M3& q) {
  const double eps = std::numeric_limits<double>::epsilon();
  double norm, n1, n2, n3, precon;
  int i;
  w[1] = std::max(std::fabs(a(1, 1)), std::fabs(a(1, 2)));
  w[2] = std::max(std::fabs(a(1, 3)), std::fabs(a(2, 2)));
  w[3] = std::max(std::fabs(a(2, 3)), std::fabs(a(3, 3)));
  precon = std::max(w[1], std::max(w[2], w[3]));
  if (precon < eps) {   // null matrix
    w[1] = w[2] = w[3] = 0.0;
    q = M3();
    q(1, 1) = q(2, 2) = q(3, 3) = 1.0;
    return;
  }
  norm = 1.0 / precon;
  a(1, 1) *= norm; a(1, 2) *= norm; a(2, 2) *= norm;
  a(1, 3) *= norm; a(2, 3) *= norm; a(3, 3) *= norm;

  eigval3x3(a, w);

  a(1, 1) -= w[1]; a(2, 2) -= w[1]; a(3, 3) -= w[1];
  q(1, 1) = a(1, 2) * a(2, 3) - a(1, 3) * a(2, 2);
  q(2, 1) = a(1, 3) * a(1, 2) - a(1, 1) * a(2, 3);
  q(3, 1) = a(1, 1) * a(2, 2) - a(1, 2) * a(1, 2);
  q(1, 2) = a(1, 2) * a(3, 3) - a(1, 3) * a(2, 3);
  q(2, 2) = a(1, 3) * a(1, 3) - a(1, 1) * a(3, 3);
  q(3, 2) = a(1, 1) * a(2, 3) - a(1, 2) * a(1, 3);
  q(1, 3) = a(2, 2) * a(3, 3) - a(2, 3) * a(2, 3);
  q(2, 3) = a(2, 3) * a(1, 3) - a(1, 2) * a(3, 3);
  q(3, 3) = a(1, 2) * a(2, 3) - a(2, 2) * a(1, 3);
  n1 = q(1, 1) * q(1, 1) + q(2, 1) * q(2, 1) + q(3, 1) * q(3, 1);
  n2 = q(1, 2) * q(1, 2) + q(2, 2) * q(2, 2) + q(3, 2) * q(3, 2);
  n3 = q(1, 3) * q(1, 3) + q(2, 3) * q(2, 3) + q(3, 3) * q(3, 3);

  // NB (f90:1150-1158): after picking column 2 the running maximum is reset to n1, not n2,
  // so column 3 wins whenever n3 > n1.  Reproduced as written.
  norm = n1; i = 1;
  if (n2 > norm) { i = 2; norm = n1; }
  if (n3 > norm) { i = 3; }

  if (i == 1) {
    norm = std::sqrt(1.0 / n1);
    q(1, 1) = q(1, 1) * norm; q(2, 1) = q(2, 1) * norm; q(3, 1) = q(3, 1) * norm;
  } else if (i == 2) {
    norm = std::sqrt(1.0 / n2);
    q(1, 1) = q(1, 2) * norm; q(2, 1) = q(2, 2) * norm; q(3, 1) = q(3, 2) * norm;
  } else {
    norm = std::sqrt(1.0 / n3);
    q(1, 1) = q(1, 3) * norm; q(2, 1) = q(2, 3) * norm; q(3, 1) = q(3, 3) * norm;
  }

  // right-handed orthonormal set (ev1, u, v)
  if (std::fabs(q(1, 1)) > std::fabs(q(2, 1))) {
    norm = std::sqrt(1.0 / (q(1, 1) * q(1, 1) + q(3, 1) * q(3, 1)));
    q(1, 2) = -q(3, 1) * norm; q(2, 2) = 0.0; q(3, 2) = +q(1, 1) * norm;
  } else {
    norm = std::sqrt(1.0 / (q(2, 1) * q(2, 1) + q(3, 1) * q(3, 1)));
    q(1, 2) = 0.0; q(2, 2) = +q(3, 1) * norm; q(3, 2) = -q(2, 1) * norm;
  }
  q(1, 3) = q(2, 1) * q(3, 2) - q(3, 1) * q(2, 2);
  q(2, 3) = q(3, 1) * q(1, 2) - q(1, 1) * q(3, 2);
  q(3, 3) = q(1, 1) * q(2, 2) - q(2, 1) * q(1, 2);

  a(1, 1) += w[1]; a(2, 2) += w[1]; a(3, 3) += w[1];

  // A*U
  n1 = a(1, 1) * q(1, 2) + a(1, 2) * q(2, 2) + a(1, 3) * q(3, 2);
  n2 = a(1, 2) * q(1, 2) + a(2, 2) * q(2, 2) + a(2, 3) * q(3, 2);
  n3 = a(1, 3) * q(1, 2) + a(2, 3) * q(2, 2) + a(3, 3) * q(3, 2);
  // A*V (out-of-order on purpose, f90:1200-1203)
  a(3, 3) = a(1, 3) * q(1, 3) + a(2, 3) * q(2, 3) + a(3, 3) * q(3, 3);
  a(1, 3) = a(1, 1) * q(1, 3) + a(1, 2) * q(2, 3) + a(1, 3) * q(3, 3);
  a(2, 3) = a(1, 2) * q(1, 3) + a(2, 2) * q(2, 3) + a(2, 3) * q(3, 3);
  // UT*(A*U) - l2*E ; UT*(A*V) ; VT*(A*V) - l2*E
  n1 = q(1, 2) * n1 + q(2, 2) * n2 + q(3, 2) * n3 - w[2];
  n2 = q(1, 2) * a(1, 3) + q(2, 2) * a(2, 3) + q(3, 2) * a(3, 3);
  n3 = q(1, 3) * a(1, 3) + q(2, 3) * a(2, 3) + q(3, 3) * a(3, 3) - w[2];

  if (std::fabs(n1) >= std::fabs(n3)) {
    norm = std::max(std::fabs(n1), std::fabs(n2));
    if (norm > eps) {
      if (std::fabs(n1) >= std::fabs(n2)) {
        n2 = n2 / n1; n1 = std::sqrt(1.0 / (1.0 + n2 * n2)); n2 = n2 * n1;
      } else {
        n1 = n1 / n2; n2 = std::sqrt(1.0 / (1.0 + n1 * n1)); n1 = n1 * n2;
      }
      q(1, 2) = n2 * q(1, 2) - n1 * q(1, 3);
      q(2, 2) = n2 * q(2, 2) - n1 * q(2, 3);
      q(3, 2) = n2 * q(3, 2) - n1 * q(3, 3);
    }
  } else {
    norm = std::max(std::fabs(n3), std::fabs(n2));
    if (norm > eps) {
      if (std::fabs(n3) >= std::fabs(n2)) {
        n2 = n2 / n3; n3 = std::sqrt(1.0 / (1.0 + n2 * n2)); n2 = n2 * n3;
      } else {
        n3 = n3 / n2; n2 = std::sqrt(1.0 / (1.0 + n3 * n3)); n3 = n3 * n2;
      }
      q(1, 2) = n3 * q(1, 2) - n2 * q(1, 3);
      q(2, 2) = n3 * q(2, 2) - n2 * q(2, 3);
      q(3, 2) = n3 * q(3, 2) - n2 * q(3, 3);
    }
  }
  q(1, 3) = q(2, 1) * q(3, 2) - q(3, 1) * q(2, 2);
  q(2, 3) = q(3, 1) * q(1, 2) - q(1, 1) * q(3, 2);
  q(3, 3) = q(1, 1) * q(2, 2) - q(2, 1) * q(1, 2);

  w[1] *= precon; w[2] *= precon; w[3] *= precon;
}

// axis_module.f90:384-412.  coord (3,nat) column-major -> x,y,z shifted to the mass centre.
inline void cma_xyz(int nat, const int* at, const double* coord, double* x, double* y, double* z) {
  double sumw = 1.0e-20, sumwx = 0.0, sumwy = 0.0, sumwz = 0.0;
  for (int i = 0; i < nat; ++i) {
    const double atmass = ams(at[i]);
    sumw = sumw + atmass;
    sumwx = sumwx + atmass * coord[3 * i + 0];
    sumwy = sumwy + atmass * coord[3 * i + 1];
    sumwz = sumwz + atmass * coord[3 * i + 2];
  }
  sumwx = sumwx / sumw; sumwy = sumwy / sumw; sumwz = sumwz / sumw;
  for (int i = 0; i < nat; ++i) {
    x[i] = coord[3 * i + 0] - sumwx;
    y[i] = coord[3 * i + 1] - sumwy;
    z[i] = coord[3 * i + 2] - sumwz;
  }
}

// axis_module.f90:96-176.  rot in MHz (A>=B>=C order follows ascending moments), evec(3,3).
inline void axis_0(int nat, const int* at, const double* coord, double rot[3], double& avmom,
                   M3& evec) {
  const double const1 = 1.66053;
  std::vector<double> x(nat, 0.0), y(nat, 0.0), z(nat, 0.0);
  cma_xyz(nat, at, coord, x.data(), y.data(), z.data());
  double t[7];
  for (int i = 1; i <= 6; ++i) t[i] = (double)i * 1.0e-10;
  for (int i = 0; i < nat; ++i) {
    const double atmass = ams(at[i]);
    t[1] = t[1] + atmass * (y[i] * y[i] + z[i] * z[i]);
    t[2] = t[2] - atmass * x[i] * y[i];
    t[3] = t[3] + atmass * (z[i] * z[i] + x[i] * x[i]);
    t[4] = t[4] - atmass * z[i] * x[i];
    t[5] = t[5] - atmass * y[i] * z[i];
    t[6] = t[6] + atmass * (x[i] * x[i] + y[i] * y[i]);
  }
  M3 a;
  a(1, 1) = t[1]; a(2, 1) = t[2]; a(1, 2) = t[2]; a(2, 2) = t[3];
  a(3, 1) = t[4]; a(1, 3) = t[4]; a(3, 2) = t[5]; a(2, 3) = t[5]; a(3, 3) = t[6];
  double eig[4] = {0, 0, 0, 0};
  evec = M3();
  eigvec3x3(a, eig, evec);
  double xyzmom[4];
  for (int i = 1; i <= 3; ++i) {
    for (int j = 1; j <= 3; ++j)
      if (std::fabs(evec(i, j)) < 1e-9) evec(i, j) = 0.0;
    if (eig[i] < 3.0e-4) { eig[i] = 0.0; rot[i - 1] = 0.0; }
    else rot[i - 1] = kIcm2MHz * kAamu2icm / eig[i];
    xyzmom[i] = eig[i] * const1;
  }
  avmom = 1.0e-47 * (xyzmom[1] + xyzmom[2] + xyzmom[3]) / 3.0;
}

// axis_module.f90:372-378
inline double calcxsum(const M3& e) {
  return e(1, 1) * (e(2, 2) * e(3, 3) - e(3, 2) * e(2, 3)) +
         e(1, 2) * (e(2, 3) * e(3, 1) - e(2, 1) * e(3, 3)) +
         e(1, 3) * (e(2, 1) * e(3, 2) - e(2, 2) * e(3, 1));
}

// axis_module.f90:249-293 + 302-316 (axis_4): shift to CMA and rotate into the
// principal-axis frame IN PLACE; chirality preserved by flipping evec(:,1) when det<0.
inline void axis_inplace(int nat, const int* at, double* coord, double rot[3]) {
  M3 evec;
  double avmom;
  axis_0(nat, at, coord, rot, avmom, evec);
  std::vector<double> x(nat), y(nat), z(nat);
  cma_xyz(nat, at, coord, x.data(), y.data(), z.data());
  double xsum = calcxsum(evec);
  if (xsum < 0.0) for (int j = 1; j <= 3; ++j) evec(j, 1) = -evec(j, 1);
  for (int i = 0; i < nat; ++i) {
    const double c[4] = {0.0, x[i], y[i], z[i]};
    for (int j = 1; j <= 3; ++j) {
      double s = 0.0;
      for (int k = 1; k <= 3; ++k) s = s + c[k] * evec(k, j);
      coord[3 * i + (j - 1)] = s;
    }
  }
}

}  // namespace crest_oracle
