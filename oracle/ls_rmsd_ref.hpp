// ORACLE — TEST INFRASTRUCTURE ONLY.  Not part of the product path.
//
// CPU restatement (C++17) of the reference's quaternion RMSD
//   /root/reference/src/ls_rmsd.f90:72-174   rmsd
//   /root/reference/src/ls_rmsd.f90:223-284  dstmev
//   /root/reference/src/ls_rmsd.f90:286-349  givens4
//   /root/reference/src/ls_rmsd.f90:351-586  svdcmp
//   /root/reference/src/ls_rmsd.f90:588-606  pythag
//
// PARITY UNPINNED: the reference ships no golden vector / known-answer test for
// this path and it is Fortran with no compiler in this image,
// so this restatement is validated only against independent numpy/LAPACK math
// (tests/test_oracle_*.py), never against the reference binary.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
// reference legs may use anything under oracle/.
#pragma once
#include <cmath>
#include <algorithm>

namespace crest_oracle {

// ls_rmsd.f90:588-606
inline double pythag(double a, double b) {
  const double absa = std::fabs(a), absb = std::fabs(b);
  if (absa > absb) {
    const double q = absb / absa;
    return absa * std::sqrt(1.0 + q * q);
  }
  if (absb == 0.0) return 0.0;
  const double q = absa / absb;
  return absb * std::sqrt(1.0 + q * q);
}

// Fortran SIGN(a,b): |a| carrying the sign of b.
inline double fsign(double a, double b) { return std::copysign(std::fabs(a), b); }

// 1-based 4x4 helper so the index arithmetic reads like the reference.
struct M4 {
  double v[5][5];
  M4() { for (auto& r : v) for (auto& x : r) x = 0.0; }
  double& operator()(int i, int j) { return v[i][j]; }
  double operator()(int i, int j) const { return v[i][j]; }
};

// Apply G=[[c,s],[-s,c]] from the left to rows (p,q), columns c0..c1:  T(p:q,c0:c1) = G*T(p:q,c0:c1)
inline void rot_rows(M4& T, int p, int q, int c0, int c1, double c, double s) {
  for (int j = c0; j <= c1; ++j) {
    const double a = T(p, j), b = T(q, j);
    T(p, j) = c * a + s * b;
    T(q, j) = -s * a + c * b;
  }
}
// T(p:q,p:q) = T(p:q,p:q) * G^T   (rows p,q; q = p+1)
inline void rot_cols_2x2(M4& T, int p, int q, double c, double s) {
  for (int i = p; i <= q; ++i) {
    const double a = T(i, p), b = T(i, q);
    T(i, p) = a * c + b * s;
    T(i, q) = a * (-s) + b * c;
  }
}

// ls_rmsd.f90:286-349 — three Givens rotations to tridiagonal form; V is the
// accumulated (transposed) similarity transform.
inline void givens4(const M4& S, M4& T, M4& V) {
  T = S;
  V = M4();
  double c1, s1, c2, s2, c3, s3;
  // zero T(4,1)/T(1,4): rotation in the 3-4 plane
  const double r1 = pythag(T(3, 1), T(4, 1));
  if (r1 != 0.0) {
    c1 = T(3, 1) / r1; s1 = T(4, 1) / r1;
    T(3, 1) = r1; T(4, 1) = 0.0;
    rot_rows(T, 3, 4, 2, 4, c1, s1);
    T(1, 3) = T(3, 1); T(1, 4) = T(4, 1); T(2, 3) = T(3, 2); T(2, 4) = T(4, 2);
    rot_cols_2x2(T, 3, 4, c1, s1);
  } else { c1 = 1.0; s1 = 0.0; }
  // zero T(3,1)/T(1,3): rotation in the 2-3 plane
  const double r2 = pythag(T(3, 1), T(2, 1));
  if (r2 != 0.0) {
    c2 = T(2, 1) / r2; s2 = T(3, 1) / r2;
    T(2, 1) = r2; T(3, 1) = 0.0;
    rot_rows(T, 2, 3, 2, 4, c2, s2);
    T(1, 2) = T(2, 1); T(1, 3) = T(3, 1); T(4, 2) = T(2, 4); T(4, 3) = T(3, 4);
    rot_cols_2x2(T, 2, 3, c2, s2);
  } else { c2 = 1.0; s2 = 0.0; }
  // zero T(4,2)/T(2,4): rotation in the 3-4 plane
  const double r3 = pythag(T(4, 2), T(3, 2));
  if (r3 != 0.0) {
    c3 = T(3, 2) / r3; s3 = T(4, 2) / r3;
    T(3, 2) = r3; T(4, 2) = 0.0;
    rot_rows(T, 3, 4, 3, 4, c3, s3);
    T(1, 3) = T(3, 1); T(1, 4) = T(4, 1); T(2, 3) = T(3, 2); T(2, 4) = T(4, 2);
    rot_cols_2x2(T, 3, 4, c3, s3);
  } else { c3 = 1.0; s3 = 0.0; }
  // net rotation (transposed), ls_rmsd.f90:341-346
  V(1, 1) = 1.0;
  for (int k = 2; k <= 4; ++k) { V(1, k) = 0.0; V(k, 1) = 0.0; }
  const double c1c2 = c1 * c2, s1c2 = s1 * c2;
  V(2, 2) = c2; V(3, 2) = c1 * s2; V(4, 2) = s1 * s2;
  V(2, 3) = -s2 * c3; V(3, 3) = c1c2 * c3 - s1 * s3; V(4, 3) = s1c2 * c3 + c1 * s3;
  V(2, 4) = s2 * s3; V(3, 4) = -c1c2 * s3 - s1 * c3; V(4, 4) = -s1c2 * s3 + c1 * c3;
}

int svdcmp4(M4& a, double* w /*1..4*/, M4& v, double* rv1 /*1..8*/);

// ls_rmsd.f90:223-284 — leading eigenvalue (and eigenvector) of a symmetric traceless 4x4.
// Returns io (0 ok, -1 SVD did not converge in 30 sweeps).
inline int dstmev(const M4& A, double& lambda, double evec[4]) {
  M4 T, V, SV;
  double SW[5] = {0, 0, 0, 0, 0}, rv1[9] = {0};
  givens4(A, T, V);
  lambda = std::min(std::min(T(1, 1) - std::fabs(T(1, 2)),
                             -std::fabs(T(2, 1)) + T(2, 2) - std::fabs(T(2, 3))),
                    std::min(-std::fabs(T(3, 2)) + T(3, 3) - std::fabs(T(3, 4)),
                             -std::fabs(T(4, 3)) + T(4, 4)));
  for (int i = 1; i <= 4; ++i) T(i, i) = T(i, i) - lambda;
  const int io = svdcmp4(T, SW, SV, rv1);
  if (io != 0) return io;
  int ml = 1;  // maxloc: first maximum
  for (int i = 2; i <= 4; ++i) if (SW[i] > SW[ml]) ml = i;
  lambda = SW[ml] + lambda;
  for (int i = 1; i <= 4; ++i) {
    double s = 0.0;
    for (int k = 1; k <= 4; ++k) s += V(i, k) * SV(k, ml);
    evec[i - 1] = s;
  }
  return 0;
}

}  // namespace crest_oracle
