// ORACLE — TEST INFRASTRUCTURE ONLY (see ls_rmsd_ref.hpp header).  PARITY UNPINNED.
//
// Restatement of /root/reference/src/ls_rmsd.f90:72-174 (rmsd, option 0/1, no gradient)
// and /root/reference/src/ls_rmsd.f90:176-221 (rotation_matrix).
#pragma once
#include <vector>
#include "ls_rmsd_ref.hpp"

namespace crest_oracle {

struct RmsdOut {
  double error;        // the RMSD (or -1 on SVD non-convergence)
  double lambda;       // leading eigenvalue of the key matrix
  double xnorm, ynorm; // centred squared norms
  double R[3][3];      // covariance R(i,j)=sum_k x(i,k) y(j,k)
  double q[4];         // leading eigenvector (quaternion)
  double xc[3], yc[3]; // centroids
};

// coord1/coord2: (3,n) column-major, i.e. c[3*k+a] = coordinate a of atom k.
// scratch x,y: 3*n doubles each (row-major by axis: x[a*n+k]) so the summation order of the
// reference (sum over atoms for a fixed axis, ls_rmsd.f90:108-126) is reproduced.
inline void rmsd_full(int n, const double* coord1, const double* coord2, double* x, double* y,
                      RmsdOut& o) {
  double x_norm = 0.0, y_norm = 0.0;
  for (int a = 0; a < 3; ++a) {
    double* xi = x + (size_t)a * n;
    double* yi = y + (size_t)a * n;
    double sx = 0.0, sy = 0.0;
    for (int k = 0; k < n; ++k) { xi[k] = coord1[3 * k + a]; sx += xi[k]; }
    for (int k = 0; k < n; ++k) { yi[k] = coord2[3 * k + a]; sy += yi[k]; }
    o.xc[a] = sx / (double)n;
    o.yc[a] = sy / (double)n;
    for (int k = 0; k < n; ++k) xi[k] = xi[k] - o.xc[a];
    for (int k = 0; k < n; ++k) yi[k] = yi[k] - o.yc[a];
    double dx = 0.0, dy = 0.0;
    for (int k = 0; k < n; ++k) dx += xi[k] * xi[k];
    for (int k = 0; k < n; ++k) dy += yi[k] * yi[k];
    x_norm = x_norm + dx;
    y_norm = y_norm + dy;
  }
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) {
      const double* xi = x + (size_t)i * n;
      const double* yj = y + (size_t)j * n;
      double s = 0.0;
      for (int k = 0; k < n; ++k) s += xi[k] * yj[k];
      o.R[i][j] = s;
    }
  const auto& R = o.R;
  M4 S;
  S(1, 1) = R[0][0] + R[1][1] + R[2][2];
  S(2, 1) = R[1][2] - R[2][1];
  S(3, 1) = R[2][0] - R[0][2];
  S(4, 1) = R[0][1] - R[1][0];
  S(1, 2) = S(2, 1);
  S(2, 2) = R[0][0] - R[1][1] - R[2][2];
  S(3, 2) = R[0][1] + R[1][0];
  S(4, 2) = R[0][2] + R[2][0];
  S(1, 3) = S(3, 1);
  S(2, 3) = S(3, 2);
  S(3, 3) = -R[0][0] + R[1][1] - R[2][2];
  S(4, 3) = R[1][2] + R[2][1];
  S(1, 4) = S(4, 1);
  S(2, 4) = S(4, 2);
  S(3, 4) = S(4, 3);
  S(4, 4) = -R[0][0] - R[1][1] + R[2][2];
  o.xnorm = x_norm;
  o.ynorm = y_norm;
  double lambda = 0.0;
  const int io = dstmev(S, lambda, o.q);
  o.lambda = lambda;
  if (io != 0) { o.error = -1.0; return; }
  o.error = std::sqrt(std::max(0.0, ((x_norm + y_norm) - 2.0 * lambda)) / (double)n);
}

inline double rmsd(int n, const double* coord1, const double* coord2, double* x, double* y) {
  RmsdOut o;
  rmsd_full(n, coord1, coord2, x, y, o);
  return o.error;
}

// ls_rmsd.f90:176-221; U row-major U[i][j] = U(i+1,j+1)
inline void rotation_matrix(const double q[4], double U[3][3]) {
  const double q0 = q[0], q1 = q[1], q2 = q[2], q3 = q[3];
  const double b0 = 2.0 * q0, b1 = 2.0 * q1, b2 = 2.0 * q2, b3 = 2.0 * q3;
  const double q00 = b0 * q0 - 1.0, q01 = b0 * q1, q02 = b0 * q2, q03 = b0 * q3;
  const double q11 = b1 * q1, q12 = b1 * q2, q13 = b1 * q3;
  const double q22 = b2 * q2, q23 = b2 * q3, q33 = b3 * q3;
  U[0][0] = q00 + q11; U[0][1] = q12 - q03; U[0][2] = q13 + q02;
  U[1][0] = q12 + q03; U[1][1] = q00 + q22; U[1][2] = q23 - q01;
  U[2][0] = q13 - q02; U[2][1] = q23 + q01; U[2][2] = q00 + q33;
}

}  // namespace crest_oracle
