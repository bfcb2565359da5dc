// ORACLE — TEST INFRASTRUCTURE ONLY (see ls_rmsd_ref.hpp header).  PARITY UNPINNED.
//
// Restatement of /root/reference/src/ls_rmsd.f90:351-586 (svdcmp, the Golub-Reinsch
// SVD the reference uses for the shifted tridiagonal 4x4), specialised to m=n=4 but
// keeping the operation order, the "(abs(x)+anorm)==anorm" negligibility tests and
// the 30-sweep cap that makes rmsd() return -1 (ls_rmsd.f90:152-155, 527-531).
#include "ls_rmsd_ref.hpp"

namespace crest_oracle {

int svdcmp4(M4& a, double* w, M4& v, double* rv1) {
  const int m = 4, n = 4;
  int i, its, j, jj, k, l = 0, nm = 0;
  double anorm, c, f, g, h, s, scale, x, y, z;

  g = 0.0; scale = 0.0; anorm = 0.0;
  // Householder reduction to bidiagonal form (f90:365-434)
  for (i = 1; i <= n; ++i) {
    l = i + 1;
    rv1[i] = scale * g;
    g = 0.0; s = 0.0; scale = 0.0;
    if (i <= m) {
      for (k = i; k <= m; ++k) scale += std::fabs(a(k, i));
      if (scale != 0.0) {
        for (k = i; k <= m; ++k) { a(k, i) = a(k, i) / scale; s += a(k, i) * a(k, i); }
        f = a(i, i);
        g = -fsign(std::sqrt(s), f);
        h = f * g - s;
        a(i, i) = f - g;
        for (j = l; j <= n; ++j) {
          s = 0.0;
          for (k = i; k <= m; ++k) s += a(k, i) * a(k, j);
          f = s / h;
          for (k = i; k <= m; ++k) a(k, j) = a(k, j) + f * a(k, i);
        }
        for (k = i; k <= m; ++k) a(k, i) = scale * a(k, i);
      }
    }
    w[i] = scale * g;
    g = 0.0; s = 0.0; scale = 0.0;
    if (i <= m && i != n) {
      for (k = l; k <= n; ++k) scale += std::fabs(a(i, k));
      if (scale != 0.0) {
        for (k = l; k <= n; ++k) { a(i, k) = a(i, k) / scale; s += a(i, k) * a(i, k); }
        f = a(i, l);
        g = -fsign(std::sqrt(s), f);
        h = f * g - s;
        a(i, l) = f - g;
        for (k = l; k <= n; ++k) rv1[k] = a(i, k) / h;
        for (j = l; j <= m; ++j) {
          s = 0.0;
          for (k = l; k <= n; ++k) s += a(j, k) * a(i, k);
          for (k = l; k <= n; ++k) a(j, k) = a(j, k) + s * rv1[k];
        }
        for (k = l; k <= n; ++k) a(i, k) = scale * a(i, k);
      }
    }
    anorm = std::max(anorm, std::fabs(w[i]) + std::fabs(rv1[i]));
  }

  // accumulation of right-hand transformations (f90:436-460)
  for (i = n; i >= 1; --i) {
    if (i < n) {
      if (g != 0.0) {
        for (j = l; j <= n; ++j) v(j, i) = (a(i, j) / a(i, l)) / g;
        for (j = l; j <= n; ++j) {
          s = 0.0;
          for (k = l; k <= n; ++k) s += a(i, k) * v(k, j);
          for (k = l; k <= n; ++k) v(k, j) = v(k, j) + s * v(k, i);
        }
      }
      for (j = l; j <= n; ++j) { v(i, j) = 0.0; v(j, i) = 0.0; }
    }
    v(i, i) = 1.0;
    g = rv1[i];
    l = i;
  }

  // accumulation of left-hand transformations (f90:462-489)
  for (i = std::min(m, n); i >= 1; --i) {
    l = i + 1;
    g = w[i];
    for (j = l; j <= n; ++j) a(i, j) = 0.0;
    if (g != 0.0) {
      g = 1.0 / g;
      for (j = l; j <= n; ++j) {
        s = 0.0;
        for (k = l; k <= m; ++k) s += a(k, i) * a(k, j);
        f = (s / a(i, i)) * g;
        for (k = i; k <= m; ++k) a(k, j) = a(k, j) + f * a(k, i);
      }
      for (j = i; j <= m; ++j) a(j, i) = a(j, i) * g;
    } else {
      for (j = i; j <= m; ++j) a(j, i) = 0.0;
    }
    a(i, i) = a(i, i) + 1.0;
  }

  // diagonalisation of the bidiagonal form (f90:491-583)
  for (k = n; k >= 1; --k) {
    for (its = 1; its <= 30; ++its) {
      bool skip_cancel = false;   // "goto 2" straight from the split test
      for (l = k; l >= 1; --l) {
        nm = l - 1;
        if ((std::fabs(rv1[l]) + anorm) == anorm) { skip_cancel = true; break; }
        if ((std::fabs(w[nm]) + anorm) == anorm) break;   // rv1[1]==0 always, so nm>=1 here
      }
      if (!skip_cancel) {
        // label 1: cancellation of rv1(l) if l>1
        c = 0.0; s = 1.0;
        for (i = l; i <= k; ++i) {
          f = s * rv1[i];
          rv1[i] = c * rv1[i];
          if ((std::fabs(f) + anorm) == anorm) break;
          g = w[i];
          h = pythag(f, g);
          w[i] = h;
          h = 1.0 / h;
          c = (g * h);
          s = -(f * h);
          for (j = 1; j <= m; ++j) {
            y = a(j, nm); z = a(j, i);
            a(j, nm) = (y * c) + (z * s);
            a(j, i) = -(y * s) + (z * c);
          }
        }
      }
      // label 2
      z = w[k];
      if (l == k) {               // convergence
        if (z < 0.0) {
          w[k] = -z;
          for (j = 1; j <= n; ++j) v(j, k) = -v(j, k);
        }
        break;                    // goto 3
      }
      if (its == 30) return -1;
      x = w[l];
      nm = k - 1;
      y = w[nm];
      g = rv1[nm];
      h = rv1[k];
      f = ((y - z) * (y + z) + (g - h) * (g + h)) / (2.0 * h * y);
      g = pythag(f, 1.0);
      f = ((x - z) * (x + z) + h * ((y / (f + fsign(g, f))) - h)) / x;
      c = 1.0; s = 1.0;
      for (j = l; j <= nm; ++j) {
        i = j + 1;
        g = rv1[i];
        y = w[i];
        h = s * g;
        g = c * g;
        z = pythag(f, h);
        rv1[j] = z;
        c = f / z;
        s = h / z;
        f = (x * c) + (g * s);
        g = -(x * s) + (g * c);
        h = y * s;
        y = y * c;
        for (jj = 1; jj <= n; ++jj) {
          x = v(jj, j); z = v(jj, i);
          v(jj, j) = (x * c) + (z * s);
          v(jj, i) = -(x * s) + (z * c);
        }
        z = pythag(f, h);
        w[j] = z;
        if (z != 0.0) {
          z = 1.0 / z;
          c = f * z;
          s = h * z;
        }
        f = (c * g) + (s * y);
        x = -(s * g) + (c * y);
        for (jj = 1; jj <= m; ++jj) {
          y = a(jj, j); z = a(jj, i);
          a(jj, j) = (y * c) + (z * s);
          a(jj, i) = -(y * s) + (z * c);
        }
      }
      rv1[l] = 0.0;
      rv1[k] = f;
      w[k] = x;
    }
  }
  return 0;
}

}  // namespace crest_oracle
