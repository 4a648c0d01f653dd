// Small dense linear algebra for the init-time vertical-mode problem of the
// split-explicit scheme.  The reference calls its vendored
// EISPACK `rg` (balanc/elmhes/eltran/hqr2/balbak, Main/netlib/eispack.F90:12694-12804)
// and LINPACK `sgefa/sgedi` (Main/netlib/linpack.F90) from `vmodes`
// (Main/mod_vmodes.F90:358-378,613-647).  Eigenvector scale and order from
// EISPACK are removed afterwards by vorder/vnorml, so any correct real
// eigen-solver reproduces the tables to round-off.
// This header is a generic numerical routine, written from scratch:
//   * eigenvalues: Householder reduction to Hessenberg form, then the shifted
//     QR iteration with Givens rotations and deflation (real spectrum: `tau`
//     has real positive eigenvalues = equivalent depths; a complex pair is
//     reported through wi != 0),
//   * eigenvectors: inverse iteration on the ORIGINAL matrix with an LU
//     factorisation with partial pivoting,
//   * inverse: LU with partial pivoting.
// Matrices are column-major n x n: a[i + n*j] = A(i,j), 0-based.
#pragma once
#include <algorithm>
#include <cmath>
#include <limits>
#include <vector>

namespace rcm_la {

// LU factorisation with partial pivoting, in place.  Returns 0, or k+1 when
// the k-th pivot is exactly zero (LINPACK sgefa `info`).
inline int lu_factor(int n, std::vector<double>& a, std::vector<int>& piv) {
  piv.assign((size_t)n, 0);
  int info = 0;
  for (int k = 0; k < n; ++k) {
    int p = k;
    double amax = std::fabs(a[k + (size_t)n * k]);
    for (int i = k + 1; i < n; ++i)
      if (std::fabs(a[i + (size_t)n * k]) > amax) { amax = std::fabs(a[i + (size_t)n * k]); p = i; }
    piv[k] = p;
    if (amax == 0.0) { if (!info) info = k + 1; continue; }
    if (p != k)
      for (int j = 0; j < n; ++j) std::swap(a[k + (size_t)n * j], a[p + (size_t)n * j]);
    const double r = 1.0 / a[k + (size_t)n * k];
    for (int i = k + 1; i < n; ++i) a[i + (size_t)n * k] *= r;
    for (int j = k + 1; j < n; ++j) {
      const double t = a[k + (size_t)n * j];
      if (t != 0.0)
        for (int i = k + 1; i < n; ++i) a[i + (size_t)n * j] -= a[i + (size_t)n * k] * t;
    }
  }
  return info;
}

// solve LU x = b (b overwritten)
inline void lu_solve(int n, const std::vector<double>& lu, const std::vector<int>& piv, double* b) {
  // rows were interchanged across the whole matrix (LAPACK style): permute b first
  for (int k = 0; k < n; ++k)
    if (piv[k] != k) std::swap(b[k], b[piv[k]]);
  for (int k = 0; k < n; ++k) {
    const double t = b[k];
    if (t != 0.0)
      for (int i = k + 1; i < n; ++i) b[i] -= lu[i + (size_t)n * k] * t;
  }
  for (int k = n - 1; k >= 0; --k) {
    b[k] /= lu[k + (size_t)n * k];
    const double t = b[k];
    for (int i = 0; i < k; ++i) b[i] -= lu[i + (size_t)n * k] * t;
  }
}

// v = a^-1.  Returns sgefa-style info (0 = ok).
inline int invert(int n, const std::vector<double>& a, std::vector<double>& v) {
  std::vector<double> lu = a;
  std::vector<int> piv;
  const int info = lu_factor(n, lu, piv);
  v.assign((size_t)n * n, 0.0);
  if (info) return info;
  std::vector<double> col((size_t)n);
  for (int j = 0; j < n; ++j) {
    std::fill(col.begin(), col.end(), 0.0);
    col[j] = 1.0;
    lu_solve(n, lu, piv, col.data());
    for (int i = 0; i < n; ++i) v[i + (size_t)n * j] = col[i];
  }
  return 0;
}

// Eigenvalues of a real general matrix by Hessenberg reduction + shifted QR.
// wr/wi: real and imaginary parts (wi = 0 for a real spectrum).  Returns 0 or
// the (1-based) index of the eigenvalue that failed to converge (like `ierr`
// of EISPACK hqr).
inline int eigenvalues(int n, const std::vector<double>& a_in, std::vector<double>& wr, std::vector<double>& wi) {
  std::vector<double> h = a_in;
  wr.assign((size_t)n, 0.0); wi.assign((size_t)n, 0.0);
  auto H = [&](int i, int j) -> double& { return h[i + (size_t)n * j]; };
  // Householder reduction to upper Hessenberg form
  std::vector<double> v((size_t)n);
  for (int k = 0; k + 2 < n; ++k) {
    double alpha = 0.0;
    for (int i = k + 1; i < n; ++i) alpha += H(i, k) * H(i, k);
    alpha = std::sqrt(alpha);
    if (alpha == 0.0) continue;
    if (H(k + 1, k) > 0.0) alpha = -alpha;
    for (int i = 0; i < n; ++i) v[i] = 0.0;
    for (int i = k + 1; i < n; ++i) v[i] = H(i, k);
    v[k + 1] -= alpha;
    double vn = 0.0;
    for (int i = k + 1; i < n; ++i) vn += v[i] * v[i];
    if (vn == 0.0) continue;
    // H <- (I - 2vv'/v'v) H (I - 2vv'/v'v)
    for (int j = 0; j < n; ++j) {
      double s = 0.0;
      for (int i = k + 1; i < n; ++i) s += v[i] * H(i, j);
      s = 2.0 * s / vn;
      for (int i = k + 1; i < n; ++i) H(i, j) -= s * v[i];
    }
    for (int i = 0; i < n; ++i) {
      double s = 0.0;
      for (int j = k + 1; j < n; ++j) s += H(i, j) * v[j];
      s = 2.0 * s / vn;
      for (int j = k + 1; j < n; ++j) H(i, j) -= s * v[j];
    }
    for (int i = k + 2; i < n; ++i) H(i, k) = 0.0;
  }
  const double eps = std::numeric_limits<double>::epsilon();
  double hnorm = 0.0;
  for (int j = 0; j < n; ++j)
    for (int i = 0; i <= std::min(n - 1, j + 1); ++i) hnorm += std::fabs(H(i, j));
  int hi = n - 1, iter = 0;
  std::vector<double> cs((size_t)n), sn((size_t)n);
  while (hi >= 0) {
    // look for a negligible sub-diagonal element
    int lo = hi;
    while (lo > 0) {
      double s = std::fabs(H(lo - 1, lo - 1)) + std::fabs(H(lo, lo));
      if (s == 0.0) s = hnorm;
      if (std::fabs(H(lo, lo - 1)) <= eps * s) { H(lo, lo - 1) = 0.0; break; }
      --lo;
    }
    if (lo == hi) { wr[hi] = H(hi, hi); wi[hi] = 0.0; --hi; iter = 0; continue; }
    // trailing 2x2 block of the active window
    const double a = H(hi - 1, hi - 1), b = H(hi - 1, hi), c = H(hi, hi - 1), d = H(hi, hi);
    const double tr = a + d, det = a * d - b * c, disc = 0.25 * tr * tr - det;
    if (lo == hi - 1) {  // isolated 2x2 block: take its two eigenvalues
      if (disc >= 0.0) {
        const double sq = std::sqrt(disc);
        const double l1 = 0.5 * tr + (tr >= 0.0 ? sq : -sq);
        const double l2 = (l1 != 0.0) ? det / l1 : 0.5 * tr - (tr >= 0.0 ? sq : -sq);
        wr[hi - 1] = l1; wr[hi] = l2; wi[hi - 1] = wi[hi] = 0.0;
      } else {
        wr[hi - 1] = wr[hi] = 0.5 * tr; wi[hi - 1] = std::sqrt(-disc); wi[hi] = -wi[hi - 1];
      }
      hi -= 2; iter = 0; continue;
    }
    if (++iter > 60 * n) return hi + 1;
    // Wilkinson shift (eigenvalue of the trailing block closer to d); an
    // exceptional shift every 10 iterations
    double mu = d;
    if (disc >= 0.0) {
      const double sq = std::sqrt(disc);
      const double l1 = 0.5 * tr + sq, l2 = 0.5 * tr - sq;
      mu = (std::fabs(l1 - d) < std::fabs(l2 - d)) ? l1 : l2;
    }
    if (iter % 10 == 0) mu = d + 0.75 * std::fabs(c);
    // QR step on the window [lo..hi]: H - mu I = QR, H <- RQ + mu I
    for (int i = lo; i <= hi; ++i) H(i, i) -= mu;
    for (int k = lo; k < hi; ++k) {
      const double x = H(k, k), y = H(k + 1, k);
      const double r = std::hypot(x, y);
      double cc = 1.0, ss = 0.0;
      if (r != 0.0) { cc = x / r; ss = y / r; }
      cs[k] = cc; sn[k] = ss;
      for (int j = k; j <= hi; ++j) {
        const double t1 = H(k, j), t2 = H(k + 1, j);
        H(k, j) = cc * t1 + ss * t2;
        H(k + 1, j) = -ss * t1 + cc * t2;
      }
    }
    for (int k = lo; k < hi; ++k) {
      const double cc = cs[k], ss = sn[k];
      for (int i = lo; i <= std::min(hi, k + 1); ++i) {
        const double t1 = H(i, k), t2 = H(i, k + 1);
        H(i, k) = cc * t1 + ss * t2;
        H(i, k + 1) = -ss * t1 + cc * t2;
      }
    }
    for (int i = lo; i <= hi; ++i) H(i, i) += mu;
  }
  return 0;
}

// Real eigenproblem A z = w z for a matrix with a real spectrum (the role of
// EISPACK `rg` with matz = 1 in vmodes).  z(:,m) is the eigenvector of wr[m]
// (arbitrary scale).  Returns 0, or an hqr-style error index.
inline int eig_real(int n, const std::vector<double>& a, std::vector<double>& wr, std::vector<double>& wi,
                    std::vector<double>& z) {
  const int ier = eigenvalues(n, a, wr, wi);
  z.assign((size_t)n * n, 0.0);
  if (ier) return ier;
  double anorm = 0.0;
  for (double x : a) anorm = std::max(anorm, std::fabs(x));
  const double eps = std::numeric_limits<double>::epsilon();
  std::vector<double> lu, x((size_t)n);
  std::vector<int> piv;
  for (int m = 0; m < n; ++m) {
    // perturb the shift so that (A - s I) is not exactly singular
    const double s = wr[m] + 8.0 * eps * std::max(std::fabs(wr[m]), 1e-3 * anorm);
    lu = a;
    for (int i = 0; i < n; ++i) lu[i + (size_t)n * i] -= s;
    lu_factor(n, lu, piv);
    for (int i = 0; i < n; ++i)
      if (lu[i + (size_t)n * i] == 0.0) lu[i + (size_t)n * i] = eps * anorm;
    for (int i = 0; i < n; ++i) x[i] = 1.0 + 0.01 * i;
    for (int it = 0; it < 4; ++it) {
      lu_solve(n, lu, piv, x.data());
      double nm = 0.0;
      for (int i = 0; i < n; ++i) nm = std::max(nm, std::fabs(x[i]));
      if (nm == 0.0) break;
      for (int i = 0; i < n; ++i) x[i] /= nm;
    }
    for (int i = 0; i < n; ++i) z[i + (size_t)n * m] = x[i];
  }
  return 0;
}

}  // namespace rcm_la
