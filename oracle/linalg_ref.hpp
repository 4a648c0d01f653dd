// ORACLE — TEST INFRASTRUCTURE ONLY.
// Dense linear algebra for the oracle's `vmodes` (Main/mod_vmodes.F90:358-378,613-647),
// standing in for the reference's vendored EISPACK `rg` and LINPACK `sgefa/sgedi`
// (Main/netlib/eispack.F90:12694-12804, Main/netlib/linpack.F90).
//
// Deliberately NOT the product's solver (regcm_b200/csrc/linalg_small.hpp): the two sides of a
// parity comparison must not share numerics.  The algorithms here are different on purpose:
//   * eigenvalues: shifted QR iteration on the FULL matrix (no Hessenberg reduction), the QR
//     factorisation by Householder reflections, Rayleigh-quotient shift, deflation of the last
//     row;
//   * eigenvectors: null vector of (A - lambda I) by Gaussian elimination with COMPLETE
//     pivoting and back substitution with the free unknown set to 1;
//   * inverse: Gauss-Jordan elimination with partial pivoting on [A | I].
// Eigenvector scale and order are removed by vorder/vnorml, so any correct solver gives the
// same tables to round-off; tests/test_hydro.py pins both solvers
// against LAPACK (numpy.linalg).
// Matrices are column-major n x n: a[i + n*j] = A(i,j), 0-based.
#pragma once
#include <algorithm>
#include <cmath>
#include <limits>
#include <vector>

namespace orc_la {

// v = a^-1 by Gauss-Jordan.  Returns 0, or k+1 when no pivot is found in column k.
inline int invert(int n, const std::vector<double>& a, std::vector<double>& v) {
  std::vector<double> m((size_t)n * 2 * n, 0.0);  // row-major [A | I]
  const int w = 2 * n;
  for (int i = 0; i < n; ++i) {
    for (int j = 0; j < n; ++j) m[(size_t)i * w + j] = a[i + (size_t)n * j];
    m[(size_t)i * w + n + i] = 1.0;
  }
  for (int k = 0; k < n; ++k) {
    int p = k;
    for (int i = k + 1; i < n; ++i)
      if (std::fabs(m[(size_t)i * w + k]) > std::fabs(m[(size_t)p * w + k])) p = i;
    if (m[(size_t)p * w + k] == 0.0) { v.assign((size_t)n * n, 0.0); return k + 1; }
    if (p != k)
      for (int j = 0; j < w; ++j) std::swap(m[(size_t)k * w + j], m[(size_t)p * w + j]);
    const double piv = m[(size_t)k * w + k];
    for (int j = 0; j < w; ++j) m[(size_t)k * w + j] /= piv;
    for (int i = 0; i < n; ++i) {
      if (i == k) continue;
      const double t = m[(size_t)i * w + k];
      if (t == 0.0) continue;
      for (int j = 0; j < w; ++j) m[(size_t)i * w + j] -= t * m[(size_t)k * w + j];
    }
  }
  v.assign((size_t)n * n, 0.0);
  for (int i = 0; i < n; ++i)
    for (int j = 0; j < n; ++j) v[i + (size_t)n * j] = m[(size_t)i * w + n + j];
  return 0;
}

// eigenvalues of a real matrix with a real spectrum: shifted QR on the full active block
inline int eigenvalues(int n, const std::vector<double>& a_in, std::vector<double>& wr, std::vector<double>& wi) {
  std::vector<double> A = a_in;
  wr.assign((size_t)n, 0.0); wi.assign((size_t)n, 0.0);
  auto a = [&](int i, int j) -> double& { return A[i + (size_t)n * j]; };
  const double eps = std::numeric_limits<double>::epsilon();
  double anorm = 0.0;
  for (double x : A) anorm += x * x;
  anorm = std::sqrt(anorm);
  std::vector<double> R((size_t)n * n), Q((size_t)n * n), u((size_t)n);
  int m = n, iter = 0;
  while (m > 0) {
    if (m == 1) { wr[0] = a(0, 0); break; }
    // deflate when the last row of the active block is negligible left of the diagonal
    double off = 0.0;
    for (int j = 0; j < m - 1; ++j) off += std::fabs(a(m - 1, j));
    if (off <= eps * (std::fabs(a(m - 1, m - 1)) + anorm * 1e-3) || off <= 1e-300) {
      wr[m - 1] = a(m - 1, m - 1); --m; iter = 0; continue;
    }
    if (++iter > 500) {
      // a complex pair shows up as a 2x2 block that never deflates
      const double p = a(m - 2, m - 2), q = a(m - 2, m - 1), r = a(m - 1, m - 2), s = a(m - 1, m - 1);
      const double tr = p + s, det = p * s - q * r, disc = 0.25 * tr * tr - det;
      double off2 = 0.0;
      for (int j = 0; j < m - 2; ++j) off2 += std::fabs(a(m - 1, j)) + std::fabs(a(m - 2, j));
      if (disc < 0.0 && off2 <= 1e-10 * anorm) {
        wr[m - 1] = wr[m - 2] = 0.5 * tr; wi[m - 2] = std::sqrt(-disc); wi[m - 1] = -wi[m - 2];
        m -= 2; iter = 0; continue;
      }
      return m;
    }
    // shift: eigenvalue of the trailing 2x2 closer to the corner element (real case), else the corner
    double mu = a(m - 1, m - 1);
    {
      const double p = a(m - 2, m - 2), q = a(m - 2, m - 1), r = a(m - 1, m - 2), s = a(m - 1, m - 1);
      const double tr = p + s, det = p * s - q * r, disc = 0.25 * tr * tr - det;
      if (disc >= 0.0) {
        const double sq = std::sqrt(disc), l1 = 0.5 * tr + sq, l2 = 0.5 * tr - sq;
        mu = (std::fabs(l1 - s) < std::fabs(l2 - s)) ? l1 : l2;
      }
      if (iter % 37 == 0) mu = s + 0.5 * std::fabs(r) + eps * anorm;  // exceptional shift
    }
    // (A_m - mu I) = Q R by Householder reflections, then A_m = R Q + mu I
    for (int j = 0; j < m; ++j)
      for (int i = 0; i < m; ++i) { R[i + (size_t)n * j] = a(i, j) - (i == j ? mu : 0.0); Q[i + (size_t)n * j] = (i == j) ? 1.0 : 0.0; }
    for (int k = 0; k < m - 1; ++k) {
      double nrm = 0.0;
      for (int i = k; i < m; ++i) nrm += R[i + (size_t)n * k] * R[i + (size_t)n * k];
      nrm = std::sqrt(nrm);
      if (nrm == 0.0) continue;
      const double alpha = (R[k + (size_t)n * k] > 0.0) ? -nrm : nrm;
      for (int i = 0; i < m; ++i) u[i] = (i < k) ? 0.0 : R[i + (size_t)n * k];
      u[k] -= alpha;
      double un = 0.0;
      for (int i = k; i < m; ++i) un += u[i] * u[i];
      if (un == 0.0) continue;
      for (int j = 0; j < m; ++j) {      // R <- H R
        double s = 0.0;
        for (int i = k; i < m; ++i) s += u[i] * R[i + (size_t)n * j];
        s = 2.0 * s / un;
        for (int i = k; i < m; ++i) R[i + (size_t)n * j] -= s * u[i];
      }
      for (int i = 0; i < m; ++i) {      // Q <- Q H
        double s = 0.0;
        for (int j = k; j < m; ++j) s += Q[i + (size_t)n * j] * u[j];
        s = 2.0 * s / un;
        for (int j = k; j < m; ++j) Q[i + (size_t)n * j] -= s * u[j];
      }
    }
    for (int j = 0; j < m; ++j)
      for (int i = 0; i < m; ++i) {
        double s = 0.0;
        for (int l = i; l < m; ++l) s += R[i + (size_t)n * l] * Q[l + (size_t)n * j];  // R is upper triangular
        a(i, j) = s + (i == j ? mu : 0.0);
      }
  }
  return 0;
}

// A z = w z for a real spectrum: z(:,q) belongs to wr[q], arbitrary scale.
inline int eig_real(int n, const std::vector<double>& a, std::vector<double>& wr, std::vector<double>& wi, std::vector<double>& z) {
  const int ier = eigenvalues(n, a, wr, wi);
  z.assign((size_t)n * n, 0.0);
  if (ier) return ier;
  std::vector<double> M((size_t)n * n), x((size_t)n), rhs((size_t)n);
  std::vector<int> cperm((size_t)n);
  for (int q = 0; q < n; ++q) {
    // eliminate (A - lambda I) with complete pivoting; the last pivot is ~0: its unknown is free
    for (int j = 0; j < n; ++j)
      for (int i = 0; i < n; ++i) M[i + (size_t)n * j] = a[i + (size_t)n * j] - (i == j ? wr[q] : 0.0);
    for (int j = 0; j < n; ++j) cperm[j] = j;
    for (int k = 0; k < n - 1; ++k) {
      int pi = k, pj = k; double best = -1.0;
      for (int j = k; j < n; ++j)
        for (int i = k; i < n; ++i)
          if (std::fabs(M[i + (size_t)n * j]) > best) { best = std::fabs(M[i + (size_t)n * j]); pi = i; pj = j; }
      if (best <= 0.0) break;
      if (pi != k) for (int j = 0; j < n; ++j) std::swap(M[k + (size_t)n * j], M[pi + (size_t)n * j]);
      if (pj != k) { for (int i = 0; i < n; ++i) std::swap(M[i + (size_t)n * k], M[i + (size_t)n * pj]); std::swap(cperm[k], cperm[pj]); }
      for (int i = k + 1; i < n; ++i) {
        const double t = M[i + (size_t)n * k] / M[k + (size_t)n * k];
        if (t == 0.0) continue;
        for (int j = k; j < n; ++j) M[i + (size_t)n * j] -= t * M[k + (size_t)n * j];
      }
    }
    // U y = 0 with y[n-1] = 1
    x[n - 1] = 1.0;
    for (int i = n - 2; i >= 0; --i) {
      double s = 0.0;
      for (int j = i + 1; j < n; ++j) s += M[i + (size_t)n * j] * x[j];
      x[i] = -s / M[i + (size_t)n * i];
    }
    for (int j = 0; j < n; ++j) z[cperm[j] + (size_t)n * q] = x[j];
  }
  return 0;
}

}  // namespace orc_la
