// Host-side set-up of the split-explicit scheme ("vmodes on
// host"): the vertical-mode problem of `vmodes` (Main/mod_vmodes.F90:86-594) and
// the table part of `spinit` (Main/mod_split.F90:75-175), evaluated once at
// rcm_dyn_init in double precision with the reference's order of operations.
// EISPACK `rg` / LINPACK `sgefa,sgedi` are replaced by linalg_small.hpp.
#include <algorithm>
#include <cmath>
#include <limits>
#include <vector>

#include "consts.h"
#include "linalg_small.hpp"
#include "split_host.h"

using namespace rcm;

namespace {
constexpr double stdpcb = 101.3250;  // Share/mod_constants.F90:232
constexpr double stdt = 288.15;      // :233
constexpr double lrate = 0.00649;    // :235

struct Mat {  // column-major n x n, 1-based
  int n; std::vector<double> d;
  explicit Mat(int n_ = 0) : n(n_), d((size_t)n_ * n_, 0.0) {}
  double& operator()(int i, int j) { return d[(size_t)(i - 1) + (size_t)n * (j - 1)]; }
  double operator()(int i, int j) const { return d[(size_t)(i - 1) + (size_t)n * (j - 1)]; }
};
Mat mul(const Mat& a, const Mat& b) {  // Fortran MATMUL
  Mat c(a.n);
  for (int j = 1; j <= a.n; ++j)
    for (int i = 1; i <= a.n; ++i) {
      double s = 0.0;
      for (int l = 1; l <= a.n; ++l) s += a(i, l) * b(l, j);
      c(i, j) = s;
    }
  return c;
}
Mat diff(const Mat& a, const Mat& b) { Mat c(a.n); for (size_t q = 0; q < c.d.size(); ++q) c.d[q] = a.d[q] - b.d[q]; return c; }
}  // namespace

int rcm_split_tables(int kz, int nsplit, int lstand, double ptop, double dtsec, const std::vector<double>& sigma /* [1..kz+1] */,
                     double xps_in, const std::vector<double>& tbarh_in /* [1..kz], lstand = 0 */, SplitTables& T) {
  const int kzp1 = kz + 1, n = kz;
  T.kz = kz; T.nsplit = nsplit; T.numerr = 0;
  std::vector<double> sigmah(kzp1 + 1, 0.0), sdsigma(kz + 1, 0.0), tbarh(kz + 1, 0.0), tbarf(kzp1 + 1, 0.0), tweigh(kz + 1, 0.0),
      hweigh(kz + 1, 0.0), dsigma(kz + 1, 0.0);
  double xps = lstand ? stdpcb : xps_in;
  const double pd = xps - ptop;
  for (int k = 1; k <= kz; ++k) {
    sigmah[k] = (sigma[k] + sigma[k + 1]) * 0.5;
    sdsigma[k] = sigma[k + 1] - sigma[k];
    dsigma[k] = (sigma[k + 1] - sigma[k]);
  }
  sigmah[kzp1] = 1.0;
  if (lstand) {  // vtlaps
    const double tstrat = 218.15, zstrat = 10769.0;
    const double p0 = stdpcb, fac = rgas * lrate * regrav;
    for (int k = 1; k <= kz; ++k) {
      const double p = sigmah[k] * pd + ptop;
      tbarh[k] = stdt * std::pow(p / p0, fac);
      const double z = (stdt - tbarh[k]) / lrate;
      if (z > zstrat) tbarh[k] = tstrat;
    }
  } else {
    for (int k = 1; k <= kz; ++k) tbarh[k] = tbarh_in[k];
  }
  {  // vchekt
    bool lstab = true;
    for (int k = 1; k <= kz - 1; ++k) {
      const double ds1 = sigma[k + 1] - sigma[k], ds2 = sigma[k + 2] - sigma[k + 1];
      const double tb = (ds1 * tbarh[k] + ds2 * tbarh[k + 1]) / (ds1 + ds2);
      const double g1 = rovcp * tb / (sigma[k + 1] + ptop / pd);
      const double g2 = (tbarh[k + 1] - tbarh[k]) / (sigmah[k + 1] - sigmah[k]);
      if (g1 - g2 < 0.0) lstab = false;
    }
    if (!lstab) T.numerr++;
  }
  for (int k = 2; k <= kz; ++k) {
    const int k1 = k - 1;
    tbarf[k] = tbarh[k1] * (sigmah[k] - sigma[k]) / (sigmah[k] - sigmah[k1]) + tbarh[k] * (sigma[k] - sigmah[k1]) / (sigmah[k] - sigmah[k1]);
  }
  tbarf[1] = 0.0; tbarf[kzp1] = 0.0;
  Mat a1(n), a2(n), a3(n), a4(n), d1(n), d2(n), e1(n), e2(n), e3(n), g1(n), g2(n), s1(n), s2(n), w1(n), w2(n), x1(n), a0(n);
  for (int l = 1; l <= kz; ++l)
    for (int k = 1; k <= kz; ++k) {
      if (l > k) e2(k, l) = 0.0;
      if (l <= k) e2(k, l) = 1.0;
      e1(k, l) = 1.0;
    }
  for (int k = 1; k <= kz; ++k) {
    d1(k, k) = sigma[k + 1] - sigma[k];
    a3(k, k) = -tbarh[k];
    d2(k, k) = rovcp * tbarh[k] / (sigmah[k] + ptop / pd);
    s1(k, k) = sigma[k];
    s2(k, k) = sigmah[k];
    x1(k, k) = 1.0;
  }
  for (int k = 1; k <= kz; ++k) {
    e3(k, k) = 1.0;
    if (k > 1) g1(k, k) = tbarf[k];
    if (k < kz) g1(k, k + 1) = -tbarf[k + 1];
    if (k < kz) e3(k, k + 1) = 1.0;
  }
  w1 = diff(e2, x1);
  w2 = mul(w1, d1);
  g2 = mul(e1, d1);
  w1 = mul(s1, g2);
  g2 = diff(w1, w2);
  w2 = Mat(n);
  for (int k = 1; k <= kz; ++k) w2(k, k) = 1.0 / d1(k, k);
  w1 = mul(g1, g2);
  a1 = mul(w2, w1);
  w1 = mul(e1, d1);
  a2 = mul(s2, w1);
  w2 = mul(e3, g2);
  for (double& x : w2.d) x = x * 0.5;
  w1 = diff(w2, a2);
  a2 = mul(d2, w1);
  w1 = mul(e1, d1);
  a4 = mul(a3, w1);
  for (double& x : a4.d) x = -x;
  for (int j = 1; j <= kz; ++j)
    for (int i = 1; i <= kz; ++i) a0(i, j) = a1(i, j) + a2(i, j) + a3(i, j) + a4(i, j);
  // hydros, hydroc
  Mat hydros(n);
  std::vector<double> hydroc((size_t)kz * kzp1, 0.0);
  auto HC = [&](int k, int l) -> double& { return hydroc[(size_t)(k - 1) + (size_t)kz * (l - 1)]; };
  for (int k = 2; k <= kz; ++k) w1(k, 1) = std::log((sigmah[k] + ptop / pd) / (sigmah[k - 1] + ptop / pd));
  for (int k = 1; k <= kz - 1; ++k)
    for (int l = k; l <= kz - 1; ++l) {
      hydros(k, l) = hydros(k, l) + w1(l + 1, 1) * sdsigma[l] / (sdsigma[l + 1] + sdsigma[l]);
      hydros(k, l + 1) = hydros(k, l + 1) + w1(l + 1, 1) * sdsigma[l + 1] / (sdsigma[l + 1] + sdsigma[l]);
    }
  for (int k = 1; k <= kz; ++k) hydros(k, kz) = hydros(k, kz) + std::log((1.0 + ptop / pd) / (sigmah[kz] + ptop / pd));
  tweigh[1] = 0.0;
  for (int l = 2; l <= kz; ++l) tweigh[l] = (tbarh[l] * sdsigma[l] + tbarh[l - 1] * sdsigma[l - 1]) / (sdsigma[l] + sdsigma[l - 1]);
  for (int l = 2; l <= kz - 1; ++l)
    for (int k = 1; k <= l - 1; ++k) HC(k, l) = tweigh[l] - tweigh[l + 1];
  for (int l = 1; l <= kz - 1; ++l) HC(l, l) = tbarh[l] - tweigh[l + 1];
  for (int k = 1; k <= kz - 1; ++k) HC(k, kz) = tweigh[kz] - tbarh[kz];
  for (int k = 1; k <= kz; ++k) HC(k, kzp1) = tbarh[kz];
  {  // self-test of the linearised hydrostatic equation
    bool bad = false;
    for (int k = 1; k <= kz; ++k) {
      double t1 = 0.0;
      for (int l = 1; l <= kz; ++l) t1 = t1 + hydros(k, l) * tbarh[l];
      double t2 = -tbarh[k] * std::log(sigmah[k] * pd + ptop);
      for (int l = 1; l <= kzp1; ++l) t2 = t2 + HC(k, l) * std::log(sigmah[l] * pd + ptop);
      const double x = std::fabs(t1 - t2) / (std::fabs(t1) + std::fabs(t2));
      if (x > 1000.0 * std::numeric_limits<double>::epsilon()) bad = true;
    }
    if (bad) T.numerr++;
  }
  // tau
  Mat tau(n);
  {
    std::vector<double> w3((size_t)(kzp1 + 1) * (kz + 1), 0.0);
    auto W3 = [&](int k, int l) -> double& { return w3[(size_t)k + (size_t)(kzp1 + 1) * l]; };
    for (int l = 1; l <= kz; ++l)
      for (int k = 1; k <= kzp1; ++k) W3(k, l) = sdsigma[l] / (1.0 + ptop / (pd * sigmah[k]));
    for (int l = 1; l <= kz; ++l)
      for (int k = 1; k <= kz; ++k) {
        w2(k, l) = 0.0;
        for (int mm = 1; mm <= kzp1; ++mm) w2(k, l) = w2(k, l) + HC(k, mm) * W3(mm, l);
      }
    w1 = mul(hydros, a0);
    for (int j = 1; j <= kz; ++j)
      for (int i = 1; i <= kz; ++i) {
        tau(i, j) = w1(i, j) - w2(i, j);
        tau(i, j) = -rgas * tau(i, j);
      }
  }
  T.tau0 = tau.d;
  // eigenproblem: rg -> vcheke -> vorder -> vnorml
  std::vector<double> wr, wi, z;
  if (rcm_la::eig_real(kz, tau.d, wr, wi, z)) T.numerr++;
  {
    int numneg = 0, nimag = 0; double emax = 0.0;
    for (int q = 0; q < kz; ++q) { if (wr[q] <= 0.0) numneg++; if (wr[q] > emax) emax = wr[q]; }
    for (int q = 0; q < kz; ++q) if (wi[q] / emax > std::numeric_limits<double>::epsilon()) nimag++;
    if (numneg + nimag != 0) T.numerr++;
  }
  Mat zmatx(n);
  T.hbar.assign(kz + 1, 0.0);
  {
    std::vector<int> used(kz, 0);
    int kmax = 0;
    for (int l = 1; l <= kz; ++l) {
      double hmax = -std::numeric_limits<double>::epsilon();
      for (int k = 0; k < kz; ++k)
        if (!used[k] && wr[k] > hmax) { hmax = wr[k]; kmax = k; }
      T.hbar[l] = hmax;
      used[kmax] = 1;
      for (int k = 1; k <= kz; ++k) zmatx(k, l) = z[(size_t)(k - 1) + (size_t)kz * kmax];
    }
  }
  {
    int kmax = 1;
    for (int l = 1; l <= kz; ++l) {
      double zmax = -1.0, v = 0.0;
      for (int k = 1; k <= kz; ++k) {
        const double a = std::fabs(zmatx(k, l));
        if (a > zmax) { zmax = a; kmax = k; }
        v = (sigma[k + 1] - sigma[k]) * a * a + v;
      }
      const double a = (zmatx(kmax, l) / zmax) / std::sqrt(v);
      for (int k = 1; k <= kz; ++k) zmatx(k, l) = a * zmatx(k, l);
    }
  }
  Mat zmatxr(n);
  if (rcm_la::invert(kz, zmatx.d, zmatxr.d)) T.numerr++;
  std::vector<double> hydror;
  if (rcm_la::invert(kz, hydros.d, hydror)) T.numerr++;
  auto HR = [&](int i, int j) { return hydror[(size_t)(i - 1) + (size_t)kz * (j - 1)]; };
  hweigh[kz] = 1.0;
  for (int k1 = 1; k1 <= kz; ++k1)
    for (int k2 = 1; k2 <= kz; ++k2) {
      w1(k2, k1) = 0.0;
      for (int k = 1; k <= kz; ++k) w1(k2, k1) = HR(k, k2) * HR(k, k1) * hweigh[k] / (tbarh[k] * tbarh[k]) + w1(k2, k1);
    }
  const double ps2 = xps * xps;
  std::vector<double> varpa1((size_t)kz * kzp1, 0.0);
  auto VP = [&](int k2, int k1) -> double& { return varpa1[(size_t)(k2 - 1) + (size_t)kz * (k1 - 1)]; };
  for (int k1 = 1; k1 <= kzp1; ++k1)
    for (int k2 = 1; k2 <= kz; ++k2) {
      VP(k2, k1) = 0.0;
      for (int k = 1; k <= kz; ++k) VP(k2, k1) = VP(k2, k1) + w1(k2, k) * HC(k, k1) * ps2;
    }
  for (int k1 = 1; k1 <= kz; ++k1)
    for (int k2 = 1; k2 <= kz; ++k2) a0(k2, k1) = a0(k2, k1) - a4(k2, k1);

  // ---- spinit, table part (Main/mod_split.F90:88-91,137-175; dtau: Main/mod_params.F90:1702-1707) ----
  T.dtau.assign(nsplit + 1, 0.0); T.aam.assign(nsplit + 1, 0.0); T.an.assign(nsplit + 1, 0.0);
  T.am.assign((size_t)kz * nsplit, 0.0);
  auto AM = [&](int k, int l) -> double& { return T.am[(size_t)(k - 1) + (size_t)kz * (l - 1)]; };
  for (int ns = 1; ns <= nsplit; ++ns) T.dtau[ns] = dtsec * (0.5 / (double)(nsplit - ns + 1));
  for (int ns = 1; ns <= nsplit; ++ns) T.aam[ns] = (double)std::lround(dtsec / T.dtau[ns]);
  for (int m = 1; m <= nsplit; ++m) {
    T.an[m] = 0.0;
    for (int l = 1; l <= kz; ++l) T.an[m] = T.an[m] + dsigma[l] * zmatx(l, m);
    for (int k = 1; k <= kz; ++k) { AM(k, m) = 0.0; tau(m, k) = 0.0; }
    for (int l = 1; l <= kz; ++l)
      for (int k = 1; k <= kz; ++k) {
        AM(k, m) = AM(k, m) + a0(k, l) * zmatx(l, m);
        tau(m, k) = tau(m, k) + rgas * zmatxr(m, l) * hydros(l, k);
      }
    for (int k = 1; k <= kzp1; ++k) VP(m, k) = 0.0;
    for (int l = 1; l <= kz; ++l)
      for (int k = 1; k <= kzp1; ++k) VP(m, k) = VP(m, k) + rgas * zmatxr(m, l) * HC(l, k);
  }
  for (int l = 1; l <= nsplit; ++l) {
    const double fac = 2.0 * dtsec / (2.0 * T.aam[l] + 1.0);
    T.an[l] = T.an[l] * fac;
    for (int k = 1; k <= kz; ++k) {
      zmatx(k, l) = zmatx(k, l) * fac;
      AM(k, l) = AM(k, l) * fac;
    }
  }
  T.zmatx = zmatx.d; T.zmatxr = zmatxr.d; T.tau = tau.d; T.varpa1 = varpa1;
  T.sigmah = sigmah; T.pd = pd; T.xps = xps;
  // per (l,k) scalars of the modal geopotential (:215-216,225-226)
  T.pdlog.assign((size_t)nsplit * kzp1 + 1, 0.0); T.eps1.assign((size_t)nsplit * kzp1 + 1, 0.0);
  for (int l = 1; l <= nsplit; ++l)
    for (int k = 1; k <= kzp1; ++k) {
      T.pdlog[l + (size_t)nsplit * (k - 1)] = VP(l, k) * std::log(sigmah[k] * pd + ptop);
      T.eps1[l + (size_t)nsplit * (k - 1)] = VP(l, k) * sigmah[k] / (sigmah[k] * pd + ptop);
    }
  return T.numerr;
}
