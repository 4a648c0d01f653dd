// ORACLE — TEST INFRASTRUCTURE ONLY.
// Split-explicit scheme of the hydrostatic core, restated loop for loop:
//   vmodes (+ vtlaps, vchekt, vcheke, vorder, vnorml, invmtrx)  Main/mod_vmodes.F90:86-647
//   spinit   Main/mod_split.F90:75-239      splitf  :243-461      spstep  :463-669
// The eigen-solver and the matrix inverse stand in for the reference's vendored
// EISPACK `rg` and LINPACK `sgefa/sgedi` (Main/netlib): oracle/linalg_ref.hpp, a solver of
// its own (NOT the product's regcm_b200/csrc/linalg_small.hpp), pinned against LAPACK in
// tests/test_hydro.py::test_vmodes_against_lapack.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <limits>

#include "linalg_ref.hpp"
#include "model.hpp"

namespace orc {

namespace {
constexpr double stdpcb = 101.3250;  // Share/mod_constants.F90:232
constexpr double stdt = 288.15;      // :233
constexpr double lrate = 0.00649;    // :235

// column-major kz x kz helper with 1-based (i,j)
struct M {
  int n; std::vector<double> d;
  explicit M(int n_) : n(n_), d((size_t)n_ * n_, 0.0) {}
  double& operator()(int i, int j) { return d[(size_t)(i - 1) + (size_t)n * (j - 1)]; }
  double operator()(int i, int j) const { return d[(size_t)(i - 1) + (size_t)n * (j - 1)]; }
};
// Fortran MATMUL (the reference's intrinsic; summation order over the inner
// index ascending)
M matmul(const M& a, const M& b) {
  M c(a.n);
  for (int j = 1; j <= a.n; ++j)
    for (int i = 1; i <= a.n; ++i) {
      double s = 0.0;
      for (int l = 1; l <= a.n; ++l) s += a(i, l) * b(l, j);
      c(i, j) = s;
    }
  return c;
}
M sub(const M& a, const M& b) { M c(a.n); for (size_t q = 0; q < c.d.size(); ++q) c.d[q] = a.d[q] - b.d[q]; return c; }
}  // namespace

void Oracle::vmodes() {
  const int n = kz;
  M a1(n), a2(n), a3(n), a4(n), d1(n), d2(n), e1(n), e2(n), e3(n), g1(n), g2(n), s1(n), s2(n), w1(n), w2(n), x1(n);
  std::vector<double> tbarf(kzp1 + 1, 0.0), thetaf(kzp1 + 1, 0.0), tweigh(kz + 1, 0.0), sdsigma(kz + 1, 0.0),
      hweigh(kz + 1, 0.0), cpfac(kz + 1, 0.0);
  vm_numerr = 0;
  if (cfg.lstand) xps = stdpcb;
  pd = xps - ptop;
  for (int k = 1; k <= kz; ++k) {
    sigmah(k) = (sigma(k) + sigma(k + 1)) * d_half;
    sdsigma[k] = sigma(k + 1) - sigma(k);
  }
  sigmah(kzp1) = d_one;
  if (cfg.lstand) {  // vtlaps :494-508
    const double tstrat = 218.15, zstrat = 10769.0;
    const double p0 = stdpcb, fac = rgas * lrate * regrav;
    for (int k = 1; k <= kz; ++k) {
      double p = sigmah(k) * pd + ptop;
      tbarh(k) = stdt * std::pow(p / p0, fac);
      double z = (stdt - tbarh(k)) / lrate;
      if (z > zstrat) tbarh(k) = tstrat;
    }
  }
  {  // vchekt :471-489
    bool lstab = true;
    for (int k = 1; k <= kz - 1; ++k) {
      double ds1 = sigma(k + 1) - sigma(k), ds2 = sigma(k + 2) - sigma(k + 1);
      double tb = (ds1 * tbarh(k) + ds2 * tbarh(k + 1)) / (ds1 + ds2);
      double g1_ = rovcp * tb / (sigma(k + 1) + ptop / pd);
      double g2_ = (tbarh(k + 1) - tbarh(k)) / (sigmah(k + 1) - sigmah(k));
      if (g1_ - g2_ < d_zero) lstab = false;
    }
    if (!lstab) vm_numerr++;
  }
  for (int k = 2; k <= kz; ++k) {
    int k1 = k - 1;
    tbarf[k] = tbarh(k1) * (sigmah(k) - sigma(k)) / (sigmah(k) - sigmah(k1)) +
               tbarh(k) * (sigma(k) - sigmah(k1)) / (sigmah(k) - sigmah(k1));
  }
  tbarf[1] = d_zero; tbarf[kzp1] = d_zero;
  for (int k = 1; k <= kzp1; ++k) {
    if (sigma(k) < dlowval) thetaf[k] = tbarf[k];
    else thetaf[k] = tbarf[k] * std::pow(sigma(k) + ptop / pd, -rovcp);
  }
  for (int l = 1; l <= kz; ++l)
    for (int k = 1; k <= kz; ++k) {
      if (l > k) e2(k, l) = d_zero;
      if (l <= k) e2(k, l) = d_one;
      e1(k, l) = d_one;
    }
  for (int k = 1; k <= kz; ++k) {
    d1(k, k) = sigma(k + 1) - sigma(k);
    a3(k, k) = -tbarh(k);
    d2(k, k) = rovcp * tbarh(k) / (sigmah(k) + ptop / pd);
    s1(k, k) = sigma(k);
    s2(k, k) = sigmah(k);
    x1(k, k) = d_one;
  }
  for (int k = 1; k <= kz; ++k) {
    for (int l = 1; l <= kz; ++l) { e3(k, l) = d_zero; g1(k, l) = d_zero; }
    e3(k, k) = d_one;
    if (k > 1) g1(k, k) = tbarf[k];
    if (k < kz) g1(k, k + 1) = -tbarf[k + 1];
    if (k < kz) e3(k, k + 1) = d_one;
  }
  w1 = sub(e2, x1);
  w2 = matmul(w1, d1);
  g2 = matmul(e1, d1);
  w1 = matmul(s1, g2);
  g2 = sub(w1, w2);
  w2 = M(n);
  for (int k = 1; k <= kz; ++k) w2(k, k) = d_one / d1(k, k);
  w1 = matmul(g1, g2);
  a1 = matmul(w2, w1);
  w1 = matmul(e1, d1);
  a2 = matmul(s2, w1);
  w2 = matmul(e3, g2);
  for (double& x : w2.d) x = x * d_half;
  w1 = sub(w2, a2);
  a2 = matmul(d2, w1);
  w1 = matmul(e1, d1);
  a4 = matmul(a3, w1);
  for (double& x : a4.d) x = -x;
  for (int j = 1; j <= kz; ++j)
    for (int i = 1; i <= kz; ++i) vm_a0(i, j) = a1(i, j) + a2(i, j) + a3(i, j) + a4(i, j);
  // geopotential matrices (:270-310)
  for (int k = 2; k <= kz; ++k) w1(k, 1) = std::log((sigmah(k) + ptop / pd) / (sigmah(k - 1) + ptop / pd));
  hydros.fill(d_zero);
  for (int k = 1; k <= kz - 1; ++k)
    for (int l = k; l <= kz - 1; ++l) {
      hydros(k, l) = hydros(k, l) + w1(l + 1, 1) * sdsigma[l] / (sdsigma[l + 1] + sdsigma[l]);
      hydros(k, l + 1) = hydros(k, l + 1) + w1(l + 1, 1) * sdsigma[l + 1] / (sdsigma[l + 1] + sdsigma[l]);
    }
  for (int k = 1; k <= kz; ++k)
    hydros(k, kz) = hydros(k, kz) + std::log((d_one + ptop / pd) / (sigmah(kz) + ptop / pd));
  hydroc.fill(d_zero);
  tweigh[1] = d_zero;
  for (int l = 2; l <= kz; ++l)
    tweigh[l] = (tbarh(l) * sdsigma[l] + tbarh(l - 1) * sdsigma[l - 1]) / (sdsigma[l] + sdsigma[l - 1]);
  for (int l = 2; l <= kz - 1; ++l)
    for (int k = 1; k <= l - 1; ++k) hydroc(k, l) = tweigh[l] - tweigh[l + 1];
  for (int l = 1; l <= kz - 1; ++l) hydroc(l, l) = tbarh(l) - tweigh[l + 1];
  for (int k = 1; k <= kz - 1; ++k) hydroc(k, kz) = tweigh[kz] - tbarh(kz);
  for (int k = 1; k <= kz; ++k) hydroc(k, kzp1) = tbarh(kz);
  {  // self-test :314-334
    bool lhydro = false;
    for (int k = 1; k <= kz; ++k) {
      double t1 = d_zero;
      for (int l = 1; l <= kz; ++l) t1 = t1 + hydros(k, l) * tbarh(l);
      double t2 = -tbarh(k) * std::log(sigmah(k) * pd + ptop);
      for (int l = 1; l <= kzp1; ++l) t2 = t2 + hydroc(k, l) * std::log(sigmah(l) * pd + ptop);
      double x = std::fabs(t1 - t2) / (std::fabs(t1) + std::fabs(t2));
      if (x > 1000.0 * std::numeric_limits<double>::epsilon()) lhydro = true;
    }
    if (lhydro) vm_numerr++;
  }
  // tau matrix (:338-353)
  {
    std::vector<double> w3((size_t)(kzp1 + 1) * (kz + 1), 0.0);
    auto W3 = [&](int k, int l) -> double& { return w3[(size_t)k + (size_t)(kzp1 + 1) * l]; };
    for (int l = 1; l <= kz; ++l)
      for (int k = 1; k <= kzp1; ++k) W3(k, l) = sdsigma[l] / (d_one + ptop / (pd * sigmah(k)));
    for (int l = 1; l <= kz; ++l)
      for (int k = 1; k <= kz; ++k) {
        w2(k, l) = d_zero;
        for (int mm = 1; mm <= kzp1; ++mm) w2(k, l) = w2(k, l) + hydroc(k, mm) * W3(mm, l);
      }
    M hs(n), a0m(n);
    for (int j = 1; j <= kz; ++j)
      for (int i = 1; i <= kz; ++i) { hs(i, j) = hydros(i, j); a0m(i, j) = vm_a0(i, j); }
    w1 = matmul(hs, a0m);
    for (int j = 1; j <= kz; ++j)
      for (int i = 1; i <= kz; ++i) {
        vm_tau(i, j) = w1(i, j) - w2(i, j);
        vm_tau(i, j) = -rgas * vm_tau(i, j);
      }
  }
  vm_tau0.d = vm_tau.d;
  // eigenproblem (:358-363): rg -> vcheke -> vorder -> vnorml
  std::vector<double> wr, wi, z;
  int ier = orc_la::eig_real(kz, vm_tau.d, wr, wi, z);
  if (ier) vm_numerr++;
  {  // vcheke :571-593
    int numneg = 0, nimag = 0; double emax = d_zero;
    for (int q = 0; q < kz; ++q) { if (wr[q] <= d_zero) numneg++; if (wr[q] > emax) emax = wr[q]; }
    for (int q = 0; q < kz; ++q) if (wi[q] / emax > std::numeric_limits<double>::epsilon()) nimag++;
    if (numneg + nimag != 0) vm_numerr++;
  }
  {  // vorder :514-540
    std::vector<int> used(kz, 0);
    int kmax = 0;
    for (int l = 1; l <= kz; ++l) {
      double hmax = -std::numeric_limits<double>::epsilon();
      for (int k = 0; k < kz; ++k)
        if (!used[k] && wr[k] > hmax) { hmax = wr[k]; kmax = k; }
      hbar(l) = hmax;
      used[kmax] = 1;
      for (int k = 1; k <= kz; ++k) zmatx(k, l) = z[(size_t)(k - 1) + (size_t)kz * kmax];
    }
  }
  {  // vnorml :546-567
    int kmax = 1;
    for (int l = 1; l <= kz; ++l) {
      double zmax = -d_one, v = d_zero;
      for (int k = 1; k <= kz; ++k) {
        double a = std::fabs(zmatx(k, l));
        if (a > zmax) { zmax = a; kmax = k; }
        v = (sigma(k + 1) - sigma(k)) * a * a + v;
      }
      double a = (zmatx(kmax, l) / zmax) / std::sqrt(v);
      for (int k = 1; k <= kz; ++k) zmatx(k, l) = a * zmatx(k, l);
    }
  }
  std::vector<double> inv;
  if (orc_la::invert(kz, zmatx.d, inv)) vm_numerr++;
  zmatxr.d = inv;
  std::vector<double> hydror;
  if (orc_la::invert(kz, hydros.d, hydror)) vm_numerr++;
  auto HR = [&](int i, int j) { return hydror[(size_t)(i - 1) + (size_t)kz * (j - 1)]; };
  // daley's variational scheme arrays (:389-416): varpa1
  for (int k = 1; k <= kz; ++k) hweigh[k] = d_zero;
  hweigh[kz] = d_one;
  for (int k1 = 1; k1 <= kz; ++k1)
    for (int k2 = 1; k2 <= kz; ++k2) {
      w1(k2, k1) = d_zero;
      for (int k = 1; k <= kz; ++k) w1(k2, k1) = HR(k, k2) * HR(k, k1) * hweigh[k] / (tbarh(k) * tbarh(k)) + w1(k2, k1);
    }
  const double ps2 = xps * xps;
  for (int k1 = 1; k1 <= kzp1; ++k1)
    for (int k2 = 1; k2 <= kz; ++k2) {
      varpa1(k2, k1) = d_zero;
      for (int k = 1; k <= kz; ++k) varpa1(k2, k1) = varpa1(k2, k1) + w1(k2, k) * hydroc(k, k1) * ps2;
    }
  // subtract a4 from a for use in computing am (:422-426)
  for (int k1 = 1; k1 <= kz; ++k1)
    for (int k2 = 1; k2 <= kz; ++k2) vm_a0(k2, k1) = vm_a0(k2, k1) - a4(k2, k1);
}

// Main/mod_split.F90:75-239 (not a restart)
void Oracle::spinit() {
  const double rdx2 = d_one / dx2;
  // dtau: Main/mod_params.F90:1702-1707
  for (int ns = 1; ns <= nsplit; ++ns) dtau(ns) = dt * (d_half / (double)(nsplit - ns + 1));
  for (int ns = 1; ns <= nsplit; ++ns) aam(ns) = (double)std::lround(dtsec / dtau(ns));
  for (int i = ice1; i <= ice2; ++i)
    for (int j = jce1; j <= jce2; ++j) map_(j, i) = d_one / (msfx(j, i) * msfx(j, i));
  if (!cfg.lstand) {
    const double rnpts_ = d_one / (double)((iy - 1) * (jx - 1));  // nicross*njcross
    double lxps = d_zero, lxms = d_zero;
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j) {
        lxms = lxms + rnpts_ * map_(j, i);
        lxps = lxps + rnpts_ * psa(j, i) * map_(j, i);
      }
    xps = lxps / lxms;
    for (int k = 1; k <= kz; ++k) {
      double ltbark = d_zero;
      for (int i = ice1; i <= ice2; ++i)
        for (int j = jce1; j <= jce2; ++j) ltbark = ltbark + rnpts_ * map_(j, i) * a1_t(j, i, k) / psa(j, i);
      tbarh(k) = ltbark / lxms;
    }
  }
  vmodes();
  for (int n = 1; n <= nsplit; ++n) {
    an_(n) = d_zero;
    for (int l = 1; l <= kz; ++l) an_(n) = an_(n) + dsigma(l) * zmatx(l, n);
    for (int k = 1; k <= kz; ++k) { am_(k, n) = d_zero; vm_tau(n, k) = d_zero; }
    for (int l = 1; l <= kz; ++l)
      for (int k = 1; k <= kz; ++k) {
        am_(k, n) = am_(k, n) + vm_a0(k, l) * zmatx(l, n);
        vm_tau(n, k) = vm_tau(n, k) + rgas * zmatxr(n, l) * hydros(l, k);
      }
    for (int k = 1; k <= kzp1; ++k) varpa1(n, k) = d_zero;
    for (int l = 1; l <= kz; ++l)
      for (int k = 1; k <= kzp1; ++k) varpa1(n, k) = varpa1(n, k) + rgas * zmatxr(n, l) * hydroc(l, k);
  }
  for (int l = 1; l <= nsplit; ++l) {
    double fac = d_two * dtsec / (d_two * aam(l) + d_one);
    an_(l) = an_(l) * fac;
    for (int k = 1; k <= kz; ++k) {
      zmatx(k, l) = zmatx(k, l) * fac;
      am_(k, l) = am_(k, l) * fac;
    }
  }
  // divergence manipulations (0)
  for (int k = 1; k <= kz; ++k)
    for (int i = ide1; i <= ide2; ++i)
      for (int j = jde1; j <= jde2; ++j) {
        uuu(j, i, k) = a2_u(j, i, k) * msfd(j, i);
        vvv(j, i, k) = a2_v(j, i, k) * msfd(j, i);
      }
  for (int l = 1; l <= nsplit; ++l)
    for (int k = 1; k <= kz; ++k)
      for (int i = ice1; i <= ice2; ++i)
        for (int j = jce1; j <= jce2; ++j)
          dstor(j, i, l) = dstor(j, i, l) + zmatxr(l, k) * map_(j, i) * rdx2 *
                           (-uuu(j, i + 1, k) + uuu(j + 1, i + 1, k) - uuu(j, i, k) + uuu(j + 1, i, k) +
                            vvv(j, i + 1, k) + vvv(j + 1, i + 1, k) - vvv(j, i, k) - vvv(j + 1, i, k));
  // geopotential manipulations
  for (int l = 1; l <= nsplit; ++l) {
    double pdlog = varpa1(l, kzp1) * std::log(sigmah(kzp1) * pd + ptop);
    double eps1 = varpa1(l, kzp1) * sigmah(kzp1) / (sigmah(kzp1) * pd + ptop);
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j) {
        double eps = eps1 * (psb(j, i) - pd);
        hstor(j, i, l) = pdlog + eps;
      }
    for (int k = 1; k <= kz; ++k) {
      pdlog = varpa1(l, k) * std::log(sigmah(k) * pd + ptop);
      eps1 = varpa1(l, k) * sigmah(k) / (sigmah(k) * pd + ptop);
      for (int i = ice1; i <= ice2; ++i)
        for (int j = jce1; j <= jce2; ++j) {
          double eps = eps1 * (psb(j, i) - pd);
          hstor(j, i, l) = hstor(j, i, l) + pdlog + vm_tau(l, k) * a2_t(j, i, k) / psb(j, i) + eps;
        }
    }
  }
}

// Main/mod_split.F90:243-461
void Oracle::splitf() {
  const double rdx2 = d_one / dx2;
  const double gnu1 = cfg.gnu1;
  deld.fill(d_zero); delh.fill(d_zero);
  psc2psd(psa, psdota);
  for (int n = 1; n <= nsplit; ++n)
    for (int i = ide1; i <= ide2; ++i)
      for (int j = jde1; j <= jde2; ++j) { deld(j, i, n, 1) = dstor(j, i, n); delh(j, i, n, 1) = hstor(j, i, n); }
  // divergence manipulations (f), then (0)
  for (int pass = 0; pass < 2; ++pass) {
    const A3& su = pass ? a2_u : a1_u; const A3& sv = pass ? a2_v : a1_v;
    const int slot = pass ? 2 : 3;
    for (int k = 1; k <= kz; ++k)
      for (int i = ide1; i <= ide2; ++i)
        for (int j = jde1; j <= jde2; ++j) { uuu(j, i, k) = su(j, i, k) * msfd(j, i); vvv(j, i, k) = sv(j, i, k) * msfd(j, i); }
    for (int l = 1; l <= nsplit; ++l) {
      for (int i = ide1; i <= ide2; ++i)
        for (int j = jde1; j <= jde2; ++j) deld(j, i, l, slot) = d_zero;
      for (int k = 1; k <= kz; ++k)
        for (int i = ice1; i <= ice2; ++i)
          for (int j = jce1; j <= jce2; ++j)
            deld(j, i, l, slot) = deld(j, i, l, slot) + zmatxr(l, k) * rdx2 * map_(j, i) *
                                  (-uuu(j, i + 1, k) + uuu(j + 1, i + 1, k) - uuu(j, i, k) + uuu(j + 1, i, k) +
                                   vvv(j, i + 1, k) + vvv(j + 1, i + 1, k) - vvv(j, i, k) - vvv(j + 1, i, k));
    }
    for (int n = 1; n <= nsplit; ++n)
      for (int i = ide1; i <= ide2; ++i)
        for (int j = jde1; j <= jde2; ++j) {
          if (pass == 0) deld(j, i, n, 3) = deld(j, i, n, 3) - deld(j, i, n, 1);
          else deld(j, i, n, 1) = deld(j, i, n, 1) - deld(j, i, n, 2);
        }
  }
  // geopotential manipulations (f), then (0)
  for (int pass = 0; pass < 2; ++pass) {
    const A3& st = pass ? a2_t : a1_t; const A2& sp = pass ? psb : psa;
    const int slot = pass ? 2 : 3;
    for (int l = 1; l <= nsplit; ++l) {
      double pdlog = varpa1(l, kzp1) * std::log(sigmah(kzp1) * pd + ptop);
      double eps1 = varpa1(l, kzp1) * sigmah(kzp1) / (sigmah(kzp1) * pd + ptop);
      for (int i = ice1; i <= ice2; ++i)
        for (int j = jce1; j <= jce2; ++j) {
          double eps = eps1 * (sp(j, i) - pd);
          delh(j, i, l, slot) = pdlog + eps;
        }
      for (int k = 1; k <= kz; ++k) {
        pdlog = varpa1(l, k) * std::log(sigmah(k) * pd + ptop);
        eps1 = varpa1(l, k) * sigmah(k) / (sigmah(k) * pd + ptop);
        for (int i = ice1; i <= ice2; ++i)
          for (int j = jce1; j <= jce2; ++j) {
            double eps = eps1 * (sp(j, i) - pd);
            delh(j, i, l, slot) = delh(j, i, l, slot) + pdlog + vm_tau(l, k) * st(j, i, k) / sp(j, i) + eps;
          }
      }
    }
    for (int n = 1; n <= nsplit; ++n)
      for (int i = ide1; i <= ide2; ++i)
        for (int j = jde1; j <= jde2; ++j) {
          if (pass == 0) delh(j, i, n, 3) = delh(j, i, n, 3) - delh(j, i, n, 1);
          else delh(j, i, n, 1) = delh(j, i, n, 1) - delh(j, i, n, 2);
        }
  }
  for (int n = 1; n <= nsplit; ++n)
    for (int i = ide1; i <= ide2; ++i)
      for (int j = jde1; j <= jde2; ++j) { dstor(j, i, n) = deld(j, i, n, 2); hstor(j, i, n) = delh(j, i, n, 2); }
  spstep();
  // corrections to t and p; u and v
  for (int l = 1; l <= nsplit; ++l) {
    double gnuan = gnu1 * an_(l);
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) {
        psa(j, i) = psa(j, i) - an_(l) * ddsum(j, i, l);
        psb(j, i) = psb(j, i) - gnuan * ddsum(j, i, l);
      }
  }
  for (int l = 1; l <= nsplit; ++l)
    for (int k = 1; k <= kz; ++k) {
      double gnuam = gnu1 * am_(k, l);
      for (int i = ici1; i <= ici2; ++i)
        for (int j = jci1; j <= jci2; ++j) {
          a1_t(j, i, k) = a1_t(j, i, k) + am_(k, l) * ddsum(j, i, l);
          a2_t(j, i, k) = a2_t(j, i, k) + gnuam * ddsum(j, i, l);
        }
    }
  for (int l = 1; l <= nsplit; ++l)
    for (int k = 1; k <= kz; ++k) {
      double gnuzm = gnu1 * zmatx(k, l);
      for (int i = idi1; i <= idi2; ++i)
        for (int j = jdi1; j <= jdi2; ++j) {
          double fac = psdota(j, i) / (dx2 * msfd(j, i));
          double x = fac * (dhsum(j, i, l) + dhsum(j, i - 1, l) - dhsum(j - 1, i, l) - dhsum(j - 1, i - 1, l));
          double y = fac * (dhsum(j, i, l) - dhsum(j, i - 1, l) + dhsum(j - 1, i, l) - dhsum(j - 1, i - 1, l));
          a1_u(j, i, k) = a1_u(j, i, k) - zmatx(k, l) * x;
          a1_v(j, i, k) = a1_v(j, i, k) - zmatx(k, l) * y;
          a2_u(j, i, k) = a2_u(j, i, k) - gnuzm * x;
          a2_v(j, i, k) = a2_v(j, i, k) - gnuzm * y;
        }
    }
}

// Main/mod_split.F90:463-669
void Oracle::spstep() {
  const double rdx2 = d_one / dx2;
  ddsum.fill(d_zero); dhsum.fill(d_zero);
  for (int ns = 1; ns <= nsplit; ++ns) {
    int n0 = 1, n1 = 2, n2 = n0;
    const int m2 = (int)aam(ns) * 2;
    const double dtau2 = dtau(ns) * d_two;
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j) { ddsum(j, i, ns) = deld(j, i, ns, n0); dhsum(j, i, ns) = delh(j, i, ns, n0); }
    // gradient of delh -> (work1, work2) -> divergence work3; `src` = time slot of delh
    auto grad_div = [&](int src) {
      for (int i = ide1; i <= ide2; ++i)
        for (int j = jde1; j <= jde2; ++j) xdelh(j, i) = delh(j, i, ns, src);
      for (int i = idi1; i <= idi2; ++i)
        for (int j = jdi1; j <= jdi2; ++j) {
          double fac = dx2 * msfx(j, i);
          work(j, i, 1) = (xdelh(j, i) + xdelh(j, i - 1) - xdelh(j - 1, i) - xdelh(j - 1, i - 1)) / fac;
          work(j, i, 2) = (xdelh(j, i) + xdelh(j - 1, i) - xdelh(j, i - 1) - xdelh(j - 1, i - 1)) / fac;
        }
      for (int nw = 1; nw <= 2; ++nw)
        for (int i = idi1; i <= idi2; ++i)
          for (int j = jdi1; j <= jdi2; ++j) work(j, i, nw) = work(j, i, nw) * psdota(j, i);
      for (int i = idi1; i <= idi2; ++i)
        for (int j = jdi1; j <= jdi2; ++j) { uu(j, i) = work(j, i, 1) * msfd(j, i); vv(j, i) = work(j, i, 2) * msfd(j, i); }
      for (int i = ici1; i <= ici2; ++i)
        for (int j = jci1; j <= jci2; ++j)
          work(j, i, 3) = rdx2 * map_(j, i) * (-uu(j, i + 1) + uu(j + 1, i + 1) - uu(j, i) + uu(j + 1, i) +
                                               vv(j, i + 1) + vv(j + 1, i + 1) - vv(j, i) - vv(j + 1, i));
    };
    // first step: forward scheme
    grad_div(n0);
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) {
        deld(j, i, ns, n1) = deld(j, i, ns, n0) - dtau(ns) * work(j, i, 3) + deld(j, i, ns, 3) / (double)m2;
        delh(j, i, ns, n1) = delh(j, i, ns, n0) - dtau(ns) * hbar(ns) * deld(j, i, ns, n0) / psa(j, i) +
                             delh(j, i, ns, 3) / (double)m2;
      }
    {
      const double fac = (aam(ns) - d_one) / aam(ns);
      for (int i = ici1; i <= ici2; ++i) delh(jce1, i, ns, n1) = delh(jce1, i, ns, n0) * fac;
      for (int i = ici1; i <= ici2; ++i) delh(jce2, i, ns, n1) = delh(jce2, i, ns, n0) * fac;
      for (int j = jce1; j <= jce2; ++j) delh(j, ice1, ns, n1) = delh(j, ice1, ns, n0) * fac;
      for (int j = jce1; j <= jce2; ++j) delh(j, ice2, ns, n1) = delh(j, ice2, ns, n0) * fac;
    }
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j) {
        ddsum(j, i, ns) = ddsum(j, i, ns) + deld(j, i, ns, n1);
        dhsum(j, i, ns) = dhsum(j, i, ns) + delh(j, i, ns, n1);
      }
    // subsequent steps: leapfrog
    for (int n = 2; n <= m2; ++n) {
      grad_div(n1);
      for (int i = ici1; i <= ici2; ++i)
        for (int j = jci1; j <= jci2; ++j) {
          deld(j, i, ns, n2) = deld(j, i, ns, n0) - dtau2 * work(j, i, 3) + deld(j, i, ns, 3) / aam(ns);
          delh(j, i, ns, n2) = delh(j, i, ns, n0) - dtau2 * hbar(ns) * deld(j, i, ns, n1) / psa(j, i) +
                               delh(j, i, ns, 3) / aam(ns);
        }
      for (int i = ici1; i <= ici2; ++i) delh(jce1, i, ns, n2) = d_two * delh(jce1, i, ns, n1) - delh(jce1, i, ns, n0);
      for (int i = ici1; i <= ici2; ++i) delh(jce2, i, ns, n2) = d_two * delh(jce2, i, ns, n1) - delh(jce2, i, ns, n0);
      for (int j = jce1; j <= jce2; ++j) delh(j, ice1, ns, n2) = d_two * delh(j, ice1, ns, n1) - delh(j, ice1, ns, n0);
      for (int j = jce1; j <= jce2; ++j) delh(j, ice2, ns, n2) = d_two * delh(j, ice2, ns, n1) - delh(j, ice2, ns, n0);
      for (int i = ice1; i <= ice2; ++i)
        for (int j = jce1; j <= jce2; ++j) {
          ddsum(j, i, ns) = ddsum(j, i, ns) + deld(j, i, ns, n2);
          dhsum(j, i, ns) = dhsum(j, i, ns) + delh(j, i, ns, n2);
        }
      n0 = n1; n1 = n2; n2 = n0;
    }
  }
}

int Oracle::get_table(int id, double* out, int n) {
  if (!hydro) { err = "tables exist for idynamic = 1 only"; return RCM_ERR_ARG; }
  const std::vector<double>* v = nullptr;
  switch (id) {
    case RCM_T_HBAR: v = &hbar.d; break;
    case RCM_T_ZMATX: v = &zmatx.d; break;
    case RCM_T_ZMATXR: v = &zmatxr.d; break;
    case RCM_T_TAU: v = &vm_tau.d; break;
    case RCM_T_VARPA1: v = &varpa1.d; break;
    case RCM_T_AM: v = &am_.d; break;
    case RCM_T_AN: v = &an_.d; break;
    case RCM_T_TAU0: v = &vm_tau0.d; break;
    default: err = "bad table id"; return RCM_ERR_ARG;
  }
  const int m = std::min<int>(n, (int)v->size());
  if (out && m > 0) std::memcpy(out, v->data(), (size_t)m * sizeof(double));
  return (int)v->size();
}

}  // namespace orc
