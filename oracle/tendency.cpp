// ORACLE — TEST INFRASTRUCTURE ONLY.
// Restates `tend` (Main/mod_tendency.F90:212-616) for idynamic = 2 with the
// physics calls bypassed ("physics off": physical_parametrizations :271,
// condtq :336-350, shallow_convection, tractend2 are not called; their
// tendency components stay zero), and its internal procedures
//   surface_pressures :815-850   decouple :852-1116   compute_omega :1118-1225
//   init_tendencies :1227-1268   advection :1270-1426 boundary :1462-1513
//   diffusion :1515-1553         adiabatic :1555-1680 curvature :1822-1884
// plus mkslice (Main/mod_slice.F90:167-263, the subset the dynamics reads).
#include <algorithm>
#include <cmath>

#include <chrono>
#include <cstdio>
#include <cstdlib>

#include "model.hpp"

namespace orc {

void Oracle::surface_pressures() {
  // idynamic == 2: computed once (linit); p* is constant so recomputing every
  // step gives the same bits.
  for (int i = ice1; i <= ice2; ++i)
    for (int j = jce1; j <= jce2; ++j) {
      rpsa(j, i) = d_one / psa(j, i);
      rpsb(j, i) = d_one / psb(j, i);
    }
  for (int i = ide1; i <= ide2; ++i)
    for (int j = jde1; j <= jde2; ++j) rpsda(j, i) = d_one / psdota(j, i);
}

void Oracle::decouple() {
  const int iboudy = cfg.iboudy;
  for (int k = 1; k <= kz; ++k)
    for (int i = ide1; i <= ide2; ++i)
      for (int j = jde1; j <= jde2; ++j) {
        ax_uc(j, i, k) = a1_u(j, i, k);
        ax_vc(j, i, k) = a1_v(j, i, k);
        ax_umc(j, i, k) = a1_u(j, i, k) * msfd(j, i);
        ax_vmc(j, i, k) = a1_v(j, i, k) * msfd(j, i);
      }
  for (int k = 1; k <= kz; ++k)
    for (int i = idi1; i <= idi2; ++i)
      for (int j = jdi1; j <= jdi2; ++j) {
        ax_ud(j, i, k) = a1_u(j, i, k) * rpsda(j, i);
        ax_vd(j, i, k) = a1_v(j, i, k) * rpsda(j, i);
      }
  // Boundary U,V points (:895-994)
  if (has_bdyleft) {
    for (int k = 1; k <= kz; ++k)
      for (int i = idi1; i <= idi2; ++i) {
        ax_ud(jdi1, i, k) = wui(i, k) * rpsda(jdi1, i);
        ax_vd(jdi1, i, k) = wvi(i, k) * rpsda(jdi1, i);
      }
    for (int k = 1; k <= kz; ++k)
      for (int i = idi1; i <= idi2; ++i) {
        ax_ud(jde1, i, k) = wue(i, k) * rpsda(jde1, i);
        ax_vd(jde1, i, k) = wve(i, k) * rpsda(jde1, i);
      }
    if (iboudy == 3 || iboudy == 4)
      for (int k = 1; k <= kz; ++k)
        for (int i = idi1; i <= idi2; ++i)
          if (a1_u(jde1, i, k) <= d_zero) {
            ax_ud(jde1, i, k) = ax_ud(jdi1, i, k);
            ax_vd(jde1, i, k) = ax_vd(jdi1, i, k);
          }
  }
  if (has_bdyright) {
    for (int k = 1; k <= kz; ++k)
      for (int i = idi1; i <= idi2; ++i) {
        ax_ud(jdi2, i, k) = eui(i, k) * rpsda(jdi2, i);
        ax_vd(jdi2, i, k) = evi(i, k) * rpsda(jdi2, i);
      }
    for (int k = 1; k <= kz; ++k)
      for (int i = idi1; i <= idi2; ++i) {
        ax_ud(jde2, i, k) = eue(i, k) * rpsda(jde2, i);
        ax_vd(jde2, i, k) = eve(i, k) * rpsda(jde2, i);
      }
    if (iboudy == 3 || iboudy == 4)
      for (int k = 1; k <= kz; ++k)
        for (int i = idi1; i <= idi2; ++i)
          if (a1_u(jde2, i, k) >= d_zero) {
            ax_ud(jde2, i, k) = ax_ud(jdi2, i, k);
            ax_vd(jde2, i, k) = ax_vd(jdi2, i, k);
          }
  }
  if (has_bdybottom) {
    for (int k = 1; k <= kz; ++k)
      for (int j = jdi1; j <= jdi2; ++j) {
        ax_ud(j, idi1, k) = sui(j, k) * rpsda(j, idi1);
        ax_vd(j, idi1, k) = svi(j, k) * rpsda(j, idi1);
      }
    for (int k = 1; k <= kz; ++k)
      for (int j = jde1; j <= jde2; ++j) {
        ax_ud(j, ide1, k) = sue(j, k) * rpsda(j, ide1);
        ax_vd(j, ide1, k) = sve(j, k) * rpsda(j, ide1);
      }
    if (iboudy == 3 || iboudy == 4)
      for (int k = 1; k <= kz; ++k)
        for (int j = jde1; j <= jde2; ++j)
          if (a1_v(j, ide1, k) >= d_zero) {
            ax_ud(j, ide1, k) = ax_ud(j, idi1, k);
            ax_vd(j, ide1, k) = ax_vd(j, idi1, k);
          }
  }
  if (has_bdytop) {
    for (int k = 1; k <= kz; ++k)
      for (int j = jdi1; j <= jdi2; ++j) {
        ax_ud(j, idi2, k) = nui(j, k) * rpsda(j, idi2);
        ax_vd(j, idi2, k) = nvi(j, k) * rpsda(j, idi2);
      }
    for (int k = 1; k <= kz; ++k)
      for (int j = jde1; j <= jde2; ++j) {
        ax_ud(j, ide2, k) = nue(j, k) * rpsda(j, ide2);
        ax_vd(j, ide2, k) = nve(j, k) * rpsda(j, ide2);
      }
    if (iboudy == 3 || iboudy == 4)
      for (int k = 1; k <= kz; ++k)
        for (int j = jde1; j <= jde2; ++j)
          if (a1_v(j, ide2, k) <= d_zero) {
            ax_ud(j, ide2, k) = ax_ud(j, idi2, k);
            ax_vd(j, ide2, k) = ax_vd(j, idi2, k);
          }
  }
  for (int k = 1; k <= kz; ++k)
    for (int i = ide1; i <= ide2; ++i)
      for (int j = jde1; j <= jde2; ++j) {
        ax_umd(j, i, k) = ax_ud(j, i, k) * msfd(j, i);
        ax_vmd(j, i, k) = ax_vd(j, i, k) * msfd(j, i);
      }
  // T, QV, QC (:1013-1025)
  for (int k = 1; k <= kz; ++k)
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j) {
        ax_t(j, i, k) = a1_t(j, i, k) * rpsa(j, i);
        ax_qx(j, i, k, iqv) = std::max(a1_qx(j, i, k, iqv) * rpsa(j, i), minqq);
      }
  for (int n = iqfrst; n <= iqlst; ++n)
    for (int k = 1; k <= kz; ++k)
      for (int i = ice1; i <= ice2; ++i)
        for (int j = jce1; j <= jce2; ++j)
          ax_qx(j, i, k, n) = std::max(a1_qx(j, i, k, n) * rpsa(j, i), d_zero);
  for (int k = 1; k <= kz; ++k)
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j)
        ax_tv(j, i, k) = ax_t(j, i, k) * (d_one + ep1 * ax_qx(j, i, k, iqv));
  if (hydro) {  // idynamic == 1 (:1036-1040, :1093-1096)
    for (int k = 1; k <= kz; ++k)
      for (int i = ice1; i <= ice2; ++i)
        for (int j = jce1; j <= jce2; ++j) {
          a1_pr(j, i, k) = (hsigma(k) * psa(j, i) + ptop) * d_1000;
          a1_rho(j, i, k) = a1_pr(j, i, k) / (rgas * ax_tv(j, i, k));
        }
    for (int k = 1; k <= kz; ++k)
      for (int i = ice1; i <= ice2; ++i)
        for (int j = jce1; j <= jce2; ++j) a2_pr(j, i, k) = (hsigma(k) * psb(j, i) + ptop) * d_1000;
    return;
  }
  // non-hydrostatic part (:1046-1065)
  for (int k = 1; k <= kz; ++k)
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j) ax_pp(j, i, k) = a1_pp(j, i, k) * rpsa(j, i);
  for (int k = 1; k <= kzp1; ++k)
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j) ax_w(j, i, k) = a1_w(j, i, k) * rpsa(j, i);
  for (int k = 1; k <= kz; ++k)
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j) {
        a1_pr(j, i, k) = a0_pr(j, i, k) + ax_pp(j, i, k);
        a1_rho(j, i, k) = a1_pr(j, i, k) / (rgas * ax_tv(j, i, k));
      }
  for (int k = 1; k <= kz; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j)
        ax_pr(j, i, k) = (ax_tv(j, i, k) - a0_t(j, i, k) -
                          ax_pp(j, i, k) / (cpd * a0_rho(j, i, k))) / ax_t(j, i, k);
  for (int k = 1; k <= kz; ++k)
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j)
        a2_pr(j, i, k) = a0_pr(j, i, k) + a2_pp(j, i, k) * rpsb(j, i);
  // ipptls == 1: qcd is a pointer to atmx%qx(:,:,:,iqc) (:117-121)
}

void Oracle::compute_omega() {
  qdot.fill(d_zero);
  A2 dummy; dummy.alloc(jde1, jde2, ide1, ide2);
  for (int i = ice1; i <= ice2; ++i)
    for (int j = jce1; j <= jce2; ++j) dummy(j, i) = d_one / (dx2 * msfx(j, i) * msfx(j, i));
  for (int k = 1; k <= kz; ++k)
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j) {
        ucc(j, i, k) = (ax_umd(j, i, k) + ax_umd(j, i + 1, k) + ax_umd(j + 1, i, k) + ax_umd(j + 1, i + 1, k));
        vcc(j, i, k) = (ax_vmd(j, i, k) + ax_vmd(j, i + 1, k) + ax_vmd(j + 1, i, k) + ax_vmd(j + 1, i + 1, k));
      }
  for (int k = 2; k <= kz; ++k)
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j)
        qdot(j, i, k) = -a0_rhof(j, i, k) * egrav * ax_w(j, i, k) / a0_ps(j, i) -
                        sigma(k) * (dpsdxm(j, i) * (twt(k, 1) * ucc(j, i, k) + twt(k, 2) * ucc(j, i, k - 1)) +
                                    dpsdym(j, i) * (twt(k, 1) * vcc(j, i, k) + twt(k, 2) * vcc(j, i, k - 1)));
  for (int k = 1; k <= kz; ++k)
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j)
        cr(j, i, k) = ((ax_umc(j + 1, i + 1, k) + ax_umc(j + 1, i, k) - ax_umc(j, i + 1, k) - ax_umc(j, i, k)) +
                       (ax_vmc(j + 1, i + 1, k) + ax_vmc(j, i + 1, k) - ax_vmc(j + 1, i, k) - ax_vmc(j, i, k))) *
                          dummy(j, i) +
                      (qdot(j, i, k + 1) - qdot(j, i, k)) * psa(j, i) / dsigma(k);
  omega.fill(d_zero);
  for (int k = 1; k <= kz; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j)
        omega(j, i, k) = -d_half * egrav * a0_rho(j, i, k) * rpsb(j, i) * (a2_w(j, i, k) + a2_w(j, i, k + 1));
}

// Main/mod_slice.F90:169-263 (fields read by diffusion / calc_coeff only)
void Oracle::mkslice() {
  for (int k = 1; k <= kz; ++k)
    for (int i = ide1; i <= ide2; ++i)
      for (int j = jde1; j <= jde2; ++j) {
        double rpsdotb = d_one / psdotb(j, i);
        ubd3d(j, i, k) = a2_u(j, i, k) * rpsdotb;
        vbd3d(j, i, k) = a2_v(j, i, k) * rpsdotb;
      }
  for (int k = 1; k <= kz; ++k)
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j) {
        double r = d_one / psb(j, i);
        tb3d(j, i, k) = a2_t(j, i, k) * r;
        qxb3d(j, i, k, iqv) = std::max(a2_qx(j, i, k, iqv) * r, minqq);
      }
  for (int n = iqfrst; n <= iqlst; ++n)
    for (int k = 1; k <= kz; ++k)
      for (int i = ice1; i <= ice2; ++i)
        for (int j = jce1; j <= jce2; ++j)
          qxb3d(j, i, k, n) = std::max(a2_qx(j, i, k, n) * (d_one / psb(j, i)), d_zero);
  for (int k = 1; k <= kz; ++k)
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j) ppb3d(j, i, k) = a2_pp(j, i, k) * (d_one / psb(j, i));
  for (int k = 1; k <= kzp1; ++k)
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j) wb3d(j, i, k) = a2_w(j, i, k) * (d_one / psb(j, i));
}

void Oracle::init_tendencies() {
  for (A3* a : {&uten, &vten, &udyn, &vdyn, &tten, &tdyn, &ppten, &ppdyn, &wten, &wdyn}) a->fill(d_zero);
  qxten.fill(d_zero); qxdyn.fill(d_zero);
}

void Oracle::advection() {
  start_advect();
  if (cfg.isladvec == 1 && (sl_rc = trajcalc_x())) return;   // :1283-1285
  hadvuv(udyn, vdyn, ax_ud, ax_vd);
  vadvuv(udyn, vdyn, ax_uc, ax_vc);
  hadv3d(ppdyn, ax_pp, 0);
  vadv3d(ppdyn, a1_pp, kz, 0);
  hadv3d(wdyn, ax_w, 1);
  vadv3d(wdyn, a1_w, kzp1, 0);
  // ithadv == 1 (:1348-1357)
  thten.fill(d_zero);
  for (int k = 1; k <= kz; ++k)
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j) {
        th(j, i, k) = ax_t(j, i, k) * std::pow(p00 / a1_pr(j, i, k), rovcp);
        tha(j, i, k) = th(j, i, k) * psa(j, i);
      }
  hadvt(thten, th);
  vadv3d(thten, tha, kz, 0);
  auto sl = [&](int n) {   // isladvec = 1 (:1361-1363,1378-1380): slhadv_x on atm2%qx, hdvg_x on atm1%qx
    slhadv_x([&](int j, int i, int k) -> double& { return qxdyn(j, i, k, n); }, [&](int j, int i, int k) { return a2_qx(j, i, k, n); });
    hdvg_x([&](int j, int i, int k) -> double& { return qxdyn(j, i, k, n); }, [&](int j, int i, int k) { return a1_qx(j, i, k, n); });
  };
  if (cfg.isladvec == 1) sl(iqv); else
  hadvqv(qxdyn, ax_qx, iqv);
  vadvqv(qxdyn, a1_qx, iqv);   // all(icup /= 1): physics off
  if (cfg.ipptls > 0) {
    if (cfg.isladvec == 1) { for (int n = iqfrst; n <= iqlst; ++n) sl(n); } else
    hadvqx(qxdyn, ax_qx, iqfrst, iqlst);
    vadv4d(qxdyn, a1_qx, iqfrst, iqlst, iqxvadv);
  }
}

void Oracle::curvature() {
  for (int k = 1; k <= kz; ++k)
    for (int i = idi1; i <= idi2; ++i)
      for (int j = jdi1; j <= jdi2; ++j) {
        double wadot = 0.125 * (a1_w(j - 1, i - 1, k) + a1_w(j - 1, i, k) + a1_w(j, i - 1, k) + a1_w(j, i, k));
        double wadotp1 = 0.125 * (a1_w(j - 1, i - 1, k + 1) + a1_w(j - 1, i, k + 1) + a1_w(j, i - 1, k + 1) + a1_w(j, i, k + 1));
        double wabar = wadot + wadotp1;
        double amfac = wabar * rpsda(j, i) * rearthrad;
        double duv = ax_uc(j, i, k) * dmdy(j, i) - ax_vc(j, i, k) * dmdx(j, i);
        udyn(j, i, k) = udyn(j, i, k) + coriol(j, i) * ax_vc(j, i, k) - ef(j, i) * ddx(j, i) * wabar +
                        ax_vmd(j, i, k) * duv - ax_uc(j, i, k) * amfac;
        vdyn(j, i, k) = vdyn(j, i, k) - coriol(j, i) * ax_uc(j, i, k) + ef(j, i) * ddy(j, i) * wabar -
                        ax_umd(j, i, k) * duv - ax_vc(j, i, k) * amfac;
      }
}

void Oracle::adiabatic() {
  // ithadv == 1 branch (:1595-1599)
  for (int k = 1; k <= kz; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) {
        thten(j, i, k) = thten(j, i, k) + th(j, i, k) * cr(j, i, k);
        tdyn(j, i, k) = tdyn(j, i, k) + a1_t(j, i, k) * thten(j, i, k) / tha(j, i, k);
      }
  for (int k = 1; k <= kz; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) ppdyn(j, i, k) = ppdyn(j, i, k) + ax_pp(j, i, k) * cr(j, i, k);
  for (int n = 1; n <= nqx; ++n)
    for (int k = 1; k <= kz; ++k)
      for (int i = ici1; i <= ici2; ++i)
        for (int j = jci1; j <= jci2; ++j)
          qxdyn(j, i, k, n) = qxdyn(j, i, k, n) + ax_qx(j, i, k, n) * cr(j, i, k);
  for (int k = 1; k <= kz; ++k)
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j) {
        ucc(j, i, k) = (ax_uc(j, i, k) + ax_uc(j, i + 1, k) + ax_uc(j + 1, i, k) + ax_uc(j + 1, i + 1, k));
        vcc(j, i, k) = (ax_vc(j, i, k) + ax_vc(j, i + 1, k) + ax_vc(j + 1, i, k) + ax_vc(j + 1, i + 1, k));
      }
  for (int k = 2; k <= kz; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) {
        double rofac = (dsigma(k - 1) * a0_rho(j, i, k) + dsigma(k) * a0_rho(j, i, k - 1)) /
                       (dsigma(k - 1) * a1_rho(j, i, k) + dsigma(k) * a1_rho(j, i, k - 1));
        double uaq = d_rfour * (twt(k, 1) * ucc(j, i, k) + twt(k, 2) * ucc(j, i, k - 1));
        double vaq = d_rfour * (twt(k, 1) * vcc(j, i, k) + twt(k, 2) * vcc(j, i, k - 1));
        wdyn(j, i, k) = wdyn(j, i, k) +
                        (twt(k, 2) * ax_pr(j, i, k - 1) + twt(k, 1) * ax_pr(j, i, k)) * rofac * egrav * psa(j, i) +
                        ex(j, i) * (uaq * crx(j, i) - vaq * cry(j, i)) +
                        (uaq * uaq + vaq * vaq) * rearthrad * rpsa(j, i) +
                        ax_w(j, i, k) * (twt(k, 1) * cr(j, i, k) + twt(k, 2) * cr(j, i, k - 1));
      }
  if (cfg.ipptls > 0)
    for (int k = 2; k <= kz; ++k)
      for (int i = ici1; i <= ici2; ++i)
        for (int j = jci1; j <= jci2; ++j)
          wdyn(j, i, k) = wdyn(j, i, k) - egrav * psa(j, i) *
                                              (twt(k, 2) * ax_qx(j, i, k - 1, iqc) + twt(k, 1) * ax_qx(j, i, k, iqc));
}

void Oracle::boundary() {
  const int ib = cfg.iboudy;
  if (ib == 1 || ib == 5) {
    nudge3d(ib, a2_t, xtb, tdyn);
    nudge4d3d(ib, a2_qx, xqb, qxdyn, iqv);
    nudgeuv(ib, a2_u, a2_v, xub, xvb, udyn, vdyn);
    nudge3d(ib, a2_pp, xppb, ppdyn);
    nudge3d(ib, a2_w, xwwb, wdyn);
  } else if (ib == 4) {   // sponge on the TOTAL tendencies (Main/mod_tendency.F90:1480-1483,1497-1499)
    sponge3d(xtb, tten);
    sponge4d(xqb, qxten, iqv);
    spongeuv(xub, xvb, uten, vten);
    sponge3d(xppb, ppten);
    sponge3d(xwwb, wten);
  }
}

void Oracle::diffusion() {
  diffu_d(udyn, vdyn, ubd3d, vbd3d);
  diffu_x3d(tdyn, tb3d);
  diffu_x4d(qxdyn, qxb3d, 1, nqx, d_one);
  diffu_x3d(ppdyn, ppb3d);
  diffu_x3df(wdyn, wb3d, d_one);
}

void Oracle::tend() {
  if (hydro) { tend_hydro(); return; }
  static const bool prof = std::getenv("ORC_PROF") != nullptr;
  auto t0 = std::chrono::steady_clock::now();
  auto lap = [&](const char* nm) {
    if (!prof) return;
    auto t1 = std::chrono::steady_clock::now();
    std::fprintf(stderr, "  [orc] %-18s %8.3f ms\n", nm, std::chrono::duration<double, std::milli>(t1 - t0).count());
    t0 = t1;
  };
  surface_pressures(); lap("surface_pressures");
  decouple(); lap("decouple");
  compute_omega(); lap("compute_omega");
  mkslice(); lap("mkslice");
  calc_coeff(); lap("calc_coeff");
  init_tendencies(); lap("init_tendencies");
  advection(); lap("advection");
  curvature(); lap("curvature");
  adiabatic(); lap("adiabatic");
  boundary(); lap("boundary");
  // physical_parametrizations: OFF
  diffusion(); lap("diffusion");
  tke_tendency();      // ibltyp = 2: TKE advection, diffusion, forecast, filter (oracle/tke.cpp)
  tracer_tendency();   // ichem = 1: advection, nudge_chi, diffusion, forecast, negative fix, filter (oracle/tracers.cpp)
  // tendency sums (:285-314,331-335): ten = ten + dyn + phy (aten(pc_physic), zero until set: oracle/physics.cpp)
  for (int k = 1; k <= kz; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) tten(j, i, k) = tten(j, i, k) + tdyn(j, i, k) + phy3(tphy, j, i, k);
  for (int k = 1; k <= kz; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j)
        qxten(j, i, k, iqv) = qxten(j, i, k, iqv) + qxdyn(j, i, k, iqv) + phy4(j, i, k, iqv);
  for (int k = 1; k <= kzp1; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) wten(j, i, k) = wten(j, i, k) + wdyn(j, i, k) + phy3(wphy, j, i, k);
  for (int k = 1; k <= kz; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) ppten(j, i, k) = ppten(j, i, k) + ppdyn(j, i, k) + phy3(ppphy, j, i, k);
  if (cfg.ipptls > 0)
    for (int n = iqfrst; n <= iqlst; ++n)
      for (int k = 1; k <= kz; ++k)
        for (int i = ici1; i <= ici2; ++i)
          for (int j = jci1; j <= jci2; ++j)
            qxten(j, i, k, n) = qxten(j, i, k, n) + qxdyn(j, i, k, n) + phy4(j, i, k, n);
  if (cfg.ifrayd == 1) {  // i_crm /= 1 (:356-364)
    raydamp3(a0_z, a2_t, tten, xtb);
    raydampqv(a0_z, a2_qx, qxten, xqb);
  }
  // forecast t, qv, qc at tau+1 (:368-393)
  for (int k = 1; k <= kz; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) ac_t(j, i, k) = a2_t(j, i, k) + dt * tten(j, i, k);
  for (int n = 1; n <= nqx; ++n)
    for (int k = 1; k <= kz; ++k)
      for (int i = ice1; i <= ice2; ++i)
        for (int j = jce1; j <= jce2; ++j) ac_qx(j, i, k, n) = a2_qx(j, i, k, n);
  for (int n = 1; n <= nqx; ++n)
    for (int k = 1; k <= kz; ++k)
      for (int i = ici1; i <= ici2; ++i)
        for (int j = jci1; j <= jci2; ++j) ac_qx(j, i, k, n) = ac_qx(j, i, k, n) + dt * qxten(j, i, k, n);
  repair_negative_qx();  // :381-393
  // (:404-411)
  for (int k = 1; k <= kz; ++k)
    for (int i = idi1; i <= idi2; ++i)
      for (int j = jdi1; j <= jdi2; ++j) {
        uten(j, i, k) = uten(j, i, k) + udyn(j, i, k) + phy3(uphy, j, i, k);
        vten(j, i, k) = vten(j, i, k) + vdyn(j, i, k) + phy3(vphy, j, i, k);
      }
  // time filters (:422-427)
  filter_ra_3d(a1_t, a2_t, ac_t, cfg.gnu1);
  filter_raw_qv(a1_qx, a2_qx, ac_qx, cfg.gnu1, 0.53, psa, psb);
  filter_raw_4d(a1_qx, a2_qx, ac_qx, cfg.gnu2, 0.53, iqfrst, iqlst, d_zero);
  // decouple before calling sound (:466-499)
  if (cfg.ifrayd == 1) {
    raydampuv(a0_z, a2_u, a2_v, uten, vten, xub, xvb);
    raydamp3(a0_z, a2_pp, ppten, xppb);
    raydamp3f(a0_zf, a2_w, wten, d_zero);
  }
  for (int k = 1; k <= kz; ++k)
    for (int i = idi1; i <= idi2; ++i)
      for (int j = jdi1; j <= jdi2; ++j) {
        uten(j, i, k) = uten(j, i, k) * rpsda(j, i);
        vten(j, i, k) = vten(j, i, k) * rpsda(j, i);
      }
  for (int k = 1; k <= kz; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) ppten(j, i, k) = ppten(j, i, k) * rpsa(j, i);
  for (int k = 1; k <= kzp1; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) wten(j, i, k) = wten(j, i, k) * rpsa(j, i);
  if (break_before_sound) return;  // test hook: leave the inputs of sound in place (oracle/taps.cpp)
  sound(); lap("sound");
  for (int k = 1; k <= kz; ++k)
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j) a2_pr(j, i, k) = a0_pr(j, i, k) + a2_pp(j, i, k) * rpsb(j, i);
  // next timestep ready (:608-616)
  timer_advance();
  if (lcount == 2) dt = dt2;
}

// Negative-moisture repair, Main/mod_tendency.F90:381-393: sequential sweep, a point
// sees the already repaired values of the neighbours that precede it in loop order.
//   call exchange(atmc%qx,1,...) ; do n ; do k ; do i ; do j
//     if ( atmc%qx(j,i,k,n) < d_zero ) atmc%qx(j,i,k,n) = 0.01 * sum(abs(atmc%qx(j-1:j+1,i-1:i+1,k,n)))/9
// On a decomposed domain every rank sweeps its own tile and its ghost cells keep the values exchanged BEFORE
// the sweep.  set_repair_tiles() lets this single-domain restatement reproduce that: neighbours that belong
// to another tile are read from a copy taken before the sweep.
void Oracle::repair_negative_qx() {
  neg_q_points = 0;
  const bool tiled = rt_j.size() > 1 || rt_i.size() > 1;
  A4 raw;
  if (tiled) raw = ac_qx;
  auto tile_of = [](const std::vector<int>& st, int x) { int q = 0; while (q + 1 < (int)st.size() && st[q + 1] <= x) ++q; return q; };
  for (int n = 1; n <= nqx; ++n)
    for (int k = 1; k <= kz; ++k)
      for (int i = ici1; i <= ici2; ++i)
        for (int j = jci1; j <= jci2; ++j)
          if (ac_qx(j, i, k, n) < d_zero) {
            // sum(abs(atmc%qx(j-1:j+1,i-1:i+1,k,n))): array-section order
            double s = 0.0;
            for (int ii = i - 1; ii <= i + 1; ++ii)
              for (int jj = j - 1; jj <= j + 1; ++jj) {
                const bool other = tiled && (tile_of(rt_j, jj) != tile_of(rt_j, j) || tile_of(rt_i, ii) != tile_of(rt_i, i));
                s += std::fabs(other ? raw(jj, ii, k, n) : ac_qx(jj, ii, k, n));
              }
            ac_qx(j, i, k, n) = 0.01 * s / 9.0;
            neg_q_points++;
          }
}

}  // namespace orc
