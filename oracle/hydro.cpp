// ORACLE — TEST INFRASTRUCTURE ONLY.
// Hydrostatic core (idynamic = 1) of the reference, restated loop for loop:
//   init_bdy / bdyin (p* part)      Main/mod_bdycod.F90:393-396,522-525,636-637,686,774-777,824-825
//   init (p* part)                  Main/mod_init.F90:140-147,173-178
//   bdyval (p* rim)                 Main/mod_bdycod.F90:1317-1342,1426-1451
//   surface_pressures               Main/mod_tendency.F90:815-850
//   decouple (hydro lines)          Main/mod_tendency.F90:858-862,1036-1040,1093-1096
//   compute_omega (hydro branch)    Main/mod_tendency.F90:1127-1156,1199-1215
//   mkslice (hydro lines)           Main/mod_slice.F90:169-240
//   new_pressure                    Main/mod_tendency.F90:1428-1460
//   nudge2d                         Main/mod_bdycod.F90:4597-4766
//   calc_coeff (hydro branch)       Main/mod_diffusion.F90:193-211
//   hadvuv (hydro branches)         Main/mod_advection.F90:142-168,203-234
//   vadv3d ind = 1 (hydro)          Main/mod_advection.F90:771-783
//   curvature / adiabatic (hydro)   Main/mod_tendency.F90:1829-1838,1561-1575
//   pressure_gradient_force ipgf=0,1  Main/mod_tendency.F90:1886-2115
//   filter_ra_2d                    Main/mod_timefilter.F90:58-73
//   tend (hydro path)               Main/mod_tendency.F90:212-616
// The split-explicit scheme (spinit, splitf, spstep, vmodes) is in split.cpp.
#include <algorithm>
#include <cmath>

#include "model.hpp"

namespace orc {

// the four zone loops in the reference's order (south, north, west, east)
#define ORC_ZONES_H(BA, ...)                                          \
  do {                                                                \
    const I2* zs[4] = {&BA.bsouth, &BA.bnorth, &BA.bwest, &BA.beast}; \
    const int zn[4] = {BA.ns, BA.nn, BA.nw, BA.ne};                   \
    for (int z = 0; z < 4; ++z) {                                     \
      if (zn[z] == 0) continue;                                       \
      const I2& zone = *zs[z];                                        \
      __VA_ARGS__                                                      \
    }                                                                 \
  } while (0)

void Oracle::allocate_hydro() {
  pten.alloc(jce1, jce2, ice1, ice2); pten_raw.alloc(jce1, jce2, ice1, ice2); rpsc.alloc(jce1, jce2, ice1, ice2);
  ptn2d.alloc(jce1, jce2, ice1, ice2); pt22d.alloc(jce1, jce2, ice1, ice2);
  phi.alloc(jce1, jce2, ice1, ice2, 1, kz); td.alloc(jce1, jce2, ice1, ice2, 1, kz);
  tvfac.alloc(jce1, jce2, ice1, ice2, 1, kz);
  pf3d.alloc(jce1, jce2, ice1, ice2, 1, kzp1); pb3d.alloc(jce1, jce2, ice1, ice2, 1, kz);
  xpsb.b0.alloc(jce1, jce2, ice1, ice2); xpsb.b1.alloc(jce1, jce2, ice1, ice2);
  xpsb.bt.alloc(jce1, jce2, ice1, ice2); nx_ps.alloc(jce1, jce2, ice1, ice2);
  // mod_vmodes.F90:56-68
  vm_a0.alloc(1, kz, 1, kz); hbar.alloc(1, kz); sigmah.alloc(1, kzp1); tbarh.alloc(1, kz);
  zmatx.alloc(1, kz, 1, kz); zmatxr.alloc(1, kz, 1, kz); vm_tau.alloc(1, kz, 1, kz); vm_tau0.alloc(1, kz, 1, kz);
  varpa1.alloc(1, kz, 1, kzp1); hydroc.alloc(1, kz, 1, kzp1); hydros.alloc(1, kz, 1, kz);
  // mod_split.F90:55-71, mod_atm_interface dstor/hstor
  aam.alloc(1, nsplit); an_.alloc(1, nsplit); dtau.alloc(1, nsplit);
  am_.alloc(1, kz, 1, nsplit); map_.alloc(jce1, jce2, ice1, ice2);
  ddsum.alloc(jde1, jde2, ide1, ide2, 1, nsplit); dhsum.alloc(jde1, jde2, ide1, ide2, 1, nsplit);
  deld.alloc(jde1, jde2, ide1, ide2, 1, nsplit, 1, 3); delh.alloc(jde1, jde2, ide1, ide2, 1, nsplit, 1, 3);
  xdelh.alloc(jde1, jde2, ide1, ide2);
  work.alloc(jdi1, jdi2, idi1, idi2, 1, 3);
  uu.alloc(jdi1, jdi2, idi1, idi2); vv.alloc(jdi1, jdi2, idi1, idi2);
  uuu.alloc(jde1, jde2, ide1, ide2, 1, kz); vvv.alloc(jde1, jde2, ide1, ide2, 1, kz);
  dstor.alloc(jde1, jde2, ide1, ide2, 1, nsplit); hstor.alloc(jde1, jde2, ide1, ide2, 1, nsplit);
}

// Main/mod_bdycod.F90:393-396,408-423,522-552,625-637.  The records arrive
// decoupled; xpsb arrives as surface pressure in hPa.
void Oracle::init_bdy_hydro() {
  xbctime = d_zero;
  const double d_r10 = 0.1;
  for (int s = 0; s < 2; ++s) {
    A2& ps = s ? xpsb.b1 : xpsb.b0;
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j) ps(j, i) = (ps(j, i) * d_r10) - ptop;
    psc2psd(ps, psdot);
    A3& u = s ? xub.b1 : xub.b0; A3& v = s ? xvb.b1 : xvb.b0;
    A3& t = s ? xtb.b1 : xtb.b0; A3& q = s ? xqb.b1 : xqb.b0;
    couple(u, psdot, jde1, jde2, ide1, ide2, 1, kz);
    couple(v, psdot, jde1, jde2, ide1, ide2, 1, kz);
    couple(t, ps, jce1, jce2, ice1, ice2, 1, kz);
    couple(q, ps, jce1, jce2, ice1, ice2, 1, kz);
  }
  timeint(xub.b1, xub.b0, xub.bt, jde1, jde2, ide1, ide2, 1, kz);
  timeint(xvb.b1, xvb.b0, xvb.bt, jde1, jde2, ide1, ide2, 1, kz);
  timeint(xtb.b1, xtb.b0, xtb.bt, jce1, jce2, ice1, ice2, 1, kz);
  timeint(xqb.b1, xqb.b0, xqb.bt, jce1, jce2, ice1, ice2, 1, kz);
  for (int i = ice1; i <= ice2; ++i)
    for (int j = jce1; j <= jce2; ++j) xpsb.bt(j, i) = (xpsb.b1(j, i) - xpsb.b0(j, i)) * rdtbdy;
}

// state part of bdyin, Main/mod_bdycod.F90:669-690,774-787,808-825
void Oracle::bdyin_hydro() {
  const double d_r10 = 0.1;
  xbctime = d_zero;
  xub.b0.d = xub.b1.d; xvb.b0.d = xvb.b1.d; xtb.b0.d = xtb.b1.d; xqb.b0.d = xqb.b1.d;
  xpsb.b0.d = xpsb.b1.d;
  xub.b1.d = nx_u.d; xvb.b1.d = nx_v.d; xtb.b1.d = nx_t.d; xqb.b1.d = nx_q.d;
  xpsb.b1.d = nx_ps.d;
  have_next = false;
  for (int i = ice1; i <= ice2; ++i)
    for (int j = jce1; j <= jce2; ++j) xpsb.b1(j, i) = (xpsb.b1(j, i) * d_r10) - ptop;
  psc2psd(xpsb.b1, psdot);
  couple(xub.b1, psdot, jde1, jde2, ide1, ide2, 1, kz);
  couple(xvb.b1, psdot, jde1, jde2, ide1, ide2, 1, kz);
  couple(xtb.b1, xpsb.b1, jce1, jce2, ice1, ice2, 1, kz);
  couple(xqb.b1, xpsb.b1, jce1, jce2, ice1, ice2, 1, kz);
  timeint(xub.b1, xub.b0, xub.bt, jde1, jde2, ide1, ide2, 1, kz);
  timeint(xvb.b1, xvb.b0, xvb.bt, jde1, jde2, ide1, ide2, 1, kz);
  timeint(xtb.b1, xtb.b0, xtb.bt, jce1, jce2, ice1, ice2, 1, kz);
  timeint(xqb.b1, xqb.b0, xqb.bt, jce1, jce2, ice1, ice2, 1, kz);
  for (int i = ice1; i <= ice2; ++i)
    for (int j = jce1; j <= jce2; ++j) xpsb.bt(j, i) = (xpsb.b1(j, i) - xpsb.b0(j, i)) * rdtbdy;
}

// Main/mod_init.F90:96-147,173-178
void Oracle::init_state_hydro() {
  for (int k = 1; k <= kz; ++k)
    for (int i = ide1; i <= ide2; ++i)
      for (int j = jde1; j <= jde2; ++j) {
        a1_u(j, i, k) = xub.b0(j, i, k); a1_v(j, i, k) = xvb.b0(j, i, k);
        a2_u(j, i, k) = xub.b0(j, i, k); a2_v(j, i, k) = xvb.b0(j, i, k);
      }
  for (int k = 1; k <= kz; ++k)
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j) {
        a1_t(j, i, k) = xtb.b0(j, i, k); a2_t(j, i, k) = xtb.b0(j, i, k);
        a1_qx(j, i, k, iqv) = xqb.b0(j, i, k); a2_qx(j, i, k, iqv) = xqb.b0(j, i, k);
      }
  for (int i = ice1; i <= ice2; ++i)
    for (int j = jce1; j <= jce2; ++j) { psa(j, i) = xpsb.b0(j, i); psb(j, i) = xpsb.b0(j, i); }
  psc2psd(psa, psdota);
  psc2psd(psb, psdotb);
}

// p* rim of bdyval, Main/mod_bdycod.F90:1317-1342 (fixed), :1426-1451 (time dependent)
void Oracle::bdyval_ps() {
  const double xt = xbctime + dt;
  auto bc = [&](int j, int i) { return (cfg.iboudy == 0) ? xpsb.b0(j, i) : xpsb.b0(j, i) + xt * xpsb.bt(j, i); };
  for (int i = ici1; i <= ici2; ++i) psa(jce1, i) = bc(jce1, i);
  for (int i = ici1; i <= ici2; ++i) psa(jce2, i) = bc(jce2, i);
  for (int j = jce1; j <= jce2; ++j) psa(j, ice1) = bc(j, ice1);
  for (int j = jce1; j <= jce2; ++j) psa(j, ice2) = bc(j, ice2);
}

// Main/mod_tendency.F90:822-834 and :858-862
void Oracle::surface_pressures_hydro() {
  for (int i = ice1; i <= ice2; ++i)
    for (int j = jce1; j <= jce2; ++j) rpsa(j, i) = d_one / psa(j, i);
  psc2psd(psa, psdota);
  for (int i = ice1; i <= ice2; ++i)
    for (int j = jce1; j <= jce2; ++j) rpsb(j, i) = d_one / psb(j, i);
  psc2psd(psb, psdotb);
  for (int i = ide1; i <= ide2; ++i)
    for (int j = jde1; j <= jde2; ++j) rpsda(j, i) = d_one / psdota(j, i);
}

// Main/mod_tendency.F90:1123-1156,1198-1215
void Oracle::compute_omega_hydro() {
  qdot.fill(d_zero);
  A2 dummy; dummy.alloc(jde1, jde2, ide1, ide2);
  for (int i = ice1; i <= ice2; ++i)
    for (int j = jce1; j <= jce2; ++j) dummy(j, i) = d_one / (dx2 * msfx(j, i) * msfx(j, i));
  pten.fill(d_zero);
  for (int k = 1; k <= kz; ++k)
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j)
        cr(j, i, k) = ((ax_umc(j + 1, i + 1, k) + ax_umc(j + 1, i, k) - ax_umc(j, i + 1, k) - ax_umc(j, i, k)) +
                       (ax_vmc(j + 1, i + 1, k) + ax_vmc(j, i + 1, k) - ax_vmc(j + 1, i, k) - ax_vmc(j, i, k))) *
                      dummy(j, i);
  for (int k = 1; k <= kz; ++k)
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j) pten(j, i) = pten(j, i) - cr(j, i, k) * dsigma(k);
  for (int k = 2; k <= kz; ++k)
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j)
        qdot(j, i, k) = qdot(j, i, k - 1) - (pten(j, i) + cr(j, i, k - 1)) * dsigma(k - 1) * rpsa(j, i);
  omega.fill(d_zero);
  for (int i = ici1; i <= ici2; ++i)
    for (int j = jci1; j <= jci2; ++j) dummy(j, i) = d_one / (dx8 * msfx(j, i));
  for (int k = 1; k <= kz; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j)
        omega(j, i, k) = d_half * (qdot(j, i, k + 1) + qdot(j, i, k)) * psa(j, i) +
                         hsigma(k) * (pten(j, i) +
                                      ((ax_ud(j, i, k) + ax_ud(j, i + 1, k) + ax_ud(j + 1, i + 1, k) + ax_ud(j + 1, i, k)) *
                                           (psa(j + 1, i) - psa(j - 1, i)) +
                                       (ax_vd(j, i, k) + ax_vd(j, i + 1, k) + ax_vd(j + 1, i + 1, k) + ax_vd(j + 1, i, k)) *
                                           (psa(j, i + 1) - psa(j, i - 1))) *
                                          dummy(j, i));
}

// Main/mod_slice.F90:169-200,226-235 (the fields the hydrostatic dynamics reads:
// ubd3d, vbd3d, tb3d, qxb3d for diffusion/calc_coeff; pb3d, pf3d for vadv3d ind=1)
void Oracle::mkslice_hydro() {
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
      for (int j = jce1; j <= jce2; ++j) pb3d(j, i, k) = (hsigma(k) * psb(j, i) + ptop) * d_1000;
  for (int k = 1; k <= kzp1; ++k)
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j) pf3d(j, i, k) = (sigma(k) * psb(j, i) + ptop) * d_1000;
}

// Main/mod_bdycod.F90:4597-4766 (ibdy = 1: fcx/gcx; otherwise hefc/hegc at k = kz)
void Oracle::nudge2d(int ibdy, const A2& f, const V2dBound& bnd, A2& ften) {
  if (!ba_cr.havebound) return;
  const double xt = xbctime + dt;
  for (int i = ice1; i <= ice2; ++i)
    for (int j = jce1; j <= jce2; ++j) fg1(j, i, 1) = (bnd.b0(j, i) + xt * bnd.bt(j, i)) - f(j, i);
  ORC_ZONES_H(ba_cr, {
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) {
        if (!zone(j, i)) continue;
        int ib = ba_cr.ibnd(j, i);
        double xf = (ibdy == 1) ? fcx(ib) : hefc(ib, kz);
        double xg = (ibdy == 1) ? gcx(ib) : hegc(ib, kz);
        double fls0 = fg1(j, i, 1), fls1 = fg1(j - 1, i, 1), fls2 = fg1(j + 1, i, 1),
               fls3 = fg1(j, i - 1, 1), fls4 = fg1(j, i + 1, 1);
        ften(j, i) = ften(j, i) + xf * fls0 - xg * (fls1 + fls2 + fls3 + fls4 - d_four * fls0);
      }
  });
}

// Main/mod_tendency.F90:1428-1460
void Oracle::new_pressure() {
  if (cfg.iboudy == 4) sponge2d(xpsb, pten);   // Main/mod_tendency.F90:1434
  else if (cfg.iboudy == 1 || cfg.iboudy == 5) nudge2d(cfg.iboudy, psb, xpsb, pten);
  for (int i = ice1; i <= ice2; ++i)
    for (int j = jce1; j <= jce2; ++j) {
      psc(j, i) = psb(j, i) + pten(j, i) * dt;
      rpsc(j, i) = d_one / psc(j, i);
    }
  ptntot = d_zero; pt2tot = d_zero;
  if (integrating()) {
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) {
        ptntot = ptntot + std::fabs(pten(j, i));
        pt2tot = pt2tot + std::fabs((psc(j, i) + psb(j, i) - d_two * psa(j, i)) / (dt * dt * d_rfour));
      }
  }
}

// Main/mod_diffusion.F90:182-211,232-250 (hydrostatic: no dw/dz term)
void Oracle::calc_coeff_hydro() {
  if (cfg.idiffu == 3) { calc_coeff6(); return; }
  const A3& ud = ubd3d; const A3& vd = vbd3d;
  xkc.fill(d_zero); xkd.fill(d_zero); xkcf.fill(d_zero);
  for (int k = 1; k <= kz; ++k)
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j) {
        double dudx = ud(j + 1, i, k) + ud(j + 1, i + 1, k) - ud(j, i, k) - ud(j, i + 1, k);
        double dvdx = vd(j + 1, i, k) + vd(j + 1, i + 1, k) - vd(j, i, k) - vd(j, i + 1, k);
        double dudy = ud(j, i + 1, k) + ud(j + 1, i + 1, k) - ud(j, i, k) - ud(j + 1, i, k);
        double dvdy = vd(j, i + 1, k) + vd(j + 1, i + 1, k) - vd(j, i, k) - vd(j + 1, i, k);
        double duv = std::sqrt((dudx - dvdy) * (dudx - dvdy) + (dvdx + dudy) * (dvdx + dudy));
        xkc(j, i, k) = std::min((hgfact(j, i) + dydc * duv), xkhmax);
      }
  if (xkc_raw.d.size() != xkc.d.size()) xkc_raw.alloc(jce1, jce2, ice1, ice2, 1, kz);
  xkc_raw.d = xkc.d;
  // xkcf (:231-235,245-247) is formed for both cores; with idynamic = 1 only the TKE diffusion reads it
  for (int i = ici1; i <= ici2; ++i)
    for (int j = jci1; j <= jci2; ++j) xkcf(j, i, 1) = xkc(j, i, 1);
  for (int k = 1; k <= kz; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) xkcf(j, i, k + 1) = xkc(j, i, k);
  for (int k = 1; k <= kz; ++k)
    for (int i = idi1; i <= idi2; ++i)
      for (int j = jdi1; j <= jdi2; ++j)
        xkd(j, i, k) = d_rfour * (xkc(j, i, k) + xkc(j - 1, i - 1, k) + xkc(j - 1, i, k) + xkc(j, i - 1, k));
  for (int k = 1; k <= kz; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) xkc(j, i, k) = xkc(j, i, k) * rdxsq * psb(j, i);
  for (int k = 1; k <= kzp1; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) xkcf(j, i, k) = xkcf(j, i, k) * rdxsq * psb(j, i);
  for (int k = 1; k <= kz; ++k)
    for (int i = idi1; i <= idi2; ++i)
      for (int j = jdi1; j <= jdi2; ++j) xkd(j, i, k) = xkd(j, i, k) * rdxsq * psdotb(j, i);
}

// Main/mod_advection.F90:142-168 (centred), :203-234 (upstream_mode)
void Oracle::hadvuv_hydro(A3& ut, A3& vt, const A3& u, const A3& v) {
  const A3& ua = ax_umc; const A3& va = ax_vmc;
  const bool up = cfg.upstream_mode != 0;
  for (int k = 1; k <= kz; ++k)
    for (int i = idi1; i <= idi2; ++i)
      for (int j = jdi1; j <= jdi2; ++j) {
        double ucmona = ua(j, i + 1, k) + d_two * ua(j, i, k) + ua(j, i - 1, k);
        double ucmonb = ua(j + 1, i + 1, k) + d_two * ua(j + 1, i, k) + ua(j + 1, i - 1, k);
        double ucmonc = ua(j - 1, i + 1, k) + d_two * ua(j - 1, i, k) + ua(j - 1, i - 1, k);
        double vcmona = va(j + 1, i, k) + d_two * va(j, i, k) + va(j - 1, i, k);
        double vcmonb = va(j + 1, i + 1, k) + d_two * va(j, i + 1, k) + va(j - 1, i + 1, k);
        double vcmonc = va(j + 1, i - 1, k) + d_two * va(j, i - 1, k) + va(j - 1, i - 1, k);
        if (up) {
          double ff1 = ul * (u(j + 1, i, k) + u(j, i, k));
          double ff2 = ul * (u(j - 1, i, k) + u(j, i, k));
          double ff3 = ul * (v(j, i + 1, k) + v(j, i, k));
          double ff4 = ul * (v(j, i - 1, k) + v(j, i, k));
          ucmonb = (d_one + ff1) * ucmona + (d_one - ff1) * ucmonb;
          ucmonc = (d_one + ff2) * ucmonc + (d_one - ff2) * ucmona;
          vcmonb = (d_one + ff3) * vcmona + (d_one - ff3) * vcmonb;
          vcmonc = (d_one + ff4) * vcmonc + (d_one - ff4) * vcmona;
        } else {
          ucmonb = ucmona + ucmonb;
          ucmonc = ucmonc + ucmona;
          vcmonb = vcmona + vcmonb;
          vcmonc = vcmonc + vcmona;
        }
        ut(j, i, k) = ut(j, i, k) - dmsf(j, i) * ((u(j + 1, i, k) + u(j, i, k)) * ucmonb - (u(j, i, k) + u(j - 1, i, k)) * ucmonc +
                                                    (u(j, i + 1, k) + u(j, i, k)) * vcmonb - (u(j, i, k) + u(j, i - 1, k)) * vcmonc);
        vt(j, i, k) = vt(j, i, k) - dmsf(j, i) * ((v(j + 1, i, k) + v(j, i, k)) * ucmonb - (v(j, i, k) + v(j - 1, i, k)) * ucmonc +
                                                    (v(j, i + 1, k) + v(j, i, k)) * vcmonb - (v(j, i, k) + v(j, i - 1, k)) * vcmonc);
      }
}

// Main/mod_advection.F90:771-783: pfs = atms%pf3d, phs = atms%pb3d (:88-89)
void Oracle::vadv3d_ind1_hydro(A3& ften, const A3& f) {
  const A3& svv = qdot;
  A3 dotqdot; dotqdot.alloc(jci1, jci2, ici1, ici2, 1, kz);
  for (int k = 2; k <= kz; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j)
        dotqdot(j, i, k) = svv(j, i, k) * (twt(k, 1) * f(j, i, k) * std::pow(pf3d(j, i, k) / pb3d(j, i, k), c287) +
                                           twt(k, 2) * f(j, i, k - 1) * std::pow(pf3d(j, i, k) / pb3d(j, i, k - 1), c287));
  for (int k = 2; k <= kz; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) {
        ften(j, i, k - 1) = ften(j, i, k - 1) - dotqdot(j, i, k) * xds(k - 1);
        ften(j, i, k) = ften(j, i, k) + dotqdot(j, i, k) * xds(k);
      }
}

// Main/mod_tendency.F90:1886-2119, both ipgf branches (ipgf = 1: the reference's first loop runs `do k = - 1 , kz`,
// :1894 — levels -1 and 0 are outside td / ttld / atm1%t; the levels 1..kz it defines are restated).  alpha_hyd = 0,
// beta_hyd = 1 - 2*alpha_hyd (Share/mod_constants.F90:319-320); qvd, qcd are
// atmx%qx(:,:,:,iqv|iqc) (:113-121, ipptls == 1).
void Oracle::pressure_gradient_force() {
  const double alpha_hyd = 0.0, beta_hyd = d_one - d_two * alpha_hyd;
  for (int k = 1; k <= kz; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) {
        double tva = a1_t(j, i, k) * (d_one + ep1 * ax_qx(j, i, k, iqv));
        double tvb = a2_t(j, i, k) * (d_one + ep1 * a2_qx(j, i, k, iqv) * rpsb(j, i));
        double tvc = ac_t(j, i, k) * (d_one + ep1 * ac_qx(j, i, k, iqv) * rpsc(j, i));
        td(j, i, k) = alpha_hyd * (tvc + tvb) + beta_hyd * tva;
      }
  for (int k = 1; k <= kz; ++k)
    for (int i = ici1; i <= ici2; ++i) td(jce1, i, k) = a1_t(jce1, i, k) * (d_one + ep1 * ax_qx(jce1, i, k, iqv));
  for (int k = 1; k <= kz; ++k)
    for (int i = ici1; i <= ici2; ++i) td(jce2, i, k) = a1_t(jce2, i, k) * (d_one + ep1 * ax_qx(jce2, i, k, iqv));
  for (int k = 1; k <= kz; ++k)
    for (int j = jce1; j <= jce2; ++j) td(j, ice1, k) = a1_t(j, ice1, k) * (d_one + ep1 * ax_qx(j, ice1, k, iqv));
  for (int k = 1; k <= kz; ++k)
    for (int j = jce1; j <= jce2; ++j) td(j, ice2, k) = a1_t(j, ice2, k) * (d_one + ep1 * ax_qx(j, ice2, k, iqv));
  const bool pg1 = (cfg.ipgf == 1);
  if (pg1) {   // ttld (:1901-1902,1909-1910,...): on every point td is defined on
    if (ttld.d.size() != td.d.size()) ttld.alloc(jce1, jce2, ice1, ice2, 1, kz);
    for (int k = 1; k <= kz; ++k)
      for (int i = ice1; i <= ice2; ++i)
        for (int j = jce1; j <= jce2; ++j)
          ttld(j, i, k) = td(j, i, k) - psa(j, i) * t00pg * std::pow((hsigma(k) * psa(j, i) + ptop) / p00pg, pgfaa1);
  }
  for (int k = 1; k <= kz; ++k)
    for (int i = idi1; i <= idi2; ++i)
      for (int j = jdi1; j <= jdi2; ++j) {
        double rtbar = d_rfour * (ax_tv(j - 1, i - 1, k) + ax_tv(j - 1, i, k) + ax_tv(j, i - 1, k) + ax_tv(j, i, k));
        if (pg1) rtbar = rtbar - t00pg * std::pow((hsigma(k) * psdota(j, i) + ptop) / p00pg, pgfaa1);
        rtbar = rgas * rtbar * psdota(j, i);
        udyn(j, i, k) = udyn(j, i, k) -
                        rtbar * (std::log(d_half * (psa(j, i) + psa(j, i - 1)) * hsigma(k) + ptop) -
                                 std::log(d_half * (psa(j - 1, i) + psa(j - 1, i - 1)) * hsigma(k) + ptop)) /
                            (dx * msfd(j, i));
        vdyn(j, i, k) = vdyn(j, i, k) -
                        rtbar * (std::log(d_half * (psa(j, i) + psa(j - 1, i)) * hsigma(k) + ptop) -
                                 std::log(d_half * (psa(j - 1, i - 1) + psa(j, i - 1)) * hsigma(k) + ptop)) /
                            (dx * msfd(j, i));
      }
  for (int k = 1; k <= kz; ++k)
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j)
        tvfac(j, i, k) = d_one / (d_one + ax_qx(j, i, k, iqc) / (d_one + ax_qx(j, i, k, iqv)));
  for (int i = ice1; i <= ice2; ++i)
    for (int j = jce1; j <= jce2; ++j) {
      const A3& tq = pg1 ? ttld : td;
      double tv = tq(j, i, kz) * rpsa(j, i) * tvfac(j, i, kz);
      phi(j, i, kz) = ht(j, i);
      if (pg1) phi(j, i, kz) = ht(j, i) + rgas * t00pg / pgfaa1 * std::pow((psa(j, i) + ptop) / p00pg, pgfaa1);   // :2046-2047
      phi(j, i, kz) = phi(j, i, kz) - rgas * tv * std::log((hsigma(kz) + ptop * rpsa(j, i)) / (d_one + ptop * rpsa(j, i)));
    }
  for (int k = 1; k <= kz - 1; ++k) {
    int lev = kz - k;
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j) {
        const A3& tq = pg1 ? ttld : td;
        double tvavg = ((tq(j, i, lev) * dsigma(lev) + tq(j, i, lev + 1) * dsigma(lev + 1)) /
                        (psa(j, i) * (dsigma(lev) + dsigma(lev + 1)))) * tvfac(j, i, lev);
        phi(j, i, lev) = phi(j, i, lev + 1) -
                         rgas * tvavg * std::log((hsigma(lev) + ptop * rpsa(j, i)) / (hsigma(lev + 1) + ptop * rpsa(j, i)));
      }
  }
  for (int k = 1; k <= kz; ++k)
    for (int i = idi1; i <= idi2; ++i)
      for (int j = jdi1; j <= jdi2; ++j) {
        udyn(j, i, k) = udyn(j, i, k) - psdota(j, i) *
                        (phi(j, i, k) + phi(j, i - 1, k) - phi(j - 1, i, k) - phi(j - 1, i - 1, k)) / (dx2 * msfd(j, i));
        vdyn(j, i, k) = vdyn(j, i, k) - psdota(j, i) *
                        (phi(j, i, k) + phi(j - 1, i, k) - phi(j, i - 1, k) - phi(j - 1, i - 1, k)) / (dx2 * msfd(j, i));
      }
}

// Main/mod_timefilter.F90:58-73
void Oracle::filter_ra_2d(A2& phin, A2& phinm1, const A2& phinp1, double alpha) {
  for (int i = ici1; i <= ici2; ++i)
    for (int j = jci1; j <= jci2; ++j) {
      double d = alpha * (phinp1(j, i) + phinm1(j, i) - d_two * phin(j, i));
      phinm1(j, i) = phin(j, i) + d;
      phin(j, i) = phinp1(j, i);
    }
}

// `tend` for idynamic = 1, physics off (Main/mod_tendency.F90:212-616)
void Oracle::tend_hydro() {
  surface_pressures_hydro();
  decouple();
  compute_omega_hydro();
  mkslice_hydro();
  pten_raw.d = pten.d;  // tap (tests only): pten before the boundary relaxation
  new_pressure();
  calc_coeff_hydro();
  // init_tendencies (:1227-1268)
  for (A3* a : {&uten, &vten, &udyn, &vdyn, &tten, &tdyn}) a->fill(d_zero);
  qxten.fill(d_zero); qxdyn.fill(d_zero);
  // advection (:1270-1392), ithadv = 0
  start_advect();
  if (cfg.isladvec == 1 && (sl_rc = trajcalc_x())) return;   // :1283-1285
  hadvuv_hydro(udyn, vdyn, ax_ud, ax_vd);
  vadvuv(udyn, vdyn, ax_uc, ax_vc);
  hadvt(tdyn, ax_t);
  vadv3d_ind1_hydro(tdyn, a1_t);
  auto sl = [&](int n) {   // isladvec = 1 (:1361-1363,1378-1380; oracle/sladvection.cpp)
    slhadv_x([&](int j, int i, int k) -> double& { return qxdyn(j, i, k, n); }, [&](int j, int i, int k) { return a2_qx(j, i, k, n); });
    hdvg_x([&](int j, int i, int k) -> double& { return qxdyn(j, i, k, n); }, [&](int j, int i, int k) { return a1_qx(j, i, k, n); });
  };
  if (cfg.isladvec == 1) sl(iqv); else
  hadvqv(qxdyn, ax_qx, iqv);
  vadvqv(qxdyn, a1_qx, iqv);
  if (cfg.ipptls > 0) {
    if (cfg.isladvec == 1) { for (int n = iqfrst; n <= iqlst; ++n) sl(n); } else
    hadvqx(qxdyn, ax_qx, iqfrst, iqlst);
    vadv4d(qxdyn, a1_qx, iqfrst, iqlst, iqxvadv);
  }
  // curvature (:1829-1838)
  for (int k = 1; k <= kz; ++k)
    for (int i = idi1; i <= idi2; ++i)
      for (int j = jdi1; j <= jdi2; ++j) {
        udyn(j, i, k) = udyn(j, i, k) + coriol(j, i) * ax_vc(j, i, k);
        vdyn(j, i, k) = vdyn(j, i, k) - coriol(j, i) * ax_uc(j, i, k);
      }
  // adiabatic (:1561-1575)
  for (int k = 1; k <= kz; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) {
        double rovcpm = rgas / cpmf(ax_qx(j, i, k, iqv));
        tdyn(j, i, k) = tdyn(j, i, k) + (omega(j, i, k) * rovcpm * ax_tv(j, i, k)) / (ptop * rpsa(j, i) + hsigma(k));
      }
  // boundary (:1462-1492)
  if (cfg.iboudy == 1 || cfg.iboudy == 5) {
    nudge3d(cfg.iboudy, a2_t, xtb, tdyn);
    nudge4d3d(cfg.iboudy, a2_qx, xqb, qxdyn, iqv);
    nudgeuv(cfg.iboudy, a2_u, a2_v, xub, xvb, udyn, vdyn);
  } else if (cfg.iboudy == 4) {   // Main/mod_tendency.F90:1480-1483
    sponge3d(xtb, tten);
    sponge4d(xqb, qxten, iqv);
    spongeuv(xub, xvb, uten, vten);
  }
  // diffusion (:1515-1530)
  diffu_d(udyn, vdyn, ubd3d, vbd3d);
  diffu_x3d(tdyn, tb3d);
  diffu_x4d(qxdyn, qxb3d, 1, nqx, d_one);
  tke_tendency();      // ibltyp = 2 (oracle/tke.cpp)
  tracer_tendency();   // ichem = 1 (oracle/tracers.cpp)
  // tendency sums (:285-294,331-335)
  for (int k = 1; k <= kz; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) tten(j, i, k) = tten(j, i, k) + tdyn(j, i, k) + phy3(tphy, j, i, k);
  for (int k = 1; k <= kz; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) qxten(j, i, k, iqv) = qxten(j, i, k, iqv) + qxdyn(j, i, k, iqv) + phy4(j, i, k, iqv);
  if (cfg.ipptls > 0)
    for (int n = iqfrst; n <= iqlst; ++n)
      for (int k = 1; k <= kz; ++k)
        for (int i = ici1; i <= ici2; ++i)
          for (int j = jci1; j <= jci2; ++j) qxten(j, i, k, n) = qxten(j, i, k, n) + qxdyn(j, i, k, n) + phy4(j, i, k, n);
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
  repair_negative_qx();  // :381-393 (tendency.cpp)
  pressure_gradient_force();
  for (int k = 1; k <= kz; ++k)
    for (int i = idi1; i <= idi2; ++i)
      for (int j = jdi1; j <= jdi2; ++j) {
        uten(j, i, k) = uten(j, i, k) + udyn(j, i, k) + phy3(uphy, j, i, k);
        vten(j, i, k) = vten(j, i, k) + vdyn(j, i, k) + phy3(vphy, j, i, k);
      }
  // time filters (:419-427)
  filter_ra_2d(psa, psb, psc, cfg.gnu1);
  filter_ra_3d(a1_t, a2_t, ac_t, cfg.gnu1);
  filter_raw_qv(a1_qx, a2_qx, ac_qx, cfg.gnu1, 0.53, psa, psb);
  filter_raw_4d(a1_qx, a2_qx, ac_qx, cfg.gnu2, 0.53, iqfrst, iqlst, d_zero);
  // forecast p*u, p*v (:433-445)
  for (int k = 1; k <= kz; ++k)
    for (int i = idi1; i <= idi2; ++i)
      for (int j = jdi1; j <= jdi2; ++j) {
        ac_u(j, i, k) = a2_u(j, i, k) + dt * uten(j, i, k);
        ac_v(j, i, k) = a2_v(j, i, k) + dt * vten(j, i, k);
      }
  filter_ra_uv(a1_u, a2_u, ac_u, a1_v, a2_v, ac_v, cfg.gnu1);
  for (int i = ice1; i <= ice2; ++i)
    for (int j = jce1; j <= jce2; ++j) rpsb(j, i) = d_one / psb(j, i);
  for (int k = 1; k <= kz; ++k)
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j) a2_pr(j, i, k) = (hsigma(k) * psb(j, i) + ptop) * d_1000;
  splitf();
  timer_advance();
  if (lcount == 2) dt = dt2;
}

}  // namespace orc
