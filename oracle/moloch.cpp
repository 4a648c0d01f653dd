// ORACLE — TEST INFRASTRUCTURE ONLY.
// Restates the MOLOCH core (idynamic = 3) of the reference, single rank,
// both projection branches (generic and iproj = 'ROTLLR'), physics off:
//   Main/mod_moloch.F90       allocate_moloch:115 init_moloch:159 moloch:224
//                             boundary:485 filt3d:549 sound:708 advection:1042
//                             wafone:1098 stagger operators :1574-1794
//   Share/mod_zita.F90:46-120 coordinate functions
//   Main/mod_params.F90:2743-2847 compute_moloch_static
//   Main/mod_init.F90:178-336,998-1013 initial state, ffilt
//   Main/mod_bdycod.F90       init_bdy/bdyin (idynamic 3 branches), bdyval
//                             (:1343-1385,1452-1491,1539-1575,1662-1699,1963-2100),
//                             monudge3d:4408 monudgeuv:3825 moraydamp:5047
//                             paicompute:5125
// Loop nests, bounds and the order of floating-point operations follow the
// reference line by line; reference quirks are kept.
// "parity unpinned" (no golden vectors in the reference); behavioural pins are
// the reference's own analytic cases moloch_static_test1/2
// (Main/mod_bdycod.F90:5157-5199), see tests/test_moloch.py.
#include <algorithm>
#include <cmath>

#include "model.hpp"

namespace orc {

// ---- Share/mod_zita.F90:46-92 ----
double Oracle::gzita(double z) const {
  const double ratio = z / hzita, a0 = cfg.mo_a0;
  return 1.0 - a0 * ratio - (3.0 - 2.0 * a0) * (ratio * ratio) + (2.0 - a0) * (ratio * ratio * ratio);
}
double Oracle::gzitap(double z) const {
  const double ratio = z / hzita, a0 = cfg.mo_a0;
  return (-a0 - (6.0 - 4.0 * a0) * ratio + (6.0 - 3.0 * a0) * (ratio * ratio)) / hzita;
}
double Oracle::bzita(double z) const { return cfg.mo_b0 + (1.0 - cfg.mo_b0) * (z / hzita); }
double Oracle::bzitap(double) const { return (1.0 - cfg.mo_b0) / hzita; }
double Oracle::md_fmz(double z, double geopot) const {
  const double zfz = 1.0 - z / hzita;
  return zfz / (bzita(z) + geopot / egrav * zfz * gzitap(z) - hzita * zfz * std::log(zfz) * bzitap(z));
}

// Share/pfesat.inc, Share/pfwsat.inc
static double pfesat(double t) {
  const double a0 = 6.11213476, a1 = 0.444007856, a2 = 0.143064234e-01, a3 = 0.264461437e-03, a4 = 0.305903558e-05,
               a5 = 0.196237241e-07, a6 = 0.892344772e-10, a7 = -0.373208410e-12, a8 = 0.209339997e-15;
  const double c0 = 6.11123516, c1 = 0.503109514, c2 = 0.188369801e-01, c3 = 0.420547422e-03, c4 = 0.614396778e-05,
               c5 = 0.602780717e-07, c6 = 0.387940929e-09, c7 = 0.149436277e-11, c8 = 0.262655803e-14;
  double td = t - tzero;
  if (td > 100.0) td = 100.0;
  if (td < -75.0) td = -75.0;
  double esat;
  if (td >= 0.0)
    esat = a0 + td * (a1 + td * (a2 + td * (a3 + td * (a4 + td * (a5 + td * (a6 + td * (a7 + td * a8)))))));
  else
    esat = c0 + td * (c1 + td * (c2 + td * (c3 + td * (c4 + td * (c5 + td * (c6 + td * (c7 + td * c8)))))));
  return esat * 100.0;
}
static double pfwsat(double t, double p) {
  const double es = pfesat(t);
  return ep2 * (es / (p - es));
}

// allocate_mod_atm_interface (mo_atm part), allocate_moloch, allocate_mod_bdycon
int Oracle::create_moloch() {
  const rcm_cfg& c = cfg;
  if (c.mo_nadv < 1 || c.mo_nsound < 1) { err = "oracle: mo_nadv, mo_nsound must be >= 1"; return RCM_ERR_ARG; }
  nadv = c.mo_nadv; nsound = c.mo_nsound;
  hzita = rgas * 280.0 / egrav;            // Share/mod_zita.F90:30-31
  mo_dz = hzita / (double)kz;              // Main/mod_params.F90:1169
  const int nspgx = c.nspgx, nspgd = c.nspgd;
  sigma.alloc(1, kzp1); hsigma.alloc(1, kz); dsigma.alloc(1, kz);
  zita.alloc(1, kzp1); zitah.alloc(1, kz); gzitak.alloc(1, kzp1); gzitakh.alloc(1, kz);
  ffilt.alloc(1, kz); ak.alloc(1, kz); bk.alloc(1, kz);
  for (A2* a : {&ht, &msfx, &msfu, &msfv, &xlat, &xlon, &ulat, &vlat, &coriou, &coriov, &hx, &hy, &mx2, &rmu, &rmv})
    a->alloc(jde1, jde2, ide1, ide2);
  p2d.alloc(jdi1, jdi2, idi1, idi2);
  psa.alloc(jce1, jce2, ice1, ice2);
  m_u.alloc(jde1, jde2, ice1, ice2, 1, kz); m_ud.alloc(jde1, jde2, ice1, ice2, 1, kz);
  m_v.alloc(jce1, jce2, ide1, ide2, 1, kz); m_vd.alloc(jce1, jce2, ide1, ide2, 1, kz);
  for (A3* a : {&m_ux, &m_vx, &m_wx, &m_t, &m_pai, &m_tetav, &m_tvirt, &m_p, &m_rho, &m_zeta, &m_fmz, &m_zdiv2})
    a->alloc(jce1, jce2, ice1, ice2, 1, kz);
  for (A3* a : {&m_w, &m_fmzf, &m_deltaw, &m_s, &m_wwkw}) a->alloc(jce1, jce2, ice1, ice2, 1, kzp1);
  m_qx.alloc(jce1, jce2, ice1, ice2, 1, kz, 1, nqx);
  m_wz.alloc(jci1, jci2, ice1, ice2, 1, kz);
  m_wfw.alloc(jci1, jci2, 1, kzp1);
  m_p0.alloc(jce1, jce2, ici1, ici2, 1, kz);
  m_zpby.alloc(jci1, jci2, ici1, ice2);
  m_zpbw.alloc(jci1, jce2, ici1, ici2);
  m_zetau.alloc(jdi1, jdi2, ici1, ici2, 1, kz);
  m_zetav.alloc(jci1, jci2, idi1, idi2, 1, kz);
  for (A3* a : {&xub.b0, &xub.b1, &xub.bt, &nx_u}) a->alloc(jde1, jde2, ice1, ice2, 1, kz);
  for (A3* a : {&xvb.b0, &xvb.b1, &xvb.bt, &nx_v}) a->alloc(jce1, jce2, ide1, ide2, 1, kz);
  for (A3* a : {&xtb.b0, &xtb.b1, &xtb.bt, &nx_t, &xqb.b0, &xqb.b1, &xqb.bt, &nx_q, &xpaib.b0, &xpaib.b1, &xpaib.bt})
    a->alloc(jce1, jce2, ice1, ice2, 1, kz);
  for (A2* a : {&xpsb.b0, &xpsb.b1, &xpsb.bt, &nx_ps}) a->alloc(jce1, jce2, ice1, ice2);
  fcx.alloc(2, nspgx - 1); gcx.alloc(2, nspgx - 1); fcd.alloc(2, nspgd - 1); gcd.alloc(2, nspgd - 1);
  hefc.alloc(2, nspgx - 1, 1, kz); hegc.alloc(2, nspgx - 1, 1, kz);
  hefd.alloc(2, nspgd - 1, 1, kz); hegd.alloc(2, nspgd - 1, 1, kz);
  fg1.alloc(jde1, jde2, ide1, ide2, 1, kzp1); fg2.alloc(jde1, jde2, ide1, ide2, 1, kz);
  wgtx.alloc(2, nspgx - 1); wgtd.alloc(2, nspgd - 1);
  setup_boundaries(false, false, ba_cr);   // Main/mod_params.F90:1788-1791
  setup_boundaries(true, false, ba_ut);
  setup_boundaries(false, true, ba_vt);
  return RCM_OK;
}

// Main/mod_params.F90:2743-2847
void Oracle::compute_moloch_static() {
  for (int i = ice1; i <= ice2; ++i)
    for (int j = jde1; j <= jde2; ++j) coriou(j, i) = eomeg2 * std::sin(ulat(j, i) * degrad);
  for (int i = ide1; i <= ide2; ++i)
    for (int j = jce1; j <= jce2; ++j) coriov(j, i) = eomeg2 * std::sin(vlat(j, i) * degrad);
  for (int i = ice1; i <= ice2; ++i)
    for (int j = jdi1; j <= jdi2; ++j) hx(j, i) = (ht(j, i) - ht(j - 1, i)) * msfu(j, i) * rdx * regrav;
  for (int i = idi1; i <= idi2; ++i)
    for (int j = jce1; j <= jce2; ++j) {
      if (cfg.mo_rotllr) hy(j, i) = (ht(j, i) - ht(j, i - 1)) * rdx * regrav;   // :2766-2771
      else hy(j, i) = (ht(j, i) - ht(j, i - 1)) * msfv(j, i) * rdx * regrav;
    }
  for (int i = ice1; i <= ice2; ++i) { hx(jde1, i) = hx(jdi1, i); hx(jde2, i) = hx(jdi2, i); }
  for (int j = jce1; j <= jce2; ++j) { hy(j, ide1) = hy(j, idi1); hy(j, ide2) = hy(j, idi2); }
  zita(kzp1) = d_zero;
  for (int k = kz; k >= 1; --k) {
    zita(k) = zita(k + 1) + mo_dz;
    zitah(k) = zita(k) - mo_dz * d_half;
  }
  for (int k = 1; k <= kzp1; ++k) sigma(k) = d_one - zita(k) / hzita;
  for (int k = 1; k <= kz; ++k) hsigma(k) = d_one - zitah(k) / hzita;
  for (int k = 1; k <= kz; ++k) {
    ak(k) = -hzita * bzita(zitah(k)) * std::log(hsigma(k));
    bk(k) = gzita(zitah(k));
  }
  for (int k = 1; k <= kz; ++k) dsigma(k) = (sigma(k + 1) - sigma(k));
  for (int k = 1; k <= kz; ++k)
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j) {
        m_zeta(j, i, k) = ak(k) + (bk(k) - d_one) * ht(j, i) * regrav;
        m_fmz(j, i, k) = md_fmz(zitah(k), ht(j, i));
      }
  for (int i = ice1; i <= ice2; ++i)
    for (int j = jce1; j <= jce2; ++j) m_fmzf(j, i, 1) = 0.0;
  for (int k = 2; k <= kzp1; ++k)
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j) m_fmzf(j, i, k) = md_fmz(zita(k), ht(j, i));
  // zetaf, dz (Main/mod_params.F90:2810-2811,2829-2840): the heights the physics sees as atms%zq, atms%dzq
  m_zetaf.alloc(jce1, jce2, ice1, ice2, 1, kzp1); m_dz.alloc(jce1, jce2, ice1, ice2, 1, kz);
  for (int i = ice1; i <= ice2; ++i)
    for (int j = jce1; j <= jce2; ++j) m_zetaf(j, i, 1) = 48446.4;
  for (int k = 2; k <= kzp1; ++k) {
    const double fak = -hzita * bzita(zita(k)) * std::log(std::max(sigma(k), 2.2250738585072014e-308));
    const double fbk = gzita(zita(k));
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j) m_zetaf(j, i, k) = fak + (fbk - d_one) * ht(j, i) * regrav;
  }
  for (int k = 1; k <= kz; ++k)
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j) m_dz(j, i, k) = m_zetaf(j, i, k) - m_zetaf(j, i, k + 1);
  double ztop = -1e300;
  for (double x : m_zeta.d) ztop = std::max(ztop, x);
  rayzd = ztop;
}

// Main/mod_bdycod.F90:5125-5155
void Oracle::paicompute(const A2& ps, const A3& t, const A3& q, A3& pai) {
  const double zz1 = -egrav * hzita * bzita(d_half * mo_dz) * std::log(d_one - d_half * mo_dz / hzita);
  for (int i = ice1; i <= ice2; ++i)
    for (int j = jce1; j <= jce2; ++j) {
      double zdelta = ht(j, i) * (gzita(d_half * mo_dz) - d_one) + zz1;
      double tv = t(j, i, kz) * (d_one + ep1 * q(j, i, kz));
      double zz2 = d_one / (rgas * tv);
      double p = ps(j, i) * std::exp(-zdelta * zz2);
      pai(j, i, kz) = std::pow(p / p00, rovcp);
    }
  for (int k = kz - 1; k >= 1; --k)
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j) {
        double tv1 = t(j, i, k) * (d_one + ep1 * q(j, i, k));
        double tv2 = t(j, i, k + 1) * (d_one + ep1 * q(j, i, k + 1));
        double zb = d_two * egrav * mo_dz / (m_fmzf(j, i, k + 1) * cpd) + tv1 - tv2;
        double zdelta = std::sqrt(zb * zb + d_four * tv2 * tv1);
        pai(j, i, k) = -pai(j, i, k + 1) / (d_two * tv2) * (zb - zdelta);
      }
}

// state part of init_bdy for idynamic 3 (Main/mod_bdycod.F90:342-356,469-482,
// 388-392,526-530,624-634,641-645)
void Oracle::init_bdy_moloch() {
  for (auto& x : xpsb.b0.d) x = x * 100.0;
  for (auto& x : xpsb.b1.d) x = x * 100.0;
  timeint(xub.b1, xub.b0, xub.bt, jde1, jde2, ice1, ice2, 1, kz);
  timeint(xvb.b1, xvb.b0, xvb.bt, jce1, jce2, ide1, ide2, 1, kz);
  timeint(xtb.b1, xtb.b0, xtb.bt, jce1, jce2, ice1, ice2, 1, kz);
  timeint(xqb.b1, xqb.b0, xqb.bt, jce1, jce2, ice1, ice2, 1, kz);
  paicompute(xpsb.b0, xtb.b0, xqb.b0, xpaib.b0);
  paicompute(xpsb.b1, xtb.b1, xqb.b1, xpaib.b1);
  timeint(xpaib.b1, xpaib.b0, xpaib.bt, jce1, jce2, ice1, ice2, 1, kz);
}

// Main/mod_bdycod.F90:654-889, idynamic 3 branches
void Oracle::bdyin_moloch() {
  xbctime = d_zero;
  xub.b0.d = xub.b1.d; xvb.b0.d = xvb.b1.d; xtb.b0.d = xtb.b1.d; xqb.b0.d = xqb.b1.d;
  xpsb.b0.d = xpsb.b1.d; xpaib.b0.d = xpaib.b1.d;
  xub.b1.d = nx_u.d; xvb.b1.d = nx_v.d; xtb.b1.d = nx_t.d; xqb.b1.d = nx_q.d; xpsb.b1.d = nx_ps.d;
  have_next = false;
  for (auto& x : xpsb.b1.d) x = x * 100.0;
  timeint(xub.b1, xub.b0, xub.bt, jde1, jde2, ice1, ice2, 1, kz);
  timeint(xvb.b1, xvb.b0, xvb.bt, jce1, jce2, ide1, ide2, 1, kz);
  timeint(xtb.b1, xtb.b0, xtb.bt, jce1, jce2, ice1, ice2, 1, kz);
  timeint(xqb.b1, xqb.b0, xqb.bt, jce1, jce2, ice1, ice2, 1, kz);
  paicompute(xpsb.b1, xtb.b1, xqb.b1, xpaib.b1);
  timeint(xpaib.b1, xpaib.b0, xpaib.bt, jce1, jce2, ice1, ice2, 1, kz);
}

// Main/mod_init.F90:178-336 (initial state from b0), :998-1013 (ffilt)
void Oracle::init_state_moloch() {
  const int ipptls = cfg.ipptls;
  for (int k = 1; k <= kz; ++k)
    for (int i = ide1; i <= ide2; ++i)
      for (int j = jce1; j <= jce2; ++j) m_v(j, i, k) = xvb.b0(j, i, k);
  for (int k = 1; k <= kz; ++k)
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jde1; j <= jde2; ++j) m_u(j, i, k) = xub.b0(j, i, k);
  for (int k = 1; k <= kz; ++k)
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j) {
        m_t(j, i, k) = xtb.b0(j, i, k);
        m_qx(j, i, k, iqv) = xqb.b0(j, i, k);
      }
  if (ipptls == 1) {   // .not. is_present_qc
    for (int k = 1; k <= kz; ++k)
      for (int i = ice1; i <= ice2; ++i)
        for (int j = jce1; j <= jce2; ++j) m_qx(j, i, k, iqc) = minqc;
  }
  for (int k = 1; k <= kz; ++k)
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j) {
        m_tvirt(j, i, k) = m_t(j, i, k) * (d_one + ep1 * m_qx(j, i, k, iqv));
        m_pai(j, i, k) = xpaib.b0(j, i, k);
      }
  for (int k = 1; k <= kz; ++k)
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j) m_p(j, i, k) = std::pow(m_pai(j, i, k), cpovr) * p00;
  const double zz1 = -egrav * hzita * bzita(d_half * mo_dz) * std::log(d_one - d_half * mo_dz / hzita);
  for (int i = ice1; i <= ice2; ++i)
    for (int j = jce1; j <= jce2; ++j) {
      double zdgz = ht(j, i) * (gzita(d_half * mo_dz) - d_one) + zz1;
      psa(j, i) = m_p(j, i, kz) * std::exp(zdgz / (rgas * m_tvirt(j, i, kz)));
    }
  for (int k = 1; k <= kz; ++k)
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j) {
        double qs = pfwsat(m_t(j, i, k), m_p(j, i, k));
        m_qx(j, i, k, iqv) = std::min(m_qx(j, i, k, iqv), qs * 1.001);   // remove excessive supersaturation
      }
  for (int k = 2; k <= kz - 1; ++k)
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j) {
        double zuh = xub.b0(j, i, k) * hx(j, i) + xub.b0(j + 1, i, k) * hx(j + 1, i) +
                     xub.b0(j, i, k + 1) * hx(j, i) + xub.b0(j + 1, i, k + 1) * hx(j + 1, i);
        double zvh = xvb.b0(j, i, k) * hy(j, i) + xvb.b0(j, i + 1, k) * hy(j, i + 1) +
                     xvb.b0(j, i, k + 1) * hy(j, i) + xvb.b0(j, i + 1, k + 1) * hy(j, i + 1);
        m_w(j, i, k) = d_rfour * (zuh + zvh);
      }
  for (int i = ice1; i <= ice2; ++i)
    for (int j = jce1; j <= jce2; ++j) {
      double zuh = xub.b0(j, i, kz) * hx(j, i) + xub.b0(j + 1, i, kz) * hx(j + 1, i);
      double zvh = xvb.b0(j, i, kz) * hy(j, i) + xvb.b0(j, i + 1, kz) * hy(j, i + 1);
      m_w(j, i, kz) = d_half * (zuh + zvh);
    }
  // :876-903
  for (int k = 1; k <= kz; ++k)
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j) {
        m_p(j, i, k) = std::pow(m_pai(j, i, k), cpovr) * p00;
        m_tvirt(j, i, k) = m_t(j, i, k) * (d_one + ep1 * m_qx(j, i, k, iqv));
      }
  // :998-1013
  if (cfg.mo_nzfilt > 0) {
    const int ntop = (int)(0.08 * (double)kz);
    const double zfilt = (double)(kzp1 - ntop + cfg.mo_nzfilt) * mo_dz;
    for (int k = 1; k <= kz; ++k) {
      if (k > ntop + cfg.mo_nzfilt - 1) ffilt(k) = d_zero;
      else {
        double zzi = (mo_dz * (double)(kzp1 - k) - zfilt) / (hzita - zfilt);
        double sn = std::sin(d_half * mathpi * zzi);
        ffilt(k) = 0.8 * (sn * sn);
      }
    }
  } else {
    for (int k = 1; k <= kz; ++k) ffilt(k) = d_zero;
  }
}

// Main/mod_moloch.F90:159-220
void Oracle::init_moloch() {
  if (cfg.ifrayd == 1) { xtoustag(m_zeta, m_zetau); xtovstag(m_zeta, m_zetav); }
  for (size_t n = 0; n < mx2.d.size(); ++n) {
    mx2.d[n] = msfx.d[n] * msfx.d[n];
    rmu.d[n] = d_one / msfu.d[n];
    rmv.d[n] = d_one / msfv.d[n];
  }
  for (int k = 1; k <= kzp1; ++k) gzitak(k) = gzita(zita(k));
  for (int k = 1; k <= kz; ++k) gzitakh(k) = gzita(zitah(k));
  m_deltaw.fill(d_zero);
}

}  // namespace orc

namespace orc {

// ---- stagger operators, Main/mod_moloch.F90:1574-1794 ----
void Oracle::wstagtox(const A3& w, A3& wx) {
  for (int k = 2; k <= kz - 1; ++k)
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j)
        wx(j, i, k) = 0.5625 * (w(j, i, k + 1) + w(j, i, k)) - 0.0625 * (w(j, i, k + 2) + w(j, i, k - 1));
  for (int i = ice1; i <= ice2; ++i)
    for (int j = jce1; j <= jce2; ++j) {
      wx(j, i, 1) = 0.5 * (w(j, i, 2) + w(j, i, 1));
      wx(j, i, kz) = 0.5 * (w(j, i, kzp1) + w(j, i, kz));
    }
}

void Oracle::xtowstag(const A3& wx, A3& w) {
  for (int k = 3; k <= kz - 1; ++k)
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j)
        w(j, i, k) = 0.5625 * (wx(j, i, k) + wx(j, i, k - 1)) - 0.0625 * (wx(j, i, k + 1) + wx(j, i, k - 2));
  for (int i = ice1; i <= ice2; ++i)
    for (int j = jce1; j <= jce2; ++j) {
      w(j, i, 2) = 0.5 * (wx(j, i, 2) + wx(j, i, 1));
      w(j, i, kz) = 0.5 * (wx(j, i, kz) + wx(j, i, kz - 1));
    }
}

void Oracle::xtoustag(const A3& ux, A3& u) {
  for (int k = 1; k <= kz; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jdii1; j <= jdii2; ++j)
        u(j, i, k) = 0.5625 * (ux(j, i, k) + ux(j - 1, i, k)) - 0.0625 * (ux(j + 1, i, k) + ux(j - 2, i, k));
  for (int k = 1; k <= kz; ++k)
    for (int i = ici1; i <= ici2; ++i) u(jdi2, i, k) = 0.5 * (ux(jci2, i, k) + ux(jce2, i, k));
  for (int k = 1; k <= kz; ++k)
    for (int i = ici1; i <= ici2; ++i) u(jdi1, i, k) = 0.5 * (ux(jci1, i, k) + ux(jce1, i, k));
}

void Oracle::xtovstag(const A3& vx, A3& v) {
  for (int k = 1; k <= kz; ++k)
    for (int i = idii1; i <= idii2; ++i)
      for (int j = jci1; j <= jci2; ++j)
        v(j, i, k) = 0.5625 * (vx(j, i, k) + vx(j, i - 1, k)) - 0.0625 * (vx(j, i + 1, k) + vx(j, i - 2, k));
  for (int k = 1; k <= kz; ++k)
    for (int j = jci1; j <= jci2; ++j) v(j, idi2, k) = 0.5 * (vx(j, ici2, k) + vx(j, ice2, k));
  for (int k = 1; k <= kz; ++k)
    for (int j = jci1; j <= jci2; ++j) v(j, idi1, k) = 0.5 * (vx(j, ici1, k) + vx(j, ice1, k));
}

void Oracle::xtouvstag() { xtoustag(m_ux, m_u); xtovstag(m_vx, m_v); }

void Oracle::uvstagtox() {
  A3 &u = m_u, &v = m_v, &ux = m_ux, &vx = m_vx;
  for (int k = 1; k <= kz; ++k)
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jci1; j <= jci2; ++j)
        ux(j, i, k) = 0.5625 * (u(j + 1, i, k) + u(j, i, k)) - 0.0625 * (u(j + 2, i, k) + u(j - 1, i, k));
  for (int k = 1; k <= kz; ++k)
    for (int i = ice1; i <= ice2; ++i) ux(jce1, i, k) = 0.5 * (u(jde1, i, k) + u(jdi1, i, k));
  for (int k = 1; k <= kz; ++k)
    for (int i = ice1; i <= ice2; ++i) ux(jce2, i, k) = 0.5 * (u(jde2, i, k) + u(jdi2, i, k));
  for (int k = 1; k <= kz; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jce1; j <= jce2; ++j)
        vx(j, i, k) = 0.5625 * (v(j, i + 1, k) + v(j, i, k)) - 0.0625 * (v(j, i + 2, k) + v(j, i - 1, k));
  for (int k = 1; k <= kz; ++k)
    for (int j = jce1; j <= jce2; ++j) vx(j, ice1, k) = 0.5 * (v(j, ide1, k) + v(j, idi1, k));
  for (int k = 1; k <= kz; ++k)
    for (int j = jce1; j <= jce2; ++j) vx(j, ice2, k) = 0.5 * (v(j, ide2, k) + v(j, idi2, k));
}

// ---- Main/mod_moloch.F90:549-569 ----
void Oracle::mo_filt3d() {
  for (int k = 1; k <= kz; ++k) {
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j)
        p2d(j, i) = 0.125 * (m_zdiv2(j - 1, i, k) + m_zdiv2(j + 1, i, k) + m_zdiv2(j, i - 1, k) + m_zdiv2(j, i + 1, k)) -
                    0.5 * m_zdiv2(j, i, k);
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) m_zdiv2(j, i, k) = m_zdiv2(j, i, k) + cfg.mo_anu2 * p2d(j, i);
  }
}

// ---- Main/mod_moloch.F90:708-1040 (do_fulleq = .false.; both projection branches) ----
void Oracle::mo_sound(double dts) {
  A3 &u = m_u, &v = m_v, &w = m_w, &s = m_s, &zdiv2 = m_zdiv2, &deltaw = m_deltaw, &wwkw = m_wwkw, &pai = m_pai,
     &tetav = m_tetav, &fmz = m_fmz, &fmzf = m_fmzf, &ud = m_ud, &vd = m_vd;
  const double zdtrdx = dts / dx, zdtrdy = dts / dx, zdtrdz = dts / mo_dz;
  const double zcs2 = (zdtrdz * zdtrdz) * rdrcv;
  const bool rot = cfg.mo_rotllr != 0;   // iproj == 'ROTLLR'
  for (int jsound = 1; jsound <= nsound; ++jsound) {
    ud.d = u.d; vd.d = v.d;
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j) {
        double zuh = u(j, i, kz) * hx(j, i) + u(j + 1, i, kz) * hx(j + 1, i);
        double zvh = v(j, i, kz) * hy(j, i) + v(j, i + 1, kz) * hy(j, i + 1);
        w(j, i, kzp1) = d_half * (zuh + zvh);
      }
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j) s(j, i, kzp1) = -w(j, i, kzp1);
    for (int k = kz; k >= 2; --k)
      for (int i = ice1; i <= ice2; ++i)
        for (int j = jce1; j <= jce2; ++j) {
          double zuh = (u(j, i, k) + u(j, i, k - 1)) * hx(j, i) + (u(j + 1, i, k) + u(j + 1, i, k - 1)) * hx(j + 1, i);
          double zvh = (v(j, i, k) + v(j, i, k - 1)) * hy(j, i) + (v(j, i + 1, k) + v(j, i + 1, k - 1)) * hy(j, i + 1);
          s(j, i, k) = -0.25 * (zuh + zvh) * gzitak(k);
        }
    for (int k = 1; k <= kz; ++k)
      for (int i = ice1; i <= ice2; ++i) {
        const int im1 = std::max(icross1, i - 1), ip1 = std::min(icross2, i + 1);
        for (int j = jce1; j <= jce2; ++j) {
          const int jm1 = std::max(jcross1, j - 1), jp1 = std::min(jcross2, j + 1);
          double zrfmzum = d_two / (fmz(j, i, k) + fmz(jm1, i, k));
          double zrfmzvm = d_two / (fmz(j, i, k) + fmz(j, im1, k));
          double zrfmzup = d_two / (fmz(j, i, k) + fmz(jp1, i, k));
          double zrfmzvp = d_two / (fmz(j, i, k) + fmz(j, ip1, k));
          if (rot) {   // :775-793
            double zum = zdtrdx * u(j, i, k) * zrfmzum;
            double zup = zdtrdx * u(j + 1, i, k) * zrfmzup;
            double zvm = zdtrdy * v(j, i, k) * zrfmzvm * rmv(j, i);
            double zvp = zdtrdy * v(j, i + 1, k) * zrfmzvp * rmv(j, i + 1);
            zdiv2(j, i, k) = (zup - zum) + (zvp - zvm);
          } else {
            double zum = u(j, i, k) * rmu(j, i) * zrfmzum;
            double zup = u(j + 1, i, k) * rmu(j + 1, i) * zrfmzup;
            double zvm = v(j, i, k) * rmv(j, i) * zrfmzvm;
            double zvp = v(j, i + 1, k) * rmv(j, i + 1) * zrfmzvp;
            zdiv2(j, i, k) = (zup - zum) * zdtrdx + (zvp - zvm) * zdtrdy;
          }
        }
      }
    mo_filt3d();
    for (int k = 1; k <= kz; ++k)
      for (int i = ice1; i <= ice2; ++i)
        for (int j = jce1; j <= jce2; ++j)
          zdiv2(j, i, k) = fmz(j, i, k) * ((rot ? msfx(j, i) : mx2(j, i)) * zdiv2(j, i, k) + zdtrdz * (s(j, i, k) - s(j, i, k + 1)));
    // new w (implicit scheme), Equation 19
    for (int k = kz; k >= 2; --k)
      for (int i = ice1; i <= ice2; ++i)
        for (int j = jce1; j <= jce2; ++j) {
          deltaw(j, i, k) = -w(j, i, k);
          double zrom1w = d_half * cpd * fmzf(j, i, k) * (tetav(j, i, k - 1) + tetav(j, i, k));
          zrom1w = zrom1w - cpd * w(j, i, k) * (fmzf(j, i, k) * fmzf(j, i, k)) * (double)jsound * zdtrdz *
                                (tetav(j, i, k - 1) - tetav(j, i, k));
          double zwexpl = w(j, i, k) - zrom1w * zdtrdz * (pai(j, i, k - 1) - pai(j, i, k)) - egrav * dts;
          zwexpl = zwexpl + rdrcv * zrom1w * zdtrdz * (pai(j, i, k - 1) * zdiv2(j, i, k - 1) - pai(j, i, k) * zdiv2(j, i, k));
          double zp = zcs2 * fmz(j, i, k - 1) * zrom1w * pai(j, i, k - 1) + ffilt(k);
          double zm = zcs2 * fmz(j, i, k) * zrom1w * pai(j, i, k) + ffilt(k);
          double zrapp = d_one / (d_one + zm + zp - zm * wwkw(j, i, k + 1));
          w(j, i, k) = zrapp * (zwexpl + zm * w(j, i, k + 1));
          wwkw(j, i, k) = zrapp * zp;
        }
    for (int k = 2; k <= kz; ++k)
      for (int i = ice1; i <= ice2; ++i)
        for (int j = jce1; j <= jce2; ++j) {
          w(j, i, k) = w(j, i, k) + wwkw(j, i, k) * w(j, i, k - 1);
          deltaw(j, i, k) = deltaw(j, i, k) + w(j, i, k);
        }
    // new Exner function
    for (int k = 1; k <= kz; ++k)
      for (int i = ici1; i <= ici2; ++i)
        for (int j = jci1; j <= jci2; ++j)
          zdiv2(j, i, k) = zdiv2(j, i, k) + zdtrdz * fmz(j, i, k) * (w(j, i, k) - w(j, i, k + 1));
    for (int k = 1; k <= kz; ++k)
      for (int i = ici1; i <= ici2; ++i)
        for (int j = jci1; j <= jci2; ++j) pai(j, i, k) = pai(j, i, k) * (d_one - rdrcv * zdiv2(j, i, k));
    // horizontal momentum equations
    for (int k = 1; k <= kz; ++k)
      for (int i = ici1; i <= ici2; ++i)
        for (int j = jdii1; j <= jdii2; ++j) {
          double zcx = zdtrdx * msfu(j, i);
          double zfz = 0.25 * (deltaw(j - 1, i, k) + deltaw(j - 1, i, k + 1) + deltaw(j, i, k) + deltaw(j, i, k + 1)) + egrav * dts;
          double zrom1u = d_half * cpd * (tetav(j - 1, i, k) + tetav(j, i, k));
          double zcor1u = coriou(j, i) * 0.25 * (vd(j, i, k) + vd(j - 1, i, k) + vd(j, i + 1, k) + vd(j - 1, i + 1, k));
          u(j, i, k) = u(j, i, k) + zcor1u * dts - zfz * hx(j, i) * gzitakh(k) - zcx * zrom1u * (pai(j, i, k) - pai(j - 1, i, k));
        }
    for (int k = 1; k <= kz; ++k)
      for (int i = idii1; i <= idii2; ++i)
        for (int j = jci1; j <= jci2; ++j) {
          double zcy = rot ? zdtrdy : zdtrdy * msfv(j, i);   // :956 / :997
          double zfz = 0.25 * (deltaw(j, i - 1, k) + deltaw(j, i - 1, k + 1) + deltaw(j, i, k) + deltaw(j, i, k + 1)) + egrav * dts;
          double zrom1v = d_half * cpd * (tetav(j, i - 1, k) + tetav(j, i, k));
          double zcor1v = coriov(j, i) * 0.25 * (ud(j, i, k) + ud(j + 1, i, k) + ud(j, i - 1, k) + ud(j + 1, i - 1, k));
          v(j, i, k) = v(j, i, k) - zcor1v * dts - zfz * hy(j, i) * gzitakh(k) - zcy * zrom1v * (pai(j, i, k) - pai(j, i - 1, k));
        }
  }
  // complete computation of generalized vertical velocity
  for (int k = 2; k <= kz; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) s(j, i, k) = (w(j, i, k) + s(j, i, k)) * fmzf(j, i, k);
  for (int i = ici1; i <= ici2; ++i)
    for (int j = jci1; j <= jci2; ++j) { s(j, i, 1) = d_zero; s(j, i, kzp1) = d_zero; }
}

static inline double rdeno(double t1, double t2, double t3, double t4) {
  const double minden = 1.0e-15;
  const double zzden = (t3 - t4);
  return (t1 - t2) / std::copysign(std::max(std::fabs(zzden), minden), zzden);
}

// ---- Main/mod_moloch.F90:1098-1361 (ROTLLR branch :1183-1270, generic branch :1272-1358) ----
void Oracle::wafone(A3& pp, double dta) {
  A3 &s = m_s, &fmz = m_fmz, &fmzf = m_fmzf, &wz = m_wz, &p0 = m_p0, &u = m_u, &v = m_v;
  A2 &wfw = m_wfw, &zpby = m_zpby, &zpbw = m_zpbw;
  const double zdtrdx = dta / dx, zdtrdy = dta / dx, zdtrdz = dta / mo_dz;
  for (int j = jci1; j <= jci2; ++j) { wfw(j, 1) = d_zero; wfw(j, kzp1) = d_zero; }
  for (int i = ici1; i <= ici2; ++i) {
    for (int k = 1; k <= kz - 1; ++k)
      for (int j = jci1; j <= jci2; ++j) {
        double zamu = s(j, i, k + 1) * zdtrdz, is;
        int k1, k1m1;
        if (zamu >= d_zero) { is = d_one; k1 = k + 1; k1m1 = k1 + 1; if (k1m1 > kz) k1m1 = kz; }
        else { is = -d_one; k1m1 = k; k1 = k - 1; if (k1 < 1) k1 = 1; }
        double r = rdeno(pp(j, i, k1), pp(j, i, k1m1), pp(j, i, k), pp(j, i, k + 1));
        double b = std::max(d_zero, std::min(d_two, std::max(r, std::min(d_two * r, d_one))));
        double zphi = is + zamu * b - is * b;
        wfw(j, k + 1) = d_half * s(j, i, k + 1) * ((d_one + zphi) * pp(j, i, k + 1) + (d_one - zphi) * pp(j, i, k));
      }
    for (int j = jci1; j <= jci2; ++j) {
      double zrfmm = zdtrdz * fmz(j, i, 1) / fmzf(j, i, 2);
      double zdv = -s(j, i, 2) * zrfmm * pp(j, i, 1);
      wz(j, i, 1) = pp(j, i, 1) + wfw(j, 2) * zrfmm + zdv;
    }
    for (int k = 2; k <= kz; ++k)
      for (int j = jci1; j <= jci2; ++j) {
        double zrfmm = zdtrdz * fmz(j, i, k) / fmzf(j, i, k + 1);
        double zrfmp = zdtrdz * fmz(j, i, k) / fmzf(j, i, k);
        double zdv = (s(j, i, k) * zrfmp - s(j, i, k + 1) * zrfmm) * pp(j, i, k);
        wz(j, i, k) = pp(j, i, k) - wfw(j, k) * zrfmp + wfw(j, k + 1) * zrfmm + zdv;
      }
  }
  for (int k = 1; k <= kz; ++k)
    for (int j = jci1; j <= jci2; ++j) { wz(j, ice1, k) = wz(j, ici1, k); wz(j, ice2, k) = wz(j, ici2, k); }
  if (cfg.mo_rotllr) {   // iproj == 'ROTLLR' (:1183-1270)
    const A2& mu = msfu; const A2& mx = msfx;
    for (int k = 1; k <= kz; ++k) {
      for (int i = ici1; i <= ice2; ++i)
        for (int j = jci1; j <= jci2; ++j) {
          double zamu = v(j, i, k) * zdtrdy, is;
          int ih;
          if (zamu > d_zero) { is = d_one; ih = i - 1; }
          else { is = -d_one; ih = std::min(i + 1, icross2); }
          const int ihm1 = std::max(ih - 1, icross1);
          double r = rdeno(wz(j, ih, k), wz(j, ihm1, k), wz(j, i, k), wz(j, i - 1, k));
          double b = std::max(d_zero, std::min(d_two, std::max(r, std::min(d_two * r, d_one))));
          double zphi = is + zamu * b - is * b;
          zpby(j, i) = d_half * v(j, i, k) * ((d_one + zphi) * wz(j, i - 1, k) + (d_one - zphi) * wz(j, i, k));
        }
      for (int i = ici1; i <= ici2; ++i)
        for (int j = jci1; j <= jci2; ++j) {
          double zhxvtn = zdtrdy * rmv(j, i + 1) * mx(j, i);
          double zhxvts = zdtrdy * rmv(j, i) * mx(j, i);
          double zrfmn = zhxvtn * d_two * fmz(j, i, k) / (fmz(j, i, k) + fmz(j, i + 1, k));
          double zrfms = zhxvts * d_two * fmz(j, i, k) / (fmz(j, i, k) + fmz(j, i - 1, k));
          double zdv = (v(j, i + 1, k) * zrfmn - v(j, i, k) * zrfms) * pp(j, i, k);
          p0(j, i, k) = wz(j, i, k) + zpby(j, i) * zrfms - zpby(j, i + 1) * zrfmn + zdv;
        }
    }
    for (int k = 1; k <= kz; ++k)
      for (int i = ici1; i <= ici2; ++i) { p0(jce1, i, k) = p0(jci1, i, k); p0(jce2, i, k) = p0(jci2, i, k); }
    for (int k = 1; k <= kz; ++k) {
      for (int i = ici1; i <= ici2; ++i)
        for (int j = jci1; j <= jce2; ++j) {
          double zamu = u(j, i, k) * mu(j, i) * zdtrdx, is;
          int jh;
          if (zamu > d_zero) { is = d_one; jh = j - 1; }
          else { is = -d_one; jh = std::min(j + 1, jcross2); }
          const int jhm1 = std::max(jh - 1, jcross1);
          double r = rdeno(p0(jh, i, k), p0(jhm1, i, k), p0(j, i, k), p0(j - 1, i, k));
          double b = std::max(d_zero, std::min(d_two, std::max(r, std::min(d_two * r, d_one))));
          double zphi = is + zamu * b - is * b;
          zpbw(j, i) = d_half * u(j, i, k) * ((d_one + zphi) * p0(j - 1, i, k) + (d_one - zphi) * p0(j, i, k));
        }
      for (int i = ici1; i <= ici2; ++i)
        for (int j = jci1; j <= jci2; ++j) {
          double zcostx = zdtrdx * mu(j, i);
          double zrfme = zcostx * d_two * fmz(j, i, k) / (fmz(j, i, k) + fmz(j + 1, i, k));
          double zrfmw = zcostx * d_two * fmz(j, i, k) / (fmz(j, i, k) + fmz(j - 1, i, k));
          double zdv = (u(j + 1, i, k) * zrfme - u(j, i, k) * zrfmw) * pp(j, i, k);
          pp(j, i, k) = p0(j, i, k) + zpbw(j, i) * zrfmw - zpbw(j + 1, i) * zrfme + zdv;
        }
    }
    return;
  }
  // Meridional advection
  for (int k = 1; k <= kz; ++k) {
    for (int i = ici1; i <= ice2; ++i)
      for (int j = jci1; j <= jci2; ++j) {
        double zamu = v(j, i, k) * rmv(j, i) * zdtrdy, is;
        int ih;
        if (zamu > d_zero) { is = d_one; ih = i - 1; }
        else { is = -d_one; ih = std::min(i + 1, icross2); }
        const int ihm1 = std::max(ih - 1, icross1);
        double r = rdeno(wz(j, ih, k), wz(j, ihm1, k), wz(j, i, k), wz(j, i - 1, k));
        double b = std::max(d_zero, std::min(d_two, std::max(r, std::min(d_two * r, d_one))));
        double zphi = is + zamu * b - is * b;
        zpby(j, i) = d_half * v(j, i, k) * rmv(j, i) * ((d_one + zphi) * wz(j, i - 1, k) + (d_one - zphi) * wz(j, i, k));
      }
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) {
        double zrfmn = zdtrdy * d_two * fmz(j, i, k) / (fmz(j, i, k) + fmz(j, i + 1, k));
        double zrfms = zdtrdy * d_two * fmz(j, i, k) / (fmz(j, i, k) + fmz(j, i - 1, k));
        double zdv = (v(j, i + 1, k) * rmv(j, i + 1) * zrfmn - v(j, i, k) * rmv(j, i) * zrfms) * pp(j, i, k);
        p0(j, i, k) = wz(j, i, k) + mx2(j, i) * (zpby(j, i) * zrfms - zpby(j, i + 1) * zrfmn + zdv);
      }
  }
  for (int k = 1; k <= kz; ++k)
    for (int i = ici1; i <= ici2; ++i) { p0(jce1, i, k) = p0(jci1, i, k); p0(jce2, i, k) = p0(jci2, i, k); }
  // Zonal advection
  for (int k = 1; k <= kz; ++k) {
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jce2; ++j) {
        double zamu = u(j, i, k) * rmu(j, i) * zdtrdx, is;
        int jh;
        if (zamu > d_zero) { is = d_one; jh = j - 1; }
        else { is = -d_one; jh = std::min(j + 1, jcross2); }
        const int jhm1 = std::max(jh - 1, jcross1);
        double r = rdeno(p0(jh, i, k), p0(jhm1, i, k), p0(j, i, k), p0(j - 1, i, k));
        double b = std::max(d_zero, std::min(d_two, std::max(r, std::min(d_two * r, d_one))));
        double zphi = is + zamu * b - is * b;
        zpbw(j, i) = d_half * u(j, i, k) * rmu(j, i) * ((d_one + zphi) * p0(j - 1, i, k) + (d_one - zphi) * p0(j, i, k));
      }
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) {
        double zrfme = zdtrdx * d_two * fmz(j, i, k) / (fmz(j, i, k) + fmz(j + 1, i, k));
        double zrfmw = zdtrdx * d_two * fmz(j, i, k) / (fmz(j, i, k) + fmz(j - 1, i, k));
        double zdv = (u(j + 1, i, k) * rmu(j + 1, i) * zrfme - u(j, i, k) * rmu(j, i) * zrfmw) * pp(j, i, k);
        pp(j, i, k) = p0(j, i, k) + mx2(j, i) * (zpbw(j, i) * zrfmw - zpbw(j + 1, i) * zrfme + zdv);
      }
  }
}

void Oracle::wafone4(A4& q, int n, double dta) {
  // assignpnt(qx,ptr,n): operate on slice n through a temporary of the same shape
  A3 tmp; tmp.alloc(q.l1, q.l1 + q.n1 - 1, q.l2, q.l2 + q.n2 - 1, q.l3, q.l3 + q.n3 - 1);
  const size_t sz = tmp.d.size();
  std::copy(q.d.begin() + (size_t)(n - q.l4) * sz, q.d.begin() + (size_t)(n - q.l4 + 1) * sz, tmp.d.begin());
  wafone(tmp, dta);
  std::copy(tmp.d.begin(), tmp.d.end(), q.d.begin() + (size_t)(n - q.l4) * sz);
}

// Main/mod_moloch.F90:1042-1088
void Oracle::mo_advection(double dta) {
  uvstagtox();
  wstagtox(m_w, m_wx);
  wafone(m_tetav, dta);
  wafone(m_pai, dta);
  wafone(m_ux, dta);
  wafone(m_vx, dta);
  wafone(m_wx, dta);
  wafone4(m_qx, iqv, dta);
  if (cfg.ipptls > 0)
    for (int n = iqfrst; n <= iqlst; ++n) wafone4(m_qx, n, dta);
  xtouvstag();
  xtowstag(m_wx, m_w);
}

}  // namespace orc

namespace orc {

// ---- Main/mod_bdycod.F90:4408-4595 ----
void Oracle::monudge3d(int ibdy, A3& f, const V3dBound& bnd) {
  if (!ba_cr.havebound) return;
  const int ns = f.l3, nk = f.u3();
  const double xt = xbctime + dt;
  for (int k = ns; k <= nk; ++k)
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j) fg1(j, i, k) = (bnd.b0(j, i, k) + xt * bnd.bt(j, i, k)) - f(j, i, k);
  const I2* masks[4] = {&ba_cr.bsouth, &ba_cr.bnorth, &ba_cr.bwest, &ba_cr.beast};
  const int cnt[4] = {ba_cr.ns, ba_cr.nn, ba_cr.nw, ba_cr.ne};
  for (int m = 0; m < 4; ++m) {
    if (cnt[m] == 0) continue;
    for (int k = ns; k <= nk; ++k)
      for (int i = ici1; i <= ici2; ++i)
        for (int j = jci1; j <= jci2; ++j) {
          if (!(*masks[m])(j, i)) continue;
          const int ib = ba_cr.ibnd(j, i);
          double xf, xg;
          if (ibdy == 1) { xf = fcx(ib); xg = gcx(ib); }
          else { xf = hefc(ib, std::min(k, kz)); xg = hegc(ib, std::min(k, kz)); }
          double fls0 = fg1(j, i, k), fls1 = fg1(j - 1, i, k), fls2 = fg1(j + 1, i, k), fls3 = fg1(j, i - 1, k),
                 fls4 = fg1(j, i + 1, k);
          f(j, i, k) = f(j, i, k) + xf * fls0 - xg * (fls1 + fls2 + fls3 + fls4 - d_four * fls0);
        }
  }
}

// qv is a 3-D pointer into qx in the reference (assignpnt); same arithmetic as monudge3d
void Oracle::monudge4d3d(int ibdy, A4& q, const V3dBound& bnd, int n) {
  A3 tmp; tmp.alloc(q.l1, q.l1 + q.n1 - 1, q.l2, q.l2 + q.n2 - 1, q.l3, q.l3 + q.n3 - 1);
  const size_t sz = tmp.d.size();
  std::copy(q.d.begin() + (size_t)(n - q.l4) * sz, q.d.begin() + (size_t)(n - q.l4 + 1) * sz, tmp.d.begin());
  monudge3d(ibdy, tmp, bnd);
  std::copy(tmp.d.begin(), tmp.d.end(), q.d.begin() + (size_t)(n - q.l4) * sz);
}

// ---- Main/mod_bdycod.F90:3825-4165 (loop bounds and coefficient tables exactly
// as written there, including the single-row south loop of the linear branch) ----
void Oracle::monudgeuv(int ibdy, A3& fu, A3& fv, const V3dBound& bu, const V3dBound& bv) {
  if (!ba_ut.havebound && !ba_vt.havebound) return;
  const double xt = xbctime + dt;
  for (int k = 1; k <= kz; ++k)
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jde1; j <= jde2; ++j) fg1(j, i, k) = (bu.b0(j, i, k) + xt * bu.bt(j, i, k)) - fu(j, i, k);
  for (int k = 1; k <= kz; ++k)
    for (int i = ide1; i <= ide2; ++i)
      for (int j = jce1; j <= jce2; ++j) fg2(j, i, k) = (bv.b0(j, i, k) + xt * bv.bt(j, i, k)) - fv(j, i, k);
  struct Pass { bool isu; const I2* mask; int cnt; int i1, i2, j1, j2; bool dcoef; };
  std::vector<Pass> P;
  if (ibdy == 1) {
    P = {{true, &ba_ut.bsouth, ba_ut.ns, ici1, ici1, jdi1, jdi2, false},
         {false, &ba_vt.bsouth, ba_vt.ns, idi1, idi2, jci1, jci2, true},
         {true, &ba_ut.bnorth, ba_ut.nn, ici1, ici2, jdi1, jdi2, false},
         {false, &ba_vt.bnorth, ba_vt.nn, idi1, idi2, jci1, jci2, true},
         {true, &ba_ut.bwest, ba_ut.nw, ici1, ici2, jdi1, jdi2, true},
         {false, &ba_vt.bwest, ba_vt.nw, idi1, idi2, jci1, jci2, false},
         {true, &ba_ut.beast, ba_ut.ne, ici1, ici2, jdi1, jdi2, true},
         {false, &ba_vt.beast, ba_vt.ne, idi1, idi2, jci1, jci2, false}};
  } else {
    P = {{true, &ba_ut.bsouth, ba_ut.ns, ici1, ici2, jdi1, jdi2, false},
         {false, &ba_vt.bsouth, ba_vt.ns, idi1, idi2, jdi1, jdi2, true},
         {true, &ba_ut.bnorth, ba_ut.nn, idi1, idi2, jdi1, jdi2, false},
         {false, &ba_vt.bnorth, ba_vt.nn, idi1, idi2, jci1, jci2, true},
         {true, &ba_ut.bwest, ba_ut.nw, ici1, ici2, jdi1, jdi2, true},
         {false, &ba_vt.bwest, ba_vt.nw, idi1, idi2, jci1, jci2, false},
         {true, &ba_ut.beast, ba_ut.ne, ici1, ici2, jdi1, jdi2, true},
         {false, &ba_vt.beast, ba_vt.ne, idi1, idi2, jci1, jci2, false}};
  }
  for (const Pass& p : P) {
    if (p.cnt == 0) continue;
    A3& f = p.isu ? fu : fv;
    A3& fg = p.isu ? fg1 : fg2;
    for (int k = 1; k <= kz; ++k)
      for (int i = p.i1; i <= p.i2; ++i)
        for (int j = p.j1; j <= p.j2; ++j) {
          if (!(*p.mask)(j, i)) continue;
          const int ib = (p.isu ? ba_ut : ba_vt).ibnd(j, i);
          double xf, xg;
          if (ibdy == 1) { xf = p.dcoef ? fcd(ib) : fcx(ib); xg = p.dcoef ? gcd(ib) : gcx(ib); }
          else { xf = p.dcoef ? hefd(ib, k) : hefc(ib, k); xg = p.dcoef ? hegd(ib, k) : hegc(ib, k); }
          double fls0 = fg(j, i, k), fls1 = fg(j - 1, i, k), fls2 = fg(j + 1, i, k), fls3 = fg(j, i - 1, k),
                 fls4 = fg(j, i + 1, k);
          f(j, i, k) = f(j, i, k) + xf * fls0 - xg * (fls1 + fls2 + fls3 + fls4 - d_four * fls0);
        }
  }
}

// ---- sponge, iboudy = 4: mosponge3d (Main/mod_bdycod.F90:2996-3063), mosponge4d (:2665-2733),
// mospongeuv (:2815-2925) ----
void Oracle::mosponge3d(A3& f, const V3dBound& bnd) {
  if (!ba_cr.havebound) return;
  const int nk = f.n3;
  const I2* masks[4] = {&ba_cr.bsouth, &ba_cr.bnorth, &ba_cr.bwest, &ba_cr.beast};
  const int cnt[4] = {ba_cr.ns, ba_cr.nn, ba_cr.nw, ba_cr.ne};
  for (int q = 0; q < 4; ++q) {
    if (cnt[q] == 0) continue;
    for (int k = 1; k <= nk; ++k)
      for (int i = ici1; i <= ici2; ++i)
        for (int j = jci1; j <= jci2; ++j) {
          if (!(*masks[q])(j, i)) continue;
          const int ib = ba_cr.ibnd(j, i);
          f(j, i, k) = f(j, i, k) + (d_one - wgtx(ib)) * bnd.bt(j, i, k);
        }
  }
}

void Oracle::mosponge4d3d(A4& q, const V3dBound& bnd, int n) {
  A3 tmp; tmp.alloc(q.l1, q.l1 + q.n1 - 1, q.l2, q.l2 + q.n2 - 1, q.l3, q.l3 + q.n3 - 1);
  const size_t sz = tmp.d.size();
  std::copy(q.d.begin() + (size_t)(n - q.l4) * sz, q.d.begin() + (size_t)(n - q.l4 + 1) * sz, tmp.d.begin());
  mosponge3d(tmp, bnd);
  std::copy(tmp.d.begin(), tmp.d.end(), q.d.begin() + (size_t)(n - q.l4) * sz);
}

void Oracle::mospongeuv(A3& fu, A3& fv, const V3dBound& bu, const V3dBound& bv) {
  if (!ba_ut.havebound && !ba_vt.havebound) return;
  struct Pass { bool isu; const I2* mask; int cnt; bool dw; };
  const Pass P[8] = {{true, &ba_ut.bsouth, ba_ut.ns, false}, {false, &ba_vt.bsouth, ba_vt.ns, true},
                     {true, &ba_ut.bnorth, ba_ut.nn, false}, {false, &ba_vt.bnorth, ba_vt.nn, true},
                     {true, &ba_ut.bwest, ba_ut.nw, true},   {false, &ba_vt.bwest, ba_vt.nw, false},
                     {true, &ba_ut.beast, ba_ut.ne, true},   {false, &ba_vt.beast, ba_vt.ne, false}};
  for (const Pass& p : P) {
    if (p.cnt == 0) continue;
    A3& f = p.isu ? fu : fv;
    const V3dBound& b = p.isu ? bu : bv;
    const int i1 = p.isu ? ici1 : idi1, i2 = p.isu ? ici2 : idi2, j1 = p.isu ? jdi1 : jci1, j2 = p.isu ? jdi2 : jci2;
    for (int k = 1; k <= kz; ++k)
      for (int i = i1; i <= i2; ++i)
        for (int j = j1; j <= j2; ++j) {
          if (!(*p.mask)(j, i)) continue;
          const int ib = (p.isu ? ba_ut : ba_vt).ibnd(j, i);
          f(j, i, k) = f(j, i, k) + (d_one - (p.dw ? wgtd(ib) : wgtx(ib))) * b.bt(j, i, k);
        }
  }
}

// ---- Main/mod_bdycod.F90:5047-5065 ----
void Oracle::moraydamp(const A3& z, A3& var, const V3dBound& bnd, int j1, int j2, int i1, int i2, int k1, int k2) {
  const double xt = xbctime + dt;
  const int k3 = std::max(k2, cfg.rayndamp);
  for (int k = k1; k <= k3; ++k)
    for (int i = i1; i <= i2; ++i)
      for (int j = j1; j <= j2; ++j) {
        double bval = bnd.b0(j, i, k) + xt * bnd.bt(j, i, k);
        var(j, i, k) = var(j, i, k) + dt * tau(z(j, i, k), z(j, i, 1)) * (bval - var(j, i, k));
      }
}

// ---- Main/mod_moloch.F90:485-547 ----
void Oracle::mo_boundary() {
  const int iboudy = cfg.iboudy;
  if (iboudy == 1 || iboudy == 5) {
    monudge3d(iboudy, m_pai, xpaib);
    monudge3d(iboudy, m_t, xtb);
    monudge4d3d(iboudy, m_qx, xqb, iqv);
    monudgeuv(iboudy, m_u, m_v, xub, xvb);
  } else if (iboudy == 4) {
    mosponge3d(m_pai, xpaib);
    mosponge3d(m_t, xtb);
    mosponge4d3d(m_qx, xqb, iqv);
    mospongeuv(m_u, m_v, xub, xvb);
  }
}

// ---- Main/mod_moloch.F90:224-481 (physics off) ----
void Oracle::moloch_step() {
  if (!mo_mid) moloch_head();
  mo_mid = false;
  moloch_tail();
}

// everything of `moloch` up to `call mkslice` (Main/mod_moloch.F90:296-393)
void Oracle::moloch_head() {
  A3 &t = m_t, &pai = m_pai, &tetav = m_tetav, &tvirt = m_tvirt, &p = m_p, &rho = m_rho;
  A4& qx = m_qx;
  const int ipptls = cfg.ipptls;
  const double dtstepa = dtsec / (double)nadv;
  const double dtsound = dtstepa / (double)nsound;
  for (int k = 1; k <= kz; ++k)
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j) {
        if (ipptls > 0) tvirt(j, i, k) = t(j, i, k) * (d_one + ep1 * qx(j, i, k, iqv) - qx(j, i, k, iqc));
        else tvirt(j, i, k) = t(j, i, k) * (d_one + ep1 * qx(j, i, k, iqv));
      }
  for (int k = 1; k <= kz; ++k)
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j) tetav(j, i, k) = tvirt(j, i, k) / pai(j, i, k);
  for (int jadv = 1; jadv <= nadv; ++jadv) {
    mo_sound(dtsound);
    mo_advection(dtstepa);
  }
  for (int k = 1; k <= kz; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) tvirt(j, i, k) = tetav(j, i, k) * pai(j, i, k);
  for (int k = 1; k <= kz; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) {
        if (ipptls > 0) t(j, i, k) = tvirt(j, i, k) / (d_one + ep1 * qx(j, i, k, iqv) - qx(j, i, k, iqc));
        else t(j, i, k) = tvirt(j, i, k) / (d_one + ep1 * qx(j, i, k, iqv));
      }
  for (int k = 1; k <= kz; ++k)
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j) {
        p(j, i, k) = std::pow(pai(j, i, k), cpovr) * p00;
        rho(j, i, k) = p(j, i, k) / (rgas * t(j, i, k));
      }
  const double zz1 = -egrav * hzita * bzita(d_half * mo_dz) * std::log(d_one - d_half * mo_dz / hzita);
  for (int i = ice1; i <= ice2; ++i)
    for (int j = jce1; j <= jce2; ++j) {
      double zdgz = ht(j, i) * (gzita(d_half * mo_dz) - d_one) + zz1;
      psa(j, i) = p(j, i, kz) * std::exp(zdgz / (rgas * tvirt(j, i, kz)));
    }
}

// physical_parametrizations' state update (:1548-1557, when tendencies were handed in), boundary, Rayleigh damping
void Oracle::moloch_tail() {
  if (mo_phys_on) {
    for (int k = 1; k <= kz; ++k)
      for (int i = ici1; i <= ici2; ++i)
        for (int j = jci1; j <= jci2; ++j) m_t(j, i, k) = m_t(j, i, k) + dtsec * m_tten(j, i, k);
    for (int k = 1; k <= kz; ++k)
      for (int i = ici1; i <= ici2; ++i)
        for (int j = jdi1; j <= jdi2; ++j) m_u(j, i, k) = m_u(j, i, k) + dtsec * m_uten(j, i, k);
    for (int k = 1; k <= kz; ++k)
      for (int i = idi1; i <= idi2; ++i)
        for (int j = jci1; j <= jci2; ++j) m_v(j, i, k) = m_v(j, i, k) + dtsec * m_vten(j, i, k);
    for (int n = 1; n <= nqx; ++n)
      for (int k = 1; k <= kz; ++k)
        for (int i = ici1; i <= ici2; ++i)
          for (int j = jci1; j <= jci2; ++j) m_qx(j, i, k, n) = m_qx(j, i, k, n) + dtsec * m_qxten(j, i, k, n);
    for (int n = iqfrst; n <= iqlst; ++n)
      for (int k = 1; k <= kz; ++k)
        for (int i = ici1; i <= ici2; ++i)
          for (int j = jci1; j <= jci2; ++j) m_qx(j, i, k, n) = std::max(m_qx(j, i, k, n), d_zero);
    for (int k = 1; k <= kz; ++k)
      for (int i = ici1; i <= ici2; ++i)
        for (int j = jci1; j <= jci2; ++j) m_qx(j, i, k, iqv) = std::max(m_qx(j, i, k, iqv), minqq);
  }
  // lateral / damping boundary condition (:417-426)
  mo_boundary();
  if (cfg.ifrayd == 1) {
    moraydamp(m_zetau, m_u, xub, jdi1, jdi2, ici1, ici2, 1, kz);
    moraydamp(m_zetav, m_v, xvb, jci1, jci2, idi1, idi2, 1, kz);
    moraydamp(m_zeta, m_t, xtb, jci1, jci2, ici1, ici2, 1, kz);
    moraydamp(m_zeta, m_pai, xpaib, jci1, jci2, ici1, ici2, 1, kz);
  }
  // :437-441
  ps_max = -1e300; ps_min = 1e300;
  for (int i = ici1; i <= ici2; ++i)
    for (int j = jci1; j <= jci2; ++j) { ps_max = std::max(ps_max, psa(j, i)); ps_min = std::min(ps_min, psa(j, i)); }
  timer_advance();
}

// ---- bdyval, idynamic 3 (Main/mod_bdycod.F90:1343-1385,1452-1491,1539-1575,
// 1662-1699,1963-2100); iboudy 0/1/5, no hydrometeor boundary data, bdyflow ----
void Oracle::bdyval_moloch() {
  const double xt = xbctime + dt;
  const bool fixed = (cfg.iboudy == 0);
  A4& qx = m_qx;
  auto bu = [&](int j, int i, int k) { return fixed ? xub.b0(j, i, k) : xub.b0(j, i, k) + xt * xub.bt(j, i, k); };
  auto bv = [&](int j, int i, int k) { return fixed ? xvb.b0(j, i, k) : xvb.b0(j, i, k) + xt * xvb.bt(j, i, k); };
  for (int k = 1; k <= kz; ++k) {
    for (int i = ice1; i <= ice2; ++i) m_u(jde1, i, k) = bu(jde1, i, k);
    for (int i = ide1; i <= ide2; ++i) m_v(jce1, i, k) = bv(jce1, i, k);
  }
  for (int k = 1; k <= kz; ++k) {
    for (int i = ice1; i <= ice2; ++i) m_u(jde2, i, k) = bu(jde2, i, k);
    for (int i = ide1; i <= ide2; ++i) m_v(jce2, i, k) = bv(jce2, i, k);
  }
  for (int k = 1; k <= kz; ++k) {
    for (int j = jde1; j <= jde2; ++j) m_u(j, ice1, k) = bu(j, ice1, k);
    for (int j = jce1; j <= jce2; ++j) m_v(j, ide1, k) = bv(j, ide1, k);
  }
  for (int k = 1; k <= kz; ++k) {
    for (int j = jde1; j <= jde2; ++j) m_u(j, ice2, k) = bu(j, ice2, k);
    for (int j = jce1; j <= jce2; ++j) m_v(j, ide2, k) = bv(j, ide2, k);
  }
  auto cell = [&](int j, int i, int k) {
    if (fixed) {
      m_t(j, i, k) = xtb.b0(j, i, k); qx(j, i, k, iqv) = xqb.b0(j, i, k); m_pai(j, i, k) = xpaib.b0(j, i, k);
    } else {
      m_t(j, i, k) = xtb.b0(j, i, k) + xt * xtb.bt(j, i, k);
      qx(j, i, k, iqv) = xqb.b0(j, i, k) + xt * xqb.bt(j, i, k);
      m_pai(j, i, k) = xpaib.b0(j, i, k) + xt * xpaib.bt(j, i, k);
    }
  };
  for (int k = 1; k <= kz; ++k) for (int i = ici1; i <= ici2; ++i) cell(jce1, i, k);
  for (int k = 1; k <= kz; ++k) for (int i = ici1; i <= ici2; ++i) cell(jce2, i, k);
  for (int k = 1; k <= kz; ++k) for (int j = jce1; j <= jce2; ++j) cell(j, ice1, k);
  for (int k = 1; k <= kz; ++k) for (int j = jce1; j <= jce2; ++j) cell(j, ice2, k);
  if (cfg.iboudy == 3 || cfg.iboudy == 4) {   // qv by inflow/outflow (:1813-1890)
    A3 &u = m_u, &v = m_v;
    for (int k = 1; k <= kz; ++k)
      for (int i = ici1; i <= ici2; ++i) {
        double qext = qx(jce1, i, k, iqv), qint = qx(jci1, i, k, iqv);
        double windavg = u(jde1, i, k) + u(jdi1, i, k);
        qx(jce1, i, k, iqv) = (windavg > d_zero) ? qext : qint;
      }
    for (int k = 1; k <= kz; ++k)
      for (int i = ici1; i <= ici2; ++i) {
        double qext = qx(jce2, i, k, iqv), qint = qx(jci2, i, k, iqv);
        double windavg = u(jde2, i, k) + u(jdi2, i, k);
        qx(jce2, i, k, iqv) = (windavg < d_zero) ? qext : qint;
      }
    for (int k = 1; k <= kz; ++k)
      for (int j = jce1; j <= jce2; ++j) {
        double qext = qx(j, ice1, k, iqv), qint = qx(j, ici1, k, iqv);
        double windavg = v(j, ide1, k) + v(j, idi1, k);
        qx(j, ice1, k, iqv) = (windavg > d_zero) ? qext : qint;
      }
    for (int k = 1; k <= kz; ++k)
      for (int j = jce1; j <= jce2; ++j) {
        double qext = qx(j, ice2, k, iqv), qint = qx(j, ici2, k, iqv);
        double windavg = v(j, ide2, k) + v(j, idi2, k);
        qx(j, ice2, k, iqv) = (windavg < d_zero) ? qext : qint;
      }
  }
  if (cfg.ipptls > 0) {   // .not. present_qc, bdyflow (:1963-2100)
    A3 &u = m_u, &v = m_v;
    for (int n = iqfrst; n <= iqlst; ++n) {
      // west
      for (int k = 1; k <= kz; ++k)
        for (int i = ici1; i <= ici2; ++i) {
          double qxint = std::max(qx(jci1, i, k, n), d_zero);
          double windavg = (u(jde1, i, k) + u(jdi1, i, k));
          qx(jce1, i, k, n) = (windavg >= d_zero) ? d_zero : qxint;
        }
      for (int k = 1; k <= kz; ++k) {
        double qxint = std::max(qx(jci1, ici1, k, n), d_zero);
        double windavg = (u(jde1, ice1, k) + u(jdi1, ice1, k) + v(jce1, ide1, k) + u(jce1, idi1, k));
        qx(jce1, ice1, k, n) = (windavg >= d_zero) ? d_zero : qxint;
      }
      for (int k = 1; k <= kz; ++k) {
        double qxint = std::max(qx(jci1, ici2, k, n), d_zero);
        double windavg = (u(jde1, ice2, k) + u(jdi1, ice2, k) + v(jce1, ide2, k) + u(jce1, idi2, k));
        qx(jce1, ice2, k, n) = (windavg >= d_zero) ? d_zero : qxint;
      }
      // east
      for (int k = 1; k <= kz; ++k)
        for (int i = ici1; i <= ici2; ++i) {
          double qxint = std::max(qx(jci2, i, k, n), d_zero);
          double windavg = (u(jde2, i, k) + u(jdi2, i, k));
          qx(jce2, i, k, n) = (windavg <= d_zero) ? d_zero : qxint;
        }
      for (int k = 1; k <= kz; ++k) {
        double qxint = std::max(qx(jci2, ici1, k, n), d_zero);
        double windavg = (u(jde2, ice1, k) + u(jdi2, ice1, k) + v(jce2, ide1, k) + u(jce2, idi1, k));
        qx(jce2, ice1, k, n) = (windavg <= d_zero) ? d_zero : qxint;
      }
      for (int k = 1; k <= kz; ++k) {
        double qxint = std::max(qx(jci2, ici2, k, n), d_zero);
        double windavg = (u(jde2, ice2, k) + u(jdi2, ice2, k) + v(jce2, ide2, k) + u(jce2, idi2, k));
        qx(jce2, ice2, k, n) = (windavg <= d_zero) ? d_zero : qxint;
      }
      // south, north
      for (int k = 1; k <= kz; ++k)
        for (int j = jci1; j <= jci2; ++j) {
          double qxint = std::max(qx(j, ici1, k, n), d_zero);
          double windavg = (v(j, ide1, k) + v(j, idi1, k));
          qx(j, ice1, k, n) = (windavg >= d_zero) ? d_zero : qxint;
        }
      for (int k = 1; k <= kz; ++k)
        for (int j = jci1; j <= jci2; ++j) {
          double qxint = std::max(qx(j, ici2, k, n), d_zero);
          double windavg = (v(j, ide2, k) + v(j, idi2, k));
          qx(j, ice2, k, n) = (windavg <= d_zero) ? d_zero : qxint;
        }
    }
  }
  xbctime = xbctime + dtsec;
}

static void mo_copy_out(const A3& src, double* a, const int32_t lb[3], const int32_t ub[3]) {
  const int n1 = ub[0] - lb[0] + 1, n2 = ub[1] - lb[1] + 1;
  for (int k = std::max(lb[2], src.l3); k <= std::min(ub[2], src.l3 + src.n3 - 1); ++k)
    for (int i = std::max(lb[1], src.l2); i <= std::min(ub[1], src.l2 + src.n2 - 1); ++i)
      for (int j = std::max(lb[0], src.l1); j <= std::min(ub[0], src.l1 + src.n1 - 1); ++j)
        a[(size_t)(j - lb[0]) + (size_t)n1 * ((size_t)(i - lb[1]) + (size_t)n2 * (size_t)(k - lb[2]))] = src(j, i, k);
}

int Oracle::mo_get_state(int field, double* a, const int32_t lb[3], const int32_t ub[3]) {
  auto q4 = [&](int n) {
    const int n1 = ub[0] - lb[0] + 1, n2 = ub[1] - lb[1] + 1;
    for (int k = std::max(lb[2], 1); k <= std::min(ub[2], kz); ++k)
      for (int i = std::max(lb[1], ice1); i <= std::min(ub[1], ice2); ++i)
        for (int j = std::max(lb[0], jce1); j <= std::min(ub[0], jce2); ++j)
          a[(size_t)(j - lb[0]) + (size_t)n1 * ((size_t)(i - lb[1]) + (size_t)n2 * (size_t)(k - lb[2]))] = m_qx(j, i, k, n);
    return RCM_OK;
  };
  switch (field) {
    case RCM_F_U: mo_copy_out(m_u, a, lb, ub); return RCM_OK;
    case RCM_F_V: mo_copy_out(m_v, a, lb, ub); return RCM_OK;
    case RCM_F_W: mo_copy_out(m_w, a, lb, ub); return RCM_OK;
    case RCM_F_T: mo_copy_out(m_t, a, lb, ub); return RCM_OK;
    case RCM_F_PAI: mo_copy_out(m_pai, a, lb, ub); return RCM_OK;
    case RCM_F_QV: return q4(iqv);
    case RCM_F_QC: return q4(iqc);
    case RCM_F_PS: {
      const int n1 = ub[0] - lb[0] + 1;
      for (int i = std::max(lb[1], ice1); i <= std::min(ub[1], ice2); ++i)
        for (int j = std::max(lb[0], jce1); j <= std::min(ub[0], jce2); ++j)
          a[(size_t)(j - lb[0]) + (size_t)n1 * (size_t)(i - lb[1])] = psa(j, i);
      return RCM_OK;
    }
  }
  err = "bad field for idynamic = 3"; return RCM_ERR_ARG;
}

int Oracle::mo_set_state(int field, const double* a, const int32_t lb[3], const int32_t ub[3]) {
  A3* dst = nullptr; int n4 = 0;
  switch (field) {
    case RCM_F_U: dst = &m_u; break;
    case RCM_F_V: dst = &m_v; break;
    case RCM_F_W: dst = &m_w; break;
    case RCM_F_T: dst = &m_t; break;
    case RCM_F_PAI: dst = &m_pai; break;
    case RCM_F_QV: n4 = iqv; break;
    case RCM_F_QC: n4 = iqc; break;
    default: err = "bad field for idynamic = 3"; return RCM_ERR_ARG;
  }
  const int n1 = ub[0] - lb[0] + 1, n2 = ub[1] - lb[1] + 1;
  if (dst) {
    for (int k = std::max(lb[2], dst->l3); k <= std::min(ub[2], dst->u3()); ++k)
      for (int i = std::max(lb[1], dst->l2); i <= std::min(ub[1], dst->l2 + dst->n2 - 1); ++i)
        for (int j = std::max(lb[0], dst->l1); j <= std::min(ub[0], dst->l1 + dst->n1 - 1); ++j)
          (*dst)(j, i, k) = a[(size_t)(j - lb[0]) + (size_t)n1 * ((size_t)(i - lb[1]) + (size_t)n2 * (size_t)(k - lb[2]))];
  } else {
    for (int k = std::max(lb[2], 1); k <= std::min(ub[2], kz); ++k)
      for (int i = std::max(lb[1], ice1); i <= std::min(ub[1], ice2); ++i)
        for (int j = std::max(lb[0], jce1); j <= std::min(ub[0], jce2); ++j)
          m_qx(j, i, k, n4) = a[(size_t)(j - lb[0]) + (size_t)n1 * ((size_t)(i - lb[1]) + (size_t)n2 * (size_t)(k - lb[2]))];
  }
  return RCM_OK;
}

int Oracle::mo_get_work(int id, double* a, const int32_t lb[3], const int32_t ub[3]) {
  switch (id) {
    case RCM_W_MO_P: mo_copy_out(m_p, a, lb, ub); return RCM_OK;
    case RCM_W_MO_TVIRT: mo_copy_out(m_tvirt, a, lb, ub); return RCM_OK;
    case RCM_W_MO_FMZ: mo_copy_out(m_fmz, a, lb, ub); return RCM_OK;
    case RCM_W_MO_FMZF: mo_copy_out(m_fmzf, a, lb, ub); return RCM_OK;
    case RCM_W_MO_ZETA: mo_copy_out(m_zeta, a, lb, ub); return RCM_OK;
    case RCM_W_MO_S: mo_copy_out(m_s, a, lb, ub); return RCM_OK;
  }
  err = "bad work id for idynamic = 3"; return RCM_ERR_ARG;
}

}  // namespace orc
