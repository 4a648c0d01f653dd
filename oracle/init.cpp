// ORACLE — TEST INFRASTRUCTURE ONLY.
// Second half of the initialisation restatement (see oracle/setup.cpp header
// for the list of reference routines and line ranges).
#include <algorithm>
#include <cmath>

#include "model.hpp"

namespace orc {

// Main/mpplib/mod_mppparam.F90:13811-13862
void Oracle::psc2psd(const A2& pc, A2& pd) {
  for (int i = idi1; i <= idi2; ++i)
    for (int j = jdi1; j <= jdi2; ++j)
      pd(j, i) = (pc(j, i) + pc(j, i - 1) + pc(j - 1, i) + pc(j - 1, i - 1)) * d_rfour;
  if (has_bdytop)
    for (int j = jdi1; j <= jdi2; ++j) pd(j, ide2) = (pc(j, ice2) + pc(j - 1, ice2)) * d_half;
  if (has_bdybottom)
    for (int j = jdi1; j <= jdi2; ++j) pd(j, ide1) = (pc(j, ice1) + pc(j - 1, ice1)) * d_half;
  if (has_bdyleft)
    for (int i = idi1; i <= idi2; ++i) pd(jde1, i) = (pc(jce1, i) + pc(jce1, i - 1)) * d_half;
  if (has_bdyright)
    for (int i = idi1; i <= idi2; ++i) pd(jde2, i) = (pc(jce2, i) + pc(jce2, i - 1)) * d_half;
  pd(jde1, ide1) = pc(jce1, ice1);
  pd(jde1, ide2) = pc(jce1, ice2);
  pd(jde2, ide1) = pc(jce2, ice1);
  pd(jde2, ide2) = pc(jce2, ice2);
}

// Share/mod_nhinterp.F90:75-106 (nhsetup values: ptoppa = ptop*1000,
// p0 = base_state_pressure, tlp = logp_lrate, st0 = base_state_ts0)
void Oracle::nhbase(int i1, int i2, int j1, int j2, int kx, const A1& sig,
                    const A2& ter, A2& ps0, A3& pr0, A3& t0, A3& rho0, A3& z0) {
  const double p0 = cfg.base_state_pressure, tlp = cfg.logp_lrate,
               st0 = cfg.base_state_ts0, ptoppa = ptop * d_1000;
  for (int i = i1; i <= i2; ++i) {
    for (int j = j1; j <= j2; ++j) {
      double ac = d_half * govr * ter(j, i) / tlp;
      double b = st0 / tlp;
      double alnp = -b + std::sqrt(b * b - d_four * ac);
      ps0(j, i) = p0 * std::exp(alnp) - ptoppa;
      for (int k = 1; k <= kx; ++k) {
        pr0(j, i, k) = ps0(j, i) * sig(k) + ptoppa;
        t0(j, i, k) = std::max(st0 + tlp * std::log(pr0(j, i, k) / p0), tiso);
        rho0(j, i, k) = pr0(j, i, k) / rgas / t0(j, i, k);
        alnp = std::log(pr0(j, i, k) / (ps0(j, i) + ptoppa));
        z0(j, i, k) = std::max(-(d_half * rovg * tlp * alnp * alnp + rovg * st0 * alnp), d_zero);
      }
    }
  }
}

// Main/mod_params.F90:1981-2010,2208-2215
void Oracle::param_static() {
  for (int i = ide1; i <= ide2; ++i)
    for (int j = jde1; j <= jde2; ++j) ht(j, i) = ht(j, i) * egrav;
  for (int i = ide1; i <= ide2; ++i)
    for (int j = jde1; j <= jde2; ++j) {
      msfd(j, i) = d_one / msfd(j, i);
      msfx(j, i) = d_one / msfx(j, i);
    }
  for (int i = idi1; i <= idi2; ++i)
    for (int j = jdi1; j <= jdi2; ++j) {
      dmsf(j, i) = d_one / (msfd(j, i) * msfd(j, i) * dx16);
      xmsf(j, i) = d_one / (msfx(j, i) * msfx(j, i) * dx4);
    }
  for (int k = 1; k <= kz; ++k) {
    hsigma(k) = (sigma(k + 1) + sigma(k)) * d_half;
    dsigma(k) = (sigma(k + 1) - sigma(k));
  }
  twt(1, 1) = d_zero; twt(1, 2) = d_zero; qcon(1) = d_zero;
  for (int k = 2; k <= kz; ++k) {
    twt(k, 1) = (sigma(k) - hsigma(k - 1)) / (hsigma(k) - hsigma(k - 1));
    twt(k, 2) = d_one - twt(k, 1);
    qcon(k) = (sigma(k) - hsigma(k)) / (hsigma(k - 1) - hsigma(k));
  }
}

// Main/mod_advection.F90:77-107
void Oracle::init_advection() {
  for (int k = 1; k <= kz; ++k) xds(k) = d_one / dsigma(k);
  dds(1) = d_zero; dds(kzp1) = d_zero;
  for (int k = 2; k <= kz; ++k) dds(k) = d_one / (dsigma(k) + dsigma(k - 1));
  ul = cfg.uoffc * d_half * dt / dx;
}

// Main/mod_bdycod.F90:183-278
void Oracle::setup_bdycon() {
  const int nspgx = cfg.nspgx, nspgd = cfg.nspgd, iboudy = cfg.iboudy;
  rdtbdy = d_one / dtbdys;
  if (iboudy == 1 || iboudy == 5) {
    fnudge = (cfg.bdy_nm > d_zero) ? cfg.bdy_nm : 0.1 / dt;
    gnudge = (cfg.bdy_dm > d_zero) ? cfg.bdy_dm : d_one / (dt * 50.0);
    if (moloch) {   // :216-219
      fnudge = fnudge * (double)nadv * (double)nsound;
      gnudge = gnudge * (double)nadv * (double)nsound;
    }
  }
  // iboudy == 1 .or. idynamic == 2
  for (int n = 2; n <= nspgx - 1; ++n) {
    double xfun = (double)(nspgx - n) / (double)(nspgx - 2);
    fcx(n) = fnudge * xfun; gcx(n) = gnudge * xfun;
  }
  for (int n = 2; n <= nspgd - 1; ++n) {
    double xfun = (double)(nspgd - n) / (double)(nspgd - 2);
    fcd(n) = fnudge * xfun; gcd(n) = gnudge * xfun;
  }
  if (iboudy == 4) {   // :237-250 (MOLOCH only in this restatement)
    for (int k = 2; k <= nspgd - 1; ++k) wgtd(k) = d_one;
    if (nspgd - 1 >= 2) wgtd(2) = 0.20;
    if (nspgd - 1 >= 3) wgtd(3) = 0.55;
    if (nspgd - 1 >= 4) wgtd(4) = 0.80;
    if (nspgd - 1 >= 5) wgtd(5) = 0.95;
    for (int k = 2; k <= nspgx - 1; ++k) wgtx(k) = d_one;
    if (nspgx - 1 >= 2) wgtx(2) = 0.4;
    if (nspgx - 1 >= 3) wgtx(3) = 0.7;
    if (nspgx - 1 >= 4) wgtx(4) = 0.9;
  }
  if (iboudy == 5) {
    std::vector<double> anudge(kz + 1);
    for (int k = 1; k <= kz; ++k) {
      if (hsigma(k) < 0.4) anudge[k] = cfg.high_nudge;
      else if (hsigma(k) < 0.8) anudge[k] = cfg.medium_nudge;
      else anudge[k] = cfg.low_nudge;
    }
    for (int k = 1; k <= kz; ++k) {
      for (int n = 2; n <= nspgx - 1; ++n) {
        double xfun = std::exp(-((double)(n - 2) / anudge[k]));
        hefc(n, k) = fnudge * xfun; hegc(n, k) = gnudge * xfun;
      }
      for (int n = 2; n <= nspgd - 1; ++n) {
        double xfun = std::exp(-((double)(n - 2) / anudge[k]));
        hefd(n, k) = fnudge * xfun; hegd(n, k) = gnudge * xfun;
      }
    }
  }
}

// Main/mod_params.F90:2613-2694
void Oracle::make_reference_atmosphere() {
  for (auto& x : ht.d) x = x * regrav;
  nhbase(ice1, ice2, jce1, jce2, kz, hsigma, ht, a0_ps, a0_pr, a0_t, a0_rho, a0_z);
  nhbase(ice1, ice2, jce1, jce2, kzp1, sigma, ht, a0_ps, a0_pf, a0_tf, a0_rhof, a0_zf);
  for (auto& x : ht.d) x = x * egrav;
  psc2psd(a0_ps, a0_psdot);
  for (int k = 1; k <= kz; ++k)
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j)
        a0_dzf(j, i, k) = a0_zf(j, i, k) - a0_zf(j, i, k + 1);
  for (int i = ice1; i <= ice2; ++i)
    for (int j = jci1; j <= jci2; ++j)
      dpsdxm(j, i) = (a0_ps(j + 1, i) - a0_ps(j - 1, i)) / (a0_ps(j, i) * dx8 * msfx(j, i));
  for (int i = ice1; i <= ice2; ++i)
    dpsdxm(jce1, i) = (a0_ps(jci1, i) - a0_ps(jce1, i)) / (a0_ps(jce1, i) * dx8 * msfx(jce1, i));
  for (int i = ice1; i <= ice2; ++i)
    dpsdxm(jce2, i) = (a0_ps(jce2, i) - a0_ps(jci2, i)) / (a0_ps(jce2, i) * dx8 * msfx(jce2, i));
  for (int i = ici1; i <= ici2; ++i)
    for (int j = jce1; j <= jce2; ++j)
      dpsdym(j, i) = (a0_ps(j, i + 1) - a0_ps(j, i - 1)) / (a0_ps(j, i) * dx8 * msfx(j, i));
  for (int j = jce1; j <= jce2; ++j)
    dpsdym(j, ice1) = (a0_ps(j, ici1) - a0_ps(j, ice1)) / (a0_ps(j, ice1) * dx8 * msfx(j, ice1));
  for (int j = jce1; j <= jce2; ++j)
    dpsdym(j, ice2) = (a0_ps(j, ice2) - a0_ps(j, ici2)) / (a0_ps(j, ice2) * dx8 * msfx(j, ice2));
  for (int k = 1; k <= kz; ++k)
    for (int i = idi1; i <= idi2; ++i)
      for (int j = jdi1; j <= jdi2; ++j) {
        a0_dprddx(j, i, k) = a0_pr(j, i, k) - a0_pr(j - 1, i, k) + a0_pr(j, i - 1, k) - a0_pr(j - 1, i - 1, k);
        a0_dprddy(j, i, k) = a0_pr(j, i, k) - a0_pr(j, i - 1, k) + a0_pr(j - 1, i, k) - a0_pr(j - 1, i - 1, k);
      }
}

// Main/mod_params.F90:2696-2741
void Oracle::compute_full_coriolis_coefficients() {
  for (int i = idi1; i <= idi2; ++i)
    for (int j = jdi1; j <= jdi2; ++j) {
      double dlat_ = dlat(j, i);
      double dlatdy = d_half * (xlat(j - 1, i) + xlat(j, i) - xlat(j - 1, i - 1) - xlat(j, i - 1));
      if (std::fabs(dlatdy) < 1.0e-8) dlatdy = std::copysign(1.0e-8, dlatdy);
      double dlondy = d_half * (xlon(j - 1, i) + xlon(j, i) - xlon(j - 1, i - 1) - xlon(j, i - 1));
      if (dlondy > 180.0) dlondy = dlondy - 360.0;
      if (dlondy < -180.0) dlondy = dlondy + 360.0;
      double rotang = -std::atan(dlondy / dlatdy * std::cos(degrad * dlat_));
      if (dlatdy < 0.0) rotang = rotang + mathpi;
      ef(j, i) = eomeg2 * std::cos(degrad * dlat_);
      ddx(j, i) = std::cos(rotang);
      ddy(j, i) = std::sin(rotang);
      dmdx(j, i) = -d_half * (msfx(j, i) + msfx(j, i - 1) - msfx(j - 1, i) - msfx(j - 1, i - 1)) /
                   (dx * msfd(j, i) * msfd(j, i));
      dmdy(j, i) = -d_half * (msfx(j, i) + msfx(j - 1, i) - msfx(j, i - 1) - msfx(j - 1, i - 1)) /
                   (dx * msfd(j, i) * msfd(j, i));
    }
  for (int i = ici1; i <= ici2; ++i)
    for (int j = jci1; j <= jci2; ++j) {
      ex(j, i) = d_rfour * (ef(j, i) + ef(j, i + 1) + ef(j + 1, i) + ef(j + 1, i + 1));
      crx(j, i) = d_rfour * (ddx(j, i) + ddx(j, i + 1) + ddx(j + 1, i) + ddx(j + 1, i + 1));
      cry(j, i) = d_rfour * (ddy(j, i) + ddy(j, i + 1) + ddy(j + 1, i) + ddy(j + 1, i + 1));
    }
}

// Main/mod_bdycod.F90:4938-4951
void Oracle::couple(A3& a, const A2& c, int j1, int j2, int i1, int i2, int k1, int k2) {
  for (int k = k1; k <= k2; ++k)
    for (int i = i1; i <= i2; ++i)
      for (int j = j1; j <= j2; ++j) a(j, i, k) = a(j, i, k) * c(j, i);
}

// Main/mod_bdycod.F90:5100-5113
void Oracle::timeint(const A3& a, const A3& b, A3& c, int j1, int j2, int i1, int i2, int k1, int k2) {
  for (int k = k1; k <= k2; ++k)
    for (int i = i1; i <= i2; ++i)
      for (int j = j1; j <= j2; ++j) c(j, i, k) = (a(j, i, k) - b(j, i, k)) * rdtbdy;
}

// state part of init_bdy, Main/mod_bdycod.F90:393-435,522-559,625-640
void Oracle::init_bdy() {
  xbctime = d_zero;
  for (int i = ice1; i <= ice2; ++i)
    for (int j = jce1; j <= jce2; ++j) xpsb_b0(j, i) = a0_ps(j, i) * d_r1000;
  for (int i = ide1; i <= ide2; ++i)
    for (int j = jde1; j <= jde2; ++j) psdot(j, i) = a0_psdot(j, i) * d_r1000;
  for (int s = 0; s < 2; ++s) {
    A3& u = s ? xub.b1 : xub.b0; A3& v = s ? xvb.b1 : xvb.b0;
    A3& t = s ? xtb.b1 : xtb.b0; A3& q = s ? xqb.b1 : xqb.b0;
    A3& pp = s ? xppb.b1 : xppb.b0; A3& ww = s ? xwwb.b1 : xwwb.b0;
    couple(u, psdot, jde1, jde2, ide1, ide2, 1, kz);
    couple(v, psdot, jde1, jde2, ide1, ide2, 1, kz);
    couple(t, xpsb_b0, jce1, jce2, ice1, ice2, 1, kz);
    couple(q, xpsb_b0, jce1, jce2, ice1, ice2, 1, kz);
    couple(pp, xpsb_b0, jce1, jce2, ice1, ice2, 1, kz);
    couple(ww, xpsb_b0, jce1, jce2, ice1, ice2, 1, kzp1);
  }
  timeint(xub.b1, xub.b0, xub.bt, jde1, jde2, ide1, ide2, 1, kz);
  timeint(xvb.b1, xvb.b0, xvb.bt, jde1, jde2, ide1, ide2, 1, kz);
  timeint(xtb.b1, xtb.b0, xtb.bt, jce1, jce2, ice1, ice2, 1, kz);
  timeint(xqb.b1, xqb.b0, xqb.bt, jce1, jce2, ice1, ice2, 1, kz);
  timeint(xppb.b1, xppb.b0, xppb.bt, jce1, jce2, ice1, ice2, 1, kz);
  timeint(xwwb.b1, xwwb.b0, xwwb.bt, jce1, jce2, ice1, ice2, 1, kzp1);
}

// Main/mod_init.F90:96-178,878-903
void Oracle::init_state() {
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
    for (int j = jce1; j <= jce2; ++j) {
      psa(j, i) = a0_ps(j, i) * d_r1000; psb(j, i) = psa(j, i); psc(j, i) = psa(j, i);
    }
  for (int k = 1; k <= kz; ++k)
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j) {
        a1_pp(j, i, k) = xppb.b0(j, i, k); a2_pp(j, i, k) = xppb.b0(j, i, k);
      }
  for (int k = 1; k <= kzp1; ++k)
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j) {
        a1_w(j, i, k) = xwwb.b0(j, i, k); a2_w(j, i, k) = xwwb.b0(j, i, k);
      }
  psc2psd(psa, psdota);
  psc2psd(psb, psdotb);
  for (int k = 1; k <= kz; ++k)
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j) {
        a1_pr(j, i, k) = a0_pr(j, i, k) + a1_pp(j, i, k) / psa(j, i);
        a1_rho(j, i, k) = a1_pr(j, i, k) /
            (rgas * a1_t(j, i, k) / psa(j, i) * (d_one + ep1 * a1_qx(j, i, k, iqv) / psa(j, i)));
        a2_pr(j, i, k) = a0_pr(j, i, k) + a2_pp(j, i, k) / psb(j, i);
      }
}

// Main/mod_diffusion.F90:93-167 (idiffu < 3)
void Oracle::initialize_diffusion() {
  xkhmax = dxsq / (64.0 * dtsec);
  dydc = cfg.adyndif * vonkar * vonkar * dx * d_rfour;
  double xkhz;
  if (hydro) {
    xkhz = cfg.ckh * 1.5e-3 * dxsq / dtsec;
  } else {
    xkhz = cfg.ckh * dx;
    xkhmax = d_two * xkhmax;
  }
  for (int i = ice1; i <= ice2; ++i)
    for (int j = jce1; j <= jce2; ++j) hgfact(j, i) = xkhz;
  if (cfg.diffu_hgtf == 1) {
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) {
        double hg1 = std::fabs((ht(j, i) - ht(j, i - 1)) / dx);
        double hg2 = std::fabs((ht(j, i) - ht(j, i + 1)) / dx);
        double hg3 = std::fabs((ht(j, i) - ht(j - 1, i)) / dx);
        double hg4 = std::fabs((ht(j, i) - ht(j + 1, i)) / dx);
        double hgmax = std::max(std::max(hg1, hg2), std::max(hg3, hg4)) * regrav * 1.0e3;
        hgfact(j, i) = xkhz / (d_one + hgmax * hgmax);
      }
  }
}

// Main/mod_sound.F90:115-161
void Oracle::init_sound() {
  rnpts = d_one / (double)((iy - 1 - 2) * (jx - 1 - 2));  // nicross = iy-1, njcross = jx-1
  if (cfg.ifupr == 1) {
    for (int i = 1; i <= 5; ++i) { fk[i] = d_two; fl[i] = d_two; }
    fk[0] = d_one; fl[0] = d_one; fk[6] = d_one; fl[6] = d_one;
    for (int i = -5; i <= 5; ++i) { fi[i + 6] = d_one; fj[i + 6] = d_one; }
    fi[0] = d_half; fj[0] = d_half; fi[12] = d_half; fj[12] = d_half;
    double loc = 0.0;  // Fortran SUM over (jci1:jci2,ici1:ici2), array order
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) loc += msfx(j, i);
    xmsf_mean = loc * rnpts;
  }
  bet = cfg.nhbet; xkd_snd = cfg.nhxkd;
  double maxt = -1e300;
  for (double x : a0_t.d) maxt = std::max(maxt, x);
  cs = std::sqrt(xgamma * rgas * maxt);
  dtsmax = dx / cs / (d_one + xkd_snd);
  bp = (d_one + bet) * d_half;
  bm = (d_one - bet) * d_half;
  bpxbp = bp * bp;
  bpxbm = bp * bm;
}

int Oracle::init() {
  if (moloch) {
    // Main/mod_params.F90:1981-1986,2005-2010 (ht -> geopotential, hsigma of the
    // DOMAIN-file sigma: what setup_bdycon sees, :2070), :2077, then init_bdy,
    // init, init_moloch and the first bdyval (Main/mod_regcm_interface.F90:146)
    for (auto& x : ht.d) x = x * egrav;
    for (int k = 1; k <= kz; ++k) hsigma(k) = (sigma(k + 1) + sigma(k)) * d_half;
    setup_bdycon();
    compute_moloch_static();
    init_bdy_moloch();
    init_state_moloch();
    init_moloch();
    bdyval_moloch();
    return RCM_OK;
  }
  param_static();
  init_advection();
  setup_bdycon();
  if (hydro) {
    init_bdy_hydro();
    init_state_hydro();
    init_bdy_tracers(); init_state_tracers();
    initialize_diffusion();
    spinit();   // Main/mod_regcm_interface.F90:141
    if (vm_numerr) { err = "vmodes: vertical mode problem reported errors"; return RCM_ERR_ARG; }
    bdyval();   // :146
    return RCM_OK;
  }
  make_reference_atmosphere();
  compute_full_coriolis_coefficients();
  init_bdy();
  init_state();
  init_bdy_tracers(); init_state_tracers();
  initialize_diffusion();
  init_sound();
  bdyval();   // Main/mod_regcm_interface.F90:146
  return RCM_OK;
}

// rcmtimer%advance + alarm checks, Main/mpplib/mod_timer.F90:152-177,306-327
void Oracle::timer_advance() {
  model_time += (int64_t)dtsec;
  lcount += 1;
  auto check = [&](int64_t& lastact, int64_t& now, int64_t actint) {
    if (now == model_time) return;
    double t1 = (double)model_time, t2 = (double)(lastact + actint);
    if (t1 >= t2) { lastact = (model_time / actint) * actint; now = model_time; }
  };
  check(day_lastact, day_now, 86400);
  check(bdy_lastact, bdy_now, (int64_t)dtbdys);
}

// Main/mod_regcm_interface.F90:172-228
int Oracle::step(int n) {
  for (int s = 0; s < n; ++s) {
    if (moloch) {
      moloch_step();
      if (!(ps_max == ps_max) || !(ps_min == ps_min)) { err = "CFL violation"; return RCM_ERR_CFL; }  // mod_moloch.F90:443-451
      if (alarm_bdy_act()) {
        if (!have_next) { err = "boundary record needed but not staged"; return RCM_ERR_BDY; }
        bdyin_moloch();
      }
      bdyval_moloch();
      continue;
    }
    tend();
    if (cfl_error) { err = "CFL violation"; return RCM_ERR_CFL; }
    if (sl_rc) return sl_rc;
    if (alarm_bdy_act()) {
      if (!have_next) { err = "boundary record needed but not staged"; return RCM_ERR_BDY; }
      if (hydro) bdyin_hydro(); else bdyin();
      if (ntr > 0) bdyin_tracers();
    }
    bdyval();
  }
  return RCM_OK;
}

static void copy_out_3d(const A3& src, double* a, const int32_t lb[3], const int32_t ub[3]) {
  const int n1 = ub[0] - lb[0] + 1, n2 = ub[1] - lb[1] + 1;
  for (int k = std::max(lb[2], src.l3); k <= std::min(ub[2], src.l3 + src.n3 - 1); ++k)
    for (int i = std::max(lb[1], src.l2); i <= std::min(ub[1], src.l2 + src.n2 - 1); ++i)
      for (int j = std::max(lb[0], src.l1); j <= std::min(ub[0], src.l1 + src.n1 - 1); ++j)
        a[(size_t)(j - lb[0]) + (size_t)n1 * ((size_t)(i - lb[1]) + (size_t)n2 * (size_t)(k - lb[2]))] = src(j, i, k);
}

static void copy_out_4d(const A4& src, int n, double* a, const int32_t lb[3], const int32_t ub[3]) {
  const int n1 = ub[0] - lb[0] + 1, n2 = ub[1] - lb[1] + 1;
  for (int k = std::max(lb[2], src.l3); k <= std::min(ub[2], src.l3 + src.n3 - 1); ++k)
    for (int i = std::max(lb[1], src.l2); i <= std::min(ub[1], src.l2 + src.n2 - 1); ++i)
      for (int j = std::max(lb[0], src.l1); j <= std::min(ub[0], src.l1 + src.n1 - 1); ++j)
        a[(size_t)(j - lb[0]) + (size_t)n1 * ((size_t)(i - lb[1]) + (size_t)n2 * (size_t)(k - lb[2]))] = src(j, i, k, n);
}

int Oracle::get_state(int field, int level, double* a, const int32_t lb[3], const int32_t ub[3]) {
  if (moloch) return mo_get_state(field, a, lb, ub);
  if (field >= RCM_F_CHI) return tracer_state(field - RCM_F_CHI, level, a, lb, ub, false);
  if (field == RCM_F_TKE) return tke_state(level, a, lb, ub, false);
  const bool A = (level == RCM_LEVEL_A);
  switch (field) {
    case RCM_F_U: copy_out_3d(A ? a1_u : a2_u, a, lb, ub); return RCM_OK;
    case RCM_F_V: copy_out_3d(A ? a1_v : a2_v, a, lb, ub); return RCM_OK;
    case RCM_F_T: copy_out_3d(A ? a1_t : a2_t, a, lb, ub); return RCM_OK;
    case RCM_F_QV: copy_out_4d(A ? a1_qx : a2_qx, iqv, a, lb, ub); return RCM_OK;
    case RCM_F_QC: copy_out_4d(A ? a1_qx : a2_qx, iqc, a, lb, ub); return RCM_OK;
    case RCM_F_PP: copy_out_3d(A ? a1_pp : a2_pp, a, lb, ub); return RCM_OK;
    case RCM_F_W: copy_out_3d(A ? a1_w : a2_w, a, lb, ub); return RCM_OK;
    case RCM_F_PS: {
      const A2& src = A ? psa : psb;
      const int n1 = ub[0] - lb[0] + 1;
      for (int i = std::max(lb[1], src.l2); i <= std::min(ub[1], src.l2 + src.n2 - 1); ++i)
        for (int j = std::max(lb[0], src.l1); j <= std::min(ub[0], src.l1 + src.n1 - 1); ++j)
          a[(size_t)(j - lb[0]) + (size_t)n1 * (size_t)(i - lb[1])] = src(j, i);
      return RCM_OK;
    }
  }
  err = "bad field"; return RCM_ERR_ARG;
}

int Oracle::get_work(int id, double* a, const int32_t lb[3], const int32_t ub[3]) {
  if (moloch) return mo_get_work(id, a, lb, ub);
  switch (id) {
    case RCM_W_CR: copy_out_3d(cr, a, lb, ub); return RCM_OK;
    case RCM_W_QDOT: copy_out_3d(qdot, a, lb, ub); return RCM_OK;
    case RCM_W_TH: copy_out_3d(th, a, lb, ub); return RCM_OK;
    case RCM_W_RHO1: copy_out_3d(a1_rho, a, lb, ub); return RCM_OK;
    case RCM_W_PR1: copy_out_3d(a1_pr, a, lb, ub); return RCM_OK;
    case RCM_W_UTEN: copy_out_3d(uten, a, lb, ub); return RCM_OK;
    case RCM_W_VTEN: copy_out_3d(vten, a, lb, ub); return RCM_OK;
    case RCM_W_PPTEN: copy_out_3d(ppten, a, lb, ub); return RCM_OK;
    case RCM_W_WTEN: copy_out_3d(wten, a, lb, ub); return RCM_OK;
    case RCM_W_XKC: copy_out_3d(xkc_raw, a, lb, ub); return RCM_OK;
    case RCM_W_OMEGA: copy_out_3d(omega, a, lb, ub); return RCM_OK;
    case RCM_W_PHI: copy_out_3d(phi, a, lb, ub); return RCM_OK;
    case RCM_W_PTEN: {
      const int n1 = ub[0] - lb[0] + 1;
      for (int i = std::max(lb[1], pten.l2); i <= std::min(ub[1], pten.l2 + pten.n2 - 1); ++i)
        for (int j = std::max(lb[0], pten.l1); j <= std::min(ub[0], pten.l1 + pten.n1 - 1); ++j)
          a[(size_t)(j - lb[0]) + (size_t)n1 * (size_t)(i - lb[1])] = pten_raw(j, i);
      return RCM_OK;
    }
  }
  err = "bad work id"; return RCM_ERR_ARG;
}

}  // namespace orc
