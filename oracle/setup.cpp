// ORACLE — TEST INFRASTRUCTURE ONLY.
// Restates the initialisation path of the reference for idynamic = 2:
//   setup_model_indexes   Main/mod_atm_interface.F90:181-381
//   setup_boundaries      Main/mod_atm_interface.F90:383-542
//   param (static part)   Main/mod_params.F90:1763-1770,1960-2074,2208-2215
//   make_reference_atmosphere / compute_full_coriolis_coefficients
//                         Main/mod_params.F90:2613-2741
//   nhbase                Share/mod_nhinterp.F90:75-106
//   psc2psd               Main/mpplib/mod_mppparam.F90:13811-13862
//   init_bdy (state part) Main/mod_bdycod.F90:393-435,522-559,625-640
//   init                  Main/mod_init.F90:96-178,878-903
//   initialize_diffusion  Main/mod_diffusion.F90:93-167
//   init_sound            Main/mod_sound.F90:115-161
#include <algorithm>
#include <cmath>
#include <cstring>

#include "model.hpp"

namespace orc {

int Oracle::create(const rcm_cfg& c) {
  cfg = c;
  if (c.struct_size != (int)sizeof(rcm_cfg)) { err = "rcm_cfg size mismatch"; return RCM_ERR_ARG; }
  if (c.idynamic < 1 || c.idynamic > 3) { err = "oracle: idynamic must be 1, 2 or 3"; return RCM_ERR_ARG; }
  hydro = (c.idynamic == 1);
  moloch = (c.idynamic == 3);
  if (hydro && (c.nsplit < 1 || c.nsplit > c.kz)) { err = "oracle: nsplit out of range"; return RCM_ERR_ARG; }
  nsplit = hydro ? c.nsplit : 0;
  if (c.isladvec != 0 && (c.isladvec != 1 || c.idynamic == 3)) { err = "oracle: isladvec must be 0 or 1 (idynamic 1, 2)"; return RCM_ERR_ARG; }
  if (c.ipgf != 0 && c.ipgf != 1) { err = "oracle: ipgf must be 0 or 1"; return RCM_ERR_ARG; }
  if (c.idiffu < 1 || c.idiffu > 3) { err = "oracle: idiffu must be 1, 2 or 3"; return RCM_ERR_ARG; }
  // 0 fixed, 1 / 5 relaxation, 4 sponge; sigma-coordinate cores also 2 (time-dependent, no relaxation) and 3 (2 + inflow/outflow)
  if (!(c.iboudy == 0 || c.iboudy == 1 || c.iboudy == 4 || c.iboudy == 5 || (c.idynamic != 3 && (c.iboudy == 2 || c.iboudy == 3)))) { err = "oracle: iboudy must be 0..5 (2, 3: idynamic 1, 2 only)"; return RCM_ERR_ARG; }
  if (c.nqx != 2) { err = "oracle: nqx must be 2"; return RCM_ERR_ARG; }
  if (c.ntr < 0 || c.ntr > RCM_MAXTR || (c.ntr > 0 && c.idynamic == 3)) { err = "oracle: ntr must be 0..8 (idynamic 1, 2)"; return RCM_ERR_ARG; }
  jx = c.jx; iy = c.iy; kz = c.kz; kzp1 = kz + 1; nqx = c.nqx;
  iqv = 1; iqc = 2; iqfrst = 2; iqlst = nqx;
  ptop = c.ptop;
  dtsec = c.dt; dt = c.dt; dt2 = d_two * dt;   // mod_params.F90:1628,1708
  dtbdys = (double)c.ibdyfrq * 3600.0;          // :1631
  idif = (c.idiffu == 2) ? 1 : (c.idiffu == 3) ? 3 : 2;   // :1968-1973
  // mod_params.F90:1763-1770
  dx = c.ds * d_1000;
  dx2 = d_two * dx; dx4 = d_four * dx; dx8 = 8.0 * dx; dx16 = 16.0 * dx;
  dxsq = dx * dx; rdx = 1.0 / dx; rdxsq = 1.0 / dxsq;
  xgamma = d_one / (d_one - rovcp);             // mod_sound.F90:75
  lcount = 0; model_time = 0;
  day_lastact = 0; day_now = 0; bdy_lastact = 0; bdy_now = 0;
  xbctime = 0; have_next = false;
  setup_model_indexes();
  if (moloch) return create_moloch();
  allocate_all();
  if (hydro) allocate_hydro();
  allocate_tracers();
  allocate_tke();
  setup_boundaries(false, false, ba_cr);   // mod_params.F90:1788-1793
  setup_boundaries(true, true, ba_dt);
  return RCM_OK;
}

void Oracle::setup_model_indexes() {
  // single rank: every side is a physical boundary, no ghost rows
  jde1 = 1; jde2 = jx; ide1 = 1; ide2 = iy;
  jdi1 = jde1 + 1; jdii1 = jde1 + 2; jdi2 = jde2 - 1; jdii2 = jde2 - 2;
  idi1 = ide1 + 1; idii1 = ide1 + 2; idi2 = ide2 - 1; idii2 = ide2 - 2;
  jce1 = 1; jce2 = jx - 1; ice1 = 1; ice2 = iy - 1;
  jci1 = jce1 + 1; jcii1 = jce1 + 2; jci2 = jce2 - 1; jcii2 = jce2 - 2;
  ici1 = ice1 + 1; icii1 = ice1 + 2; ici2 = ice2 - 1; icii2 = ice2 - 2;
  jcross1 = 1; jcross2 = jx - 1; icross1 = 1; icross2 = iy - 1;
}

void Oracle::setup_boundaries(bool ldotx, bool ldoty, BoundArea& ba) {
  ba.ibnd.alloc(jde1, jde2, ide1, ide2, -1);
  ba.bsouth.alloc(jde1, jde2, ide1, ide2, 0);
  ba.bnorth.alloc(jde1, jde2, ide1, ide2, 0);
  ba.beast.alloc(jde1, jde2, ide1, ide2, 0);
  ba.bwest.alloc(jde1, jde2, ide1, ide2, 0);
  int icx = ldotx ? 0 : 1, icy = ldoty ? 0 : 1;
  ba.nsp = (ldotx || ldoty) ? cfg.nspgd : cfg.nspgx;
  int igbb1 = 2, igbb2 = ba.nsp - 1, jgbl1 = 2, jgbl2 = ba.nsp - 1;
  int igbt1 = iy - icy - ba.nsp + 2, igbt2 = iy - icy - 1;
  int jgbr1 = jx - icx - ba.nsp + 2, jgbr2 = jx - icx - 1;
  int i1 = ide1, i2 = ide2, j1 = jde1, j2 = jde2;
  ba.ns = ba.nn = ba.nw = ba.ne = 0;
  // South
  for (int i = i1; i <= i2; ++i) {
    if (i >= igbb1 && i <= igbb2) {
      for (int j = j1; j <= j2; ++j) {
        if (j >= jgbl1 && j <= jgbr2) {
          if (j <= jgbl2 && i >= j) continue;
          if (j >= jgbr1 && i >= (jgbr2 - j + 2)) continue;
          ba.ibnd(j, i) = i - igbb1 + 2; ba.bsouth(j, i) = 1; ba.ns++;
        }
      }
    }
  }
  // North
  for (int i = i1; i <= i2; ++i) {
    if (i >= igbt1 && i <= igbt2) {
      for (int j = j1; j <= j2; ++j) {
        if (j >= jgbl1 && j <= jgbr2) {
          if (j <= jgbl2 && (igbt2 - i + 2) >= j) continue;
          if (j >= jgbr1 && (igbt2 - i) >= (jgbr2 - j)) continue;
          ba.ibnd(j, i) = igbt2 - i + 2; ba.bnorth(j, i) = 1; ba.nn++;
        }
      }
    }
  }
  // West
  for (int i = i1; i <= i2; ++i) {
    if (i < igbb1 || i > igbt2) continue;
    for (int j = j1; j <= j2; ++j) {
      if (j >= jgbl1 && j <= jgbl2) {
        if (i < igbb2 && j > i) continue;
        if (i > igbt1 && j > (igbt2 - i + 2)) continue;
        ba.ibnd(j, i) = j - jgbl1 + 2; ba.bwest(j, i) = 1; ba.nw++;
      }
    }
  }
  // East
  for (int i = i1; i <= i2; ++i) {
    if (i < igbb1 || i > igbt2) continue;
    for (int j = j1; j <= j2; ++j) {
      if (j >= jgbr1 && j <= jgbr2) {
        if (i < igbb2 && (jgbr2 - j + 2) > i) continue;
        if (i > igbt1 && (jgbr2 - j) > (igbt2 - i)) continue;
        ba.ibnd(j, i) = jgbr2 - j + 2; ba.beast(j, i) = 1; ba.ne++;
      }
    }
  }
  ba.havebound = (ba.ns != 0 || ba.nn != 0 || ba.nw != 0 || ba.ne != 0);
}

void Oracle::allocate_all() {
  const int nspgx = cfg.nspgx, nspgd = cfg.nspgd;
  sigma.alloc(1, kzp1); hsigma.alloc(1, kz); dsigma.alloc(1, kz);
  qcon.alloc(1, kz); dds.alloc(1, kzp1); xds.alloc(1, kz);
  twt.alloc(1, kz, 1, 2);
  // mddom (dot extent)
  for (A2* a : {&ht, &msfx, &msfd, &coriol, &xlat, &xlon, &dlat, &xmsf, &dmsf,
                &ef, &ddx, &ddy, &dmdx, &dmdy, &ex, &crx, &cry})
    a->alloc(jde1, jde2, ide1, ide2);
  dpsdxm.alloc(jce1, jce2, ice1, ice2); dpsdym.alloc(jce1, jce2, ice1, ice2);
  psa.alloc(jce1, jce2, ice1, ice2); psb.alloc(jce1, jce2, ice1, ice2);
  psc.alloc(jce1, jce2, ice1, ice2);
  psdota.alloc(jde1, jde2, ide1, ide2); psdotb.alloc(jde1, jde2, ide1, ide2);
  rpsa.alloc(jce1, jce2, ice1, ice2); rpsb.alloc(jce1, jce2, ice1, ice2);
  rpsda.alloc(jde1, jde2, ide1, ide2);
  // atm0 (mod_atm_interface.F90:758-774)
  a0_ps.alloc(jce1, jce2, ice1, ice2); a0_psdot.alloc(jde1, jde2, ide1, ide2);
  for (A3* a : {&a0_t, &a0_pr, &a0_rho, &a0_z})
    a->alloc(jce1, jce2, ice1, ice2, 1, kz);
  for (A3* a : {&a0_tf, &a0_pf, &a0_rhof, &a0_zf, &a0_dzf})
    a->alloc(jce1, jce2, ice1, ice2, 1, kzp1);
  a0_dprddx.alloc(jdi1, jdi2, idi1, idi2, 1, kz);
  a0_dprddy.alloc(jdi1, jdi2, idi1, idi2, 1, kz);
  // atm1 / atm2 / atmc (:617-696)
  for (A3* a : {&a1_u, &a1_v, &a2_u, &a2_v, &ac_u, &ac_v})
    a->alloc(jde1, jde2, ide1, ide2, 1, kz);
  for (A3* a : {&a1_t, &a1_pp, &a1_pr, &a1_rho, &a2_t, &a2_pp, &a2_pr, &ac_t,
                &ac_pp})
    a->alloc(jce1, jce2, ice1, ice2, 1, kz);
  for (A3* a : {&a1_w, &a2_w, &ac_w}) a->alloc(jce1, jce2, ice1, ice2, 1, kzp1);
  for (A4* a : {&a1_qx, &a2_qx, &ac_qx})
    a->alloc(jce1, jce2, ice1, ice2, 1, kz, 1, nqx);
  // atmx (:698-728)
  for (A3* a : {&ax_uc, &ax_vc, &ax_umc, &ax_vmc, &ax_ud, &ax_vd, &ax_umd,
                &ax_vmd})
    a->alloc(jde1, jde2, ide1, ide2, 1, kz);
  for (A3* a : {&ax_t, &ax_tv, &ax_pr, &ax_pp})
    a->alloc(jce1, jce2, ice1, ice2, 1, kz);
  ax_w.alloc(jce1, jce2, ice1, ice2, 1, kzp1);
  ax_qx.alloc(jce1, jce2, ice1, ice2, 1, kz, 1, nqx);
  // aten (:730-756)
  for (A3* a : {&uten, &vten, &udyn, &vdyn}) a->alloc(jdi1, jdi2, idi1, idi2, 1, kz);
  for (A3* a : {&tten, &tdyn, &ppten, &ppdyn}) a->alloc(jci1, jci2, ici1, ici2, 1, kz);
  for (A3* a : {&wten, &wdyn}) a->alloc(jci1, jci2, ici1, ici2, 1, kzp1);
  for (A4* a : {&qxten, &qxdyn}) a->alloc(jci1, jci2, ici1, ici2, 1, kz, 1, nqx);
  // atms slices
  ubd3d.alloc(jde1, jde2, ide1, ide2, 1, kz); vbd3d.alloc(jde1, jde2, ide1, ide2, 1, kz);
  tb3d.alloc(jce1, jce2, ice1, ice2, 1, kz); ppb3d.alloc(jce1, jce2, ice1, ice2, 1, kz);
  wb3d.alloc(jce1, jce2, ice1, ice2, 1, kzp1);
  qxb3d.alloc(jce1, jce2, ice1, ice2, 1, kz, 1, nqx);
  // work
  qdot.alloc(jce1, jce2, ice1, ice2, 1, kzp1);
  cr.alloc(jce1, jce2, ice1, ice2, 1, kz);
  omega.alloc(jci1, jci2, ici1, ici2, 1, kz);
  ucc.alloc(jce1, jce2, ice1, ice2, 1, kz); vcc.alloc(jce1, jce2, ice1, ice2, 1, kz);
  th.alloc(jce1, jce2, ice1, ice2, 1, kz); tha.alloc(jce1, jce2, ice1, ice2, 1, kz);
  thten.alloc(jci1, jci2, ici1, ici2, 1, kz);
  for (A3* a : {&fg, &uavg1, &uavg2, &vavg1, &vavg2})
    a->alloc(jci1, jci2, ici1, ici2, 1, kz);
  // mod_diffusion.F90:83-91
  hgfact.alloc(jce1, jce2, ice1, ice2);
  xkc.alloc(jce1, jce2, ice1, ice2, 1, kz);
  xkd.alloc(jdi1, jdi2, idi1, idi2, 1, kz);
  xkcf.alloc(jci1, jci2, ici1, ici2, 1, kzp1);
  // boundary structures (mod_atm_interface.F90:544-573)
  for (V3dBound* b : {&xub, &xvb}) {
    b->b0.alloc(jde1, jde2, ide1, ide2, 1, kz);
    b->b1.alloc(jde1, jde2, ide1, ide2, 1, kz);
    b->bt.alloc(jde1, jde2, ide1, ide2, 1, kz);
  }
  for (V3dBound* b : {&xtb, &xqb, &xppb}) {
    b->b0.alloc(jce1, jce2, ice1, ice2, 1, kz);
    b->b1.alloc(jce1, jce2, ice1, ice2, 1, kz);
    b->bt.alloc(jce1, jce2, ice1, ice2, 1, kz);
  }
  xwwb.b0.alloc(jce1, jce2, ice1, ice2, 1, kzp1);
  xwwb.b1.alloc(jce1, jce2, ice1, ice2, 1, kzp1);
  xwwb.bt.alloc(jce1, jce2, ice1, ice2, 1, kzp1);
  nx_u.alloc(jde1, jde2, ide1, ide2, 1, kz); nx_v.alloc(jde1, jde2, ide1, ide2, 1, kz);
  nx_t.alloc(jce1, jce2, ice1, ice2, 1, kz); nx_q.alloc(jce1, jce2, ice1, ice2, 1, kz);
  nx_pp.alloc(jce1, jce2, ice1, ice2, 1, kz); nx_ww.alloc(jce1, jce2, ice1, ice2, 1, kzp1);
  xpsb_b0.alloc(jce1, jce2, ice1, ice2);
  psdot.alloc(jde1, jde2, ide1, ide2);
  // mod_bdycod.F90:134-181
  for (A2* a : {&wue, &wui, &eue, &eui, &wve, &wvi, &eve, &evi}) a->alloc(ide1, ide2, 1, kz);
  for (A2* a : {&sue, &sui, &nue, &nui, &sve, &svi, &nve, &nvi}) a->alloc(jde1, jde2, 1, kz);
  fcx.alloc(2, nspgx - 1); gcx.alloc(2, nspgx - 1);
  fcd.alloc(2, nspgd - 1); gcd.alloc(2, nspgd - 1);
  wgtx.alloc(2, nspgx - 1); wgtd.alloc(2, nspgd - 1);   // sponge weights (iboudy = 4, Main/mod_bdycod.F90:237-250)
  hefc.alloc(2, nspgx - 1, 1, kz); hegc.alloc(2, nspgx - 1, 1, kz);
  hefd.alloc(2, nspgd - 1, 1, kz); hegd.alloc(2, nspgd - 1, 1, kz);
  fg1.alloc(jde1, jde2, ide1, ide2, 1, kzp1);
  fg2.alloc(jde1, jde2, ide1, ide2, 1, kz);
  // mod_sound.F90:82-113
  for (A3* a : {&s_aa, &s_b, &s_c, &s_rhs, &s_ca, &s_g1, &s_g2})
    a->alloc(jci1, jci2, ici1, ici2, 2, kz);
  for (A3* a : {&s_sigdot, &s_wo}) a->alloc(jci1, jci2, ici1, ici2, 1, kzp1);
  for (A3* a : {&s_e, &s_f, &s_ptend, &s_pxup, &s_pyvp, &s_tk, &s_cc, &s_cdd,
                &s_cj, &s_pi, &s_ucrs, &s_vcrs})
    a->alloc(jci1, jci2, ici1, ici2, 1, kz);
  s_astore.alloc(jci1, jci2, ici1, ici2);
  s_estore.alloc(jce1, jce2, ice1, ice2);
  s_estore_g.alloc(jcross1, jcross2, icross1, icross2);
  s_wpval.alloc(jci1, jci2, ici1, ici2);
  s_rpsb.alloc(jce1, jce2, ice1, ice2);
  tmask.alloc(-6, 6, -6, 6);
}

static int copy_in_2d(A2& dst, const double* a, const int32_t lb[3], const int32_t ub[3]) {
  const int n1 = ub[0] - lb[0] + 1;
  for (int i = std::max(lb[1], dst.l2); i <= std::min(ub[1], dst.l2 + dst.n2 - 1); ++i)
    for (int j = std::max(lb[0], dst.l1); j <= std::min(ub[0], dst.l1 + dst.n1 - 1); ++j)
      dst(j, i) = a[(size_t)(j - lb[0]) + (size_t)n1 * (size_t)(i - lb[1])];
  return RCM_OK;
}

static int copy_in_3d(A3& dst, const double* a, const int32_t lb[3], const int32_t ub[3]) {
  const int n1 = ub[0] - lb[0] + 1, n2 = ub[1] - lb[1] + 1;
  for (int k = std::max(lb[2], dst.l3); k <= std::min(ub[2], dst.l3 + dst.n3 - 1); ++k)
    for (int i = std::max(lb[1], dst.l2); i <= std::min(ub[1], dst.l2 + dst.n2 - 1); ++i)
      for (int j = std::max(lb[0], dst.l1); j <= std::min(ub[0], dst.l1 + dst.n1 - 1); ++j)
        dst(j, i, k) = a[(size_t)(j - lb[0]) + (size_t)n1 * ((size_t)(i - lb[1]) + (size_t)n2 * (size_t)(k - lb[2]))];
  return RCM_OK;
}

int Oracle::set_static(int id, const double* a, const int32_t lb[3], const int32_t ub[3]) {
  switch (id) {
    case RCM_S_HT: return copy_in_2d(ht, a, lb, ub);
    case RCM_S_MSFX: return copy_in_2d(msfx, a, lb, ub);
    case RCM_S_MSFD: return moloch ? RCM_OK : copy_in_2d(msfd, a, lb, ub);
    case RCM_S_CORIOL: return moloch ? RCM_OK : copy_in_2d(coriol, a, lb, ub);
    case RCM_S_XLAT: return copy_in_2d(xlat, a, lb, ub);
    case RCM_S_XLON: return copy_in_2d(xlon, a, lb, ub);
    case RCM_S_DLAT: return moloch ? RCM_OK : copy_in_2d(dlat, a, lb, ub);
    case RCM_S_MSFU: return copy_in_2d(msfu, a, lb, ub);
    case RCM_S_MSFV: return copy_in_2d(msfv, a, lb, ub);
    case RCM_S_ULAT: return copy_in_2d(ulat, a, lb, ub);
    case RCM_S_VLAT: return copy_in_2d(vlat, a, lb, ub);
    case RCM_S_SIGMA:
      for (int k = lb[2]; k <= ub[2]; ++k) sigma(k) = a[k - lb[2]];
      return RCM_OK;
  }
  err = "bad static id"; return RCM_ERR_ARG;
}

int Oracle::set_bdy(int field, int slot, const double* a, const int32_t lb[3], const int32_t ub[3]) {
  if (field >= RCM_F_CHI) { if (slot == RCM_BDY_NEXT) have_next = true; return tracer_bdy(field - RCM_F_CHI, slot, a, lb, ub); }
  if (field == RCM_F_PS) {
    if (!hydro && !moloch) { err = "ps boundary data exist for idynamic = 1 and 3 only"; return RCM_ERR_ARG; }
    if (slot == RCM_BDY_B0) return copy_in_2d(xpsb.b0, a, lb, ub);
    if (slot == RCM_BDY_B1) return copy_in_2d(xpsb.b1, a, lb, ub);
    if (slot == RCM_BDY_NEXT) { have_next = true; return copy_in_2d(nx_ps, a, lb, ub); }
    err = "bad slot"; return RCM_ERR_ARG;
  }
  V3dBound* b = nullptr; A3* nx = nullptr;
  switch (field) {
    case RCM_F_U: b = &xub; nx = &nx_u; break;
    case RCM_F_V: b = &xvb; nx = &nx_v; break;
    case RCM_F_T: b = &xtb; nx = &nx_t; break;
    case RCM_F_QV: b = &xqb; nx = &nx_q; break;
    case RCM_F_PP: if (!moloch) { b = &xppb; nx = &nx_pp; } break;
    case RCM_F_W: if (!moloch) { b = &xwwb; nx = &nx_ww; } break;
    default: break;
  }
  if (!b) { err = "no boundary data for this field"; return RCM_ERR_ARG; }
  if (slot == RCM_BDY_B0) return copy_in_3d(b->b0, a, lb, ub);
  if (slot == RCM_BDY_B1) return copy_in_3d(b->b1, a, lb, ub);
  if (slot == RCM_BDY_NEXT) { have_next = true; return copy_in_3d(*nx, a, lb, ub); }
  err = "bad slot"; return RCM_ERR_ARG;
}

// state import (tests only): the coupled atm1/atm2 fields
int Oracle::set_state(int field, int level, const double* a, const int32_t lb[3], const int32_t ub[3]) {
  if (moloch) return mo_set_state(field, a, lb, ub);
  if (field >= RCM_F_CHI) return tracer_state(field - RCM_F_CHI, level, const_cast<double*>(a), lb, ub, true);
  if (field == RCM_F_TKE) return tke_state(level, const_cast<double*>(a), lb, ub, true);
  const bool A = (level == RCM_LEVEL_A);
  auto in4 = [&](A4& dst, int n) {
    const int n1 = ub[0] - lb[0] + 1, n2 = ub[1] - lb[1] + 1;
    for (int k = std::max(lb[2], dst.l3); k <= std::min(ub[2], dst.l3 + dst.n3 - 1); ++k)
      for (int i = std::max(lb[1], dst.l2); i <= std::min(ub[1], dst.l2 + dst.n2 - 1); ++i)
        for (int j = std::max(lb[0], dst.l1); j <= std::min(ub[0], dst.l1 + dst.n1 - 1); ++j)
          dst(j, i, k, n) = a[(size_t)(j - lb[0]) + (size_t)n1 * ((size_t)(i - lb[1]) + (size_t)n2 * (size_t)(k - lb[2]))];
    return RCM_OK;
  };
  switch (field) {
    case RCM_F_U: return copy_in_3d(A ? a1_u : a2_u, a, lb, ub);
    case RCM_F_V: return copy_in_3d(A ? a1_v : a2_v, a, lb, ub);
    case RCM_F_T: return copy_in_3d(A ? a1_t : a2_t, a, lb, ub);
    case RCM_F_QV: return in4(A ? a1_qx : a2_qx, iqv);
    case RCM_F_QC: return in4(A ? a1_qx : a2_qx, iqc);
    case RCM_F_PP: return copy_in_3d(A ? a1_pp : a2_pp, a, lb, ub);
    case RCM_F_W: return copy_in_3d(A ? a1_w : a2_w, a, lb, ub);
    case RCM_F_PS: return copy_in_2d(A ? psa : psb, a, lb, ub);
  }
  err = "bad field"; return RCM_ERR_ARG;
}

}  // namespace orc
