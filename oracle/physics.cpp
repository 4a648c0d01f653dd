// ORACLE — TEST INFRASTRUCTURE ONLY.
// Physics coupling boundary of the sigma-coordinate cores:
//   * the slices `mkslice` hands to the physical parametrizations, Main/mod_slice.F90:167-308 (idynamic /= 3),
//     evaluated on demand from the state `tend` sees at the point where it calls physical_parametrizations
//     (Main/mod_tendency.F90:271);
//   * aten(:,:,:,pc_physic): the tendencies the parametrizations return, summed at Main/mod_tendency.F90:285-314,
//     331-335,404-411 (tten = tten + tdyn + tphy, ...).
#include <algorithm>
#include <cmath>

#include "model.hpp"

namespace orc {

namespace {
// Share/pfesat.inc, Share/pfwsat.inc
double pfesat(double t) {
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
double pfwsat(double t, double p) {
  const double es = pfesat(t);
  return ep2 * (es / (p - es));
}
constexpr double rhmax = 1.01, rhmin = 0.01;   // Main/mod_params.F90:331-332
}  // namespace

// everything `tend` does before physical_parametrizations that the slices depend on; idempotent (a function of the
// prognostic state only), so that step() afterwards repeats it harmlessly
void Oracle::pre_physics() {
  if (moloch) {   // everything up to `call mkslice`; the step itself then resumes behind it
    if (!mo_mid) { moloch_head(); mo_mid = true; }
    return;
  }
  if (hydro) {
    surface_pressures_hydro();
    decouple();
    compute_omega_hydro();
    mkslice_hydro();
  } else {
    surface_pressures();
    decouple();
    compute_omega();
    mkslice();
  }
}

void Oracle::ensure_phys() {
  if (phys_on) return;
  for (A3* a : {&uphy, &vphy}) a->alloc(jdi1, jdi2, idi1, idi2, 1, kz);
  for (A3* a : {&tphy, &ppphy}) a->alloc(jci1, jci2, ici1, ici2, 1, kz);
  wphy.alloc(jci1, jci2, ici1, ici2, 1, kzp1);
  qxphy.alloc(jci1, jci2, ici1, ici2, 1, kz, 1, nqx);
  phys_on = true;
}

// MOLOCH: mo_atm%tten, uten, vten, qxten (Main/mod_moloch.F90:1548-1557), not coupled with anything
int Oracle::mo_set_phys(int field, const double* a, const int32_t lb[3], const int32_t ub[3]) {
  if (!mo_phys_on) {
    m_tten.alloc(jci1, jci2, ici1, ici2, 1, kz); m_uten.alloc(jdi1, jdi2, ici1, ici2, 1, kz); m_vten.alloc(jci1, jci2, idi1, idi2, 1, kz);
    m_qxten.alloc(jci1, jci2, ici1, ici2, 1, kz, 1, nqx);
    mo_phys_on = true;
  }
  A3* dst = nullptr;
  int n4 = 0;
  switch (field) {
    case RCM_F_T: dst = &m_tten; break;
    case RCM_F_U: dst = &m_uten; break;
    case RCM_F_V: dst = &m_vten; break;
    case RCM_F_QV: n4 = iqv; break;
    case RCM_F_QC: n4 = iqc; break;
    default: err = "oracle: no physics tendency for this field"; return RCM_ERR_ARG;
  }
  const long long n1 = ub[0] - lb[0] + 1, n2 = ub[1] - lb[1] + 1;
  const int l1 = dst ? dst->l1 : m_qxten.l1, l2 = dst ? dst->l2 : m_qxten.l2, m1 = dst ? dst->n1 : m_qxten.n1, m2 = dst ? dst->n2 : m_qxten.n2;
  for (int k = std::max(lb[2], 1); k <= std::min(ub[2], kz); ++k)
    for (int i = std::max(lb[1], l2); i <= std::min(ub[1], l2 + m2 - 1); ++i)
      for (int j = std::max(lb[0], l1); j <= std::min(ub[0], l1 + m1 - 1); ++j) {
        const double v = a[(j - lb[0]) + n1 * ((i - lb[1]) + n2 * (long long)(k - lb[2]))];
        if (dst) (*dst)(j, i, k) = v; else m_qxten(j, i, k, n4) = v;
      }
  return RCM_OK;
}

// the members of `atms` the MOLOCH core hands to the physics: aliases of mo_atm (Main/mod_slice.F90:82-100) and what
// mkslice computes for idynamic = 3 (:107-146)
int Oracle::mo_get_slice(int id, double* a, const int32_t lb[3], const int32_t ub[3]) {
  if (id < 0 || id >= RCM_SL_COUNT) { err = "oracle: bad slice id"; return RCM_ERR_ARG; }
  A3 out; out.alloc(jce1, jce2, ice1, ice2, 1, kzp1);
  int nk = kz;
  if (id == RCM_SL_PF3D || id == RCM_SL_WB3D || id == RCM_SL_ZQ) nk = kzp1;
  if (id == RCM_SL_PS2D) nk = 1;
  auto inner = [&](int j, int i) { return j >= jci1 && j <= jci2 && i >= ici1 && i <= ici2; };
  for (int i = ice1; i <= ice2; ++i)
    for (int j = jce1; j <= jce2; ++j)
      for (int k = 1; k <= nk; ++k) {
        double v = d_zero;
        const double qvv = inner(j, i) ? std::max(m_qx(j, i, std::min(k, kz), iqv), minqq) : m_qx(j, i, std::min(k, kz), iqv);
        switch (id) {
          case RCM_SL_TB3D: v = m_t(j, i, k); break;
          case RCM_SL_QVB3D: v = qvv; break;
          case RCM_SL_QCB3D: v = inner(j, i) ? std::max(m_qx(j, i, k, iqc), d_zero) : m_qx(j, i, k, iqc); break;
          case RCM_SL_UBX3D: v = m_ux(j, i, k); break;
          case RCM_SL_VBX3D: v = m_vx(j, i, k); break;
          case RCM_SL_PB3D: v = m_p(j, i, k); break;
          case RCM_SL_PF3D:
            if (k == 1) v = 100.0;
            else if (k == kzp1) v = psa(j, i);
            else v = p00 * std::pow(d_half * (m_pai(j, i, k) + m_pai(j, i, k - 1)), cpovr);
            break;
          case RCM_SL_PS2D: v = psa(j, i); break;
          case RCM_SL_RHOB3D: v = m_rho(j, i, k); break;
          case RCM_SL_TH3D: v = m_t(j, i, k) * std::pow(p00 / m_p(j, i, k), rovcp); break;
          case RCM_SL_TV3D: v = m_tvirt(j, i, k); break;
          case RCM_SL_TP3D: v = inner(j, i) ? m_t(j, i, k) * std::pow(psa(j, i) / m_p(j, i, k), rovcp) : d_zero; break;
          case RCM_SL_WPX3D: v = inner(j, i) ? -d_half * egrav * m_rho(j, i, k) * (m_w(j, i, k) + m_w(j, i, k + 1)) : d_zero; break;
          case RCM_SL_WB3D: v = m_w(j, i, k); break;
          case RCM_SL_ZQ: v = m_zetaf(j, i, k); break;
          case RCM_SL_ZA: v = m_zeta(j, i, k); break;
          case RCM_SL_DZQ: v = m_dz(j, i, k); break;
          case RCM_SL_QSB3D: v = pfwsat(m_t(j, i, k), m_p(j, i, k)); break;
          case RCM_SL_RHB3D: v = inner(j, i) ? std::min(std::max(qvv / pfwsat(m_t(j, i, k), m_p(j, i, k)), rhmin), rhmax) : d_zero; break;
        }
        out(j, i, k) = v;
      }
  const long long n1 = ub[0] - lb[0] + 1, n2 = ub[1] - lb[1] + 1;
  for (int k = std::max(lb[2], 1); k <= std::min(ub[2], nk); ++k)
    for (int i = std::max(lb[1], ice1); i <= std::min(ub[1], ice2); ++i)
      for (int j = std::max(lb[0], jce1); j <= std::min(ub[0], jce2); ++j)
        a[(j - lb[0]) + n1 * ((i - lb[1]) + n2 * (long long)(k - lb[2]))] = out(j, i, k);
  return RCM_OK;
}

int Oracle::set_phys(int field, const double* a, const int32_t lb[3], const int32_t ub[3]) {
  if (moloch) return mo_set_phys(field, a, lb, ub);
  ensure_phys();
  A3* dst = nullptr;
  int n4 = 0;
  switch (field) {
    case RCM_F_U: dst = &uphy; break;
    case RCM_F_V: dst = &vphy; break;
    case RCM_F_T: dst = &tphy; break;
    case RCM_F_PP: dst = &ppphy; break;
    case RCM_F_W: dst = &wphy; break;
    case RCM_F_TKE: if (!uw_on) { err = "oracle: TKE is carried with ibltyp = 2 only"; return RCM_ERR_ARG; } dst = &tkephy; break;
    case RCM_F_QV: n4 = iqv; break;
    case RCM_F_QC: n4 = iqc; break;
    default: err = "oracle: no physics tendency for this field"; return RCM_ERR_ARG;
  }
  if (hydro && (field == RCM_F_PP || field == RCM_F_W)) { err = "oracle: p', w tendencies exist for idynamic = 2 only"; return RCM_ERR_ARG; }
  const long long n1 = ub[0] - lb[0] + 1, n2 = ub[1] - lb[1] + 1;
  const int l1 = dst ? dst->l1 : qxphy.l1, l2 = dst ? dst->l2 : qxphy.l2, m1 = dst ? dst->n1 : qxphy.n1, m2 = dst ? dst->n2 : qxphy.n2,
            m3 = dst ? dst->n3 : qxphy.n3;
  for (int k = std::max(lb[2], 1); k <= std::min(ub[2], m3); ++k)
    for (int i = std::max(lb[1], l2); i <= std::min(ub[1], l2 + m2 - 1); ++i)
      for (int j = std::max(lb[0], l1); j <= std::min(ub[0], l1 + m1 - 1); ++j) {
        const double v = a[(j - lb[0]) + n1 * ((i - lb[1]) + n2 * (long long)(k - lb[2]))];
        if (dst) (*dst)(j, i, k) = v; else qxphy(j, i, k, n4) = v;
      }
  return RCM_OK;
}

void Oracle::clear_phys() {
  if (moloch) { if (mo_phys_on) { m_tten.fill(d_zero); m_uten.fill(d_zero); m_vten.fill(d_zero); m_qxten.fill(d_zero); } return; }
  if (!phys_on) return;
  for (A3* a : {&uphy, &vphy, &tphy, &ppphy, &wphy}) a->fill(d_zero);
  if (uw_on) tkephy.fill(d_zero);
  qxphy.fill(d_zero);
}

// One slice of Main/mod_slice.F90:167-308 on the cross grid (values outside the range the reference defines are 0).
int Oracle::get_slice(int id, double* a, const int32_t lb[3], const int32_t ub[3]) {
  if (moloch) return mo_get_slice(id, a, lb, ub);
  A3 out; out.alloc(jce1, jce2, ice1, ice2, 1, kzp1);
  const double ptoppa = ptop * d_1000;
  auto rps = [&](int j, int i) { return d_one / psb(j, i); };
  auto tb = [&](int j, int i, int k) { return a2_t(j, i, k) * rps(j, i); };
  auto qv = [&](int j, int i, int k) { return std::max(a2_qx(j, i, k, iqv) * rps(j, i), minqq); };
  auto qc = [&](int j, int i, int k) { return std::max(a2_qx(j, i, k, iqc) * rps(j, i), d_zero); };
  auto ppb = [&](int j, int i, int k) { return a2_pp(j, i, k) * rps(j, i); };
  auto pb = [&](int j, int i, int k) {
    if (hydro) return (hsigma(k) * psb(j, i) + ptop) * d_1000;
    if (k == 1) return std::max(a0_pr(j, i, 1) + ppb(j, i, 1), ptoppa + 1.0);
    return a0_pr(j, i, k) + ppb(j, i, k);
  };
  auto ps2 = [&](int j, int i) { return hydro ? (psb(j, i) + ptop) * d_1000 : a0_ps(j, i) + ptoppa + ppb(j, i, kz); };
  auto pf = [&](int j, int i, int k) {
    if (hydro) return (sigma(k) * psb(j, i) + ptop) * d_1000;
    if (k == 1) return ptoppa;
    if (k == kzp1) return ps2(j, i);
    return a0_pf(j, i, k) + d_half * (ppb(j, i, k - 1) + ppb(j, i, k));
  };
  auto tv = [&](int j, int i, int k) { return tb(j, i, k) * (d_one + ep1 * qv(j, i, k) - qc(j, i, k)); };
  auto rhob = [&](int j, int i, int k) { return pb(j, i, k) / (rgas * tb(j, i, k)); };
  auto wpx = [&](int j, int i, int k) { return hydro ? omega(j, i, k) * d_1000 : omega(j, i, k); };
  auto ubd = [&](const A3& u, int j, int i, int k) { return u(j, i, k) * (d_one / psdotb(j, i)); };
  const bool interior_only = (id == RCM_SL_WPX3D) || (hydro && id == RCM_SL_WB3D);
  const int j1 = interior_only ? jci1 : jce1, j2 = interior_only ? jci2 : jce2, i1 = interior_only ? ici1 : ice1, i2 = interior_only ? ici2 : ice2;
  int nk = kz;
  if (id == RCM_SL_PF3D || id == RCM_SL_WB3D || id == RCM_SL_ZQ) nk = kzp1;
  if (id == RCM_SL_PS2D) nk = 1;
  if (id < 0 || id >= RCM_SL_COUNT) { err = "oracle: bad slice id"; return RCM_ERR_ARG; }
  for (int i = i1; i <= i2; ++i)
    for (int j = j1; j <= j2; ++j) {
      if (hydro && (id == RCM_SL_ZQ || id == RCM_SL_ZA || id == RCM_SL_DZQ)) {   // :274-292
        std::vector<double> zq(kzp1 + 2, d_zero);
        const double cell = ptop * rps(j, i);
        for (int k = kz; k >= 1; --k) zq[k] = zq[k + 1] + rovg * tv(j, i, k) * std::log((sigma(k + 1) + cell) / (sigma(k) + cell));
        for (int k = 1; k <= nk; ++k)
          out(j, i, k) = id == RCM_SL_ZQ ? zq[k] : id == RCM_SL_ZA ? d_half * (zq[k] + zq[k + 1]) : zq[k] - zq[k + 1];
        continue;
      }
      for (int k = 1; k <= nk; ++k) {
        double v = d_zero;
        switch (id) {
          case RCM_SL_TB3D: v = tb(j, i, k); break;
          case RCM_SL_QVB3D: v = qv(j, i, k); break;
          case RCM_SL_QCB3D: v = qc(j, i, k); break;
          case RCM_SL_UBX3D: v = d_rfour * (ubd(a2_u, j, i, k) + ubd(a2_u, j, i + 1, k) + ubd(a2_u, j + 1, i, k) + ubd(a2_u, j + 1, i + 1, k)); break;
          case RCM_SL_VBX3D: v = d_rfour * (ubd(a2_v, j, i, k) + ubd(a2_v, j, i + 1, k) + ubd(a2_v, j + 1, i, k) + ubd(a2_v, j + 1, i + 1, k)); break;
          case RCM_SL_PB3D: v = pb(j, i, k); break;
          case RCM_SL_PF3D: v = pf(j, i, k); break;
          case RCM_SL_PS2D: v = ps2(j, i); break;
          case RCM_SL_RHOB3D: v = rhob(j, i, k); break;
          case RCM_SL_TH3D: v = tb(j, i, k) * std::pow(p00 / pb(j, i, k), rovcp); break;
          case RCM_SL_TV3D: v = tv(j, i, k); break;
          case RCM_SL_TP3D: v = tb(j, i, k) * std::pow(ps2(j, i) / pb(j, i, k), rovcp); break;
          case RCM_SL_WPX3D: v = wpx(j, i, k); break;
          case RCM_SL_WB3D:
            if (!hydro) v = a2_w(j, i, k) * rps(j, i);
            else if (k >= 2 && k <= kz) v = -d_half * regrav * (wpx(j, i, k - 1) / rhob(j, i, k - 1) + wpx(j, i, k) / rhob(j, i, k));
            break;
          case RCM_SL_ZQ: v = a0_zf(j, i, k); break;      // constant for the non-hydrostatic core
          case RCM_SL_ZA: v = a0_z(j, i, k); break;
          case RCM_SL_DZQ: v = a0_dzf(j, i, k); break;
          case RCM_SL_QSB3D: v = pfwsat(tb(j, i, k), pb(j, i, k)); break;
          case RCM_SL_RHB3D: v = std::min(std::max(qv(j, i, k) / pfwsat(tb(j, i, k), pb(j, i, k)), rhmin), rhmax); break;
        }
        out(j, i, k) = v;
      }
    }
  const long long n1 = ub[0] - lb[0] + 1, n2 = ub[1] - lb[1] + 1;
  for (int k = std::max(lb[2], 1); k <= std::min(ub[2], nk); ++k)
    for (int i = std::max(lb[1], ice1); i <= std::min(ub[1], ice2); ++i)
      for (int j = std::max(lb[0], jce1); j <= std::min(ub[0], jce2); ++j)
        a[(j - lb[0]) + n1 * ((i - lb[1]) + n2 * (long long)(k - lb[2]))] = out(j, i, k);
  return RCM_OK;
}

}  // namespace orc

extern "C" {
int orc_step_pre_physics(void* h) { ((orc::Oracle*)h)->pre_physics(); return RCM_OK; }
int orc_step_post_physics(void* h) { return ((orc::Oracle*)h)->step(1); }   // the head is idempotent: step() repeats it
int orc_set_phys_tend(void* h, int field, const double* a, const int32_t* lb, const int32_t* ub) { return ((orc::Oracle*)h)->set_phys(field, a, lb, ub); }
int orc_clear_phys_tend(void* h) { ((orc::Oracle*)h)->clear_phys(); return RCM_OK; }
int orc_get_slice(void* h, int id, double* a, const int32_t* lb, const int32_t* ub) { return ((orc::Oracle*)h)->get_slice(id, a, lb, ub); }
}
