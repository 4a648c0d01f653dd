// ORACLE — TEST INFRASTRUCTURE ONLY.
// Named access to the oracle's internal arrays and single-routine entry points, so that the
// independent NumPy restatements in tests/numpy_ref.py (written from the Fortran text, not from
// this C++) can be fed the same inputs and compared with the same outputs.
#include <cstring>
#include <string>

#include "model.hpp"

using orc::Oracle;

namespace {
struct View { double* p; int32_t lb[4]; int32_t n[4]; };
void v1(View& v, orc::A1& a) { v = {a.d.data(), {a.l1, 1, 1, 1}, {a.n1, 1, 1, 1}}; }
void v2(View& v, orc::A2& a) { v = {a.d.data(), {a.l1, a.l2, 1, 1}, {a.n1, a.n2, 1, 1}}; }
void v3(View& v, orc::A3& a) { v = {a.d.data(), {a.l1, a.l2, a.l3, 1}, {a.n1, a.n2, a.n3, 1}}; }
void v4(View& v, orc::A4& a) { v = {a.d.data(), {a.l1, a.l2, a.l3, a.l4}, {a.n1, a.n2, a.n3, a.n4}}; }
}  // namespace

extern "C" {

// -> pointer into the oracle's own storage (column-major, lower bounds lb, extents n)
int orc_array(void* h, const char* name, double** data, int32_t lb[4], int32_t n[4]) {
  Oracle* o = (Oracle*)h;
  const std::string s(name);
  View v{nullptr, {1, 1, 1, 1}, {0, 0, 0, 0}};
#define T1(x) if (s == #x) v1(v, o->x); else
#define T2(x) if (s == #x) v2(v, o->x); else
#define T3(x) if (s == #x) v3(v, o->x); else
#define T4(x) if (s == #x) v4(v, o->x); else
  T1(sigma) T1(hsigma) T1(dsigma) T1(xds) T1(dds) T1(qcon) T2(twt)
  T2(msfx) T2(msfd) T2(xmsf) T2(dmsf) T2(ht) T2(psa) T2(psb) T2(psdota) T2(psdotb) T2(rpsa) T2(rpsb) T2(rpsda)
  T2(a0_ps) T3(a0_t) T3(a0_pr) T3(a0_rho) T3(a0_rhof) T3(a0_dprddx) T3(a0_dprddy) T2(dpsdxm) T2(dpsdym)
  T3(a1_u) T3(a1_v) T3(a1_t) T3(a1_pp) T3(a1_w) T3(a1_pr) T3(a1_rho) T4(a1_qx)
  T3(a2_u) T3(a2_v) T3(a2_t) T3(a2_pp) T3(a2_w) T3(a2_pr) T4(a2_qx)
  T3(ac_u) T3(ac_v) T3(ac_t) T3(ac_pp) T3(ac_w) T4(ac_qx)
  T3(ax_uc) T3(ax_vc) T3(ax_umc) T3(ax_vmc) T3(ax_ud) T3(ax_vd) T3(ax_umd) T3(ax_vmd)
  T3(uten) T3(vten) T3(ppten) T3(wten) T3(udyn) T3(vdyn) T3(qdot) T3(cr)
  T2(s_estore) T2(s_astore) T2(s_wpval) T2(tmask)
  T3(m_u) T3(m_v) T3(m_w) T3(m_s) T3(m_fmz) T3(m_fmzf) T3(m_tetav) T3(m_pai) T3(m_t) T4(m_qx) T3(m_ux) T3(m_vx) T3(m_wx)
  T2(mx2) T2(rmu) T2(rmv) T2(msfu) T2(msfv) T3(xkc) T3(th) T3(tha) T3(thten)
  if (s == "chia0" && o->ntr > 0) v3(v, o->chia[0]); else if (s == "chib0" && o->ntr > 0) v3(v, o->chib[0]); else
  if (s == "chic0" && o->ntr > 0) v3(v, o->chic[0]); else
  if (o->uw_on && s == "tke_a") v3(v, o->tke_a); else if (o->uw_on && s == "tke_b") v3(v, o->tke_b); else
  if (o->uw_on && s == "tke_c") v3(v, o->tke_c); else if (s == "xkcf") v3(v, o->xkcf); else
  return RCM_ERR_ARG;
#undef T1
#undef T2
#undef T3
#undef T4
  *data = v.p;
  for (int q = 0; q < 4; ++q) { lb[q] = v.lb[q]; n[q] = v.n[q]; }
  return RCM_OK;
}

// scalars the restatements need
int orc_scalar(void* h, const char* name, double* out) {
  Oracle* o = (Oracle*)h;
  const std::string s(name);
  if (s == "ul") *out = o->ul; else if (s == "dt") *out = o->dt; else if (s == "dtsec") *out = o->dtsec;
  else if (s == "dx") *out = o->dx; else if (s == "dtsmax") *out = o->dtsmax; else if (s == "xkd") *out = o->xkd_snd;
  else if (s == "bet") *out = o->bet; else if (s == "xgamma") *out = o->xgamma; else if (s == "rnpts") *out = o->rnpts;
  else if (s == "lcount") *out = (double)o->lcount; else if (s == "mo_dz") *out = o->mo_dz;
  else if (s == "alarm_day") *out = o->alarm_day_act() ? 1.0 : 0.0;
  else return RCM_ERR_ARG;
  return RCM_OK;
}

// make the next tend() return just before `call sound` (Main/mod_tendency.F90:500), leaving its inputs in place
int orc_break_before_sound(void* h, int on) { ((Oracle*)h)->break_before_sound = on != 0; return RCM_OK; }

// run ONE routine on the current state
int orc_call(void* h, const char* routine, double arg) {
  Oracle* o = (Oracle*)h;
  const std::string s(routine);
  if (s == "advect_uv") {   // Main/mod_tendency.F90:1274,1293,1304 on zeroed udyn, vdyn
    o->udyn.fill(0.0); o->vdyn.fill(0.0);
    o->start_advect();
    o->hadvuv(o->udyn, o->vdyn, o->ax_ud, o->ax_vd);
    o->vadvuv(o->udyn, o->vdyn, o->ax_uc, o->ax_vc);
  } else if (s == "advect_theta") {   // ithadv = 1 (Main/mod_tendency.F90:1348-1357): hadvt + vadv3d on theta of the last tend()
    o->thten.fill(0.0);
    o->start_advect();
    o->hadvt(o->thten, o->th);
    o->vadv3d(o->thten, o->tha, o->kz, 0);
  } else if (s == "sound") {
    o->sound();
  } else if (s == "wafone_tetav") {   // one wafone call on mo_atm%tetav with time step arg (Main/mod_moloch.F90:1098-1361)
    o->wafone(o->m_tetav, arg);
  } else {
    return RCM_ERR_ARG;
  }
  return RCM_OK;
}

}  // extern "C"
