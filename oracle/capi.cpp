// ORACLE — TEST INFRASTRUCTURE ONLY.
// C entry points (ctypes) mirroring include/regcm_dyn.h one-to-one, so that the
// parity tests drive the CUDA library and the CPU restatement with the same
// calls.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
// --impl reference legs may load this library.
#include <cstring>

#include "model.hpp"

using orc::Oracle;

extern "C" {

int orc_create(const rcm_cfg* cfg, void** h) {
  if (!cfg || !h) return RCM_ERR_ARG;
  Oracle* o = new Oracle();
  int rc = o->create(*cfg);
  if (rc != RCM_OK) { std::fprintf(stderr, "orc_create: %s\n", o->err.c_str()); delete o; *h = nullptr; return rc; }
  *h = o;
  return RCM_OK;
}
int orc_set_static(void* h, int id, const double* a, const int32_t* lb, const int32_t* ub) {
  return ((Oracle*)h)->set_static(id, a, lb, ub);
}
int orc_set_bdy(void* h, int field, int slot, const double* a, const int32_t* lb, const int32_t* ub) {
  return ((Oracle*)h)->set_bdy(field, slot, a, lb, ub);
}
int orc_init(void* h) { return ((Oracle*)h)->init(); }
int orc_step(void* h, int n) { return ((Oracle*)h)->step(n); }
int orc_get_state(void* h, int field, int level, double* a, const int32_t* lb, const int32_t* ub) {
  return ((Oracle*)h)->get_state(field, level, a, lb, ub);
}
int orc_set_state(void* h, int field, int level, const double* a, const int32_t* lb, const int32_t* ub) {
  return ((Oracle*)h)->set_state(field, level, a, lb, ub);
}
int orc_get_work(void* h, int id, double* a, const int32_t* lb, const int32_t* ub) {
  return ((Oracle*)h)->get_work(id, a, lb, ub);
}
int orc_diag(void* h, rcm_diag* d) {
  Oracle* o = (Oracle*)h;
  std::memset(d, 0, sizeof(*d));
  d->lcount = o->lcount; d->xbctime = o->xbctime; d->dt = o->dt;
  d->max_cfl = o->max_cfl_step; d->istep = o->last_istep;
  d->neg_q_points = o->neg_q_points;
  d->neg_chi_points = o->neg_chi_points;
  d->ptntot = o->ptntot; d->pt2tot = o->pt2tot;
  d->ps_max = o->ps_max; d->ps_min = o->ps_min;
  return RCM_OK;
}
int orc_last_error(void* h, char* buf, int len) {
  Oracle* o = (Oracle*)h;
  std::strncpy(buf, o->err.c_str(), (size_t)len - 1); buf[len - 1] = 0;
  return RCM_OK;
}
// 2-D/1-D setup products, for unit tests of the initialisation path
int orc_get_aux(void* h, const char* name, double* out, int n) {
  Oracle* o = (Oracle*)h;
  const std::vector<double>* v = nullptr;
  std::string s(name);
  if (s == "a0_ps") v = &o->a0_ps.d; else if (s == "a0_pr") v = &o->a0_pr.d;
  else if (s == "a0_t") v = &o->a0_t.d; else if (s == "a0_rho") v = &o->a0_rho.d;
  else if (s == "a0_z") v = &o->a0_z.d; else if (s == "a0_zf") v = &o->a0_zf.d;
  else if (s == "tmask") v = &o->tmask.d; else if (s == "twt") v = &o->twt.d;
  else if (s == "hefc") v = &o->hefc.d; else if (s == "dpsdxm") v = &o->dpsdxm.d;
  else if (s == "psdota") v = &o->psdota.d; else if (s == "ex") v = &o->ex.d;
  else if (s == "crx") v = &o->crx.d; else if (s == "dmdx") v = &o->dmdx.d;
  else return RCM_ERR_ARG;
  if ((size_t)n < v->size()) return (int)v->size();
  std::memcpy(out, v->data(), v->size() * sizeof(double));
  return (int)v->size();
}
int orc_get_table(void* h, int id, double* out, int n) { return ((Oracle*)h)->get_table(id, out, n); }
int orc_get_zone(void* h, int dot, int32_t* out) {
  // packed zone map: bits 0..3 = south,north,west,east; bits 8.. = ibnd
  Oracle* o = (Oracle*)h;
  orc::BoundArea& ba = dot ? o->ba_dt : o->ba_cr;
  for (int i = 1; i <= o->iy; ++i)
    for (int j = 1; j <= o->jx; ++j) {
      int v = (ba.bsouth(j, i) ? 1 : 0) | (ba.bnorth(j, i) ? 2 : 0) | (ba.bwest(j, i) ? 4 : 0) | (ba.beast(j, i) ? 8 : 0);
      int ib = ba.ibnd(j, i); if (ib < 0) ib = 0;
      out[(j - 1) + (size_t)o->jx * (i - 1)] = v | (ib << 8);
    }
  return RCM_OK;
}
// test hook: make the negative-moisture sweep behave as on a decomposed domain (see repair_negative_qx)
int orc_set_repair_tiles(void* h, int ncj, const int32_t* jstart, int nci, const int32_t* istart) {
  Oracle* o = (Oracle*)h;
  o->rt_j.assign(jstart, jstart + ncj);
  o->rt_i.assign(istart, istart + nci);
  return RCM_OK;
}
int orc_destroy(void* h) { delete (Oracle*)h; return RCM_OK; }

}  // extern "C"
