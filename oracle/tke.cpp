// ORACLE — TEST INFRASTRUCTURE ONLY.
// UW-PBL coupling of the transport, sigma-coordinate cores (ibltyp = 2):
//   TKE state         atm1%tke, atm2%tke, atmc%tke on (jce1:jce2, ice1:ice2, 1:kzp1), NOT coupled with p*
//                     (Main/mod_atm_interface.F90:604-606,625-626,656-658); tkemin at start (Main/mod_init.F90:378-380)
//   advection         hadv3d(tkedyn, atm1%tke, ind = 1); tkeps = atm1%tke * psa; vadv3d(tkedyn, tkeps, kzp1, 0)
//                     (Main/mod_tendency.F90:1414-1425; Main/mod_advection.F90:395-515,755-766)
//   diffusion         diffu_x3df(tkedyn, atm2%tke, nuk)   (Main/mod_tendency.F90:1544-1548; Main/mod_diffusion.F90:523-601)
//   forecast, filter  tketen = tketen + tkedyn * rpsa + tkephy; atmc%tke = max(tkemin, atm2%tke + dt * tketen);
//                     filter_ra_3d(atm1%tke, atm2%tke, atmc%tke, gnu2)   (Main/mod_tendency.F90:518-544)
//   bdyval            atm2 rim <- atm1 rim when integrating (Main/mod_bdycod.F90:1166-1172,1212-1218,1258-1264,1301-1307);
//                     start: rim = tkemin; afterwards inflow -> tkemin, outflow -> interior neighbour, on levels 3..kzp1,
//                     level 1 = tkemin (:2412-2496; bdyflow = .true., :118)
//   vadv4d ind = 3    (iuwvadv = 1) for qc and the tracers: twt interpolation, reshaped around the PBL top kpbl
//                     (Main/mod_advection.F90:915-957; Main/mod_tendency.F90:148-156)
#include <algorithm>
#include <cmath>

#include "model.hpp"

namespace orc {

void Oracle::allocate_tke() {
  uw_on = (cfg.ibltyp == 2) && !moloch;
  iqxvadv = 1; itrvadv = 2;
  if (!uw_on) return;
  if (cfg.iuwvadv == 1) { iqxvadv = 3; itrvadv = 3; }
  tkemin = cfg.uw_tkemin;
  for (A3* a : {&tke_a, &tke_b, &tke_c, &tkeps}) a->alloc(jce1, jce2, ice1, ice2, 1, kzp1);
  for (A3* a : {&tkedyn, &tketen, &tkephy}) a->alloc(jci1, jci2, ici1, ici2, 1, kzp1);
  kpb.assign((size_t)(jci2 - jci1 + 1) * (ici2 - ici1 + 1), 0);   // getmem2d leaves zeros (Main/mod_atm_interface.F90:1135)
  tke_a.fill(tkemin); tke_b.fill(tkemin); tke_c.fill(d_zero);
}

int Oracle::set_kpbl(const int32_t* k, const int32_t lb[2], const int32_t ub[2]) {
  if (!uw_on) { err = "oracle: kpbl exists with ibltyp = 2 only"; return RCM_ERR_ARG; }
  const int n1 = ub[0] - lb[0] + 1, nj = jci2 - jci1 + 1;
  for (int i = std::max(lb[1], ici1); i <= std::min(ub[1], ici2); ++i)
    for (int j = std::max(lb[0], jci1); j <= std::min(ub[0], jci2); ++j) {
      const int v = k[(j - lb[0]) + (size_t)n1 * (i - lb[1])];
      if (v > kz) { err = "kpbl is greater than kz"; return RCM_ERR_ARG; }
      kpb[(j - jci1) + (size_t)nj * (i - ici1)] = v;
    }
  return RCM_OK;
}

// the flux array of vadv4d ind = 3 (Main/mod_advection.F90:915-957) for one species `f` (p*-coupled), svv included
void Oracle::vadv_ind3(const std::function<double(int, int, int)>& f) {
  const A3& svv = qdot;
  const int nj = jci2 - jci1 + 1;
  fg.fill(d_zero);
  for (int k = 2; k <= kz; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) fg(j, i, k) = twt(k, 1) * f(j, i, k) + twt(k, 2) * f(j, i, k - 1);
  for (int i = ici1; i <= ici2; ++i)
    for (int j = jci1; j <= jci2; ++j) {
      const int kp = kpb[(j - jci1) + (size_t)nj * (i - ici1)];
      if (kp >= 4) {
        int k = kp - 2;
        double slope;
        if ((f(j, i, k + 1) - f(j, i, k)) > d_zero && (f(j, i, k) - f(j, i, k - 1)) > d_zero)
          slope = std::min((f(j, i, k + 1) - f(j, i, k)) / (hsigma(k + 1) - hsigma(k)), (f(j, i, k) - f(j, i, k - 1)) / (hsigma(k) - hsigma(k - 1)));
        else if ((f(j, i, k + 1) - f(j, i, k)) < d_zero && (f(j, i, k) - f(j, i, k - 1)) < d_zero)
          slope = std::max((f(j, i, k + 1) - f(j, i, k)) / (hsigma(k + 1) - hsigma(k)), (f(j, i, k) - f(j, i, k - 1)) / (hsigma(k) - hsigma(k - 1)));
        else
          slope = d_zero;
        k = kp;
        fg(j, i, k - 1) = f(j, i, k - 2) + slope * (sigma(k - 1) - hsigma(k - 2));
        if (std::fabs(f(j, i, k - 2) + slope * (hsigma(k - 1) - hsigma(k - 2)) - f(j, i, k)) > std::fabs(f(j, i, k - 1) - f(j, i, k)))
          fg(j, i, k) = f(j, i, k);
        else
          fg(j, i, k) = f(j, i, k - 2) + slope * (sigma(k) - hsigma(k - 2));
      }
    }
  for (int k = 2; k <= kz; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) fg(j, i, k) = fg(j, i, k) * svv(j, i, k);
}

void Oracle::tke_tendency() {
  if (!uw_on) return;
  tkedyn.fill(d_zero); tketen.fill(d_zero);          // aten%tke = 0 (Main/mod_tendency.F90:1244-1246)
  hadv3d(tkedyn, tke_a, 1);
  for (int k = 1; k <= kzp1; ++k)
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j) tkeps(j, i, k) = tke_a(j, i, k) * psa(j, i);
  vadv3d(tkedyn, tkeps, kzp1, 0);
  diffu_x3df(tkedyn, tke_b, cfg.uw_nuk);
  for (int k = 1; k <= kzp1; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) tketen(j, i, k) = tketen(j, i, k) + tkedyn(j, i, k) * rpsa(j, i) + tkephy(j, i, k);
  for (int k = 1; k <= kzp1; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) tke_c(j, i, k) = std::max(tkemin, tke_b(j, i, k) + dt * tketen(j, i, k));
  // filter_ra_3d (Main/mod_timefilter.F90:92-110), nk = kzp1
  const double alpha = cfg.gnu2;
  for (int k = 1; k <= kzp1; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) {
        const double d = alpha * (tke_c(j, i, k) + tke_b(j, i, k) - d_two * tke_a(j, i, k));
        tke_b(j, i, k) = tke_a(j, i, k) + d;
        tke_a(j, i, k) = tke_c(j, i, k);
      }
}

// the copy atm2 rim <- atm1 rim (first part of bdyval, when integrating)
void Oracle::bdyval_tke_copy() {
  if (!uw_on) return;
  for (int k = 1; k <= kzp1; ++k) for (int i = ici1; i <= ici2; ++i) tke_b(jce1, i, k) = tke_a(jce1, i, k);
  for (int k = 1; k <= kzp1; ++k) for (int i = ici1; i <= ici2; ++i) tke_b(jce2, i, k) = tke_a(jce2, i, k);
  for (int k = 1; k <= kzp1; ++k) for (int j = jce1; j <= jce2; ++j) tke_b(j, ice1, k) = tke_a(j, ice1, k);
  for (int k = 1; k <= kzp1; ++k) for (int j = jce1; j <= jce2; ++j) tke_b(j, ice2, k) = tke_a(j, ice2, k);
}

// Main/mod_bdycod.F90:2412-2496 (idynamic /= 3)
void Oracle::bdyval_tke() {
  if (!uw_on) return;
  if (!integrating()) {    // rcmtimer%start()
    for (A3* a : {&tke_a, &tke_b})
      for (int k = 1; k <= kzp1; ++k) {
        for (int i = ice1; i <= ice2; ++i) { (*a)(jce1, i, k) = tkemin; (*a)(jce2, i, k) = tkemin; }
        for (int j = jce1; j <= jce2; ++j) { (*a)(j, ice2, k) = tkemin; (*a)(j, ice1, k) = tkemin; }
      }
    return;
  }
  // west
  for (int i = ice1; i <= ice2; ++i) { tke_a(jce1, i, 1) = tkemin; tke_b(jce1, i, 1) = tkemin; }
  for (int k = 2; k <= kz; ++k)
    for (int i = ice1; i <= ice2; ++i) {
      const double tkeint = tke_a(jci1, i, k + 1);
      const double windavg = wue(i, k) + wue(i + 1, k) + wui(i, k) + wui(i + 1, k) + wue(i, k - 1) + wue(i + 1, k - 1) + wui(i, k - 1) + wui(i + 1, k - 1);
      tke_a(jce1, i, k + 1) = (windavg > d_zero) ? tkemin : tkeint;
    }
  // east
  for (int i = ice1; i <= ice2; ++i) { tke_a(jce2, i, 1) = tkemin; tke_b(jce2, i, 1) = tkemin; }
  for (int k = 2; k <= kz; ++k)
    for (int i = ice1; i <= ice2; ++i) {
      const double tkeint = tke_a(jci2, i, k + 1);
      const double windavg = eue(i, k) + eue(i + 1, k) + eui(i, k) + eui(i + 1, k) + eue(i, k - 1) + eue(i + 1, k - 1) + eui(i, k - 1) + eui(i + 1, k - 1);
      tke_a(jce2, i, k + 1) = (windavg < d_zero) ? tkemin : tkeint;
    }
  // south
  for (int j = jce1; j <= jce2; ++j) { tke_a(j, ice1, 1) = tkemin; tke_b(j, ice1, 1) = tkemin; }
  for (int k = 2; k <= kz; ++k)
    for (int j = jci1; j <= jci2; ++j) {
      const double tkeint = tke_a(j, ici1, k + 1);
      const double windavg = sve(j, k) + sve(j + 1, k) + svi(j, k) + svi(j + 1, k) + sve(j, k - 1) + sve(j + 1, k - 1) + svi(j, k - 1) + svi(j + 1, k - 1);
      tke_a(j, ice1, k + 1) = (windavg > d_zero) ? tkemin : tkeint;
    }
  // north
  for (int j = jce1; j <= jce2; ++j) { tke_a(j, ice2, 1) = tkemin; tke_b(j, ice2, 1) = tkemin; }
  for (int k = 2; k <= kz; ++k)
    for (int j = jci1; j <= jci2; ++j) {
      const double tkeint = tke_a(j, ici2, k + 1);
      const double windavg = nve(j, k) + nve(j + 1, k) + nvi(j, k) + nvi(j + 1, k) + nve(j, k - 1) + nve(j + 1, k - 1) + nvi(j, k - 1) + nvi(j + 1, k - 1);
      tke_a(j, ice2, k + 1) = (windavg < d_zero) ? tkemin : tkeint;
    }
}

int Oracle::tke_state(int level, double* a, const int32_t lb[3], const int32_t ub[3], bool in) {
  if (!uw_on) { err = "TKE is carried with ibltyp = 2 only"; return RCM_ERR_ARG; }
  A3& dst = level == RCM_LEVEL_A ? tke_a : tke_b;
  const int n1 = ub[0] - lb[0] + 1, n2 = ub[1] - lb[1] + 1;
  for (int k = std::max(lb[2], dst.l3); k <= std::min(ub[2], dst.l3 + dst.n3 - 1); ++k)
    for (int i = std::max(lb[1], dst.l2); i <= std::min(ub[1], dst.l2 + dst.n2 - 1); ++i)
      for (int j = std::max(lb[0], dst.l1); j <= std::min(ub[0], dst.l1 + dst.n1 - 1); ++j) {
        const size_t q = (size_t)(j - lb[0]) + (size_t)n1 * ((size_t)(i - lb[1]) + (size_t)n2 * (size_t)(k - lb[2]));
        if (in) dst(j, i, k) = a[q]; else a[q] = dst(j, i, k);
      }
  return RCM_OK;
}

}  // namespace orc

extern "C" int orc_set_kpbl(void* h, const int32_t* k, const int32_t* lb, const int32_t* ub) { return ((orc::Oracle*)h)->set_kpbl(k, lb, ub); }
