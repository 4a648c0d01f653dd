// ORACLE — TEST INFRASTRUCTURE ONLY.
// Restates Main/mod_sladvection.F90 (isladvec = 1) as `tend` uses it for the moisture species and the
// tracers (Main/mod_tendency.F90:995-1000,1283-1285,1361-1363,1378-1380,1397-1399): adv_velocity(.false.) :67-119,
// trajcalc_x :121-229, slhadv_x3d/x4d :335-479 on atm2 (time level b, p*-coupled), hdvg_x3d/x4d :545-660 on atm1.
// ua, va = atmx%umd, atmx%vmd; mapfx, mapfd = mddom%msfx, mddom%msfd as stored (:58-65).
#include <algorithm>
#include <cmath>

#include "model.hpp"

namespace orc {

int Oracle::trajcalc_x() {
  const A3& ua = ax_umd; const A3& va = ax_vmd;
  const A2& mapfx = msfx;
  if (sl_xn.empty()) {
    for (A3* a : {&sl_alf, &sl_am2, &sl_am1, &sl_a0, &sl_ap1, &sl_bm2, &sl_bm1, &sl_b0, &sl_bp1}) a->alloc(jci1, jci2, ici1, ici2, 1, kz);
    for (auto* a : {&sl_xn, &sl_xm1, &sl_xm2, &sl_xp1, &sl_yn, &sl_ym1, &sl_ym2, &sl_yp1}) a->assign((size_t)(jci2 - jci1 + 1) * (ici2 - ici1 + 1) * kz, 0);
  }
  const double dtsq = dt * dt, dtcb = dt * dt * dt;      // Main/mod_params.F90:1709-1710, Main/mod_tendency.F90:614-615
  const double ddx = dx, ddy = dx;
  const int nj = jci2 - jci1 + 1, ni = ici2 - ici1 + 1;
  for (int k = 1; k <= kz; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) {
        // adv_velocity, cross points (:92-114; the last term of vadym1_x repeats va(j+1,i,k) as the reference does)
        const double uadvx = 0.25 * (ua(j, i, k) + ua(j, i + 1, k) + ua(j + 1, i + 1, k) + ua(j + 1, i, k)) / mapfx(j, i);
        const double uadxp1 = 0.25 * (ua(j + 1, i, k) + ua(j + 1, i + 1, k) + ua(j + 2, i + 1, k) + ua(j + 2, i, k)) / mapfx(j + 1, i);
        const double uadxm1 = 0.25 * (ua(j, i, k) + ua(j, i + 1, k) + ua(j - 1, i + 1, k) + ua(j - 1, i, k)) / mapfx(j - 1, i);
        const double vadvy = 0.25 * (va(j, i, k) + va(j, i + 1, k) + va(j + 1, i + 1, k) + va(j + 1, i, k)) / mapfx(j, i);
        const double vadyp1 = 0.25 * (va(j, i + 1, k) + va(j + 1, i + 1, k) + va(j + 1, i + 2, k) + va(j, i + 2, k)) / mapfx(j, i + 1);
        const double vadym1 = 0.25 * (va(j, i, k) + va(j, i - 1, k) + va(j + 1, i, k) + va(j + 1, i, k)) / mapfx(j, i - 1);
        const size_t q = (size_t)(j - jci1) + (size_t)nj * ((size_t)(i - ici1) + (size_t)ni * (size_t)(k - 1));
        const double ux = 0.5 * (uadxp1 - uadxm1) / ddx;
        const double uxx = (uadxp1 - 2.0 * uadvx + uadxm1) / (ddx * ddx);
        const double xdis = -uadvx * dt + 0.5 * (dtsq * uadvx * ux) - (dtcb * uadvx) * (ux * ux + uadvx * uxx) / 6.0;
        const double xn = xdis / ddx;
        const int xnp = (int)xn;
        if (std::abs(xnp) > 1) { err = "SL advection: departure point more than one cell away in the WE direction"; return RCM_ERR_CFL; }
        const double alfax = std::fabs((xnp * ddx - xdis) / ddx);
        const int xsn = (int)std::copysign(d_one, xn);
        int x0 = j + xnp, xm1 = x0 + xsn, xm2 = xm1 + xsn, xp1 = x0 - xsn;
        for (int* p : {&x0, &xm1, &xm2, &xp1}) { if (has_bdyleft && *p < jce1) *p = jce1; if (has_bdyright && *p > jce2) *p = jce2; }
        const double vy = 0.5 * (vadyp1 - vadym1) / ddy;
        const double vyy = (vadyp1 - 2.0 * vadvy + vadym1) / (ddy * ddy);
        const double ydis = -vadvy * dt + 0.5 * (dtsq * vadvy * vy) - (dtcb * vadvy) * (vy * vy + vadvy * vyy) / 6.0;
        const double yn = ydis / ddy;
        const int ynp = (int)yn;
        if (std::abs(ynp) > 1) { err = "SL advection: departure point more than one cell away in the SN direction"; return RCM_ERR_CFL; }
        const double betay = std::fabs((ynp * ddy - ydis) / ddy);
        const int ysn = (int)std::copysign(d_one, yn);
        int y0 = i + ynp, ym1 = y0 + ysn, ym2 = ym1 + ysn, yp1 = y0 - ysn;
        for (int* p : {&y0, &ym1, &ym2, &yp1}) { if (has_bdybottom && *p < ice1) *p = ice1; if (has_bdytop && *p > ice2) *p = ice2; }
        sl_xn[q] = x0; sl_xm1[q] = xm1; sl_xm2[q] = xm2; sl_xp1[q] = xp1;
        sl_yn[q] = y0; sl_ym1[q] = ym1; sl_ym2[q] = ym2; sl_yp1[q] = yp1;
        sl_alf(j, i, k) = alfax;
        sl_am2(j, i, k) = -(alfax * (1.0 - alfax * alfax)) / 6.0;
        sl_am1(j, i, k) = (alfax * (1.0 + alfax) * (2.0 - alfax)) / 2.0;
        sl_a0(j, i, k) = ((1.0 - alfax * alfax) * (2.0 - alfax)) / 2.0;
        sl_ap1(j, i, k) = -(alfax * (1.0 - alfax) * (2.0 - alfax)) / 6.0;
        sl_bm2(j, i, k) = -(betay * (1.0 - betay * betay)) / 6.0;
        sl_bm1(j, i, k) = (betay * (1.0 + betay) * (2.0 - betay)) / 2.0;
        sl_b0(j, i, k) = ((1.0 - betay * betay) * (2.0 - betay)) / 2.0;
        sl_bp1(j, i, k) = -(betay * (1.0 - betay) * (2.0 - betay)) / 6.0;
      }
  return RCM_OK;
}

// slhadv_x3d / one species of slhadv_x4d: `var` is read as var(j, i, k)
void Oracle::slhadv_x(const std::function<double&(int, int, int)>& ften, const std::function<double(int, int, int)>& var) {
  const int nj = jci2 - jci1 + 1, ni = ici2 - ici1 + 1;
  for (int k = 1; k <= kz; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) {
        const size_t q = (size_t)(j - jci1) + (size_t)nj * ((size_t)(i - ici1) + (size_t)ni * (size_t)(k - 1));
        const int x0 = sl_xn[q], xm1 = sl_xm1[q], xm2 = sl_xm2[q], xp1 = sl_xp1[q];
        const int y0 = sl_yn[q], ym1 = sl_ym1[q], ym2 = sl_ym2[q], yp1 = sl_yp1[q];
        const double alf = sl_alf(j, i, k);
        const double bl1 = alf * var(xm1, yp1, k) + (d_one - alf) * var(x0, yp1, k);
        const double bl2 = alf * var(xm1, ym2, k) + (d_one - alf) * var(x0, ym2, k);
        const double cb1 = sl_am2(j, i, k) * var(xm2, y0, k) + sl_am1(j, i, k) * var(xm1, y0, k) + sl_a0(j, i, k) * var(x0, y0, k) +
                           sl_ap1(j, i, k) * var(xp1, y0, k);
        const double cb2 = sl_am2(j, i, k) * var(xm2, ym1, k) + sl_am1(j, i, k) * var(xm1, ym1, k) + sl_a0(j, i, k) * var(x0, ym1, k) +
                           sl_ap1(j, i, k) * var(xp1, ym1, k);
        const double tbadp = sl_bm2(j, i, k) * bl2 + sl_bm1(j, i, k) * cb2 + sl_b0(j, i, k) * cb1 + sl_bp1(j, i, k) * bl1;
        double tsla = tbadp;
        if (cfg.iqmsl == 1) {
          const double tbmax = std::max(std::max(var(x0, y0, k), var(x0, ym1, k)), std::max(var(xm1, y0, k), var(xm1, ym1, k)));
          const double tbmin = std::min(std::min(var(x0, y0, k), var(x0, ym1, k)), std::min(var(xm1, y0, k), var(xm1, ym1, k)));
          if (tbadp > tbmax) tsla = tbmax;
          else if (tbadp < tbmin) tsla = tbmin;
        }
        if (std::fabs(tsla - var(j, i, k)) > dlowval) ften(j, i, k) = ften(j, i, k) + (tsla - var(j, i, k)) / dt;
      }
}

// hdvg_x3d / one species of hdvg_x4d
void Oracle::hdvg_x(const std::function<double&(int, int, int)>& ften, const std::function<double(int, int, int)>& var) {
  const A3& ua = ax_umd; const A3& va = ax_vmd;
  const A2& mapfx = msfx; const A2& mapfd = msfd;
  for (int k = 1; k <= kz; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) {
        const double ucapf = (ua(j + 1, i + 1, k) * mapfd(j + 1, i + 1) + ua(j + 1, i, k) * mapfd(j + 1, i)) * d_half;
        const double ucapi = (ua(j, i + 1, k) * mapfd(j, i + 1) + ua(j, i, k) * mapfd(j, i)) * d_half;
        const double vcapf = (va(j + 1, i + 1, k) * mapfd(j + 1, i + 1) + va(j, i + 1, k) * mapfd(j, i + 1)) * d_half;
        const double vcapi = (va(j + 1, i, k) * mapfd(j + 1, i) + va(j, i, k) * mapfd(j, i)) * d_half;
        const double ducapdx = (ucapf - ucapi) / dx;
        const double dvcapdy = (vcapf - vcapi) / dx;
        const double hdvg = (ducapdx + dvcapdy) / (mapfx(j, i) * mapfx(j, i));
        const double t = (var(j, i, k) > 2.220446049250313e-16) ? var(j, i, k) * hdvg : d_zero;   // epsilon(d_one)
        ften(j, i, k) = ften(j, i, k) - t;
      }
}

}  // namespace orc
