// ORACLE — TEST INFRASTRUCTURE ONLY.
// Restates Main/mod_diffusion.F90 for idiffu = 1 (4th order, 2nd order on the boundary ring) and
// idiffu = 2 (the nine-point "LeVeque" operator on every interior point):
// calc_coeff :169-251, diffu_d :266-411, diffu_x3df :523-601,
// diffu_x3d :658-735, diffu_x4d/diffu_x4d3d :792-894.
// Module aliases (:161-166): mpd = mddom%msfd, pc = sfs%psb, pd = sfs%psdotb,
// ud = atms%ubd3d, vd = atms%vbd3d, wx = atms%wb3d.
#include <algorithm>
#include <cmath>

#include "model.hpp"

namespace orc {

static constexpr double z4_c1 = 1.0, z4_c2 = -4.0, z4_c3 = 12.0;  // :69-71
static constexpr double o4_c1 = 4.0 / 6.0, o4_c2 = 1.0 / 6.0, o4_c3 = -20.0 / 6.0;  // :63-65

void Oracle::calc_coeff() {
  const A3& ud = ubd3d; const A3& vd = vbd3d; const A3& wx = wb3d;
  xkc.fill(d_zero); xkd.fill(d_zero); xkcf.fill(d_zero);
  for (int k = 1; k <= kz; ++k)
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j) {
        double dudx = ud(j + 1, i, k) + ud(j + 1, i + 1, k) - ud(j, i, k) - ud(j, i + 1, k);
        double dvdx = vd(j + 1, i, k) + vd(j + 1, i + 1, k) - vd(j, i, k) - vd(j, i + 1, k);
        double dudy = ud(j, i + 1, k) + ud(j + 1, i + 1, k) - ud(j, i, k) - ud(j + 1, i, k);
        double dvdy = vd(j, i + 1, k) + vd(j + 1, i + 1, k) - vd(j, i, k) - vd(j + 1, i, k);
        double dwdz = wx(j, i, k) - wx(j, i, k + 1);
        double duv = std::sqrt(std::max((dudx - dvdy) * (dudx - dvdy) + (dvdx + dudy) * (dvdx + dudy) - dwdz * dwdz, d_zero));
        xkc(j, i, k) = std::min((hgfact(j, i) + dydc * duv), xkhmax);
      }
  if (xkc_raw.d.size() != xkc.d.size()) xkc_raw.alloc(jce1, jce2, ice1, ice2, 1, kz);
  xkc_raw.d = xkc.d;
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

void Oracle::diffu_d(A3& ut, A3& vt, const A3& u, const A3& v) {
  const A2& mpd = msfd;
  if (cfg.idiffu == 2) {   // :386-411
    for (int k = 1; k <= kz; ++k)
      for (int i = idi1; i <= idi2; ++i)
        for (int j = jdi1; j <= jdi2; ++j) {
          ut(j, i, k) = ut(j, i, k) + xkd(j, i, k) *
              (o4_c1 * (u(j + 1, i, k) / mpd(j + 1, i) + u(j - 1, i, k) / mpd(j - 1, i) +
                        u(j, i + 1, k) / mpd(j, i + 1) + u(j, i - 1, k) / mpd(j, i - 1)) +
               o4_c2 * (u(j + 1, i + 1, k) / mpd(j + 1, i + 1) + u(j - 1, i - 1, k) / mpd(j - 1, i - 1) +
                        u(j - 1, i + 1, k) / mpd(j - 1, i + 1) + u(j + 1, i - 1, k) / mpd(j + 1, i - 1)) +
               o4_c3 * (u(j, i, k) / mpd(j, i)));
          vt(j, i, k) = vt(j, i, k) + xkd(j, i, k) *
              (o4_c1 * (v(j + 1, i, k) / mpd(j + 1, i) + v(j - 1, i, k) / mpd(j - 1, i) +
                        v(j, i + 1, k) / mpd(j, i + 1) + v(j, i - 1, k) / mpd(j, i - 1)) +
               o4_c2 * (v(j + 1, i + 1, k) / mpd(j + 1, i + 1) + v(j - 1, i - 1, k) / mpd(j - 1, i - 1) +
                        v(j - 1, i + 1, k) / mpd(j - 1, i + 1) + v(j + 1, i - 1, k) / mpd(j + 1, i - 1)) +
               o4_c3 * (v(j, i, k) / mpd(j, i)));
        }
    return;
  }
  for (int k = 1; k <= kz; ++k)
    for (int i = idii1; i <= idii2; ++i)
      for (int j = jdii1; j <= jdii2; ++j) {
        ut(j, i, k) = ut(j, i, k) - xkd(j, i, k) *
            (z4_c1 * (u(j + 2, i, k) / mpd(j + 2, i) + u(j - 2, i, k) / mpd(j - 2, i) +
                      u(j, i + 2, k) / mpd(j, i + 2) + u(j, i - 2, k) / mpd(j, i - 2)) +
             z4_c2 * (u(j + 1, i, k) / mpd(j + 1, i) + u(j - 1, i, k) / mpd(j - 1, i) +
                      u(j, i + 1, k) / mpd(j, i + 1) + u(j, i - 1, k) / mpd(j, i - 1)) +
             z4_c3 * (u(j, i, k) / mpd(j, i)));
        vt(j, i, k) = vt(j, i, k) - xkd(j, i, k) *
            (z4_c1 * (v(j + 2, i, k) / mpd(j + 2, i) + v(j - 2, i, k) / mpd(j - 2, i) +
                      v(j, i + 2, k) / mpd(j, i + 2) + v(j, i - 2, k) / mpd(j, i - 2)) +
             z4_c2 * (v(j + 1, i, k) / mpd(j + 1, i) + v(j - 1, i, k) / mpd(j - 1, i) +
                      v(j, i + 1, k) / mpd(j, i + 1) + v(j, i - 1, k) / mpd(j, i - 1)) +
             z4_c3 * (v(j, i, k) / mpd(j, i)));
      }
  auto ring = [&](int j, int i, int k) {
    ut(j, i, k) = ut(j, i, k) + xkd(j, i, k) *
        (z4_c1 * (u(j + 1, i, k) / mpd(j + 1, i) + u(j - 1, i, k) / mpd(j - 1, i) +
                  u(j, i + 1, k) / mpd(j, i + 1) + u(j, i - 1, k) / mpd(j, i - 1)) +
         z4_c2 * (u(j, i, k) / mpd(j, i)));
    vt(j, i, k) = vt(j, i, k) + xkd(j, i, k) *
        (z4_c1 * (v(j + 1, i, k) / mpd(j + 1, i) + v(j - 1, i, k) / mpd(j - 1, i) +
                  v(j, i + 1, k) / mpd(j, i + 1) + v(j, i - 1, k) / mpd(j, i - 1)) +
         z4_c2 * (v(j, i, k) / mpd(j, i)));
  };
  if (has_bdyleft)
    for (int k = 1; k <= kz; ++k) for (int i = idi1; i <= idi2; ++i) ring(jdi1, i, k);
  if (has_bdyright)
    for (int k = 1; k <= kz; ++k) for (int i = idi1; i <= idi2; ++i) ring(jdi2, i, k);
  if (has_bdybottom)
    for (int k = 1; k <= kz; ++k) for (int j = jdi1; j <= jdi2; ++j) ring(j, idi1, k);
  if (has_bdytop)
    for (int k = 1; k <= kz; ++k) for (int j = jdi1; j <= jdi2; ++j) ring(j, idi2, k);
}

void Oracle::diffu_x3d(A3& ften, const A3& f) {
  if (cfg.idiffu == 2) {   // :726-735
    for (int k = 1; k <= kz; ++k)
      for (int i = ici1; i <= ici2; ++i)
        for (int j = jci1; j <= jci2; ++j)
          ften(j, i, k) = ften(j, i, k) + xkc(j, i, k) *
              (o4_c1 * (f(j + 1, i, k) + f(j - 1, i, k) + f(j, i + 1, k) + f(j, i - 1, k)) +
               o4_c2 * (f(j + 1, i + 1, k) + f(j - 1, i - 1, k) + f(j - 1, i + 1, k) + f(j + 1, i - 1, k)) +
               o4_c3 * (f(j, i, k)));
    return;
  }
  for (int k = 1; k <= kz; ++k)
    for (int i = icii1; i <= icii2; ++i)
      for (int j = jcii1; j <= jcii2; ++j)
        ften(j, i, k) = ften(j, i, k) - xkc(j, i, k) *
            (z4_c1 * (f(j + 2, i, k) + f(j - 2, i, k) + f(j, i + 2, k) + f(j, i - 2, k)) +
             z4_c2 * (f(j + 1, i, k) + f(j - 1, i, k) + f(j, i + 1, k) + f(j, i - 1, k)) +
             z4_c3 * (f(j, i, k)));
  auto ring = [&](int j, int i, int k) {
    ften(j, i, k) = ften(j, i, k) + xkc(j, i, k) *
        (z4_c1 * (f(j + 1, i, k) + f(j - 1, i, k) + f(j, i + 1, k) + f(j, i - 1, k)) +
         z4_c2 * (f(j, i, k)));
  };
  if (has_bdyleft)
    for (int k = 1; k <= kz; ++k) for (int i = ici1; i <= ici2; ++i) ring(jci1, i, k);
  if (has_bdyright)
    for (int k = 1; k <= kz; ++k) for (int i = ici1; i <= ici2; ++i) ring(jci2, i, k);
  if (has_bdybottom)
    for (int k = 1; k <= kz; ++k) for (int j = jci1; j <= jci2; ++j) ring(j, ici1, k);
  if (has_bdytop)
    for (int k = 1; k <= kz; ++k) for (int j = jci1; j <= jci2; ++j) ring(j, ici2, k);
}

void Oracle::diffu_x3df(A3& ften, const A3& f, double fac) {
  if (cfg.idiffu == 2) {   // :592-601
    for (int k = 1; k <= kzp1; ++k)
      for (int i = ici1; i <= ici2; ++i)
        for (int j = jci1; j <= jci2; ++j)
          ften(j, i, k) = ften(j, i, k) + fac * xkcf(j, i, k) *
              (o4_c1 * (f(j + 1, i, k) + f(j - 1, i, k) + f(j, i + 1, k) + f(j, i - 1, k)) +
               o4_c2 * (f(j + 1, i + 1, k) + f(j - 1, i - 1, k) + f(j - 1, i + 1, k) + f(j + 1, i - 1, k)) +
               o4_c3 * (f(j, i, k)));
    return;
  }
  for (int k = 1; k <= kzp1; ++k)
    for (int i = icii1; i <= icii2; ++i)
      for (int j = jcii1; j <= jcii2; ++j)
        ften(j, i, k) = ften(j, i, k) - fac * xkcf(j, i, k) *
            (z4_c1 * (f(j + 2, i, k) + f(j - 2, i, k) + f(j, i + 2, k) + f(j, i - 2, k)) +
             z4_c2 * (f(j + 1, i, k) + f(j - 1, i, k) + f(j, i + 1, k) + f(j, i - 1, k)) +
             z4_c3 * (f(j, i, k)));
  auto ring = [&](int j, int i, int k) {
    ften(j, i, k) = ften(j, i, k) + fac * xkcf(j, i, k) *
        (z4_c1 * (f(j + 1, i, k) + f(j - 1, i, k) + f(j, i + 1, k) + f(j, i - 1, k)) +
         z4_c2 * (f(j, i, k)));
  };
  if (has_bdyleft)
    for (int k = 1; k <= kzp1; ++k) for (int i = ici1; i <= ici2; ++i) ring(jci1, i, k);
  if (has_bdyright)
    for (int k = 1; k <= kzp1; ++k) for (int i = ici1; i <= ici2; ++i) ring(jci2, i, k);
  if (has_bdybottom)
    for (int k = 1; k <= kzp1; ++k) for (int j = jci1; j <= jci2; ++j) ring(j, ici1, k);
  if (has_bdytop)
    for (int k = 1; k <= kzp1; ++k) for (int j = jci1; j <= jci2; ++j) ring(j, ici2, k);
}

void Oracle::diffu_x4d(A4& ften, const A4& f, int n1, int n2, double fac) {
  for (int n = n1; n <= n2; ++n) {
    if (cfg.idiffu == 2) {   // :881-894
      for (int k = 1; k <= kz; ++k)
        for (int i = ici1; i <= ici2; ++i)
          for (int j = jci1; j <= jci2; ++j)
            ften(j, i, k, n) = ften(j, i, k, n) + fac * xkc(j, i, k) *
                (o4_c1 * (f(j + 1, i, k, n) + f(j - 1, i, k, n) + f(j, i + 1, k, n) + f(j, i - 1, k, n)) +
                 o4_c2 * (f(j + 1, i + 1, k, n) + f(j - 1, i - 1, k, n) + f(j - 1, i + 1, k, n) + f(j + 1, i - 1, k, n)) +
                 o4_c3 * f(j, i, k, n));
      continue;
    }
    for (int k = 1; k <= kz; ++k)
      for (int i = icii1; i <= icii2; ++i)
        for (int j = jcii1; j <= jcii2; ++j)
          ften(j, i, k, n) = ften(j, i, k, n) - fac * xkc(j, i, k) *
              (z4_c1 * (f(j + 2, i, k, n) + f(j - 2, i, k, n) + f(j, i + 2, k, n) + f(j, i - 2, k, n)) +
               z4_c2 * (f(j + 1, i, k, n) + f(j - 1, i, k, n) + f(j, i + 1, k, n) + f(j, i - 1, k, n)) +
               z4_c3 * f(j, i, k, n));
    auto ring = [&](int j, int i, int k) {
      ften(j, i, k, n) = ften(j, i, k, n) + fac * xkc(j, i, k) *
          (z4_c1 * (f(j + 1, i, k, n) + f(j - 1, i, k, n) + f(j, i + 1, k, n) + f(j, i - 1, k, n)) +
           z4_c2 * f(j, i, k, n));
    };
    if (has_bdyleft)
      for (int k = 1; k <= kz; ++k) for (int i = ici1; i <= ici2; ++i) ring(jci1, i, k);
    if (has_bdyright)
      for (int k = 1; k <= kz; ++k) for (int i = ici1; i <= ici2; ++i) ring(jci2, i, k);
    if (has_bdybottom)
      for (int k = 1; k <= kz; ++k) for (int j = jci1; j <= jci2; ++j) ring(j, ici1, k);
    if (has_bdytop)
      for (int k = 1; k <= kz; ++k) for (int j = jci1; j <= jci2; ++j) ring(j, ici2, k);
  }
}

}  // namespace orc
