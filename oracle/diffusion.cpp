// ORACLE — TEST INFRASTRUCTURE ONLY.
// Restates Main/mod_diffusion.F90 for idiffu = 1 (4th order, 2nd order on the boundary ring),
// idiffu = 2 (the nine-point "LeVeque" operator on every interior point) and idiffu = 3 (flux-limited 6th order; in this
// snapshot of the reference every one of its loops runs over the single column j = jdi2 / jci2, :421,611,745,902, with the
// stencil indices clamped to the domain — restated as written):
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
static constexpr double h4_c1 = 10.0, h4_c2 = -5.0, h4_c3 = 1.0;                   // :75-77

// the idiffu = 3 operator at (j, i): F = the diffused quantity, G = the quantity whose difference signs the flux
// (F itself for the winds, which are passed divided by the map factor; F / mpd for the scalars); jmax, imax = the upper
// clamp of the stencil indices (:414-426,604-616)
template <class FF, class GG>
static double diff6(FF F, GG G, int j, int i, int jmax, int imax) {
  const int im1 = std::max(i - 1, 1), im2 = std::max(i - 2, 1), im3 = std::max(i - 3, 1);
  const int ip1 = std::min(i + 1, imax), ip2 = std::min(i + 2, imax), ip3 = std::min(i + 3, imax);
  const int jm1 = std::max(j - 1, 1), jm2 = std::max(j - 2, 1), jm3 = std::max(j - 3, 1);
  const int jp1 = std::min(j + 1, jmax), jp2 = std::min(j + 2, jmax), jp3 = std::min(j + 3, jmax);
  double x0 = (h4_c1 * (F(j, i) - F(jm1, i)) + h4_c2 * (F(jp1, i) - F(jm2, i)) + h4_c3 * (F(jp2, i) - F(jm3, i)));
  if (x0 * (G(j, i) - G(jm1, i)) <= d_zero) x0 = d_zero;
  double x1 = (h4_c1 * (F(jp1, i) - F(j, i)) + h4_c2 * (F(jp2, i) - F(jm1, i)) + h4_c3 * (F(jp3, i) - F(jm2, i)));
  if (x1 * (G(jp1, i) - G(j, i)) <= d_zero) x1 = d_zero;
  double y0 = (h4_c1 * (F(j, i) - F(j, im1)) + h4_c2 * (F(j, ip1) - F(j, im2)) + h4_c3 * (F(j, ip2) - F(j, im3)));
  if (y0 * (G(j, i) - G(j, im1)) <= d_zero) y0 = d_zero;
  double y1 = (h4_c1 * (F(j, ip1) - F(j, i)) + h4_c2 * (F(j, ip2) - F(j, im1)) + h4_c3 * (F(j, ip3) - F(j, im2)));
  if (y1 * (G(j, ip1) - G(j, i)) <= d_zero) y1 = d_zero;
  return ((x1 - x0) + (y1 - y0));
}

// :174-183
void Oracle::calc_coeff6() {
  const double coef = 0.12 * 0.015625 / (2.0 * dtsec);   // diff_6th_factor (:78), :154
  if (xkc_raw.d.size() != xkc.d.size()) { xkc_raw.alloc(jce1, jce2, ice1, ice2, 1, kz); }
  for (int k = 1; k <= kz; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) xkc(j, i, k) = coef * psb(j, i);
  for (int k = 1; k <= kzp1; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) xkcf(j, i, k) = coef * psb(j, i);
  for (int k = 1; k <= kz; ++k)
    for (int i = idi1; i <= idi2; ++i)
      for (int j = jdi1; j <= jdi2; ++j) xkd(j, i, k) = coef * psdotb(j, i);
}

void Oracle::calc_coeff() {
  if (cfg.idiffu == 3) { calc_coeff6(); return; }
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
  if (cfg.idiffu == 3) {   // :412-517: the loop over j is `do j = jdi2 , jdi2`
    const int j = jdi2;
    for (int k = 1; k <= kz; ++k)
      for (int i = idi1; i <= idi2; ++i) {
        auto U = [&](int jj, int ii) { return u(jj, ii, k) / mpd(jj, ii); };
        auto V = [&](int jj, int ii) { return v(jj, ii, k) / mpd(jj, ii); };
        ut(j, i, k) = ut(j, i, k) + xkd(j, i, k) * diff6(U, U, j, i, jx, iy);
        vt(j, i, k) = vt(j, i, k) + xkd(j, i, k) * diff6(V, V, j, i, jx, iy);
      }
    return;
  }
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
  if (cfg.idiffu == 3) {   // :736-790, j = jci2 only; the flux sign test divides by mpd = msfd at the cross indices
    const int j = jci2;
    for (int k = 1; k <= kz; ++k)
      for (int i = ici1; i <= ici2; ++i) {
        auto F = [&](int jj, int ii) { return f(jj, ii, k); };
        auto G = [&](int jj, int ii) { return f(jj, ii, k) / msfd(jj, ii); };
        ften(j, i, k) = ften(j, i, k) + xkc(j, i, k) * diff6(F, G, j, i, jx - 1, iy - 1);
      }
    return;
  }
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
  if (cfg.idiffu == 3) {   // :602-652, j = jci2 only.  The reference multiplies by xkc(j,i,k) for k up to kzp1, one level beyond
    const int j = jci2;    // the array (:646); xkc does not depend on k (:176), so level kzp1 takes the value of the other levels
    for (int k = 1; k <= kzp1; ++k)
      for (int i = ici1; i <= ici2; ++i) {
        auto F = [&](int jj, int ii) { return f(jj, ii, k); };
        auto G = [&](int jj, int ii) { return f(jj, ii, k) / msfd(jj, ii); };
        ften(j, i, k) = ften(j, i, k) + fac * xkc(j, i, std::min(k, kz)) * diff6(F, G, j, i, jx - 1, iy - 1);
      }
    return;
  }
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
    if (cfg.idiffu == 3) {   // :895-960, j = jci2 only
      const int j = jci2;
      for (int k = 1; k <= kz; ++k)
        for (int i = ici1; i <= ici2; ++i) {
          auto F = [&](int jj, int ii) { return f(jj, ii, k, n); };
          auto G = [&](int jj, int ii) { return f(jj, ii, k, n) / msfd(jj, ii); };
          ften(j, i, k, n) = ften(j, i, k, n) + fac * xkc(j, i, k) * diff6(F, G, j, i, jx - 1, iy - 1);
        }
      continue;
    }
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
