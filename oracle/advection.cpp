// ORACLE — TEST INFRASTRUCTURE ONLY.
// Restates Main/mod_advection.F90: start_advect :111-120, hadvuv :124-269
// (idynamic = 2 branches), vadvuv :271-303, hadvt :311-393, hadv3d :395-515,
// hadvqv :517-603, hadvqx :607-662, vadv3d :730-809 (ind = 0), vadvqv
// :811-836, vadv4d :847-966 (ind = 0,1).  Module aliases: ua = atmx%umc,
// va = atmx%vmc, ps = sfs%psa, xmapf = mddom%xmsf, dmapf = mddom%dmsf,
// svv = qdot, divx = mdv%cr (:82-92).
#include <algorithm>
#include <cmath>

#include "model.hpp"

namespace orc {

void Oracle::start_advect() {
  const A3& ua = ax_umc; const A3& va = ax_vmc;
  for (int k = 1; k <= kz; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) {
        uavg1(j, i, k) = ua(j, i + 1, k) + ua(j, i, k);
        uavg2(j, i, k) = ua(j + 1, i + 1, k) + ua(j + 1, i, k);
        vavg1(j, i, k) = va(j + 1, i, k) + va(j, i, k);
        vavg2(j, i, k) = va(j + 1, i + 1, k) + va(j, i + 1, k);
      }
}

void Oracle::hadvuv(A3& ut, A3& vt, const A3& u, const A3& v) {
  const A3& ua = ax_umc; const A3& va = ax_vmc; const A3& divx = cr;
  const bool up = cfg.upstream_mode != 0;
  for (int k = 1; k <= kz; ++k)
    for (int i = idi1; i <= idi2; ++i)
      for (int j = jdi1; j <= jdi2; ++j) {
        double divd = d_rfour * (divx(j, i, k) + divx(j, i - 1, k) + divx(j - 1, i, k) + divx(j - 1, i - 1, k));
        double ucmona = ua(j, i + 1, k) + d_two * ua(j, i, k) + ua(j, i - 1, k);
        double ucmonb = ua(j + 1, i + 1, k) + d_two * ua(j + 1, i, k) + ua(j + 1, i - 1, k);
        double ucmonc = ua(j - 1, i + 1, k) + d_two * ua(j - 1, i, k) + ua(j - 1, i - 1, k);
        double vcmona = va(j + 1, i, k) + d_two * va(j, i, k) + va(j - 1, i, k);
        double vcmonb = va(j + 1, i + 1, k) + d_two * va(j, i + 1, k) + va(j - 1, i + 1, k);
        double vcmonc = va(j + 1, i - 1, k) + d_two * va(j, i - 1, k) + va(j - 1, i - 1, k);
        double diag = divd - dmsf(j, i) * ((ucmonb - ucmonc) + (vcmonb - vcmonc));
        if (up) {
          double ff1 = ul * (u(j + 1, i, k) + u(j, i, k));
          double ff2 = ul * (u(j - 1, i, k) + u(j, i, k));
          double ff3 = ul * (v(j, i + 1, k) + v(j, i, k));
          double ff4 = ul * (v(j, i - 1, k) + v(j, i, k));
          ucmonb = (d_one + ff1) * ucmona + (d_one - ff1) * ucmonb;
          ucmonc = (d_one + ff2) * ucmonc + (d_one - ff2) * ucmona;
          vcmonb = (d_one + ff3) * vcmona + (d_one - ff3) * vcmonb;
          vcmonc = (d_one + ff4) * vcmonc + (d_one - ff4) * vcmona;
        } else {
          ucmonb = ucmona + ucmonb;
          ucmonc = ucmonc + ucmona;
          vcmonb = vcmona + vcmonb;
          vcmonc = vcmonc + vcmona;
        }
        ut(j, i, k) = ut(j, i, k) + u(j, i, k) * diag -
                      dmsf(j, i) * (u(j + 1, i, k) * ucmonb - u(j - 1, i, k) * ucmonc +
                                    u(j, i + 1, k) * vcmonb - u(j, i - 1, k) * vcmonc);
        vt(j, i, k) = vt(j, i, k) + v(j, i, k) * diag -
                      dmsf(j, i) * (v(j + 1, i, k) * ucmonb - v(j - 1, i, k) * ucmonc +
                                    v(j, i + 1, k) * vcmonb - v(j, i - 1, k) * vcmonc);
      }
}

void Oracle::vadvuv(A3& ut, A3& vt, const A3& ud, const A3& vd) {
  const A3& svv = qdot;
  for (int k = 2; k <= kz; ++k)
    for (int i = idi1; i <= idi2; ++i)
      for (int j = jdi1; j <= jdi2; ++j) {
        double qq = d_rfour * (svv(j, i, k) + svv(j, i - 1, k) + svv(j - 1, i, k) + svv(j - 1, i - 1, k));
        double uu = qq * (twt(k, 1) * ud(j, i, k) + twt(k, 2) * ud(j, i, k - 1));
        double vv = qq * (twt(k, 1) * vd(j, i, k) + twt(k, 2) * vd(j, i, k - 1));
        ut(j, i, k - 1) = ut(j, i, k - 1) - uu * xds(k - 1);
        ut(j, i, k) = ut(j, i, k) + uu * xds(k);
        vt(j, i, k - 1) = vt(j, i, k - 1) - vv * xds(k - 1);
        vt(j, i, k) = vt(j, i, k) + vv * xds(k);
      }
}

// common flux-form kernel of hadvt/hadv3d(ind=0)/hadvqv/hadvqx
#define ORC_HFLUX(F)                                                              \
  (up ? [&]() {                                                                   \
    double f1 = d_half * ul * (uavg2(j, i, k) + uavg1(j, i, k)) / psa(j, i);      \
    double f2 = d_half * ul * (vavg2(j, i, k) + vavg1(j, i, k)) / psa(j, i);      \
    double fx1 = (d_one + f1) * F(j - 1, i) + (d_one - f1) * F(j, i);             \
    double fx2 = (d_one + f1) * F(j, i) + (d_one - f1) * F(j + 1, i);             \
    double fy1 = (d_one + f2) * F(j, i - 1) + (d_one - f2) * F(j, i);             \
    double fy2 = (d_one + f2) * F(j, i) + (d_one - f2) * F(j, i + 1);             \
    return -xmsf(j, i) * (uavg2(j, i, k) * fx2 - uavg1(j, i, k) * fx1 +           \
                          vavg2(j, i, k) * fy2 - vavg1(j, i, k) * fy1);           \
  }() : [&]() {                                                                   \
    double fx1 = F(j - 1, i) + F(j, i);                                           \
    double fx2 = F(j, i) + F(j + 1, i);                                           \
    double fy1 = F(j, i - 1) + F(j, i);                                           \
    double fy2 = F(j, i) + F(j, i + 1);                                           \
    return -xmsf(j, i) * (uavg2(j, i, k) * fx2 - uavg1(j, i, k) * fx1 +           \
                          vavg2(j, i, k) * fy2 - vavg1(j, i, k) * fy1);           \
  }())

void Oracle::hadvt(A3& ften, const A3& f) {
  const bool up = cfg.upstream_mode != 0;
  const double t_extrema = cfg.t_extrema;
  for (int k = 1; k <= kz; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) {
        auto F = [&](int jj, int ii) { return f(jj, ii, k); };
        fg(j, i, k) = ORC_HFLUX(F);
      }
  if (cfg.stability_enhance) {
    for (int k = 1; k <= kz; ++k)
      for (int i = ici1; i <= ici2; ++i)
        for (int j = jci1; j <= jci2; ++j) {
          if (std::fabs(f(j, i + 1, k) + f(j, i - 1, k) - d_two * f(j, i, k)) / psa(j, i) > t_extrema) {
            if ((f(j, i, k) > f(j, i + 1, k)) && (f(j, i, k) > f(j, i - 1, k)))
              fg(j, i, k) = std::min(fg(j, i, k), d_zero);
            else if ((f(j, i, k) < f(j, i + 1, k)) && (f(j, i, k) < f(j, i - 1, k)))
              fg(j, i, k) = std::max(fg(j, i, k), d_zero);
          }
          if (std::fabs(f(j + 1, i, k) + f(j - 1, i, k) - d_two * f(j, i, k)) / psa(j, i) > t_extrema) {
            if ((f(j, i, k) > f(j + 1, i, k)) && (f(j, i, k) > f(j - 1, i, k)))
              fg(j, i, k) = std::min(fg(j, i, k), d_zero);
            else if ((f(j, i, k) < f(j + 1, i, k)) && (f(j, i, k) < f(j - 1, i, k)))
              fg(j, i, k) = std::max(fg(j, i, k), d_zero);
          }
        }
  }
  for (int k = 1; k <= kz; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) ften(j, i, k) = ften(j, i, k) + fg(j, i, k);
}

void Oracle::hadv3d(A3& ften, const A3& f, int ind) {
  const bool up = cfg.upstream_mode != 0;
  if (ind == 0) {
    for (int k = 1; k <= kz; ++k)
      for (int i = ici1; i <= ici2; ++i)
        for (int j = jci1; j <= jci2; ++j) {
          auto F = [&](int jj, int ii) { return f(jj, ii, k); };
          // ften = ften - xmapf*(...): a + (-x) is bitwise a - x
          ften(j, i, k) = ften(j, i, k) + ORC_HFLUX(F);
        }
  } else {
    for (int k = 2; k <= kz; ++k)
      for (int i = ici1; i <= ici2; ++i)
        for (int j = jci1; j <= jci2; ++j) {
          double uaz1 = (twt(k, 1) * uavg1(j, i, k) + twt(k, 2) * uavg1(j, i, k - 1));
          double uaz2 = (twt(k, 1) * uavg2(j, i, k) + twt(k, 2) * uavg2(j, i, k - 1));
          double vaz1 = (twt(k, 1) * vavg1(j, i, k) + twt(k, 2) * vavg1(j, i, k - 1));
          double vaz2 = (twt(k, 1) * vavg2(j, i, k) + twt(k, 2) * vavg2(j, i, k - 1));
          double fx1, fx2, fy1, fy2;
          if (up) {
            double f1 = d_half * ul * (uavg2(j, i, k) + uavg1(j, i, k)) / psa(j, i);
            double f2 = d_half * ul * (vavg2(j, i, k) + vavg1(j, i, k)) / psa(j, i);
            fx1 = (d_one + f1) * f(j - 1, i, k) + (d_one - f1) * f(j, i, k);
            fx2 = (d_one + f1) * f(j, i, k) + (d_one - f1) * f(j + 1, i, k);
            fy1 = (d_one + f2) * f(j, i - 1, k) + (d_one - f2) * f(j, i, k);
            fy2 = (d_one + f2) * f(j, i, k) + (d_one - f2) * f(j, i + 1, k);
          } else {
            fx1 = f(j - 1, i, k) + f(j, i, k);
            fx2 = f(j, i, k) + f(j + 1, i, k);
            fy1 = f(j, i - 1, k) + f(j, i, k);
            fy2 = f(j, i, k) + f(j, i + 1, k);
          }
          ften(j, i, k) = ften(j, i, k) - xmsf(j, i) * (uaz2 * fx2 - uaz1 * fx1 + vaz2 * fy2 - vaz1 * fy1);
        }
  }
}

void Oracle::hadvqv(A4& ften, const A4& f, int iv) {
  const bool up = cfg.upstream_mode != 0;
  const double q_rel_extrema = cfg.q_rel_extrema;
  for (int k = 1; k <= kz; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) {
        auto F = [&](int jj, int ii) { return f(jj, ii, k, iv); };
        fg(j, i, k) = ORC_HFLUX(F);
      }
  if (cfg.stability_enhance) {
    for (int k = 1; k <= kz; ++k)
      for (int i = ici1; i <= ici2; ++i)
        for (int j = jci1; j <= jci2; ++j) {
          const double f0 = f(j, i, k, iv);
          if (std::fabs(f(j, i + 1, k, iv) + f(j, i - 1, k, iv) - d_two * f0) / std::max(f0, dlowval) > q_rel_extrema) {
            if ((f0 > f(j, i + 1, k, iv)) && (f0 > f(j, i - 1, k, iv)))
              fg(j, i, k) = std::min(fg(j, i, k), d_zero);
            else if ((f0 < f(j, i + 1, k, iv)) && (f0 < f(j, i - 1, k, iv)))
              fg(j, i, k) = std::max(fg(j, i, k), d_zero);
          }
          if (std::fabs(f(j + 1, i, k, iv) + f(j - 1, i, k, iv) - d_two * f0) / std::max(f0, dlowval) > q_rel_extrema) {
            if ((f0 > f(j + 1, i, k, iv)) && (f0 > f(j - 1, i, k, iv)))
              fg(j, i, k) = std::min(fg(j, i, k), d_zero);
            else if ((f0 < f(j + 1, i, k, iv)) && (f0 < f(j - 1, i, k, iv)))
              fg(j, i, k) = std::max(fg(j, i, k), d_zero);
          }
        }
  }
  for (int k = 1; k <= kz; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) ften(j, i, k, iv) = ften(j, i, k, iv) + fg(j, i, k);
}

void Oracle::hadvqx(A4& ften, const A4& f, int n1, int n2) {
  const bool up = cfg.upstream_mode != 0;
  for (int n = n1; n <= n2; ++n) {
    for (int k = 1; k <= kz; ++k)
      for (int i = ici1; i <= ici2; ++i)
        for (int j = jci1; j <= jci2; ++j) {
          auto F = [&](int jj, int ii) { return f(jj, ii, k, n); };
          fg(j, i, k) = ORC_HFLUX(F);
        }
    for (int k = 1; k <= kz; ++k)
      for (int i = ici1; i <= ici2; ++i)
        for (int j = jci1; j <= jci2; ++j) ften(j, i, k, n) = ften(j, i, k, n) + fg(j, i, k);
  }
}

void Oracle::vadv3d(A3& ften, const A3& f, int nk, int ind) {
  const A3& svv = qdot;
  if (ind != 0) std::abort();  // ind = 1 (hydrostatic T) not on the NH path
  if (nk == kz) {
    for (int k = 2; k <= kz; ++k)
      for (int i = ici1; i <= ici2; ++i)
        for (int j = jci1; j <= jci2; ++j) {
          double fx = svv(j, i, k) * (twt(k, 1) * f(j, i, k) + twt(k, 2) * f(j, i, k - 1));
          ften(j, i, k - 1) = ften(j, i, k - 1) - fx * xds(k - 1);
          ften(j, i, k) = ften(j, i, k) + fx * xds(k);
        }
  } else {
    for (int k = 1; k <= kz; ++k)
      for (int i = ici1; i <= ici2; ++i)
        for (int j = jci1; j <= jci2; ++j) {
          double qq = d_half * (svv(j, i, k) + svv(j, i, k + 1));
          double fx = qq * ((f(j, i, k) + f(j, i, k + 1)));
          ften(j, i, k + 1) = ften(j, i, k + 1) + fx * dds(k + 1);
          ften(j, i, k) = ften(j, i, k) - fx * dds(k);
        }
  }
}

void Oracle::vadvqv(A4& ften, const A4& f, int n) {
  const A3& svv = qdot;
  fg.fill(d_zero);
  for (int k = 2; k <= kz; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j)
        if (f(j, i, k, n) > minqq * psa(j, i) && f(j, i, k - 1, n) > minqq * psa(j, i))
          fg(j, i, k) = f(j, i, k, n) * std::pow(f(j, i, k - 1, n) / f(j, i, k, n), qcon(k));
  for (int k = 2; k <= kz; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) {
        ften(j, i, k - 1, n) = ften(j, i, k - 1, n) - svv(j, i, k) * fg(j, i, k) * xds(k - 1);
        ften(j, i, k, n) = ften(j, i, k, n) + svv(j, i, k) * fg(j, i, k) * xds(k);
      }
}

void Oracle::vadv4d(A4& ften, const A4& f, int n1, int n2, int ind) {
  const A3& svv = qdot;
  for (int n = n1; n <= n2; ++n) {
    fg.fill(d_zero);
    if (ind == 0) {
      for (int k = 2; k <= kz; ++k)
        for (int i = ici1; i <= ici2; ++i)
          for (int j = jci1; j <= jci2; ++j)
            fg(j, i, k) = (svv(j, i, k) > d_zero) ? svv(j, i, k) * f(j, i, k - 1, n) : svv(j, i, k) * f(j, i, k, n);
    } else if (ind == 1) {
      for (int k = 2; k <= kz; ++k)
        for (int i = ici1; i <= ici2; ++i)
          for (int j = jci1; j <= jci2; ++j) {
            if (svv(j, i, k) > d_zero) {
              if (f(j, i, k - 1, n) > minqq * minqq * psa(j, i))
                fg(j, i, k) = svv(j, i, k) * (twt(k, 1) * f(j, i, k, n) + twt(k, 2) * f(j, i, k - 1, n));
              else
                fg(j, i, k) = d_zero;
            } else {
              if (f(j, i, k, n) > minqq * minqq * psa(j, i))
                fg(j, i, k) = svv(j, i, k) * (twt(k, 1) * f(j, i, k, n) + twt(k, 2) * f(j, i, k - 1, n));
              else
                fg(j, i, k) = d_zero;
            }
          }
    } else if (ind == 3) {   // :915-957 (oracle/tke.cpp)
      vadv_ind3([&](int j, int i, int k) { return f(j, i, k, n); });
    } else {
      std::abort();
    }
    for (int k = 2; k <= kz; ++k)
      for (int i = ici1; i <= ici2; ++i)
        for (int j = jci1; j <= jci2; ++j) {
          ften(j, i, k - 1, n) = ften(j, i, k - 1, n) - fg(j, i, k) * xds(k - 1);
          ften(j, i, k, n) = ften(j, i, k, n) + fg(j, i, k) * xds(k);
        }
  }
}

}  // namespace orc
