// ORACLE — TEST INFRASTRUCTURE ONLY.
// Restates the dynamics part of Main/mod_bdycod.F90 for idynamic = 2:
// bdyin (state part) :669-690,772-831, bdyuv :896-1094, bdyval :1109-2571
// (non-chem, non-tke, iboudy 0/1/5 branches), nudge4d3d :3206-3392,
// nudgeuv :3581-3823, nudge3d :4218-4406, raydamp* :4953-5085, tau :5115-5123.
#include <algorithm>
#include <cmath>

#include "model.hpp"

namespace orc {

double Oracle::tau(double z, double zmax) const {
  if (z > zmax - cfg.rayhd) {
    double s = std::sin(halfpi * (d_one - (zmax - z) / cfg.rayhd));
    return cfg.rayalpha0 * (s * s);
  }
  return d_zero;
}

// the four zone loops (south, north, west, east) share one body; they are
// executed in the reference's order so that a point flagged in two zones
// (does not happen for nspg <= min(jx,iy)/2) would be treated identically.
#define ORC_ZONES(BA, ...)                                          \
  do {                                                              \
    const I2* zs[4] = {&BA.bsouth, &BA.bnorth, &BA.bwest, &BA.beast}; \
    const int zn[4] = {BA.ns, BA.nn, BA.nw, BA.ne};                 \
    for (int z = 0; z < 4; ++z) {                                   \
      if (zn[z] == 0) continue;                                     \
      const I2& zone = *zs[z];                                      \
      __VA_ARGS__                                                    \
    }                                                               \
  } while (0)

void Oracle::nudge3d(int ibdy, const A3& f, const V3dBound& bnd, A3& ften) {
  if (!ba_cr.havebound) return;
  const int ns = f.l3, nk = f.u3();
  const double xt = xbctime + dt;
  for (int k = ns; k <= nk; ++k)
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j)
        fg1(j, i, k) = (bnd.b0(j, i, k) + xt * bnd.bt(j, i, k)) - f(j, i, k);
  ORC_ZONES(ba_cr, {
    for (int k = ns; k <= nk; ++k)
      for (int i = ici1; i <= ici2; ++i)
        for (int j = jci1; j <= jci2; ++j) {
          if (!zone(j, i)) continue;
          int ib = ba_cr.ibnd(j, i);
          double xf = (ibdy == 1) ? fcx(ib) : hefc(ib, std::min(k, kz));
          double xg = (ibdy == 1) ? gcx(ib) : hegc(ib, std::min(k, kz));
          double fls0 = fg1(j, i, k), fls1 = fg1(j - 1, i, k), fls2 = fg1(j + 1, i, k),
                 fls3 = fg1(j, i - 1, k), fls4 = fg1(j, i + 1, k);
          ften(j, i, k) = ften(j, i, k) + xf * fls0 - xg * (fls1 + fls2 + fls3 + fls4 - d_four * fls0);
        }
  });
}

void Oracle::nudge4d3d(int ibdy, const A4& f, const V3dBound& bnd, A4& ften, int n) {
  if (!ba_cr.havebound) return;
  const double nfac = 1.0e3, rfac = d_one / nfac;
  const double xt = xbctime + dt;
  for (int k = 1; k <= kz; ++k)
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j)
        fg1(j, i, k) = nfac * (bnd.b0(j, i, k) + xt * bnd.bt(j, i, k)) - nfac * f(j, i, k, n);
  ORC_ZONES(ba_cr, {
    for (int k = 1; k <= kz; ++k)
      for (int i = ici1; i <= ici2; ++i)
        for (int j = jci1; j <= jci2; ++j) {
          if (!zone(j, i)) continue;
          int ib = ba_cr.ibnd(j, i);
          double xf = (ibdy == 1) ? fcx(ib) : hefc(ib, k);
          double xg = (ibdy == 1) ? gcx(ib) : hegc(ib, k);
          double fls0 = fg1(j, i, k), fls1 = fg1(j - 1, i, k), fls2 = fg1(j + 1, i, k),
                 fls3 = fg1(j, i - 1, k), fls4 = fg1(j, i + 1, k);
          ften(j, i, k, n) = ften(j, i, k, n) + rfac * (xf * fls0 - xg * (fls1 + fls2 + fls3 + fls4 - d_four * fls0));
        }
  });
}

void Oracle::nudgeuv(int ibdy, const A3& fu, const A3& fv, const V3dBound& bu, const V3dBound& bv, A3& ftu, A3& ftv) {
  if (!ba_dt.havebound) return;
  const double xt = xbctime + dt;
  for (int k = 1; k <= kz; ++k)
    for (int i = ide1; i <= ide2; ++i)
      for (int j = jde1; j <= jde2; ++j) {
        fg1(j, i, k) = ((bu.b0(j, i, k) + xt * bu.bt(j, i, k)) - fu(j, i, k));
        fg2(j, i, k) = ((bv.b0(j, i, k) + xt * bv.bt(j, i, k)) - fv(j, i, k));
      }
  ORC_ZONES(ba_dt, {
    for (int k = 1; k <= kz; ++k)
      for (int i = idi1; i <= idi2; ++i)
        for (int j = jdi1; j <= jdi2; ++j) {
          if (!zone(j, i)) continue;
          int ib = ba_dt.ibnd(j, i);
          double xf = (ibdy == 1) ? fcd(ib) : hefd(ib, k);
          double xg = (ibdy == 1) ? gcd(ib) : hegd(ib, k);
          double fls0 = fg1(j, i, k), fls1 = fg1(j - 1, i, k), fls2 = fg1(j + 1, i, k),
                 fls3 = fg1(j, i - 1, k), fls4 = fg1(j, i + 1, k);
          ftu(j, i, k) = ftu(j, i, k) + xf * fls0 - xg * (fls1 + fls2 + fls3 + fls4 - d_four * fls0);
          fls0 = fg2(j, i, k); fls1 = fg2(j - 1, i, k); fls2 = fg2(j + 1, i, k);
          fls3 = fg2(j, i - 1, k); fls4 = fg2(j, i + 1, k);
          ftv(j, i, k) = ftv(j, i, k) + xf * fls0 - xg * (fls1 + fls2 + fls3 + fls4 - d_four * fls0);
        }
  });
}

void Oracle::raydampuv(const A3& z, const A3& u, const A3& v, A3& ut, A3& vt, const V3dBound& ub, const V3dBound& vb) {
  const double xt = xbctime + dt;
  const int kt = std::min(kz, cfg.rayndamp);
  for (int k = 1; k <= kt; ++k)
    for (int i = idi1; i <= idi2; ++i)
      for (int j = jdi1; j <= jdi2; ++j) {
        double bval = ub.b0(j, i, k) + xt * ub.bt(j, i, k);
        double zz = d_rfour * (z(j, i, k) + z(j - 1, i, k) + z(j, i - 1, k) + z(j - 1, i - 1, k));
        double zm = d_rfour * (z(j, i, 1) + z(j - 1, i, 1) + z(j, i - 1, 1) + z(j - 1, i - 1, 1));
        ut(j, i, k) = ut(j, i, k) + tau(zz, zm) * (bval - u(j, i, k));
      }
  for (int k = 1; k <= kt; ++k)
    for (int i = idi1; i <= idi2; ++i)
      for (int j = jdi1; j <= jdi2; ++j) {
        double bval = vb.b0(j, i, k) + xt * vb.bt(j, i, k);
        double zz = d_rfour * (z(j, i, k) + z(j - 1, i, k) + z(j, i - 1, k) + z(j - 1, i - 1, k));
        double zm = d_rfour * (z(j, i, 1) + z(j - 1, i, 1) + z(j, i - 1, 1) + z(j - 1, i - 1, 1));
        vt(j, i, k) = vt(j, i, k) + tau(zz, zm) * (bval - v(j, i, k));
      }
}

void Oracle::raydamp3f(const A3& z, const A3& var, A3& vten, double sval) {
  const int kt = std::min(kzp1, cfg.rayndamp);
  for (int k = 1; k <= kt; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j)
        vten(j, i, k) = vten(j, i, k) + tau(z(j, i, k), z(j, i, 1)) * (sval - var(j, i, k));
}

void Oracle::raydamp3(const A3& z, const A3& var, A3& vten, const V3dBound& bnd) {
  const double xt = xbctime + dt;
  const int kt = std::min(kz, cfg.rayndamp);
  for (int k = 1; k <= kt; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) {
        double bval = bnd.b0(j, i, k) + xt * bnd.bt(j, i, k);
        vten(j, i, k) = vten(j, i, k) + tau(z(j, i, k), z(j, i, 1)) * (bval - var(j, i, k));
      }
}

void Oracle::raydampqv(const A3& z, const A4& var, A4& vten, const V3dBound& bnd) {
  const double xt = xbctime + dt;
  const int kt = std::min(kz, cfg.rayndamp);
  for (int k = 1; k <= kt; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) {
        double bval = bnd.b0(j, i, k) + xt * bnd.bt(j, i, k);
        vten(j, i, k, iqv) = vten(j, i, k, iqv) + tau(z(j, i, k), z(j, i, 1)) * (bval - var(j, i, k, iqv));
      }
}

// state part of bdyin; the staged record (decoupled) plays the role of the
// file read into x?b%b1
void Oracle::bdyin() {
  xbctime = d_zero;
  xub.b0.d = xub.b1.d; xvb.b0.d = xvb.b1.d; xtb.b0.d = xtb.b1.d; xqb.b0.d = xqb.b1.d;
  xppb.b0.d = xppb.b1.d; xwwb.b0.d = xwwb.b1.d;
  xub.b1.d = nx_u.d; xvb.b1.d = nx_v.d; xtb.b1.d = nx_t.d; xqb.b1.d = nx_q.d;
  xppb.b1.d = nx_pp.d; xwwb.b1.d = nx_ww.d;
  have_next = false;
  couple(xub.b1, psdot, jde1, jde2, ide1, ide2, 1, kz);
  couple(xvb.b1, psdot, jde1, jde2, ide1, ide2, 1, kz);
  couple(xtb.b1, xpsb_b0, jce1, jce2, ice1, ice2, 1, kz);   // xpsb%b1 == xpsb%b0 (:398-400)
  couple(xqb.b1, xpsb_b0, jce1, jce2, ice1, ice2, 1, kz);
  couple(xppb.b1, xpsb_b0, jce1, jce2, ice1, ice2, 1, kz);
  couple(xwwb.b1, xpsb_b0, jce1, jce2, ice1, ice2, 1, kzp1);
  timeint(xub.b1, xub.b0, xub.bt, jde1, jde2, ide1, ide2, 1, kz);
  timeint(xvb.b1, xvb.b0, xvb.bt, jde1, jde2, ide1, ide2, 1, kz);
  timeint(xtb.b1, xtb.b0, xtb.bt, jce1, jce2, ice1, ice2, 1, kz);
  timeint(xqb.b1, xqb.b0, xqb.bt, jce1, jce2, ice1, ice2, 1, kz);
  timeint(xppb.b1, xppb.b0, xppb.bt, jce1, jce2, ice1, ice2, 1, kz);
  timeint(xwwb.b1, xwwb.b0, xwwb.bt, jce1, jce2, ice1, ice2, 1, kzp1);
}

void Oracle::bdyuv(double xt) {
  for (int k = 1; k <= kz; ++k)
    for (int i = idi1; i <= idi2; ++i) {
      wui(i, k) = a1_u(jdi1, i, k); wvi(i, k) = a1_v(jdi1, i, k);
      eui(i, k) = a1_u(jdi2, i, k); evi(i, k) = a1_v(jdi2, i, k);
    }
  for (int k = 1; k <= kz; ++k)
    for (int j = jdi1; j <= jdi2; ++j) {
      sui(j, k) = a1_u(j, idi1, k); svi(j, k) = a1_v(j, idi1, k);
      nui(j, k) = a1_u(j, idi2, k); nvi(j, k) = a1_v(j, idi2, k);
    }
  if (cfg.iboudy == 0) {
    for (int k = 1; k <= kz; ++k)
      for (int i = ide1; i <= ide2; ++i) {
        wue(i, k) = xub.b0(jde1, i, k); wve(i, k) = xvb.b0(jde1, i, k);
        eue(i, k) = xub.b0(jde2, i, k); eve(i, k) = xvb.b0(jde2, i, k);
      }
    for (int k = 1; k <= kz; ++k)
      for (int j = jde1; j <= jde2; ++j) {
        sue(j, k) = xub.b0(j, ide1, k); sve(j, k) = xvb.b0(j, ide1, k);
        nue(j, k) = xub.b0(j, ide2, k); nve(j, k) = xvb.b0(j, ide2, k);
      }
  } else {
    for (int k = 1; k <= kz; ++k)
      for (int i = idi1; i <= idi2; ++i) {
        wue(i, k) = (xub.b0(jde1, i, k) + xt * xub.bt(jde1, i, k));
        wve(i, k) = (xvb.b0(jde1, i, k) + xt * xvb.bt(jde1, i, k));
        eue(i, k) = (xub.b0(jde2, i, k) + xt * xub.bt(jde2, i, k));
        eve(i, k) = (xvb.b0(jde2, i, k) + xt * xvb.bt(jde2, i, k));
      }
    for (int k = 1; k <= kz; ++k)
      for (int j = jde1; j <= jde2; ++j) {
        sue(j, k) = (xub.b0(j, ide1, k) + xt * xub.bt(j, ide1, k));
        sve(j, k) = (xvb.b0(j, ide1, k) + xt * xvb.bt(j, ide1, k));
        nue(j, k) = (xub.b0(j, ide2, k) + xt * xub.bt(j, ide2, k));
        nve(j, k) = (xvb.b0(j, ide2, k) + xt * xvb.bt(j, ide2, k));
      }
  }
  for (int k = 1; k <= kz; ++k) {
    // topleft, bottomleft, topright, bottomright (:1030-1061)
    wui(ide2, k) = nue(jdi1, k); wvi(ide2, k) = nve(jdi1, k);
    nui(jde1, k) = wue(idi2, k); nvi(jde1, k) = wve(idi2, k);
    wui(ide1, k) = sue(jdi1, k); wvi(ide1, k) = sve(jdi1, k);
    sui(jde1, k) = wue(idi1, k); svi(jde1, k) = wve(idi1, k);
    eui(ide2, k) = nue(jdi2, k); evi(ide2, k) = nve(jdi2, k);
    nui(jde2, k) = eue(idi2, k); nvi(jde2, k) = eve(idi2, k);
    eui(ide1, k) = sue(jdi2, k); evi(ide1, k) = sve(jdi2, k);
    sui(jde2, k) = eue(idi1, k); svi(jde2, k) = eve(idi1, k);
  }
}

void Oracle::bdyval() {
  const int iboudy = cfg.iboudy;
  const double xt = xbctime + dt;
  if (integrating()) {
    // fill boundary values of xxb from xxa (:1126-1310)
    for (int k = 1; k <= kz; ++k)
      for (int i = idi1; i <= idi2; ++i) {
        a2_u(jde1, i, k) = a1_u(jde1, i, k); a2_v(jde1, i, k) = a1_v(jde1, i, k);
        a2_u(jde2, i, k) = a1_u(jde2, i, k); a2_v(jde2, i, k) = a1_v(jde2, i, k);
      }
    for (int k = 1; k <= kz; ++k)
      for (int i = ici1; i <= ici2; ++i) {
        a2_t(jce1, i, k) = a1_t(jce1, i, k); a2_t(jce2, i, k) = a1_t(jce2, i, k);
        a2_pp(jce1, i, k) = a1_pp(jce1, i, k); a2_pp(jce2, i, k) = a1_pp(jce2, i, k);
        for (int n = 1; n <= nqx; ++n) {
          a2_qx(jce1, i, k, n) = a1_qx(jce1, i, k, n); a2_qx(jce2, i, k, n) = a1_qx(jce2, i, k, n);
        }
      }
    for (int k = 1; k <= kzp1; ++k)
      for (int i = ici1; i <= ici2; ++i) {
        a2_w(jce1, i, k) = a1_w(jce1, i, k); a2_w(jce2, i, k) = a1_w(jce2, i, k);
      }
    for (int k = 1; k <= kz; ++k)
      for (int j = jde1; j <= jde2; ++j) {
        a2_u(j, ide1, k) = a1_u(j, ide1, k); a2_v(j, ide1, k) = a1_v(j, ide1, k);
        a2_u(j, ide2, k) = a1_u(j, ide2, k); a2_v(j, ide2, k) = a1_v(j, ide2, k);
      }
    for (int k = 1; k <= kz; ++k)
      for (int j = jce1; j <= jce2; ++j) {
        a2_t(j, ice1, k) = a1_t(j, ice1, k); a2_t(j, ice2, k) = a1_t(j, ice2, k);
        a2_pp(j, ice1, k) = a1_pp(j, ice1, k); a2_pp(j, ice2, k) = a1_pp(j, ice2, k);
        for (int n = 1; n <= nqx; ++n) {
          a2_qx(j, ice1, k, n) = a1_qx(j, ice1, k, n); a2_qx(j, ice2, k, n) = a1_qx(j, ice2, k, n);
        }
      }
    for (int k = 1; k <= kzp1; ++k)
      for (int j = jce1; j <= jce2; ++j) {
        a2_w(j, ice1, k) = a1_w(j, ice1, k); a2_w(j, ice2, k) = a1_w(j, ice2, k);
      }
  }
  if (hydro) bdyval_ps();  // p* rim first (:1317-1342,1426-1451)
  // boundary values of p*u, p*v (:1393-1425 fixed, :1494-1525 time dependent)
  const double xtu = (iboudy == 0) ? d_zero : xt;
  auto bc = [&](const V3dBound& b, int j, int i, int k) {
    return (iboudy == 0) ? b.b0(j, i, k) : b.b0(j, i, k) + xtu * b.bt(j, i, k);
  };
  for (int k = 1; k <= kz; ++k)
    for (int i = idi1; i <= idi2; ++i) {
      a1_u(jde1, i, k) = bc(xub, jde1, i, k); a1_v(jde1, i, k) = bc(xvb, jde1, i, k);
      a1_u(jde2, i, k) = bc(xub, jde2, i, k); a1_v(jde2, i, k) = bc(xvb, jde2, i, k);
    }
  for (int k = 1; k <= kz; ++k)
    for (int j = jde1; j <= jde2; ++j) {
      a1_u(j, ide1, k) = bc(xub, j, ide1, k); a1_v(j, ide1, k) = bc(xvb, j, ide1, k);
      a1_u(j, ide2, k) = bc(xub, j, ide2, k); a1_v(j, ide2, k) = bc(xvb, j, ide2, k);
    }
  bdyuv(xt);
  // p*t, p*qv, p*pp, p*w (:1577-1656 fixed, :1708-1793 time dependent)
  for (int side = 0; side < 2; ++side) {
    const int jb = side ? jce2 : jce1, jin = side ? jci2 : jci1;
    for (int k = 1; k <= kz; ++k)
      for (int i = ici1; i <= ici2; ++i) {
        a1_t(jb, i, k) = bc(xtb, jb, i, k);
        a1_qx(jb, i, k, iqv) = bc(xqb, jb, i, k);
      }
    for (int k = 1; k <= kz; ++k)
      for (int i = ici1; i <= ici2; ++i) a1_pp(jb, i, k) = bc(xppb, jb, i, k);
    for (int k = 1; k <= kzp1; ++k)
      for (int i = ici1; i <= ici2; ++i) a1_w(jb, i, k) = bc(xwwb, jb, i, k);
    if (iboudy != 0)
      for (int i = ici1; i <= ici2; ++i) a1_w(jb, i, 1) = a1_w(jin, i, 1);
  }
  for (int side = 0; side < 2; ++side) {
    const int ib_ = side ? ice2 : ice1, iin = side ? ici2 : ici1;
    for (int k = 1; k <= kz; ++k)
      for (int j = jce1; j <= jce2; ++j) {
        a1_t(j, ib_, k) = bc(xtb, j, ib_, k);
        a1_qx(j, ib_, k, iqv) = bc(xqb, j, ib_, k);
      }
    for (int k = 1; k <= kz; ++k)
      for (int j = jce1; j <= jce2; ++j) a1_pp(j, ib_, k) = bc(xppb, j, ib_, k);
    for (int k = 1; k <= kzp1; ++k)
      for (int j = jce1; j <= jce2; ++j) a1_w(j, ib_, k) = bc(xwwb, j, ib_, k);
    if (iboudy != 0)
      for (int j = jce1; j <= jce2; ++j) a1_w(j, ib_, 1) = a1_w(j, iin, 1);
  }
  // iboudy 3, 4: water vapour on the boundary cells by inflow / outflow (Main/mod_bdycod.F90:1879-1947):
  // inflow keeps the boundary value, outflow takes the interior neighbour; both pass through (q/p*)*p*
  if (iboudy == 3 || iboudy == 4) {
    for (int k = 1; k <= kz; ++k)
      for (int i = ici1; i <= ici2; ++i) {
        double qext = a1_qx(jce1, i, k, iqv) / psa(jce1, i), qint = a1_qx(jci1, i, k, iqv) / psa(jci1, i);
        double windavg = wue(i, k) + wue(i + 1, k) + wui(i, k) + wui(i + 1, k);
        a1_qx(jce1, i, k, iqv) = (windavg > d_zero ? qext : qint) * psa(jce1, i);
      }
    for (int k = 1; k <= kz; ++k)
      for (int i = ici1; i <= ici2; ++i) {
        double qext = a1_qx(jce2, i, k, iqv) / psa(jce2, i), qint = a1_qx(jci2, i, k, iqv) / psa(jci2, i);
        double windavg = eue(i, k) + eue(i + 1, k) + eui(i, k) + eui(i + 1, k);
        a1_qx(jce2, i, k, iqv) = (windavg < d_zero ? qext : qint) * psa(jce2, i);
      }
    for (int k = 1; k <= kz; ++k)
      for (int j = jce1; j <= jce2; ++j) {
        double qext = a1_qx(j, ice1, k, iqv) / psa(j, ice1), qint = a1_qx(j, ici1, k, iqv) / psa(j, ici1);
        double windavg = sve(j, k) + sve(j + 1, k) + svi(j, k) + svi(j + 1, k);
        a1_qx(j, ice1, k, iqv) = (windavg > d_zero ? qext : qint) * psa(j, ice1);
      }
    for (int k = 1; k <= kz; ++k)
      for (int j = jce1; j <= jce2; ++j) {
        double qext = a1_qx(j, ice2, k, iqv) / psa(j, ice2), qint = a1_qx(j, ici2, k, iqv) / psa(j, ici2);
        double windavg = nve(j, k) + nve(j + 1, k) + nvi(j, k) + nvi(j + 1, k);
        a1_qx(j, ice2, k, iqv) = (windavg < d_zero ? qext : qint) * psa(j, ice2);
      }
  }
  // hydrometeors: no boundary data (.not. present_qc), bdyflow = .true.
  // inflow -> 0, outflow -> interior value (:2153-2221)
  for (int n = iqfrst; n <= iqlst; ++n) {
    for (int k = 1; k <= kz; ++k)
      for (int i = ice1; i <= ice2; ++i) {
        double qxint = a1_qx(jci1, i, k, n) / psa(jci1, i);
        double windavg = wue(i, k) + wue(i + 1, k) + wui(i, k) + wui(i + 1, k);
        if (windavg > d_zero) a1_qx(jce1, i, k, n) = d_zero;
        else a1_qx(jce1, i, k, n) = qxint * psa(jce1, i);
      }
    for (int k = 1; k <= kz; ++k)
      for (int i = ice1; i <= ice2; ++i) {
        double qxint = a1_qx(jci2, i, k, n) / psa(jci2, i);
        double windavg = eue(i, k) + eue(i + 1, k) + eui(i, k) + eui(i + 1, k);
        if (windavg < d_zero) a1_qx(jce2, i, k, n) = d_zero;
        else a1_qx(jce2, i, k, n) = qxint * psa(jce2, i);
      }
    for (int k = 1; k <= kz; ++k)
      for (int j = jci1; j <= jci2; ++j) {
        double qxint = a1_qx(j, ici1, k, n) / psa(j, ici1);
        double windavg = sve(j, k) + sve(j + 1, k) + svi(j, k) + svi(j + 1, k);
        if (windavg > d_zero) a1_qx(j, ice1, k, n) = d_zero;
        else a1_qx(j, ice1, k, n) = qxint * psa(j, ice1);
      }
    for (int k = 1; k <= kz; ++k)
      for (int j = jci1; j <= jci2; ++j) {
        double qxint = a1_qx(j, ici2, k, n) / psa(j, ici2);
        double windavg = nve(j, k) + nve(j + 1, k) + nvi(j, k) + nvi(j + 1, k);
        if (windavg < d_zero) a1_qx(j, ice2, k, n) = d_zero;
        else a1_qx(j, ice2, k, n) = qxint * psa(j, ice2);
      }
  }
  if (uw_on) {   // TKE (ibltyp = 2): the rim copy of :1166-1307 and the block of :2412-2496
    if (integrating()) bdyval_tke_copy();
    bdyval_tke();
  }
  if (ntr > 0) bdyval_tracers(xbctime + dt);   // chem_bdyval (Main/mod_bdycod.F90:2558-2564)
  xbctime = xbctime + dtsec;
}

// ---- sponge, iboudy = 4, sigma-coordinate cores: sponge3d (Main/mod_bdycod.F90:2919-2994), sponge4d (:2591-2663),
//      spongeuv (:2735-2813), sponge2d (:3065-3122).  ften = wgt(ib)*ften + (1-wgt(ib))*bt in the four zones
//      (the zones are disjoint: every point is visited at most once). ----
void Oracle::sponge3d(const V3dBound& bnd, A3& ften) {
  const int nk = ften.n3;
  for (int k = 1; k <= nk; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) {
        if (!(ba_cr.bsouth(j, i) || ba_cr.bnorth(j, i) || ba_cr.bwest(j, i) || ba_cr.beast(j, i))) continue;
        const int ib = ba_cr.ibnd(j, i);
        ften(j, i, k) = wgtx(ib) * ften(j, i, k) + (d_one - wgtx(ib)) * bnd.bt(j, i, k);
      }
}
void Oracle::sponge4d(const V3dBound& bnd, A4& ften, int m) {
  for (int k = 1; k <= kz; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) {
        if (!(ba_cr.bsouth(j, i) || ba_cr.bnorth(j, i) || ba_cr.bwest(j, i) || ba_cr.beast(j, i))) continue;
        const int ib = ba_cr.ibnd(j, i);
        ften(j, i, k, m) = wgtx(ib) * ften(j, i, k, m) + (d_one - wgtx(ib)) * bnd.bt(j, i, k);
      }
}
void Oracle::spongeuv(const V3dBound& bu, const V3dBound& bv, A3& ftu, A3& ftv) {
  for (int k = 1; k <= kz; ++k)
    for (int i = idi1; i <= idi2; ++i)
      for (int j = jdi1; j <= jdi2; ++j) {
        if (!(ba_dt.bsouth(j, i) || ba_dt.bnorth(j, i) || ba_dt.bwest(j, i) || ba_dt.beast(j, i))) continue;
        const int ib = ba_dt.ibnd(j, i);
        ftu(j, i, k) = wgtd(ib) * ftu(j, i, k) + (d_one - wgtd(ib)) * bu.bt(j, i, k);
        ftv(j, i, k) = wgtd(ib) * ftv(j, i, k) + (d_one - wgtd(ib)) * bv.bt(j, i, k);
      }
}
void Oracle::sponge2d(const V2dBound& bnd, A2& ften) {
  for (int i = ici1; i <= ici2; ++i)
    for (int j = jci1; j <= jci2; ++j) {
      if (!(ba_cr.bsouth(j, i) || ba_cr.bnorth(j, i) || ba_cr.bwest(j, i) || ba_cr.beast(j, i))) continue;
      const int ib = ba_cr.ibnd(j, i);
      ften(j, i) = wgtx(ib) * ften(j, i) + (d_one - wgtx(ib)) * bnd.bt(j, i);
    }
}

}  // namespace orc
