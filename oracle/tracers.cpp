// ORACLE — TEST INFRASTRUCTURE ONLY.  Never on the product path.
// Tracer transport of the sigma-coordinate cores: CPU restatement of what the reference does
// with atm1%chi / atm2%chi when ichem = 1 and physics is off, time-dependent chemical boundary conditions
// (ichebdy = 1):
//   decouple                  Main/mod_tendency.F90:1029-1034   atmx%chi = atm1%chi * rpsa
//   mkslice                   Main/mod_slice.F90:198            chib3d = max(atm2%chi * rpsb, 0)
//   hadv  -> hadvtr           Main/mod_advection.F90:666-720
//   vadv  -> vadv4d(ind = 2)  Main/mod_advection.F90:895-914    (itrvadv = 2, Main/mod_tendency.F90:149)
//   nudge_chi -> nudge_chiten Main/chemlib/mod_che_bdyco.F90:792-986 (ichebdy = 1 branch :898-981),
//                             coefficients setup_che_bdycon :1177-1217
//   diffu_x -> diffu_x4d      Main/mod_diffusion.F90 (same operator as for qx)
//   sums, forecast, negative fix (factor 0.1), RAW filter (gnu2, 0.53, floor 0)   Main/mod_tendency.F90:548-586
//   chem_bdyval_coupled       Main/chemlib/mod_che_bdyco.F90:565-760 (ichebdy = 1 branch :697-754)
//   che_init_bdy / chem_bdyin Main/chemlib/mod_che_bdyco.F90:118-300,302-420: couple with xpsb%b0/b1, chibt
// Not restated: chemistry itself, emissions, cumulus transport (ichcumtra), ichebdy = 0, vadv ind = 3 (needs the
// PBL height of the UW scheme), TKE (ibltyp = 2 needs the PBL scheme's state).
#include <algorithm>
#include <cmath>

#include "model.hpp"

namespace orc {

void Oracle::allocate_tracers() {
  ntr = cfg.ntr;
  if (ntr <= 0) return;
  auto c3 = [&](A3& a) { a.alloc(jce1 - 2, jce2 + 2, ice1 - 2, ice2 + 2, 1, kz); };
  chia.resize(ntr); chib.resize(ntr); chic.resize(ntr); chix.resize(ntr); chib3d.resize(ntr);
  chidyn.resize(ntr); chiten.resize(ntr); xchib.resize(ntr); nx_chi.resize(ntr);
  for (int n = 0; n < ntr; ++n) {
    for (A3* a : {&chia[n], &chib[n], &chic[n], &chix[n], &chib3d[n], &chidyn[n], &chiten[n], &xchib[n].b0, &xchib[n].b1,
                  &xchib[n].bt, &nx_chi[n]})
      c3(*a);
  }
  cefc.alloc(1, cfg.nspgx, 1, kz); cegc.alloc(1, cfg.nspgx, 1, kz);
}

// che_init_bdy (:257-290): couple both records with the boundary p*, tendency by DIVISION by dtbdys
void Oracle::init_bdy_tracers() {
  if (ntr <= 0) return;
  // setup_che_bdycon (:1177-1217): always the exponential profile; dt2 = 2 * dtsec (Main/mod_params.F90:1708)
  const double chdt2 = d_two * dtsec;
  const double fn = (cfg.bdy_nm > d_zero) ? cfg.bdy_nm : 0.1 / chdt2;
  const double gn = (cfg.bdy_dm > d_zero) ? cfg.bdy_dm : d_one / (50.0 * chdt2);
  for (int k = 1; k <= kz; ++k) {
    const double an = hsigma(k) < 0.4 ? cfg.high_nudge : (hsigma(k) < 0.8 ? cfg.medium_nudge : cfg.low_nudge);
    for (int n = 2; n <= cfg.nspgx - 1; ++n) {
      const double xfune = std::exp(-(double)(n - 2) / an);
      cefc(n, k) = fn * xfune; cegc(n, k) = gn * xfune;
    }
  }
  const A2& p0 = hydro ? xpsb.b0 : xpsb_b0;
  const A2& p1 = hydro ? xpsb.b1 : xpsb_b0;
  for (int n = 0; n < ntr; ++n) {
    couple(xchib[n].b0, p0, jce1, jce2, ice1, ice2, 1, kz);
    couple(xchib[n].b1, p1, jce1, jce2, ice1, ice2, 1, kz);
    for (int k = 1; k <= kz; ++k)
      for (int i = ice1; i <= ice2; ++i)
        for (int j = jce1; j <= jce2; ++j) xchib[n].bt(j, i, k) = (xchib[n].b1(j, i, k) - xchib[n].b0(j, i, k)) / dtbdys;
  }
}

// initial tracer state = the first boundary record (the reference reads it from the same ICBC/CHBC file)
void Oracle::init_state_tracers() {
  for (int n = 0; n < ntr; ++n)
    for (int k = 1; k <= kz; ++k)
      for (int i = ice1; i <= ice2; ++i)
        for (int j = jce1; j <= jce2; ++j) { chia[n](j, i, k) = xchib[n].b0(j, i, k); chib[n](j, i, k) = xchib[n].b0(j, i, k); }
}

// chem_bdyin (:326-413)
void Oracle::bdyin_tracers() {
  const A2& p1 = hydro ? xpsb.b1 : xpsb_b0;   // called after the meteorological bdyin: xpsb%b1 is the new record
  for (int n = 0; n < ntr; ++n) {
    xchib[n].b0.d = xchib[n].b1.d;
    xchib[n].b1.d = nx_chi[n].d;
    couple(xchib[n].b1, p1, jce1, jce2, ice1, ice2, 1, kz);
    for (int k = 1; k <= kz; ++k)
      for (int i = ice1; i <= ice2; ++i)
        for (int j = jce1; j <= jce2; ++j) xchib[n].bt(j, i, k) = (xchib[n].b1(j, i, k) - xchib[n].b0(j, i, k)) / dtbdys;
  }
}

// chem_bdyval_coupled, ichebdy = 1 (:580-626,697-754); xt = xbctime + dt with the values bdyval sees
void Oracle::bdyval_tracers(double xt) {
  for (int n = 0; n < ntr; ++n) {
    A3& a = chia[n]; A3& b = chib[n];
    const V3dBound& bd = xchib[n];
    if (integrating()) {
      for (int k = 1; k <= kz; ++k) for (int i = ice1; i <= ice2; ++i) b(jce1, i, k) = a(jce1, i, k);
      for (int k = 1; k <= kz; ++k) for (int i = ice1; i <= ice2; ++i) b(jce2, i, k) = a(jce2, i, k);
      for (int k = 1; k <= kz; ++k) for (int j = jci1; j <= jci2; ++j) b(j, ice1, k) = a(j, ice1, k);
      for (int k = 1; k <= kz; ++k) for (int j = jci1; j <= jci2; ++j) b(j, ice2, k) = a(j, ice2, k);
    }
    for (int k = 1; k <= kz; ++k) for (int i = ici1; i <= ici2; ++i) a(jce1, i, k) = bd.b0(jce1, i, k) + xt * bd.bt(jce1, i, k);
    for (int k = 1; k <= kz; ++k) for (int i = ici1; i <= ici2; ++i) a(jce2, i, k) = bd.b0(jce2, i, k) + xt * bd.bt(jce2, i, k);
    for (int k = 1; k <= kz; ++k) for (int j = jce1; j <= jce2; ++j) a(j, ice1, k) = bd.b0(j, ice1, k) + xt * bd.bt(j, ice1, k);
    for (int k = 1; k <= kz; ++k) for (int j = jce1; j <= jce2; ++j) a(j, ice2, k) = bd.b0(j, ice2, k) + xt * bd.bt(j, ice2, k);
  }
}

// everything `tend` does with the tracers, in the reference's order of accumulation
void Oracle::tracer_tendency() {
  if (ntr <= 0) return;
  const bool up = cfg.upstream_mode != 0;
  const A3& svv = qdot;
  const double xt = xbctime + dt;
  const bool tiled = rt_j.size() > 1 || rt_i.size() > 1;
  auto tile_of = [](const std::vector<int>& st, int x) { int q = 0; while (q + 1 < (int)st.size() && st[q + 1] <= x) ++q; return q; };
  neg_chi_points = 0;
  for (int n = 0; n < ntr; ++n) {
    A3& a1 = chia[n]; A3& a2 = chib[n]; A3& ac = chic[n]; A3& ax = chix[n]; A3& b3 = chib3d[n];
    A3& dyn = chidyn[n]; A3& ten = chiten[n];
    dyn.fill(d_zero); ten.fill(d_zero);
    for (int k = 1; k <= kz; ++k)
      for (int i = ice1; i <= ice2; ++i)
        for (int j = jce1; j <= jce2; ++j) {
          ax(j, i, k) = a1(j, i, k) * rpsa(j, i);
          b3(j, i, k) = std::max(a2(j, i, k) * rpsb(j, i), d_zero);
        }
    if (cfg.isladvec == 1) {   // Main/mod_tendency.F90:1397-1399
      fg.fill(d_zero);
      slhadv_x([&](int j, int i, int k) -> double& { return fg(j, i, k); }, [&](int j, int i, int k) { return a2(j, i, k); });
      hdvg_x([&](int j, int i, int k) -> double& { return fg(j, i, k); }, [&](int j, int i, int k) { return a1(j, i, k); });
    } else
    // hadvtr (:666-720)
    for (int k = 1; k <= kz; ++k)
      for (int i = ici1; i <= ici2; ++i)
        for (int j = jci1; j <= jci2; ++j) {
          double fx1, fx2, fy1, fy2;
          if (!up) {
            fx1 = ax(j - 1, i, k) + ax(j, i, k); fx2 = ax(j, i, k) + ax(j + 1, i, k);
            fy1 = ax(j, i - 1, k) + ax(j, i, k); fy2 = ax(j, i, k) + ax(j, i + 1, k);
          } else {
            const double f1 = d_half * ul * (uavg2(j, i, k) + uavg1(j, i, k)) / psa(j, i);
            const double f2 = d_half * ul * (vavg2(j, i, k) + vavg1(j, i, k)) / psa(j, i);
            fx1 = (d_one + f1) * ax(j - 1, i, k) + (d_one - f1) * ax(j, i, k);
            fx2 = (d_one + f1) * ax(j, i, k) + (d_one - f1) * ax(j + 1, i, k);
            fy1 = (d_one + f2) * ax(j, i - 1, k) + (d_one - f2) * ax(j, i, k);
            fy2 = (d_one + f2) * ax(j, i, k) + (d_one - f2) * ax(j, i + 1, k);
          }
          fg(j, i, k) = -xmsf(j, i) * (uavg2(j, i, k) * fx2 - uavg1(j, i, k) * fx1 + vavg2(j, i, k) * fy2 - vavg1(j, i, k) * fy1);
        }
    for (int k = 1; k <= kz; ++k)
      for (int i = ici1; i <= ici2; ++i)
        for (int j = jci1; j <= jci2; ++j) dyn(j, i, k) = dyn(j, i, k) + fg(j, i, k);
    // vadv4d, ind = 2 (:895-914): on p*-coupled values, threshold mintr * ps
    fg.fill(d_zero);
    if (itrvadv == 3) vadv_ind3([&](int j, int i, int k) { return a1(j, i, k); });   // iuwvadv = 1 (Main/mod_tendency.F90:150-154)
    else
    for (int k = 2; k <= kz; ++k)
      for (int i = ici1; i <= ici2; ++i)
        for (int j = jci1; j <= jci2; ++j) {
          const double thr = mintr * psa(j, i);
          const bool ok = (svv(j, i, k) > d_zero) ? (a1(j, i, k - 1) > thr) : (a1(j, i, k) > thr);
          fg(j, i, k) = ok ? svv(j, i, k) * (twt(k, 1) * a1(j, i, k) + twt(k, 2) * a1(j, i, k - 1)) : d_zero;
        }
    for (int k = 2; k <= kz; ++k)
      for (int i = ici1; i <= ici2; ++i)
        for (int j = jci1; j <= jci2; ++j) {
          dyn(j, i, k - 1) = dyn(j, i, k - 1) - fg(j, i, k) * xds(k - 1);
          dyn(j, i, k) = dyn(j, i, k) + fg(j, i, k) * xds(k);
        }
    // nudge_chiten, ichebdy = 1 (:898-981): the four zones use the same expression (their neighbour order differs:
    // south/north take j-1, j+1, i-1, i+1; west/east take i-1, i+1, j-1, j+1)
    if ((cfg.iboudy == 1 || cfg.iboudy == 5) && ba_cr.havebound) {
      auto F = [&](int jj, int ii, int k) { return (xchib[n].b0(jj, ii, k) + xt * xchib[n].bt(jj, ii, k)) - a2(jj, ii, k); };
      const I2* zs[4] = {&ba_cr.bsouth, &ba_cr.bnorth, &ba_cr.bwest, &ba_cr.beast};
      const int zn[4] = {ba_cr.ns, ba_cr.nn, ba_cr.nw, ba_cr.ne};
      for (int z = 0; z < 4; ++z) {
        if (zn[z] == 0) continue;
        for (int k = 1; k <= kz; ++k)
          for (int i = ici1; i <= ici2; ++i)
            for (int j = jci1; j <= jci2; ++j) {
              if (!(*zs[z])(j, i)) continue;
              const int ib = ba_cr.ibnd(j, i);
              const double xf = cefc(ib, k), xg = cegc(ib, k);
              const double fls0 = F(j, i, k);
              double fls1, fls2, fls3, fls4;
              if (z < 2) { fls1 = F(j - 1, i, k); fls2 = F(j + 1, i, k); fls3 = F(j, i - 1, k); fls4 = F(j, i + 1, k); }
              else { fls1 = F(j, i - 1, k); fls2 = F(j, i + 1, k); fls3 = F(j - 1, i, k); fls4 = F(j + 1, i, k); }
              dyn(j, i, k) = dyn(j, i, k) + xf * fls0 - xg * (fls1 + fls2 + fls3 + fls4 - d_four * fls0);
            }
      }
    }
    // diffu_x4d(chidyn, chib3d, 1, ntr, 1.0)
    if (cfg.idiffu == 3) {   // Main/mod_diffusion.F90:895-960: j = jci2 only, clamped stencil, flux sign from b3 / msfd
      const int j = jci2;
      for (int k = 1; k <= kz; ++k)
        for (int i = ici1; i <= ici2; ++i) {
          auto cl = [&](int v, int hi) { return std::min(std::max(v, 1), hi); };
          auto F = [&](int jj, int ii) { return b3(jj, ii, k); };
          auto G = [&](int jj, int ii) { return b3(jj, ii, k) / msfd(jj, ii); };
          const int jm1 = cl(j - 1, jx - 1), jm2 = cl(j - 2, jx - 1), jm3 = cl(j - 3, jx - 1), jp1 = cl(j + 1, jx - 1), jp2 = cl(j + 2, jx - 1), jp3 = cl(j + 3, jx - 1);
          const int im1 = cl(i - 1, iy - 1), im2 = cl(i - 2, iy - 1), im3 = cl(i - 3, iy - 1), ip1 = cl(i + 1, iy - 1), ip2 = cl(i + 2, iy - 1), ip3 = cl(i + 3, iy - 1);
          double x0 = (10.0 * (F(j, i) - F(jm1, i)) + (-5.0) * (F(jp1, i) - F(jm2, i)) + 1.0 * (F(jp2, i) - F(jm3, i)));
          if (x0 * (G(j, i) - G(jm1, i)) <= d_zero) x0 = d_zero;
          double x1 = (10.0 * (F(jp1, i) - F(j, i)) + (-5.0) * (F(jp2, i) - F(jm1, i)) + 1.0 * (F(jp3, i) - F(jm2, i)));
          if (x1 * (G(jp1, i) - G(j, i)) <= d_zero) x1 = d_zero;
          double y0 = (10.0 * (F(j, i) - F(j, im1)) + (-5.0) * (F(j, ip1) - F(j, im2)) + 1.0 * (F(j, ip2) - F(j, im3)));
          if (y0 * (G(j, i) - G(j, im1)) <= d_zero) y0 = d_zero;
          double y1 = (10.0 * (F(j, ip1) - F(j, i)) + (-5.0) * (F(j, ip2) - F(j, im1)) + 1.0 * (F(j, ip3) - F(j, im2)));
          if (y1 * (G(j, ip1) - G(j, i)) <= d_zero) y1 = d_zero;
          dyn(j, i, k) = dyn(j, i, k) + d_one * xkc(j, i, k) * ((x1 - x0) + (y1 - y0));
        }
    } else
    if (cfg.idiffu == 2) {   // Main/mod_diffusion.F90:881-894
      const double o4_c1 = 4.0 / 6.0, o4_c2 = 1.0 / 6.0, o4_c3 = -20.0 / 6.0;
      for (int k = 1; k <= kz; ++k)
        for (int i = ici1; i <= ici2; ++i)
          for (int j = jci1; j <= jci2; ++j)
            dyn(j, i, k) = dyn(j, i, k) + d_one * xkc(j, i, k) *
                (o4_c1 * (b3(j + 1, i, k) + b3(j - 1, i, k) + b3(j, i + 1, k) + b3(j, i - 1, k)) +
                 o4_c2 * (b3(j + 1, i + 1, k) + b3(j - 1, i - 1, k) + b3(j - 1, i + 1, k) + b3(j + 1, i - 1, k)) +
                 o4_c3 * b3(j, i, k));
    } else {
    for (int k = 1; k <= kz; ++k)
      for (int i = icii1; i <= icii2; ++i)
        for (int j = jcii1; j <= jcii2; ++j)
          dyn(j, i, k) = dyn(j, i, k) - d_one * xkc(j, i, k) *
              (1.0 * (b3(j + 2, i, k) + b3(j - 2, i, k) + b3(j, i + 2, k) + b3(j, i - 2, k)) +
               (-4.0) * (b3(j + 1, i, k) + b3(j - 1, i, k) + b3(j, i + 1, k) + b3(j, i - 1, k)) + 12.0 * b3(j, i, k));
    auto ring = [&](int j, int i, int k) {
      dyn(j, i, k) = dyn(j, i, k) + d_one * xkc(j, i, k) *
          (1.0 * (b3(j + 1, i, k) + b3(j - 1, i, k) + b3(j, i + 1, k) + b3(j, i - 1, k)) + (-4.0) * b3(j, i, k));
    };
    if (has_bdyleft) for (int k = 1; k <= kz; ++k) for (int i = ici1; i <= ici2; ++i) ring(jci1, i, k);
    if (has_bdyright) for (int k = 1; k <= kz; ++k) for (int i = ici1; i <= ici2; ++i) ring(jci2, i, k);
    if (has_bdybottom) for (int k = 1; k <= kz; ++k) for (int j = jci1; j <= jci2; ++j) ring(j, ici1, k);
    if (has_bdytop) for (int k = 1; k <= kz; ++k) for (int j = jci1; j <= jci2; ++j) ring(j, ici2, k);
    }
    // chiten = chiten + chidyn + chiphy (physics off); forecast (:548-569)
    for (int k = 1; k <= kz; ++k)
      for (int i = ici1; i <= ici2; ++i)
        for (int j = jci1; j <= jci2; ++j) ten(j, i, k) = ten(j, i, k) + dyn(j, i, k) + d_zero;
    for (int k = 1; k <= kz; ++k)
      for (int i = ice1; i <= ice2; ++i)
        for (int j = jce1; j <= jce2; ++j) ac(j, i, k) = a2(j, i, k);
    for (int k = 1; k <= kz; ++k)
      for (int i = ici1; i <= ici2; ++i)
        for (int j = jci1; j <= jci2; ++j) ac(j, i, k) = ac(j, i, k) + dt * ten(j, i, k);
    // negative fix (:570-582): sequential sweep, factor 0.1 (the moisture fix uses 0.01)
    A3 raw;
    if (tiled) raw = ac;
    for (int k = 1; k <= kz; ++k)
      for (int i = ici1; i <= ici2; ++i)
        for (int j = jci1; j <= jci2; ++j)
          if (ac(j, i, k) < d_zero) {
            double s = 0.0;
            for (int ii = i - 1; ii <= i + 1; ++ii)
              for (int jj = j - 1; jj <= j + 1; ++jj) {
                const bool other = tiled && (tile_of(rt_j, jj) != tile_of(rt_j, j) || tile_of(rt_i, ii) != tile_of(rt_i, i));
                s += std::fabs(other ? raw(jj, ii, k) : ac(jj, ii, k));
              }
            ac(j, i, k) = 0.1 * s / 9.0;
            neg_chi_points++;
          }
    // timefilter_apply(atm1%chi, atm2%chi, atmc%chi, gnu2, 0.53, 1, ntr, 0) = filter_raw_4d (Main/mod_timefilter.F90:231-278)
    const double alpha = cfg.gnu2, beta = 0.53;
    for (int k = 1; k <= kz; ++k)
      for (int i = ici1; i <= ici2; ++i)
        for (int j = jci1; j <= jci2; ++j) {
          const double d = alpha * (ac(j, i, k) + a2(j, i, k) - d_two * a1(j, i, k));
          a2(j, i, k) = a1(j, i, k) + beta * d;
          a1(j, i, k) = ac(j, i, k) + (beta - d_one) * d;
          if (a2(j, i, k) < d_zero) a2(j, i, k) = d_zero;
          if (a1(j, i, k) < d_zero) a1(j, i, k) = d_zero;
        }
  }
}

static int tr_copy(A3& dst, const double* a, const int32_t lb[3], const int32_t ub[3], bool in, double* out) {
  const int n1 = ub[0] - lb[0] + 1, n2 = ub[1] - lb[1] + 1;
  for (int k = std::max(lb[2], dst.l3); k <= std::min(ub[2], dst.l3 + dst.n3 - 1); ++k)
    for (int i = std::max(lb[1], dst.l2); i <= std::min(ub[1], dst.l2 + dst.n2 - 1); ++i)
      for (int j = std::max(lb[0], dst.l1); j <= std::min(ub[0], dst.l1 + dst.n1 - 1); ++j) {
        const size_t q = (size_t)(j - lb[0]) + (size_t)n1 * ((size_t)(i - lb[1]) + (size_t)n2 * (size_t)(k - lb[2]));
        if (in) dst(j, i, k) = a[q]; else out[q] = dst(j, i, k);
      }
  return RCM_OK;
}

int Oracle::tracer_state(int n, int level, double* a, const int32_t lb[3], const int32_t ub[3], bool in) {
  if (n < 0 || n >= ntr) { err = "no such tracer"; return RCM_ERR_ARG; }
  return tr_copy(level == RCM_LEVEL_A ? chia[n] : chib[n], a, lb, ub, in, a);
}

int Oracle::tracer_bdy(int n, int slot, const double* a, const int32_t lb[3], const int32_t ub[3]) {
  if (n < 0 || n >= ntr) { err = "no such tracer"; return RCM_ERR_ARG; }
  A3* dst = slot == RCM_BDY_B0 ? &xchib[n].b0 : slot == RCM_BDY_B1 ? &xchib[n].b1 : slot == RCM_BDY_NEXT ? &nx_chi[n] : nullptr;
  if (!dst) { err = "bad slot"; return RCM_ERR_ARG; }
  return tr_copy(*dst, a, lb, ub, true, nullptr);
}

}  // namespace orc
