// Semi-Lagrangian horizontal advection of the moisture species and the tracers (isladvec = 1).
//   k_sl   one thread per (j,i) column marching k: the departure geometry of the point once, then for every species
//          (qv, qc, the tracers) the horizontal advection tendency slhadv_x(.., atm2) + hdvg_x(.., atm1) of
//          Main/mod_tendency.F90:1361-1363,1378-1380,1397-1399 into a scratch field that k_tq / k_h_tq / k_tr read in
//          place of their flux-form term.
// A kernel of its own so that the default path of the hot kernels carries one predicated load and nothing else
// (a call inside them cost 100 B of spills per thread in k_tq_tma and k_h_tq; measured with ptxas -v).
#include "dev_common.cuh"
#include "launch.h"

using namespace rcm;

namespace {

// Semi-Lagrangian horizontal advection of a p*-coupled scalar (isladvec = 1; Main/mod_sladvection.F90 as `tend` uses it,
// Main/mod_tendency.F90:1283-1285,1361-1363,1378-1380,1397-1399): adv_velocity(.false.) :92-114 (the last term of vadym1_x
// repeats va(j+1,i,k) as the reference does), trajcalc_x :139-221, slhadv_x :347-393 on time level b (`vb`), hdvg_x
// :568-588 on the level-a value `va_c`: tendency 0 + (value - vb(j,i,k))/dt - va_c*hdvg.  Single rank (the stencil reaches
// three cells out, clamped to the cross grid on the physical boundary, :166-177,200-211).  A departure point more than one
// cell away sets *flag = 2 (the reference calls fatal, :152-157,186-191).
struct SlTraj {
  int x0, xm1, xm2, xp1, y0, ym1, ym2, yp1;
  double al, am2, am1, a0, ap1, bm2, bm1, b0, bp1;
  double hdvg;
};

// departure geometry and weights of one (j,i,k) (adv_velocity, trajcalc_x) and the divergence factor of hdvg_x: shared by
// every species advected at this point
// the divisors of a column and their reciprocals for div_r (shared-reciprocal division: same quotients, fewer instructions)
struct SlCol {
  double mx_c, mx_jp, mx_jm, mx_ip, mx_im, mx2;   // msfx at the point, its four neighbours; msfx(j,i)^2
  double r_c, r_jp, r_jm, r_ip, r_im, r_mx2, r_dx, r_dx2, r_6;
  double dx2;
};
__device__ __forceinline__ void sl_col(const double* __restrict__ msfx, const int si, const double dx, const int j, const int i, SlCol& c) {
  c.mx_c = L2(msfx, j, i); c.mx_jp = L2(msfx, j + 1, i); c.mx_jm = L2(msfx, j - 1, i); c.mx_ip = L2(msfx, j, i + 1); c.mx_im = L2(msfx, j, i - 1);
  c.mx2 = c.mx_c * c.mx_c; c.dx2 = dx * dx;
  c.r_c = recip_for_div(c.mx_c); c.r_jp = recip_for_div(c.mx_jp); c.r_jm = recip_for_div(c.mx_jm); c.r_ip = recip_for_div(c.mx_ip);
  c.r_im = recip_for_div(c.mx_im); c.r_mx2 = recip_for_div(c.mx2); c.r_dx = recip_for_div(dx); c.r_dx2 = recip_for_div(c.dx2);
  c.r_6 = recip_for_div(6.0);
}

__device__ __forceinline__ void sl_traj(const double* __restrict__ au, const double* __restrict__ av, const double* __restrict__ rpsda,
                                        const double* __restrict__ msfd, const SlCol& c, const int si, const long long sk,
                                        const int jx, const int iy, const int io, const double dt, const double dx, int* flag,
                                        const int j, const int i, const int k, SlTraj& t) {
  // atmx%umd / vmd = (u * rpsda) * msfd with the inflow/outflow rule of `decouple` on the outermost dot ring (iboudy 3, 4)
  auto wind = [&](const double* a, int jj, int ii) {
    if (io) {
      int i2 = ii, j2 = jj;
      if (ii == 1 && L3(av, jj, 1, k) >= 0.0) i2 = 2;
      else if (ii == iy && L3(av, jj, iy, k) <= 0.0) i2 = iy - 1;
      if (i2 >= 2 && i2 <= iy - 1) {
        if (jj == 1 && L3(au, 1, i2, k) <= 0.0) j2 = 2;
        else if (jj == jx && L3(au, jx, i2, k) >= 0.0) j2 = jx - 1;
      }
      return (L3(a, j2, i2, k) * L2(rpsda, j2, i2)) * L2(msfd, jj, ii);
    }
    return (L3(a, jj, ii, k) * L2(rpsda, jj, ii)) * L2(msfd, jj, ii);
  };
  const double u00 = wind(au, j, i), u01 = wind(au, j, i + 1), u10 = wind(au, j + 1, i), u11 = wind(au, j + 1, i + 1);
  const double v00 = wind(av, j, i), v01 = wind(av, j, i + 1), v10 = wind(av, j + 1, i), v11 = wind(av, j + 1, i + 1);
  const double uadvx = div_r(0.25 * (u00 + u01 + u11 + u10), c.mx_c, c.r_c);
  const double uadxp1 = div_r(0.25 * (u10 + u11 + wind(au, j + 2, i + 1) + wind(au, j + 2, i)), c.mx_jp, c.r_jp);
  const double uadxm1 = div_r(0.25 * (u00 + u01 + wind(au, j - 1, i + 1) + wind(au, j - 1, i)), c.mx_jm, c.r_jm);
  const double vadvy = div_r(0.25 * (v00 + v01 + v11 + v10), c.mx_c, c.r_c);
  const double vadyp1 = div_r(0.25 * (v01 + v11 + wind(av, j + 1, i + 2) + wind(av, j, i + 2)), c.mx_ip, c.r_ip);
  const double vadym1 = div_r(0.25 * (v00 + wind(av, j, i - 1) + v10 + v10), c.mx_im, c.r_im);
  const double dtsq = dt * dt, dtcb = dt * dt * dt;
  const int jce2 = jx - 1, ice2 = iy - 1;
  double al, be;
  {
    const double ux = div_r(0.5 * (uadxp1 - uadxm1), dx, c.r_dx);
    const double uxx = div_r(uadxp1 - 2.0 * uadvx + uadxm1, c.dx2, c.r_dx2);
    const double xdis = -uadvx * dt + 0.5 * (dtsq * uadvx * ux) - div_r((dtcb * uadvx) * (ux * ux + uadvx * uxx), 6.0, c.r_6);
    const double xn = div_r(xdis, dx, c.r_dx);
    int xnp = (int)xn;
    if (abs(xnp) > 1 || !(xn == xn)) { *flag = 2; xnp = 0; }
    al = fabs(div_r(xnp * dx - xdis, dx, c.r_dx));
    const int xsn = (int)copysign(1.0, xn);
    const int x0 = j + xnp, xm1 = x0 + xsn, xm2 = xm1 + xsn, xp1 = x0 - xsn;
    t.x0 = min(max(x0, 1), jce2); t.xm1 = min(max(xm1, 1), jce2); t.xm2 = min(max(xm2, 1), jce2); t.xp1 = min(max(xp1, 1), jce2);
  }
  {
    const double vy = div_r(0.5 * (vadyp1 - vadym1), dx, c.r_dx);
    const double vyy = div_r(vadyp1 - 2.0 * vadvy + vadym1, c.dx2, c.r_dx2);
    const double ydis = -vadvy * dt + 0.5 * (dtsq * vadvy * vy) - div_r((dtcb * vadvy) * (vy * vy + vadvy * vyy), 6.0, c.r_6);
    const double yn = div_r(ydis, dx, c.r_dx);
    int ynp = (int)yn;
    if (abs(ynp) > 1 || !(yn == yn)) { *flag = 2; ynp = 0; }
    be = fabs(div_r(ynp * dx - ydis, dx, c.r_dx));
    const int ysn = (int)copysign(1.0, yn);
    const int y0 = i + ynp, ym1 = y0 + ysn, ym2 = ym1 + ysn, yp1 = y0 - ysn;
    t.y0 = min(max(y0, 1), ice2); t.ym1 = min(max(ym1, 1), ice2); t.ym2 = min(max(ym2, 1), ice2); t.yp1 = min(max(yp1, 1), ice2);
  }
  t.al = al;
  t.am2 = div_r(-(al * (1.0 - al * al)), 6.0, c.r_6); t.am1 = (al * (1.0 + al) * (2.0 - al)) / 2.0;
  t.a0 = ((1.0 - al * al) * (2.0 - al)) / 2.0; t.ap1 = div_r(-(al * (1.0 - al) * (2.0 - al)), 6.0, c.r_6);
  t.bm2 = div_r(-(be * (1.0 - be * be)), 6.0, c.r_6); t.bm1 = (be * (1.0 + be) * (2.0 - be)) / 2.0;
  t.b0 = ((1.0 - be * be) * (2.0 - be)) / 2.0; t.bp1 = div_r(-(be * (1.0 - be) * (2.0 - be)), 6.0, c.r_6);
  // hdvg_x
  const double m00 = L2(msfd, j, i), m01 = L2(msfd, j, i + 1), m10 = L2(msfd, j + 1, i), m11 = L2(msfd, j + 1, i + 1);
  const double ucapf = (u11 * m11 + u10 * m10) * 0.5, ucapi = (u01 * m01 + u00 * m00) * 0.5;
  const double vcapf = (v11 * m11 + v01 * m01) * 0.5, vcapi = (v10 * m10 + v00 * m00) * 0.5;
  t.hdvg = div_r(div_r(ucapf - ucapi, dx, c.r_dx) + div_r(vcapf - vcapi, dx, c.r_dx), c.mx2, c.r_mx2);
}

// slhadv_x on time level b minus hdvg_x on the level-a value, for one species
__device__ __forceinline__ double sl_apply(const SlTraj& t, const double* __restrict__ vb, const double va_c, const int si,
                                           const long long sk, const int iqmsl, const double dt, const double r_dt, const int j,
                                           const int i, const int k) {
#define SLV(x, y) L3(vb, x, y, k)
  const double v_00 = SLV(t.x0, t.y0), v_0m = SLV(t.x0, t.ym1), v_m0 = SLV(t.xm1, t.y0), v_mm = SLV(t.xm1, t.ym1);
  const double bl1 = t.al * SLV(t.xm1, t.yp1) + (1.0 - t.al) * SLV(t.x0, t.yp1);
  const double bl2 = t.al * SLV(t.xm1, t.ym2) + (1.0 - t.al) * SLV(t.x0, t.ym2);
  const double cb1 = t.am2 * SLV(t.xm2, t.y0) + t.am1 * v_m0 + t.a0 * v_00 + t.ap1 * SLV(t.xp1, t.y0);
  const double cb2 = t.am2 * SLV(t.xm2, t.ym1) + t.am1 * v_mm + t.a0 * v_0m + t.ap1 * SLV(t.xp1, t.ym1);
#undef SLV
  const double tbadp = t.bm2 * bl2 + t.bm1 * cb2 + t.b0 * cb1 + t.bp1 * bl1;
  double tsla = tbadp;
  if (iqmsl == 1) {
    const double tbmax = fmax(fmax(v_00, v_0m), fmax(v_m0, v_mm)), tbmin = fmin(fmin(v_00, v_0m), fmin(v_m0, v_mm));
    if (tbadp > tbmax) tsla = tbmax;
    else if (tbadp < tbmin) tsla = tbmin;
  }
  const double vb_c = L3(vb, j, i, k);
  double ten = 0.0;
  if (fabs(tsla - vb_c) > dlowval) ten = ten + div_r(tsla - vb_c, dt, r_dt);
  const double tat = (va_c > 2.220446049250313e-16) ? va_c * t.hdvg : 0.0;
  return ten - tat;
}

// the species of one launch: qv, qc, then the tracers
struct SlList {
  int n;
  const double* vb[2 + RCM_MAXTR];
  const double* va[2 + RCM_MAXTR];
  double* out[2 + RCM_MAXTR];
};

__global__ void __launch_bounds__(RCM_MAXT)
k_sl(const Geom g, const Fields f, const SlList sp, const StepScalars s) {
  const int si = g.si;
  const long long sk = g.sk;
  const int j = g.jci1 + blockIdx.x * blockDim.x + threadIdx.x;
  const int i = g.ici1 + blockIdx.y * blockDim.y + threadIdx.y;
  if (j > g.jci2 || i > g.ici2) return;
  const int io = (s.iboudy == 3 || s.iboudy == 4) ? 1 : 0;
  SlCol c;
  sl_col(f.msfx, si, s.dx, j, i, c);
  const double r_dt = recip_for_div(s.dt);
  for (int k = 1; k <= g.kz; ++k) {
    SlTraj t;
    sl_traj(f.au, f.av, f.rpsda, f.msfd, c, si, sk, g.jx, g.iy, io, s.dt, s.dx, f.nanflag, j, i, k, t);
    for (int n = 0; n < sp.n; ++n)
      P3(sp.out[n], j, i, k) = sl_apply(t, sp.vb[n], L3(sp.va[n], j, i, k), si, sk, s.iqmsl, s.dt, r_dt, j, i, k);
  }
}

}  // namespace

void launch_sl(const Geom& g, const Fields& f, const TrFields& tr, const StepScalars& s, cudaStream_t st) {
  SlList sp{};
  sp.n = 2 + tr.ntr;
  sp.vb[0] = f.bqv; sp.va[0] = f.aqv; sp.out[0] = f.slqv;
  sp.vb[1] = f.bqc; sp.va[1] = f.aqc; sp.out[1] = f.slqc;
  for (int n = 0; n < tr.ntr; ++n) { sp.vb[2 + n] = tr.b[n]; sp.va[2 + n] = tr.a[n]; sp.out[2 + n] = tr.sl[n]; }
  const dim3 b = rcm::col_block();
  const dim3 gr((g.jci2 - g.jci1 + b.x) / b.x, (g.ici2 - g.ici1 + b.y) / b.y);
  RCM_LAUNCH(k_sl, gr, b, st, g, f, sp, s);
}
