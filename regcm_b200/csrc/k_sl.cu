// Semi-Lagrangian horizontal advection of the moisture species and the tracers (isladvec = 1).
//   k_sl   one thread per (j,i) column and species (grid.z: 0 = qv, 1 = qc, 2.. = tracers) marching k: the horizontal
//          advection tendency slhadv_x(.., atm2) + hdvg_x(.., atm1) of Main/mod_tendency.F90:1361-1363,1378-1380,
//          1397-1399 into a scratch field that k_tq / k_h_tq / k_tr read in place of their flux-form term.
// A kernel of its own so that the default path of the hot kernels carries one predicated load and nothing else
// (a call inside them cost 100 B of spills per thread in k_tq_tma and k_h_tq; measured with ptxas -v).
#include "dev_common.cuh"
#include "launch.h"

using namespace rcm;

namespace {

// Semi-Lagrangian horizontal advection of a p*-coupled scalar (isladvec = 1; Main/mod_sladvection.F90 as `tend` uses it,
// Main/mod_tendency.F90:1283-1285,1361-1363,1378-1380,1397-1399): adv_velocity(.false.) :92-114 (the last term of vadym1_x
// repeats va(j+1,i,k) as the reference does), trajcalc_x :139-221, slhadv_x :347-393 on time level b (`vb`), hdvg_x
// :568-588 on the level-a value `va_c`.  Returns the tendency 0 + (value - vb(j,i,k))/dt - va_c*hdvg.  The departure
// geometry is recomputed per species (an optional path).  Single rank (the stencil reaches three cells out, clamped
// to the cross grid on the physical boundary, :166-177,200-211).  A departure point more than one cell away sets
// *flag = 2 (the reference calls fatal, :152-157,186-191).
__device__ __forceinline__ double sl_hadv(const double* __restrict__ au, const double* __restrict__ av, const double* __restrict__ rpsda,
                                       const double* __restrict__ msfd, const double* __restrict__ msfx,
                                       const double* __restrict__ vb, const double va_c, const int si, const long long sk,
                                       const int jx, const int iy, const int io, const int iqmsl, const double dt, const double dx,
                                       int* flag, const int j, const int i, const int k) {
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
  const double uadvx = 0.25 * (u00 + u01 + u11 + u10) / L2(msfx, j, i);
  const double uadxp1 = 0.25 * (u10 + u11 + wind(au, j + 2, i + 1) + wind(au, j + 2, i)) / L2(msfx, j + 1, i);
  const double uadxm1 = 0.25 * (u00 + u01 + wind(au, j - 1, i + 1) + wind(au, j - 1, i)) / L2(msfx, j - 1, i);
  const double vadvy = 0.25 * (v00 + v01 + v11 + v10) / L2(msfx, j, i);
  const double vadyp1 = 0.25 * (v01 + v11 + wind(av, j + 1, i + 2) + wind(av, j, i + 2)) / L2(msfx, j, i + 1);
  const double vadym1 = 0.25 * (v00 + wind(av, j, i - 1) + v10 + v10) / L2(msfx, j, i - 1);
  const double dtsq = dt * dt, dtcb = dt * dt * dt;
  const int jce2 = jx - 1, ice2 = iy - 1;
  int x0, xm1, xm2, xp1, y0, ym1, ym2, yp1;
  double al, be;
  {
    const double ux = 0.5 * (uadxp1 - uadxm1) / dx;
    const double uxx = (uadxp1 - 2.0 * uadvx + uadxm1) / (dx * dx);
    const double xdis = -uadvx * dt + 0.5 * (dtsq * uadvx * ux) - (dtcb * uadvx) * (ux * ux + uadvx * uxx) / 6.0;
    const double xn = xdis / dx;
    int xnp = (int)xn;
    if (abs(xnp) > 1 || !(xn == xn)) { *flag = 2; xnp = 0; }
    al = fabs((xnp * dx - xdis) / dx);
    const int xsn = (int)copysign(1.0, xn);
    x0 = j + xnp; xm1 = x0 + xsn; xm2 = xm1 + xsn; xp1 = x0 - xsn;
    x0 = min(max(x0, 1), jce2); xm1 = min(max(xm1, 1), jce2); xm2 = min(max(xm2, 1), jce2); xp1 = min(max(xp1, 1), jce2);
  }
  {
    const double vy = 0.5 * (vadyp1 - vadym1) / dx;
    const double vyy = (vadyp1 - 2.0 * vadvy + vadym1) / (dx * dx);
    const double ydis = -vadvy * dt + 0.5 * (dtsq * vadvy * vy) - (dtcb * vadvy) * (vy * vy + vadvy * vyy) / 6.0;
    const double yn = ydis / dx;
    int ynp = (int)yn;
    if (abs(ynp) > 1 || !(yn == yn)) { *flag = 2; ynp = 0; }
    be = fabs((ynp * dx - ydis) / dx);
    const int ysn = (int)copysign(1.0, yn);
    y0 = i + ynp; ym1 = y0 + ysn; ym2 = ym1 + ysn; yp1 = y0 - ysn;
    y0 = min(max(y0, 1), ice2); ym1 = min(max(ym1, 1), ice2); ym2 = min(max(ym2, 1), ice2); yp1 = min(max(yp1, 1), ice2);
  }
  const double am2 = -(al * (1.0 - al * al)) / 6.0, am1 = (al * (1.0 + al) * (2.0 - al)) / 2.0,
               a0 = ((1.0 - al * al) * (2.0 - al)) / 2.0, ap1 = -(al * (1.0 - al) * (2.0 - al)) / 6.0;
  const double bm2 = -(be * (1.0 - be * be)) / 6.0, bm1 = (be * (1.0 + be) * (2.0 - be)) / 2.0,
               b0 = ((1.0 - be * be) * (2.0 - be)) / 2.0, bp1 = -(be * (1.0 - be) * (2.0 - be)) / 6.0;
#define SLV(x, y) L3(vb, x, y, k)
  const double v_00 = SLV(x0, y0), v_0m = SLV(x0, ym1), v_m0 = SLV(xm1, y0), v_mm = SLV(xm1, ym1);
  const double bl1 = al * SLV(xm1, yp1) + (1.0 - al) * SLV(x0, yp1);
  const double bl2 = al * SLV(xm1, ym2) + (1.0 - al) * SLV(x0, ym2);
  const double cb1 = am2 * SLV(xm2, y0) + am1 * v_m0 + a0 * v_00 + ap1 * SLV(xp1, y0);
  const double cb2 = am2 * SLV(xm2, ym1) + am1 * v_mm + a0 * v_0m + ap1 * SLV(xp1, ym1);
#undef SLV
  const double tbadp = bm2 * bl2 + bm1 * cb2 + b0 * cb1 + bp1 * bl1;
  double tsla = tbadp;
  if (iqmsl == 1) {
    const double tbmax = fmax(fmax(v_00, v_0m), fmax(v_m0, v_mm)), tbmin = fmin(fmin(v_00, v_0m), fmin(v_m0, v_mm));
    if (tbadp > tbmax) tsla = tbmax;
    else if (tbadp < tbmin) tsla = tbmin;
  }
  const double vb_c = L3(vb, j, i, k);
  double ten = 0.0;
  if (fabs(tsla - vb_c) > dlowval) ten = ten + (tsla - vb_c) / dt;
  // hdvg_x
  const double m00 = L2(msfd, j, i), m01 = L2(msfd, j, i + 1), m10 = L2(msfd, j + 1, i), m11 = L2(msfd, j + 1, i + 1);
  const double ucapf = (u11 * m11 + u10 * m10) * 0.5, ucapi = (u01 * m01 + u00 * m00) * 0.5;
  const double vcapf = (v11 * m11 + v01 * m01) * 0.5, vcapi = (v10 * m10 + v00 * m00) * 0.5;
  const double hdvg = ((ucapf - ucapi) / dx + (vcapf - vcapi) / dx) / (L2(msfx, j, i) * L2(msfx, j, i));
  const double tat = (va_c > 2.220446049250313e-16) ? va_c * hdvg : 0.0;
  return ten - tat;
}

__global__ void __launch_bounds__(RCM_MAXT)
k_sl(const Geom g, const Fields f, const TrFields tr, const StepScalars s) {
  const int si = g.si;
  const long long sk = g.sk;
  const int j = g.jci1 + blockIdx.x * blockDim.x + threadIdx.x;
  const int i = g.ici1 + blockIdx.y * blockDim.y + threadIdx.y;
  const int n = blockIdx.z;
  if (j > g.jci2 || i > g.ici2) return;
  const double* __restrict__ vb = n == 0 ? f.bqv : n == 1 ? f.bqc : tr.b[n - 2];
  const double* __restrict__ va = n == 0 ? f.aqv : n == 1 ? f.aqc : tr.a[n - 2];
  double* __restrict__ out = n == 0 ? f.slqv : n == 1 ? f.slqc : tr.sl[n - 2];
  const int io = (s.iboudy == 3 || s.iboudy == 4) ? 1 : 0;
  for (int k = 1; k <= g.kz; ++k)
    P3(out, j, i, k) = sl_hadv(f.au, f.av, f.rpsda, f.msfd, f.msfx, vb, L3(va, j, i, k), si, sk, g.jx, g.iy, io, s.iqmsl, s.dt, s.dx,
                               f.nanflag, j, i, k);
}

}  // namespace

void launch_sl(const Geom& g, const Fields& f, const TrFields& tr, const StepScalars& s, cudaStream_t st) {
  const dim3 b = rcm::col_block();
  const dim3 gr((g.jci2 - g.jci1 + b.x) / b.x, (g.ici2 - g.ici1 + b.y) / b.y, 2 + tr.ntr);
  RCM_LAUNCH(k_sl, gr, b, st, g, f, tr, s);
}
