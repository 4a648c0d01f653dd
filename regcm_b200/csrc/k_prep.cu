// K_prepare / K_sigdot_div / K_smag.
//
// k_prep fuses, per cross-point column marched in k:
//   decouple  (Main/mod_tendency.F90:1013-1065): T, qv floors, Tv, p', atm1%pr,
//             atm1%rho — only pr1, rho1 are materialised; everything else is
//             recomputed in registers downstream;
//   theta     (Main/mod_tendency.F90:1349-1353);
//   compute_omega NH branch (Main/mod_tendency.F90:1158-1191): ucc/vcc, qdot
//             (sigma-dot) and the mass divergence mdv%cr.
// The decoupled winds ud = atm1%u / p*dot never exist in memory: the boundary
// strips wue/wui/... the reference decouples from (Main/mod_tendency.F90:895-994)
// are, for iboudy 0/1/5, copies of atm1%u on the same rows made by bdyuv at
// the end of the previous step (Main/mod_bdycod.F90:905-1028), so
// ud == atm1%u * rpsda on every dot point.
//
// k_xkc is calc_coeff (Main/mod_diffusion.F90:212-230): Smagorinsky
// coefficient on cross points from the level-b decoupled winds; xkd, xkcf and
// the rdxsq*p* scaling (:236-249) are applied by the consumers.
#include "dev_common.cuh"
#include "launch.h"

using namespace rcm;

// k_prep measured 10 % faster with 64-bit element indices, k_xkc 10 % faster with 32-bit ones (dev_common.cuh)
#ifndef RCM_IDX32_ALL   // (tuning builds only)
#undef P3
#undef P2
#define P3(p, j, i, k) P3_64(p, j, i, k)
#define P2(p, j, i) P2_64(p, j, i)
#endif
__global__ void __launch_bounds__(RCM_MAXT, RCM_MB_PREP)
k_prep(const Geom g, const Fields f, const StepScalars s) {
  const int si = g.si;
  const long long sk = g.sk;
  const int j = (g.jce1 - g.gl) + blockIdx.x * blockDim.x + threadIdx.x;
  const int i = (g.ice1 - g.gb) + blockIdx.y * blockDim.y + threadIdx.y;
  if (j > g.jce2 + g.gr || i > g.ice2 + g.gt) return;
  const int kz = g.kz;

  const double rpsa = L2(f.rpsa, j, i), psa = L2(f.psa, j, i), ps0 = L2(f.ps0, j, i);
  const double rdiv_ps0 = recip_for_div(ps0);   // shared-denominator divisions (div_r)
  const double msfx = L2(f.msfx, j, i);
  const double dummy = 1.0 / (s.dx2 * msfx * msfx);
  const double dpsdxm = L2(f.dpsdxm, j, i), dpsdym = L2(f.dpsdym, j, i);
  // dots around the cross point: 00=(j,i) 01=(j,i+1) 10=(j+1,i) 11=(j+1,i+1)
  const double m00 = L2(f.msfd, j, i), m01 = L2(f.msfd, j, i + 1), m10 = L2(f.msfd, j + 1, i),
               m11 = L2(f.msfd, j + 1, i + 1);
  const double r00 = L2(f.rpsda, j, i), r01 = L2(f.rpsda, j, i + 1), r10 = L2(f.rpsda, j + 1, i),
               r11 = L2(f.rpsda, j + 1, i + 1);

  double u00 = L3(f.au, j, i, 1), u01 = L3(f.au, j, i + 1, 1), u10 = L3(f.au, j + 1, i, 1),
         u11 = L3(f.au, j + 1, i + 1, 1);
  double v00 = L3(f.av, j, i, 1), v01 = L3(f.av, j, i + 1, 1), v10 = L3(f.av, j + 1, i, 1),
         v11 = L3(f.av, j + 1, i + 1, 1);
  // iboudy 3, 4: cells that touch the outermost dot ring see the inflow/outflow-adjusted decoupled winds (dwind_io)
  const bool io_ring = (s.iboudy == 3 || s.iboudy == 4) && (j == 1 || j + 1 == g.jx || i == 1 || i + 1 == g.iy);
  auto ucc_of = [&](const double* a, int kk, double a00, double a01, double a10, double a11) {
    // sum of umd = (u*rpsda)*msfd over the 4 dots, reference order
    if (io_ring)
      return dwind_io(g, f, a, j, i, kk) * m00 + dwind_io(g, f, a, j, i + 1, kk) * m01 + dwind_io(g, f, a, j + 1, i, kk) * m10 +
             dwind_io(g, f, a, j + 1, i + 1, kk) * m11;
    return (a00 * r00) * m00 + (a01 * r01) * m01 + (a10 * r10) * m10 + (a11 * r11) * m11;
  };
  double ucc_k = ucc_of(f.au, 1, u00, u01, u10, u11);
  double vcc_k = ucc_of(f.av, 1, v00, v01, v10, v11);
  double qd_k = 0.0;
  P3(f.qdot, j, i, 1) = 0.0;

  // request level kk of every streamed field into L2 (see PF3)
  auto prefetch = [&](int kk) {
    PF3(f.au, j, i, kk); PF3(f.av, j, i, kk); PF3(f.at, j, i, kk); PF3(f.aqv, j, i, kk); PF3(f.app, j, i, kk); PF3(f.aw, j, i, kk); PF3(f.pr0, j, i, kk); PF3(f.rhof0, j, i, kk);
  };
  const int pfd = g.pfd;
  if (pfd > 0)
    for (int kk = 1; kk <= min(pfd, kz); ++kk) prefetch(kk);
  for (int k = 1; k <= kz; ++k) {
    if (pfd > 0 && k + pfd <= kz) prefetch(k + pfd);
    // ---- decouple + theta ----
    const double t = L3(f.at, j, i, k) * rpsa;
    const double qv = fmax(L3(f.aqv, j, i, k) * rpsa, minqq);
    const double tv = t * (1.0 + ep1 * qv);
    const double pp = L3(f.app, j, i, k) * rpsa;
    const double pr1 = L3(f.pr0, j, i, k) + pp;
    P3(f.pr1, j, i, k) = pr1;
    P3(f.rho1, j, i, k) = pr1 / (rgas * tv);
    P3(f.th, j, i, k) = t * pow_pos(p00 / pr1, rovcp);
    // ---- sigma-dot at the next full level, divergence at this half level ----
    double qd_n = 0.0, ucc_n = 0.0, vcc_n = 0.0;
    double n00 = 0, n01 = 0, n10 = 0, n11 = 0, w00 = 0, w01 = 0, w10 = 0, w11 = 0;
    if (k < kz) {
      const int kn = k + 1;
      n00 = L3(f.au, j, i, kn); n01 = L3(f.au, j, i + 1, kn); n10 = L3(f.au, j + 1, i, kn); n11 = L3(f.au, j + 1, i + 1, kn);
      w00 = L3(f.av, j, i, kn); w01 = L3(f.av, j, i + 1, kn); w10 = L3(f.av, j + 1, i, kn); w11 = L3(f.av, j + 1, i + 1, kn);
      ucc_n = ucc_of(f.au, kn, n00, n01, n10, n11);
      vcc_n = ucc_of(f.av, kn, w00, w01, w10, w11);
      const double w = L3(f.aw, j, i, kn) * rpsa;
      const double t1 = LD1(&f.twt1[kn]), t2 = LD1(&f.twt2[kn]);
      qd_n = -div_r(L3(f.rhof0, j, i, kn) * egrav * w, ps0, rdiv_ps0) -
             LD1(&f.sigma[kn]) * (dpsdxm * (t1 * ucc_n + t2 * ucc_k) + dpsdym * (t1 * vcc_n + t2 * vcc_k));
    }
    P3(f.qdot, j, i, k + 1) = qd_n;
    P3(f.cr, j, i, k) = ((u11 * m11 + u10 * m10 - u01 * m01 - u00 * m00) +
                         (v11 * m11 + v01 * m01 - v10 * m10 - v00 * m00)) * dummy +
                        (qd_n - qd_k) * psa / LD1(&f.dsigma[k]);
    u00 = n00; u01 = n01; u10 = n10; u11 = n11;
    v00 = w00; v01 = w01; v10 = w10; v11 = w11;
    ucc_k = ucc_n; vcc_k = vcc_n; qd_k = qd_n;
  }
}

#undef P3
#undef P2
#define P3(p, j, i, k) P3_32(p, j, i, k)
#define P2(p, j, i) P2_32(p, j, i)

// Smagorinsky coefficient on cross points (calc_coeff) and, for diffu_d, the
// level-b winds decoupled and divided by the dot-point map factor,
// (u*(1/psdotb))/msfd (Main/mod_diffusion.F90:266-385 reads ud/mapfd at 9
// stencil points; evaluating the quotient once per dot point here instead of at
// every stencil reference removes 18 divisions per point from k_uv and gives
// the same bits).  1/psdotb and 1/psb are the host tables rpsda, rpsb (p* is
// constant for idynamic = 2).
__global__ void __launch_bounds__(RCM_MAXT, RCM_MB_XKC)
k_xkc(const Geom g, const Fields f, const StepScalars s) {
  const int si = g.si;
  const long long sk = g.sk;
  const int j = (g.jde1 - 2 * g.gl) + blockIdx.x * blockDim.x + threadIdx.x;
  const int i = (g.ide1 - 2 * g.gb) + blockIdx.y * blockDim.y + threadIdx.y;
  const int k = 1 + blockIdx.z;
  if (j > g.jde2 + 2 * g.gr || i > g.ide2 + 2 * g.gt) return;
  // ubd3d = atm2%u * (1/psdotb), wb3d = atm2%w * (1/psb)  (Main/mod_slice.F90:169-263)
  const double u00 = L3(f.bu, j, i, k) * L2(f.rpsdb, j, i);
  const double v00 = L3(f.bv, j, i, k) * L2(f.rpsdb, j, i);
  const double md = L2(f.msfd, j, i);
  const double rmd = recip_for_div(md);
  P3(f.dub, j, i, k) = div_r(u00, md, rmd);
  P3(f.dvb, j, i, k) = div_r(v00, md, rmd);
  if (j < g.jce1 - g.gl || j > g.jce2 + g.gr || i < g.ice1 - g.gb || i > g.ice2 + g.gt) return;
  const double r01 = L2(f.rpsdb, j, i + 1), r10 = L2(f.rpsdb, j + 1, i), r11 = L2(f.rpsdb, j + 1, i + 1);
  const double rpsb = L2(f.rpsb, j, i);
  const double u01 = L3(f.bu, j, i + 1, k) * r01, u10 = L3(f.bu, j + 1, i, k) * r10, u11 = L3(f.bu, j + 1, i + 1, k) * r11;
  const double v01 = L3(f.bv, j, i + 1, k) * r01, v10 = L3(f.bv, j + 1, i, k) * r10, v11 = L3(f.bv, j + 1, i + 1, k) * r11;
  const double dudx = u10 + u11 - u00 - u01;
  const double dvdx = v10 + v11 - v00 - v01;
  const double dudy = u01 + u11 - u00 - u10;
  const double dvdy = v01 + v11 - v00 - v10;
  // hydrostatic core: no dw/dz term (Main/mod_diffusion.F90:193-211)
  const double dwdz = s.hydro ? 0.0 : L3(f.bw, j, i, k) * rpsb - L3(f.bw, j, i, k + 1) * rpsb;
  const double duv = sqrt(fmax((dudx - dvdy) * (dudx - dvdy) + (dvdx + dudy) * (dvdx + dudy) - dwdz * dwdz, 0.0));
  P3(f.xkc, j, i, k) = fmin((L2(f.hgfact, j, i) + s.dydc * duv), s.xkhmax);
}

void launch_prep(const Geom& g, const Fields& f, const StepScalars& s, cudaStream_t st) {
  const int nj = (g.jce2 + g.gr) - (g.jce1 - g.gl) + 1, ni = (g.ice2 + g.gt) - (g.ice1 - g.gb) + 1;
  const dim3 b = rcm::col_block(); dim3 gr((nj + b.x - 1) / b.x, (ni + b.y - 1) / b.y);
  RCM_LAUNCH(k_prep, gr, b, st, g, f, s);
}
void launch_xkc(const Geom& g, const Fields& f, const StepScalars& s, cudaStream_t st) {
  const int nj = (g.jde2 + 2 * g.gr) - (g.jde1 - 2 * g.gl) + 1, ni = (g.ide2 + 2 * g.gt) - (g.ide1 - 2 * g.gb) + 1;
  const dim3 b = rcm::col_block(); dim3 gr3((nj + b.x - 1) / b.x, (ni + b.y - 1) / b.y, g.kz);
  RCM_LAUNCH(k_xkc, gr3, b, st, g, f, s);
}
