// K_scalar_slow<pp,w> (SURVEY.md §8 a10,a11,a13,a15,a12,a16): complete slow
// tendencies of p*p' (half levels 1..kz) and p*w (full levels 1..kz+1) on one
// cross-point column marched in k, handed to the acoustic loop decoupled and
// premultiplied by dts.  Fuses, in the reference's order of accumulation:
//   hadv3d(pp,ind=0), vadv3d(pp,kz,0)    Main/mod_advection.F90:395-480,746-754
//   hadv3d(w,ind=1),  vadv3d(w,kzp1,0)   Main/mod_advection.F90:481-507,756-765
//   adiabatic: pp*div; w buoyancy, vertical Coriolis, curvature, w*div, water
//              loading                    Main/mod_tendency.F90:1612-1671
//   nudge3d(pp), nudge3d(w)              Main/mod_bdycod.F90:4218-4406
//   diffu_x3d(pp), diffu_x3df(w)         Main/mod_diffusion.F90:658-725,523-591
//   raydamp3(pp), raydamp3f(w)           Main/mod_bdycod.F90:4953-5085
//   *rpsa, *dts                          Main/mod_tendency.F90:488-499, mod_sound.F90:238-246
#include "dev_common.cuh"
#include "launch.h"

using namespace rcm;

__global__ void __launch_bounds__(RCM_MAXT, RCM_MB_PPW)
k_ppw(const Geom g, const Fields f, const StepScalars s) {
  const int si = g.si;
  const long long sk = g.sk;
  const int j = g.jci1 + blockIdx.x * blockDim.x + threadIdx.x;
  const int i = g.ici1 + blockIdx.y * blockDim.y + threadIdx.y;
  if (j > g.jci2 || i > g.ici2) return;
  const int kz = g.kz, kzp1 = g.kzp1;
  const bool up = s.upstream_mode != 0;
  const double ul = s.ul, xt = s.xt;

  const double psa = L2(f.psa, j, i), psb = L2(f.psb, j, i);
  const double rpsa_c = L2(f.rpsa, j, i), rpsa_jm = L2(f.rpsa, j - 1, i), rpsa_jp = L2(f.rpsa, j + 1, i),
               rpsa_im = L2(f.rpsa, j, i - 1), rpsa_ip = L2(f.rpsa, j, i + 1);
  const double rpsb_c = L2(f.rpsb, j, i), rpsb_jm = L2(f.rpsb, j - 1, i), rpsb_jp = L2(f.rpsb, j + 1, i),
               rpsb_im = L2(f.rpsb, j, i - 1), rpsb_ip = L2(f.rpsb, j, i + 1);
  const double xmsf = L2(f.xmsf, j, i);
  const double m00 = L2(f.msfd, j, i), m01 = L2(f.msfd, j, i + 1), m10 = L2(f.msfd, j + 1, i), m11 = L2(f.msfd, j + 1, i + 1);
  const double ex = L2(f.ex, j, i), crx = L2(f.crx, j, i), cry = L2(f.cry, j, i);
  const int ib = (s.iboudy == 1 || s.iboudy == 5) ? L2(f.zone_c, j, i) : 0;
  const bool dint = diff_interior(g, j, i, false);
  const int nring = ring_count(g, j, i, false);

  // ---- carried state (level k-1 values and interface fluxes) ----
  double uavg1_m = 0, uavg2_m = 0, vavg1_m = 0, vavg2_m = 0;  // uavg at k-1
  double ucc_m = 0, vcc_m = 0;                                // raw 4-dot sums at k-1
  double axpr_m = 0, rho0_m = 0, rho1_m = 0, qc_m = 0, cr_m = 0;
  double app_m = 0;        // atm1%pp(k-1)
  double fx_pp_k = 0.0;    // vertical pp flux at full level k
  double fxw_m = 0.0;      // vertical w flux of loop index k-1
  double xkc_m = 0.0;      // raw xkc(k-1)
  double aw_k = L3(f.aw, j, i, 1);

  // request level kk of every streamed field into L2 (see PF3)
  auto prefetch = [&](int kk) {
    PF3(f.au, j, i, kk); PF3(f.av, j, i, kk); PF3(f.aw, j, i, kk); PF3(f.app, j, i, kk); PF3(f.bpp, j, i, kk); PF3(f.bw, j, i, kk); PF3(f.qdot, j, i, kk); PF3(f.cr, j, i, kk); PF3(f.xkc, j, i, kk); PF3(f.at, j, i, kk); PF3(f.aqv, j, i, kk); PF3(f.aqc, j, i, kk); PF3(f.rho0, j, i, kk); PF3(f.rho1, j, i, kk); PF3(f.t0, j, i, kk);
  };
  const int pfd = g.pfd;
  if (pfd > 0)
    for (int kk = 1; kk <= min(pfd, kz); ++kk) prefetch(kk);
  for (int k = 1; k <= kzp1; ++k) {
    if (pfd > 0 && k + pfd <= kz) prefetch(k + pfd);
    const bool half = (k <= kz);  // a half level k exists (pp, T, u, v live there)
    double uavg1 = 0, uavg2 = 0, vavg1 = 0, vavg2 = 0, ucc = 0, vcc = 0, f1 = 0, f2 = 0;
    double cr = 0, xkc_raw = 0;
    if (half) {
      const double u00 = L3(f.au, j, i, k), u01 = L3(f.au, j, i + 1, k), u10 = L3(f.au, j + 1, i, k), u11 = L3(f.au, j + 1, i + 1, k);
      const double v00 = L3(f.av, j, i, k), v01 = L3(f.av, j, i + 1, k), v10 = L3(f.av, j + 1, i, k), v11 = L3(f.av, j + 1, i + 1, k);
      const double ua00 = u00 * m00, ua01 = u01 * m01, ua10 = u10 * m10, ua11 = u11 * m11;
      const double va00 = v00 * m00, va01 = v01 * m01, va10 = v10 * m10, va11 = v11 * m11;
      uavg1 = ua01 + ua00; uavg2 = ua11 + ua10;
      vavg1 = va10 + va00; vavg2 = va11 + va01;
      f1 = 0.5 * ul * (uavg2 + uavg1) / psa;
      f2 = 0.5 * ul * (vavg2 + vavg1) / psa;
      // adiabatic's ucc/vcc are sums of the COUPLED winds (atmx%uc), :1630-1638
      ucc = (u00 + u01 + u10 + u11);
      vcc = (v00 + v01 + v10 + v11);
      cr = L3(f.cr, j, i, k);
      xkc_raw = L3(f.xkc, j, i, k);
    }
    const int kk = half ? k : kz;  // min(k,kz)
    double xf = 0.0, xg = 0.0;
    if (ib > 0) { xf = LD1(&f.hefc[ib + s.nspgx * (kk - 1)]); xg = LD1(&f.hegc[ib + s.nspgx * (kk - 1)]); }
    const double qdot_k = L3(f.qdot, j, i, k);
    const double qdot_n = half ? L3(f.qdot, j, i, k + 1) : 0.0;

    // =================== pressure perturbation (half level k) ===================
    double app_c = 0.0;
    if (half) {
      app_c = L3(f.app, j, i, k);
      const double p_c = app_c * rpsa_c, p_jm = L3(f.app, j - 1, i, k) * rpsa_jm, p_jp = L3(f.app, j + 1, i, k) * rpsa_jp,
                   p_im = L3(f.app, j, i - 1, k) * rpsa_im, p_ip = L3(f.app, j, i + 1, k) * rpsa_ip;
      double ten = hflux(up, f1, f2, xmsf, uavg1, uavg2, vavg1, vavg2, p_c, p_jm, p_jp, p_im, p_ip);
      const double xdsk = LD1(&f.xds[k]);
      if (k >= 2) ten = ten + fx_pp_k * xdsk;
      double fx_n = 0.0;
      if (k < kz) {
        fx_n = qdot_n * (LD1(&f.twt1[k + 1]) * L3(f.app, j, i, k + 1) + LD1(&f.twt2[k + 1]) * app_c);
        ten = ten - fx_n * xdsk;
      }
      fx_pp_k = fx_n;
      ten = ten + p_c * cr;
      const double b_c = L3(f.bpp, j, i, k);
      const double b_jm = L3(f.bpp, j - 1, i, k), b_jp = L3(f.bpp, j + 1, i, k), b_im = L3(f.bpp, j, i - 1, k), b_ip = L3(f.bpp, j, i + 1, k);
      if (ib > 0) {
#define FG(jj, ii, bval) ((L3(f.xpb0, jj, ii, k) + xt * L3(f.xpbt, jj, ii, k)) - (bval))
        const double fls0 = FG(j, i, b_c), fls1 = FG(j - 1, i, b_jm), fls2 = FG(j + 1, i, b_jp), fls3 = FG(j, i - 1, b_im), fls4 = FG(j, i + 1, b_ip);
#undef FG
        ten = ten + xf * fls0 - xg * (fls1 + fls2 + fls3 + fls4 - 4.0 * fls0);
      }
      {
        const double xkc = xkc_raw * s.rdxsq * psb;
        const double d_c = b_c * rpsb_c, d_jm = b_jm * rpsb_jm, d_jp = b_jp * rpsb_jp, d_im = b_im * rpsb_im, d_ip = b_ip * rpsb_ip;
        if (dint) {
          const double d_jp2 = L3(f.bpp, j + 2, i, k) * L2(f.rpsb, j + 2, i), d_jm2 = L3(f.bpp, j - 2, i, k) * L2(f.rpsb, j - 2, i);
          const double d_ip2 = L3(f.bpp, j, i + 2, k) * L2(f.rpsb, j, i + 2), d_im2 = L3(f.bpp, j, i - 2, k) * L2(f.rpsb, j, i - 2);
          ten = ten - xkc * (1.0 * (d_jp2 + d_jm2 + d_ip2 + d_im2) + (-4.0) * (d_jp + d_jm + d_ip + d_im) + 12.0 * (d_c));
        }
        for (int r = 0; r < nring; ++r) ten = ten + xkc * (1.0 * (d_jp + d_jm + d_ip + d_im) + (-4.0) * (d_c));
      }
      if (s.ifrayd == 1 && k <= s.kt_ray)
        ten = ten + L3(f.tau_c, j, i, k) * ((L3(f.xpb0, j, i, k) + xt * L3(f.xpbt, j, i, k)) - b_c);
      P3(f.ppten, j, i, k) = (ten * rpsa_c) * s.dts;
    }

    // =================== vertical velocity (full level k) ===================
    {
      // decoupled state needed by the buoyancy term at half level k
      double axpr = 0, rho0 = 0, rho1 = 0, qc = 0;
      if (half) {
        const double t = L3(f.at, j, i, k) * rpsa_c;
        const double qv = fmax(L3(f.aqv, j, i, k) * rpsa_c, minqq);
        const double tv = t * (1.0 + ep1 * qv);
        const double pp = app_c * rpsa_c;
        rho0 = L3(f.rho0, j, i, k);
        rho1 = L3(f.rho1, j, i, k);
        axpr = (tv - L3(f.t0, j, i, k) - pp / (cpd * rho0)) / t;
        qc = fmax(L3(f.aqc, j, i, k) * rpsa_c, 0.0);
      }
      const double aw_n = half ? L3(f.aw, j, i, k + 1) : 0.0;
      double ten = 0.0;
      const double w_c = aw_k * rpsa_c;
      if (k >= 2 && k <= kz) {
        const double t1 = LD1(&f.twt1[k]), t2 = LD1(&f.twt2[k]);
        const double uaz1 = (t1 * uavg1 + t2 * uavg1_m);
        const double uaz2 = (t1 * uavg2 + t2 * uavg2_m);
        const double vaz1 = (t1 * vavg1 + t2 * vavg1_m);
        const double vaz2 = (t1 * vavg2 + t2 * vavg2_m);
        const double w_jm = L3(f.aw, j - 1, i, k) * rpsa_jm, w_jp = L3(f.aw, j + 1, i, k) * rpsa_jp,
                     w_im = L3(f.aw, j, i - 1, k) * rpsa_im, w_ip = L3(f.aw, j, i + 1, k) * rpsa_ip;
        double fx1, fx2, fy1, fy2;
        if (up) {
          fx1 = (1.0 + f1) * w_jm + (1.0 - f1) * w_c;
          fx2 = (1.0 + f1) * w_c + (1.0 - f1) * w_jp;
          fy1 = (1.0 + f2) * w_im + (1.0 - f2) * w_c;
          fy2 = (1.0 + f2) * w_c + (1.0 - f2) * w_ip;
        } else {
          fx1 = w_jm + w_c; fx2 = w_c + w_jp; fy1 = w_im + w_c; fy2 = w_c + w_ip;
        }
        ten = ten - xmsf * (uaz2 * fx2 - uaz1 * fx1 + vaz2 * fy2 - vaz1 * fy1);
      }
      // vadv3d nk = kzp1: loop index k-1 adds, loop index k subtracts
      const double ddsk = LD1(&f.dds[k]);
      if (k >= 2) ten = ten + fxw_m * ddsk;
      double fxw = 0.0;
      if (half) {
        const double qq = 0.5 * (qdot_k + qdot_n);
        fxw = qq * ((aw_k + aw_n));
        ten = ten - fxw * ddsk;
      }
      if (k >= 2 && k <= kz) {
        const double t1 = LD1(&f.twt1[k]), t2 = LD1(&f.twt2[k]);
        const double dsk = LD1(&f.dsigma[k]), dsm = LD1(&f.dsigma[k - 1]);
        const double rofac = (dsm * rho0 + dsk * rho0_m) / (dsm * rho1 + dsk * rho1_m);
        const double uaq = 0.25 * (t1 * ucc + t2 * ucc_m);
        const double vaq = 0.25 * (t1 * vcc + t2 * vcc_m);
        ten = ten + (t2 * axpr_m + t1 * axpr) * rofac * egrav * psa + ex * (uaq * crx - vaq * cry) +
              (uaq * uaq + vaq * vaq) * rearthrad * rpsa_c + w_c * (t1 * cr + t2 * cr_m);
        if (s.ipptls > 0) ten = ten - egrav * psa * (t2 * qc_m + t1 * qc);
      }
      const double b_c = L3(f.bw, j, i, k);
      const double b_jm = L3(f.bw, j - 1, i, k), b_jp = L3(f.bw, j + 1, i, k), b_im = L3(f.bw, j, i - 1, k), b_ip = L3(f.bw, j, i + 1, k);
      if (ib > 0) {
#define FG(jj, ii, bval) ((L3(f.xwb0, jj, ii, k) + xt * L3(f.xwbt, jj, ii, k)) - (bval))
        const double fls0 = FG(j, i, b_c), fls1 = FG(j - 1, i, b_jm), fls2 = FG(j + 1, i, b_jp), fls3 = FG(j, i - 1, b_im), fls4 = FG(j, i + 1, b_ip);
#undef FG
        ten = ten + xf * fls0 - xg * (fls1 + fls2 + fls3 + fls4 - 4.0 * fls0);
      }
      {
        // xkcf(1) = xkc(1), xkcf(k+1) = xkc(k)   (Main/mod_diffusion.F90:231-235)
        const double xkcf = ((k == 1) ? xkc_raw : xkc_m) * s.rdxsq * psb;
        const double d_c = b_c * rpsb_c, d_jm = b_jm * rpsb_jm, d_jp = b_jp * rpsb_jp, d_im = b_im * rpsb_im, d_ip = b_ip * rpsb_ip;
        if (dint) {
          const double d_jp2 = L3(f.bw, j + 2, i, k) * L2(f.rpsb, j + 2, i), d_jm2 = L3(f.bw, j - 2, i, k) * L2(f.rpsb, j - 2, i);
          const double d_ip2 = L3(f.bw, j, i + 2, k) * L2(f.rpsb, j, i + 2), d_im2 = L3(f.bw, j, i - 2, k) * L2(f.rpsb, j, i - 2);
          ten = ten - 1.0 * xkcf * (1.0 * (d_jp2 + d_jm2 + d_ip2 + d_im2) + (-4.0) * (d_jp + d_jm + d_ip + d_im) + 12.0 * (d_c));
        }
        for (int r = 0; r < nring; ++r) ten = ten + 1.0 * xkcf * (1.0 * (d_jp + d_jm + d_ip + d_im) + (-4.0) * (d_c));
      }
      if (s.ifrayd == 1 && k <= s.ktf_ray) ten = ten + L3(f.tau_f, j, i, k) * (0.0 - b_c);
      P3(f.wten, j, i, k) = (ten * rpsa_c) * s.dts;

      // shift the carried state
      uavg1_m = uavg1; uavg2_m = uavg2; vavg1_m = vavg1; vavg2_m = vavg2;
      ucc_m = ucc; vcc_m = vcc;
      axpr_m = axpr; rho0_m = rho0; rho1_m = rho1; qc_m = qc; cr_m = cr;
      fxw_m = fxw; xkc_m = xkc_raw; aw_k = aw_n;
      app_m = app_c;
    }
  }
  (void)app_m;
}

void launch_ppw(const Geom& g, const Fields& f, const StepScalars& s, cudaStream_t st) {
  const int nj = g.jci2 - g.jci1 + 1, ni = g.ici2 - g.ici1 + 1;
  const dim3 b = rcm::col_block(); dim3 gr((nj + b.x - 1) / b.x, (ni + b.y - 1) / b.y);
  RCM_LAUNCH(k_ppw, gr, b, st, g, f, s);
}
