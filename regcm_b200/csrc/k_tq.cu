// K_scalar_slow<T,qv,qc> (SURVEY.md §8 a10,a11,a13,a15,a12,a16,a17,a18): the
// complete slow tendency, leapfrog forecast and Asselin filter of p*T, p*qv,
// p*qc on one cross-point column, marched in k.  Fuses, in the reference's
// order of accumulation:
//   start_advect                         Main/mod_advection.F90:111-120
//   hadvt(theta) + limiter               Main/mod_advection.F90:311-393
//   vadv3d(theta*p*, kz, ind=0)          Main/mod_advection.F90:746-754
//   hadvqv + limiter, vadvqv             Main/mod_advection.F90:517-603,811-836
//   hadvqx, vadv4d(ind=1)                Main/mod_advection.F90:607-662,873-894
//   adiabatic (ithadv=1)                 Main/mod_tendency.F90:1595-1617
//   nudge3d / nudge4d3d                  Main/mod_bdycod.F90:4218-4406,3206-3392
//   diffu_x3d / diffu_x4d                Main/mod_diffusion.F90:658-725,792-880
//   raydamp3 / raydampqv                 Main/mod_bdycod.F90:4953-5085
//   forecast + negative-q guard          Main/mod_tendency.F90:368-393
//   timefilter RA(t), RAW(qv), RAW(qc)   Main/mod_timefilter.F90:92-110,231-278
// The filtered levels a and b are written to the double-buffered "next"
// copies (nat/nbt, ...) because the stencils above still read the current
// level-b neighbours; the buffers are swapped at the end of the step.
#include "dev_common.cuh"
#include "launch.h"

using namespace rcm;

__global__ void __launch_bounds__(RCM_MAXT, RCM_MB_TQ)
k_tq(const Geom g, const Fields f, const StepScalars s) {
  const int si = g.si;
  const long long sk = g.sk;
  const int j = g.jci1 + blockIdx.x * blockDim.x + threadIdx.x;
  const int i = g.ici1 + blockIdx.y * blockDim.y + threadIdx.y;
  if (j > g.jci2 || i > g.ici2) return;
  const int kz = g.kz;
  const bool up = s.upstream_mode != 0;
  const bool se = s.stability_enhance != 0;
  const double ul = s.ul, xt = s.xt, dt = s.dt;

  const double psa = L2(f.psa, j, i), psb = L2(f.psb, j, i);
  const double rpsa_c = L2(f.rpsa, j, i), rpsa_jm = L2(f.rpsa, j - 1, i), rpsa_jp = L2(f.rpsa, j + 1, i),
               rpsa_im = L2(f.rpsa, j, i - 1), rpsa_ip = L2(f.rpsa, j, i + 1);
  const double xmsf = L2(f.xmsf, j, i);
  const double m00 = L2(f.msfd, j, i), m01 = L2(f.msfd, j, i + 1), m10 = L2(f.msfd, j + 1, i), m11 = L2(f.msfd, j + 1, i + 1);
  const int ib = (s.iboudy == 1 || s.iboudy == 5) ? L2(f.zone_c, j, i) : 0;
  const bool dint = diff_interior(g, j, i, false);
  const int nring = ring_count(g, j, i, false);
  const double qthr = minqq * psa;           // vadvqv threshold
  const double cthr = minqq * minqq * psa;   // vadv4d ind=1 threshold
  const double beta = 0.53;

  // vertical fluxes at full level k (zero at k = 1)
  double fx_th_k = 0.0, fl_qv_k = 0.0, fl_qc_k = 0.0;
  double tha_k = L3(f.th, j, i, 1) * psa;

  // request level kk of every streamed field into L2 (see PF3)
  auto prefetch = [&](int kk) {
    PF3(f.au, j, i, kk); PF3(f.av, j, i, kk); PF3(f.cr, j, i, kk); PF3(f.qdot, j, i, kk); PF3(f.xkc, j, i, kk); PF3(f.th, j, i, kk); PF3(f.at, j, i, kk); PF3(f.bt, j, i, kk); PF3(f.aqv, j, i, kk); PF3(f.bqv, j, i, kk); PF3(f.aqc, j, i, kk); PF3(f.bqc, j, i, kk);
  };
  const int pfd = g.pfd;
  if (pfd > 0)
    for (int kk = 1; kk <= min(pfd, kz); ++kk) prefetch(kk);
  for (int k = 1; k <= kz; ++k) {
    if (pfd > 0 && k + pfd <= kz) prefetch(k + pfd);
    // ---- start_advect: ua = umc = u*msfd at the four surrounding dots ----
    const double ua00 = L3(f.au, j, i, k) * m00, ua01 = L3(f.au, j, i + 1, k) * m01, ua10 = L3(f.au, j + 1, i, k) * m10,
                 ua11 = L3(f.au, j + 1, i + 1, k) * m11;
    const double va00 = L3(f.av, j, i, k) * m00, va01 = L3(f.av, j, i + 1, k) * m01, va10 = L3(f.av, j + 1, i, k) * m10,
                 va11 = L3(f.av, j + 1, i + 1, k) * m11;
    const double uavg1 = ua01 + ua00, uavg2 = ua11 + ua10;
    const double vavg1 = va10 + va00, vavg2 = va11 + va01;
    const double f1 = 0.5 * ul * (uavg2 + uavg1) / psa;
    const double f2 = 0.5 * ul * (vavg2 + vavg1) / psa;
    const double cr = L3(f.cr, j, i, k);
    const double xdsk = LD1(&f.xds[k]);
    const double svv_n = (k < kz) ? L3(f.qdot, j, i, k + 1) : 0.0;
    const int kk = k;  // min(k,kz) for the nudging table
    double xf = 0.0, xg = 0.0;
    if (ib > 0) { xf = LD1(&f.hefc[ib + s.nspgx * (kk - 1)]); xg = LD1(&f.hegc[ib + s.nspgx * (kk - 1)]); }
    double xkc = L3(f.xkc, j, i, k);
    xkc = xkc * s.rdxsq * psb;
    const double tauc = (s.ifrayd == 1 && k <= s.kt_ray) ? L3(f.tau_c, j, i, k) : 0.0;
    const bool ray = (s.ifrayd == 1 && k <= s.kt_ray);

    // =================== temperature ===================
    {
      const double th_c = L3(f.th, j, i, k), th_jm = L3(f.th, j - 1, i, k), th_jp = L3(f.th, j + 1, i, k),
                   th_im = L3(f.th, j, i - 1, k), th_ip = L3(f.th, j, i + 1, k);
      double fg = hflux(up, f1, f2, xmsf, uavg1, uavg2, vavg1, vavg2, th_c, th_jm, th_jp, th_im, th_ip);
      if (se) fg = extrema_limit(fg, th_c, th_jm, th_jp, th_im, th_ip, psa, s.t_extrema);
      double thten = fg;
      if (k >= 2) thten = thten + fx_th_k * xdsk;
      double fx_th_n = 0.0, tha_n = 0.0;
      if (k < kz) {
        tha_n = L3(f.th, j, i, k + 1) * psa;
        fx_th_n = svv_n * (LD1(&f.twt1[k + 1]) * tha_n + LD1(&f.twt2[k + 1]) * tha_k);
        thten = thten - fx_th_n * xdsk;
      }
      thten = thten + th_c * cr;
      const double at = L3(f.at, j, i, k);
      double ten = at * thten / tha_k;
      const double bt_c = L3(f.bt, j, i, k);
      const double bt_jm = L3(f.bt, j - 1, i, k), bt_jp = L3(f.bt, j + 1, i, k), bt_im = L3(f.bt, j, i - 1, k), bt_ip = L3(f.bt, j, i + 1, k);
      if (ib > 0) {
#define FG(jj, ii, bval) ((L3(f.xtb0, jj, ii, k) + xt * L3(f.xtbt, jj, ii, k)) - (bval))
        const double fls0 = FG(j, i, bt_c), fls1 = FG(j - 1, i, bt_jm), fls2 = FG(j + 1, i, bt_jp), fls3 = FG(j, i - 1, bt_im), fls4 = FG(j, i + 1, bt_ip);
#undef FG
        ten = ten + xf * fls0 - xg * (fls1 + fls2 + fls3 + fls4 - 4.0 * fls0);
      }
      {
        const double d_c = bt_c * L2(f.rpsb, j, i), d_jm = bt_jm * L2(f.rpsb, j - 1, i), d_jp = bt_jp * L2(f.rpsb, j + 1, i),
                     d_im = bt_im * L2(f.rpsb, j, i - 1), d_ip = bt_ip * L2(f.rpsb, j, i + 1);
        if (dint) {
          const double d_jp2 = L3(f.bt, j + 2, i, k) * L2(f.rpsb, j + 2, i), d_jm2 = L3(f.bt, j - 2, i, k) * L2(f.rpsb, j - 2, i);
          const double d_ip2 = L3(f.bt, j, i + 2, k) * L2(f.rpsb, j, i + 2), d_im2 = L3(f.bt, j, i - 2, k) * L2(f.rpsb, j, i - 2);
          ten = ten - xkc * (1.0 * (d_jp2 + d_jm2 + d_ip2 + d_im2) + (-4.0) * (d_jp + d_jm + d_ip + d_im) + 12.0 * (d_c));
        }
        for (int r = 0; r < nring; ++r) ten = ten + xkc * (1.0 * (d_jp + d_jm + d_ip + d_im) + (-4.0) * (d_c));
      }
      if (ray) ten = ten + tauc * ((L3(f.xtb0, j, i, k) + xt * L3(f.xtbt, j, i, k)) - bt_c);
      const double ct = bt_c + dt * ten;
      const double d = s.gnu1 * (ct + bt_c - 2.0 * at);
      P3(f.nbt, j, i, k) = at + d;
      P3(f.nat, j, i, k) = ct;
      fx_th_k = fx_th_n; tha_k = tha_n;
    }

    // =================== water vapour ===================
    {
      const double aq_c = L3(f.aqv, j, i, k);
      const double q_c = fmax(aq_c * rpsa_c, minqq), q_jm = fmax(L3(f.aqv, j - 1, i, k) * rpsa_jm, minqq),
                   q_jp = fmax(L3(f.aqv, j + 1, i, k) * rpsa_jp, minqq), q_im = fmax(L3(f.aqv, j, i - 1, k) * rpsa_im, minqq),
                   q_ip = fmax(L3(f.aqv, j, i + 1, k) * rpsa_ip, minqq);
      double fg = hflux(up, f1, f2, xmsf, uavg1, uavg2, vavg1, vavg2, q_c, q_jm, q_jp, q_im, q_ip);
      if (se) fg = extrema_limit(fg, q_c, q_jm, q_jp, q_im, q_ip, fmax(q_c, dlowval), s.q_rel_extrema);
      double ten = fg;
      if (k >= 2) ten = ten + fl_qv_k * xdsk;
      double fl_n = 0.0, aqv_n = 0.0;
      if (k < kz) {
        aqv_n = L3(f.aqv, j, i, k + 1);
        double fgv = 0.0;
        if (aqv_n > qthr && aq_c > qthr) fgv = aqv_n * pow(aq_c / aqv_n, LD1(&f.qcon[k + 1]));
        fl_n = svv_n * fgv;
        ten = ten - fl_n * xdsk;
      }
      ten = ten + q_c * cr;
      const double b_c = L3(f.bqv, j, i, k);
      const double b_jm = L3(f.bqv, j - 1, i, k), b_jp = L3(f.bqv, j + 1, i, k), b_im = L3(f.bqv, j, i - 1, k), b_ip = L3(f.bqv, j, i + 1, k);
      if (ib > 0) {
        const double nfac = 1.0e3, rfac = 1.0 / nfac;
#define FG(jj, ii, bval) (nfac * (L3(f.xqb0, jj, ii, k) + xt * L3(f.xqbt, jj, ii, k)) - nfac * (bval))
        const double fls0 = FG(j, i, b_c), fls1 = FG(j - 1, i, b_jm), fls2 = FG(j + 1, i, b_jp), fls3 = FG(j, i - 1, b_im), fls4 = FG(j, i + 1, b_ip);
#undef FG
        ten = ten + rfac * (xf * fls0 - xg * (fls1 + fls2 + fls3 + fls4 - 4.0 * fls0));
      }
      {
        const double d_c = fmax(b_c * L2(f.rpsb, j, i), minqq), d_jm = fmax(b_jm * L2(f.rpsb, j - 1, i), minqq),
                     d_jp = fmax(b_jp * L2(f.rpsb, j + 1, i), minqq), d_im = fmax(b_im * L2(f.rpsb, j, i - 1), minqq),
                     d_ip = fmax(b_ip * L2(f.rpsb, j, i + 1), minqq);
        if (dint) {
          const double d_jp2 = fmax(L3(f.bqv, j + 2, i, k) * L2(f.rpsb, j + 2, i), minqq), d_jm2 = fmax(L3(f.bqv, j - 2, i, k) * L2(f.rpsb, j - 2, i), minqq);
          const double d_ip2 = fmax(L3(f.bqv, j, i + 2, k) * L2(f.rpsb, j, i + 2), minqq), d_im2 = fmax(L3(f.bqv, j, i - 2, k) * L2(f.rpsb, j, i - 2), minqq);
          ten = ten - 1.0 * xkc * (1.0 * (d_jp2 + d_jm2 + d_ip2 + d_im2) + (-4.0) * (d_jp + d_jm + d_ip + d_im) + 12.0 * d_c);
        }
        for (int r = 0; r < nring; ++r) ten = ten + 1.0 * xkc * (1.0 * (d_jp + d_jm + d_ip + d_im) + (-4.0) * d_c);
      }
      if (ray) ten = ten + tauc * ((L3(f.xqb0, j, i, k) + xt * L3(f.xqbt, j, i, k)) - b_c);
      double cq = b_c + dt * ten;
      if (cq < 0.0) {  // Main/mod_tendency.F90:382-393 (clipped, counted)
        atomicAdd(f.negq, 1ULL);
        cq = 0.0;
      }
      const double d = s.gnu1 * (cq + b_c - 2.0 * aq_c);
      P3(f.nbqv, j, i, k) = fmax(aq_c + beta * d, minqq * psa);
      P3(f.naqv, j, i, k) = fmax(cq + (beta - 1.0) * d, minqq * psb);
      fl_qv_k = fl_n;
    }

    // =================== cloud water ===================
    {
      const double aq_c = L3(f.aqc, j, i, k);
      const double q_c = fmax(aq_c * rpsa_c, 0.0);
      double ten = 0.0;
      double fl_n = 0.0, aqc_n = 0.0;
      if (k < kz) aqc_n = L3(f.aqc, j, i, k + 1);
      if (s.ipptls > 0) {
        const double q_jm = fmax(L3(f.aqc, j - 1, i, k) * rpsa_jm, 0.0), q_jp = fmax(L3(f.aqc, j + 1, i, k) * rpsa_jp, 0.0),
                     q_im = fmax(L3(f.aqc, j, i - 1, k) * rpsa_im, 0.0), q_ip = fmax(L3(f.aqc, j, i + 1, k) * rpsa_ip, 0.0);
        ten = hflux(up, f1, f2, xmsf, uavg1, uavg2, vavg1, vavg2, q_c, q_jm, q_jp, q_im, q_ip);
        if (k >= 2) ten = ten + fl_qc_k * xdsk;
        if (k < kz) {
          const double t1 = LD1(&f.twt1[k + 1]), t2 = LD1(&f.twt2[k + 1]);
          if (svv_n > 0.0) {
            if (aq_c > cthr) fl_n = svv_n * (t1 * aqc_n + t2 * aq_c);
          } else {
            if (aqc_n > cthr) fl_n = svv_n * (t1 * aqc_n + t2 * aq_c);
          }
          ten = ten - fl_n * xdsk;
        }
      }
      ten = ten + q_c * cr;
      const double b_c = L3(f.bqc, j, i, k);
      {
        const double d_c = fmax(b_c * L2(f.rpsb, j, i), 0.0), d_jm = fmax(L3(f.bqc, j - 1, i, k) * L2(f.rpsb, j - 1, i), 0.0),
                     d_jp = fmax(L3(f.bqc, j + 1, i, k) * L2(f.rpsb, j + 1, i), 0.0), d_im = fmax(L3(f.bqc, j, i - 1, k) * L2(f.rpsb, j, i - 1), 0.0),
                     d_ip = fmax(L3(f.bqc, j, i + 1, k) * L2(f.rpsb, j, i + 1), 0.0);
        if (dint) {
          const double d_jp2 = fmax(L3(f.bqc, j + 2, i, k) * L2(f.rpsb, j + 2, i), 0.0), d_jm2 = fmax(L3(f.bqc, j - 2, i, k) * L2(f.rpsb, j - 2, i), 0.0);
          const double d_ip2 = fmax(L3(f.bqc, j, i + 2, k) * L2(f.rpsb, j, i + 2), 0.0), d_im2 = fmax(L3(f.bqc, j, i - 2, k) * L2(f.rpsb, j, i - 2), 0.0);
          ten = ten - 1.0 * xkc * (1.0 * (d_jp2 + d_jm2 + d_ip2 + d_im2) + (-4.0) * (d_jp + d_jm + d_ip + d_im) + 12.0 * d_c);
        }
        for (int r = 0; r < nring; ++r) ten = ten + 1.0 * xkc * (1.0 * (d_jp + d_jm + d_ip + d_im) + (-4.0) * d_c);
      }
      if (s.ipptls <= 0) ten = 0.0;  // Main/mod_tendency.F90:331-335
      double cq = b_c + dt * ten;
      if (cq < 0.0) {
        atomicAdd(f.negq, 1ULL);
        cq = 0.0;
      }
      const double d = s.gnu2 * (cq + b_c - 2.0 * aq_c);
      double nb = aq_c + beta * d;
      double na = cq + (beta - 1.0) * d;
      if (nb < 0.0) nb = 0.0;
      if (na < 0.0) na = 0.0;
      P3(f.nbqc, j, i, k) = nb;
      P3(f.naqc, j, i, k) = na;
      fl_qc_k = fl_n;
    }
  }
}

void launch_tq(const Geom& g, const Fields& f, const StepScalars& s, cudaStream_t st) {
  const int nj = g.jci2 - g.jci1 + 1, ni = g.ici2 - g.ici1 + 1;
  const dim3 b = rcm::col_block(); dim3 gr((nj + b.x - 1) / b.x, (ni + b.y - 1) / b.y);
  RCM_LAUNCH(k_tq, gr, b, st, g, f, s);
}
