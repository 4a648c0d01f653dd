// K_mom_slow (SURVEY.md §8 a8,a9,a14,a15,a12,a16): the complete slow tendency
// of the coupled horizontal momentum p*u, p*v on one dot-point column, marched
// in k with the vertical fluxes kept in registers.  Fuses, in the reference's
// order of accumulation (so that the result is bit-comparable without FMA):
//   hadvuv  NH branch        Main/mod_advection.F90:235-265
//   vadvuv                   Main/mod_advection.F90:271-303
//   curvature NH branch      Main/mod_tendency.F90:1840-1879
//   nudgeuv (iboudy 1|5)     Main/mod_bdycod.F90:3581-3823
//   diffu_d (idiffu = 1)     Main/mod_diffusion.F90:266-385
//   raydampuv                Main/mod_bdycod.F90:4990-5026
//   uten*rpsda, *dts         Main/mod_tendency.F90:478-487, mod_sound.F90:228-231
// Output: uten, vten (the acoustic tendencies), one write each.
#include "dev_common.cuh"
#include "launch.h"

using namespace rcm;

__global__ void __launch_bounds__(RCM_MAXT, RCM_MB_UV)
k_uv(const Geom g, const Fields f, const StepScalars s) {
  const int si = g.si;
  const long long sk = g.sk;
  // blockIdx.x = (tile column) * nchunk + (k chunk): the chunks of one tile are
  // neighbours in launch order, so the level shared by two chunks is read from L2
  const int nch = (g.kz + g.kc_uv - 1) / g.kc_uv;
  const int j = g.jdi1 + (blockIdx.x / nch) * blockDim.x + threadIdx.x;
  const int i = g.idi1 + blockIdx.y * blockDim.y + threadIdx.y;
  if (j > g.jdi2 || i > g.idi2) return;
  const int kz = g.kz;
  const int k0 = 1 + (blockIdx.x % nch) * g.kc_uv, k1 = min(kz, k0 + g.kc_uv - 1);
  const bool up = s.upstream_mode != 0;
  const double ul = s.ul;

  // ---- per-column 2-D constants ----
  double md[3][3];  // msfd(j+a-1, i+b-1)
#pragma unroll
  for (int a = 0; a < 3; ++a)
#pragma unroll
    for (int b = 0; b < 3; ++b) md[a][b] = L2(f.msfd, j + a - 1, i + b - 1);
  const double rp_c = L2(f.rpsda, j, i), rp_jm = L2(f.rpsda, j - 1, i), rp_jp = L2(f.rpsda, j + 1, i),
               rp_im = L2(f.rpsda, j, i - 1), rp_ip = L2(f.rpsda, j, i + 1);
  const double dmsf = L2(f.dmsf, j, i), coriol = L2(f.coriol, j, i), ef = L2(f.ef, j, i);
  const double ddx = L2(f.ddx, j, i), ddy = L2(f.ddy, j, i), dmdx = L2(f.dmdx, j, i), dmdy = L2(f.dmdy, j, i);
  const double psdotb = L2(f.psdotb, j, i);
  const int ib = (s.iboudy == 1 || s.iboudy == 5) ? L2(f.zone_d, j, i) : 0;
  const bool dint = diff_interior(g, j, i, true);
  const int nring = ring_count(g, j, i, true);

  // vertical-advection flux at the upper interface of level k (full level k)
  double uu_k = 0.0, vv_k = 0.0;  // uu(1) = 0: loop starts at k = 2
  if (k0 >= 2) {  // this thread starts mid-column: flux through its upper interface
    const double qq = 0.25 * (L3(f.qdot, j, i, k0) + L3(f.qdot, j, i - 1, k0) + L3(f.qdot, j - 1, i, k0) + L3(f.qdot, j - 1, i - 1, k0));
    const double t1 = LD1(&f.twt1[k0]), t2 = LD1(&f.twt2[k0]);
    uu_k = qq * (t1 * L3(f.au, j, i, k0) + t2 * L3(f.au, j, i, k0 - 1));
    vv_k = qq * (t1 * L3(f.av, j, i, k0) + t2 * L3(f.av, j, i, k0 - 1));
  }
  double wadot_k = 0.125 * (L3(f.aw, j - 1, i - 1, k0) + L3(f.aw, j - 1, i, k0) + L3(f.aw, j, i - 1, k0) + L3(f.aw, j, i, k0));

  // request level kk of every streamed field into L2 (see PF3)
  auto prefetch = [&](int kk) {
    PF3(f.au, j, i, kk); PF3(f.av, j, i, kk); PF3(f.cr, j, i, kk); PF3(f.qdot, j, i, kk); PF3(f.aw, j, i, kk); PF3(f.xkc, j, i, kk); PF3(f.dub, j, i, kk); PF3(f.dvb, j, i, kk);
  };
  const int pfd = g.pfd;
  if (pfd > 0)
    for (int kk = k0; kk <= min(k0 + pfd - 1, k1); ++kk) prefetch(kk);
  for (int k = k0; k <= k1; ++k) {
    if (pfd > 0 && k + pfd <= k1) prefetch(k + pfd);
    // ---- level-a raw momentum on the 3x3 box ----
    double au[3][3], av[3][3];
#pragma unroll
    for (int a = 0; a < 3; ++a)
#pragma unroll
      for (int b = 0; b < 3; ++b) {
        au[a][b] = L3(f.au, j + a - 1, i + b - 1, k);
        av[a][b] = L3(f.av, j + a - 1, i + b - 1, k);
      }
    // ua = umc = u*msfd ; [a][b] = (j+a-1, i+b-1)
#define UA(a, b) (au[a][b] * md[a][b])
#define VA(a, b) (av[a][b] * md[a][b])
    const double divd = 0.25 * (L3(f.cr, j, i, k) + L3(f.cr, j, i - 1, k) + L3(f.cr, j - 1, i, k) + L3(f.cr, j - 1, i - 1, k));
    const double ucmona = UA(1, 2) + 2.0 * UA(1, 1) + UA(1, 0);
    double ucmonb = UA(2, 2) + 2.0 * UA(2, 1) + UA(2, 0);
    double ucmonc = UA(0, 2) + 2.0 * UA(0, 1) + UA(0, 0);
    const double vcmona = VA(2, 1) + 2.0 * VA(1, 1) + VA(0, 1);
    double vcmonb = VA(2, 2) + 2.0 * VA(1, 2) + VA(0, 2);
    double vcmonc = VA(2, 0) + 2.0 * VA(1, 0) + VA(0, 0);
    const double diag = divd - dmsf * ((ucmonb - ucmonc) + (vcmonb - vcmonc));
    // decoupled winds ud = u*rpsda on the 5-point cross
    const double u_c = au[1][1] * rp_c, u_jp = au[2][1] * rp_jp, u_jm = au[0][1] * rp_jm, u_ip = au[1][2] * rp_ip,
                 u_im = au[1][0] * rp_im;
    const double v_c = av[1][1] * rp_c, v_jp = av[2][1] * rp_jp, v_jm = av[0][1] * rp_jm, v_ip = av[1][2] * rp_ip,
                 v_im = av[1][0] * rp_im;
    if (up) {
      const double ff1 = ul * (u_jp + u_c);
      const double ff2 = ul * (u_jm + u_c);
      const double ff3 = ul * (v_ip + v_c);
      const double ff4 = ul * (v_im + v_c);
      ucmonb = (1.0 + ff1) * ucmona + (1.0 - ff1) * ucmonb;
      ucmonc = (1.0 + ff2) * ucmonc + (1.0 - ff2) * ucmona;
      vcmonb = (1.0 + ff3) * vcmona + (1.0 - ff3) * vcmonb;
      vcmonc = (1.0 + ff4) * vcmonc + (1.0 - ff4) * vcmona;
    } else {
      ucmonb = ucmona + ucmonb;
      ucmonc = ucmonc + ucmona;
      vcmonb = vcmona + vcmonb;
      vcmonc = vcmonc + vcmona;
    }
    double ut = u_c * diag - dmsf * (u_jp * ucmonb - u_jm * ucmonc + u_ip * vcmonb - u_im * vcmonc);
    double vt = v_c * diag - dmsf * (v_jp * ucmonb - v_jm * ucmonc + v_ip * vcmonb - v_im * vcmonc);
    // ---- vadvuv: + flux(k)*xds(k) then - flux(k+1)*xds(k) ----
    const double uc = au[1][1], vc = av[1][1];
    const double xdsk = LD1(&f.xds[k]);
    if (k >= 2) {
      ut = ut + uu_k * xdsk;
      vt = vt + vv_k * xdsk;
    }
    double uu_n = 0.0, vv_n = 0.0;
    if (k < kz) {
      const int kn = k + 1;
      const double qq = 0.25 * (L3(f.qdot, j, i, kn) + L3(f.qdot, j, i - 1, kn) + L3(f.qdot, j - 1, i, kn) + L3(f.qdot, j - 1, i - 1, kn));
      const double t1 = LD1(&f.twt1[kn]), t2 = LD1(&f.twt2[kn]);
      uu_n = qq * (t1 * L3(f.au, j, i, kn) + t2 * uc);
      vv_n = qq * (t1 * L3(f.av, j, i, kn) + t2 * vc);
      ut = ut - uu_n * xdsk;
      vt = vt - vv_n * xdsk;
    }
    // ---- curvature / Coriolis (NH) ----
    const double wadot_n = 0.125 * (L3(f.aw, j - 1, i - 1, k + 1) + L3(f.aw, j - 1, i, k + 1) + L3(f.aw, j, i - 1, k + 1) + L3(f.aw, j, i, k + 1));
    {
      const double wabar = wadot_k + wadot_n;
      const double amfac = wabar * rp_c * rearthrad;
      const double duv = uc * dmdy - vc * dmdx;
      const double umd = u_c * md[1][1], vmd = v_c * md[1][1];
      ut = ut + coriol * vc - ef * ddx * wabar + vmd * duv - uc * amfac;
      vt = vt - coriol * uc + ef * ddy * wabar - umd * duv - vc * amfac;
    }
    // ---- relaxation toward the boundary fields in the lateral zone ----
    const double xt = s.xt;
    if (ib > 0) {
      const double bu_c = L3(f.bu, j, i, k), bv_c = L3(f.bv, j, i, k);
      const double bu_jm = L3(f.bu, j - 1, i, k), bu_jp = L3(f.bu, j + 1, i, k), bu_im = L3(f.bu, j, i - 1, k), bu_ip = L3(f.bu, j, i + 1, k);
      const double bv_jm = L3(f.bv, j - 1, i, k), bv_jp = L3(f.bv, j + 1, i, k), bv_im = L3(f.bv, j, i - 1, k), bv_ip = L3(f.bv, j, i + 1, k);
      const double xf = LD1(&f.hefd[ib + s.nspgd * (k - 1)]), xg = LD1(&f.hegd[ib + s.nspgd * (k - 1)]);
#define FGU(jj, ii, bval) ((L3(f.xub0, jj, ii, k) + xt * L3(f.xubt, jj, ii, k)) - (bval))
#define FGV(jj, ii, bval) ((L3(f.xvb0, jj, ii, k) + xt * L3(f.xvbt, jj, ii, k)) - (bval))
      double fls0 = FGU(j, i, bu_c), fls1 = FGU(j - 1, i, bu_jm), fls2 = FGU(j + 1, i, bu_jp), fls3 = FGU(j, i - 1, bu_im), fls4 = FGU(j, i + 1, bu_ip);
      ut = ut + xf * fls0 - xg * (fls1 + fls2 + fls3 + fls4 - 4.0 * fls0);
      fls0 = FGV(j, i, bv_c); fls1 = FGV(j - 1, i, bv_jm); fls2 = FGV(j + 1, i, bv_jp); fls3 = FGV(j, i - 1, bv_im); fls4 = FGV(j, i + 1, bv_ip);
      vt = vt + xf * fls0 - xg * (fls1 + fls2 + fls3 + fls4 - 4.0 * fls0);
#undef FGU
#undef FGV
    }
    // ---- 4th-order diffusion on level-b decoupled winds, each neighbour
    //      divided by its own (inverted) map factor: the quotients
    //      (ub/psdotb)/msfd come precomputed from k_xkc ----
    {
      double xkd = 0.25 * (L3(f.xkc, j, i, k) + L3(f.xkc, j - 1, i - 1, k) + L3(f.xkc, j - 1, i, k) + L3(f.xkc, j, i - 1, k));
      xkd = xkd * s.rdxsq * psdotb;
      const double du_c = L3(f.dub, j, i, k), du_jm = L3(f.dub, j - 1, i, k), du_jp = L3(f.dub, j + 1, i, k), du_im = L3(f.dub, j, i - 1, k), du_ip = L3(f.dub, j, i + 1, k);
      const double dv_c = L3(f.dvb, j, i, k), dv_jm = L3(f.dvb, j - 1, i, k), dv_jp = L3(f.dvb, j + 1, i, k), dv_im = L3(f.dvb, j, i - 1, k), dv_ip = L3(f.dvb, j, i + 1, k);
      if (dint) {
        const double du_jp2 = L3(f.dub, j + 2, i, k), du_jm2 = L3(f.dub, j - 2, i, k);
        const double du_ip2 = L3(f.dub, j, i + 2, k), du_im2 = L3(f.dub, j, i - 2, k);
        const double dv_jp2 = L3(f.dvb, j + 2, i, k), dv_jm2 = L3(f.dvb, j - 2, i, k);
        const double dv_ip2 = L3(f.dvb, j, i + 2, k), dv_im2 = L3(f.dvb, j, i - 2, k);
        ut = ut - xkd * (1.0 * (du_jp2 + du_jm2 + du_ip2 + du_im2) + (-4.0) * (du_jp + du_jm + du_ip + du_im) + 12.0 * (du_c));
        vt = vt - xkd * (1.0 * (dv_jp2 + dv_jm2 + dv_ip2 + dv_im2) + (-4.0) * (dv_jp + dv_jm + dv_ip + dv_im) + 12.0 * (dv_c));
      }
      for (int r = 0; r < nring; ++r) {
        ut = ut + xkd * (1.0 * (du_jp + du_jm + du_ip + du_im) + (-4.0) * (du_c));
        vt = vt + xkd * (1.0 * (dv_jp + dv_jm + dv_ip + dv_im) + (-4.0) * (dv_c));
      }
    }
    // ---- Rayleigh damping toward the boundary fields in the top levels ----
    if (s.ifrayd == 1 && k <= s.kt_ray) {
      const double tau = L3(f.tau_d, j, i, k);
      ut = ut + tau * ((L3(f.xub0, j, i, k) + xt * L3(f.xubt, j, i, k)) - L3(f.bu, j, i, k));
      vt = vt + tau * ((L3(f.xvb0, j, i, k) + xt * L3(f.xvbt, j, i, k)) - L3(f.bv, j, i, k));
    }
    // ---- hand over to the acoustic loop: decouple and premultiply by dts ----
    P3(f.uten, j, i, k) = (ut * rp_c) * s.dts;
    P3(f.vten, j, i, k) = (vt * rp_c) * s.dts;
    uu_k = uu_n; vv_k = vv_n; wadot_k = wadot_n;
#undef UA
#undef VA
  }
}

void launch_uv(const Geom& g, const Fields& f, const StepScalars& s, cudaStream_t st) {
  const int nj = g.jdi2 - g.jdi1 + 1, ni = g.idi2 - g.idi1 + 1;
  const int nch = (g.kz + g.kc_uv - 1) / g.kc_uv;
  const dim3 b = rcm::col_block(); dim3 gr(((nj + b.x - 1) / b.x) * nch, (ni + b.y - 1) / b.y);
  RCM_LAUNCH(k_uv, gr, b, st, g, f, s);
}
