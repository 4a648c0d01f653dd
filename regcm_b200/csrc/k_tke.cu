// TKE transport of the sigma-coordinate cores (ibltyp = 2, the UW PBL's prognostic TKE):
//   k_tke          the dynamical TKE tendency of `tend` and the forecast, one thread per (j,i) column marching the
//                  kz+1 full levels in the reference's order of accumulation:
//                    hadv -> hadv3d(ind = 1)   Main/mod_advection.F90:395-515   on atm1%tke (uncoupled), winds
//                                              interpolated to full levels with twt
//                    vadv -> vadv3d(kzp1, 0)   Main/mod_advection.F90:755-766   on tkeps = atm1%tke * psa
//                                              (Main/mod_tendency.F90:1414-1425)
//                    diffu_x -> diffu_x3df     Main/mod_diffusion.F90:523-601   on atm2%tke with fac = nuk (:1544-1548)
//                    tketen = tkedyn*rpsa + tkephy; atmc%tke = max(tkemin, atm2%tke + dt*tketen)   (:518-541)
//   k_tke_filter   timefilter_apply(atm1%tke, atm2%tke, atmc%tke, gnu2) = filter_ra_3d (Main/mod_timefilter.F90:92-110),
//                  in place (nothing reads TKE between the two kernels)
//   k_tke_bdy      bdyval: atm2 rim <- atm1 rim when integrating (Main/mod_bdycod.F90:1166-1172,1212-1218,1258-1264,
//                  1301-1307); rim = tkemin at start; afterwards level 1 = tkemin and, on levels 3..kz+1, inflow -> tkemin,
//                  outflow -> the interior neighbour (:2412-2496).  West/east cells first, south/north cells in a second
//                  launch: the west/east rule reads the south/north rows before they change, as the sequential code does.
// Both cores share the kernel: they provide au, av, msfd (uavg/vavg), qdot, xkc and p* under the same names.
#include "dev_common.cuh"
#include "launch.h"

using namespace rcm;

namespace {

__global__ void __launch_bounds__(RCM_MAXT)
k_tke(const Geom g, const Fields f, const UwFields uw, const StepScalars s) {
  const int si = g.si;
  const long long sk = g.sk;
  const int j = g.jci1 + blockIdx.x * blockDim.x + threadIdx.x;
  const int i = g.ici1 + blockIdx.y * blockDim.y + threadIdx.y;
  if (j > g.jci2 || i > g.ici2) return;
  const int kz = g.kz, kzp1 = g.kzp1;
  const bool up = s.upstream_mode != 0;
  const double ul = s.ul, dt = s.dt;
  const double* __restrict__ a1 = uw.a;
  const double* __restrict__ a2 = uw.b;

  const double psa = L2(f.psa, j, i), psb = L2(f.psb, j, i);
  const double rpsa_c = L2(f.rpsa, j, i);
  const double xmsf = L2(f.xmsf, j, i);
  const double m00 = L2(f.msfd, j, i), m01 = L2(f.msfd, j, i + 1), m10 = L2(f.msfd, j + 1, i), m11 = L2(f.msfd, j + 1, i + 1);
  const bool d9 = (s.idiffu == 2);
  const bool d6on = (s.idiffu == 3) && j == g.jci2;   // the one column idiffu = 3 acts on (k_diff6)
  const bool dint = (s.idiffu == 1) && diff_interior(g, j, i, false);
  const int nring = (s.idiffu == 1) ? ring_count(g, j, i, false) : 0;

  double uavg1_m = 0, uavg2_m = 0, vavg1_m = 0, vavg2_m = 0;   // half level k-1
  double fx_m = 0.0;                                           // vadv3d flux of loop index k-1
  double xkc_m = 0.0;
  double tke_k = L3(a1, j, i, 1);
  auto prefetch = [&](int kk) {   // level kk of every streamed field into L2 (see PF3)
    PF3(f.au, j, i, kk); PF3(f.av, j, i, kk); PF3(f.au, j, i + 1, kk); PF3(f.av, j, i + 1, kk); PF3(f.qdot, j, i, kk); PF3(f.xkc, j, i, kk);
    PF3(a1, j, i, kk); PF3(a2, j, i, kk); PF3(uw.phy, j, i, kk);
  };
  const int pfd = g.pfd;
  if (pfd > 0)
    for (int kk = 1; kk <= min(pfd, kzp1); ++kk) prefetch(kk);
  for (int k = 1; k <= kzp1; ++k) {
    if (pfd > 0 && k + pfd <= kzp1) prefetch(k + pfd);
    const bool half = (k <= kz);
    double uavg1 = 0, uavg2 = 0, vavg1 = 0, vavg2 = 0, xkc_raw = 0;
    if (half) {
      const double ua00 = L3(f.au, j, i, k) * m00, ua01 = L3(f.au, j, i + 1, k) * m01, ua10 = L3(f.au, j + 1, i, k) * m10,
                   ua11 = L3(f.au, j + 1, i + 1, k) * m11;
      const double va00 = L3(f.av, j, i, k) * m00, va01 = L3(f.av, j, i + 1, k) * m01, va10 = L3(f.av, j + 1, i, k) * m10,
                   va11 = L3(f.av, j + 1, i + 1, k) * m11;
      uavg1 = ua01 + ua00; uavg2 = ua11 + ua10;
      vavg1 = va10 + va00; vavg2 = va11 + va01;
      xkc_raw = L3(f.xkc, j, i, k);
    }
    double ten = 0.0;
    // hadv3d, ind = 1: full levels 2..kz
    if (k >= 2 && k <= kz) {
      const double t1 = LD1(&f.twt1[k]), t2 = LD1(&f.twt2[k]);
      const double uaz1 = (t1 * uavg1 + t2 * uavg1_m);
      const double uaz2 = (t1 * uavg2 + t2 * uavg2_m);
      const double vaz1 = (t1 * vavg1 + t2 * vavg1_m);
      const double vaz2 = (t1 * vavg2 + t2 * vavg2_m);
      const double q_jm = L3(a1, j - 1, i, k), q_jp = L3(a1, j + 1, i, k), q_im = L3(a1, j, i - 1, k), q_ip = L3(a1, j, i + 1, k);
      double fx1, fx2, fy1, fy2;
      if (up) {
        const double f1 = 0.5 * ul * (uavg2 + uavg1) / psa;
        const double f2 = 0.5 * ul * (vavg2 + vavg1) / psa;
        fx1 = (1.0 + f1) * q_jm + (1.0 - f1) * tke_k;
        fx2 = (1.0 + f1) * tke_k + (1.0 - f1) * q_jp;
        fy1 = (1.0 + f2) * q_im + (1.0 - f2) * tke_k;
        fy2 = (1.0 + f2) * tke_k + (1.0 - f2) * q_ip;
      } else {
        fx1 = q_jm + tke_k; fx2 = tke_k + q_jp; fy1 = q_im + tke_k; fy2 = tke_k + q_ip;
      }
      ten = ten - xmsf * (uaz2 * fx2 - uaz1 * fx1 + vaz2 * fy2 - vaz1 * fy1);
    }
    // vadv3d nk = kzp1 on tkeps: loop index k-1 adds, loop index k subtracts
    const double ddsk = LD1(&f.dds[k]);
    if (k >= 2) ten = ten + fx_m * ddsk;
    double fx = 0.0, tke_n = 0.0;
    if (half) {
      tke_n = L3(a1, j, i, k + 1);
      const double qq = 0.5 * (L3(f.qdot, j, i, k) + L3(f.qdot, j, i, k + 1));
      fx = qq * ((tke_k * psa + tke_n * psa));
      ten = ten - fx * ddsk;
    }
    // diffu_x3df(tkedyn, atm2%tke, nuk): xkcf(1) = xkc(1), xkcf(k+1) = xkc(k)   (Main/mod_diffusion.F90:231-235)
    const double b_c = L3(a2, j, i, k);
    {
      const double xkcf = ((k == 1) ? xkc_raw : xkc_m) * s.rdxsq * psb;
      const double d_jm = L3(a2, j - 1, i, k), d_jp = L3(a2, j + 1, i, k), d_im = L3(a2, j, i - 1, k), d_ip = L3(a2, j, i + 1, k);
      if (dint) {
        const double d_jp2 = L3(a2, j + 2, i, k), d_jm2 = L3(a2, j - 2, i, k), d_ip2 = L3(a2, j, i + 2, k), d_im2 = L3(a2, j, i - 2, k);
        ten = ten - uw.nuk * xkcf * (1.0 * (d_jp2 + d_jm2 + d_ip2 + d_im2) + (-4.0) * (d_jp + d_jm + d_ip + d_im) + 12.0 * (b_c));
      }
      for (int r = 0; r < nring; ++r) ten = ten + uw.nuk * xkcf * (1.0 * (d_jp + d_jm + d_ip + d_im) + (-4.0) * (b_c));
      if (d9) ten = ten + uw.nuk * xkcf * lap9(b_c, d_jm, d_jp, d_im, d_ip, L3(a2, j + 1, i + 1, k), L3(a2, j - 1, i - 1, k),
                                               L3(a2, j - 1, i + 1, k), L3(a2, j + 1, i - 1, k));
      if (d6on) ten = ten + D6(RCM_D6_TKE, i, k);
    }
    const double tsum = 0.0 + ten * rpsa_c + L3(uw.phy, j, i, k);
    P3(uw.c, j, i, k) = fmax(uw.tkemin, b_c + dt * tsum);
    uavg1_m = uavg1; uavg2_m = uavg2; vavg1_m = vavg1; vavg2_m = vavg2;
    fx_m = fx; xkc_m = xkc_raw; tke_k = tke_n;
  }
}

__global__ void __launch_bounds__(RCM_MAXT)
k_tke_filter(const Geom g, const UwFields uw, double alpha) {
  const int si = g.si;
  const long long sk = g.sk;
  const int j = g.jci1 + blockIdx.x * blockDim.x + threadIdx.x;
  const int i = g.ici1 + blockIdx.y * blockDim.y + threadIdx.y;
  if (j > g.jci2 || i > g.ici2) return;
  for (int k = 1; k <= g.kzp1; ++k) {
    const double a = P3(uw.a, j, i, k), b = P3(uw.b, j, i, k), c = P3(uw.c, j, i, k);
    const double d = alpha * (c + b - 2.0 * a);
    P3(uw.b, j, i, k) = a + d;
    P3(uw.a, j, i, k) = c;
  }
}

// pass 0: every rim cell copies b <- a (integrating) and sets level 1; west/east cells apply the rule.
// pass 1: south/north cells with j in jci1..jci2 apply the rule.  One thread per (rim cell, full level kk).
__global__ void k_tke_bdy(const Geom g, const Fields f, const UwFields uw, const int2* rim, int nrim, int integrating, int iboudy,
                          int pass) {
  const int si = g.si;
  const long long sk = g.sk;
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  const int kk = 1 + blockIdx.y;
  if (p >= nrim) return;
  const int j = rim[p].x, i = rim[p].y;
  const int gjce1 = 1, gjce2 = g.jx - 1, gice1 = 1, gice2 = g.iy - 1;
  const int gjci1 = 2, gjci2 = g.jx - 2, gici1 = 2, gici2 = g.iy - 2;
  const bool we = (j == gjce1 || j == gjce2);
  if (!integrating) {   // rcmtimer%start()
    if (pass == 0) { P3(uw.a, j, i, kk) = uw.tkemin; P3(uw.b, j, i, kk) = uw.tkemin; }
    return;
  }
  if (pass == 0) {
    // west/east rows i = ici1..ici2, south/north rows whole (the corners belong to them)
    if (!we || (i >= gici1 && i <= gici2) || i == gice1 || i == gice2) P3(uw.b, j, i, kk) = P3(uw.a, j, i, kk);
    if (kk == 1) { P3(uw.a, j, i, 1) = uw.tkemin; P3(uw.b, j, i, 1) = uw.tkemin; }
  }
  if (kk < 3) return;
  const int k = kk - 1;   // the reference's loop index: half levels k and k-1 of the boundary winds, TKE level k+1
  const bool tdep = (iboudy != 0);
  if (pass == 0 && we) {
    const bool west = (j == gjce1);
    const int jin = west ? gjci1 : gjci2, jde = west ? 1 : g.jx, jdi = west ? 2 : g.jx - 1;
    const double tkeint = P3(uw.a, jin, i, kk);
    // wue/eue are never assigned at the two corner rows in the time-dependent branch (Main/mod_bdycod.F90:1000-1027)
    const bool z0 = tdep && (i < 2 || i > g.iy - 1), z1 = tdep && (i + 1 < 2 || i + 1 > g.iy - 1);
    const double e0 = z0 ? 0.0 : P3(f.au, jde, i, k), e1 = z1 ? 0.0 : P3(f.au, jde, i + 1, k);
    const double e0m = z0 ? 0.0 : P3(f.au, jde, i, k - 1), e1m = z1 ? 0.0 : P3(f.au, jde, i + 1, k - 1);
    const double windavg = e0 + e1 + P3(f.au, jdi, i, k) + P3(f.au, jdi, i + 1, k) + e0m + e1m + P3(f.au, jdi, i, k - 1) +
                           P3(f.au, jdi, i + 1, k - 1);
    const bool inflow = west ? (windavg > 0.0) : (windavg < 0.0);
    P3(uw.a, j, i, kk) = inflow ? uw.tkemin : tkeint;
  } else if (pass == 1 && !we && j >= gjci1 && j <= gjci2) {
    const bool south = (i == gice1);
    const int iin = south ? gici1 : gici2, ide = south ? 1 : g.iy, idi = south ? 2 : g.iy - 1;
    const double tkeint = P3(uw.a, j, iin, kk);
    const double windavg = P3(f.av, j, ide, k) + P3(f.av, j + 1, ide, k) + P3(f.av, j, idi, k) + P3(f.av, j + 1, idi, k) +
                           P3(f.av, j, ide, k - 1) + P3(f.av, j + 1, ide, k - 1) + P3(f.av, j, idi, k - 1) + P3(f.av, j + 1, idi, k - 1);
    const bool inflow = south ? (windavg > 0.0) : (windavg < 0.0);
    P3(uw.a, j, i, kk) = inflow ? uw.tkemin : tkeint;
  }
}

}  // namespace

void launch_tke(const Geom& g, const Fields& f, const UwFields& uw, const StepScalars& s, cudaStream_t st) {
  const dim3 b = rcm::col_block();
  const dim3 gr((g.jci2 - g.jci1 + b.x) / b.x, (g.ici2 - g.ici1 + b.y) / b.y);
  RCM_LAUNCH(k_tke, gr, b, st, g, f, uw, s);
  RCM_LAUNCH(k_tke_filter, gr, b, st, g, uw, s.gnu2);
}
void launch_tke_bdy(const Geom& g, const Fields& f, const UwFields& uw, const int2* rim, int nrim, int integrating, int iboudy,
                    cudaStream_t st) {
  if (nrim <= 0) return;
  dim3 b(128), gr((nrim + 127) / 128, g.kzp1);
  RCM_LAUNCH(k_tke_bdy, gr, b, st, g, f, uw, rim, nrim, integrating, iboudy, 0);
  if (integrating) RCM_LAUNCH(k_tke_bdy, gr, b, st, g, f, uw, rim, nrim, integrating, iboudy, 1);
}
