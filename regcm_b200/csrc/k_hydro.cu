// Hydrostatic core (idynamic = 1), slow part of `tend` (SURVEY.md §8 a3,a4,a5,
// a8-a15,a17,a18,a20,a21).  One thread owns one (j,i) column and marches k, as
// in the non-hydrostatic kernels; each kernel evaluates the COMPLETE tendency
// of its prognostics in the reference's order of accumulation so that, with FMA
// contraction off, results are bit-comparable with the CPU restatement except
// where libm pow/log differ.
//   k_h_ps      surface_pressures + psc2psd   Main/mod_tendency.F90:815-850,858-862;
//                                             Main/mpplib/mod_mppparam.F90:13811-13862
//   k_h_rimcopy carries the physical-boundary rows into the double-buffered "next" levels
//   k_h_prep    compute_omega (cr, pten, qdot) :1127-1156; geopotential phi of
//               pressure_gradient_force :2033-2097 (ipgf = 0, 1); new_pressure :1428-1460
//               with nudge2d Main/mod_bdycod.F90:4597-4766
//   k_h_uv      hadvuv (hydro) Main/mod_advection.F90:203-234, vadvuv :271-303,
//               curvature :1829-1838, nudgeuv, diffu_d, pressure_gradient_force
//               :1996-2025,2103-2115, forecast :433-440, filter_ra_uv
//   k_h_tq      hadvt, vadv3d(ind=1) Main/mod_advection.F90:771-783, omega :1199-1215,
//               adiabatic :1561-1575, hadvqv/vadvqv, hadvqx/vadv4d, nudge3d/4d3d,
//               diffu_x3d/x4d, forecast, negative-q guard, filters RA(t) RAW(qv) RAW(qc)
//   k_h_psfilt  filter_ra_2d(psa,psb,psc) Main/mod_timefilter.F90:58-73, then psc2psd(psa)
//               as at the head of splitf (Main/mod_split.F90:259-260)
#ifndef RCM_IDX32_ALL   // (tuning builds only)
#define RCM_IDX64   // 64-bit element indices measured faster for the kernels of this file (dev_common.cuh)
#endif
#include "dev_common.cuh"
#include "launch.h"

using namespace rcm;

// psc2psd on global indices: the physical-boundary rules key on the global row/column, so the same code serves
// every tile of a decomposition (the cross-point p* it reads lies in the exchanged ghost rows)
__device__ __forceinline__ double psd_at(const Geom& g, const double* pc, int j, int i) {
  const int si = g.si;
  const bool jl = (j == 1), jr = (j == g.jx), ib = (i == 1), it = (i == g.iy);
  if ((jl || jr) && (ib || it)) return P2(pc, jl ? 1 : g.jx - 1, ib ? 1 : g.iy - 1);
  if (it) return (P2(pc, j, g.iy - 1) + P2(pc, j - 1, g.iy - 1)) * 0.5;
  if (ib) return (P2(pc, j, 1) + P2(pc, j - 1, 1)) * 0.5;
  if (jl) return (P2(pc, 1, i) + P2(pc, 1, i - 1)) * 0.5;
  if (jr) return (P2(pc, g.jx - 1, i) + P2(pc, g.jx - 1, i - 1)) * 0.5;
  return (P2(pc, j, i) + P2(pc, j, i - 1) + P2(pc, j - 1, i) + P2(pc, j - 1, i - 1)) * 0.25;
}

// generic: pd = psc2psd(pc) on the dot extent; optional reciprocals
// gw = ghost rows computed on the HIGH sides that have a neighbour rank (0: the tile's own dot points)
__global__ void k_psc2psd(const Geom g, const double* pc, double* pd, double* rpd, const int gw) {
  const int si = g.si;
  const int j = g.jde1 + blockIdx.x * blockDim.x + threadIdx.x;
  const int i = g.ide1 + blockIdx.y * blockDim.y + threadIdx.y;
  if (j > g.jde2 + gw * g.gr || i > g.ide2 + gw * g.gt) return;
  const double v = psd_at(g, pc, j, i);
  P2(pd, j, i) = v;
  if (rpd) P2(rpd, j, i) = 1.0 / v;
}

// xpsb record: surface pressure [hPa] -> p* [cb] (Main/mod_bdycod.F90:394)
__global__ void k_ps_convert(const Geom g, double* ps, double ptop) {   // whole host window: tile + RCM_HALO ghost rows
  const int si = g.si;
  const int j = (g.jce1 - RCM_HALO * g.gl) + blockIdx.x * blockDim.x + threadIdx.x;
  const int i = (g.ice1 - RCM_HALO * g.gb) + blockIdx.y * blockDim.y + threadIdx.y;
  if (j > g.jce2 + RCM_HALO * g.gr || i > g.ice2 + RCM_HALO * g.gt) return;
  P2(ps, j, i) = (P2(ps, j, i) * 0.1) - ptop;
}

// On a decomposed domain 1/p* on cross points is also computed on the RCM_HALO ghost rows (p* was exchanged that
// wide at the end of the previous step); the dot-point tables need cross-point p* one row further out than the
// ghost rows reach, so they are computed on the tile's own points and exchanged (the reference does the same:
// psc2psd, then exchange; Main/mod_tendency.F90:858-869).
__global__ void k_h_ps(const Geom g, const Fields f) {
  const int si = g.si;
  const int j = (g.jde1 - 2 * g.gl) + blockIdx.x * blockDim.x + threadIdx.x;
  const int i = (g.ide1 - 2 * g.gb) + blockIdx.y * blockDim.y + threadIdx.y;
  if (j > g.jde2 + 2 * g.gr || i > g.ide2 + 2 * g.gt) return;
  if (j >= g.jde1 && j <= g.jde2 && i >= g.ide1 && i <= g.ide2) {
    const double pa = psd_at(g, f.psa, j, i), pb = psd_at(g, f.psb, j, i);
    P2(f.psdota, j, i) = pa; P2(f.rpsda, j, i) = 1.0 / pa;
    P2(f.psdotb, j, i) = pb; P2(f.rpsdb, j, i) = 1.0 / pb;
  }
  if (j <= g.jce2 + 2 * g.gr && i <= g.ice2 + 2 * g.gt) {
    P2(f.rpsa, j, i) = 1.0 / P2(f.psa, j, i);
    P2(f.rpsb, j, i) = 1.0 / P2(f.psb, j, i);
  }
}

// The slow kernels write interior points of the "next" buffers only; the rim
// rows (set by the last bdyval) are carried over here, because splitf reads
// atm1/atm2 on the whole tile (Main/mod_split.F90:271-274,354-359).
__global__ void k_h_rimcopy(const Geom g, const Fields f, const int2* rim_d, int nrim_d, const int2* rim_c, int nrim_c) {
  const int si = g.si;
  const long long sk = g.sk;
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  const int k = 1 + blockIdx.y;
  if (p < nrim_d) {
    const int j = rim_d[p].x, i = rim_d[p].y;
    P3(f.nau, j, i, k) = P3(f.au, j, i, k); P3(f.nav, j, i, k) = P3(f.av, j, i, k);
    P3(f.nbu, j, i, k) = P3(f.bu, j, i, k); P3(f.nbv, j, i, k) = P3(f.bv, j, i, k);
  }
  if (p < nrim_c) {
    const int j = rim_c[p].x, i = rim_c[p].y;
    P3(f.nat, j, i, k) = P3(f.at, j, i, k); P3(f.nbt, j, i, k) = P3(f.bt, j, i, k);
    P3(f.naqv, j, i, k) = P3(f.aqv, j, i, k); P3(f.nbqv, j, i, k) = P3(f.bqv, j, i, k);
    P3(f.naqc, j, i, k) = P3(f.aqc, j, i, k); P3(f.nbqc, j, i, k) = P3(f.bqc, j, i, k);
  }
}

// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(RCM_MAXT, RCM_MB_PREP)
k_h_prep(const Geom g, const Fields f, const StepScalars s) {
  const int si = g.si;
  const long long sk = g.sk;
  // tile + one ghost ring towards neighbour ranks: qdot, phi and the forecast p* are read one row out by k_h_uv,
  // k_h_psfilt and psc2psd; recomputing that ring (bit-identical to the owner's values) replaces three exchanges
  const int j = (g.jce1 - g.gl) + blockIdx.x * blockDim.x + threadIdx.x;
  const int i = (g.ice1 - g.gb) + blockIdx.y * blockDim.y + threadIdx.y;
  if (j > g.jce2 + g.gr || i > g.ice2 + g.gt) return;
  const int kz = g.kz;
  const double rpsa = L2(f.rpsa, j, i), psa = L2(f.psa, j, i), psb = L2(f.psb, j, i);
  const double msfx = L2(f.msfx, j, i);
  const double dummy = 1.0 / (s.dx2 * msfx * msfx);
  const double m00 = L2(f.msfd, j, i), m01 = L2(f.msfd, j, i + 1), m10 = L2(f.msfd, j + 1, i),
               m11 = L2(f.msfd, j + 1, i + 1);
  // ---- mass divergence and p* tendency (compute_omega :1131-1148) ----
  // request the next levels of the streamed fields into L2 (see PF3 in dev_common.cuh)
  const int pfd = g.pfd;
  auto pf_uv = [&](int kk) { PF3(f.au, j, i, kk); PF3(f.av, j, i, kk); PF3(f.au, j, i + 1, kk); PF3(f.av, j, i + 1, kk); };
  auto pf_t = [&](int kk) { PF3(f.at, j, i, kk); PF3(f.aqv, j, i, kk); PF3(f.aqc, j, i, kk); };
  if (pfd > 0) {
    for (int kk = 1; kk <= min(pfd, kz); ++kk) pf_uv(kk);
    for (int kk = kz; kk >= max(kz - pfd + 1, 1); --kk) pf_t(kk);
  }
  double pten = 0.0;
  for (int k = 1; k <= kz; ++k) {
    if (pfd > 0 && k + pfd <= kz) pf_uv(k + pfd);
    const double u00 = L3(f.au, j, i, k) * m00, u01 = L3(f.au, j, i + 1, k) * m01, u10 = L3(f.au, j + 1, i, k) * m10,
                 u11 = L3(f.au, j + 1, i + 1, k) * m11;
    const double v00 = L3(f.av, j, i, k) * m00, v01 = L3(f.av, j, i + 1, k) * m01, v10 = L3(f.av, j + 1, i, k) * m10,
                 v11 = L3(f.av, j + 1, i + 1, k) * m11;
    const double cr = ((u11 + u10 - u01 - u00) + (v11 + v01 - v10 - v00)) * dummy;
    P3(f.cr, j, i, k) = cr;
    pten = pten - cr * LD1(&f.dsigma[k]);
  }
  P2(f.pten, j, i) = pten;
  // ---- sigma-dot by downward recurrence (:1149-1156) ----
  double qd = 0.0;
  P3(f.qdot, j, i, 1) = 0.0;
  for (int k = 2; k <= kz; ++k) {
    qd = qd - (pten + P3(f.cr, j, i, k - 1)) * LD1(&f.dsigma[k - 1]) * rpsa;
    P3(f.qdot, j, i, k) = qd;
  }
  P3(f.qdot, j, i, kz + 1) = 0.0;
  // ---- geopotential at half levels, bottom-up (pressure_gradient_force :2033-2097;
  //      alpha_hyd = 0, beta_hyd = 1 => td = p*.Tv of level a everywhere) ----
  {
    const double ptr = s.ptop * rpsa;
    double td_n = 0.0, phi_n = 0.0, hs_n = 0.0, ds_n = 0.0;
    for (int k = kz; k >= 1; --k) {
      if (pfd > 0 && k - pfd >= 1) pf_t(k - pfd);
      const double qvd = fmax(L3(f.aqv, j, i, k) * rpsa, minqq);
      const double qcd = fmax(L3(f.aqc, j, i, k) * rpsa, 0.0);
      double td = L3(f.at, j, i, k) * (1.0 + ep1 * qvd);
      const double tvfac = 1.0 / (1.0 + qcd / (1.0 + qvd));
      const double hs = (LD1(&f.sigma[k + 1]) + LD1(&f.sigma[k])) * 0.5;
      // ipgf = 1: ttld = td - p*.t00pg.(p/p00pg)^pgfaa1 takes the place of td (:1901-1902,2044,2060-2062)
      if (s.ipgf == 1) td = td - psa * t00pg * pow_pos((hs * psa + s.ptop) / p00pg, pgfaa1);
      const double ds = LD1(&f.dsigma[k]);
      double phi;
      if (k == kz) {
        const double tv = td * rpsa * tvfac;
        phi = L2(f.ht, j, i);
        if (s.ipgf == 1) phi = L2(f.ht, j, i) + rgas * t00pg / pgfaa1 * pow_pos((psa + s.ptop) / p00pg, pgfaa1);   // :2046-2047
        phi = phi - rgas * tv * log((hs + ptr) / (1.0 + ptr));
      } else {
        const double tvavg = ((td * ds + td_n * ds_n) / (psa * (ds + ds_n))) * tvfac;
        phi = phi_n - rgas * tvavg * log((hs + ptr) / (hs_n + ptr));
      }
      P3(f.phi, j, i, k) = phi;
      td_n = td; phi_n = phi; hs_n = hs; ds_n = ds;
    }
  }
  // ---- new_pressure (:1428-1460): relaxation of pten, forecast p* ----
  const bool own = (j >= g.jci1 && j <= g.jci2 && i >= g.ici1 && i <= g.ici2);        // this rank's interior points
  const bool interior = (j >= 2 && j <= g.jx - 2 && i >= 2 && i <= g.iy - 2);         // interior of the global domain
  double ptn = pten;
  if (interior && (s.iboudy == 1 || s.iboudy == 5)) {
    const int ib = L2(f.zone_c, j, i);
    if (ib > 0) {
      const double xt = s.xt;
      const double xf = LD1(&f.hefc[ib + s.nspgx * (kz - 1)]), xg = LD1(&f.hegc[ib + s.nspgx * (kz - 1)]);
#define FG(jj, ii) ((L2(f.xpsb0, jj, ii) + xt * L2(f.xpsbt, jj, ii)) - L2(f.psb, jj, ii))
      const double fls0 = FG(j, i), fls1 = FG(j - 1, i), fls2 = FG(j + 1, i), fls3 = FG(j, i - 1), fls4 = FG(j, i + 1);
#undef FG
      ptn = ptn + xf * fls0 - xg * (fls1 + fls2 + fls3 + fls4 - 4.0 * fls0);
    }
  }
  if (interior && s.iboudy == 4) {   // sponge2d (Main/mod_tendency.F90:1434, mod_bdycod.F90:3065-3122)
    const int ib = L2(f.zone_c, j, i);
    if (ib > 0) ptn = LD1(&f.hegc[ib]) * ptn + LD1(&f.hefc[ib]) * L2(f.xpsbt, j, i);
  }
  const double psc = psb + ptn * s.dt;
  P2(f.psc, j, i) = psc;
  if (own) {
    P2(f.ptn2d, j, i) = s.integrating ? fabs(ptn) : 0.0;
    P2(f.pt22d, j, i) = s.integrating ? fabs((psc + psb - 2.0 * psa) / (s.dt * s.dt * 0.25)) : 0.0;
  }
}

// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(RCM_MAXT, RCM_MB_UV)
k_h_uv(const Geom g, const Fields f, const StepScalars s) {
  const int si = g.si;
  const long long sk = g.sk;
  const int j = g.jdi1 + blockIdx.x * blockDim.x + threadIdx.x;
  const int i = g.idi1 + blockIdx.y * blockDim.y + threadIdx.y;
  if (j > g.jdi2 || i > g.idi2) return;
  const int kz = g.kz;
  const bool up = s.upstream_mode != 0;
  const double ul = s.ul, xt = s.xt;

  double md[3][3];  // msfd(j+a-1, i+b-1)
#pragma unroll
  for (int a = 0; a < 3; ++a)
#pragma unroll
    for (int b = 0; b < 3; ++b) md[a][b] = L2(f.msfd, j + a - 1, i + b - 1);
  const double rp_c = L2(f.rpsda, j, i), rp_jm = L2(f.rpsda, j - 1, i), rp_jp = L2(f.rpsda, j + 1, i),
               rp_im = L2(f.rpsda, j, i - 1), rp_ip = L2(f.rpsda, j, i + 1);
  const double dmsf = L2(f.dmsf, j, i), coriol = L2(f.coriol, j, i);
  const double psdota = L2(f.psdota, j, i), psdotb = L2(f.psdotb, j, i);
  const int ib = (s.iboudy == 1 || s.iboudy == 5) ? L2(f.zone_d, j, i) : 0;
  const int ibs = (s.iboudy == 4) ? L2(f.zone_d, j, i) : 0;   // sponge zone (iboudy = 4)
  const bool d9 = (s.idiffu == 2);
  const bool d6on = (s.idiffu == 3) && j == g.jdi2;   // the one column idiffu = 3 acts on (k_diff6)   // nine-point operator on every interior point, no boundary ring
  const bool dint = (s.idiffu == 1) && diff_interior(g, j, i, true);
  const int nring = (s.idiffu == 1) ? ring_count(g, j, i, true) : 0;
  // cross points around the dot: 11=(j,i) 10=(j,i-1) 01=(j-1,i) 00=(j-1,i-1)
  const double pa11 = L2(f.psa, j, i), pa10 = L2(f.psa, j, i - 1), pa01 = L2(f.psa, j - 1, i), pa00 = L2(f.psa, j - 1, i - 1);
  const double ra11 = L2(f.rpsa, j, i), ra10 = L2(f.rpsa, j, i - 1), ra01 = L2(f.rpsa, j - 1, i), ra00 = L2(f.rpsa, j - 1, i - 1);
  const double dxm = s.dx * md[1][1], dx2m = s.dx2 * md[1][1];
  const double rdxm = recip_for_div(dxm), rdx2m = recip_for_div(dx2m);   // shared-denominator divisions (div_r)

  double uu_k = 0.0, vv_k = 0.0;  // vertical-advection flux at the upper interface of level k
  const int pfd = g.pfd;
  auto prefetch = [&](int kk) {   // the rows above and below this thread's are requested by the neighbouring threads
    PF3(f.au, j, i, kk); PF3(f.av, j, i, kk); PF3(f.bu, j, i, kk); PF3(f.bv, j, i, kk); PF3(f.dub, j, i, kk); PF3(f.dvb, j, i, kk);
    PF3(f.qdot, j, i, kk); PF3(f.xkc, j, i, kk); PF3(f.at, j, i, kk); PF3(f.aqv, j, i, kk); PF3(f.phi, j, i, kk);
  };
  if (pfd > 0)
    for (int kk = 1; kk <= min(pfd, kz); ++kk) prefetch(kk);
  for (int k = 1; k <= kz; ++k) {
    if (pfd > 0 && k + pfd <= kz) prefetch(k + pfd);
    double au[3][3], av[3][3];
#pragma unroll
    for (int a = 0; a < 3; ++a)
#pragma unroll
      for (int b = 0; b < 3; ++b) {
        au[a][b] = L3(f.au, j + a - 1, i + b - 1, k);
        av[a][b] = L3(f.av, j + a - 1, i + b - 1, k);
      }
#define UA(a, b) (au[a][b] * md[a][b])
#define VA(a, b) (av[a][b] * md[a][b])
    const double ucmona = UA(1, 2) + 2.0 * UA(1, 1) + UA(1, 0);
    double ucmonb = UA(2, 2) + 2.0 * UA(2, 1) + UA(2, 0);
    double ucmonc = UA(0, 2) + 2.0 * UA(0, 1) + UA(0, 0);
    const double vcmona = VA(2, 1) + 2.0 * VA(1, 1) + VA(0, 1);
    double vcmonb = VA(2, 2) + 2.0 * VA(1, 2) + VA(0, 2);
    double vcmonc = VA(2, 0) + 2.0 * VA(1, 0) + VA(0, 0);
#undef UA
#undef VA
    const double u_c = au[1][1] * rp_c, v_c = av[1][1] * rp_c;
    double u_jp = au[2][1] * rp_jp, u_jm = au[0][1] * rp_jm, u_ip = au[1][2] * rp_ip, u_im = au[1][0] * rp_im;
    double v_jp = av[2][1] * rp_jp, v_jm = av[0][1] * rp_jm, v_ip = av[1][2] * rp_ip, v_im = av[1][0] * rp_im;
    if (s.iboudy == 3 || s.iboudy == 4) {   // inflow/outflow rule of `decouple` (Main/mod_tendency.F90:909-983), see k_uv_level.inc
      if (j == 2 && au[0][1] <= 0.0) { u_jm = u_c; v_jm = v_c; }
      if (j == g.jx - 1 && au[2][1] >= 0.0) { u_jp = u_c; v_jp = v_c; }
      if (i == 2 && av[1][0] >= 0.0) { u_im = u_c; v_im = v_c; }
      if (i == g.iy - 1 && av[1][2] <= 0.0) { u_ip = u_c; v_ip = v_c; }
    }
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
    double ut = 0.0 - dmsf * ((u_jp + u_c) * ucmonb - (u_c + u_jm) * ucmonc + (u_ip + u_c) * vcmonb - (u_c + u_im) * vcmonc);
    double vt = 0.0 - dmsf * ((v_jp + v_c) * ucmonb - (v_c + v_jm) * ucmonc + (v_ip + v_c) * vcmonb - (v_c + v_im) * vcmonc);
    // ---- vadvuv ----
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
    // ---- Coriolis (hydrostatic curvature, Main/mod_tendency.F90:1829-1838) ----
    ut = ut + coriol * vc;
    vt = vt - coriol * uc;
    // ---- relaxation toward the boundary fields in the lateral zone ----
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
    // ---- 4th-order diffusion (operands (ub/psdotb)/msfd precomputed by k_xkc) ----
    {
      double xkd = 0.25 * (L3(f.xkc, j, i, k) + L3(f.xkc, j - 1, i - 1, k) + L3(f.xkc, j - 1, i, k) + L3(f.xkc, j, i - 1, k));
      xkd = xkd * s.rdxsq * psdotb;
#define DU(dj, di) L3(f.dub, j + (dj), i + (di), k)
#define DV(dj, di) L3(f.dvb, j + (dj), i + (di), k)
      const double du_c = DU(0, 0), du_jm = DU(-1, 0), du_jp = DU(1, 0), du_im = DU(0, -1), du_ip = DU(0, 1);
      const double dv_c = DV(0, 0), dv_jm = DV(-1, 0), dv_jp = DV(1, 0), dv_im = DV(0, -1), dv_ip = DV(0, 1);
      if (dint) {
        const double du_jp2 = DU(2, 0), du_jm2 = DU(-2, 0), du_ip2 = DU(0, 2), du_im2 = DU(0, -2);
        const double dv_jp2 = DV(2, 0), dv_jm2 = DV(-2, 0), dv_ip2 = DV(0, 2), dv_im2 = DV(0, -2);
        ut = ut - xkd * (1.0 * (du_jp2 + du_jm2 + du_ip2 + du_im2) + (-4.0) * (du_jp + du_jm + du_ip + du_im) + 12.0 * (du_c));
        vt = vt - xkd * (1.0 * (dv_jp2 + dv_jm2 + dv_ip2 + dv_im2) + (-4.0) * (dv_jp + dv_jm + dv_ip + dv_im) + 12.0 * (dv_c));
      }
      for (int r = 0; r < nring; ++r) {
        ut = ut + xkd * (1.0 * (du_jp + du_jm + du_ip + du_im) + (-4.0) * (du_c));
        vt = vt + xkd * (1.0 * (dv_jp + dv_jm + dv_ip + dv_im) + (-4.0) * (dv_c));
      }
      if (d9) {
        ut = ut + xkd * lap9(du_c, du_jm, du_jp, du_im, du_ip, DU(1, 1), DU(-1, -1), DU(-1, 1), DU(1, -1));
        vt = vt + xkd * lap9(dv_c, dv_jm, dv_jp, dv_im, dv_ip, DV(1, 1), DV(-1, -1), DV(-1, 1), DV(1, -1));
      }
      if (d6on) { ut = ut + D6(RCM_D6_U, i, k); vt = vt + D6(RCM_D6_V, i, k); }
#undef DU
#undef DV
    }
    // ---- pressure gradient force, ipgf = 0 / 1 (Main/mod_tendency.F90:1942-1964,1996-2025,2103-2115) ----
    {
      const double hs = (LD1(&f.sigma[k + 1]) + LD1(&f.sigma[k])) * 0.5;
#define TV(jj, ii, ra) ((L3(f.at, jj, ii, k) * (ra)) * (1.0 + ep1 * fmax(L3(f.aqv, jj, ii, k) * (ra), minqq)))
      double rtbar = 0.25 * (TV(j - 1, i - 1, ra00) + TV(j - 1, i, ra01) + TV(j, i - 1, ra10) + TV(j, i, ra11));
#undef TV
      if (s.ipgf == 1) rtbar = rtbar - t00pg * pow_pos((hs * psdota + s.ptop) / p00pg, pgfaa1);   // :1945-1947
      rtbar = rgas * rtbar * psdota;
      ut = ut - div_r(rtbar * (log(0.5 * (pa11 + pa10) * hs + s.ptop) - log(0.5 * (pa01 + pa00) * hs + s.ptop)), dxm, rdxm);
      vt = vt - div_r(rtbar * (log(0.5 * (pa11 + pa01) * hs + s.ptop) - log(0.5 * (pa00 + pa10) * hs + s.ptop)), dxm, rdxm);
      const double p11 = L3(f.phi, j, i, k), p10 = L3(f.phi, j, i - 1, k), p01 = L3(f.phi, j - 1, i, k), p00 = L3(f.phi, j - 1, i - 1, k);
      ut = ut - div_r(psdota * (p11 + p10 - p01 - p00), dx2m, rdx2m);
      vt = vt - div_r(psdota * (p11 + p01 - p10 - p00), dx2m, rdx2m);
    }
    // ---- forecast p*u, p*v and Robert-Asselin filter (:433-445) ----
    const double bu_c = L3(f.bu, j, i, k), bv_c = L3(f.bv, j, i, k);
    if (ibs > 0) {   // sponge (iboudy = 4): (1 - wgtd(ib)) * bt added to the still-zero total tendency (spongeuv)
      const double xs = LD1(&f.hefd[ibs]);
      ut = ut + xs * L3(f.xubt, j, i, k); vt = vt + xs * L3(f.xvbt, j, i, k);
    }
    if (f.phy_u) { ut = ut + L3(f.phy_u, j, i, k); vt = vt + L3(f.phy_v, j, i, k); }   // Main/mod_tendency.F90:404-411
    const double cu = bu_c + s.dt * ut, cv = bv_c + s.dt * vt;
    double d = s.gnu1 * (cu + bu_c - 2.0 * uc);
    P3(f.nbu, j, i, k) = uc + d;
    P3(f.nau, j, i, k) = cu;
    d = s.gnu1 * (cv + bv_c - 2.0 * vc);
    P3(f.nbv, j, i, k) = vc + d;
    P3(f.nav, j, i, k) = cv;
    uu_k = uu_n; vv_k = vv_n;
  }
}

// ---------------------------------------------------------------------------
template <int CQ_MODE>   // StepScalars::cq_mode (negative-moisture repair), compile-time
#ifndef RCM_MB_HTQ
#define RCM_MB_HTQ 3
#endif
__global__ void __launch_bounds__(RCM_MAXT, RCM_MB_HTQ)  // 3: 168 registers, no spills (80 spilled 300 B)
k_h_tq(const Geom g, const Fields f, const StepScalars s) {
  const int si = g.si;
  const long long sk = g.sk;
  const int j = g.jci1 + blockIdx.x * blockDim.x + threadIdx.x;
  const int i = g.ici1 + blockIdx.y * blockDim.y + threadIdx.y;
  if (j > g.jci2 || i > g.ici2) return;
  if (CQ_MODE == 1 && *f.negq == 0ULL) return;  // second pass of the negative-moisture repair: nothing went negative
  const bool cq_edge = tile_edge(g, j, i);
  const bool out = (CQ_MODE != 1);              // the second pass only stores the forecast moisture
  const int kz = g.kz;
  const bool up = s.upstream_mode != 0;
  const bool se = s.stability_enhance != 0;
  const double ul = s.ul, xt = s.xt, dt = s.dt;

  const double psa = L2(f.psa, j, i), psb = L2(f.psb, j, i), psc = L2(f.psc, j, i);
  const double rdiv_psa = recip_for_div(psa);
  const double rpsa_c = L2(f.rpsa, j, i), rpsa_jm = L2(f.rpsa, j - 1, i), rpsa_jp = L2(f.rpsa, j + 1, i),
               rpsa_im = L2(f.rpsa, j, i - 1), rpsa_ip = L2(f.rpsa, j, i + 1);
  const double xmsf = L2(f.xmsf, j, i);
  const double m00 = L2(f.msfd, j, i), m01 = L2(f.msfd, j, i + 1), m10 = L2(f.msfd, j + 1, i), m11 = L2(f.msfd, j + 1, i + 1);
  const double r00 = L2(f.rpsda, j, i), r01 = L2(f.rpsda, j, i + 1), r10 = L2(f.rpsda, j + 1, i), r11 = L2(f.rpsda, j + 1, i + 1);
  const int ib = (s.iboudy == 1 || s.iboudy == 5) ? L2(f.zone_c, j, i) : 0;
  const int ibs = (s.iboudy == 4) ? L2(f.zone_c, j, i) : 0;   // sponge zone (iboudy = 4)
  const bool d9 = (s.idiffu == 2);
  const bool d6on = (s.idiffu == 3) && j == g.jci2;   // the one column idiffu = 3 acts on (k_diff6)   // nine-point operator on every interior point, no boundary ring
  const bool dint = (s.idiffu == 1) && diff_interior(g, j, i, false);
  const int nring = (s.idiffu == 1) ? ring_count(g, j, i, false) : 0;
  const double beta = 0.53;
  // p* after filter_ra_2d (Main/mod_tendency.F90:420): floors of the qv filter
  const double dps = s.gnu1 * (psc + psb - 2.0 * psa);
  const double npsb = psa + dps, npsa = psc;
  const double qthr = minqq * psa;          // vadvqv threshold (ps = sfs%psa before the filter)
  const double cthr = minqq * minqq * psa;  // vadv4d ind=1 threshold
  const int kpb_c = f.kpb ? L2(f.kpb, j, i) : 0;   // top of the PBL for vadv4d ind = 3
  // omega (:1199-1215)
  const double pten = L2(f.pten, j, i);
  const double dummy8 = 1.0 / (s.dx8 * L2(f.msfx, j, i));
  const double dpsx = L2(f.psa, j + 1, i) - L2(f.psa, j - 1, i), dpsy = L2(f.psa, j, i + 1) - L2(f.psa, j, i - 1);
#define RPSB(dj, di) L2(f.rpsb, j + (dj), i + (di))

  double dq_k = 0.0, fl_qv_k = 0.0, fl_qc_k = 0.0;  // vertical fluxes at full level k (zero at k = 1)
  double qdot_k = 0.0;
  const int pfd = g.pfd;
  auto prefetch = [&](int kk) {
    PF3(f.au, j, i, kk); PF3(f.av, j, i, kk); PF3(f.au, j, i + 1, kk); PF3(f.av, j, i + 1, kk); PF3(f.qdot, j, i, kk); PF3(f.xkc, j, i, kk);
    PF3(f.at, j, i, kk); PF3(f.bt, j, i, kk); PF3(f.aqv, j, i, kk); PF3(f.bqv, j, i, kk); PF3(f.aqc, j, i, kk); PF3(f.bqc, j, i, kk);
  };
  if (pfd > 0)
    for (int kk = 1; kk <= min(pfd, kz); ++kk) prefetch(kk);
  for (int k = 1; k <= kz; ++k) {
    if (pfd > 0 && k + pfd <= kz) prefetch(k + pfd);
    const double au00 = L3(f.au, j, i, k), au01 = L3(f.au, j, i + 1, k), au10 = L3(f.au, j + 1, i, k), au11 = L3(f.au, j + 1, i + 1, k);
    const double av00 = L3(f.av, j, i, k), av01 = L3(f.av, j, i + 1, k), av10 = L3(f.av, j + 1, i, k), av11 = L3(f.av, j + 1, i + 1, k);
    const double ua00 = au00 * m00, ua01 = au01 * m01, ua10 = au10 * m10, ua11 = au11 * m11;
    const double va00 = av00 * m00, va01 = av01 * m01, va10 = av10 * m10, va11 = av11 * m11;
    const double uavg1 = ua01 + ua00, uavg2 = ua11 + ua10;
    const double vavg1 = va10 + va00, vavg2 = va11 + va01;
    const double f1 = div_r(0.5 * ul * (uavg2 + uavg1), psa, rdiv_psa);
    const double f2 = div_r(0.5 * ul * (vavg2 + vavg1), psa, rdiv_psa);
    const double xdsk = LD1(&f.xds[k]);
    const double svv_n = (k < kz) ? L3(f.qdot, j, i, k + 1) : 0.0;
    const double hs = (LD1(&f.sigma[k + 1]) + LD1(&f.sigma[k])) * 0.5;
    double xf = 0.0, xg = 0.0;
    if (ib > 0) { xf = LD1(&f.hefc[ib + s.nspgx * (k - 1)]); xg = LD1(&f.hegc[ib + s.nspgx * (k - 1)]); }
    double xkc = L3(f.xkc, j, i, k);
    xkc = xkc * s.rdxsq * psb;
    const double aq_c = L3(f.aqv, j, i, k);
    const double qv_c = fmax(aq_c * rpsa_c, minqq);

    // =================== temperature ===================
    {
      const double at = L3(f.at, j, i, k);
      const double t_c = at * rpsa_c, t_jm = L3(f.at, j - 1, i, k) * rpsa_jm, t_jp = L3(f.at, j + 1, i, k) * rpsa_jp,
                   t_im = L3(f.at, j, i - 1, k) * rpsa_im, t_ip = L3(f.at, j, i + 1, k) * rpsa_ip;
      double fg = hflux(up, f1, f2, xmsf, uavg1, uavg2, vavg1, vavg2, t_c, t_jm, t_jp, t_im, t_ip);
      if (se) fg = extrema_limit(fg, t_c, t_jm, t_jp, t_im, t_ip, psa, rdiv_psa, s.t_extrema);
      double ten = 0.0 + fg;
      // vadv3d ind = 1: potential-temperature-like interpolation to full levels, level-b pressures
      if (k >= 2) ten = ten + dq_k * xdsk;
      double dq_n = 0.0;
      if (k < kz) {
        const int kn = k + 1;
        const double pfs = (LD1(&f.sigma[kn]) * psb + s.ptop) * 1000.0;
        const double hs_n = (LD1(&f.sigma[kn + 1]) + LD1(&f.sigma[kn])) * 0.5;
        const double phs_n = (hs_n * psb + s.ptop) * 1000.0, phs_k = (hs * psb + s.ptop) * 1000.0;
        dq_n = svv_n * (LD1(&f.twt1[kn]) * L3(f.at, j, i, kn) * pow_pos(pfs / phs_n, c287) +
                        LD1(&f.twt2[kn]) * at * pow_pos(pfs / phs_k, c287));
        ten = ten - dq_n * xdsk;
      }
      // adiabatic term with omega evaluated in place
      {
        const double tv = t_c * (1.0 + ep1 * qv_c);
        const double omega = 0.5 * (svv_n + qdot_k) * psa +
                             hs * (pten + ((au00 * r00 + au01 * r01 + au11 * r11 + au10 * r10) * dpsx +
                                           (av00 * r00 + av01 * r01 + av11 * r11 + av10 * r10) * dpsy) * dummy8);
        if (out) P3(f.omega, j, i, k) = omega;
        const double rovcpm = rgas / cpmf(qv_c);
        ten = ten + (omega * rovcpm * tv) / (s.ptop * rpsa_c + hs);
      }
      const double bt_c = L3(f.bt, j, i, k);
      const double bt_jm = L3(f.bt, j - 1, i, k), bt_jp = L3(f.bt, j + 1, i, k), bt_im = L3(f.bt, j, i - 1, k), bt_ip = L3(f.bt, j, i + 1, k);
      if (ib > 0) {
#define FG(jj, ii, bval) ((L3(f.xtb0, jj, ii, k) + xt * L3(f.xtbt, jj, ii, k)) - (bval))
        const double fls0 = FG(j, i, bt_c), fls1 = FG(j - 1, i, bt_jm), fls2 = FG(j + 1, i, bt_jp), fls3 = FG(j, i - 1, bt_im), fls4 = FG(j, i + 1, bt_ip);
#undef FG
        ten = ten + xf * fls0 - xg * (fls1 + fls2 + fls3 + fls4 - 4.0 * fls0);
      }
      {
        const double d_c = bt_c * RPSB(0, 0), d_jm = bt_jm * RPSB(-1, 0), d_jp = bt_jp * RPSB(1, 0),
                     d_im = bt_im * RPSB(0, -1), d_ip = bt_ip * RPSB(0, 1);
        if (dint) {
          const double d_jp2 = L3(f.bt, j + 2, i, k) * RPSB(2, 0), d_jm2 = L3(f.bt, j - 2, i, k) * RPSB(-2, 0);
          const double d_ip2 = L3(f.bt, j, i + 2, k) * RPSB(0, 2), d_im2 = L3(f.bt, j, i - 2, k) * RPSB(0, -2);
          ten = ten - xkc * (1.0 * (d_jp2 + d_jm2 + d_ip2 + d_im2) + (-4.0) * (d_jp + d_jm + d_ip + d_im) + 12.0 * (d_c));
        }
        for (int r = 0; r < nring; ++r) ten = ten + xkc * (1.0 * (d_jp + d_jm + d_ip + d_im) + (-4.0) * (d_c));
        if (d9) ten = ten + xkc * lap9(d_c, d_jm, d_jp, d_im, d_ip, L3(f.bt, j + 1, i + 1, k) * RPSB(1, 1), L3(f.bt, j - 1, i - 1, k) * RPSB(-1, -1),
                                       L3(f.bt, j - 1, i + 1, k) * RPSB(-1, 1), L3(f.bt, j + 1, i - 1, k) * RPSB(1, -1));
        if (d6on) ten = ten + D6(RCM_D6_T, i, k);
      }
      if (ibs > 0) ten = ten + LD1(&f.hefc[ibs]) * L3(f.xtbt, j, i, k);   // sponge3d on tten (iboudy = 4)
      if (f.phy_t) ten = ten + L3(f.phy_t, j, i, k);   // Main/mod_tendency.F90:285-287
      const double ct = bt_c + dt * ten;
      const double d = s.gnu1 * (ct + bt_c - 2.0 * at);
      if (out) P3(f.nbt, j, i, k) = at + d;
      if (out) P3(f.nat, j, i, k) = ct;
      dq_k = dq_n;
    }

    // =================== water vapour ===================
    {
      const double q_c = qv_c, q_jm = fmax(L3(f.aqv, j - 1, i, k) * rpsa_jm, minqq),
                   q_jp = fmax(L3(f.aqv, j + 1, i, k) * rpsa_jp, minqq), q_im = fmax(L3(f.aqv, j, i - 1, k) * rpsa_im, minqq),
                   q_ip = fmax(L3(f.aqv, j, i + 1, k) * rpsa_ip, minqq);
      double fg;
      if (s.isladvec == 1) {   // slhadv_x(qxdyn, atm2%qx, iqv); hdvg_x(qxdyn, atm1%qx, iqv) (Main/mod_tendency.F90:1361-1363)
        fg = L3(f.slqv, j, i, k);   // left by k_sl
      } else {
        fg = hflux(up, f1, f2, xmsf, uavg1, uavg2, vavg1, vavg2, q_c, q_jm, q_jp, q_im, q_ip);
        if (se) { const double qs = fmax(q_c, dlowval); fg = extrema_limit(fg, q_c, q_jm, q_jp, q_im, q_ip, qs, recip_for_div(qs), s.q_rel_extrema); }
      }
      double ten = fg;
      if (k >= 2) ten = ten + fl_qv_k * xdsk;
      double fl_n = 0.0;
      if (k < kz) {
        const double aqv_n = L3(f.aqv, j, i, k + 1);
        double fgv = 0.0;
        if (aqv_n > qthr && aq_c > qthr) fgv = aqv_n * pow_pos(aq_c / aqv_n, LD1(&f.qcon[k + 1]));
        fl_n = svv_n * fgv;
        ten = ten - fl_n * xdsk;
      }
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
        const double d_c = fmax(b_c * RPSB(0, 0), minqq), d_jm = fmax(b_jm * RPSB(-1, 0), minqq),
                     d_jp = fmax(b_jp * RPSB(1, 0), minqq), d_im = fmax(b_im * RPSB(0, -1), minqq),
                     d_ip = fmax(b_ip * RPSB(0, 1), minqq);
        if (dint) {
          const double d_jp2 = fmax(L3(f.bqv, j + 2, i, k) * RPSB(2, 0), minqq), d_jm2 = fmax(L3(f.bqv, j - 2, i, k) * RPSB(-2, 0), minqq);
          const double d_ip2 = fmax(L3(f.bqv, j, i + 2, k) * RPSB(0, 2), minqq), d_im2 = fmax(L3(f.bqv, j, i - 2, k) * RPSB(0, -2), minqq);
          ten = ten - 1.0 * xkc * (1.0 * (d_jp2 + d_jm2 + d_ip2 + d_im2) + (-4.0) * (d_jp + d_jm + d_ip + d_im) + 12.0 * d_c);
        }
        for (int r = 0; r < nring; ++r) ten = ten + 1.0 * xkc * (1.0 * (d_jp + d_jm + d_ip + d_im) + (-4.0) * d_c);
        if (d9) ten = ten + 1.0 * xkc * lap9(d_c, d_jm, d_jp, d_im, d_ip, fmax(L3(f.bqv, j + 1, i + 1, k) * RPSB(1, 1), minqq), fmax(L3(f.bqv, j - 1, i - 1, k) * RPSB(-1, -1), minqq),
                                             fmax(L3(f.bqv, j - 1, i + 1, k) * RPSB(-1, 1), minqq), fmax(L3(f.bqv, j + 1, i - 1, k) * RPSB(1, -1), minqq));
        if (d6on) ten = ten + D6(RCM_D6_QV, i, k);
      }
      if (ibs > 0) ten = ten + LD1(&f.hefc[ibs]) * L3(f.xqbt, j, i, k);   // sponge4d on qxten(iqv)
      if (f.phy_qv) ten = ten + L3(f.phy_qv, j, i, k);   // :292-294
      double cq = b_c + dt * ten;
      // Main/mod_tendency.F90:381-393: a negative forecast is replaced by 1 % of the 3x3 mean of |q|; the value
      // filtered here for such a point is provisional and is redone by k_negq_repair (k_negq.cu)
      negq_note<CQ_MODE>(g, f, cq, j, i, k, 0, cq_edge);
      if (cq < 0.0) cq = 0.0;
      const double d = s.gnu1 * (cq + b_c - 2.0 * aq_c);
      if (out) P3(f.nbqv, j, i, k) = fmax(aq_c + beta * d, minqq * npsa);
      if (out) P3(f.naqv, j, i, k) = fmax(cq + (beta - 1.0) * d, minqq * npsb);
      fl_qv_k = fl_n;
    }

    // =================== cloud water ===================
    {
      const double ac_c = L3(f.aqc, j, i, k);
      const double q_c = fmax(ac_c * rpsa_c, 0.0);
      double ten = 0.0;
      double fl_n = 0.0;
      if (s.ipptls > 0) {
        const double q_jm = fmax(L3(f.aqc, j - 1, i, k) * rpsa_jm, 0.0), q_jp = fmax(L3(f.aqc, j + 1, i, k) * rpsa_jp, 0.0),
                     q_im = fmax(L3(f.aqc, j, i - 1, k) * rpsa_im, 0.0), q_ip = fmax(L3(f.aqc, j, i + 1, k) * rpsa_ip, 0.0);
        if (s.isladvec == 1) ten = L3(f.slqc, j, i, k);   // :1378-1380, left by k_sl
        else ten = hflux(up, f1, f2, xmsf, uavg1, uavg2, vavg1, vavg2, q_c, q_jm, q_jp, q_im, q_ip);
        if (k >= 2) ten = ten + fl_qc_k * xdsk;
        if (k < kz) {
          const double aqc_n = L3(f.aqc, j, i, k + 1);
          const double t1 = LD1(&f.twt1[k + 1]), t2 = LD1(&f.twt2[k + 1]);
          if (f.kpb) {   // iqxvadv = 3 (Main/mod_tendency.F90:150-153)
            fl_n = vadv3_face(g, f, f.aqc, j, i, k + 1, kpb_c, aqc_n, ac_c) * svv_n;
          } else if (svv_n > 0.0) {
            if (ac_c > cthr) fl_n = svv_n * (t1 * aqc_n + t2 * ac_c);
          } else {
            if (aqc_n > cthr) fl_n = svv_n * (t1 * aqc_n + t2 * ac_c);
          }
          ten = ten - fl_n * xdsk;
        }
      }
      const double b_c = L3(f.bqc, j, i, k);
      {
        const double d_c = fmax(b_c * RPSB(0, 0), 0.0), d_jm = fmax(L3(f.bqc, j - 1, i, k) * RPSB(-1, 0), 0.0),
                     d_jp = fmax(L3(f.bqc, j + 1, i, k) * RPSB(1, 0), 0.0), d_im = fmax(L3(f.bqc, j, i - 1, k) * RPSB(0, -1), 0.0),
                     d_ip = fmax(L3(f.bqc, j, i + 1, k) * RPSB(0, 1), 0.0);
        if (dint) {
          const double d_jp2 = fmax(L3(f.bqc, j + 2, i, k) * RPSB(2, 0), 0.0), d_jm2 = fmax(L3(f.bqc, j - 2, i, k) * RPSB(-2, 0), 0.0);
          const double d_ip2 = fmax(L3(f.bqc, j, i + 2, k) * RPSB(0, 2), 0.0), d_im2 = fmax(L3(f.bqc, j, i - 2, k) * RPSB(0, -2), 0.0);
          ten = ten - 1.0 * xkc * (1.0 * (d_jp2 + d_jm2 + d_ip2 + d_im2) + (-4.0) * (d_jp + d_jm + d_ip + d_im) + 12.0 * d_c);
        }
        for (int r = 0; r < nring; ++r) ten = ten + 1.0 * xkc * (1.0 * (d_jp + d_jm + d_ip + d_im) + (-4.0) * d_c);
        if (d9) ten = ten + 1.0 * xkc * lap9(d_c, d_jm, d_jp, d_im, d_ip, fmax(L3(f.bqc, j + 1, i + 1, k) * RPSB(1, 1), 0.0), fmax(L3(f.bqc, j - 1, i - 1, k) * RPSB(-1, -1), 0.0),
                                             fmax(L3(f.bqc, j - 1, i + 1, k) * RPSB(-1, 1), 0.0), fmax(L3(f.bqc, j + 1, i - 1, k) * RPSB(1, -1), 0.0));
        if (d6on) ten = ten + D6(RCM_D6_QC, i, k);
      }
      if (f.phy_qc) ten = ten + L3(f.phy_qc, j, i, k);
      if (s.ipptls <= 0) ten = 0.0;  // Main/mod_tendency.F90:331-335
      double cq = b_c + dt * ten;
      negq_note<CQ_MODE>(g, f, cq, j, i, k, 1, cq_edge);
      if (cq < 0.0) cq = 0.0;
      const double d = s.gnu2 * (cq + b_c - 2.0 * ac_c);
      double nb = ac_c + beta * d;
      double na = cq + (beta - 1.0) * d;
      if (nb < 0.0) nb = 0.0;
      if (na < 0.0) na = 0.0;
      if (out) P3(f.nbqc, j, i, k) = nb;
      if (out) P3(f.naqc, j, i, k) = na;
      fl_qc_k = fl_n;
    }
    qdot_k = svv_n;
  }
#undef RPSB
}

// ---------------------------------------------------------------------------
// filter_ra_2d(psa, psb, psc) on the interior (Main/mod_timefilter.F90:58-73)
__global__ void k_h_psfilt(const Geom g, const Fields f, const StepScalars s) {
  const int si = g.si;
  // + one ghost ring towards neighbour ranks (psc2psd of the filtered p* reads it; see k_h_prep)
  const int j = (g.jci1 - g.gl) + blockIdx.x * blockDim.x + threadIdx.x;
  const int i = (g.ici1 - g.gb) + blockIdx.y * blockDim.y + threadIdx.y;
  if (j > g.jci2 + g.gr || i > g.ici2 + g.gt) return;
  const double psa = P2(f.psa, j, i), psb = P2(f.psb, j, i), psc = P2(f.psc, j, i);
  const double d = s.gnu1 * (psc + psb - 2.0 * psa);
  P2(f.psb, j, i) = psa + d;
  P2(f.psa, j, i) = psc;
}

// p* rim of bdyval (Main/mod_bdycod.F90:1317-1342,1426-1451)
__global__ void k_bdyval_ps(const Geom g, const Fields f, const int2* rim, int nrim, double xt, int iboudy) {
  const int si = g.si;
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= nrim) return;
  const int j = rim[p].x, i = rim[p].y;
  // the left/right loops run over ici1:ici2, bottom/top over jce1:jce2: every rim point is covered
  P2(f.psa, j, i) = (iboudy == 0) ? P2(f.xpsb0, j, i) : P2(f.xpsb0, j, i) + xt * P2(f.xpsbt, j, i);
}

// ---------------------------------------------------------------------------
static dim3 grid2(int nj, int ni, dim3 b) { return dim3((nj + b.x - 1) / b.x, (ni + b.y - 1) / b.y); }

void launch_psc2psd(const Geom& g, const double* pc, double* pd, double* rpd, int gw, cudaStream_t st) {
  dim3 b(32, 8);
  RCM_LAUNCH(k_psc2psd, grid2(g.jde2 - g.jde1 + 1 + gw * g.gr, g.ide2 - g.ide1 + 1 + gw * g.gt, b), b, st, g, pc, pd, rpd, gw);
}
void launch_ps_convert(const Geom& g, double* ps, double ptop, cudaStream_t st) {
  dim3 b(32, 8);
  RCM_LAUNCH(k_ps_convert, grid2(g.jce2 - g.jce1 + 1 + RCM_HALO * (g.gl + g.gr), g.ice2 - g.ice1 + 1 + RCM_HALO * (g.gb + g.gt), b), b, st, g, ps, ptop);
}
void launch_h_ps(const Geom& g, const Fields& f, cudaStream_t st) {
  dim3 b(32, 8);
  RCM_LAUNCH(k_h_ps, grid2(g.jde2 - g.jde1 + 1 + 2 * (g.gl + g.gr), g.ide2 - g.ide1 + 1 + 2 * (g.gb + g.gt), b), b, st, g, f);
}
void launch_h_rimcopy(const Geom& g, const Fields& f, const int2* rim_d, int nrim_d, const int2* rim_c, int nrim_c, cudaStream_t st) {
  const int n = nrim_d > nrim_c ? nrim_d : nrim_c;
  if (n <= 0) return;
  dim3 b(128), gr((n + 127) / 128, g.kz);
  RCM_LAUNCH(k_h_rimcopy, gr, b, st, g, f, rim_d, nrim_d, rim_c, nrim_c);
}
void launch_h_prep(const Geom& g, const Fields& f, const StepScalars& s, cudaStream_t st) {
  const dim3 b = rcm::col_block();
  RCM_LAUNCH(k_h_prep, grid2(g.jce2 - g.jce1 + 1 + g.gl + g.gr, g.ice2 - g.ice1 + 1 + g.gb + g.gt, b), b, st, g, f, s);
}
void launch_h_uv(const Geom& g, const Fields& f, const StepScalars& s, cudaStream_t st) {
  const dim3 b = rcm::col_block();
  RCM_LAUNCH(k_h_uv, grid2(g.jdi2 - g.jdi1 + 1, g.idi2 - g.idi1 + 1, b), b, st, g, f, s);
}
void launch_h_tq(const Geom& g, const Fields& f, const StepScalars& s, cudaStream_t st) {
  const dim3 b = rcm::col_block();
  const dim3 gr = grid2(g.jci2 - g.jci1 + 1, g.ici2 - g.ici1 + 1, b);
  if (s.cq_mode == 0) RCM_LAUNCH(k_h_tq<0>, gr, b, st, g, f, s);
  else if (s.cq_mode == 1) RCM_LAUNCH(k_h_tq<1>, gr, b, st, g, f, s);
  else RCM_LAUNCH(k_h_tq<2>, gr, b, st, g, f, s);
}
void launch_h_psfilt(const Geom& g, const Fields& f, const StepScalars& s, cudaStream_t st) {
  dim3 b(32, 8);
  RCM_LAUNCH(k_h_psfilt, grid2(g.jci2 - g.jci1 + 1 + g.gl + g.gr, g.ici2 - g.ici1 + 1 + g.gb + g.gt, b), b, st, g, f, s);
}
void launch_bdyval_ps(const Geom& g, const Fields& f, const int2* rim_c, int nrim_c, double xt, int iboudy, cudaStream_t st) {
  if (nrim_c <= 0) return;
  dim3 b(128), gr((nrim_c + 127) / 128);
  RCM_LAUNCH(k_bdyval_ps, gr, b, st, g, f, rim_c, nrim_c, xt, iboudy);
}
