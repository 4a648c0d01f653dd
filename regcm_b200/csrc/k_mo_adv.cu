// MOLOCH core (idynamic = 3): advection step.
//   k_mo_destag  uvstagtox + wstagtox                         Main/mod_moloch.F90:1739-1794,1574-1597
//   k_mo_wafz / k_mo_wafy / k_mo_wafx   wafone (WAF/TVD, directional split z -> y -> x), both projection branches
//                                                             :1098-1181,1274-1358
//   k_mo_restag  xtouvstag + xtowstag                         :1680-1737,1599-1618
// The reference calls wafone once per advected field (tetav, pai, ux, vx, wx, qv[, qc]); the advecting
// velocities, the metric ratios and the upwind selection do not depend on the field, so here one thread
// computes them once per point and loops over the fields (RCM_MO_MAXF pointers in MoFields::adv).
// Per field the arithmetic and its order are the reference's.
#include "dev_common.cuh"
#include "launch.h"

using namespace rcm;

// min/max with the selection rule of the reference's host compilers for ordinary numbers
// ((a < b) ? b : a): a compare and two selects instead of the NaN-propagating fmin/fmax sequences
__device__ __forceinline__ double mo_max(const double a, const double b) { return (a < b) ? b : a; }
__device__ __forceinline__ double mo_min(const double a, const double b) { return (b < a) ? b : a; }
__device__ __forceinline__ double mo_rdeno(const double t1, const double t2, const double t3, const double t4) {
  const double zzden = (t3 - t4);
  return (t1 - t2) / copysign(mo_max(fabs(zzden), 1.0e-15), zzden);
}
__device__ __forceinline__ double mo_phi(const double r, const double is, const double zamu) {
  const double b = mo_max(0.0, mo_min(2.0, mo_max(r, mo_min(2.0 * r, 1.0))));
  return is + zamu * b - is * b;
}

__global__ void k_mo_destag(const Geom g, const MoFields f) {
  const int si = g.si; const long long sk = g.sk;
  const int j = g.jce1 + blockIdx.x * blockDim.x + threadIdx.x;
  const int i = g.ice1 + blockIdx.y * blockDim.y + threadIdx.y;
  const int k = 1 + blockIdx.z;
  if (j > g.jce2 || i > g.ice2) return;
  const int kz = g.kz;
  double ux;
  if (j >= g.jci1 && j <= g.jci2) ux = 0.5625 * (L3(f.u, j + 1, i, k) + L3(f.u, j, i, k)) - 0.0625 * (L3(f.u, j + 2, i, k) + L3(f.u, j - 1, i, k));
  else if (j == g.jce1) ux = 0.5 * (L3(f.u, g.jde1, i, k) + L3(f.u, g.jdi1, i, k));
  else ux = 0.5 * (L3(f.u, g.jde2, i, k) + L3(f.u, g.jdi2, i, k));
  P3(f.ux, j, i, k) = ux;
  double vx;
  if (i >= g.ici1 && i <= g.ici2) vx = 0.5625 * (L3(f.v, j, i + 1, k) + L3(f.v, j, i, k)) - 0.0625 * (L3(f.v, j, i + 2, k) + L3(f.v, j, i - 1, k));
  else if (i == g.ice1) vx = 0.5 * (L3(f.v, j, g.ide1, k) + L3(f.v, j, g.idi1, k));
  else vx = 0.5 * (L3(f.v, j, g.ide2, k) + L3(f.v, j, g.idi2, k));
  P3(f.vx, j, i, k) = vx;
  double wx;
  if (k >= 2 && k <= kz - 1) wx = 0.5625 * (L3(f.w, j, i, k + 1) + L3(f.w, j, i, k)) - 0.0625 * (L3(f.w, j, i, k + 2) + L3(f.w, j, i, k - 1));
  else if (k == 1) wx = 0.5 * (L3(f.w, j, i, 2) + L3(f.w, j, i, 1));
  else wx = 0.5 * (L3(f.w, j, i, kz + 1) + L3(f.w, j, i, kz));
  P3(f.wx, j, i, k) = wx;
}

// vertical pass: one thread marches k once for a group of up to NF fields of one interior column
// (blockIdx.z selects the group: fields z*NF .. z*NF+NF-1); the 4-level window pp(k-1..k+2) of every
// field lives in registers, the interface velocity and the metric ratios are evaluated once per level
#ifndef RCM_MB_WAFZ
#define RCM_MB_WAFZ 3
#endif
#ifndef RCM_MB_WAFH
#define RCM_MB_WAFH 4
#endif
template <int NF>
__global__ void __launch_bounds__(RCM_MAXT, RCM_MB_WAFZ) k_mo_wafz(const Geom g, const MoFields f, const MoScalars s) {
  const int si = g.si; const long long sk = g.sk;
  const int j = g.jci1 + blockIdx.x * blockDim.x + threadIdx.x;
  const int i = g.ici1 + blockIdx.y * blockDim.y + threadIdx.y;
  if (j > g.jci2 || i > g.ici2) return;
  const int kz = g.kz;
  const double zdtrdz = s.adtrdz;
  const bool rs = (i == g.ici1), rn = (i == g.ici2);
  const int n0 = blockIdx.z * NF;
  const double* adv[NF]; double* wzp[NF];
#pragma unroll
  for (int n = 0; n < NF; ++n) { const int q = min(n0 + n, s.nf - 1); adv[n] = f.adv[q]; wzp[n] = f.wz[q]; }   // a short last group repeats its last field
  double pkm1[NF], pk[NF], pkp1[NF], wfw_k[NF];
#pragma unroll
  for (int n = 0; n < NF; ++n) {
    pkm1[n] = 0.0; wfw_k[n] = 0.0;
    pk[n] = L3(adv[n], j, i, 1); pkp1[n] = L3(adv[n], j, i, 2);
  }
  double s_k = 0.0;
  double fmzf_k = 0.0;   // fmzf(k); not used for k = 1
  const int pfd = g.pfd;   // L2 prefetch distance beyond the two levels already held in registers (PF3, dev_common.cuh)
  auto prefetch = [&](int kk) {
#pragma unroll
    for (int n = 0; n < NF; ++n) PF3(adv[n], j, i, kk + 2);
    PF3(f.s, j, i, kk + 1); PF3(f.fmz, j, i, kk); PF3(f.fmzf, j, i, kk + 1);
  };
  if (pfd > 0)
    for (int kk = 1; kk <= min(pfd, kz - 2); ++kk) prefetch(kk);
  for (int k = 1; k <= kz; ++k) {
    if (pfd > 0 && k + pfd <= kz - 2) prefetch(k + pfd);
    double pkp2[NF];
#pragma unroll
    for (int n = 0; n < NF; ++n) pkp2[n] = (k + 2 <= kz) ? L3(adv[n], j, i, k + 2) : 0.0;
    const double s_kp1 = (k <= kz - 1) ? L3(f.s, j, i, k + 1) : 0.0;
    const double fmz = L3(f.fmz, j, i, k);
    const double fmzf_kp1 = L3(f.fmzf, j, i, k + 1);
    const double zamu = s_kp1 * zdtrdz;
    const bool up = (zamu >= 0.0);
    const double is = up ? 1.0 : -1.0;
    const double zrfmm = zdtrdz * fmz / fmzf_kp1;
    const double zrfmp = (k > 1) ? zdtrdz * fmz / fmzf_k : 0.0;
#pragma unroll
    for (int n = 0; n < NF; ++n) {
      double wfw_kp1 = 0.0;
      if (k <= kz - 1) {
        double p1, p1m1;
        if (up) { p1 = pkp1[n]; p1m1 = (k + 2 > kz) ? pkp1[n] : pkp2[n]; }   // k1 = k+1, k1m1 = min(k+2, kz)
        else { p1m1 = pk[n]; p1 = (k - 1 < 1) ? pk[n] : pkm1[n]; }           // k1m1 = k, k1 = max(k-1, 1)
        const double r = mo_rdeno(p1, p1m1, pk[n], pkp1[n]);
        const double zphi = mo_phi(r, is, zamu);
        wfw_kp1 = 0.5 * s_kp1 * ((1.0 + zphi) * pkp1[n] + (1.0 - zphi) * pk[n]);
      }
      double out;
      if (k == 1) {
        const double zdv = -s_kp1 * zrfmm * pk[n];
        out = pk[n] + wfw_kp1 * zrfmm + zdv;
      } else {
        const double zdv = (s_k * zrfmp - s_kp1 * zrfmm) * pk[n];
        out = pk[n] - wfw_k[n] * zrfmp + wfw_kp1 * zrfmm + zdv;
      }
      double* wz = wzp[n];
      P3(wz, j, i, k) = out;
      if (rs) P3(wz, j, g.ice1, k) = out;
      if (rn) P3(wz, j, g.ice2, k) = out;
      wfw_k[n] = wfw_kp1;
      pkm1[n] = pk[n]; pk[n] = pkp1[n]; pkp1[n] = pkp2[n];
    }
    s_k = s_kp1; fmzf_k = fmzf_kp1;
  }
}

// One horizontal pass (DIR 0: meridional, wz -> p0; DIR 1: zonal, p0 -> the field itself, in place: each
// thread reads only its own point of it).  The upwind choice, the face velocities, the metric ratios and
// all element offsets are field-independent and computed once; the field loop is unrolled (NF is a
// template parameter) so that the per-field work is loads at precomputed offsets plus the limiter.
template <int NF, int DIR, bool ROT>
__global__ void __launch_bounds__(RCM_MAXT, RCM_MB_WAFH) k_mo_wafh(const Geom g, const MoFields f, const MoScalars s) {
  const int si = g.si; const long long sk = g.sk;
  const int j = g.jci1 + blockIdx.x * blockDim.x + threadIdx.x;
  const int i = g.ici1 + blockIdx.y * blockDim.y + threadIdx.y;
  const int k = 1 + blockIdx.z;
  if (j > g.jci2 || i > g.ici2) return;
  const long long st = DIR ? 1 : (long long)si;              // element stride along the advection direction
  const long long o = (long long)j + (long long)i * si + (long long)k * sk;
  const int c = DIR ? j : i, c1 = DIR ? g.jcross1 : g.icross1, c2 = DIR ? g.jcross2 : g.icross2;
  const double zdtrd = DIR ? s.adtrdx : s.adtrdy;
  const double* vel = DIR ? f.u : f.v;
  const double* rm = DIR ? f.rmu : f.rmv;
  const long long o2 = (long long)j + (long long)i * si;
  // the two faces c (low side) and c+1 (high side)
  double vr[2], zamu[2], is[2];
  long long oh[2], ohm1[2];
#pragma unroll
  for (int m = 0; m < 2; ++m) {
    const int cc = c + m;
    if (!ROT) {
      vr[m] = LD1(vel + o + m * st) * LD1(rm + o2 + m * st);
      zamu[m] = vr[m] * zdtrd;
    } else {   // ROTLLR (:1190-1203, 1243-1256): the flux carries the wind itself; zonal Courant number with mu
      vr[m] = LD1(vel + o + m * st);
      zamu[m] = DIR ? vr[m] * LD1(f.msfu + o2 + m) * zdtrd : vr[m] * zdtrd;
    }
    int ch;
    if (zamu[m] > 0.0) { is[m] = 1.0; ch = cc - 1; }
    else { is[m] = -1.0; ch = min(cc + 1, c2); }
    const int chm1 = max(ch - 1, c1);
    oh[m] = o + (long long)(ch - c) * st; ohm1[m] = o + (long long)(chm1 - c) * st;
  }
  const double fmz = LD1(f.fmz + o);
  // metric ratios: zdtrd (:1298-1299,1350-1351); ROTLLR: zhxvtn / zhxvts (:1209-1212), zcostx (:1262-1264)
  const double chi = !ROT ? zdtrd : (DIR ? zdtrd * LD1(f.msfu + o2) : zdtrd * LD1(f.rmv + o2 + si) * LD1(f.mx + o2));
  const double clo = !ROT ? zdtrd : (DIR ? zdtrd * LD1(f.msfu + o2) : zdtrd * LD1(f.rmv + o2) * LD1(f.mx + o2));
  const double zrfhi = chi * 2.0 * fmz / (fmz + LD1(f.fmz + o + st));   // zrfmn / zrfme
  const double zrflo = clo * 2.0 * fmz / (fmz + LD1(f.fmz + o - st));   // zrfms / zrfmw
  const double mx2 = ROT ? 0.0 : LD1(f.mx2 + o2);
  const double dvf = (vr[1] * zrfhi - vr[0] * zrflo);
  const bool cw = (!DIR) && (j == g.jci1), ce = (!DIR) && (j == g.jci2);
  const long long ow = (long long)g.jce1 + (long long)i * si + (long long)k * sk, oe = ow + (g.jce2 - g.jce1);
#pragma unroll
  for (int n = 0; n < NF; ++n) {
    const double* src = DIR ? f.p0[n] : f.wz[n];
    const double wm0 = LD1(src + o - st), wc0 = LD1(src + o), wc1 = LD1(src + o + st);
    double flux[2];
#pragma unroll
    for (int m = 0; m < 2; ++m) {
      const double wc = m ? wc1 : wc0, wm = m ? wc0 : wm0;
      const double r = mo_rdeno(LD1(src + oh[m]), LD1(src + ohm1[m]), wc, wm);
      const double zphi = mo_phi(r, is[m], zamu[m]);
      flux[m] = 0.5 * vr[m] * ((1.0 + zphi) * wm + (1.0 - zphi) * wc);
    }
    double* pp = f.adv[n];
    const double zdv = dvf * pp[o];
    const double out = ROT ? (wc0 + flux[0] * zrflo - flux[1] * zrfhi + zdv)
                           : (wc0 + mx2 * (flux[0] * zrflo - flux[1] * zrfhi + zdv));
    if (DIR) {
      pp[o] = out;
    } else {
      double* p0 = f.p0[n];
      p0[o] = out;
      if (cw) p0[ow] = out;
      if (ce) p0[oe] = out;
    }
  }
}

// Variants of the horizontal passes in which one thread owns RCM_MO_NC consecutive cells along the
// advection direction: the NC+1 face fluxes are evaluated once each (the one-cell kernels above evaluate
// every face twice, once from either side), and the 4-point stencils of neighbouring cells share loads.
#define RCM_MO_NC 4
__global__ void __launch_bounds__(RCM_MAXT, 3) k_mo_wafy4(const Geom g, const MoFields f, const MoScalars s) {
  const int si = g.si; const long long sk = g.sk;
  const int j = g.jci1 + blockIdx.x * blockDim.x + threadIdx.x;
  const int i0 = g.ici1 + (blockIdx.y * blockDim.y + threadIdx.y) * RCM_MO_NC;
  const int k = 1 + blockIdx.z;
  if (j > g.jci2 || i0 > g.ici2) return;
  const int nc = min(RCM_MO_NC, g.ici2 - i0 + 1);
  const double zdtrdy = s.adtrdy;
  const int nf = s.nf;
  double vr[RCM_MO_NC + 1], zamu[RCM_MO_NC + 1], is[RCM_MO_NC + 1];
  bool up[RCM_MO_NC + 1], cl[RCM_MO_NC + 1];     // upwind side; index clamp (ihm1 at icross1 / ih at icross2)
#pragma unroll
  for (int m = 0; m <= RCM_MO_NC; ++m) {
    const int ii = min(i0 + m, g.ice2);          // faces beyond the last owned cell are not used
    vr[m] = L3(f.v, j, ii, k) * L2(f.rmv, j, ii);
    zamu[m] = vr[m] * zdtrdy;
    up[m] = (zamu[m] > 0.0);                     // ih = ii-1, ihm1 = max(ii-2, icross1)  |  ih = min(ii+1, icross2), ihm1 = ih-1
    is[m] = up[m] ? 1.0 : -1.0;
    cl[m] = up[m] ? (ii - 2 < g.icross1) : (ii + 1 > g.icross2);
  }
  double fm[RCM_MO_NC + 2];                      // fmz(i0-1 .. i0+NC)
#pragma unroll
  for (int m = 0; m < RCM_MO_NC + 2; ++m) fm[m] = L3(f.fmz, j, min(i0 - 1 + m, g.ice2), k);
  double zrfmn[RCM_MO_NC], zrfms[RCM_MO_NC], dvf[RCM_MO_NC], mx2[RCM_MO_NC];
#pragma unroll
  for (int c = 0; c < RCM_MO_NC; ++c) {
    zrfmn[c] = zdtrdy * 2.0 * fm[c + 1] / (fm[c + 1] + fm[c + 2]);
    zrfms[c] = zdtrdy * 2.0 * fm[c + 1] / (fm[c + 1] + fm[c]);
    dvf[c] = (vr[c + 1] * zrfmn[c] - vr[c] * zrfms[c]);
    mx2[c] = L2(f.mx2, j, min(i0 + c, g.ice2));
  }
  const bool cw = (j == g.jci1), ce = (j == g.jci2);
  for (int n = 0; n < nf; ++n) {
    const double* wz = f.wz[n];
    double w[RCM_MO_NC + 4];                     // wz(i0-2 .. i0+NC+1), clamped to the rows that exist
#pragma unroll
    for (int m = 0; m < RCM_MO_NC + 4; ++m) w[m] = L3(wz, j, min(max(i0 - 2 + m, g.ice1), g.ice2), k);
    double zpby[RCM_MO_NC + 1];
#pragma unroll
    for (int m = 0; m <= RCM_MO_NC; ++m) {
      const double wc = w[m + 2], wm = w[m + 1];
      // upstream pair from the register window: w[q] holds row i0-2+q, face m sits between q = m+1 and m+2
      const double wh = up[m] ? wm : (cl[m] ? wc : w[m + 3]);
      const double whm1 = up[m] ? (cl[m] ? wm : w[m]) : (cl[m] ? wm : wc);
      const double r = mo_rdeno(wh, whm1, wc, wm);
      const double zphi = mo_phi(r, is[m], zamu[m]);
      zpby[m] = 0.5 * vr[m] * ((1.0 + zphi) * wm + (1.0 - zphi) * wc);
    }
    double* p0 = f.p0[n];
    const double* pp = f.adv[n];
#pragma unroll
    for (int c = 0; c < RCM_MO_NC; ++c) {
      if (c < nc) {
        const int i = i0 + c;
        const double zdv = dvf[c] * L3(pp, j, i, k);
        const double out = w[c + 2] + mx2[c] * (zpby[c] * zrfms[c] - zpby[c + 1] * zrfmn[c] + zdv);
        P3(p0, j, i, k) = out;
        if (cw) P3(p0, g.jce1, i, k) = out;
        if (ce) P3(p0, g.jce2, i, k) = out;
      }
    }
  }
}

__global__ void __launch_bounds__(RCM_MAXT, 3) k_mo_wafx4(const Geom g, const MoFields f, const MoScalars s) {
  const int si = g.si; const long long sk = g.sk;
  const int j0 = g.jci1 + (blockIdx.x * blockDim.x + threadIdx.x) * RCM_MO_NC;
  const int i = g.ici1 + blockIdx.y * blockDim.y + threadIdx.y;
  const int k = 1 + blockIdx.z;
  if (j0 > g.jci2 || i > g.ici2) return;
  const int nc = min(RCM_MO_NC, g.jci2 - j0 + 1);
  const double zdtrdx = s.adtrdx;
  const int nf = s.nf;
  double ur[RCM_MO_NC + 1], zamu[RCM_MO_NC + 1], is[RCM_MO_NC + 1];
  bool up[RCM_MO_NC + 1], cl[RCM_MO_NC + 1];
#pragma unroll
  for (int m = 0; m <= RCM_MO_NC; ++m) {
    const int jj = min(j0 + m, g.jce2);
    ur[m] = L3(f.u, jj, i, k) * L2(f.rmu, jj, i);
    zamu[m] = ur[m] * zdtrdx;
    up[m] = (zamu[m] > 0.0);
    is[m] = up[m] ? 1.0 : -1.0;
    cl[m] = up[m] ? (jj - 2 < g.jcross1) : (jj + 1 > g.jcross2);
  }
  double fm[RCM_MO_NC + 2];
#pragma unroll
  for (int m = 0; m < RCM_MO_NC + 2; ++m) fm[m] = L3(f.fmz, min(j0 - 1 + m, g.jce2), i, k);
  double zrfme[RCM_MO_NC], zrfmw[RCM_MO_NC], dvf[RCM_MO_NC], mx2[RCM_MO_NC];
#pragma unroll
  for (int c = 0; c < RCM_MO_NC; ++c) {
    zrfme[c] = zdtrdx * 2.0 * fm[c + 1] / (fm[c + 1] + fm[c + 2]);
    zrfmw[c] = zdtrdx * 2.0 * fm[c + 1] / (fm[c + 1] + fm[c]);
    dvf[c] = (ur[c + 1] * zrfme[c] - ur[c] * zrfmw[c]);
    mx2[c] = L2(f.mx2, min(j0 + c, g.jce2), i);
  }
  for (int n = 0; n < nf; ++n) {
    const double* p0 = f.p0[n];
    double w[RCM_MO_NC + 4];
#pragma unroll
    for (int m = 0; m < RCM_MO_NC + 4; ++m) w[m] = L3(p0, min(max(j0 - 2 + m, g.jce1), g.jce2), i, k);
    double zpbw[RCM_MO_NC + 1];
#pragma unroll
    for (int m = 0; m <= RCM_MO_NC; ++m) {
      const double pc = w[m + 2], pm = w[m + 1];
      const double ph = up[m] ? pm : (cl[m] ? pc : w[m + 3]);
      const double phm1 = up[m] ? (cl[m] ? pm : w[m]) : (cl[m] ? pm : pc);
      const double r = mo_rdeno(ph, phm1, pc, pm);
      const double zphi = mo_phi(r, is[m], zamu[m]);
      zpbw[m] = 0.5 * ur[m] * ((1.0 + zphi) * pm + (1.0 - zphi) * pc);
    }
    double* pp = f.adv[n];
#pragma unroll
    for (int c = 0; c < RCM_MO_NC; ++c) {
      if (c < nc) {
        const int j = j0 + c;
        const double zdv = dvf[c] * P3(pp, j, i, k);
        P3(pp, j, i, k) = w[c + 2] + mx2[c] * (zpbw[c] * zrfmw[c] - zpbw[c + 1] * zrfme[c] + zdv);
      }
    }
  }
}

__global__ void k_mo_restag(const Geom g, const MoFields f) {
  const int si = g.si; const long long sk = g.sk;
  const int j = g.jde1 + blockIdx.x * blockDim.x + threadIdx.x;
  const int i = g.ide1 + blockIdx.y * blockDim.y + threadIdx.y;
  const int k = 1 + blockIdx.z;
  if (j > g.jde2 || i > g.ide2) return;
  const int kz = g.kz;
  if (i >= g.ici1 && i <= g.ici2) {
    if (j >= g.jdii1 && j <= g.jdii2)
      P3(f.u, j, i, k) = 0.5625 * (L3(f.ux, j, i, k) + L3(f.ux, j - 1, i, k)) - 0.0625 * (L3(f.ux, j + 1, i, k) + L3(f.ux, j - 2, i, k));
    else if (j == g.jdi2) P3(f.u, j, i, k) = 0.5 * (L3(f.ux, g.jci2, i, k) + L3(f.ux, g.jce2, i, k));
    else if (j == g.jdi1) P3(f.u, j, i, k) = 0.5 * (L3(f.ux, g.jci1, i, k) + L3(f.ux, g.jce1, i, k));
  }
  if (j >= g.jci1 && j <= g.jci2) {
    if (i >= g.idii1 && i <= g.idii2)
      P3(f.v, j, i, k) = 0.5625 * (L3(f.vx, j, i, k) + L3(f.vx, j, i - 1, k)) - 0.0625 * (L3(f.vx, j, i + 1, k) + L3(f.vx, j, i - 2, k));
    else if (i == g.idi2) P3(f.v, j, i, k) = 0.5 * (L3(f.vx, j, g.ici2, k) + L3(f.vx, j, g.ice2, k));
    else if (i == g.idi1) P3(f.v, j, i, k) = 0.5 * (L3(f.vx, j, g.ici1, k) + L3(f.vx, j, g.ice1, k));
  }
  if (j <= g.jce2 && i <= g.ice2 && k >= 2) {
    if (k >= 3 && k <= kz - 1)
      P3(f.w, j, i, k) = 0.5625 * (L3(f.wx, j, i, k) + L3(f.wx, j, i, k - 1)) - 0.0625 * (L3(f.wx, j, i, k + 1) + L3(f.wx, j, i, k - 2));
    else if (k == 2) P3(f.w, j, i, 2) = 0.5 * (L3(f.wx, j, i, 2) + L3(f.wx, j, i, 1));
    else if (k == kz) P3(f.w, j, i, kz) = 0.5 * (L3(f.wx, j, i, kz) + L3(f.wx, j, i, kz - 1));
  }
}

static dim3 grid3(int nj, int ni, int nk, dim3 b) { return dim3((nj + b.x - 1) / b.x, (ni + b.y - 1) / b.y, nk); }

void launch_mo_destag(const Geom& g, const MoFields& f, cudaStream_t st) {
  dim3 b(32, 4);
  RCM_LAUNCH(k_mo_destag, grid3(g.jce2 - g.jce1 + 1, g.ice2 - g.ice1 + 1, g.kz, b), b, st, g, f);
}
void launch_mo_waf(const Geom& g, const MoFields& f, const MoScalars& s, int passes, cudaStream_t st) {
  dim3 bc = col_block(), b(32, 4);
  const int nj = g.jci2 - g.jci1 + 1, ni = g.ici2 - g.ici1 + 1;
  // RCM_MO_WAFZ = fields per thread of the vertical pass (default 7 = all fields in one thread: measured fastest on B200, gpurun_out/mo_abz_*.json)
  static int zg = -1;
  if (zg < 0) { const char* e = std::getenv("RCM_MO_WAFZ"); zg = e ? std::atoi(e) : 7; }
  if (!(passes & 1)) {}
  else if (zg >= 7) RCM_LAUNCH(k_mo_wafz<7>, grid3(nj, ni, 1, bc), bc, st, g, f, s);
  else if (zg == 2) RCM_LAUNCH(k_mo_wafz<2>, grid3(nj, ni, (s.nf + 1) / 2, bc), bc, st, g, f, s);
  else if (zg == 3) RCM_LAUNCH(k_mo_wafz<3>, grid3(nj, ni, (s.nf + 2) / 3, bc), bc, st, g, f, s);
  else RCM_LAUNCH(k_mo_wafz<4>, grid3(nj, ni, (s.nf + 3) / 4, bc), bc, st, g, f, s);
  // RCM_MO_WAF: bit 0 = one-cell meridional kernel, bit 1 = one-cell zonal kernel (default 3: measured faster on B200 than the 4-cell kernels, profiles/r1_moloch.md)
  static int mode = -1;
  if (mode < 0) { const char* e = std::getenv("RCM_MO_WAF"); mode = e ? std::atoi(e) : 3; }
  const int md = ((g.gl | g.gr | g.gb | g.gt) || s.rotllr) ? 3 : mode;   // the 4-cell kernels: single rank, non-ROTLLR only
  if (!(passes & 2)) {}
  else if (md & 1) {
    if (s.rotllr) {
      if (s.nf == 7) RCM_LAUNCH((k_mo_wafh<7, 0, true>), grid3(nj, ni, g.kz, b), b, st, g, f, s);
      else RCM_LAUNCH((k_mo_wafh<6, 0, true>), grid3(nj, ni, g.kz, b), b, st, g, f, s);
    } else if (s.nf == 7) RCM_LAUNCH((k_mo_wafh<7, 0, false>), grid3(nj, ni, g.kz, b), b, st, g, f, s);
    else RCM_LAUNCH((k_mo_wafh<6, 0, false>), grid3(nj, ni, g.kz, b), b, st, g, f, s);
  } else RCM_LAUNCH(k_mo_wafy4, grid3(nj, (ni + RCM_MO_NC - 1) / RCM_MO_NC, g.kz, b), b, st, g, f, s);
  if (!(passes & 4)) {}
  else if (md & 2) {
    if (s.rotllr) {
      if (s.nf == 7) RCM_LAUNCH((k_mo_wafh<7, 1, true>), grid3(nj, ni, g.kz, b), b, st, g, f, s);
      else RCM_LAUNCH((k_mo_wafh<6, 1, true>), grid3(nj, ni, g.kz, b), b, st, g, f, s);
    } else if (s.nf == 7) RCM_LAUNCH((k_mo_wafh<7, 1, false>), grid3(nj, ni, g.kz, b), b, st, g, f, s);
    else RCM_LAUNCH((k_mo_wafh<6, 1, false>), grid3(nj, ni, g.kz, b), b, st, g, f, s);
  } else RCM_LAUNCH(k_mo_wafx4, grid3((nj + RCM_MO_NC - 1) / RCM_MO_NC, ni, g.kz, b), b, st, g, f, s);
}
void launch_mo_restag(const Geom& g, const MoFields& f, cudaStream_t st) {
  dim3 b(32, 4);
  RCM_LAUNCH(k_mo_restag, grid3(g.jde2 - g.jde1 + 1, g.ide2 - g.ide1 + 1, g.kz, b), b, st, g, f);
}
