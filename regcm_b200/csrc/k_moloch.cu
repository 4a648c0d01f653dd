// MOLOCH core (idynamic = 3): sound sub-steps.
//   k_mo_pre    tvirt, tetav at the head of `moloch`        Main/mod_moloch.F90:251-301
//   k_mo_div    horizontal part of the divergence (Eq. 16)  :775-793 (ROTLLR), :806-824
//   k_mo_w      filt3d + full divergence + implicit w (tridiagonal, Eq. 19) + new Exner function
//                                                           :549-569,743-768,826-925,1022-1038
//   k_mo_uv     horizontal momentum (Eq. 17, 18)            :975-1012
//   k_mo_close  tvirt, t, p, ps at the end of the step      :321-383
// One thread owns one (j,i) column in k_mo_w (the vertical recurrences); the other kernels are
// elementwise over (j,i,k).  u and v are double-buffered (u,v -> un,vn) so that the Coriolis average
// reads the pre-update winds, which is what the reference's ud/vd copies are for (:735-736).
#ifndef RCM_IDX32_ALL   // (tuning builds only)
#define RCM_IDX64   // 64-bit element indices measured faster for the kernels of this file (dev_common.cuh)
#endif
#include "dev_common.cuh"
#include "launch.h"

using namespace rcm;

__global__ void k_mo_pre(const Geom g, const MoFields f, const MoScalars s) {
  const int si = g.si; const long long sk = g.sk;
  const int j = g.jce1 + blockIdx.x * blockDim.x + threadIdx.x;
  const int i = g.ice1 + blockIdx.y * blockDim.y + threadIdx.y;
  const int k = 1 + blockIdx.z;
  if (j > g.jce2 || i > g.ice2) return;
  double tv;
  if (s.ipptls > 0) tv = P3(f.t, j, i, k) * (1.0 + ep1 * P3(f.qv, j, i, k) - P3(f.qc, j, i, k));
  else tv = P3(f.t, j, i, k) * (1.0 + ep1 * P3(f.qv, j, i, k));
  P3(f.tetav, j, i, k) = tv / P3(f.pai, j, i, k);
}

// Evaluated on the owned cells plus one ghost ring towards neighbouring ranks (g.gl/gr/gb/gt) — two rings on the
// low sides — so that the 5-point filter of k_mo_w finds its neighbours without a separate exchange of the
// divergence (the reference exchanges zdiv2, Main/mod_moloch.F90:553; here u and v carry two ghost rows instead),
// also on the low-side ghost ring that k_mo_w advances itself.
__global__ void k_mo_div(const Geom g, const MoFields f, const MoScalars s) {
  const int si = g.si; const long long sk = g.sk;
  const int j = g.jce1 - 2 * g.gl + blockIdx.x * blockDim.x + threadIdx.x;
  const int i = g.ice1 - 2 * g.gb + blockIdx.y * blockDim.y + threadIdx.y;
  const int k = 1 + blockIdx.z;
  if (j > g.jce2 + g.gr || i > g.ice2 + g.gt) return;
  if (s.rotllr) {   // :775-793
    const double zum = s.zdtrdx * L3(f.u, j, i, k) * L3(f.zru, j, i, k);
    const double zup = s.zdtrdx * L3(f.u, j + 1, i, k) * L3(f.zru, j + 1, i, k);
    const double zvm = s.zdtrdy * L3(f.v, j, i, k) * L3(f.zrv, j, i, k) * L2(f.rmv, j, i);
    const double zvp = s.zdtrdy * L3(f.v, j, i + 1, k) * L3(f.zrv, j, i + 1, k) * L2(f.rmv, j, i + 1);
    P3(f.zdiv, j, i, k) = (zup - zum) + (zvp - zvm);
    return;
  }
  const double zum = L3(f.u, j, i, k) * L2(f.rmu, j, i) * L3(f.zru, j, i, k);
  const double zup = L3(f.u, j + 1, i, k) * L2(f.rmu, j + 1, i) * L3(f.zru, j + 1, i, k);
  const double zvm = L3(f.v, j, i, k) * L2(f.rmv, j, i) * L3(f.zrv, j, i, k);
  const double zvp = L3(f.v, j, i + 1, k) * L2(f.rmv, j, i + 1) * L3(f.zrv, j, i + 1, k);
  P3(f.zdiv, j, i, k) = (zup - zum) * s.zdtrdx + (zvp - zvm) * s.zdtrdy;
}

// filtered horizontal divergence of level k at (j,i): filt3d acts on interior points only
__device__ __forceinline__ double mo_zd(const MoFields& f, const int si, const long long sk, const bool interior,
                                        const double anu2, const int j, const int i, const int k) {
  const double c = L3(f.zdiv, j, i, k);
  if (!interior) return c;
  const double p2d = 0.125 * (L3(f.zdiv, j - 1, i, k) + L3(f.zdiv, j + 1, i, k) + L3(f.zdiv, j, i - 1, k) + L3(f.zdiv, j, i + 1, k)) - 0.5 * c;
  return c + anu2 * p2d;
}

// Inputs of one level of the downward sweep.  They do not depend on the recurrence, so the loads of
// level k-1 are issued before the stores of level k (software pipelining by hand: the compiler cannot
// move them across the stores itself because the arrays may alias).
struct MoWIn {
  double um0, um1, vm0, vm1;   // winds of level k-2 (for s(k-1))
  double gz;                   // gzitak(k-1)
  double fmz_km1, zd_km1, pai_km1, tet_km1, fmzf_k, w_k, ff;
};
__device__ __forceinline__ MoWIn mo_w_load(const Geom& g, const MoFields& f, const double anu2, const bool interior,
                                           const int si, const long long sk, const int j, const int i, const int k) {
  MoWIn r;
  if (k - 1 >= 2) {
    r.um0 = L3(f.u, j, i, k - 2); r.um1 = L3(f.u, j + 1, i, k - 2); r.vm0 = L3(f.v, j, i, k - 2); r.vm1 = L3(f.v, j, i + 1, k - 2);
    r.gz = LD1(f.gzitak + k - 1);
  } else {
    r.um0 = r.um1 = r.vm0 = r.vm1 = 0.0; r.gz = 0.0;
  }
  r.fmz_km1 = L3(f.fmz, j, i, k - 1);
  r.zd_km1 = mo_zd(f, si, sk, interior, anu2, j, i, k - 1);
  r.pai_km1 = P3(f.pai, j, i, k - 1);
  r.tet_km1 = L3(f.tetav, j, i, k - 1);
  r.fmzf_k = L3(f.fmzf, j, i, k);
  r.w_k = P3(f.w, j, i, k);
  r.ff = LD1(f.ffilt + k);
  return r;
}

// `zdf` receives the full divergence (a second array: the raw one is still being read by the
// neighbours' filter); `jsound` is the 1-based sub-step index; `last`: write s (Eq. 10 completed).
//
// SM = true (GPU only): the three per-level scratch values of the column (intermediate w, elimination
// coefficient, full divergence) stay in shared memory between the two sweeps instead of making a round
// trip through HBM: slot (a*(kz+1) + k)*blockThreads + tid of the dynamic shared array, a = 0,1,2.
#ifdef RCM_EMU
#define MOW_LD(a, kk, gl) (gl)
#define MOW_ST(a, kk, gl, v) (gl) = (v)
#else
#define MOW_LD(a, kk, gl) (SM ? smw[((a) * kzp1 + (kk)) * nt + tid] : (gl))
#define MOW_ST(a, kk, gl, v) do { if (SM) smw[((a) * kzp1 + (kk)) * nt + tid] = (v); else (gl) = (v); } while (0)
#endif
#ifndef RCM_MB_MOW
#define RCM_MB_MOW 4
#endif
template <bool SM>
__global__ void __launch_bounds__(RCM_MAXT, RCM_MB_MOW) k_mo_w(const Geom g, const MoFields f, const MoScalars s, double* zdf,
                                                      const int jsound, const int last) {
#ifndef RCM_EMU
  extern __shared__ double smw[];
  const int nt = blockDim.x * blockDim.y, tid = threadIdx.x + blockDim.x * threadIdx.y;
#endif
  const int si = g.si; const long long sk = g.sk;
  // owned columns plus the first ghost row / column on the LOW sides that have a neighbour rank: k_mo_uv reads pai and
  // deltaw at (j-1, i) and (j, i-1), and advancing those columns here (bit-identical to their owner's values)
  // replaces the exchange of pai and deltaw in every sound sub-step (the reference: Main/mod_moloch.F90:929-930)
  const int j = g.jce1 - g.gl + blockIdx.x * blockDim.x + threadIdx.x;
  const int i = g.ice1 - g.gb + blockIdx.y * blockDim.y + threadIdx.y;
  if (j > g.jce2 || i > g.ice2) return;
  const int kz = g.kz, kzp1 = g.kzp1;
  const bool interior = (j >= 2 && j <= g.jx - 2 && i >= 2 && i <= g.iy - 2);   // interior of the global domain
  const double hx0 = L2(f.hx, j, i), hx1 = L2(f.hx, j + 1, i), hy0 = L2(f.hy, j, i), hy1 = L2(f.hy, j, i + 1);
  const double mx2 = s.rotllr ? L2(f.mx, j, i) : L2(f.mx2, j, i);   // mx (ROTLLR, :798) or mx**2 (:829)
  const double zdtrdz = s.zdtrdz, zcs2 = s.zcs2, egdts = s.egdts, rjs = (double)jsound;
  // winds of level k (carried) and k-1
  double u0 = L3(f.u, j, i, kz), u1 = L3(f.u, j + 1, i, kz), v0 = L3(f.v, j, i, kz), v1 = L3(f.v, j, i + 1, kz);
  double um0 = L3(f.u, j, i, kz - 1), um1 = L3(f.u, j + 1, i, kz - 1), vm0 = L3(f.v, j, i, kz - 1), vm1 = L3(f.v, j, i + 1, kz - 1);
  MoWIn cur = mo_w_load(g, f, s.anu2, interior, si, sk, j, i, kz);
  double fmz_k = L3(f.fmz, j, i, kz);
  const double zd_kz = mo_zd(f, si, sk, interior, s.anu2, j, i, kz);
  double pai_k = P3(f.pai, j, i, kz), tet_k = L3(f.tetav, j, i, kz);
  double wkp1;  // w(k+1) of the downward sweep
  {
    const double zuh = u0 * hx0 + u1 * hx1, zvh = v0 * hy0 + v1 * hy1;
    wkp1 = 0.5 * (zuh + zvh);
    P3(f.w, j, i, kzp1) = wkp1;
  }
  const double w_bot = wkp1;
  double s_kp1 = -wkp1;
  if (last) P3(f.s, j, i, kzp1) = interior ? 0.0 : s_kp1;
  double s_k;
  {
    const double zuh = (u0 + um0) * hx0 + (u1 + um1) * hx1, zvh = (v0 + vm0) * hy0 + (v1 + vm1) * hy1;
    s_k = -0.25 * (zuh + zvh) * LD1(f.gzitak + kz);
  }
  double zdiv_k = fmz_k * (mx2 * zd_kz + zdtrdz * (s_k - s_kp1));
  MOW_ST(2, kz, P3(zdf, j, i, kz), zdiv_k);
  double wwkw_kp1 = 0.0;
  // request the operands of level k-1-pfd into L2 while level k is worked on (see PF3 in dev_common.cuh)
  const int pfd = g.pfd;
  auto prefetch = [&](int kk) {
    PF3(f.u, j, i, kk - 1); PF3(f.v, j, i, kk - 1); PF3(f.v, j, i + 1, kk - 1); PF3(f.fmz, j, i, kk); PF3(f.zdiv, j, i, kk);
    PF3(f.pai, j, i, kk); PF3(f.tetav, j, i, kk); PF3(f.fmzf, j, i, kk + 1); PF3(f.w, j, i, kk + 1);
  };
  if (pfd > 0)
    for (int kk = kz - 1; kk >= max(kz - 1 - pfd, 2); --kk) prefetch(kk);
  for (int k = kz; k >= 2; --k) {
    if (pfd > 0 && k - 2 - pfd >= 2) prefetch(k - 2 - pfd);
    MoWIn nxt = cur;
    if (k > 2) nxt = mo_w_load(g, f, s.anu2, interior, si, sk, j, i, k - 1);
    // level k-1 quantities: s(k-1) (0 for k-1 = 1), zdiv(k-1)
    u0 = um0; u1 = um1; v0 = vm0; v1 = vm1;
    double s_km1 = 0.0;
    if (k - 1 >= 2) {
      um0 = cur.um0; um1 = cur.um1; vm0 = cur.vm0; vm1 = cur.vm1;
      const double zuh = (u0 + um0) * hx0 + (u1 + um1) * hx1, zvh = (v0 + vm0) * hy0 + (v1 + vm1) * hy1;
      s_km1 = -0.25 * (zuh + zvh) * cur.gz;
    }
    const double fmz_km1 = cur.fmz_km1;
    const double zdiv_km1 = fmz_km1 * (mx2 * cur.zd_km1 + zdtrdz * (s_km1 - s_k));
    MOW_ST(2, k - 1, P3(zdf, j, i, k - 1), zdiv_km1);
    const double pai_km1 = cur.pai_km1, tet_km1 = cur.tet_km1;
    const double fmzf_k = cur.fmzf_k;
    const double w_k = cur.w_k;
    P3(f.deltaw, j, i, k) = -w_k;
    double zrom1w = 0.5 * cpd * fmzf_k * (tet_km1 + tet_k);
    zrom1w = zrom1w - cpd * w_k * (fmzf_k * fmzf_k) * rjs * zdtrdz * (tet_km1 - tet_k);
    double zwexpl = w_k - zrom1w * zdtrdz * (pai_km1 - pai_k) - egdts;
    zwexpl = zwexpl + rdrcv * zrom1w * zdtrdz * (pai_km1 * zdiv_km1 - pai_k * zdiv_k);
    const double ff = cur.ff;
    const double zp = zcs2 * fmz_km1 * zrom1w * pai_km1 + ff;
    const double zm = zcs2 * fmz_k * zrom1w * pai_k + ff;
    const double zrapp = 1.0 / (1.0 + zm + zp - zm * wwkw_kp1);
    const double wn = zrapp * (zwexpl + zm * wkp1);
    wwkw_kp1 = zrapp * zp;
    MOW_ST(0, k, P3(f.w, j, i, k), wn);
    MOW_ST(1, k, P3(f.wwkw, j, i, k), wwkw_kp1);
    if (last) P3(f.s, j, i, k) = s_k;
    wkp1 = wn;
    s_k = s_km1; zdiv_k = zdiv_km1; fmz_k = fmz_km1; pai_k = pai_km1; tet_k = tet_km1;
    cur = nxt;
  }
  // 2nd loop of the tridiagonal inversion, new Exner function, completed s; the operands of level
  // k+1 are loaded before the stores of level k
  double w_km1 = P3(f.w, j, i, 1);
  double a_w = MOW_LD(0, 2, P3(f.w, j, i, 2)), a_ww = MOW_LD(1, 2, P3(f.wwkw, j, i, 2)), a_dw = P3(f.deltaw, j, i, 2);
  double a_zd = 0.0, a_fmz = 0.0, a_pai = 0.0, a_s = 0.0, a_fzf = 0.0;
  if (interior) {
    a_zd = MOW_LD(2, 1, P3(zdf, j, i, 1)); a_fmz = L3(f.fmz, j, i, 1); a_pai = P3(f.pai, j, i, 1);
    if (last) { a_s = P3(f.s, j, i, 2); a_fzf = L3(f.fmzf, j, i, 2); }
  }
  for (int k = 2; k <= kz; ++k) {
    double b_w = 0.0, b_ww = 0.0, b_dw = 0.0, b_zd = 0.0, b_fmz = 0.0, b_pai = 0.0, b_s = 0.0, b_fzf = 0.0;
    if (k < kz) {
      b_w = MOW_LD(0, k + 1, P3(f.w, j, i, k + 1)); b_ww = MOW_LD(1, k + 1, P3(f.wwkw, j, i, k + 1)); b_dw = P3(f.deltaw, j, i, k + 1);
    }
    if (interior) {   // level k (the Exner function is updated one level behind w); for k = kz these feed the tail
      b_zd = MOW_LD(2, k, P3(zdf, j, i, k)); b_fmz = L3(f.fmz, j, i, k); b_pai = P3(f.pai, j, i, k);
      if (last && k < kz) { b_s = P3(f.s, j, i, k + 1); b_fzf = L3(f.fmzf, j, i, k + 1); }
    }
    const double wk = a_w + a_ww * w_km1;
    P3(f.w, j, i, k) = wk;
    P3(f.deltaw, j, i, k) = a_dw + wk;
    if (interior) {
      const double zd = a_zd + zdtrdz * a_fmz * (w_km1 - wk);
      P3(f.pai, j, i, k - 1) = a_pai * (1.0 - rdrcv * zd);
      if (last) P3(f.s, j, i, k) = (wk + a_s) * a_fzf;
    }
    w_km1 = wk;
    a_w = b_w; a_ww = b_ww; a_dw = b_dw; a_zd = b_zd; a_fmz = b_fmz; a_pai = b_pai; a_s = b_s; a_fzf = b_fzf;
  }
  if (interior) {
    const double zd = a_zd + zdtrdz * a_fmz * (w_km1 - w_bot);
    P3(f.pai, j, i, kz) = a_pai * (1.0 - rdrcv * zd);
    if (last) P3(f.s, j, i, 1) = 0.0;
  }
}

__global__ void k_mo_uv(const Geom g, const MoFields f, const MoScalars s) {
  const int si = g.si; const long long sk = g.sk;
  const int j = g.jde1 + blockIdx.x * blockDim.x + threadIdx.x;
  const int i = g.ide1 + blockIdx.y * blockDim.y + threadIdx.y;
  const int k = 1 + blockIdx.z;
  if (j > g.jde2 || i > g.ide2) return;
  const double gk = LD1(f.gzitakh + k);
  if (i <= g.ice2) {  // u point
    double un = L3(f.u, j, i, k);
    if (j >= g.jdii1 && j <= g.jdii2 && i >= g.ici1 && i <= g.ici2) {
      const double zcx = s.zdtrdx * L2(f.msfu, j, i);
      const double zfz = 0.25 * (L3(f.deltaw, j - 1, i, k) + L3(f.deltaw, j - 1, i, k + 1) + L3(f.deltaw, j, i, k) + L3(f.deltaw, j, i, k + 1)) + s.egdts;
      const double zrom1u = 0.5 * cpd * (L3(f.tetav, j - 1, i, k) + L3(f.tetav, j, i, k));
      const double zcor1u = L2(f.coriou, j, i) * 0.25 * (L3(f.v, j, i, k) + L3(f.v, j - 1, i, k) + L3(f.v, j, i + 1, k) + L3(f.v, j - 1, i + 1, k));
      un = un + zcor1u * s.dts - zfz * L2(f.hx, j, i) * gk - zcx * zrom1u * (L3(f.pai, j, i, k) - L3(f.pai, j - 1, i, k));
    }
    P3(f.un, j, i, k) = un;
  }
  if (j <= g.jce2) {  // v point
    double vn = L3(f.v, j, i, k);
    if (i >= g.idii1 && i <= g.idii2 && j >= g.jci1 && j <= g.jci2) {
      const double zcy = s.rotllr ? s.zdtrdy : s.zdtrdy * L2(f.msfv, j, i);   // :956 / :997
      const double zfz = 0.25 * (L3(f.deltaw, j, i - 1, k) + L3(f.deltaw, j, i - 1, k + 1) + L3(f.deltaw, j, i, k) + L3(f.deltaw, j, i, k + 1)) + s.egdts;
      const double zrom1v = 0.5 * cpd * (L3(f.tetav, j, i - 1, k) + L3(f.tetav, j, i, k));
      const double zcor1v = L2(f.coriov, j, i) * 0.25 * (L3(f.u, j, i, k) + L3(f.u, j + 1, i, k) + L3(f.u, j, i - 1, k) + L3(f.u, j + 1, i - 1, k));
      vn = vn - zcor1v * s.dts - zfz * L2(f.hy, j, i) * gk - zcy * zrom1v * (L3(f.pai, j, i, k) - L3(f.pai, j, i - 1, k));
    }
    P3(f.vn, j, i, k) = vn;
  }
}

__global__ void k_mo_close(const Geom g, const MoFields f, const MoScalars s) {
  const int si = g.si; const long long sk = g.sk;
  const int j = g.jce1 + blockIdx.x * blockDim.x + threadIdx.x;
  const int i = g.ice1 + blockIdx.y * blockDim.y + threadIdx.y;
  const bool inrange = (j <= g.jce2 && i <= g.ice2);
  const bool interior = (j >= g.jci1 && j <= g.jci2 && i >= g.ici1 && i <= g.ici2);
  unsigned long long bmax = 0ULL, bmin = 0ULL;
  if (inrange) {
    double tv = 0.0, pk = 0.0;
    for (int k = 1; k <= g.kz; ++k) {
      const double qv = P3(f.qv, j, i, k);
      const double den = (s.ipptls > 0) ? (1.0 + ep1 * qv - P3(f.qc, j, i, k)) : (1.0 + ep1 * qv);
      const double pai = P3(f.pai, j, i, k);
      if (interior) {
        tv = P3(f.tetav, j, i, k) * pai;
        P3(f.t, j, i, k) = tv / den;
      } else {
        tv = P3(f.t, j, i, k) * den;
      }
      P3(f.tvirt, j, i, k) = tv;
      pk = pow_pos(pai, cpovr) * p00;
      P3(f.p, j, i, k) = pk;
    }
    const double zdgz = L2(f.ht, j, i) * s.gzm1 + s.zz1;
    const double ps = pk * exp(zdgz / (rgas * tv));
    P2(f.ps, j, i) = ps;
    if (!(ps == ps)) *f.nanflag = 1;
    else if (interior && ps > 0.0) {   // positive doubles order like their bit patterns
      bmax = (unsigned long long)__double_as_longlong(ps);
      bmin = ~bmax;
    }
  }
  // max / min of ps over the interior (Main/mod_moloch.F90:437-441): one atomic pair per warp
#ifndef RCM_EMU
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const unsigned long long a = __shfl_xor_sync(0xffffffffu, bmax, o), b = __shfl_xor_sync(0xffffffffu, bmin, o);
    bmax = a > bmax ? a : bmax; bmin = b > bmin ? b : bmin;
  }
  if (((threadIdx.x + blockDim.x * threadIdx.y) & 31) != 0) return;
#endif
  if (bmax == 0ULL) return;
  atomicMax(f.psbits, bmax);
  atomicMax(f.psbits + 1, bmin);
}

static dim3 grid3(const Geom& g, int nj, int ni, int nk, dim3 b) { return dim3((nj + b.x - 1) / b.x, (ni + b.y - 1) / b.y, nk); }

void launch_mo_pre(const Geom& g, const MoFields& f, const MoScalars& s, cudaStream_t st) {
  dim3 b(32, 4);
  RCM_LAUNCH(k_mo_pre, grid3(g, g.jce2 - g.jce1 + 1, g.ice2 - g.ice1 + 1, g.kz, b), b, st, g, f, s);
}
void launch_mo_div(const Geom& g, const MoFields& f, const MoScalars& s, cudaStream_t st) {
  dim3 b(32, 4);
  RCM_LAUNCH(k_mo_div, grid3(g, g.jce2 - g.jce1 + 1 + 2 * g.gl + g.gr, g.ice2 - g.ice1 + 1 + 2 * g.gb + g.gt, g.kz, b), b, st, g, f, s);
}
void launch_mo_w(const Geom& g, const MoFields& f, const MoScalars& s, double* zdf, int jsound, int last, cudaStream_t st) {
  dim3 b = col_block();
  const dim3 gr = grid3(g, g.jce2 - g.jce1 + 1 + g.gl, g.ice2 - g.ice1 + 1 + g.gb, 1, b);
#ifndef RCM_EMU
  // RCM_MO_WSM=1 selects the shared-memory scratch (24 B x (kz+1) per thread); default: global-memory scratch
  static int mode = -1;
  static size_t maxsm = 0;
  if (mode < 0) {
    const char* e = std::getenv("RCM_MO_WSM");
    mode = e ? std::atoi(e) : 0;   // measured on B200 (C5): 5.43 ms/step with the shared-memory scratch (8 warps/SM) against 4.61 ms without
    int dev = 0, v = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&v, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    maxsm = (size_t)v;
    if (mode) cudaFuncSetAttribute(k_mo_w<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)maxsm);
  }
  const size_t need = (size_t)3 * g.kzp1 * b.x * b.y * sizeof(double);
  if (mode && need <= maxsm) {
    k_mo_w<true><<<gr, b, need, st>>>(g, f, s, zdf, jsound, last);
    return;
  }
#endif
  RCM_LAUNCH(k_mo_w<false>, gr, b, st, g, f, s, zdf, jsound, last);
}
void launch_mo_uv(const Geom& g, const MoFields& f, const MoScalars& s, cudaStream_t st) {
  dim3 b(32, 4);
  RCM_LAUNCH(k_mo_uv, grid3(g, g.jde2 - g.jde1 + 1, g.ide2 - g.ide1 + 1, g.kz, b), b, st, g, f, s);
}
void launch_mo_close(const Geom& g, const MoFields& f, const MoScalars& s, cudaStream_t st) {
  dim3 b = col_block();
  RCM_LAUNCH(k_mo_close, grid3(g, g.jce2 - g.jce1 + 1, g.ice2 - g.ice1 + 1, 1, b), b, st, g, f, s);
}
