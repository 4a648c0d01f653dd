// Acoustic sub-stepping: `sound`, Main/mod_sound.F90:163-718.
//
//   k_snd_init   :221-246  decouple level b into the short-step state
//   k_snd_uv     :270-299  explicit u,v (pressure-gradient) step
//   k_snd_wfwd   :317-476  w/p' coefficients, tridiagonal forward elimination
//                          (bottom-up), estore/astore for the radiative BC
//   k_snd_wback  :551-561  13x13 upper-radiative-BC convolution (per column), then
//                :565-689  back substitution, CFL, new p', T correction,
//                          divergence damping + d p'/d p0 for the next sub-step
//   k_snd_final  :696-712  recouple, Robert-Asselin filters, tiny-w flush
//
// The ~25 work arrays of the reference (cc,cdd,cj,ca,g1,g2,aa,b,c,rhs,ptend,
// pxup,pyvp,tk,wo,...) live in registers; only the recurrence coefficients e,f
// cross HBM, because the radiative upper BC couples all columns between the
// two sweeps.  The short-step pressure is carried in two arrays: cpp (p') and
// cppd (p' + xkd*pi, the divergence-damped value the u,v step and d p'/d p0
// read, :252-256/:302-306), which reproduces the reference's add/subtract of
// xkd*pi bit for bit.
#include "dev_common.cuh"
#include "launch.h"

using namespace rcm;

// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k_snd_init(const Geom g, const Fields f, const StepScalars s) {
  const int si = g.si;
  const long long sk = g.sk;
  const int j = (g.jde1 - g.gl) + blockIdx.x * 32 + threadIdx.x;
  const int i = (g.ide1 - g.gb) + blockIdx.y * 8 + threadIdx.y;
  if (j > g.jde2 + g.gr || i > g.ide2 + g.gt) return;
  const int kz = g.kz;
  const double psdotb = L2(f.psdotb, j, i);
  for (int k = 1; k <= kz; ++k) {
    const double rpsd = recip_for_div(psdotb);
    P3(f.cu, j, i, k) = div_r(L3(f.bu, j, i, k), psdotb, rpsd);
    P3(f.cv, j, i, k) = div_r(L3(f.bv, j, i, k), psdotb, rpsd);
  }
  if (j <= g.jce2 + g.gr && i <= g.ice2 + g.gt) {
    const double rpsb = 1.0 / L2(f.psb, j, i);
    double pm = 0.0, pc = L3(f.bpp, j, i, 1) * rpsb, pn;
    for (int k = 1; k <= kz; ++k) {
      pn = (k < kz) ? L3(f.bpp, j, i, k + 1) * rpsb : 0.0;
      P3(f.cpp, j, i, k) = pc;
      P3(f.cppd, j, i, k) = pc;
      const int km1 = max(1, k - 1), kp1 = min(kz, k + 1);
      const double a = (k == 1) ? pc : pm, b = (k == kz) ? pc : pn;
      P3(f.ct, j, i, k) = (a - b) / (L3(f.pr0, j, i, km1) - L3(f.pr0, j, i, kp1));
      P3(f.cw, j, i, k) = L3(f.bw, j, i, k) * rpsb;
      pm = pc; pc = pn;
    }
    P3(f.cw, j, i, kz + 1) = L3(f.bw, j, i, kz + 1) * rpsb;
  }
}

// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k_snd_uv(const Geom g, const Fields f, const StepScalars s) {
  const int si = g.si;
  const long long sk = g.sk;
  const int j = g.jdi1 + blockIdx.x * 32 + threadIdx.x;
  const int i = g.idi1 + blockIdx.y * 4 + threadIdx.y;
  const int k = 1 + blockIdx.z * 2 + threadIdx.z;
  // the first ghost row/column on the high side is advanced redundantly (same
  // inputs, same arithmetic as the owning rank): k_snd_wfwd/k_snd_wback read
  // u,v at (j..j+1, i..i+1), and this saves one halo exchange per sub-step
  if (j > g.jdi2 + g.gr || i > g.idi2 + g.gt || k > g.kz) return;
  const double rho = 0.25 * (L3(f.rho1, j, i, k) + L3(f.rho1, j - 1, i, k) + L3(f.rho1, j, i - 1, k) + L3(f.rho1, j - 1, i - 1, k));
  const double dppdp0 = 0.25 * (L3(f.ct, j, i, k) + L3(f.ct, j - 1, i, k) + L3(f.ct, j, i - 1, k) + L3(f.ct, j - 1, i - 1, k));
  const double chh = 0.5 * s.dts / (rho * s.dx) / L2(f.msfd, j, i);
  const double p11 = L3(f.cppd, j, i, k), p01 = L3(f.cppd, j - 1, i, k), p10 = L3(f.cppd, j, i - 1, k), p00_ = L3(f.cppd, j - 1, i - 1, k);
  // atm0%dprddx/dprddy (Main/mod_params.F90:2680-2693) recomputed from pr0
  const double r11 = L3(f.pr0, j, i, k), r01 = L3(f.pr0, j - 1, i, k), r10 = L3(f.pr0, j, i - 1, k), r00 = L3(f.pr0, j - 1, i - 1, k);
  const double dprddx = r11 - r01 + r10 - r00;
  const double dprddy = r11 - r10 + r01 - r00;
  double u = P3(f.cu, j, i, k), v = P3(f.cv, j, i, k);
  u = u - chh * (p11 - p01 + p10 - p00_ - dprddx * dppdp0);
  v = v - chh * (p11 - p10 + p01 - p00_ - dprddy * dppdp0);
  P3(f.cu, j, i, k) = u + L3(f.uten, j, i, k);
  P3(f.cv, j, i, k) = v + L3(f.vten, j, i, k);
}

// ---------------------------------------------------------------------------
// One level of per-column coefficients (mod_sound.F90:335-372 for k = 1,
// :373-416 for k >= 2 — the k-local part).
// FIRST = first sub-step of a model step: the coefficients that do not change
// during the step (cc, cdd and ca: 4 of the 10 divisions per level) are computed
// and stored; the later sub-steps load them instead of pr1 and rho1.
// 64-bit element indices measured 4 % faster for this kernel alone (5.58 against 5.80 ms on C4; dev_common.cuh)
#ifndef RCM_IDX32_ALL   // (tuning builds only)
#undef P3
#undef P2
#define P3(p, j, i, k) P3_64(p, j, i, k)
#define P2(p, j, i) P2_64(p, j, i)
#endif
template <bool FIRST>
__global__ void __launch_bounds__(RCM_MAXT, RCM_MB_WFWD)
k_snd_wfwd(const Geom g, const Fields f, const StepScalars s) {
  const int si = g.si;
  const long long sk = g.sk;
  const int j = g.jci1 + blockIdx.x * blockDim.x + threadIdx.x;
  const int i = g.ici1 + blockIdx.y * blockDim.y + threadIdx.y;
  if (j > g.jci2 || i > g.ici2) return;
  const int kz = g.kz, kzp1 = g.kzp1;
  const double dts = s.dts, dx = s.dx, xgamma = s.xgamma, bp = s.bp, bm = s.bm, bpxbp = s.bpxbp, bpxbm = s.bpxbm;

  // request level kk of every streamed field into L2 (see PF3)
  auto prefetch = [&](int kk) {
    PF3(f.cu, j, i, kk); PF3(f.cv, j, i, kk); PF3(f.rho0, j, i, kk);
    PF3(f.pr0, j, i, kk); PF3(f.t0, j, i, kk); PF3(f.nbt, j, i, kk); PF3(f.ppten, j, i, kk);
    PF3(f.cpp, j, i, kk); PF3(f.cw, j, i, kk); PF3(f.wten, j, i, kk);
    if (FIRST) { PF3(f.pr1, j, i, kk); PF3(f.rho1, j, i, kk); }
    else { PF3(f.scc, j, i, kk); PF3(f.scdd, j, i, kk); PF3(f.sca, j, i, kk); }
  };
  const int pfd = g.pfd_wfwd;
  if (pfd > 0)
    for (int kk = kz; kk >= max(kz - pfd, 1); --kk) prefetch(kk);

#if defined(RCM_WF_PIPE) && RCM_WF_PIPE
#define WF_PIPELINED
#endif
#define WF_OK true
#define WF_ITER_BEGIN(k) if (pfd > 0 && (k) - 1 - pfd >= 1) prefetch((k) - 1 - pfd)
#define WF_LOADUV_ITER(U, V) LOADUV(U, V, k - 2)
#define WF_LOADRAW_ITER(r) r = loadraw(km1)
#define WF_P0N L3(f.pr0, j, i, k - 2)
#define WF_WTEN L3(f.wten, j, i, k)
#define WF_CA L3(f.sca, j, i, k)
#include "k_snd_wfwd_body.inc"
#undef WF_OK
#undef WF_ITER_BEGIN
#undef WF_LOADUV_ITER
#undef WF_LOADRAW_ITER
#undef WF_P0N
#undef WF_WTEN
#undef WF_CA
#undef WF_PIPELINED
}
#undef P3
#undef P2
#define P3(p, j, i, k) P3_32(p, j, i, k)
#define P2(p, j, i) P2_32(p, j, i)

// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(RCM_MAXT, RCM_MB_WBACK)
k_snd_wback(const Geom g, const Fields f, const StepScalars s) {
  const int si = g.si;
  const long long sk = g.sk;
  const int j = g.jci1 + blockIdx.x * blockDim.x + threadIdx.x;
  const int i = g.ici1 + blockIdx.y * blockDim.y + threadIdx.y;
  const bool valid = (j <= g.jci2 && i <= g.ici2);
  double cfl = 0.0;
  if (valid) {
  const int kz = g.kz;
  const double dts = s.dts, xgamma = s.xgamma, bp = s.bp, xkd = s.xkd_snd;
  const double ps0 = L2(f.ps0, j, i), psb = L2(f.psb, j, i);
  const double rpsb = 1.0 / psb;
  const double dpsdxm = L2(f.dpsdxm, j, i), dpsdym = L2(f.dpsdym, j, i);
  const bool last = (s.it == s.istep);

  // Inputs of one level, loaded as ONE batch at the top of the iteration (the
  // loop is otherwise a chain of ~8 serialized memory round trips per level, ncu:
  // 88% long-scoreboard); level k+pfd is requested into L2 at the same time.
  struct In {
    double se, sf, u0, u1, u2, u3, v0, v1, v2, v3, rhof0, cpp, cppd, pr1, rho0, rho1, ppten, nbqv, nbt, nat, pr0;
    double t1, t2, sig, ds;
  };
  auto load = [&](int k) {
    In r;
    r.se = L3(f.se, j, i, k); r.sf = L3(f.sf, j, i, k);
    r.u0 = L3(f.cu, j, i, k); r.u1 = L3(f.cu, j, i + 1, k); r.u2 = L3(f.cu, j + 1, i, k); r.u3 = L3(f.cu, j + 1, i + 1, k);
    r.v0 = L3(f.cv, j, i, k); r.v1 = L3(f.cv, j, i + 1, k); r.v2 = L3(f.cv, j + 1, i, k); r.v3 = L3(f.cv, j + 1, i + 1, k);
    r.rhof0 = L3(f.rhof0, j, i, k);
    r.cpp = P3(f.cpp, j, i, k); r.cppd = P3(f.cppd, j, i, k);
    r.pr1 = L3(f.scdd, j, i, k); r.rho0 = L3(f.rho0, j, i, k); r.rho1 = L3(f.rho1, j, i, k);  // pr1 <- cdd (k_snd_wfwd)
    r.ppten = L3(f.ppten, j, i, k); r.nbqv = L3(f.nbqv, j, i, k);
    r.nbt = P3(f.nbt, j, i, k); r.nat = P3(f.nat, j, i, k);
    r.pr0 = L3(f.pr0, j, i, k);
    r.t1 = LD1(&f.twt1[k]); r.t2 = LD1(&f.twt2[k]); r.sig = LD1(&f.sigma[k]); r.ds = LD1(&f.dsigma[k]);
    return r;
  };

  // request level kk of every streamed field into L2
  auto prefetch = [&](int kk) {
    PF3(f.se, j, i, kk); PF3(f.sf, j, i, kk); PF3(f.cu, j, i, kk); PF3(f.cv, j, i, kk);
    PF3(f.rhof0, j, i, kk); PF3(f.cpp, j, i, kk); PF3(f.cppd, j, i, kk); PF3(f.scdd, j, i, kk);
    PF3(f.rho0, j, i, kk); PF3(f.rho1, j, i, kk); PF3(f.ppten, j, i, kk); PF3(f.nbqv, j, i, kk);
    PF3(f.nbt, j, i, kk); PF3(f.nat, j, i, kk); PF3(f.pr0, j, i, kk);
  };
  const int pfd = g.pfd_wback;
  if (pfd > 0)
    for (int kk = 1; kk <= min(pfd, kz); ++kk) prefetch(kk);
  // ---- upper radiative boundary condition (:497-523): w at the model top is
  //      the 13x13 tmask convolution of estore (clamped at the domain edge) ----
  double w_k = 0.0;
  if (s.ifupr == 1) {
    for (int nsi = -6; nsi <= 6; ++nsi) {
      const int inn = min(max(g.icross1 + 1, i + nsi), g.icross2 - 1);
      const double* erow = f.estore + (long long)inn * g.esi;
      const double* trow = f.tmask + 13 * (nsi + 6) + 6;
#pragma unroll
      for (int nsj = -6; nsj <= 6; ++nsj) {
        const int jnn = min(max(g.jcross1 + 1, j + nsj), g.jcross2 - 1);
        w_k = w_k + erow[jnn] * LD1(trow + nsj);
      }
    }
  }
  P3(f.cw, j, i, 1) = w_k;
  double ucrs_m = 0.0, vcrs_m = 0.0;
  double ppd_m2 = 0.0, ppd_m1 = 0.0;  // damped p' at k-2, k-1 (for ct of level k-1)
  double pr0_m1 = 0.0, pr0_m2 = 0.0, ds_m = 1.0;
  for (int k = 1; k <= kz; ++k) {
    if (pfd > 0 && k + pfd <= kz) prefetch(k + pfd);
    const In c = load(k);
    const double w_n = c.se * w_k + c.sf;
    P3(f.cw, j, i, k + 1) = w_n;
    // ---- CFL diagnostic (:613-633) ----
    const double ucrs = c.u0 + c.u1 + c.u2 + c.u3;
    const double vcrs = c.v0 + c.v1 + c.v2 + c.v3;
    if (k >= 2) {
      const double sigdot = -c.rhof0 * egrav * w_k / ps0 -
                            c.sig * (dpsdxm * (c.t1 * ucrs + c.t2 * ucrs_m) + dpsdym * (c.t1 * vcrs + c.t2 * vcrs_m));
      const double check = fabs(sigdot) * s.dt / (c.ds + ds_m);
      cfl = fmax(check, cfl);
    }
    ucrs_m = ucrs; vcrs_m = vcrs;
    // ---- new pressure and temperature correction (:669-689) ----
    const double ppold = c.cpp;
    const double cddtmp = c.pr1;  // xgamma*pr1*rho0*egrav*dts/(ps0*dsigma), stored by k_snd_wfwd
    const double cjtmp = c.rho0 * egrav * dts * 0.5;
    const double ppn = c.cppd + (cjtmp * (w_n + w_k) + cddtmp * (w_n - w_k)) * bp;
    const double pi = ppn - ppold - c.ppten;
    const double cpm = cpmf(c.nbqv * rpsb);
    const double dpterm = psb * (ppn - ppold) / (cpm * c.rho1);
    P3(f.nbt, j, i, k) = c.nbt + s.gnu1 * dpterm;
    P3(f.nat, j, i, k) = c.nat + dpterm;
    // ---- prepare the next sub-step: divergence damping (:252-256,302-306) and
    //      d p'/d p0 (:257-264) ----
    if (!last) {
      const double ppd = ppn + xkd * pi;
      P3(f.cppd, j, i, k) = ppd;
      P3(f.cpp, j, i, k) = ppd - xkd * pi;
      if (k >= 2) {
        // ct(m), m = k-1: (ppd(max(m-1,1)) - ppd(m+1)) / (pr0(max(m-1,1)) - pr0(m+1))
        const double a = (k == 2) ? ppd_m1 : ppd_m2;
        const double pa = (k == 2) ? pr0_m1 : pr0_m2;
        P3(f.ct, j, i, k - 1) = (a - ppd) / (pa - c.pr0);
      }
      if (k == kz) {
        const double a = (kz == 1) ? ppd : ppd_m1;
        const double pa = (kz == 1) ? c.pr0 : pr0_m1;
        P3(f.ct, j, i, kz) = (a - ppd) / (pa - c.pr0);
      }
      ppd_m2 = ppd_m1; ppd_m1 = ppd;
    } else {
      P3(f.cpp, j, i, k) = ppn;
    }
    pr0_m2 = pr0_m1; pr0_m1 = c.pr0; ds_m = c.ds;
    w_k = w_n;
  }
  }  // valid
  rcm::commit_cfl(f, cfl, valid);
}

// ---------------------------------------------------------------------------
// Epilogue: recouple with p*, Robert-Asselin filters of u,v,pp (gnu1) and
// w (gnu2), flush |w| < 1e-20 to zero on levels a,b,c.
__global__ void __launch_bounds__(256)
k_snd_final(const Geom g, const Fields f, const StepScalars s) {
  const int si = g.si;
  const long long sk = g.sk;
  const int j = g.jde1 + blockIdx.x * 32 + threadIdx.x;
  const int i = g.ide1 + blockIdx.y * 4 + threadIdx.y;
  const int k = 1 + blockIdx.z * 2 + threadIdx.z;
  if (j > g.jde2 || i > g.ide2 || k > g.kzp1) return;
  const int kz = g.kz;
  if (k <= kz && j >= g.jdi1 && j <= g.jdi2 && i >= g.idi1 && i <= g.idi2) {
    const double psdotb = L2(f.psdotb, j, i);
    const double cu = psdotb * L3(f.cu, j, i, k), cv = psdotb * L3(f.cv, j, i, k);
    const double au = P3(f.au, j, i, k), av = P3(f.av, j, i, k);
    double d = s.gnu1 * (cu + P3(f.bu, j, i, k) - 2.0 * au);
    P3(f.bu, j, i, k) = au + d;
    P3(f.au, j, i, k) = cu;
    d = s.gnu1 * (cv + P3(f.bv, j, i, k) - 2.0 * av);
    P3(f.bv, j, i, k) = av + d;
    P3(f.av, j, i, k) = cv;
  }
  if (j <= g.jce2 && i <= g.ice2) {
    const bool interior = (j >= g.jci1 && j <= g.jci2 && i >= g.ici1 && i <= g.ici2);
    const double psb = L2(f.psb, j, i);
    if (interior && k <= kz) {
      const double cp = psb * L3(f.cpp, j, i, k);
      const double ap = P3(f.app, j, i, k);
      const double d = s.gnu1 * (cp + P3(f.bpp, j, i, k) - 2.0 * ap);
      P3(f.bpp, j, i, k) = ap + d;
      P3(f.app, j, i, k) = cp;
    }
    double aw = P3(f.aw, j, i, k), bw = P3(f.bw, j, i, k);
    if (interior) {
      double cw = L3(f.cw, j, i, k);
      if (fabs(cw) < dlowval) cw = 0.0;
      cw = psb * cw;
      const double d = s.gnu2 * (cw + bw - 2.0 * aw);
      bw = aw + d;
      aw = cw;
    }
    if (fabs(bw) < dlowval) bw = 0.0;
    if (fabs(aw) < dlowval) aw = 0.0;
    P3(f.bw, j, i, k) = bw;
    P3(f.aw, j, i, k) = aw;
  }
}

// ---------------------------------------------------------------------------
void launch_snd_init(const Geom& g, const Fields& f, const StepScalars& s, cudaStream_t st) {
  const int nj = (g.jde2 + g.gr) - (g.jde1 - g.gl) + 1, ni = (g.ide2 + g.gt) - (g.ide1 - g.gb) + 1;
  dim3 b(32, 8), gr((nj + 31) / 32, (ni + 7) / 8);
  RCM_LAUNCH(k_snd_init, gr, b, st, g, f, s);
}
void launch_snd_uv(const Geom& g, const Fields& f, const StepScalars& s, cudaStream_t st) {
  const int nj = g.jdi2 + g.gr - g.jdi1 + 1, ni = g.idi2 + g.gt - g.idi1 + 1;
  dim3 b(32, 4, 2), gr((nj + 31) / 32, (ni + 3) / 4, (g.kz + 1) / 2);
  RCM_LAUNCH(k_snd_uv, gr, b, st, g, f, s);
}
void launch_snd_wfwd(const Geom& g, const Fields& f, const StepScalars& s, cudaStream_t st) {
  const int nj = g.jci2 - g.jci1 + 1, ni = g.ici2 - g.ici1 + 1;
  const dim3 b = rcm::col_block(); dim3 gr((nj + b.x - 1) / b.x, (ni + b.y - 1) / b.y);
  if (s.it == 1) RCM_LAUNCH(k_snd_wfwd<true>, gr, b, st, g, f, s);
  else RCM_LAUNCH(k_snd_wfwd<false>, gr, b, st, g, f, s);
}
void launch_snd_wback(const Geom& g, const Fields& f, const StepScalars& s, cudaStream_t st) {
  const int nj = g.jci2 - g.jci1 + 1, ni = g.ici2 - g.ici1 + 1;
  const dim3 b = rcm::col_block(); dim3 gr((nj + b.x - 1) / b.x, (ni + b.y - 1) / b.y);
  RCM_LAUNCH(k_snd_wback, gr, b, st, g, f, s);
}
void launch_snd_final(const Geom& g, const Fields& f, const StepScalars& s, cudaStream_t st) {
  const int nj = g.jde2 - g.jde1 + 1, ni = g.ide2 - g.ide1 + 1;
  dim3 b(32, 4, 2), gr((nj + 31) / 32, (ni + 3) / 4, (g.kzp1 + 1) / 2);
  RCM_LAUNCH(k_snd_final, gr, b, st, g, f, s);
}
