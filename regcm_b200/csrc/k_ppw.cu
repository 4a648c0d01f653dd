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
  const double rdiv_psa = recip_for_div(psa);
  const double rpsa_c = L2(f.rpsa, j, i), rpsa_jm = L2(f.rpsa, j - 1, i), rpsa_jp = L2(f.rpsa, j + 1, i),
               rpsa_im = L2(f.rpsa, j, i - 1), rpsa_ip = L2(f.rpsa, j, i + 1);
  const double rpsb_c = L2(f.rpsb, j, i), rpsb_jm = L2(f.rpsb, j - 1, i), rpsb_jp = L2(f.rpsb, j + 1, i),
               rpsb_im = L2(f.rpsb, j, i - 1), rpsb_ip = L2(f.rpsb, j, i + 1);
  const double xmsf = L2(f.xmsf, j, i);
  const double m00 = L2(f.msfd, j, i), m01 = L2(f.msfd, j, i + 1), m10 = L2(f.msfd, j + 1, i), m11 = L2(f.msfd, j + 1, i + 1);
  const double ex = L2(f.ex, j, i), crx = L2(f.crx, j, i), cry = L2(f.cry, j, i);
  const int ib = (s.iboudy == 1 || s.iboudy == 5) ? L2(f.zone_c, j, i) : 0;
  const int ibs = (s.iboudy == 4) ? L2(f.zone_c, j, i) : 0;   // sponge zone (iboudy = 4)
  const bool d9 = (s.idiffu == 2);
  const bool d6on = (s.idiffu == 3) && j == g.jci2;   // the one column idiffu = 3 acts on (k_diff6)   // nine-point operator on every interior point, no boundary ring
  const bool dint = (s.idiffu == 1) && diff_interior(g, j, i, false);
  const int nring = (s.idiffu == 1) ? ring_count(g, j, i, false) : 0;

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
#define TK_AU(dj, di) L3(f.au, j + (dj), i + (di), k)
#define TK_AV(dj, di) L3(f.av, j + (dj), i + (di), k)
#define TK_APP(dj, di) L3(f.app, j + (dj), i + (di), k)
#define TK_AW(dj, di) L3(f.aw, j + (dj), i + (di), k)
#define TK_BPP(dj, di) L3(f.bpp, j + (dj), i + (di), k)
#define TK_BW(dj, di) L3(f.bw, j + (dj), i + (di), k)
#define TN_APP(dj, di) L3(f.app, j + (dj), i + (di), k + 1)
#define TN_AW(dj, di) L3(f.aw, j + (dj), i + (di), k + 1)
#define RPSB(dj, di) L2(f.rpsb, j + (dj), i + (di))
#define CK_AT L3(f.at, j, i, k)
#define CK_AQV L3(f.aqv, j, i, k)
#define CK_AQC L3(f.aqc, j, i, k)
#define CK_RHO0 L3(f.rho0, j, i, k)
#define CK_RHO1 L3(f.rho1, j, i, k)
#define CK_T0 L3(f.t0, j, i, k)
#define CK_CR L3(f.cr, j, i, k)
#define CK_XKC L3(f.xkc, j, i, k)
#define CK_QDOT L3(f.qdot, j, i, k)
#define CN_QDOT L3(f.qdot, j, i, k + 1)
#define RCM_STORE_OK true
#include "k_ppw_level.inc"
#undef RCM_STORE_OK
#undef RPSB
  }
  (void)app_m;
}

void launch_ppw(const Geom& g, const Fields& f, const StepScalars& s, cudaStream_t st) {
  const int nj = g.jci2 - g.jci1 + 1, ni = g.ici2 - g.ici1 + 1;
  const dim3 b = rcm::col_block(); dim3 gr((nj + b.x - 1) / b.x, (ni + b.y - 1) / b.y);
  RCM_LAUNCH(k_ppw, gr, b, st, g, f, s);
}
