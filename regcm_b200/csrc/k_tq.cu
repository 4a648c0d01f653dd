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

template <int CQ_MODE>   // StepScalars::cq_mode (negative-moisture repair), compile-time
__global__ void __launch_bounds__(RCM_MAXT, RCM_MB_TQ)
k_tq(const Geom g, const Fields f, const StepScalars s) {
  const int si = g.si;
  const long long sk = g.sk;
  const int j = g.jci1 + blockIdx.x * blockDim.x + threadIdx.x;
  const int i = g.ici1 + blockIdx.y * blockDim.y + threadIdx.y;
  if (j > g.jci2 || i > g.ici2) return;
  if (CQ_MODE == 1 && *f.negq == 0ULL) return;  // second pass of the negative-moisture repair: nothing went negative
  const bool cq_edge = tile_edge(g, j, i);
  const int kz = g.kz;
  const bool up = s.upstream_mode != 0;
  const bool se = s.stability_enhance != 0;
  const double ul = s.ul, xt = s.xt, dt = s.dt;

  const double psa = L2(f.psa, j, i), psb = L2(f.psb, j, i);
  const double rdiv_psa = recip_for_div(psa);
  const double rpsa_c = L2(f.rpsa, j, i), rpsa_jm = L2(f.rpsa, j - 1, i), rpsa_jp = L2(f.rpsa, j + 1, i),
               rpsa_im = L2(f.rpsa, j, i - 1), rpsa_ip = L2(f.rpsa, j, i + 1);
  const double xmsf = L2(f.xmsf, j, i);
  const double m00 = L2(f.msfd, j, i), m01 = L2(f.msfd, j, i + 1), m10 = L2(f.msfd, j + 1, i), m11 = L2(f.msfd, j + 1, i + 1);
  const int ib = (s.iboudy == 1 || s.iboudy == 5) ? L2(f.zone_c, j, i) : 0;
  const int ibs = (s.iboudy == 4) ? L2(f.zone_c, j, i) : 0;   // sponge zone (iboudy = 4)
  const bool d9 = (s.idiffu == 2);
  const bool d6on = (s.idiffu == 3) && j == g.jci2;   // the one column idiffu = 3 acts on (k_diff6)   // nine-point operator on every interior point, no boundary ring
  const bool dint = (s.idiffu == 1) && diff_interior(g, j, i, false);
  const int nring = (s.idiffu == 1) ? ring_count(g, j, i, false) : 0;
  const double qthr = minqq * psa;           // vadvqv threshold
  const double cthr = minqq * minqq * psa;   // vadv4d ind=1 threshold
  const double beta = 0.53;
  const int kpb_c = f.kpb ? L2(f.kpb, j, i) : 0;   // top of the PBL for vadv4d ind = 3

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
#define TK_AU(dj, di) L3(f.au, j + (dj), i + (di), k)
#define TK_AV(dj, di) L3(f.av, j + (dj), i + (di), k)
#define TK_TH(dj, di) L3(f.th, j + (dj), i + (di), k)
#define TK_AQV(dj, di) L3(f.aqv, j + (dj), i + (di), k)
#define TK_AQC(dj, di) L3(f.aqc, j + (dj), i + (di), k)
#define TK_BT(dj, di) L3(f.bt, j + (dj), i + (di), k)
#define TK_BQV(dj, di) L3(f.bqv, j + (dj), i + (di), k)
#define TK_BQC(dj, di) L3(f.bqc, j + (dj), i + (di), k)
#define TN_TH(dj, di) L3(f.th, j + (dj), i + (di), k + 1)
#define TN_AQV(dj, di) L3(f.aqv, j + (dj), i + (di), k + 1)
#define TN_AQC(dj, di) L3(f.aqc, j + (dj), i + (di), k + 1)
#define RCM_STORE_OK true
#define RPSB(dj, di) L2(f.rpsb, j + (dj), i + (di))
#define CK_CR L3(f.cr, j, i, k)
#define CK_XKC L3(f.xkc, j, i, k)
#define CK_AT L3(f.at, j, i, k)
#define CN_QDOT L3(f.qdot, j, i, k + 1)
#include "k_tq_level.inc"
#undef RCM_STORE_OK
#undef RPSB
  }
}

void launch_tq(const Geom& g, const Fields& f, const StepScalars& s, cudaStream_t st) {
  const int nj = g.jci2 - g.jci1 + 1, ni = g.ici2 - g.ici1 + 1;
  const dim3 b = rcm::col_block(); dim3 gr((nj + b.x - 1) / b.x, (ni + b.y - 1) / b.y);
  if (s.cq_mode == 0) RCM_LAUNCH(k_tq<0>, gr, b, st, g, f, s);
  else if (s.cq_mode == 1) RCM_LAUNCH(k_tq<1>, gr, b, st, g, f, s);
  else RCM_LAUNCH(k_tq<2>, gr, b, st, g, f, s);
}
