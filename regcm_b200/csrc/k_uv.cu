// K_mom_slow (SURVEY.md §8 a8,a9,a14,a15,a12,a16): the complete slow tendency
// of the coupled horizontal momentum p*u, p*v on one dot-point column, marched
// in k with the vertical fluxes kept in registers.  Fuses, in the reference's
// order of accumulation (so that the result is bit-comparable without FMA):
//   hadvuv  NH branch        Main/mod_advection.F90:235-265
//   vadvuv                   Main/mod_advection.F90:271-303
//   curvature NH branch      Main/mod_tendency.F90:1840-1879
//   nudgeuv (iboudy 1|5)     Main/mod_bdycod.F90:3581-3823
//   diffu_d (idiffu = 1, 2)  Main/mod_diffusion.F90:266-411
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
  const int ibs = (s.iboudy == 4) ? L2(f.zone_d, j, i) : 0;   // sponge zone (iboudy = 4)
  const bool d9 = (s.idiffu == 2);
  const bool d6on = (s.idiffu == 3) && j == g.jdi2;   // the one column idiffu = 3 acts on (k_diff6)   // nine-point operator on every interior point, no boundary ring
  const bool dint = (s.idiffu == 1) && diff_interior(g, j, i, true);
  const int nring = (s.idiffu == 1) ? ring_count(g, j, i, true) : 0;

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
#define TK_AU(dj, di) L3(f.au, j + (dj), i + (di), k)
#define TK_AV(dj, di) L3(f.av, j + (dj), i + (di), k)
#define TK_CR(dj, di) L3(f.cr, j + (dj), i + (di), k)
#define TK_XKC(dj, di) L3(f.xkc, j + (dj), i + (di), k)
#define TK_DUB(dj, di) L3(f.dub, j + (dj), i + (di), k)
#define TK_DVB(dj, di) L3(f.dvb, j + (dj), i + (di), k)
#define TN_AU(dj, di) L3(f.au, j + (dj), i + (di), k + 1)
#define TN_AV(dj, di) L3(f.av, j + (dj), i + (di), k + 1)
#define TN_QDOT(dj, di) L3(f.qdot, j + (dj), i + (di), k + 1)
#define TN_AW(dj, di) L3(f.aw, j + (dj), i + (di), k + 1)
#include "k_uv_level.inc"
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
