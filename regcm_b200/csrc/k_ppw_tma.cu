// k_ppw with ALL its per-level operands staged through shared memory by the TMA
// engine (see k_uv_tma.cu for the scheme).  Halo tiles (12 rows): au, av, app, aw,
// bpp, bw; centre tiles (8 rows): at, aqv, aqc, rho0, rho1, t0, cr, xkc, qdot; the
// 2-D 1/p* table is staged once.  Two stages (level k computed, k+1 in flight);
// the three level-(k+1) centre values (app, aw, qdot) are plain global loads issued
// at the top of the iteration.  The loop runs to kz+1 (w lives on full levels).
#include "dev_common.cuh"
#include "launch.h"
#include "tma_stage.cuh"

#ifndef RCM_EMU
using namespace rcm;

namespace {
constexpr int kBY = 8, kTW = 40, kTH = kBY + 4, kTileH = kTW * kTH, kTileC = kTW * kBY, kNH = 6, kNC = 9, kStages = 2;
constexpr int kRowBytes = kTW * 8, kStageDbl = kNH * kTileH + kNC * kTileC, kStageBytes = kStageDbl * 8;
constexpr int kCopies = kNH * kTH + kNC * kBY;  // 144 row copies per level
enum { T_AU, T_AV, T_APP, T_AW, T_BPP, T_BW };
enum { C_AT, C_AQV, C_AQC, C_RHO0, C_RHO1, C_T0, C_CR, C_XKC, C_QDOT };

__global__ void __launch_bounds__(256, RCM_MB_STAGED)
k_ppw_tma(const Geom g, const Fields f, const StepScalars s) {
  extern __shared__ __align__(128) unsigned char smraw[];
  double* tiles = reinterpret_cast<double*>(smraw);
  double* rtile = tiles + (size_t)kStages * kStageDbl;  // rpsb, 2-D, halo rows
  uint64_t* full = reinterpret_cast<uint64_t*>(rtile + kTileH);  // kStages + 1 barriers
  const int si = g.si;
  const long long sk = g.sk;
  const int tx = threadIdx.x, ty = threadIdx.y, tid = ty * 32 + tx;
  const int j0 = g.jci1 + blockIdx.x * 32, i0 = g.ici1 + blockIdx.y * kBY;
  const bool valid = (j0 + tx <= g.jci2) && (i0 + ty <= g.ici2);
  const int j = min(j0 + tx, g.jci2), i = min(i0 + ty, g.ici2);
  const int kz = g.kz, kzp1 = g.kzp1;
  const int jlo = g.jde1 - g.hw;
  int jstart = (j0 - 2) - (((j0 - 2) - jlo) & 1);
  jstart = min(jstart, jlo + si - kTW);
  const int ihi = g.ide2 + RCM_HALO;
  const int to = (i - (i0 - 2)) * kTW + (j - jstart);  // own cell in a halo tile
  const int tc = (i - i0) * kTW + (j - jstart);        // own cell in a centre tile

  if (tid == 0) {
    for (int q = 0; q < kStages + 1; ++q) mbar_init(full + q, 1);
    mbar_init_fence();
  }
  __syncthreads();
  // row copy number tid of a level
  const double* csrc = nullptr;
  int cdst = 0;
  if (tid < kNH * kTH) {
    const int q = tid / kTH, r = tid % kTH;
    const double* fld = q == T_AU ? f.au : q == T_AV ? f.av : q == T_APP ? f.app : q == T_AW ? f.aw : q == T_BPP ? f.bpp : f.bw;
    csrc = &P3(fld, jstart, min(i0 - 2 + r, ihi), 0);
    cdst = q * kTileH + r * kTW;
  } else if (tid < kCopies) {
    const int c = tid - kNH * kTH, q = c / kBY, r = c % kBY;
    const double* fld = q == C_AT ? f.at : q == C_AQV ? f.aqv : q == C_AQC ? f.aqc : q == C_RHO0 ? f.rho0 : q == C_RHO1 ? f.rho1
                        : q == C_T0 ? f.t0 : q == C_CR ? f.cr : q == C_XKC ? f.xkc : f.qdot;
    csrc = &P3(fld, jstart, min(i0 + r, ihi), 0);
    cdst = kNH * kTileH + q * kTileC + r * kTW;
  }
  auto issue = [&](int level) {  // level = 1..kz+1
    const int st = (level - 1) % kStages;
    if (tid == 0) mbar_expect_tx(full + st, kStageBytes);
    if (tid < kCopies) bulk_load(tiles + (size_t)st * kStageDbl + cdst, csrc + (long long)level * sk, kRowBytes, full + st);
  };
  if (tid == 0) mbar_expect_tx(full + kStages, kTileH * 8);
  if (tid < kTH) bulk_load(rtile + tid * kTW, &P2(f.rpsb, jstart, min(i0 - 2 + tid, ihi)), kRowBytes, full + kStages);
  issue(1);

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
  mbar_wait(full + kStages, 0);
  const double* rp = rtile + to;

  for (int k = 1; k <= kzp1; ++k) {
    __syncthreads();  // everyone is done with level k-1: its stage can take level k+1
    if (k + 1 <= kzp1) issue(k + 1);
    // level k+1 centre values straight from global memory, ahead of the wait
    const double n_app = (k < kz) ? L3(f.app, j, i, k + 1) : 0.0;
    const double n_aw = (k <= kz) ? L3(f.aw, j, i, k + 1) : 0.0;
    const double n_qdot = (k <= kz) ? L3(f.qdot, j, i, k + 1) : 0.0;
    mbar_wait(full + (k - 1) % kStages, ((k - 1) / kStages) & 1);  // level k
    const double* tk = tiles + (size_t)((k - 1) % kStages) * kStageDbl + to;
    const double* ck = tiles + (size_t)((k - 1) % kStages) * kStageDbl + kNH * kTileH + tc;
#define TK_AU(dj, di) tk[T_AU * kTileH + (di) * kTW + (dj)]
#define TK_AV(dj, di) tk[T_AV * kTileH + (di) * kTW + (dj)]
#define TK_APP(dj, di) tk[T_APP * kTileH + (di) * kTW + (dj)]
#define TK_AW(dj, di) tk[T_AW * kTileH + (di) * kTW + (dj)]
#define TK_BPP(dj, di) tk[T_BPP * kTileH + (di) * kTW + (dj)]
#define TK_BW(dj, di) tk[T_BW * kTileH + (di) * kTW + (dj)]
#define TN_APP(dj, di) n_app
#define TN_AW(dj, di) n_aw
#define CN_QDOT n_qdot
#define CK_AT ck[C_AT * kTileC]
#define CK_AQV ck[C_AQV * kTileC]
#define CK_AQC ck[C_AQC * kTileC]
#define CK_RHO0 ck[C_RHO0 * kTileC]
#define CK_RHO1 ck[C_RHO1 * kTileC]
#define CK_T0 ck[C_T0 * kTileC]
#define CK_CR ck[C_CR * kTileC]
#define CK_XKC ck[C_XKC * kTileC]
#define CK_QDOT ck[C_QDOT * kTileC]
#define RPSB(dj, di) rp[(di) * kTW + (dj)]
#define RCM_STORE_OK valid
#include "k_ppw_level.inc"
#undef RCM_STORE_OK
#undef RPSB
  }
  (void)app_m;
}
constexpr int kSmem = kStages * kStageBytes + kTileH * 8 + 64;
}  // namespace

bool ppw_tma_setup() {
  if (cudaFuncSetAttribute(k_ppw_tma, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem) != cudaSuccess) { cudaGetLastError(); return false; }
  return true;
}

void launch_ppw_tma(const Geom& g, const Fields& f, const StepScalars& s, cudaStream_t st) {
  const int nj = g.jci2 - g.jci1 + 1, ni = g.ici2 - g.ici1 + 1;
  dim3 b(32, kBY), gr((nj + 31) / 32, (ni + kBY - 1) / kBY);
  k_ppw_tma<<<gr, b, kSmem, st>>>(g, f, s);
}
#else
bool ppw_tma_setup() { return false; }
void launch_ppw_tma(const Geom&, const Fields&, const StepScalars&, cudaStream_t) {}
#endif
