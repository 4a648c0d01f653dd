// k_tq with its per-level operands staged through shared memory by the TMA engine
// (see k_uv_tma.cu for the scheme).  Halo tiles (12 rows): au, av, th, aqv, aqc
// (1 ghost cell used), bt, bqv, bqc (2 ghost cells, 4th-order diffusion); centre
// tiles (8 rows): cr, xkc, at; the 2-D 1/p* table is staged once.  Two stages
// (level k computed, k+1 in flight); the four level-(k+1) centre values (th, aqv,
// aqc, qdot) are plain global loads issued at the top of the iteration.
#include "dev_common.cuh"
#include "launch.h"
#include "tma_stage.cuh"

#ifndef RCM_EMU
using namespace rcm;

namespace {
constexpr int kBY = 8, kTW = 40, kTH = kBY + 4, kTileH = kTW * kTH, kTileC = kTW * kBY, kNH = 8, kNC = 3, kStages = 2;
constexpr int kRowBytes = kTW * 8, kStageDbl = kNH * kTileH + kNC * kTileC, kStageBytes = kStageDbl * 8;
constexpr int kCopies = kNH * kTH + kNC * kBY;  // 120 row copies per level
enum { T_AU, T_AV, T_TH, T_AQV, T_AQC, T_BT, T_BQV, T_BQC };
enum { C_CR, C_XKC, C_AT };

template <int CQ_MODE>   // StepScalars::cq_mode (negative-moisture repair), compile-time
__global__ void __launch_bounds__(256, RCM_MB_STAGED)
k_tq_tma(const Geom g, const Fields f, const StepScalars s) {
  if (CQ_MODE == 1 && *f.negq == 0ULL) return;  // second pass of the negative-moisture repair: nothing went negative
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
  const int kz = g.kz;
  const int jlo = g.jde1 - g.hw;
  int jstart = (j0 - 2) - (((j0 - 2) - jlo) & 1);
  jstart = min(jstart, jlo + si - kTW);
  const int ihi = g.ide2 + RCM_HALO;
  const int to = (i - (i0 - 2)) * kTW + (j - jstart);  // own cell in a halo tile
  const int tc = (i - i0) * kTW + (j - jstart);        // own cell in a centre tile
  const bool cq_edge = tile_edge(g, j, i);

  if (tid == 0) {
    for (int q = 0; q < kStages + 1; ++q) mbar_init(full + q, 1);
    mbar_init_fence();
  }
  __syncthreads();
  const double* csrc = nullptr;
  int cdst = 0;
  if (tid < kNH * kTH) {
    const int q = tid / kTH, r = tid % kTH;
    const double* fld = q == T_AU ? f.au : q == T_AV ? f.av : q == T_TH ? f.th : q == T_AQV ? f.aqv
                        : q == T_AQC ? f.aqc : q == T_BT ? f.bt : q == T_BQV ? f.bqv : f.bqc;
    csrc = &P3(fld, jstart, min(i0 - 2 + r, ihi), 0);
    cdst = q * kTileH + r * kTW;
  } else if (tid < kCopies) {
    const int c = tid - kNH * kTH, q = c / kBY, r = c % kBY;
    const double* fld = q == C_CR ? f.cr : q == C_XKC ? f.xkc : f.at;
    csrc = &P3(fld, jstart, min(i0 + r, ihi), 0);
    cdst = kNH * kTileH + q * kTileC + r * kTW;
  }
  auto issue = [&](int level) {  // level = 1..kz
    const int st = (level - 1) % kStages;
    if (tid == 0) mbar_expect_tx(full + st, kStageBytes);
    if (tid < kCopies) bulk_load(tiles + (size_t)st * kStageDbl + cdst, csrc + (long long)level * sk, kRowBytes, full + st);
  };
  if (tid == 0) mbar_expect_tx(full + kStages, kTileH * 8);
  if (tid < kTH) bulk_load(rtile + tid * kTW, &P2(f.rpsb, jstart, min(i0 - 2 + tid, ihi)), kRowBytes, full + kStages);
  issue(1);

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

  double fx_th_k = 0.0, fl_qv_k = 0.0, fl_qc_k = 0.0;
  double tha_k = L3(f.th, j, i, 1) * psa;
  mbar_wait(full + kStages, 0);
  const double* rp = rtile + to;

  for (int k = 1; k <= kz; ++k) {
    __syncthreads();  // everyone is done with level k-1: its stage can take level k+1
    if (k + 1 <= kz) issue(k + 1);
    // level k+1 centre values straight from global memory, ahead of the wait
    const bool nx = k < kz;
    const double n_th = nx ? L3(f.th, j, i, k + 1) : 0.0, n_aqv = nx ? L3(f.aqv, j, i, k + 1) : 0.0,
                 n_aqc = nx ? L3(f.aqc, j, i, k + 1) : 0.0, n_qdot = nx ? L3(f.qdot, j, i, k + 1) : 0.0;
    mbar_wait(full + (k - 1) % kStages, ((k - 1) / kStages) & 1);  // level k
    const double* tk = tiles + (size_t)((k - 1) % kStages) * kStageDbl + to;
    const double* ck = tiles + (size_t)((k - 1) % kStages) * kStageDbl + kNH * kTileH + tc;
#define TK_AU(dj, di) tk[T_AU * kTileH + (di) * kTW + (dj)]
#define TK_AV(dj, di) tk[T_AV * kTileH + (di) * kTW + (dj)]
#define TK_TH(dj, di) tk[T_TH * kTileH + (di) * kTW + (dj)]
#define TK_AQV(dj, di) tk[T_AQV * kTileH + (di) * kTW + (dj)]
#define TK_AQC(dj, di) tk[T_AQC * kTileH + (di) * kTW + (dj)]
#define TK_BT(dj, di) tk[T_BT * kTileH + (di) * kTW + (dj)]
#define TK_BQV(dj, di) tk[T_BQV * kTileH + (di) * kTW + (dj)]
#define TK_BQC(dj, di) tk[T_BQC * kTileH + (di) * kTW + (dj)]
#define TN_TH(dj, di) n_th
#define TN_AQV(dj, di) n_aqv
#define TN_AQC(dj, di) n_aqc
#define CN_QDOT n_qdot
#define CK_CR ck[C_CR * kTileC]
#define CK_XKC ck[C_XKC * kTileC]
#define CK_AT ck[C_AT * kTileC]
#define RPSB(dj, di) rp[(di) * kTW + (dj)]
#define RCM_STORE_OK valid
#include "k_tq_level.inc"
#undef RCM_STORE_OK
#undef RPSB
  }
}
constexpr int kSmem = kStages * kStageBytes + kTileH * 8 + 64;
}  // namespace

bool tq_tma_setup() {
  if (cudaFuncSetAttribute(k_tq_tma<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem) != cudaSuccess ||
      cudaFuncSetAttribute(k_tq_tma<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem) != cudaSuccess ||
      cudaFuncSetAttribute(k_tq_tma<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem) != cudaSuccess) { cudaGetLastError(); return false; }
  return true;
}

void launch_tq_tma(const Geom& g, const Fields& f, const StepScalars& s, cudaStream_t st) {
  const int nj = g.jci2 - g.jci1 + 1, ni = g.ici2 - g.ici1 + 1;
  dim3 b(32, kBY), gr((nj + 31) / 32, (ni + kBY - 1) / kBY);
  if (s.cq_mode == 0) k_tq_tma<0><<<gr, b, kSmem, st>>>(g, f, s);
  else if (s.cq_mode == 1) k_tq_tma<1><<<gr, b, kSmem, st>>>(g, f, s);
  else k_tq_tma<2><<<gr, b, kSmem, st>>>(g, f, s);
}
#else
bool tq_tma_setup() { return false; }
void launch_tq_tma(const Geom&, const Fields&, const StepScalars&, cudaStream_t) {}
#endif
