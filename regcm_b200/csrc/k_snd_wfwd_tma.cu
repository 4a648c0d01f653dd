// k_snd_wfwd (sub-steps 2..istep, which load the step-invariant coefficients) with
// the operands of every loop iteration staged through shared memory by the TMA
// engine (see k_uv_tma.cu for the scheme).  One stage = one ITERATION k of the
// upward elimination: halo tiles (10 rows) cu, cv at level k-2 and pr0 at k-1;
// centre tiles (8 rows) pr0 at k-2, rho0, scc, scdd, t0, nbt, ppten, cpp, cw at
// k-1, wten and sca at k.  Two stages: iteration k computes while k-1 is in
// flight.  The prologue (level kz) uses plain loads.  Arithmetic = the shared
// k_snd_wfwd_body.inc, so the result is bit-identical to the plain kernel.
// OPT-IN (RCM_WFWD_TMA=1): on C4 it measured 6.1-8.0 ms/step over block shapes,
// 2-4 stages and 1-4 blocks/SM against 5.6 ms for the plain kernel - the sweep is
// 11 centre streams to 3 stencil streams, so there is little reuse to gain and the
// ring adds a block-wide barrier per level.
#include "dev_common.cuh"
#include "launch.h"
#include "tma_stage.cuh"

#ifndef RCM_EMU
using namespace rcm;

#ifndef RCM_WF_BY
#define RCM_WF_BY 8
#endif
#ifndef RCM_WF_STAGES
#define RCM_WF_STAGES 2
#endif
#ifndef RCM_WF_MB
#define RCM_WF_MB 2
#endif

namespace {
constexpr int kBY = RCM_WF_BY, kTW = 40, kTH = kBY + 2, kTileH = kTW * kTH, kTileC = kTW * kBY, kNH = 3, kNC = 11, kStages = RCM_WF_STAGES;
constexpr int kRowBytes = kTW * 8, kStageDbl = kNH * kTileH + kNC * kTileC, kStageBytes = kStageDbl * 8;
constexpr int kCopies = kNH * kTH + kNC * kBY;  // row copies per iteration (118 at kBY = 8)
enum { T_CU, T_CV, T_PR0 };
enum { C_PR0N, C_RHO0, C_SCC, C_SCDD, C_T0, C_NBT, C_PPTEN, C_CPP, C_CW, C_WTEN, C_SCA };
constexpr int kSmem = kStages * kStageBytes + 64;
static_assert(kCopies <= 32 * kBY, "one row copy per thread");

__global__ void __launch_bounds__(32 * kBY, RCM_WF_MB)
k_snd_wfwd_tma(const Geom g, const Fields f, const StepScalars s) {
  constexpr bool FIRST = false;
  extern __shared__ __align__(128) unsigned char smraw[];
  double* tiles = reinterpret_cast<double*>(smraw);
  uint64_t* full = reinterpret_cast<uint64_t*>(tiles + (size_t)kStages * kStageDbl);
  const int si = g.si;
  const long long sk = g.sk;
  const int tx = threadIdx.x, ty = threadIdx.y, tid = ty * 32 + tx;
  const int j0 = g.jci1 + blockIdx.x * 32, i0 = g.ici1 + blockIdx.y * kBY;
  const bool valid = (j0 + tx <= g.jci2) && (i0 + ty <= g.ici2);
  const int j = min(j0 + tx, g.jci2), i = min(i0 + ty, g.ici2);
  const int kz = g.kz, kzp1 = g.kzp1;
  const double dts = s.dts, dx = s.dx, xgamma = s.xgamma, bp = s.bp, bm = s.bm, bpxbp = s.bpxbp, bpxbm = s.bpxbm;
  const int jlo = g.jde1 - g.hw;
  int jstart = (j0 - 1) - (((j0 - 1) - jlo) & 1);
  jstart = min(jstart, jlo + si - kTW);
  const int ihi = g.ide2 + RCM_HALO;
  const int to = (i - (i0 - 1)) * kTW + (j - jstart);  // own cell in a halo tile
  const int tc = (i - i0) * kTW + (j - jstart);        // own cell in a centre tile

  if (tid == 0) {
    for (int q = 0; q < kStages; ++q) mbar_init(full + q, 1);
    mbar_init_fence();
  }
  __syncthreads();
  // row copy number tid of an iteration: source row at level 0, level offset dk from k
  const double* csrc = nullptr;
  int cdst = 0, cdk = 0;
  if (tid < kNH * kTH) {
    const int q = tid / kTH, r = tid % kTH;
    const double* fld = q == T_CU ? f.cu : q == T_CV ? f.cv : f.pr0;
    cdk = (q == T_PR0) ? -1 : -2;
    csrc = &P3(fld, jstart, min(i0 - 1 + r, ihi), 0);
    cdst = q * kTileH + r * kTW;
  } else if (tid < kCopies) {
    const int c = tid - kNH * kTH, q = c / kBY, r = c % kBY;
    const double* fld = q == C_PR0N ? f.pr0 : q == C_RHO0 ? f.rho0 : q == C_SCC ? f.scc : q == C_SCDD ? f.scdd : q == C_T0 ? f.t0
                        : q == C_NBT ? f.nbt : q == C_PPTEN ? f.ppten : q == C_CPP ? f.cpp : q == C_CW ? f.cw
                        : q == C_WTEN ? f.wten : f.sca;
    cdk = (q == C_PR0N) ? -2 : (q == C_WTEN || q == C_SCA) ? 0 : -1;
    csrc = &P3(fld, jstart, min(i0 + r, ihi), 0);
    cdst = kNH * kTileH + q * kTileC + r * kTW;
  }
  auto issue = [&](int k) {  // iteration k = kz..2
    const int st = (kz - k) % kStages;
    if (tid == 0) mbar_expect_tx(full + st, kStageBytes);
    if (tid < kCopies)
      bulk_load(tiles + (size_t)st * kStageDbl + cdst, csrc + (long long)max(k + cdk, 1) * sk, kRowBytes, full + st);
  };
  for (int q = 0; q < kStages - 1; ++q)
    if (kz - q >= 2) issue(kz - q);

#define WF_OK valid
#define WF_ITER_BEGIN(k)                                                               \
  __syncthreads(); /* everyone is done with iteration k+1: its stage is free */        \
  if ((k) - (kStages - 1) >= 2) issue((k) - (kStages - 1));                            \
  mbar_wait(full + (kz - (k)) % kStages, ((kz - (k)) / kStages) & 1);                  \
  const double* th = tiles + (size_t)((kz - (k)) % kStages) * kStageDbl + to;          \
  const double* tcn = tiles + (size_t)((kz - (k)) % kStages) * kStageDbl + kNH * kTileH + tc
#define WF_LOADUV_ITER(U, V)                                                                        \
  U[0] = th[T_CU * kTileH]; U[1] = th[T_CU * kTileH + 1]; U[2] = th[T_CU * kTileH + kTW]; U[3] = th[T_CU * kTileH + kTW + 1]; \
  V[0] = th[T_CV * kTileH]; V[1] = th[T_CV * kTileH + 1]; V[2] = th[T_CV * kTileH + kTW]; V[3] = th[T_CV * kTileH + kTW + 1]
#define WF_LOADRAW_ITER(r)                                                                  \
  r.rho0 = tcn[C_RHO0 * kTileC]; r.pr1 = tcn[C_SCC * kTileC]; r.rho1 = tcn[C_SCDD * kTileC]; \
  r.t0 = tcn[C_T0 * kTileC]; r.nbt = tcn[C_NBT * kTileC]; r.ppten = tcn[C_PPTEN * kTileC];   \
  r.pp = tcn[C_CPP * kTileC]; r.wo = tcn[C_CW * kTileC];                                     \
  r.p0w = th[T_PR0 * kTileH - 1]; r.p0e = th[T_PR0 * kTileH + 1];                            \
  r.p0s = th[T_PR0 * kTileH - kTW]; r.p0n = th[T_PR0 * kTileH + kTW];                        \
  r.ds = LD1(&f.dsigma[km1])
#define WF_P0N tcn[C_PR0N * kTileC]
#define WF_WTEN tcn[C_WTEN * kTileC]
#define WF_CA tcn[C_SCA * kTileC]
#include "k_snd_wfwd_body.inc"
#undef WF_OK
#undef WF_ITER_BEGIN
#undef WF_LOADUV_ITER
#undef WF_LOADRAW_ITER
#undef WF_P0N
#undef WF_WTEN
#undef WF_CA
  (void)dx;
}
}  // namespace

bool wfwd_tma_setup() {
  if (cudaFuncSetAttribute(k_snd_wfwd_tma, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem) != cudaSuccess) { cudaGetLastError(); return false; }
  return true;
}

void launch_snd_wfwd_tma(const Geom& g, const Fields& f, const StepScalars& s, cudaStream_t st) {
  const int nj = g.jci2 - g.jci1 + 1, ni = g.ici2 - g.ici1 + 1;
  dim3 b(32, kBY), gr((nj + 31) / 32, (ni + kBY - 1) / kBY);
  k_snd_wfwd_tma<<<gr, b, kSmem, st>>>(g, f, s);
}
#else
bool wfwd_tma_setup() { return false; }
void launch_snd_wfwd_tma(const Geom&, const Fields&, const StepScalars&, cudaStream_t) {}
#endif
