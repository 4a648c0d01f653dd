// k_tq with its stencil operands staged through shared memory by the TMA engine
// (see k_uv_tma.cu for the scheme).  Tiles per level: au, av, th, aqv, aqc (1 ghost
// cell used), bt, bqv, bqc (2 ghost cells, 4th-order diffusion); the 2-D 1/p* table
// the diffusion operands are scaled with is staged once per block.
#include "dev_common.cuh"
#include "launch.h"
#include "tma_stage.cuh"

#ifndef RCM_EMU
using namespace rcm;

namespace {
#ifndef RCM_TQ_STAGES
#define RCM_TQ_STAGES 3  // 3: level k+1 operands from the next tile; 2: from global memory (more blocks per SM)
#endif
constexpr int kBY = 8, kTW = 40, kTH = kBY + 4, kTile = kTW * kTH, kNT = 8, kStages = RCM_TQ_STAGES;
constexpr int kAhead = kStages - 1;  // levels in flight ahead of the one being computed
constexpr int kRowBytes = kTW * 8, kStageBytes = kNT * kTile * 8;
enum { T_AU, T_AV, T_TH, T_AQV, T_AQC, T_BT, T_BQV, T_BQC };

__global__ void __launch_bounds__(256, kStages == 2 ? 3 : 2)
k_tq_tma(const Geom g, const Fields f, const StepScalars s) {
  extern __shared__ __align__(128) unsigned char smraw[];
  double* tiles = reinterpret_cast<double*>(smraw);
  double* rtile = tiles + (size_t)kStages * kNT * kTile;  // rpsb, 2-D
  uint64_t* full = reinterpret_cast<uint64_t*>(rtile + kTile);  // kStages + 1 barriers
  const int si = g.si;
  const long long sk = g.sk;
  const int tx = threadIdx.x, ty = threadIdx.y, tid = ty * 32 + tx;
  const int j0 = g.jci1 + blockIdx.x * 32, i0 = g.ici1 + blockIdx.y * kBY;
  const bool valid = (j0 + tx <= g.jci2) && (i0 + ty <= g.ici2);
  const int j = min(j0 + tx, g.jci2), i = min(i0 + ty, g.ici2);
  const int kz = g.kz;
  const int jlo = g.jde1 - RCM_HALO;
  int jstart = (j0 - 2) - (((j0 - 2) - jlo) & 1);
  jstart = min(jstart, jlo + si - kTW);
  const int ihi = g.ide2 + RCM_HALO;
  const int to = (i - (i0 - 2)) * kTW + (j - jstart);

  if (tid == 0) {
    for (int q = 0; q < kStages + 1; ++q) mbar_init(full + q, 1);
    mbar_init_fence();
  }
  __syncthreads();
  const int cq = tid / kTH, cr_ = tid % kTH;
  const double* csrc = nullptr;
  if (tid < kNT * kTH) {
    const double* fld = cq == T_AU ? f.au : cq == T_AV ? f.av : cq == T_TH ? f.th : cq == T_AQV ? f.aqv
                        : cq == T_AQC ? f.aqc : cq == T_BT ? f.bt : cq == T_BQV ? f.bqv : f.bqc;
    csrc = &P3(fld, jstart, min(i0 - 2 + cr_, ihi), 0);
  }
  auto issue = [&](int level) {  // level = 1..kz
    const int st = (level - 1) % kStages;
    if (tid == 0) mbar_expect_tx(full + st, kStageBytes);
    if (tid < kNT * kTH)
      bulk_load(tiles + (size_t)(st * kNT + cq) * kTile + cr_ * kTW, csrc + (long long)level * sk, kRowBytes, full + st);
  };
  // 2-D 1/p* tile
  if (tid == 0) mbar_expect_tx(full + kStages, kTile * 8);
  if (tid < kTH) bulk_load(rtile + tid * kTW, &P2(f.rpsb, jstart, min(i0 - 2 + tid, ihi)), kRowBytes, full + kStages);
  issue(1);
  if (kAhead >= 2 && kz >= 2) issue(2);

  const bool up = s.upstream_mode != 0;
  const bool se = s.stability_enhance != 0;
  const double ul = s.ul, xt = s.xt, dt = s.dt;
  const double psa = L2(f.psa, j, i), psb = L2(f.psb, j, i);
  const double rpsa_c = L2(f.rpsa, j, i), rpsa_jm = L2(f.rpsa, j - 1, i), rpsa_jp = L2(f.rpsa, j + 1, i),
               rpsa_im = L2(f.rpsa, j, i - 1), rpsa_ip = L2(f.rpsa, j, i + 1);
  const double xmsf = L2(f.xmsf, j, i);
  const double m00 = L2(f.msfd, j, i), m01 = L2(f.msfd, j, i + 1), m10 = L2(f.msfd, j + 1, i), m11 = L2(f.msfd, j + 1, i + 1);
  const int ib = (s.iboudy == 1 || s.iboudy == 5) ? L2(f.zone_c, j, i) : 0;
  const bool dint = diff_interior(g, j, i, false);
  const int nring = ring_count(g, j, i, false);
  const double qthr = minqq * psa;           // vadvqv threshold
  const double cthr = minqq * minqq * psa;   // vadv4d ind=1 threshold
  const double beta = 0.53;

  double fx_th_k = 0.0, fl_qv_k = 0.0, fl_qc_k = 0.0;
  mbar_wait(full + kStages, 0);
  mbar_wait(full + 0, 0);  // level 1
  double tha_k = tiles[T_TH * kTile + to] * psa;
  const double* rp = rtile + to;

  for (int k = 1; k <= kz; ++k) {
    __syncthreads();  // everyone is done with level k-1: its stage can take level k+kAhead
    if (k + kAhead <= kz) issue(k + kAhead);
    mbar_wait(full + (k - 1) % kStages, ((k - 1) / kStages) & 1);  // level k
    if (kStages >= 3 && k < kz) mbar_wait(full + k % kStages, (k / kStages) & 1);  // level k+1
    const double* tk = tiles + (size_t)((k - 1) % kStages) * kNT * kTile + to;
    const double* tn = tiles + (size_t)(k % kStages) * kNT * kTile + to;
    (void)tn;
#define TK_AU(dj, di) tk[T_AU * kTile + (di) * kTW + (dj)]
#define TK_AV(dj, di) tk[T_AV * kTile + (di) * kTW + (dj)]
#define TK_TH(dj, di) tk[T_TH * kTile + (di) * kTW + (dj)]
#define TK_AQV(dj, di) tk[T_AQV * kTile + (di) * kTW + (dj)]
#define TK_AQC(dj, di) tk[T_AQC * kTile + (di) * kTW + (dj)]
#define TK_BT(dj, di) tk[T_BT * kTile + (di) * kTW + (dj)]
#define TK_BQV(dj, di) tk[T_BQV * kTile + (di) * kTW + (dj)]
#define TK_BQC(dj, di) tk[T_BQC * kTile + (di) * kTW + (dj)]
#if RCM_TQ_STAGES >= 3
#define TN_TH(dj, di) tn[T_TH * kTile + (di) * kTW + (dj)]
#define TN_AQV(dj, di) tn[T_AQV * kTile + (di) * kTW + (dj)]
#define TN_AQC(dj, di) tn[T_AQC * kTile + (di) * kTW + (dj)]
#else
#define TN_TH(dj, di) L3(f.th, j + (dj), i + (di), k + 1)
#define TN_AQV(dj, di) L3(f.aqv, j + (dj), i + (di), k + 1)
#define TN_AQC(dj, di) L3(f.aqc, j + (dj), i + (di), k + 1)
#endif
#define RPSB(dj, di) rp[(di) * kTW + (dj)]
#define RCM_STORE_OK valid
#include "k_tq_level.inc"
#undef RCM_STORE_OK
#undef RPSB
  }
}
constexpr int kSmem = kStages * kStageBytes + kTile * 8 + 64;
}  // namespace

bool tq_tma_setup() {
  if (cudaFuncSetAttribute(k_tq_tma, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem) != cudaSuccess) { cudaGetLastError(); return false; }
  return true;
}

void launch_tq_tma(const Geom& g, const Fields& f, const StepScalars& s, cudaStream_t st) {
  const int nj = g.jci2 - g.jci1 + 1, ni = g.ici2 - g.ici1 + 1;
  dim3 b(32, kBY), gr((nj + 31) / 32, (ni + kBY - 1) / kBY);
  k_tq_tma<<<gr, b, kSmem, st>>>(g, f, s);
}
#else
bool tq_tma_setup() { return false; }
void launch_tq_tma(const Geom&, const Fields&, const StepScalars&, cudaStream_t) {}
#endif
