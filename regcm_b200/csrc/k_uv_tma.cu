// k_uv with its stencil operands staged through shared memory by the TMA engine
// (cp.async.bulk, SASS UBLKCP, mbarrier completion).  Same arithmetic as k_uv
// (the level body is the shared k_uv_level.inc); what changes is how the 8 fields
// with horizontal neighbours reach the threads: per level, for each of au, av,
// cr, xkc, dub, dvb, qdot, aw, the 12 rows x 40 columns around the block's 32x8
// columns (2 ghost cells on every side, start rounded down to a 16-byte
// boundary) are fetched as 12 row-wise bulk copies into a 3-stage ring (levels
// k and k+1 ready, k+2 in flight).  96 threads issue one copy each; the ~60
// per-thread global loads of a level become shared-memory reads and the HBM
// latency leaves the dependent chain.  (The tensor-map form, UTMALDG, raises
// "illegal instruction" on this driver even for a minimal rank-2 byte tensor -
// scripts/tma_test.cu - so the rows are copied one by one.)
#include "dev_common.cuh"
#include "launch.h"
#include "tma_stage.cuh"

#ifndef RCM_EMU
#include <cstdint>
#include <cstdio>

using namespace rcm;

namespace {
constexpr int kBY = 8, kTW = 40, kTH = kBY + 4, kTile = kTW * kTH, kNT = 8, kStages = 3;
constexpr int kRowBytes = kTW * 8, kStageBytes = kNT * kTile * 8;
enum { T_AU, T_AV, T_CR, T_XKC, T_DUB, T_DVB, T_QDOT, T_AW };

__global__ void __launch_bounds__(256, RCM_MB_STAGED)
k_uv_tma(const Geom g, const Fields f, const StepScalars s) {
  extern __shared__ __align__(128) unsigned char smraw[];
  double* tiles = reinterpret_cast<double*>(smraw);
  uint64_t* full = reinterpret_cast<uint64_t*>(smraw + kStages * kStageBytes);
  const int si = g.si;
  const long long sk = g.sk;
  const int tx = threadIdx.x, ty = threadIdx.y, tid = ty * 32 + tx;
  const int j0 = g.jdi1 + blockIdx.x * 32, i0 = g.idi1 + blockIdx.y * kBY;
  const bool valid = (j0 + tx <= g.jdi2) && (i0 + ty <= g.idi2);
  // threads beyond the tile edge stay alive for the barriers and work on a clamped column
  const int j = min(j0 + tx, g.jdi2), i = min(i0 + ty, g.idi2);
  const int kz = g.kz;
  // first column of the tile rows: j0-2 rounded down to an even array index (16-byte
  // aligned source), and kept inside the pitched row
  const int jlo = g.jde1 - g.hw;
  int jstart = (j0 - 2) - (((j0 - 2) - jlo) & 1);
  jstart = min(jstart, jlo + si - kTW);
  const int ihi = g.ide2 + RCM_HALO;
  const int to = (i - (i0 - 2)) * kTW + (j - jstart);  // this thread's own cell inside a tile

  if (tid == 0) {
    for (int q = 0; q < kStages; ++q) mbar_init(full + q, 1);
    mbar_init_fence();
  }
  __syncthreads();
  // copy number tid (0..95) of a level: field q = tid / 12, tile row r = tid % 12
  const int cq = tid / kTH, cr_ = tid % kTH;
  const double* csrc = nullptr;
  if (tid < kNT * kTH) {
    const double* fld = cq == T_AU ? f.au : cq == T_AV ? f.av : cq == T_CR ? f.cr : cq == T_XKC ? f.xkc
                        : cq == T_DUB ? f.dub : cq == T_DVB ? f.dvb : cq == T_QDOT ? f.qdot : f.aw;
    csrc = &P3(fld, jstart, min(i0 - 2 + cr_, ihi), 0);
  }
  auto issue = [&](int level) {  // level = 1..kz+1
    const int st = (level - 1) % kStages;
    if (tid == 0) mbar_expect_tx(full + st, kStageBytes);
    if (tid < kNT * kTH)
      bulk_load(tiles + (size_t)(st * kNT + cq) * kTile + cr_ * kTW, csrc + (long long)level * sk, kRowBytes, full + st);
  };
  issue(1);
  issue(2);

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

  double uu_k = 0.0, vv_k = 0.0;  // uu(1) = 0
  mbar_wait(full + 0, 0);         // level 1
  double wadot_k;
  {
    const double* t1 = tiles + T_AW * kTile + to;
    wadot_k = 0.125 * (t1[-kTW - 1] + t1[-1] + t1[-kTW] + t1[0]);  // aw(j-1,i-1), (j-1,i), (j,i-1), (j,i) at k = 1
  }

  for (int k = 1; k <= kz; ++k) {
    __syncthreads();  // everyone is done with level k-1: its stage can take level k+2
    if (k + 2 <= kz + 1) issue(k + 2);
    mbar_wait(full + (k - 1) % kStages, ((k - 1) / kStages) & 1);  // level k
    mbar_wait(full + k % kStages, (k / kStages) & 1);              // level k+1
    const double* tk = tiles + (size_t)((k - 1) % kStages) * kNT * kTile + to;
    const double* tn = tiles + (size_t)(k % kStages) * kNT * kTile + to;
#define TK_AU(dj, di) tk[T_AU * kTile + (di) * kTW + (dj)]
#define TK_AV(dj, di) tk[T_AV * kTile + (di) * kTW + (dj)]
#define TK_CR(dj, di) tk[T_CR * kTile + (di) * kTW + (dj)]
#define TK_XKC(dj, di) tk[T_XKC * kTile + (di) * kTW + (dj)]
#define TK_DUB(dj, di) tk[T_DUB * kTile + (di) * kTW + (dj)]
#define TK_DVB(dj, di) tk[T_DVB * kTile + (di) * kTW + (dj)]
#define TN_AU(dj, di) tn[T_AU * kTile + (di) * kTW + (dj)]
#define TN_AV(dj, di) tn[T_AV * kTile + (di) * kTW + (dj)]
#define TN_QDOT(dj, di) tn[T_QDOT * kTile + (di) * kTW + (dj)]
#define TN_AW(dj, di) tn[T_AW * kTile + (di) * kTW + (dj)]
#define RCM_UV_STORE_GUARD valid
#include "k_uv_level.inc"
#undef RCM_UV_STORE_GUARD
#undef UA
#undef VA
  }
}
}  // namespace

// Returns a non-null token when the staged kernel can be used on this device.
void* uv_tma_setup(const Geom&, const Fields&, long long, int) {
  const int smem = kStages * kStageBytes + 64;
  if (cudaFuncSetAttribute(k_uv_tma, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) { cudaGetLastError(); return nullptr; }
  static int token = 1;
  return &token;
}

void uv_tma_destroy(void*) {}

void launch_uv_tma(const Geom& g, const Fields& f, const StepScalars& s, cudaStream_t st, const void*) {
  const int nj = g.jdi2 - g.jdi1 + 1, ni = g.idi2 - g.idi1 + 1;
  dim3 b(32, kBY), gr((nj + 31) / 32, (ni + kBY - 1) / kBY);
  k_uv_tma<<<gr, b, kStages * kStageBytes + 64, st>>>(g, f, s);
}
#else
void* uv_tma_setup(const Geom&, const Fields&, long long, int) { return nullptr; }
void uv_tma_destroy(void*) {}
void launch_uv_tma(const Geom&, const Fields&, const StepScalars&, cudaStream_t, const void*) {}
#endif
