// Device-side helpers shared by all kernels of the dyn-core.
#pragma once
#include "consts.h"
#include "dyn_types.h"
#include "launch.h"

// origin-shifted addressing with the reference's global 1-based indices
// The element index is computed in 32-bit arithmetic: one IMAD.WIDE per access and an index shared by every field
// read at the same point, instead of a 64-bit multiply-add chain per access (which was HALF of the issue slots of the
// column kernels: 230 integer instructions for 40 memory operations per level in k_snd_wfwd).  rcm_dyn_create
// rejects geometries whose largest index would not fit in 31 bits.
// Measured per kernel on B200 (profiles/r2_measurements.md): most kernels gain (k_ppw -10 %, k_xkc -10 %, k_snd_init -8 %,
// the other acoustic kernels -1..3 %), a few lose (k_prep +10 %, k_mo_uv +15 %, k_h_tq +4 %, k_snd_wfwd +4 %): those define RCM_IDX64
// before including this header (or switch with RCM_USE_IDX64 / RCM_USE_IDX32 around one kernel).
#define P3_32(p, j, i, k) (p)[(int)(j) + (int)(i) * (int)si + (int)(k) * (int)sk]
#define P2_32(p, j, i) (p)[(int)(j) + (int)(i) * (int)si]
#define P3_64(p, j, i, k) (p)[(long long)(j) + (long long)(i) * si + (long long)(k) * sk]
#define P2_64(p, j, i) (p)[(long long)(j) + (long long)(i) * si]
#ifdef RCM_IDX64
#define P3(p, j, i, k) P3_64(p, j, i, k)
#define P2(p, j, i) P2_64(p, j, i)
#else
#define P3(p, j, i, k) P3_32(p, j, i, k)
#define P2(p, j, i) P2_32(p, j, i)
#endif
// read-only loads (ld.global.nc): the array is not written by the running
// kernel, so the compiler may hoist these above stores and pipeline them
// across the k loop
#ifdef RCM_EMU
#define L3(p, j, i, k) P3(p, j, i, k)
#define L2(p, j, i) P2(p, j, i)
#define LD1(p) (*(p))
#define PF3(p, j, i, k) ((void)0)
#else
#define L3(p, j, i, k) __ldg(&P3(p, j, i, k))
#define L2(p, j, i) __ldg(&P2(p, j, i))
#define LD1(p) __ldg(p)
// L2 prefetch of the 128-B line holding (j,i,k): the column-marching kernels
// request level k+pfd while they work on level k, so that the demand loads hit
// L2 instead of paying a DRAM round trip per dependent load phase
#define PF3(p, j, i, k) asm volatile("prefetch.global.L2 [%0];" ::"l"(&P3(p, j, i, k)))
#endif

// Column kernels run 32x4-thread blocks; RCM_MB_* = minimum resident blocks per
// SM asked of ptxas (register cap 65536 / (128 * RCM_MB)), tuned per kernel on
// B200 (scripts/build_variant.sh): the slow-tendency kernels are latency-bound
// and want occupancy, the tridiagonal forward sweep wants registers.
#define RCM_MAXT 128
#ifndef RCM_MB_PREP
#define RCM_MB_PREP 4
#endif
#ifndef RCM_MB_XKC
#define RCM_MB_XKC 8
#endif
#ifndef RCM_MB_UV
#define RCM_MB_UV 4
#endif
#ifndef RCM_MB_TQ
#define RCM_MB_TQ 6
#endif
#ifndef RCM_MB_PPW
#define RCM_MB_PPW 4
#endif
#ifndef RCM_MB_WFWD   // 2: ptxas then takes 160 registers (174 in the first sub-step) without spilling and 3 blocks still fit;
#define RCM_MB_WFWD 2   // under the 170-register cap of 3 it spilled and ran 5.6 % slower (profiles/r2_measurements.md, call N)
#endif
#ifndef RCM_MB_STAGED   // the TMA-staged slow-tendency kernels (256 threads per block)
#define RCM_MB_STAGED 2
#endif
#ifndef RCM_MB_WBACK
#define RCM_MB_WBACK 4
#endif
#include <cstdio>
#include <cstdlib>
namespace rcm {
// thread-block shape of the column-marching kernels (x along j, y along i);
// RCM_BLOCK="bx,by" overrides it for tuning runs
inline dim3 col_block() {
  static int bx = 0, by = 0;
  if (bx == 0) {
    bx = 32; by = 4;
    const char* e = std::getenv("RCM_BLOCK");
    if (e) { int a = 0, b = 0; if (std::sscanf(e, "%d,%d", &a, &b) == 2 && a > 0 && b > 0 && a * b <= RCM_MAXT) { bx = a; by = b; } }
  }
  return dim3(bx, by);
}
}  // namespace rcm

namespace rcm {

// x**c for x > 0 in the hot kernels.  The device library's pow() handles every special case of the C standard and
// costs several hundred instructions per call — a third of k_h_tq's and k_prep's issue slots (profiles/r2_measurements.md);
// every base here is a ratio of positive pressures or mixing ratios, for which exp(c*log(x)) is accurate to ~2 ulp
// (|c*log(x)| < 3 at all call sites).  The host harness keeps the host's pow so that it stays bit-comparable with
// the CPU restatement used by the tests.
__device__ __forceinline__ double pow_pos(const double x, const double c) {
#ifdef RCM_EMU
  return pow(x, c);
#else
  return exp(c * log(x));
#endif
}

// n / d when several numerators share one denominator: r = 1/d is computed once (one IEEE division) and every
// quotient costs three fused operations, q0 = n*r; rem = n - d*q0 (exact); q = q0 + rem*r — Markstein's division step,
// which returns the correctly rounded quotient when r is the correctly rounded reciprocal and q0 is within one ulp.
// Device only (RCM_NO_DIVS switches it off for A/B runs): the host harness divides, so that it stays bit-comparable with
// the CPU restatement used by the tests.
#ifndef RCM_NO_DIVS
#define RCM_DIVS 1
#endif
__device__ __forceinline__ double div_r(const double n, const double d, const double r) {
#if defined(RCM_EMU) || !defined(RCM_DIVS)
  (void)r;
  return n / d;
#else
  const double q0 = n * r;
  const double rem = fma(-d, q0, n);
  return fma(rem, r, q0);
#endif
}
__device__ __forceinline__ double recip_for_div(const double d) {
#if defined(RCM_EMU) || !defined(RCM_DIVS)
  (void)d;
  return 0.0;
#else
  return 1.0 / d;
#endif
}

// flux-form horizontal advection of a cross-point scalar
// (Main/mod_advection.F90:336-351 upstream branch, :352-357 centred branch).
// f1,f2 are the upstream weights 0.5*ul*(uavg2+uavg1)/ps and the v analogue.
__device__ __forceinline__ double hflux(const bool up, const double f1, const double f2, const double xmsf,
                                        const double ua1, const double ua2, const double va1, const double va2,
                                        const double fc, const double fjm, const double fjp, const double fim,
                                        const double fip) {
  double fx1, fx2, fy1, fy2;
  if (up) {
    fx1 = (1.0 + f1) * fjm + (1.0 - f1) * fc;
    fx2 = (1.0 + f1) * fc + (1.0 - f1) * fjp;
    fy1 = (1.0 + f2) * fim + (1.0 - f2) * fc;
    fy2 = (1.0 + f2) * fc + (1.0 - f2) * fip;
  } else {
    fx1 = fjm + fc;
    fx2 = fc + fjp;
    fy1 = fim + fc;
    fy2 = fc + fip;
  }
  return -xmsf * (ua2 * fx2 - ua1 * fx1 + va2 * fy2 - va1 * fy1);
}

// local-extrema limiter of hadvt / hadvqv (Main/mod_advection.F90:359-386,
// 566-596): i-direction first, then j-direction (order matters).
__device__ __forceinline__ double extrema_limit(double fg, const double fc, const double fjm, const double fjp,
                                                const double fim, const double fip, const double scale,
                                                const double rscale /* recip_for_div(scale) */, const double thr) {
  if (div_r(fabs(fip + fim - 2.0 * fc), scale, rscale) > thr) {
    if ((fc > fip) && (fc > fim)) fg = fmin(fg, 0.0);
    else if ((fc < fip) && (fc < fim)) fg = fmax(fg, 0.0);
  }
  if (div_r(fabs(fjp + fjm - 2.0 * fc), scale, rscale) > thr) {
    if ((fc > fjp) && (fc > fjm)) fg = fmin(fg, 0.0);
    else if ((fc < fjp) && (fc < fjm)) fg = fmax(fg, 0.0);
  }
  return fg;
}

// number of times the 2nd-order boundary-ring form of diffu_* applies to a
// point (corner points of the ring receive it twice,
// Main/mod_diffusion.F90:689-725), for cross (dot=false) or dot points.
__device__ __forceinline__ int ring_count(const Geom& g, int j, int i, bool dot) {
  const int j1 = dot ? g.jdi1 : g.jci1, j2 = dot ? g.jdi2 : g.jci2;
  const int i1 = dot ? g.idi1 : g.ici1, i2 = dot ? g.idi2 : g.ici2;
  int n = 0;
  if (g.has_bdyleft && j == j1) n++;
  if (g.has_bdyright && j == j2) n++;
  if (g.has_bdybottom && i == i1) n++;
  if (g.has_bdytop && i == i2) n++;
  return n;
}

// Fold one thread's CFL number (>= 0, so the bit pattern orders like the value)
// into the two running maxima: [0] over the sub-steps of this step, [1] over
// the whole rcm_dyn_step call.  Warp-reduced first: one same-address atomic per
// warp instead of one per column (max is order-independent => exact).  Must be
// reached by every thread of the block.
__device__ __forceinline__ void commit_cfl(const Fields& f, const double cfl, const bool valid) {
  unsigned long long bits = valid ? (unsigned long long)__double_as_longlong(cfl) : 0ULL;
  if (valid && cfl != cfl) { *f.nanflag = 1; bits = 0x7ff8000000000000ULL; }
#ifndef RCM_EMU
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const unsigned long long other = __shfl_xor_sync(0xffffffffu, bits, o);
    bits = other > bits ? other : bits;
  }
  if (((threadIdx.x + blockDim.x * threadIdx.y) & 31) != 0) return;
#endif
  if (bits == 0ULL) return;
  atomicMax(f.cflbits, bits);
  atomicMax(f.cflbits + 1, bits);
}

// Forecast of one moisture species at one point (Main/mod_tendency.F90:368-380), before the negative-value
// repair of :381-393: store it where k_negq_repair or a neighbour rank will need it, count / list negatives.
// n = 0 (qv) or 1 (qc); edge = the point lies on the tile perimeter next to a neighbour rank.
// MODE = StepScalars::cq_mode as a compile-time constant (the kernels are instantiated per mode, so that the
// common mode 0 carries no list code).
template <int MODE>
__device__ __forceinline__ void negq_note(const Geom& g, const Fields& f, const double cq, const int j,
                                          const int i, const int k, const int n, const bool edge) {
  const int si = g.si;
  const long long sk = g.sk;
  double* cqf = n == 0 ? f.cqv : f.cqc;
  if (MODE != 0 || edge) P3(cqf, j, i, k) = cq;
  if (cq < 0.0) {
    if (MODE != 1) atomicAdd(f.negq, 1ULL);
    if (MODE != 0) {
      const int plane = n * g.kz + (k - 1);
      const int slot = atomicAdd(f.nq_cnt + plane, 1);
      f.nq_list[(long long)plane * sk + slot] = (j - (g.jde1 - g.hw)) + (i - (g.ide1 - g.hw)) * si;
    }
  }
}
// Decoupled wind component `a` (au or av) at dot (j,i,k) as `decouple` leaves it for iboudy = 3, 4
// (Main/mod_tendency.F90:887-994): on outflow the outermost dot row takes the value of the row next to it — west/east
// on rows idi1..idi2 first, then south/north over the whole row, which may pick up a west/east-modified corner value.
// Identity (a * rpsda) everywhere else.  Global index tests, so it is valid on any tile.
__device__ __forceinline__ double dwind_io(const Geom& g, const Fields& f, const double* a, const int j, const int i, const int k) {
  const int si = g.si;
  const long long sk = g.sk;
  int ii = i, jj = j;
  if (i == 1 && L3(f.av, j, 1, k) >= 0.0) ii = 2;
  else if (i == g.iy && L3(f.av, j, g.iy, k) <= 0.0) ii = g.iy - 1;
  if (ii >= 2 && ii <= g.iy - 1) {
    if (j == 1 && L3(f.au, 1, ii, k) <= 0.0) jj = 2;
    else if (j == g.jx && L3(f.au, g.jx, ii, k) >= 0.0) jj = g.jx - 1;
  }
  return L3(a, jj, ii, k) * L2(f.rpsda, jj, ii);
}

__device__ __forceinline__ bool tile_edge(const Geom& g, const int j, const int i) {
  return (g.gl && j == g.jci1) || (g.gr && j == g.jci2) || (g.gb && i == g.ici1) || (g.gt && i == g.ici2);
}

// vadv4d ind = 3 (Main/mod_advection.F90:915-957; iuwvadv = 1): the value of the p*-coupled scalar `q` carried through
// interface kk (2..kz) before the multiplication by sigma-dot: the twt interpolation, replaced at the two interfaces
// around the PBL top kp = kpbl(j,i) >= 4 by an extrapolation along the limited slope of the layer above.
// q_k = q(kk), q_km1 = q(kk-1): the values the caller already holds.
__device__ __forceinline__ double vadv3_face(const Geom& g, const Fields& f, const double* __restrict__ q, int j, int i, int kk,
                                             int kp, double q_k, double q_km1) {
  const int si = g.si;
  const long long sk = g.sk;
  double fg = LD1(&f.twt1[kk]) * q_k + LD1(&f.twt2[kk]) * q_km1;
  if (kp >= 4 && (kk == kp - 1 || kk == kp)) {
    auto hs = [&](int k) { return (LD1(&f.sigma[k + 1]) + LD1(&f.sigma[k])) * 0.5; };   // hsigma (Main/mod_params.F90:2008)
    const int k = kp - 2;
    const double fp = L3(q, j, i, k + 1), fc = L3(q, j, i, k), fm = L3(q, j, i, k - 1);
    double slope;
    if ((fp - fc) > 0.0 && (fc - fm) > 0.0)
      slope = fmin((fp - fc) / (hs(k + 1) - hs(k)), (fc - fm) / (hs(k) - hs(k - 1)));
    else if ((fp - fc) < 0.0 && (fc - fm) < 0.0)
      slope = fmax((fp - fc) / (hs(k + 1) - hs(k)), (fc - fm) / (hs(k) - hs(k - 1)));
    else
      slope = 0.0;
    if (kk == kp - 1) {
      fg = fc + slope * (LD1(&f.sigma[kp - 1]) - hs(kp - 2));
    } else {
      const double fk = L3(q, j, i, kp);
      if (fabs(fc + slope * (hs(kp - 1) - hs(kp - 2)) - fk) > fabs(fp - fk)) fg = fk;
      else fg = fc + slope * (LD1(&f.sigma[kp]) - hs(kp - 2));
    }
  }
  return fg;
}

// idiffu = 3: the tendency k_diff6 left for field `fld` at (jdi2 | jci2, i, k)
#define D6(fld, i, k) LD1(&f.d6[((long long)(fld) * g.kzp1 + ((k) - 1)) * (g.iy + 2) + (i)])

// idiffu = 2: the nine-point operator o4_c1*(sides) + o4_c2*(corners) + o4_c3*centre with the reference's order of
// the sums (Main/mod_diffusion.F90:63-65,386-411,592-601,726-735,881-894); pp = (j+1,i+1), mm = (j-1,i-1),
// mp = (j-1,i+1), pm = (j+1,i-1)
__device__ __forceinline__ double lap9(double c, double jm, double jp, double im, double ip, double pp, double mm,
                                       double mp, double pm) {
  return (4.0 / 6.0) * (jp + jm + ip + im) + (1.0 / 6.0) * (pp + mm + mp + pm) + (-20.0 / 6.0) * (c);
}

__device__ __forceinline__ bool diff_interior(const Geom& g, int j, int i, bool dot) {
  if (dot) return j >= g.jdii1 && j <= g.jdii2 && i >= g.idii1 && i <= g.idii2;
  return j >= g.jcii1 && j <= g.jcii2 && i >= g.icii1 && i <= g.icii2;
}

}  // namespace rcm
