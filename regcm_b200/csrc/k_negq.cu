// Negative-moisture repair after the leapfrog forecast:
//
//   Main/mod_tendency.F90:381-393
//     call exchange(atmc%qx,1,...)
//     do n ; do k ; do i = ici1,ici2 ; do j = jci1,jci2
//       if ( atmc%qx(j,i,k,n) < 0 ) atmc%qx(j,i,k,n) = 0.01 * sum(abs(atmc%qx(j-1:j+1,i-1:i+1,k,n))) / 9
//
// The sweep is sequential: a point sees the already repaired values of the neighbours that precede it in
// loop order — (j-1,i), (j-1,i-1), (j,i-1), (j+1,i-1) — and the unrepaired values of the four that follow.
// Ghost cells hold what the neighbour rank forecast (exchanged before the sweep, never repaired locally) and
// the physical boundary cells hold atm2%qx (:374-376), exactly as in the reference.
//
// Device form.  The fused slow kernel (k_tq / k_h_tq) stores the unfiltered forecast in cqv / cqc and appends
// every negative point to the list of its (species, level) plane.  One thread block per plane then resolves
// the list in rounds: a listed point whose four predecessors are final (not negative, or not owned) is
// repaired; points are independent within a round because a point's successors cannot become ready before
// the point itself is written.  Typical inputs (isolated negatives) finish in one round; a patch needs as
// many rounds as its longest dependency chain.  Afterwards the block redoes the Asselin filter
// (Main/mod_timefilter.F90:231-278) of the repaired points.  Planes without negatives cost one counter read.
#include "dev_common.cuh"
#include "launch.h"

using namespace rcm;

namespace {

#ifdef RCM_EMU
#define NQ_LOAD(p) (*(p))
#else
#define NQ_LOAD(p) (*(const volatile double*)(p))
#endif

struct NqPlane {
  double* cq;          // forecast of this species
  const double* bq;    // level b (atm2): value of atmc%qx on the physical boundary cells
  const int* list;
  int cnt, k, jl0, il0;
};

// value of atmc%qx at (jj,ii) as the reference's sweep sees it
__device__ __forceinline__ double nq_value(const Geom& g, const NqPlane& P, const int jj, const int ii) {
  const int si = g.si;
  const long long sk = g.sk;
  const bool bdy = (jj <= g.jcross1 || jj >= g.jcross2 || ii <= g.icross1 || ii >= g.icross2);
  return bdy ? P3(P.bq, jj, ii, P.k) : NQ_LOAD(&P3(P.cq, jj, ii, P.k));
}

// a predecessor is pending while it is an owned interior point whose forecast is still negative
__device__ __forceinline__ bool nq_pending(const Geom& g, const NqPlane& P, const int jj, const int ii) {
  const int si = g.si;
  const long long sk = g.sk;
  if (jj < g.jci1 || jj > g.jci2 || ii < g.ici1 || ii > g.ici2) return false;
  return NQ_LOAD(&P3(P.cq, jj, ii, P.k)) < 0.0;
}

// returns 1 while entry e must wait for a predecessor, 0 once it is (or already was) repaired
__device__ __forceinline__ int nq_try(const Geom& g, const NqPlane& P, const int e) {
  const int si = g.si;
  const long long sk = g.sk;
  const int off = P.list[e];
  const int j = P.jl0 + off % si, i = P.il0 + off / si;
  if (!(NQ_LOAD(&P3(P.cq, j, i, P.k)) < 0.0)) return 0;
  if (nq_pending(g, P, j - 1, i) || nq_pending(g, P, j - 1, i - 1) || nq_pending(g, P, j, i - 1) || nq_pending(g, P, j + 1, i - 1))
    return 1;
  double sum = 0.0;   // sum(abs(section)): array-element order, j fastest
  for (int ii = i - 1; ii <= i + 1; ++ii)
    for (int jj = j - 1; jj <= j + 1; ++jj) sum += fabs(nq_value(g, P, jj, ii));
  P3(P.cq, j, i, P.k) = 0.01 * sum / 9.0;
  return 0;
}

// Asselin filter of one repaired point: filter_raw_qv for qv (floors minqq*p*), filter_raw_4d for qc (floor 0)
__device__ __forceinline__ void nq_filter(const Geom& g, const Fields& f, const StepScalars& s, const NqPlane& P, const int n,
                                          const int e) {
  const int si = g.si;
  const long long sk = g.sk;
  const int off = P.list[e];
  const int j = P.jl0 + off % si, i = P.il0 + off / si, k = P.k;
  const double beta = 0.53;
  const double cq = P3(P.cq, j, i, k), b_c = P3(P.bq, j, i, k);
  if (n == 0) {
    const double aq = P3(f.aqv, j, i, k);
    const double d = s.gnu1 * (cq + b_c - 2.0 * aq);
    // idynamic = 1: this runs after filter_ra_2d of p* (k_h_psfilt), as in the reference (:420-425)
    P3(f.nbqv, j, i, k) = fmax(aq + beta * d, minqq * P2(f.psa, j, i));
    P3(f.naqv, j, i, k) = fmax(cq + (beta - 1.0) * d, minqq * P2(f.psb, j, i));
  } else {
    const double aq = P3(f.aqc, j, i, k);
    const double d = s.gnu2 * (cq + b_c - 2.0 * aq);
    double nb = aq + beta * d, na = cq + (beta - 1.0) * d;
    if (nb < 0.0) nb = 0.0;
    if (na < 0.0) na = 0.0;
    P3(f.nbqc, j, i, k) = nb;
    P3(f.naqc, j, i, k) = na;
  }
}

__global__ void __launch_bounds__(256) k_negq_repair(const Geom g, const Fields f, const StepScalars s) {
  const int plane = blockIdx.x, n = plane / g.kz;
  NqPlane P;
  P.cnt = f.nq_cnt[plane];
  if (P.cnt == 0) return;
  P.k = plane % g.kz + 1;
  P.cq = n == 0 ? f.cqv : f.cqc;
  P.bq = n == 0 ? f.bqv : f.bqc;
  P.list = f.nq_list + (long long)plane * g.sk;
  P.jl0 = g.jde1 - g.hw; P.il0 = g.ide1 - g.hw;
#ifdef RCM_EMU
  // host harness: threads run one after the other, so thread 0 plays the whole block
  if (threadIdx.x != 0) return;
  for (;;) {
    int pending = 0;
    for (int e = 0; e < P.cnt; ++e) pending |= nq_try(g, P, e);
    if (!pending) break;
  }
  for (int e = 0; e < P.cnt; ++e) nq_filter(g, f, s, P, n, e);
#else
  for (;;) {
    int pending = 0;
    for (int e = threadIdx.x; e < P.cnt; e += blockDim.x) pending |= nq_try(g, P, e);
    __threadfence_block();
    if (!__syncthreads_or(pending)) break;
  }
  for (int e = threadIdx.x; e < P.cnt; e += blockDim.x) nq_filter(g, f, s, P, n, e);
#endif
}

}  // namespace

void launch_negq_repair(const Geom& g, const Fields& f, const StepScalars& s, cudaStream_t st) {
  dim3 b(256), gr(2 * g.kz);
  RCM_LAUNCH(k_negq_repair, gr, b, st, g, f, s);
}
