// Tracer transport of the sigma-coordinate cores (ichem = 1, physics off, ichebdy = 1):
//   k_tr          the complete tracer tendency of `tend`, forecast and RAW filter, one thread per (j,i) column and
//                 species, marching k in the reference's order of accumulation:
//                   decouple            Main/mod_tendency.F90:1029-1034   chi*rpsa
//                   hadv -> hadvtr      Main/mod_advection.F90:666-720
//                   vadv -> vadv4d      Main/mod_advection.F90:895-957    ind = 2, or 3 with iuwvadv = 1 (itrvadv, Main/mod_tendency.F90:149-154)
//                   nudge_chi           Main/chemlib/mod_che_bdyco.F90:898-981   (iboudy 1, 5; exponential tables of
//                                       setup_che_bdycon :1177-1217; west/east zones sum their neighbours i first)
//                   diffu_x4d           on chib3d = max(atm2%chi*rpsb, 0)  (Main/mod_slice.F90:198)
//                   forecast, filter    Main/mod_tendency.F90:548-586     (gnu2, 0.53, floor 0)
//   k_tr_repair   the negative fix of :570-582 — same sequential 3x3 sweep as for moisture (k_negq.cu), factor 0.1
//   k_tr_bdyval   chem_bdyval_coupled   Main/chemlib/mod_che_bdyco.F90:580-626,697-754
//   k_tr_timeint  chibt = (chib1 - chib0)/dtbdys   :279-290,395-403   (a division, unlike timeint)
// The same kernel serves idynamic = 1 and 2: both cores provide uavg/vavg operands (au, av, msfd), sigma-dot (qdot),
// the Smagorinsky coefficient (xkc) and p* of both time levels under the same names.
#include "dev_common.cuh"
#include "launch.h"

using namespace rcm;

namespace {

__global__ void __launch_bounds__(RCM_MAXT)
k_tr(const Geom g, const Fields f, const TrFields tr, const StepScalars s) {
  const int si = g.si;
  const long long sk = g.sk;
  const int j = g.jci1 + blockIdx.x * blockDim.x + threadIdx.x;
  const int i = g.ici1 + blockIdx.y * blockDim.y + threadIdx.y;
  const int n = blockIdx.z;
  if (j > g.jci2 || i > g.ici2) return;
  const int kz = g.kz;
  const bool up = s.upstream_mode != 0;
  const double ul = s.ul, xt = s.xt, dt = s.dt;
  const double* __restrict__ a1 = tr.a[n];
  const double* __restrict__ a2 = tr.b[n];
  const double* __restrict__ b0 = tr.b0[n];
  const double* __restrict__ bt = tr.bt[n];

  const double psa = L2(f.psa, j, i), psb = L2(f.psb, j, i);
  const double rdiv_psa = recip_for_div(psa);
  const double rpsa_c = L2(f.rpsa, j, i), rpsa_jm = L2(f.rpsa, j - 1, i), rpsa_jp = L2(f.rpsa, j + 1, i),
               rpsa_im = L2(f.rpsa, j, i - 1), rpsa_ip = L2(f.rpsa, j, i + 1);
  const double xmsf = L2(f.xmsf, j, i);
  const double m00 = L2(f.msfd, j, i), m01 = L2(f.msfd, j, i + 1), m10 = L2(f.msfd, j + 1, i), m11 = L2(f.msfd, j + 1, i + 1);
  const int ib = (s.iboudy == 1 || s.iboudy == 5) ? L2(f.zone_c, j, i) : 0;
  const bool we = ib > 0 && L2(tr.ztype, j, i) != 0;
  const bool d9 = (s.idiffu == 2);
  const bool d6on = (s.idiffu == 3) && j == g.jci2;   // the one column idiffu = 3 acts on (k_diff6)   // nine-point operator on every interior point, no boundary ring
  const bool dint = (s.idiffu == 1) && diff_interior(g, j, i, false);
  const int nring = (s.idiffu == 1) ? ring_count(g, j, i, false) : 0;
  const double beta = 0.53;
  const double thr = mintr * psa;
  const int kpb_c = f.kpb ? L2(f.kpb, j, i) : 0;   // itrvadv = 3 (iuwvadv = 1, Main/mod_tendency.F90:150-154)
#define RPSB(dj, di) L2(f.rpsb, j + (dj), i + (di))

  double fl_k = 0.0;   // vertical flux at the upper interface of level k
  // request level kk of every streamed field into L2 (see PF3)
  auto prefetch = [&](int kk) {
    PF3(f.au, j, i, kk); PF3(f.av, j, i, kk); PF3(f.au, j, i + 1, kk); PF3(f.av, j, i + 1, kk); PF3(f.qdot, j, i, kk); PF3(f.xkc, j, i, kk);
    PF3(a1, j, i, kk); PF3(a2, j, i, kk);
  };
  const int pfd = g.pfd;
  if (pfd > 0)
    for (int kk = 1; kk <= min(pfd, kz); ++kk) prefetch(kk);
  for (int k = 1; k <= kz; ++k) {
    if (pfd > 0 && k + pfd <= kz) prefetch(k + pfd);
    const double ua00 = L3(f.au, j, i, k) * m00, ua01 = L3(f.au, j, i + 1, k) * m01, ua10 = L3(f.au, j + 1, i, k) * m10,
                 ua11 = L3(f.au, j + 1, i + 1, k) * m11;
    const double va00 = L3(f.av, j, i, k) * m00, va01 = L3(f.av, j, i + 1, k) * m01, va10 = L3(f.av, j + 1, i, k) * m10,
                 va11 = L3(f.av, j + 1, i + 1, k) * m11;
    const double uavg1 = ua01 + ua00, uavg2 = ua11 + ua10;
    const double vavg1 = va10 + va00, vavg2 = va11 + va01;
    const double f1 = div_r(0.5 * ul * (uavg2 + uavg1), psa, rdiv_psa);
    const double f2 = div_r(0.5 * ul * (vavg2 + vavg1), psa, rdiv_psa);
    const double xdsk = LD1(&f.xds[k]);
    const double svv_n = (k < kz) ? L3(f.qdot, j, i, k + 1) : 0.0;
    const double ac_c = L3(a1, j, i, k);
    // hadvtr on atmx%chi
    const double q_c = ac_c * rpsa_c, q_jm = L3(a1, j - 1, i, k) * rpsa_jm, q_jp = L3(a1, j + 1, i, k) * rpsa_jp,
                 q_im = L3(a1, j, i - 1, k) * rpsa_im, q_ip = L3(a1, j, i + 1, k) * rpsa_ip;
    double ten;
    if (s.isladvec == 1) ten = 0.0 + L3(tr.sl[n], j, i, k);   // left by k_sl: slhadv_x(chidyn, atm2%chi); hdvg_x(chidyn, atm1%chi) (Main/mod_tendency.F90:1397-1399)
    else ten = 0.0 + hflux(up, f1, f2, xmsf, uavg1, uavg2, vavg1, vavg2, q_c, q_jm, q_jp, q_im, q_ip);
    // vadv4d, ind = 2
    if (k >= 2) ten = ten + fl_k * xdsk;
    double fl_n = 0.0;
    if (k < kz) {
      const double a_n = L3(a1, j, i, k + 1);
      const bool ok = (svv_n > 0.0) ? (ac_c > thr) : (a_n > thr);
      if (f.kpb) fl_n = vadv3_face(g, f, a1, j, i, k + 1, kpb_c, a_n, ac_c) * svv_n;
      else if (ok) fl_n = svv_n * (LD1(&f.twt1[k + 1]) * a_n + LD1(&f.twt2[k + 1]) * ac_c);
      ten = ten - fl_n * xdsk;
    }
    // nudge_chiten
    const double b_c = L3(a2, j, i, k);
    const double b_jm = L3(a2, j - 1, i, k), b_jp = L3(a2, j + 1, i, k), b_im = L3(a2, j, i - 1, k), b_ip = L3(a2, j, i + 1, k);
    if (ib > 0) {
      const double xf = LD1(&tr.cefc[ib + s.nspgx * (k - 1)]), xg = LD1(&tr.cegc[ib + s.nspgx * (k - 1)]);
#define FG(jj, ii, bval) ((L3(b0, jj, ii, k) + xt * L3(bt, jj, ii, k)) - (bval))
      const double fls0 = FG(j, i, b_c), fjm = FG(j - 1, i, b_jm), fjp = FG(j + 1, i, b_jp), fim = FG(j, i - 1, b_im), fip = FG(j, i + 1, b_ip);
#undef FG
      const double sum = we ? (fim + fip + fjm + fjp) : (fjm + fjp + fim + fip);
      ten = ten + xf * fls0 - xg * (sum - 4.0 * fls0);
    }
    // diffu_x4d on chib3d
    {
      double xkc = L3(f.xkc, j, i, k);
      xkc = xkc * s.rdxsq * psb;
      const double d_c = fmax(b_c * RPSB(0, 0), 0.0), d_jm = fmax(b_jm * RPSB(-1, 0), 0.0), d_jp = fmax(b_jp * RPSB(1, 0), 0.0),
                   d_im = fmax(b_im * RPSB(0, -1), 0.0), d_ip = fmax(b_ip * RPSB(0, 1), 0.0);
      if (dint) {
        const double d_jp2 = fmax(L3(a2, j + 2, i, k) * RPSB(2, 0), 0.0), d_jm2 = fmax(L3(a2, j - 2, i, k) * RPSB(-2, 0), 0.0);
        const double d_ip2 = fmax(L3(a2, j, i + 2, k) * RPSB(0, 2), 0.0), d_im2 = fmax(L3(a2, j, i - 2, k) * RPSB(0, -2), 0.0);
        ten = ten - 1.0 * xkc * (1.0 * (d_jp2 + d_jm2 + d_ip2 + d_im2) + (-4.0) * (d_jp + d_jm + d_ip + d_im) + 12.0 * d_c);
      }
      for (int r = 0; r < nring; ++r) ten = ten + 1.0 * xkc * (1.0 * (d_jp + d_jm + d_ip + d_im) + (-4.0) * d_c);
      if (d9) ten = ten + 1.0 * xkc * lap9(d_c, d_jm, d_jp, d_im, d_ip, fmax(L3(a2, j + 1, i + 1, k) * RPSB(1, 1), 0.0), fmax(L3(a2, j - 1, i - 1, k) * RPSB(-1, -1), 0.0),
                                           fmax(L3(a2, j - 1, i + 1, k) * RPSB(-1, 1), 0.0), fmax(L3(a2, j + 1, i - 1, k) * RPSB(1, -1), 0.0));
      if (d6on) ten = ten + D6(RCM_D6_CHI + n, i, k);
    }
    // chiten = chiten + chidyn + chiphy; forecast; the value filtered for a negative forecast is provisional
    // (k_tr_repair redoes it)
    const double tsum = 0.0 + ten + 0.0;
    double cq = b_c + dt * tsum;
    P3(tr.c[n], j, i, k) = cq;
    if (cq < 0.0) {
      atomicAdd(tr.negcnt, 1ULL);
      const int plane = n * kz + (k - 1);
      const int slot = atomicAdd(tr.nq_cnt + plane, 1);
      tr.nq_list[(long long)plane * sk + slot] = (j - (g.jde1 - g.hw)) + (i - (g.ide1 - g.hw)) * si;
      cq = 0.0;
    }
    const double d = s.gnu2 * (cq + b_c - 2.0 * ac_c);
    double nb = ac_c + beta * d, na = cq + (beta - 1.0) * d;
    if (nb < 0.0) nb = 0.0;
    if (na < 0.0) na = 0.0;
    P3(tr.nb[n], j, i, k) = nb;
    P3(tr.na[n], j, i, k) = na;
    fl_k = fl_n;
  }
#undef RPSB
}

#ifdef RCM_EMU
#define TQ_LOAD(p) (*(p))
#else
#define TQ_LOAD(p) (*(const volatile double*)(p))
#endif

// one block per (species, level) plane; see k_negq.cu for the scheme
__global__ void __launch_bounds__(256) k_tr_repair(const Geom g, const TrFields tr, const StepScalars s) {
  const int si = g.si;
  const long long sk = g.sk;
  const int plane = blockIdx.x, n = plane / g.kz, k = plane % g.kz + 1;
  const int cnt = tr.nq_cnt[plane];
  if (cnt == 0) return;
  double* cq = tr.c[n];
  const double* bq = tr.b[n];
  const int* list = tr.nq_list + (long long)plane * sk;
  const int jl0 = g.jde1 - g.hw, il0 = g.ide1 - g.hw;
  auto value = [&](int jj, int ii) {   // atmc%chi as the sweep sees it: atm2%chi on the physical boundary cells (:564-569)
    const bool bdy = (jj <= g.jcross1 || jj >= g.jcross2 || ii <= g.icross1 || ii >= g.icross2);
    return bdy ? P3(bq, jj, ii, k) : TQ_LOAD(&P3(cq, jj, ii, k));
  };
  auto pending = [&](int jj, int ii) {
    if (jj < g.jci1 || jj > g.jci2 || ii < g.ici1 || ii > g.ici2) return false;
    return TQ_LOAD(&P3(cq, jj, ii, k)) < 0.0;
  };
  auto attempt = [&](int e) {
    const int off = list[e];
    const int j = jl0 + off % si, i = il0 + off / si;
    if (!(TQ_LOAD(&P3(cq, j, i, k)) < 0.0)) return 0;
    if (pending(j - 1, i) || pending(j - 1, i - 1) || pending(j, i - 1) || pending(j + 1, i - 1)) return 1;
    double sum = 0.0;
    for (int ii = i - 1; ii <= i + 1; ++ii)
      for (int jj = j - 1; jj <= j + 1; ++jj) sum += fabs(value(jj, ii));
    P3(cq, j, i, k) = 0.1 * sum / 9.0;
    return 0;
  };
  auto refilter = [&](int e) {
    const int off = list[e];
    const int j = jl0 + off % si, i = il0 + off / si;
    const double beta = 0.53;
    const double c = P3(cq, j, i, k), b_c = P3(bq, j, i, k), aq = P3(tr.a[n], j, i, k);
    const double d = s.gnu2 * (c + b_c - 2.0 * aq);
    double nb = aq + beta * d, na = c + (beta - 1.0) * d;
    if (nb < 0.0) nb = 0.0;
    if (na < 0.0) na = 0.0;
    P3(tr.nb[n], j, i, k) = nb;
    P3(tr.na[n], j, i, k) = na;
  };
#ifdef RCM_EMU
  if (threadIdx.x != 0) return;
  for (;;) {
    int wait = 0;
    for (int e = 0; e < cnt; ++e) wait |= attempt(e);
    if (!wait) break;
  }
  for (int e = 0; e < cnt; ++e) refilter(e);
#else
  for (;;) {
    int wait = 0;
    for (int e = threadIdx.x; e < cnt; e += blockDim.x) wait |= attempt(e);
    __threadfence_block();
    if (!__syncthreads_or(wait)) break;
  }
  for (int e = threadIdx.x; e < cnt; e += blockDim.x) refilter(e);
#endif
}

// rim of the cross grid: atm2%chi <- atm1%chi of the previous bdyval (when integrating), atm1%chi <- chib0 + xt*chibt.
// `oa` = the level-a buffers that hold the previous rim (swapped out at the end of `tend`).
__global__ void k_tr_bdyval(const Geom g, const TrFields tr, const int2* rim, int nrim, double xt, int integrating, int use_old) {
  const int si = g.si;
  const long long sk = g.sk;
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  const int k = 1 + blockIdx.y, n = blockIdx.z;
  if (p >= nrim) return;
  const int j = rim[p].x, i = rim[p].y;
  const double* oa = use_old ? tr.na[n] : tr.a[n];
  if (integrating) P3(tr.b[n], j, i, k) = P3(oa, j, i, k);
  P3(tr.a[n], j, i, k) = P3(tr.b0[n], j, i, k) + xt * P3(tr.bt[n], j, i, k);
}

__global__ void k_tr_timeint(const Geom g, const double* b1, const double* b0, double* bt, double dtbdys, int nk, int j1, int j2,
                             int i1, int i2) {
  const int si = g.si;
  const long long sk = g.sk;
  const int j = j1 + blockIdx.x * 32 + threadIdx.x;
  const int i = i1 + blockIdx.y * 8 + threadIdx.y;
  if (j > j2 || i > i2) return;
  for (int k = 1; k <= nk; ++k) P3(bt, j, i, k) = (P3(b1, j, i, k) - P3(b0, j, i, k)) / dtbdys;
}

}  // namespace

void launch_tr(const Geom& g, const Fields& f, const TrFields& tr, const StepScalars& s, cudaStream_t st) {
  const dim3 b = rcm::col_block();
  const dim3 gr((g.jci2 - g.jci1 + b.x) / b.x, (g.ici2 - g.ici1 + b.y) / b.y, tr.ntr);
  RCM_LAUNCH(k_tr, gr, b, st, g, f, tr, s);
}
void launch_tr_repair(const Geom& g, const TrFields& tr, const StepScalars& s, cudaStream_t st) {
  dim3 b(256), gr(tr.ntr * g.kz);
  RCM_LAUNCH(k_tr_repair, gr, b, st, g, tr, s);
}
void launch_tr_bdyval(const Geom& g, const TrFields& tr, const int2* rim, int nrim, double xt, int integrating, int use_old,
                      cudaStream_t st) {
  if (nrim <= 0) return;
  dim3 b(128), gr((nrim + 127) / 128, g.kz, tr.ntr);
  RCM_LAUNCH(k_tr_bdyval, gr, b, st, g, tr, rim, nrim, xt, integrating, use_old);
}
void launch_tr_timeint(const Geom& g, const double* b1, const double* b0, double* bt, double dtbdys, int nk, int j1, int j2, int i1,
                       int i2, cudaStream_t st) {
  dim3 b(32, 8), gr((j2 - j1 + 32) / 32, (i2 - i1 + 8) / 8);
  RCM_LAUNCH(k_tr_timeint, gr, b, st, g, b1, b0, bt, dtbdys, nk, j1, j2, i1, i2);
}
