// Split-explicit gravity-wave correction of the hydrostatic core:
//   k_sp_init   spinit, storage part      Main/mod_split.F90:192-235
//   k_sp_proj   splitf, projections       Main/mod_split.F90:254-409 (+ the head of spstep :475-492)
//   k_sp_step   spstep, one sub-step      Main/mod_split.F90:494-663
//   k_sp_corr_c / k_sp_corr_d  corrections to p*, T / u, v   Main/mod_split.F90:417-457
// Modal fields are 2-D planes: deld/delh hold plane (slot-1)*nsplit + l for the
// three time slots of the reference's deld(:,:,l,slot); ddsum/dhsum/dstor/hstor
// hold plane l.  The reference's gradient -> divergence chain of spstep
// (work1/work2 -> uu/vv -> work3, two halo exchanges) is evaluated per cross
// point from the 3x3 neighbourhood of delh, so one launch advances one sub-step
// of every mode that is still active.
#include "dev_common.cuh"
#include "launch.h"

using namespace rcm;

#define SP(p, j, i, l) P3(p, j, i, l)                                   /* plane l = 1..nsplit */
#define SL(p, j, i, l, slot) P3(p, j, i, ((slot) - 1) * ns + (l))       /* slot = 1..3 */

// divergence of (p*u, p*v)*msfd at a cross point, reference order (:204-206,289-291)
__device__ __forceinline__ double div_uv(const Geom& g, const Fields& f, const double* u, const double* v, int j, int i, int k,
                                         double m00, double m01, double m10, double m11) {
  const int si = g.si;
  const long long sk = g.sk;
  const double u00 = P3(u, j, i, k) * m00, u01 = P3(u, j, i + 1, k) * m01, u10 = P3(u, j + 1, i, k) * m10, u11 = P3(u, j + 1, i + 1, k) * m11;
  const double v00 = P3(v, j, i, k) * m00, v01 = P3(v, j, i + 1, k) * m01, v10 = P3(v, j + 1, i, k) * m10, v11 = P3(v, j + 1, i + 1, k) * m11;
  return (-u01 + u11 - u00 + u10 + v01 + v11 - v00 - v10);
}

// modal geopotential of one time level (:343-361): t = coupled temperature, ps = p*
__device__ __forceinline__ double modal_h(const Geom& g, const Fields& f, const StepScalars& s, const double* t, double ps, int j,
                                          int i, int l) {
  const int si = g.si;
  const long long sk = g.sk;
  const int ns = s.nsplit, kz = g.kz;
  const double dp = ps - s.sp_pd;
  double h = LD1(&f.sp_pdlog[l + ns * kz]) + LD1(&f.sp_eps1[l + ns * kz]) * dp;  // k = kz+1 term
  for (int k = 1; k <= kz; ++k) {
    const double eps = LD1(&f.sp_eps1[l + ns * (k - 1)]) * dp;
    h = h + LD1(&f.sp_pdlog[l + ns * (k - 1)]) + LD1(&f.sp_tau[l + ns * (k - 1)]) * P3(t, j, i, k) / ps + eps;
  }
  return h;
}

// spinit: dstor, hstor from the initial level-b state
__global__ void k_sp_init(const Geom g, const Fields f, const StepScalars s) {
  const int si = g.si;
  const long long sk = g.sk;
  const int j = g.jce1 + blockIdx.x * blockDim.x + threadIdx.x;
  const int i = g.ice1 + blockIdx.y * blockDim.y + threadIdx.y;
  if (j > g.jce2 || i > g.ice2) return;
  const int ns = s.nsplit, kz = g.kz;
  const double rdx2 = 1.0 / s.dx2;
  const double msfx = P2(f.msfx, j, i);
  const double map = 1.0 / (msfx * msfx);
  P2(f.map, j, i) = map;
  const double m00 = P2(f.msfd, j, i), m01 = P2(f.msfd, j, i + 1), m10 = P2(f.msfd, j + 1, i), m11 = P2(f.msfd, j + 1, i + 1);
  const double psb = P2(f.psb, j, i);
  for (int l = 1; l <= ns; ++l) {
    double d = 0.0;  // dstor is zero-initialised by the allocator
    for (int k = 1; k <= kz; ++k)
      d = d + LD1(&f.sp_zr[l + ns * (k - 1)]) * map * rdx2 * div_uv(g, f, f.bu, f.bv, j, i, k, m00, m01, m10, m11);
    SP(f.dstor, j, i, l) = d;
    SP(f.hstor, j, i, l) = modal_h(g, f, s, f.bt, psb, j, i, l);
  }
}

// splitf: deld/delh slots from the new time levels (the "next" buffers), storage update
__global__ void k_sp_proj(const Geom g, const Fields f, const StepScalars s) {
  const int si = g.si;
  const long long sk = g.sk;
  const int j = g.jce1 + blockIdx.x * blockDim.x + threadIdx.x;
  const int i = g.ice1 + blockIdx.y * blockDim.y + threadIdx.y;
  if (j > g.jce2 || i > g.ice2) return;
  const int ns = s.nsplit, kz = g.kz;
  const double rdx2 = 1.0 / s.dx2;
  const double map = P2(f.map, j, i);
  const double m00 = P2(f.msfd, j, i), m01 = P2(f.msfd, j, i + 1), m10 = P2(f.msfd, j + 1, i), m11 = P2(f.msfd, j + 1, i + 1);
  const double psa = P2(f.psa, j, i), psb = P2(f.psb, j, i);
  double d3[RCM_MAXSPLIT], d2[RCM_MAXSPLIT];
  for (int l = 0; l < ns; ++l) { d3[l] = 0.0; d2[l] = 0.0; }
  for (int k = 1; k <= kz; ++k) {
    const double diva = div_uv(g, f, f.nau, f.nav, j, i, k, m00, m01, m10, m11);
    const double divb = div_uv(g, f, f.nbu, f.nbv, j, i, k, m00, m01, m10, m11);
    for (int l = 1; l <= ns; ++l) {
      const double zr = LD1(&f.sp_zr[l + ns * (k - 1)]);
      d3[l - 1] = d3[l - 1] + zr * rdx2 * map * diva;
      d2[l - 1] = d2[l - 1] + zr * rdx2 * map * divb;
    }
  }
  for (int l = 1; l <= ns; ++l) {
    const double dst = SP(f.dstor, j, i, l), hst = SP(f.hstor, j, i, l);
    const double h3 = modal_h(g, f, s, f.nat, psa, j, i, l);
    const double h2 = modal_h(g, f, s, f.nbt, psb, j, i, l);
    const double dd3 = d3[l - 1] - dst, dd1 = dst - d2[l - 1];
    const double hh3 = h3 - hst, hh1 = hst - h2;
    SL(f.deld, j, i, l, 3) = dd3; SL(f.deld, j, i, l, 2) = d2[l - 1]; SL(f.deld, j, i, l, 1) = dd1;
    SL(f.delh, j, i, l, 3) = hh3; SL(f.delh, j, i, l, 2) = h2; SL(f.delh, j, i, l, 1) = hh1;
    SP(f.dstor, j, i, l) = d2[l - 1];
    SP(f.hstor, j, i, l) = h2;
    // head of spstep: the sums start from slot n0 = 1
    SP(f.ddsum, j, i, l) = dd1;
    SP(f.dhsum, j, i, l) = hh1;
  }
}

// one sub-step n (1-based) of spstep for every mode with n <= m2(l)
__global__ void k_sp_step(const Geom g, const Fields f, const StepScalars s, const int n) {
  const int si = g.si;
  const long long sk = g.sk;
  const int j = g.jce1 + blockIdx.x * blockDim.x + threadIdx.x;
  const int i = g.ice1 + blockIdx.y * blockDim.y + threadIdx.y;
  if (j > g.jce2 || i > g.ice2) return;
  const int ns = s.nsplit;
  const double rdx2 = 1.0 / s.dx2;
  const bool interior = (j >= g.jci1 && j <= g.jci2 && i >= g.ici1 && i <= g.ici2);
  const bool first = (n == 1);
  // time slots: n = 1: n0 = 1, n1 = 2 (forward, writes n1); n >= 2: writes n2 = n0
  const int n0 = first ? 1 : ((n & 1) ? 2 : 1);
  const int n1 = first ? 2 : ((n & 1) ? 1 : 2);
  const int src = first ? n0 : n1;   // slot whose gradient is taken
  const int dst = first ? n1 : n0;
  for (int l = 1; l <= ns; ++l) {
    if (n > s.sp_m2[l - 1]) continue;
    const double aam = s.sp_aam[l - 1];
    double dnew, hnew;
    if (interior) {
      // gradient of delh on the four dots around the cross point, then divergence
      double uu[2][2], vv[2][2];
#pragma unroll
      for (int a = 0; a < 2; ++a)
#pragma unroll
        for (int b = 0; b < 2; ++b) {
          const int jd = j + a, id = i + b;
          const double x11 = SL(f.delh, jd, id, l, src), x10 = SL(f.delh, jd, id - 1, l, src), x01 = SL(f.delh, jd - 1, id, l, src),
                       x00 = SL(f.delh, jd - 1, id - 1, l, src);
          const double fac = s.dx2 * P2(f.msfx, jd, id);  // sic: the cross map factor at the dot index (:502)
          double w1 = (x11 + x10 - x01 - x00) / fac;
          double w2 = (x11 + x01 - x10 - x00) / fac;
          const double pd = P2(f.psdota, jd, id), md = P2(f.msfd, jd, id);
          w1 = w1 * pd; w2 = w2 * pd;
          uu[a][b] = w1 * md; vv[a][b] = w2 * md;
        }
      const double w3 = rdx2 * P2(f.map, j, i) * (-uu[0][1] + uu[1][1] - uu[0][0] + uu[1][0] + vv[0][1] + vv[1][1] - vv[0][0] - vv[1][0]);
      const double psa = P2(f.psa, j, i);
      if (first) {
        const double m2 = (double)s.sp_m2[l - 1];
        dnew = SL(f.deld, j, i, l, n0) - s.sp_dtau[l - 1] * w3 + SL(f.deld, j, i, l, 3) / m2;
        hnew = SL(f.delh, j, i, l, n0) - s.sp_dtau[l - 1] * s.sp_hbar[l - 1] * SL(f.deld, j, i, l, n0) / psa + SL(f.delh, j, i, l, 3) / m2;
      } else {
        const double dtau2 = s.sp_dtau[l - 1] * 2.0;
        dnew = SL(f.deld, j, i, l, n0) - dtau2 * w3 + SL(f.deld, j, i, l, 3) / aam;
        hnew = SL(f.delh, j, i, l, n0) - dtau2 * s.sp_hbar[l - 1] * SL(f.deld, j, i, l, n1) / psa + SL(f.delh, j, i, l, 3) / aam;
      }
      SL(f.deld, j, i, l, dst) = dnew;
    } else {
      // boundary rows ("not in Madala (1987)", :547-567,632-651); deld keeps the slot's value
      if (first) hnew = SL(f.delh, j, i, l, n0) * ((aam - 1.0) / aam);
      else hnew = 2.0 * SL(f.delh, j, i, l, n1) - SL(f.delh, j, i, l, n0);
      dnew = SL(f.deld, j, i, l, dst);
    }
    SL(f.delh, j, i, l, dst) = hnew;
    SP(f.ddsum, j, i, l) = SP(f.ddsum, j, i, l) + dnew;
    SP(f.dhsum, j, i, l) = SP(f.dhsum, j, i, l) + hnew;
  }
}

// corrections to p* and T on interior cross points (:417-436)
__global__ void k_sp_corr_c(const Geom g, const Fields f, const StepScalars s) {
  const int si = g.si;
  const long long sk = g.sk;
  const int j = g.jci1 + blockIdx.x * blockDim.x + threadIdx.x;
  const int i = g.ici1 + blockIdx.y * blockDim.y + threadIdx.y;
  if (j > g.jci2 || i > g.ici2) return;
  const int ns = s.nsplit, kz = g.kz;
  double dd[RCM_MAXSPLIT];
  double psa = P2(f.psa, j, i), psb = P2(f.psb, j, i);
  for (int l = 1; l <= ns; ++l) {
    dd[l - 1] = SP(f.ddsum, j, i, l);
    const double an = s.sp_an[l - 1], gnuan = s.gnu1 * an;
    psa = psa - an * dd[l - 1];
    psb = psb - gnuan * dd[l - 1];
  }
  P2(f.psa, j, i) = psa; P2(f.psb, j, i) = psb;
  for (int k = 1; k <= kz; ++k) {
    double ta = P3(f.nat, j, i, k), tb = P3(f.nbt, j, i, k);
    for (int l = 1; l <= ns; ++l) {
      const double am = LD1(&f.sp_am[l + ns * (k - 1)]), gnuam = s.gnu1 * am;
      ta = ta + am * dd[l - 1];
      tb = tb + gnuam * dd[l - 1];
    }
    P3(f.nat, j, i, k) = ta; P3(f.nbt, j, i, k) = tb;
  }
}

// corrections to p*u, p*v on interior dot points (:440-457)
__global__ void k_sp_corr_d(const Geom g, const Fields f, const StepScalars s) {
  const int si = g.si;
  const long long sk = g.sk;
  const int j = g.jdi1 + blockIdx.x * blockDim.x + threadIdx.x;
  const int i = g.idi1 + blockIdx.y * blockDim.y + threadIdx.y;
  if (j > g.jdi2 || i > g.idi2) return;
  const int ns = s.nsplit, kz = g.kz;
  const double fac = P2(f.psdota, j, i) / (s.dx2 * P2(f.msfd, j, i));
  double x[RCM_MAXSPLIT], y[RCM_MAXSPLIT];
  for (int l = 1; l <= ns; ++l) {
    const double h11 = SP(f.dhsum, j, i, l), h10 = SP(f.dhsum, j, i - 1, l), h01 = SP(f.dhsum, j - 1, i, l), h00 = SP(f.dhsum, j - 1, i - 1, l);
    x[l - 1] = fac * (h11 + h10 - h01 - h00);
    y[l - 1] = fac * (h11 - h10 + h01 - h00);
  }
  for (int k = 1; k <= kz; ++k) {
    double ua = P3(f.nau, j, i, k), va = P3(f.nav, j, i, k), ub = P3(f.nbu, j, i, k), vb = P3(f.nbv, j, i, k);
    for (int l = 1; l <= ns; ++l) {
      const double zm = LD1(&f.sp_zm[l + ns * (k - 1)]), gnuzm = s.gnu1 * zm;
      ua = ua - zm * x[l - 1];
      va = va - zm * y[l - 1];
      ub = ub - gnuzm * x[l - 1];
      vb = vb - gnuzm * y[l - 1];
    }
    P3(f.nau, j, i, k) = ua; P3(f.nav, j, i, k) = va; P3(f.nbu, j, i, k) = ub; P3(f.nbv, j, i, k) = vb;
  }
}

static dim3 grid2(int nj, int ni, dim3 b) { return dim3((nj + b.x - 1) / b.x, (ni + b.y - 1) / b.y); }

void launch_sp_init(const Geom& g, const Fields& f, const StepScalars& s, cudaStream_t st) {
  const dim3 b = rcm::col_block();
  RCM_LAUNCH(k_sp_init, grid2(g.jce2 - g.jce1 + 1, g.ice2 - g.ice1 + 1, b), b, st, g, f, s);
}
void launch_sp_proj(const Geom& g, const Fields& f, const StepScalars& s, cudaStream_t st) {
  const dim3 b = rcm::col_block();
  RCM_LAUNCH(k_sp_proj, grid2(g.jce2 - g.jce1 + 1, g.ice2 - g.ice1 + 1, b), b, st, g, f, s);
}
void launch_sp_step(const Geom& g, const Fields& f, const StepScalars& s, int n, cudaStream_t st) {
  dim3 b(32, 8);
  RCM_LAUNCH(k_sp_step, grid2(g.jce2 - g.jce1 + 1, g.ice2 - g.ice1 + 1, b), b, st, g, f, s, n);
}
void launch_sp_corr(const Geom& g, const Fields& f, const StepScalars& s, cudaStream_t st) {
  const dim3 b = rcm::col_block();
  RCM_LAUNCH(k_sp_corr_c, grid2(g.jci2 - g.jci1 + 1, g.ici2 - g.ici1 + 1, b), b, st, g, f, s);
  RCM_LAUNCH(k_sp_corr_d, grid2(g.jdi2 - g.jdi1 + 1, g.idi2 - g.idi1 + 1, b), b, st, g, f, s);
}
