// idiffu = 3: the flux-limited 6th-order horizontal diffusion of Main/mod_diffusion.F90:412-517 (winds),
// 602-652 (w, TKE: diffu_x3df), 736-790 (T, p': diffu_x3d), 895-960 (qx, tracers: diffu_x4d3d) AS THIS SNAPSHOT HAS IT: every
// loop over j runs over the single column j = jdi2 (dot) / jci2 (cross) (:421,611,745,902), the stencil indices are clamped
// to the grid (:414-426), the coefficients are xkc = xkcf = diff_6th_coef * p*, xkd = diff_6th_coef * p*dot (:154,174-183).
//   k_diff6   one thread per (i, level, field): the tendency fac * xk * ((x1 - x0) + (y1 - y0)) of that one column into
//             Fields::d6; k_uv / k_tq / k_ppw / k_tr / k_tke / k_h_* add it where their diffusion term stands (D6).
// One rank: jdi2 / jci2 are per-tile bounds in the reference, so under MPI its operator acts on the last column of every
// tile and the answer depends on the decomposition; rcm_dyn_create refuses that.  diffu_x3df multiplies by xkc(j,i,k) for k up to kzp1, one level beyond the
// array (:646); xkc does not depend on k, so level kzp1 takes the value of the other levels.
#include "dev_common.cuh"
#include "launch.h"

using namespace rcm;

namespace {

__global__ void k_diff6(const Geom g, const Fields f, const TrFields tr, const UwFields uw, const StepScalars s, const int nfld,
                        const int has_tke, const double coef) {
  const int si = g.si;
  const long long sk = g.sk;
  const int i = 1 + blockIdx.x * blockDim.x + threadIdx.x;
  const int k = 1 + blockIdx.y;
  const int fld = blockIdx.z;
  if (i > g.iy || fld >= nfld) return;
  const bool dot = (fld == RCM_D6_U || fld == RCM_D6_V);
  const bool full = (fld == RCM_D6_W || fld == RCM_D6_TKE);           // kz + 1 levels
  if (k > (full ? g.kzp1 : g.kz)) return;
  if (fld == RCM_D6_TKE && !has_tke) return;
  if ((fld == RCM_D6_PP || fld == RCM_D6_W) && s.hydro) return;
  const int j = dot ? g.jdi2 : g.jci2;
  const int i1 = dot ? g.idi1 : g.ici1, i2 = dot ? g.idi2 : g.ici2;
  double out = 0.0;
  if (i >= i1 && i <= i2) {
    const int jmax = dot ? g.jx : g.jx - 1, imax = dot ? g.iy : g.iy - 1;
    const double* src = fld == RCM_D6_U ? f.dub : fld == RCM_D6_V ? f.dvb : fld == RCM_D6_T ? f.bt : fld == RCM_D6_QV ? f.bqv
                      : fld == RCM_D6_QC ? f.bqc : fld == RCM_D6_PP ? f.bpp : fld == RCM_D6_W ? f.bw : fld == RCM_D6_TKE ? uw.b
                      : tr.b[fld - RCM_D6_CHI];
    // the diffused quantity: the decoupled level-b field (Main/mod_slice.F90:169-200); the winds come divided by the map factor
    auto F = [&](int jj, int ii) {
      const double v = L3(src, jj, ii, k);
      if (dot || fld == RCM_D6_TKE) return v;
      const double d = v * L2(f.rpsb, jj, ii);
      if (fld == RCM_D6_QV) return fmax(d, minqq);
      if (fld == RCM_D6_QC || fld >= RCM_D6_CHI) return fmax(d, 0.0);
      return d;
    };
    // the quantity whose difference signs the flux: F for the winds, F / mpd (mpd = msfd at the same indices) for the scalars
    auto G = [&](int jj, int ii) { return dot ? F(jj, ii) : F(jj, ii) / L2(f.msfd, jj, ii); };
    const int im1 = max(i - 1, 1), im2 = max(i - 2, 1), im3 = max(i - 3, 1);
    const int ip1 = min(i + 1, imax), ip2 = min(i + 2, imax), ip3 = min(i + 3, imax);
    const int jm1 = max(j - 1, 1), jm2 = max(j - 2, 1), jm3 = max(j - 3, 1);
    const int jp1 = min(j + 1, jmax), jp2 = min(j + 2, jmax), jp3 = min(j + 3, jmax);
    double x0 = (10.0 * (F(j, i) - F(jm1, i)) + (-5.0) * (F(jp1, i) - F(jm2, i)) + 1.0 * (F(jp2, i) - F(jm3, i)));
    if (x0 * (G(j, i) - G(jm1, i)) <= 0.0) x0 = 0.0;
    double x1 = (10.0 * (F(jp1, i) - F(j, i)) + (-5.0) * (F(jp2, i) - F(jm1, i)) + 1.0 * (F(jp3, i) - F(jm2, i)));
    if (x1 * (G(jp1, i) - G(j, i)) <= 0.0) x1 = 0.0;
    double y0 = (10.0 * (F(j, i) - F(j, im1)) + (-5.0) * (F(j, ip1) - F(j, im2)) + 1.0 * (F(j, ip2) - F(j, im3)));
    if (y0 * (G(j, i) - G(j, im1)) <= 0.0) y0 = 0.0;
    double y1 = (10.0 * (F(j, ip1) - F(j, i)) + (-5.0) * (F(j, ip2) - F(j, im1)) + 1.0 * (F(j, ip3) - F(j, im2)));
    if (y1 * (G(j, ip1) - G(j, i)) <= 0.0) y1 = 0.0;
    const double xk = coef * (dot ? L2(f.psdotb, j, i) : L2(f.psb, j, i));
    const double fac = (fld == RCM_D6_TKE) ? uw.nuk : 1.0;
    out = (fac * xk) * ((x1 - x0) + (y1 - y0));   // fac = 1 is exact: xkc * (...) of diffu_x3d, fac * xkc * (...) of the others
  }
  f.d6[((long long)fld * g.kzp1 + (k - 1)) * (g.iy + 2) + i] = out;
}

}  // namespace

void launch_diff6(const Geom& g, const Fields& f, const TrFields& tr, const UwFields& uw, const StepScalars& s, int has_tke, double coef,
                  cudaStream_t st) {
  const int nfld = RCM_D6_CHI + tr.ntr;
  dim3 b(64), gr((g.iy + 63) / 64, g.kzp1, nfld);
  RCM_LAUNCH(k_diff6, gr, b, st, g, f, tr, uw, s, nfld, has_tke, coef);
}
