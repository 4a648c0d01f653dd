// k_slice — the slices `mkslice` prepares for the physical parametrizations (type `slice`,
// Main/mod_slice.F90:167-308, idynamic = 1, 2), evaluated ON DEMAND for one slice id from the level-b state: the
// dynamics never materialise them (consumers divide by p* on the fly), so a physics package pays for
// exactly the slices it asks for.  One thread per cross-point column, marching k.
#include "dev_common.cuh"
#include "launch.h"
#include "../../include/regcm_dyn.h"

using namespace rcm;

namespace {
struct SliceArgs {
  int id, hydro;
  double ptop;        // [cb]
  double p0, tlp, ts0;  // reference-atmosphere parameters (heights of the non-hydrostatic core)
  double dx8;           // 8 dx (omega of the hydrostatic core)
};

// Share/pfesat.inc, Share/pfwsat.inc
__device__ __forceinline__ double sl_pfwsat(const double t, const double p) {
  double td = t - tzero;
  if (td > 100.0) td = 100.0;
  if (td < -75.0) td = -75.0;
  double esat;
  if (td >= 0.0)
    esat = 6.11213476 + td * (0.444007856 + td * (0.143064234e-01 + td * (0.264461437e-03 + td * (0.305903558e-05 + td * (0.196237241e-07 + td * (0.892344772e-10 + td * (-0.373208410e-12 + td * 0.209339997e-15)))))));
  else
    esat = 6.11123516 + td * (0.503109514 + td * (0.188369801e-01 + td * (0.420547422e-03 + td * (0.614396778e-05 + td * (0.602780717e-07 + td * (0.387940929e-09 + td * (0.149436277e-11 + td * 0.262655803e-14)))))));
  const double es = esat * 100.0;
  return ep2 * (es / (p - es));
}

__global__ void __launch_bounds__(128) k_slice(const Geom g, const Fields f, const SliceArgs a, double* __restrict__ out) {
  const int si = g.si;
  const long long sk = g.sk;
  const int j = g.jce1 + blockIdx.x * blockDim.x + threadIdx.x;
  const int i = g.ice1 + blockIdx.y * blockDim.y + threadIdx.y;
  if (j > g.jce2 || i > g.ice2) return;
  const int kz = g.kz, kzp1 = g.kzp1, id = a.id;
  const bool hydro = a.hydro != 0;
  const bool interior = (j >= g.jci1 && j <= g.jci2 && i >= g.ici1 && i <= g.ici2);
  const double psb = L2(f.psb, j, i), rpsb = 1.0 / psb;
  const double ptoppa = a.ptop * 1000.0;
  auto tb = [&](int k) { return L3(f.bt, j, i, k) * rpsb; };
  auto qv = [&](int k) { return fmax(L3(f.bqv, j, i, k) * rpsb, minqq); };
  auto qc = [&](int k) { return fmax(L3(f.bqc, j, i, k) * rpsb, 0.0); };
  auto ppb = [&](int k) { return L3(f.bpp, j, i, k) * rpsb; };
  auto hsig = [&](int k) { return (LD1(&f.sigma[k + 1]) + LD1(&f.sigma[k])) * 0.5; };
  auto pb = [&](int k) {
    if (hydro) return (hsig(k) * psb + a.ptop) * 1000.0;
    if (k == 1) return fmax(L3(f.pr0, j, i, 1) + ppb(1), ptoppa + 1.0);
    return L3(f.pr0, j, i, k) + ppb(k);
  };
  const double ps0 = hydro ? 0.0 : L2(f.ps0, j, i);
  const double ps2 = hydro ? (psb + a.ptop) * 1000.0 : ps0 + ptoppa + ppb(kz);
  auto pf = [&](int k) {
    if (hydro) return (LD1(&f.sigma[k]) * psb + a.ptop) * 1000.0;
    if (k == 1) return ptoppa;
    if (k == kzp1) return ps2;
    return (ps0 * LD1(&f.sigma[k]) + ptoppa) + 0.5 * (ppb(k - 1) + ppb(k));   // atm0%pf = ps0*sigma + ptop (nhbase)
  };
  auto tv = [&](int k) { return tb(k) * (1.0 + ep1 * qv(k) - qc(k)); };
  auto rhob = [&](int k) { return pb(k) / (rgas * tb(k)); };
  auto wpx = [&](int k) {   // omega (Main/mod_tendency.F90:1199-1223) [Pa/s]
    if (hydro) {   // omega [cb/s] as k_h_tq evaluates it (Main/mod_tendency.F90:1199-1215), here before the tendencies
      const double psa = L2(f.psa, j, i);
      const double q_k = (k >= 2) ? L3(f.qdot, j, i, k) : 0.0, q_n = (k < kz) ? L3(f.qdot, j, i, k + 1) : 0.0;
      const double dummy8 = 1.0 / (a.dx8 * L2(f.msfx, j, i));
      const double dpsx = L2(f.psa, j + 1, i) - L2(f.psa, j - 1, i), dpsy = L2(f.psa, j, i + 1) - L2(f.psa, j, i - 1);
      const double r00 = L2(f.rpsda, j, i), r01 = L2(f.rpsda, j, i + 1), r10 = L2(f.rpsda, j + 1, i), r11 = L2(f.rpsda, j + 1, i + 1);
      const double us = L3(f.au, j, i, k) * r00 + L3(f.au, j, i + 1, k) * r01 + L3(f.au, j + 1, i + 1, k) * r11 + L3(f.au, j + 1, i, k) * r10;
      const double vs = L3(f.av, j, i, k) * r00 + L3(f.av, j, i + 1, k) * r01 + L3(f.av, j + 1, i + 1, k) * r11 + L3(f.av, j + 1, i, k) * r10;
      const double om = 0.5 * (q_n + q_k) * psa + hsig(k) * (L2(f.pten, j, i) + (us * dpsx + vs * dpsy) * dummy8);
      return om * 1000.0;
    }
    return -0.5 * egrav * L3(f.rho0, j, i, k) * rpsb * (L3(f.bw, j, i, k) + L3(f.bw, j, i, k + 1));
  };
  auto z0 = [&](double sig) {   // nhbase height of a reference level (Share/mod_nhinterp.F90:100-104)
    const double p = ps0 * sig + ptoppa;
    const double alnp = log(p / (ps0 + ptoppa));
    return fmax(-(0.5 * rovg * a.tlp * alnp * alnp + rovg * a.ts0 * alnp), 0.0);
  };
  int nk = kz;
  if (id == RCM_SL_PF3D || id == RCM_SL_WB3D || id == RCM_SL_ZQ) nk = kzp1;
  if (id == RCM_SL_PS2D) nk = 1;
  if (hydro && (id == RCM_SL_ZQ || id == RCM_SL_ZA || id == RCM_SL_DZQ)) {   // Main/mod_slice.F90:283-298, bottom-up
    const double cell = a.ptop * rpsb;
    double zq_n = 0.0;   // zq(k+1)
    if (id == RCM_SL_ZQ) P3(out, j, i, kzp1) = 0.0;
    for (int k = kz; k >= 1; --k) {
      const double zq_k = zq_n + rovg * tv(k) * log((LD1(&f.sigma[k + 1]) + cell) / (LD1(&f.sigma[k]) + cell));
      P3(out, j, i, k) = id == RCM_SL_ZQ ? zq_k : id == RCM_SL_ZA ? 0.5 * (zq_k + zq_n) : zq_k - zq_n;
      zq_n = zq_k;
    }
    return;
  }
  for (int k = 1; k <= nk; ++k) {
    double v = 0.0;
    switch (id) {
      case RCM_SL_TB3D: v = tb(k); break;
      case RCM_SL_QVB3D: v = qv(k); break;
      case RCM_SL_QCB3D: v = qc(k); break;
      case RCM_SL_UBX3D:
      case RCM_SL_VBX3D: {
        const double* u = id == RCM_SL_UBX3D ? f.bu : f.bv;
        v = 0.25 * (L3(u, j, i, k) * (1.0 / L2(f.psdotb, j, i)) + L3(u, j, i + 1, k) * (1.0 / L2(f.psdotb, j, i + 1)) +
                    L3(u, j + 1, i, k) * (1.0 / L2(f.psdotb, j + 1, i)) + L3(u, j + 1, i + 1, k) * (1.0 / L2(f.psdotb, j + 1, i + 1)));
        break;
      }
      case RCM_SL_PB3D: v = pb(k); break;
      case RCM_SL_PF3D: v = pf(k); break;
      case RCM_SL_PS2D: v = ps2; break;
      case RCM_SL_RHOB3D: v = rhob(k); break;
      case RCM_SL_TH3D: v = tb(k) * pow(p00 / pb(k), rovcp); break;
      case RCM_SL_TV3D: v = tv(k); break;
      case RCM_SL_TP3D: v = tb(k) * pow(ps2 / pb(k), rovcp); break;
      case RCM_SL_WPX3D: v = interior ? wpx(k) : 0.0; break;
      case RCM_SL_WB3D:
        if (!hydro) v = L3(f.bw, j, i, k) * rpsb;
        else if (interior && k >= 2 && k <= kz) v = -0.5 * regrav * (wpx(k - 1) / rhob(k - 1) + wpx(k) / rhob(k));
        break;
      case RCM_SL_ZQ: v = z0(LD1(&f.sigma[k])); break;
      case RCM_SL_ZA: v = z0(hsig(k)); break;
      case RCM_SL_DZQ: v = z0(LD1(&f.sigma[k])) - z0(LD1(&f.sigma[k + 1])); break;
      case RCM_SL_QSB3D: v = sl_pfwsat(tb(k), pb(k)); break;
      case RCM_SL_RHB3D: v = fmin(fmax(qv(k) / sl_pfwsat(tb(k), pb(k)), 0.01), 1.01); break;   // rhmin, rhmax: Main/mod_params.F90:331-332
    }
    P3(out, j, i, k) = v;
  }
}
}  // namespace

void launch_slice(const Geom& g, const Fields& f, int id, int hydro, double ptop, double p0, double tlp, double ts0, double dx8,
                  double* out, cudaStream_t st) {
  SliceArgs a{id, hydro, ptop, p0, tlp, ts0, dx8};
  const int nj = g.jce2 - g.jce1 + 1, ni = g.ice2 - g.ice1 + 1;
  dim3 b(32, 4), gr((nj + 31) / 32, (ni + 3) / 4);
  RCM_LAUNCH(k_slice, gr, b, st, g, f, a, out);
}
