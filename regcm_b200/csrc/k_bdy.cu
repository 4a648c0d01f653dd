// Lateral-boundary bookkeeping kernels:
//   k_couple_pair / k_timeint   couple + timeint       Main/mod_bdycod.F90:4938-4951,5100-5113
//   k_bdyval_dot / k_bdyval_crs bdyval + bdyuv         Main/mod_bdycod.F90:1109-2571,896-1094
// Only rim points are touched: the host builds, once, the list of (j,i) rim
// points of the dot and of the cross grid owned by this rank.
#include "dev_common.cuh"
#include "launch.h"

using namespace rcm;

__global__ void k_scale2d(const Geom g, double* a, const double* c, int nk, int j1, int j2, int i1, int i2) {
  const int si = g.si;
  const long long sk = g.sk;
  const int j = j1 + blockIdx.x * 32 + threadIdx.x;
  const int i = i1 + blockIdx.y * 8 + threadIdx.y;
  if (j > j2 || i > i2) return;
  const double cc = P2(c, j, i);
  for (int k = 1; k <= nk; ++k) P3(a, j, i, k) = P3(a, j, i, k) * cc;
}

__global__ void k_timeint(const Geom g, const double* b1, const double* b0, double* bt, double rdtbdy, int nk, int j1,
                          int j2, int i1, int i2) {
  const int si = g.si;
  const long long sk = g.sk;
  const int j = j1 + blockIdx.x * 32 + threadIdx.x;
  const int i = i1 + blockIdx.y * 8 + threadIdx.y;
  if (j > j2 || i > i2) return;
  for (int k = 1; k <= nk; ++k) P3(bt, j, i, k) = (P3(b1, j, i, k) - P3(b0, j, i, k)) * rdtbdy;
}

void launch_couple(const Geom& g, double* a, const double* c, int nk, int j1, int j2, int i1, int i2, cudaStream_t st) {
  dim3 b(32, 8), gr((j2 - j1 + 32) / 32, (i2 - i1 + 8) / 8);
  RCM_LAUNCH(k_scale2d, gr, b, st, g, a, c, nk, j1, j2, i1, i2);
}
void launch_timeint(const Geom& g, const double* b1, const double* b0, double* bt, double rdtbdy, int nk, int j1, int j2,
                    int i1, int i2, cudaStream_t st) {
  dim3 b(32, 8), gr((j2 - j1 + 32) / 32, (i2 - i1 + 8) / 8);
  RCM_LAUNCH(k_timeint, gr, b, st, g, b1, b0, bt, rdtbdy, nk, j1, j2, i1, i2);
}

// ---------------------------------------------------------------------------
// dot-grid rim: atm2 rim <- atm1 rim (when integrating), atm1 rim <- b0 + xt*bt
// (Main/mod_bdycod.F90:1126-1310, 1393-1425 / 1494-1525)
__global__ void k_bdyval_dot(const Geom g, const Fields f, const int2* rim, int nrim, double xt, int integrating,
                             int iboudy) {
  const int si = g.si;
  const long long sk = g.sk;
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  const int k = 1 + blockIdx.y;
  if (p >= nrim) return;
  const int j = rim[p].x, i = rim[p].y;
  if (integrating) {
    P3(f.bu, j, i, k) = P3(f.au, j, i, k);
    P3(f.bv, j, i, k) = P3(f.av, j, i, k);
  }
  if (iboudy == 0) {
    P3(f.au, j, i, k) = P3(f.xub0, j, i, k);
    P3(f.av, j, i, k) = P3(f.xvb0, j, i, k);
  } else {
    P3(f.au, j, i, k) = P3(f.xub0, j, i, k) + xt * P3(f.xubt, j, i, k);
    P3(f.av, j, i, k) = P3(f.xvb0, j, i, k) + xt * P3(f.xvbt, j, i, k);
  }
}

// cross-grid rim.  `oa*` are the level-a arrays that hold the rim values of
// the PREVIOUS bdyval (the buffers swapped out at the end of `tend`; equal to
// the current ones on the very first call).
__global__ void k_bdyval_crs(const Geom g, const Fields f, const double* oat, const double* oaqv, const double* oaqc,
                             const int2* rim, int nrim, double xt, int integrating, int iboudy) {
  const int si = g.si;
  const long long sk = g.sk;
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  const int k = 1 + blockIdx.y;
  if (p >= nrim) return;
  const int j = rim[p].x, i = rim[p].y;
  const int kz = g.kz;
  const bool half = (k <= kz);
  // global rim geometry
  const int gjce1 = 1, gjce2 = g.jx - 1, gice1 = 1, gice2 = g.iy - 1;
  const int gjci1 = 2, gjci2 = g.jx - 2, gici1 = 2, gici2 = g.iy - 2;
  if (integrating) {
    if (half) {
      P3(f.bt, j, i, k) = P3(oat, j, i, k);
      P3(f.bqv, j, i, k) = P3(oaqv, j, i, k);
      P3(f.bqc, j, i, k) = P3(oaqc, j, i, k);
      P3(f.bpp, j, i, k) = P3(f.app, j, i, k);
    }
    P3(f.bw, j, i, k) = P3(f.aw, j, i, k);
  }
#define BC(b0, bt_) ((iboudy == 0) ? P3(b0, j, i, k) : P3(b0, j, i, k) + xt * P3(bt_, j, i, k))
  if (half) {
    P3(f.at, j, i, k) = BC(f.xtb0, f.xtbt);
    P3(f.aqv, j, i, k) = BC(f.xqb0, f.xqbt);
    P3(f.app, j, i, k) = BC(f.xpb0, f.xpbt);
  }
  double w = BC(f.xwb0, f.xwbt);
  if (iboudy != 0 && k == 1) {
    const int jn = min(max(j, gjci1), gjci2), in = min(max(i, gici1), gici2);
    w = P3(f.aw, jn, in, 1);
  }
  P3(f.aw, j, i, k) = w;
#undef BC
  if (!half) return;
  // iboudy 3, 4: water vapour on the rim by inflow/outflow (Main/mod_bdycod.F90:1879-1947): inflow keeps the boundary
  // value, outflow takes the interior neighbour, both through (q/p*)*p*.  West/east rows ici1..ici2 first, then the
  // whole south/north rows, whose corner cells read the west/east result of the cell next to them.
  if (iboudy == 3 || iboudy == 4) {
    auto bcq = [&](int jj, int ii) { return P3(f.xqb0, jj, ii, k) + xt * P3(f.xqbt, jj, ii, k); };
    // result of the west/east rule at rim cell (jr, ii), ii in ici1..ici2
    auto qv_we = [&](int jr, int ii) {
      const bool west = (jr == gjce1);
      const int jin = west ? gjci1 : gjci2, jde = west ? 1 : g.jx, jdi = west ? 2 : g.jx - 1;
      const double qext = bcq(jr, ii) / P2(f.psa, jr, ii), qint = P3(f.aqv, jin, ii, k) / P2(f.psa, jin, ii);
      const double windavg = P3(f.au, jde, ii, k) + P3(f.au, jde, ii + 1, k) + P3(f.au, jdi, ii, k) + P3(f.au, jdi, ii + 1, k);
      const bool keep = west ? (windavg > 0.0) : (windavg < 0.0);
      return (keep ? qext : qint) * P2(f.psa, jr, ii);
    };
    double qn;
    if (i == gice1 || i == gice2) {
      const bool south = (i == gice1);
      const int iin = south ? gici1 : gici2, ide = south ? 1 : g.iy, idi = south ? 2 : g.iy - 1;
      const double qext = bcq(j, i) / P2(f.psa, j, i);
      const double qin_v = (j == gjce1 || j == gjce2) ? qv_we(j, iin) : P3(f.aqv, j, iin, k);
      const double qint = qin_v / P2(f.psa, j, iin);
      const double windavg = P3(f.av, j, ide, k) + P3(f.av, j + 1, ide, k) + P3(f.av, j, idi, k) + P3(f.av, j + 1, idi, k);
      const bool keep = south ? (windavg > 0.0) : (windavg < 0.0);
      qn = (keep ? qext : qint) * P2(f.psa, j, i);
    } else {
      qn = qv_we(j, i);
    }
    P3(f.aqv, j, i, k) = qn;
  }
  // hydrometeors without boundary data: zero on inflow, interior value on
  // outflow (Main/mod_bdycod.F90:2153-2221); strips as left by bdyuv.
  const bool tdep = (iboudy != 0);
  double q;
  if (j == gjce1 || j == gjce2) {
    const bool west = (j == gjce1);
    const int jin = west ? gjci1 : gjci2;        // cross column next to the rim
    const int jde = west ? 1 : g.jx;             // exterior dot column
    const int jdi = west ? 2 : g.jx - 1;         // interior dot column
    const double* src = (i == gice1 || i == gice2) ? oaqc : f.aqc;
    const double qxint = P3(src, jin, i, k) / P2(f.psa, jin, i);
    // wue/eue are never assigned at the two corner rows in the time-dependent branch
    const double e0 = (tdep && (i < 2 || i > g.iy - 1)) ? 0.0 : P3(f.au, jde, i, k);
    const double e1 = (tdep && (i + 1 < 2 || i + 1 > g.iy - 1)) ? 0.0 : P3(f.au, jde, i + 1, k);
    const double windavg = e0 + e1 + P3(f.au, jdi, i, k) + P3(f.au, jdi, i + 1, k);
    const bool inflow = west ? (windavg > 0.0) : (windavg < 0.0);
    q = inflow ? 0.0 : qxint * P2(f.psa, j, i);
  } else {
    const bool south = (i == gice1);
    const int iin = south ? gici1 : gici2;
    const int ide = south ? 1 : g.iy;
    const int idi = south ? 2 : g.iy - 1;
    const double qxint = P3(f.aqc, j, iin, k) / P2(f.psa, j, iin);
    const double windavg = P3(f.av, j, ide, k) + P3(f.av, j + 1, ide, k) + P3(f.av, j, idi, k) + P3(f.av, j + 1, idi, k);
    const bool inflow = south ? (windavg > 0.0) : (windavg < 0.0);
    q = inflow ? 0.0 : qxint * P2(f.psa, j, i);
  }
  P3(f.aqc, j, i, k) = q;
}

void launch_bdyval(const Geom& g, const Fields& f, const double* oat, const double* oaqv, const double* oaqc,
                   const int2* rim_d, int nrim_d, const int2* rim_c, int nrim_c, double xt, int integrating, int iboudy,
                   cudaStream_t st) {
  if (nrim_d > 0) {
    dim3 b(128), gr((nrim_d + 127) / 128, g.kz);
    RCM_LAUNCH(k_bdyval_dot, gr, b, st, g, f, rim_d, nrim_d, xt, integrating, iboudy);
  }
  if (nrim_c > 0) {
    dim3 b(128), gr((nrim_c + 127) / 128, g.kzp1);
    RCM_LAUNCH(k_bdyval_crs, gr, b, st, g, f, oat, oaqv, oaqc, rim_c, nrim_c, xt, integrating, iboudy);
  }
}
