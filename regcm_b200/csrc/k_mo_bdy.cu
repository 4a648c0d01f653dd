// MOLOCH core (idynamic = 3): lateral boundary and set-up kernels.
//   k_mo_nudge_a/_b  monudge3d (pai, t, qv) and monudgeuv      Main/mod_bdycod.F90:4408-4595,3825-4165
//   k_mo_raydamp     moraydamp (u, v, t, pai)                   :5047-5065
//   k_mo_bdyval_uv / k_mo_bdyval_c   bdyval, idynamic 3         :1343-1385,1452-1491,1539-1575,1662-1699,1963-2100
//   k_mo_paicompute  hydrostatic Exner function of a record     :5125-5155
//   k_mo_init_state  initial state from record b0               Main/mod_init.F90:178-336
// The relaxation updates a field in place from differences to the boundary record taken at the point and
// its four neighbours BEFORE any update (the reference fills fg1/fg2 first): phase a evaluates the new
// values of the relaxation-zone points into a compact buffer, phase b stores them.
#include "dev_common.cuh"
#include "launch.h"
#include "../../include/regcm_dyn.h"

using namespace rcm;

// which: 0 = cross fields (pai, t, qv: three slots), 1 = u, 2 = v
__device__ __forceinline__ double mo_fg(const double* b0, const double* bt, const double* fld, const double xt,
                                        const int si, const long long sk, const int j, const int i, const int k) {
  return (L3(b0, j, i, k) + xt * L3(bt, j, i, k)) - L3(fld, j, i, k);
}

__device__ __forceinline__ double mo_nudged(const double* b0, const double* bt, const double* fld, const double xt,
                                            const double xf, const double xg, const int si, const long long sk,
                                            const int j, const int i, const int k) {
  const double fls0 = mo_fg(b0, bt, fld, xt, si, sk, j, i, k), fls1 = mo_fg(b0, bt, fld, xt, si, sk, j - 1, i, k),
               fls2 = mo_fg(b0, bt, fld, xt, si, sk, j + 1, i, k), fls3 = mo_fg(b0, bt, fld, xt, si, sk, j, i - 1, k),
               fls4 = mo_fg(b0, bt, fld, xt, si, sk, j, i + 1, k);
  return L3(fld, j, i, k) + xf * fls0 - xg * (fls1 + fls2 + fls3 + fls4 - 4.0 * fls0);
}

__global__ void k_mo_nudge_a(const Geom g, const MoFields f, const MoScalars s, const int2* rim, const int nrim,
                             const int which) {
  const int si = g.si; const long long sk = g.sk;
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  const int k = 1 + blockIdx.y;
  if (p >= nrim) return;
  const int j = rim[p].x, i = rim[p].y;
  const int z = (which == 0 ? f.zone_c : which == 1 ? f.zone_u : f.zone_v)[(long long)j + (long long)i * si];
  const int ib = z >> 8;
  // coefficient table: cross fields use the "x" tables; u uses them on the south/north sides and the
  // "d" tables on the west/east sides, v the other way round (as written in monudgeuv)
  const bool sn = (z & 3) != 0;
  const bool dtab = (which == 1) ? !sn : (which == 2) ? sn : false;
  const int nsp = dtab ? s.nspgd : s.nspgx;
  const double xf = LD1((dtab ? f.xfd : f.xfc) + ib + nsp * (k - 1)), xg = LD1((dtab ? f.xgd : f.xgc) + ib + nsp * (k - 1));
  const long long o = ((long long)(k - 1) * nrim + p) * 3;
  if (s.iboudy == 4) {   // sponge: f + (1 - wgt(ib)) * bt, the tables hold wgt (mosponge*, Main/mod_bdycod.F90:2665-3063)
    const double w1 = 1.0 - xf;
    if (which == 0) {
      f.rimbuf[o + 0] = P3(f.pai, j, i, k) + w1 * L3(f.xpbt, j, i, k);
      f.rimbuf[o + 1] = P3(f.t, j, i, k) + w1 * L3(f.xtbt, j, i, k);
      f.rimbuf[o + 2] = P3(f.qv, j, i, k) + w1 * L3(f.xqbt, j, i, k);
    } else if (which == 1) {
      f.rimbuf[o] = P3(f.u, j, i, k) + w1 * L3(f.xubt, j, i, k);
    } else {
      f.rimbuf[o] = P3(f.v, j, i, k) + w1 * L3(f.xvbt, j, i, k);
    }
    return;
  }
  if (which == 0) {
    f.rimbuf[o + 0] = mo_nudged(f.xpb0, f.xpbt, f.pai, s.xt, xf, xg, si, sk, j, i, k);
    f.rimbuf[o + 1] = mo_nudged(f.xtb0, f.xtbt, f.t, s.xt, xf, xg, si, sk, j, i, k);
    f.rimbuf[o + 2] = mo_nudged(f.xqb0, f.xqbt, f.qv, s.xt, xf, xg, si, sk, j, i, k);
  } else if (which == 1) {
    f.rimbuf[o] = mo_nudged(f.xub0, f.xubt, f.u, s.xt, xf, xg, si, sk, j, i, k);
  } else {
    f.rimbuf[o] = mo_nudged(f.xvb0, f.xvbt, f.v, s.xt, xf, xg, si, sk, j, i, k);
  }
}

__global__ void k_mo_nudge_b(const Geom g, const MoFields f, const int2* rim, const int nrim, const int which) {
  const int si = g.si; const long long sk = g.sk;
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  const int k = 1 + blockIdx.y;
  if (p >= nrim) return;
  const int j = rim[p].x, i = rim[p].y;
  const long long o = ((long long)(k - 1) * nrim + p) * 3;
  if (which == 0) {
    P3(f.pai, j, i, k) = f.rimbuf[o + 0]; P3(f.t, j, i, k) = f.rimbuf[o + 1]; P3(f.qv, j, i, k) = f.rimbuf[o + 2];
  } else if (which == 1) {
    P3(f.u, j, i, k) = f.rimbuf[o];
  } else {
    P3(f.v, j, i, k) = f.rimbuf[o];
  }
}

// levels 1..kray (the deepest level where any tau is non-zero)
__global__ void k_mo_raydamp(const Geom g, const MoFields f, const MoScalars s) {
  const int si = g.si; const long long sk = g.sk;
  const int j = g.jde1 + blockIdx.x * blockDim.x + threadIdx.x;
  const int i = g.ide1 + blockIdx.y * blockDim.y + threadIdx.y;
  const int k = 1 + blockIdx.z;
  if (j > g.jde2 || i > g.ide2) return;
  const double dt = s.dt, xt = s.xt;
  if (j >= g.jdi1 && j <= g.jdi2 && i >= g.ici1 && i <= g.ici2) {
    const double bval = L3(f.xub0, j, i, k) + xt * L3(f.xubt, j, i, k), var = P3(f.u, j, i, k);
    P3(f.u, j, i, k) = var + dt * L3(f.tau_u, j, i, k) * (bval - var);
  }
  if (j >= g.jci1 && j <= g.jci2 && i >= g.idi1 && i <= g.idi2) {
    const double bval = L3(f.xvb0, j, i, k) + xt * L3(f.xvbt, j, i, k), var = P3(f.v, j, i, k);
    P3(f.v, j, i, k) = var + dt * L3(f.tau_v, j, i, k) * (bval - var);
  }
  if (j >= g.jci1 && j <= g.jci2 && i >= g.ici1 && i <= g.ici2) {
    const double tau = L3(f.tau_c, j, i, k);
    double bval = L3(f.xtb0, j, i, k) + xt * L3(f.xtbt, j, i, k), var = P3(f.t, j, i, k);
    P3(f.t, j, i, k) = var + dt * tau * (bval - var);
    bval = L3(f.xpb0, j, i, k) + xt * L3(f.xpbt, j, i, k); var = P3(f.pai, j, i, k);
    P3(f.pai, j, i, k) = var + dt * tau * (bval - var);
  }
}

// outermost rows of u (rim list of the u grid) and of v
__global__ void k_mo_bdyval_uv(const Geom g, const MoFields f, const MoScalars s, const int2* rim, const int nrim,
                               const int isv) {
  const int si = g.si; const long long sk = g.sk;
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  const int k = 1 + blockIdx.y;
  if (p >= nrim) return;
  const int j = rim[p].x, i = rim[p].y;
  const double* b0 = isv ? f.xvb0 : f.xub0;
  const double* bt = isv ? f.xvbt : f.xubt;
  double* fld = isv ? f.v : f.u;
  P3(fld, j, i, k) = (s.iboudy == 0) ? L3(b0, j, i, k) : L3(b0, j, i, k) + s.xt * L3(bt, j, i, k);
}

// outermost rows of the cross grid: t, qv, pai from the record; hydrometeors by inflow/outflow
__global__ void k_mo_bdyval_c(const Geom g, const MoFields f, const MoScalars s, const int2* rim, const int nrim) {
  const int si = g.si; const long long sk = g.sk;
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  const int k = 1 + blockIdx.y;
  if (p >= nrim) return;
  const int j = rim[p].x, i = rim[p].y;
  if (s.iboudy == 0) {
    P3(f.t, j, i, k) = L3(f.xtb0, j, i, k); P3(f.qv, j, i, k) = L3(f.xqb0, j, i, k); P3(f.pai, j, i, k) = L3(f.xpb0, j, i, k);
  } else {
    P3(f.t, j, i, k) = L3(f.xtb0, j, i, k) + s.xt * L3(f.xtbt, j, i, k);
    P3(f.qv, j, i, k) = L3(f.xqb0, j, i, k) + s.xt * L3(f.xqbt, j, i, k);
    P3(f.pai, j, i, k) = L3(f.xpb0, j, i, k) + s.xt * L3(f.xpbt, j, i, k);
  }
  if (s.iboudy == 3 || s.iboudy == 4) {   // qv by inflow/outflow (:1813-1890): west/east rows first, then the full south/north rows
    const bool west = g.has_bdyleft && (j == g.jce1), east = g.has_bdyright && (j == g.jce2),
               south = g.has_bdybottom && (i == g.ice1), north = g.has_bdytop && (i == g.ice2);
    // value of a west/east rim point of row ii after the first pass (ii is an interior row)
    auto we = [&](const int ii) {
      const int jin = west ? g.jci1 : g.jci2, jd = west ? g.jde1 : g.jde2, jdn = west ? g.jdi1 : g.jdi2;
      const double qext = (s.iboudy == 0) ? L3(f.xqb0, j, ii, k) : L3(f.xqb0, j, ii, k) + s.xt * L3(f.xqbt, j, ii, k);
      const double qint = P3(f.qv, jin, ii, k);
      const double windavg = P3(f.u, jd, ii, k) + P3(f.u, jdn, ii, k);
      return (west ? (windavg > 0.0) : (windavg < 0.0)) ? qext : qint;
    };
    if (south || north) {
      const int iin = south ? g.ici1 : g.ici2, idd = south ? g.ide1 : g.ide2, idn = south ? g.idi1 : g.idi2;
      const double qext = P3(f.qv, j, i, k);                                    // just set from the record above
      const double qint = (west || east) ? we(iin) : P3(f.qv, j, iin, k);       // corner: the west/east pass came first
      const double windavg = P3(f.v, j, idd, k) + P3(f.v, j, idn, k);
      P3(f.qv, j, i, k) = (south ? (windavg > 0.0) : (windavg < 0.0)) ? qext : qint;
    } else {
      P3(f.qv, j, i, k) = we(i);
    }
  }
  if (s.ipptls > 0) {
    const bool west = g.has_bdyleft && (j == g.jce1), east = g.has_bdyright && (j == g.jce2),
               south = g.has_bdybottom && (i == g.ice1), north = g.has_bdytop && (i == g.ice2);
    double qxint, windavg; bool zero;
    if (west || east) {
      const int jin = west ? g.jci1 : g.jci2, jd = west ? g.jde1 : g.jde2, jdn = west ? g.jdi1 : g.jdi2;
      if (south || north) {   // corners: the fourth term reads u where v is meant (kept as in the reference)
        const int iin = south ? g.ici1 : g.ici2, idd = south ? g.ide1 : g.ide2, idn = south ? g.idi1 : g.idi2;
        qxint = fmax(P3(f.qc, jin, iin, k), 0.0);
        windavg = (P3(f.u, jd, i, k) + P3(f.u, jdn, i, k) + P3(f.v, j, idd, k) + P3(f.u, j, idn, k));
      } else {
        qxint = fmax(P3(f.qc, jin, i, k), 0.0);
        windavg = (P3(f.u, jd, i, k) + P3(f.u, jdn, i, k));
      }
      zero = west ? (windavg >= 0.0) : (windavg <= 0.0);
    } else {
      const int iin = south ? g.ici1 : g.ici2, idd = south ? g.ide1 : g.ide2, idn = south ? g.idi1 : g.idi2;
      qxint = fmax(P3(f.qc, j, iin, k), 0.0);
      windavg = (P3(f.v, j, idd, k) + P3(f.v, j, idn, k));
      zero = south ? (windavg >= 0.0) : (windavg <= 0.0);
    }
    P3(f.qc, j, i, k) = zero ? 0.0 : qxint;
  }
}

// pai of one boundary record from ps [Pa], t, qv (column march upward)
__global__ void k_mo_paicompute(const Geom g, const MoFields f, const MoScalars s, const double* ps, const double* t,
                                const double* q, double* pai, const int j1, const int j2, const int i1, const int i2) {
  const int si = g.si; const long long sk = g.sk;
  const int j = j1 + blockIdx.x * blockDim.x + threadIdx.x;
  const int i = i1 + blockIdx.y * blockDim.y + threadIdx.y;
  if (j > j2 || i > i2) return;
  const int kz = g.kz;
  const double zdelta0 = L2(f.ht, j, i) * s.gzm1 + s.zz1;
  double tv2 = L3(t, j, i, kz) * (1.0 + ep1 * L3(q, j, i, kz));
  const double zz2 = 1.0 / (rgas * tv2);
  const double p = P2(ps, j, i) * exp(-zdelta0 * zz2);
  double pk = pow(p / p00, rovcp);
  P3(pai, j, i, kz) = pk;
  for (int k = kz - 1; k >= 1; --k) {
    const double tv1 = L3(t, j, i, k) * (1.0 + ep1 * L3(q, j, i, k));
    const double zb = 2.0 * egrav * s.dz / (L3(f.fmzf, j, i, k + 1) * cpd) + tv1 - tv2;
    const double zdelta = sqrt(zb * zb + 4.0 * tv2 * tv1);
    pk = -pk / (2.0 * tv2) * (zb - zdelta);
    P3(pai, j, i, k) = pk;
    tv2 = tv1;
  }
}

__device__ __forceinline__ double mo_pfwsat(const double t, const double p) {   // Share/pfesat.inc, Share/pfwsat.inc
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

__global__ void k_mo_init_state(const Geom g, const MoFields f, const MoScalars s) {
  const int si = g.si; const long long sk = g.sk;
  const int j = g.jde1 + blockIdx.x * blockDim.x + threadIdx.x;
  const int i = g.ide1 + blockIdx.y * blockDim.y + threadIdx.y;
  if (j > g.jde2 || i > g.ide2) return;
  const int kz = g.kz;
  if (j <= g.jce2) for (int k = 1; k <= kz; ++k) P3(f.v, j, i, k) = L3(f.xvb0, j, i, k);
  if (i <= g.ice2) for (int k = 1; k <= kz; ++k) P3(f.u, j, i, k) = L3(f.xub0, j, i, k);
  if (j > g.jce2 || i > g.ice2) return;
  double tv = 0.0, pk = 0.0;
  for (int k = 1; k <= kz; ++k) {
    const double t = L3(f.xtb0, j, i, k);
    double qv = L3(f.xqb0, j, i, k);
    const double pai = L3(f.xpb0, j, i, k);
    tv = t * (1.0 + ep1 * qv);
    pk = pow(pai, cpovr) * p00;
    const double qs = mo_pfwsat(t, pk);
    qv = fmin(qv, qs * 1.001);
    P3(f.t, j, i, k) = t; P3(f.qv, j, i, k) = qv; P3(f.pai, j, i, k) = pai;
    P3(f.qc, j, i, k) = (s.ipptls == 1) ? minqc : 0.0;
    P3(f.p, j, i, k) = pk;
    P3(f.tvirt, j, i, k) = t * (1.0 + ep1 * qv);
  }
  const double zdgz = L2(f.ht, j, i) * s.gzm1 + s.zz1;
  P2(f.ps, j, i) = pk * exp(zdgz / (rgas * tv));
  const double hx0 = L2(f.hx, j, i), hx1 = L2(f.hx, j + 1, i), hy0 = L2(f.hy, j, i), hy1 = L2(f.hy, j, i + 1);
  for (int k = 2; k <= kz - 1; ++k) {
    const double zuh = L3(f.xub0, j, i, k) * hx0 + L3(f.xub0, j + 1, i, k) * hx1 + L3(f.xub0, j, i, k + 1) * hx0 + L3(f.xub0, j + 1, i, k + 1) * hx1;
    const double zvh = L3(f.xvb0, j, i, k) * hy0 + L3(f.xvb0, j, i + 1, k) * hy1 + L3(f.xvb0, j, i, k + 1) * hy0 + L3(f.xvb0, j, i + 1, k + 1) * hy1;
    P3(f.w, j, i, k) = 0.25 * (zuh + zvh);
  }
  {
    const double zuh = L3(f.xub0, j, i, kz) * hx0 + L3(f.xub0, j + 1, i, kz) * hx1;
    const double zvh = L3(f.xvb0, j, i, kz) * hy0 + L3(f.xvb0, j, i + 1, kz) * hy1;
    P3(f.w, j, i, kz) = 0.5 * (zuh + zvh);
  }
}

__global__ void k_mo_scale2d(const Geom g, double* a, const double c, const int j1, const int j2, const int i1, const int i2) {
  const int si = g.si;
  const int j = j1 + blockIdx.x * blockDim.x + threadIdx.x;
  const int i = i1 + blockIdx.y * blockDim.y + threadIdx.y;
  if (j > j2 || i > i2) return;
  P2(a, j, i) = P2(a, j, i) * c;
}

static dim3 grid2(int nj, int ni, dim3 b) { return dim3((nj + b.x - 1) / b.x, (ni + b.y - 1) / b.y, 1); }

void launch_mo_nudge(const Geom& g, const MoFields& f, const MoScalars& s, const int2* rim, int nrim, int which, cudaStream_t st) {
  if (nrim <= 0) return;
  dim3 b(128), gr((nrim + 127) / 128, g.kz);
  RCM_LAUNCH(k_mo_nudge_a, gr, b, st, g, f, s, rim, nrim, which);
  RCM_LAUNCH(k_mo_nudge_b, gr, b, st, g, f, rim, nrim, which);
}
void launch_mo_raydamp(const Geom& g, const MoFields& f, const MoScalars& s, cudaStream_t st) {
  if (s.kray < 1) return;
  dim3 b(32, 4), gr((g.jde2 - g.jde1 + 32) / 32, (g.ide2 - g.ide1 + 4) / 4, s.kray);
  RCM_LAUNCH(k_mo_raydamp, gr, b, st, g, f, s);
}
void launch_mo_bdyval(const Geom& g, const MoFields& f, const MoScalars& s, const int2* rim_u, int nu, const int2* rim_v, int nv,
                      const int2* rim_c, int nc, cudaStream_t st) {
  dim3 b(128);
  if (nu > 0) RCM_LAUNCH(k_mo_bdyval_uv, dim3((nu + 127) / 128, g.kz), b, st, g, f, s, rim_u, nu, 0);
  if (nv > 0) RCM_LAUNCH(k_mo_bdyval_uv, dim3((nv + 127) / 128, g.kz), b, st, g, f, s, rim_v, nv, 1);
  if (nc > 0) RCM_LAUNCH(k_mo_bdyval_c, dim3((nc + 127) / 128, g.kz), b, st, g, f, s, rim_c, nc);
}
void launch_mo_paicompute(const Geom& g, const MoFields& f, const MoScalars& s, const double* ps, const double* t, const double* q,
                          double* pai, int j1, int j2, int i1, int i2, cudaStream_t st) {
  dim3 b = col_block();
  RCM_LAUNCH(k_mo_paicompute, grid2(j2 - j1 + 1, i2 - i1 + 1, b), b, st, g, f, s, ps, t, q, pai, j1, j2, i1, i2);
}
void launch_mo_init_state(const Geom& g, const MoFields& f, const MoScalars& s, cudaStream_t st) {
  dim3 b = col_block();
  RCM_LAUNCH(k_mo_init_state, grid2(g.jde2 - g.jde1 + 1, g.ide2 - g.ide1 + 1, b), b, st, g, f, s);
}
void launch_mo_scale2d(const Geom& g, double* a, double c, int j1, int j2, int i1, int i2, cudaStream_t st) {
  dim3 b(32, 4);
  RCM_LAUNCH(k_mo_scale2d, grid2(j2 - j1 + 1, i2 - i1 + 1, b), b, st, g, a, c, j1, j2, i1, i2);
}

// ---------------------------------------------------------------------------------------------------------------
// Physics coupling of the MOLOCH core.
//   k_mo_phys    the state update at the end of physical_parametrizations (Main/mod_moloch.F90:1548-1557): t, u, v, qx
//                advance by dtsec times the tendencies handed in; hydrometeors floored at 0, qv at minqq
//   k_mo_slice   one member of `atms` on request: the aliases of mo_atm (Main/mod_slice.F90:82-100) and what mkslice
//                computes for idynamic = 3 (:107-146)
__global__ void k_mo_phys(const Geom g, const MoFields f, const double dtsec) {
  const int si = g.si; const long long sk = g.sk;
  const int j = g.jde1 + blockIdx.x * blockDim.x + threadIdx.x;
  const int i = g.ide1 + blockIdx.y * blockDim.y + threadIdx.y;
  const int k = 1 + blockIdx.z;
  if (j > g.jde2 || i > g.ide2) return;
  const bool ci = (i >= g.ici1 && i <= g.ici2), cj = (j >= g.jci1 && j <= g.jci2);
  if (cj && ci) {
    P3(f.t, j, i, k) = P3(f.t, j, i, k) + dtsec * L3(f.phy_t, j, i, k);
    const double qv = P3(f.qv, j, i, k) + dtsec * L3(f.phy_qv, j, i, k);
    const double qc = P3(f.qc, j, i, k) + dtsec * L3(f.phy_qc, j, i, k);
    P3(f.qc, j, i, k) = fmax(qc, 0.0);
    P3(f.qv, j, i, k) = fmax(qv, minqq);
  }
  if (j >= g.jdi1 && j <= g.jdi2 && ci) P3(f.u, j, i, k) = P3(f.u, j, i, k) + dtsec * L3(f.phy_u, j, i, k);
  if (cj && i >= g.idi1 && i <= g.idi2) P3(f.v, j, i, k) = P3(f.v, j, i, k) + dtsec * L3(f.phy_v, j, i, k);
}

__global__ void __launch_bounds__(128) k_mo_slice(const Geom g, const MoFields f, const int id, double* __restrict__ out) {
  const int si = g.si; const long long sk = g.sk;
  const int j = g.jce1 + blockIdx.x * blockDim.x + threadIdx.x;
  const int i = g.ice1 + blockIdx.y * blockDim.y + threadIdx.y;
  if (j > g.jce2 || i > g.ice2) return;
  const int kz = g.kz, kzp1 = g.kzp1;
  const bool inner = (j >= g.jci1 && j <= g.jci2 && i >= g.ici1 && i <= g.ici2);
  const double ps = L2(f.ps, j, i);
  int nk = kz;
  if (id == RCM_SL_PF3D || id == RCM_SL_WB3D || id == RCM_SL_ZQ) nk = kzp1;
  if (id == RCM_SL_PS2D) nk = 1;
  for (int k = 1; k <= nk; ++k) {
    const int kh = min(k, kz);
    const double qvv = inner ? fmax(L3(f.qv, j, i, kh), minqq) : L3(f.qv, j, i, kh);
    double v = 0.0;
    switch (id) {
      case RCM_SL_TB3D: v = L3(f.t, j, i, k); break;
      case RCM_SL_QVB3D: v = qvv; break;
      case RCM_SL_QCB3D: v = inner ? fmax(L3(f.qc, j, i, k), 0.0) : L3(f.qc, j, i, k); break;
      case RCM_SL_UBX3D: v = L3(f.ux, j, i, k); break;
      case RCM_SL_VBX3D: v = L3(f.vx, j, i, k); break;
      case RCM_SL_PB3D: v = L3(f.p, j, i, k); break;
      case RCM_SL_PF3D:
        if (k == 1) v = 100.0;
        else if (k == kzp1) v = ps;
        else v = p00 * pow(0.5 * (L3(f.pai, j, i, k) + L3(f.pai, j, i, k - 1)), cpovr);
        break;
      case RCM_SL_PS2D: v = ps; break;
      case RCM_SL_RHOB3D: v = L3(f.p, j, i, k) / (rgas * L3(f.t, j, i, k)); break;
      case RCM_SL_TH3D: v = L3(f.t, j, i, k) * pow(p00 / L3(f.p, j, i, k), rovcp); break;
      case RCM_SL_TV3D: v = L3(f.tvirt, j, i, k); break;
      case RCM_SL_TP3D: v = inner ? L3(f.t, j, i, k) * pow(ps / L3(f.p, j, i, k), rovcp) : 0.0; break;
      case RCM_SL_WPX3D: v = inner ? -0.5 * egrav * (L3(f.p, j, i, k) / (rgas * L3(f.t, j, i, k))) * (L3(f.w, j, i, k) + L3(f.w, j, i, k + 1)) : 0.0; break;
      case RCM_SL_WB3D: v = L3(f.w, j, i, k); break;
      case RCM_SL_ZQ: v = L3(f.zetaf, j, i, k); break;
      case RCM_SL_ZA: v = L3(f.zeta, j, i, k); break;
      case RCM_SL_DZQ: v = L3(f.zetaf, j, i, k) - L3(f.zetaf, j, i, k + 1); break;
      case RCM_SL_QSB3D: v = mo_pfwsat(L3(f.t, j, i, k), L3(f.p, j, i, k)); break;
      case RCM_SL_RHB3D: v = inner ? fmin(fmax(qvv / mo_pfwsat(L3(f.t, j, i, k), L3(f.p, j, i, k)), 0.01), 1.01) : 0.0; break;
    }
    P3(out, j, i, k) = v;
  }
}

void launch_mo_phys(const Geom& g, const MoFields& f, double dtsec, cudaStream_t st) {
  dim3 b(32, 4), gr((g.jde2 - g.jde1 + 32) / 32, (g.ide2 - g.ide1 + 4) / 4, g.kz);
  RCM_LAUNCH(k_mo_phys, gr, b, st, g, f, dtsec);
}
void launch_mo_slice(const Geom& g, const MoFields& f, int id, double* out, cudaStream_t st) {
  dim3 b(32, 4), gr((g.jce2 - g.jce1 + 32) / 32, (g.ice2 - g.ice1 + 4) / 4);
  RCM_LAUNCH(k_mo_slice, gr, b, st, g, f, id, out);
}
