// Shared host/device types of the B200 dyn-core library (libregcm_dyn.so).
//
// Data layout in HBM.  Every 3-D field, whatever grid it lives on (dot or
// cross, half or full levels), is stored in ONE uniform padded geometry per
// rank: (j,i,k) with j fastest as in the reference
// (Main/mod_atm_interface.F90:617-641), pitch `si` doubles per i-row (tile
// width + 2*HALO, rounded up to 16 doubles = 128 B), `sk` doubles per k-plane,
// kz+1 planes.  Ghost rows are always allocated (HALO = 2 on every side), so a
// stencil never needs a bounds test on the array, only on the index family it
// loops over.  Kernels receive ORIGIN-SHIFTED pointers: element (j,i,k) in the
// reference's GLOBAL 1-based indices is p[j + si*i + sk*k].  2-D fields use
// the same (j,i) geometry: p[j + si*i].
#pragma once
#include <stdint.h>

#define RCM_HALO 2
#define RCM_EHALO 6 /* estore halo for the upper radiative BC, mod_sound.F90:551-561 */
#define RCM_MAXSPLIT 4 /* upper bound on nsplit (split-explicit vertical modes, idynamic = 1) */

struct Geom {
  int jx, iy, kz, kzp1;
  // index families of Main/mod_atm_interface.F90:181-381 (global indices)
  int jde1, jde2, ide1, ide2, jdi1, jdi2, idi1, idi2, jdii1, jdii2, idii1, idii2;
  int jce1, jce2, ice1, ice2, jci1, jci2, ici1, ici2, jcii1, jcii2, icii1, icii2;
  int jcross1, jcross2, icross1, icross2;
  int has_bdyleft, has_bdyright, has_bdybottom, has_bdytop;
  // ghost ring actually usable on each side: 1 where a neighbour rank exists,
  // 0 on a physical boundary ("ga" suffix of the reference)
  int gl, gr, gb, gt;
  int si;        // doubles per i-row
  long long sk;  // doubles per k-plane
  int esi;       // pitch of the wide-halo 2-D estore array
  int kc_uv, kc_tq, kc_ppw;  // levels per thread of the slow-tendency kernels (>= kz: one thread marches the whole column)
  int pfd;       // L2 prefetch distance (levels ahead) of the column-marching kernels; 0 = off
  int pfd_wfwd, pfd_wback;   // the same for the two acoustic sweeps, which want different distances (profiles/r2_measurements.md, call O)
  int hw;        // ghost rows ALLOCATED around the tile: RCM_HALO, or 4 on a decomposed domain with isladvec = 1 (the departure
                 // stencil of Main/mod_sladvection.F90:139-221 reaches three cells out); the exchanges of the state stay RCM_HALO wide
};

// per-step scalars
struct StepScalars {
  double dt;     // current leapfrog dt (dtsec, then 2*dtsec)
  double dtsec;
  double xt;     // xbctime + dt (boundary interpolation time)
  double dts;    // acoustic sub-step
  double ul;     // uoffc*0.5*dt0/dx (start-up dt, mod_advection.F90:106)
  double dx, dx2, dx4, dx8, dx16, dxsq, rdx, rdxsq;
  double gnu1, gnu2;
  double t_extrema, q_rel_extrema;
  double dydc, xkhmax;
  double xgamma, bp, bm, bpxbp, bpxbm, xkd_snd;
  double ptop;
  int upstream_mode, stability_enhance, ipptls, iboudy, ifrayd, ifupr;
  int isladvec, iqmsl;   // semi-Lagrangian horizontal advection of qv, qc, tracers and its quasi-monotonic limiter
  int idiffu;   // 1: 4th order with the 2nd-order boundary ring; 2: nine-point operator (Main/mod_diffusion.F90)
  int kt_ray;   // min(kz, rayndamp)
  int ktf_ray;  // min(kzp1, rayndamp)
  int nspgx, nspgd;
  int it, istep;
  // negative-moisture repair: 0 = fused pass (count negatives; store the forecast on the tile perimeter only),
  // 1 = second pass after negatives were found (store the forecast everywhere, list the negative points, no
  // other output; exits at once when the count is zero), 2 = single pass that stores and lists everything
  // (chosen when the previous call saw negatives)
  int cq_mode;
  // ---- hydrostatic core (idynamic = 1) ----
  int hydro;     // 1: idynamic = 1
  int ipgf;      // pressure gradient force scheme of the hydrostatic core (0, 1)
  int integrating;  // rcmtimer%integrating(): lcount > 0
  // split-explicit scheme (Main/mod_split.F90): per mode l = 0..nsplit-1
  int nsplit;
  int sp_m2[RCM_MAXSPLIT];        // 2*int(aam)
  double sp_aam[RCM_MAXSPLIT], sp_dtau[RCM_MAXSPLIT], sp_hbar[RCM_MAXSPLIT], sp_an[RCM_MAXSPLIT];
  double sp_pd;                   // vmodes pd = xps - ptop
};

// Device pointers (origin-shifted).  Names follow the reference's structures:
// a* = atm1 (time level a), b* = atm2 (level b), c* = atmc, n?* = the
// double-buffered "next" a/b copies of t,qv,qc written by the fused slow
// kernels and swapped in at the end of the step.
// Tracers (atm1%chi / atm2%chi): one padded 3-D block per species, handed to the tracer kernels only
#ifndef RCM_MAXTR
#define RCM_MAXTR 8
#endif
struct TrFields {
  int ntr;
  double *a[RCM_MAXTR], *b[RCM_MAXTR];      // atm1%chi, atm2%chi (p*-coupled)
  double *na[RCM_MAXTR], *nb[RCM_MAXTR];    // the filtered time levels of the step being computed (swapped in at its end)
  double *c[RCM_MAXTR];                     // atmc%chi: unfiltered forecast (negative fix)
  double *b0[RCM_MAXTR], *bt[RCM_MAXTR];    // chib0, chibt (coupled)
  const double *cefc, *cegc;                // setup_che_bdycon tables [ib + nspgx*(k-1)]
  const int* ztype;                         // 1 where the relaxation zone is the west/east one (neighbour order of nudge_chiten)
  int* nq_list;                             // negative forecasts per (species, level) plane, as in Fields::nq_list
  int* nq_cnt;
  unsigned long long* negcnt;               // number of negative forecasts of the step
  double* sl[RCM_MAXTR];                    // isladvec = 1: semi-Lagrangian advection tendency left by k_sl
};

// UW-PBL coupling of the transport (ibltyp = 2): TKE on full levels, NOT coupled with p*
// (atm1%tke, atm2%tke, atmc%tke; Main/mod_atm_interface.F90:604-606,625-626,656-658)
struct UwFields {
  double *a, *b, *c;        // time levels a, b and the unfiltered forecast
  double* phy;              // aten%tke(:,:,:,pc_physic), zero until set
  double tkemin, nuk;
};

struct Fields {
  // prognostic state
  double *au, *av, *at, *aqv, *aqc, *app, *aw;
  double *bu, *bv, *bt, *bqv, *bqc, *bpp, *bw;
  double *nat, *naqv, *naqc, *nbt, *nbqv, *nbqc;
  // boundary records b0, bt (coupled) per field: u v t qv pp w
  double *xub0, *xubt, *xvb0, *xvbt, *xtb0, *xtbt, *xqb0, *xqbt, *xpb0, *xpbt, *xwb0, *xwbt;
  // atm0 reference atmosphere
  double *pr0, *t0, *rho0, *rhof0;
  // 2-D
  double *psa, *rpsa, *psb, *rpsb, *psdota, *rpsda, *psdotb;
  double *msfx, *msfd, *xmsf, *dmsf, *coriol, *ef, *ddx, *ddy, *dmdx, *dmdy, *ex, *crx, *cry;
  double *dpsdxm, *dpsdym, *ht, *ps0, *hgfact;
  int *zone_c, *zone_d;  // relaxation-zone distance index ibnd (0 = outside)
  // 1-D vertical tables (index k = 1-based, element [k])
  double *sigma, *dsigma, *twt1, *twt2, *qcon, *xds, *dds;
  double *hefc, *hegc, *hefd, *hegd;  // [ib + nspg*(k-1)] ; iboudy=1: same value for all k
  // Rayleigh coefficients tau(z) precomputed from the static heights
  double *tau_c, *tau_d, *tau_f;  // same geometry, levels 1..rayndamp
  // slow-part work fields
  double *pr1, *rho1, *th, *qdot, *cr, *xkc;
  double *scc, *scdd, *sca;  // per-step invariants of the implicit w solve (k_snd_wfwd, first sub-step)
  double *dub, *dvb;  // level-b winds decoupled and divided by the dot map factor (operands of diffu_d)
  double *uten, *vten, *ppten, *wten;
  // acoustic work fields
  double *cu, *cv, *cpp, *cppd, *cw, *ct, *se, *sf;
  double *estore, *astore, *wpval, *tmask;  // estore: wide-halo 2-D; tmask 13x13
  // ---- hydrostatic core (idynamic = 1); null / aliases for idynamic = 2 ----
  double *rpsdb;                 // 1/psdotb (alias of rpsda for idynamic = 2)
  double *psc, *pten;            // forecast p*, p* tendency before the boundary relaxation (2-D)
  double *ptn2d, *pt22d;         // per-point terms of the Bleck noise parameters (2-D)
  double *xpsb0, *xpsbt;         // boundary record of p* (2-D): b0, bt
  double *phi, *omega;           // geopotential (3-D); omega tap (3-D)
  double *nau, *nav, *nbu, *nbv; // double-buffered next levels of p*u, p*v
  // split-explicit work: planes are (slot-1)*nsplit + l for deld/delh (slot 1..3), l for the others
  double *deld, *delh, *ddsum, *dhsum, *dstor, *hstor, *map;
  // vertical-mode tables: [l + nsplit*(k-1)] for zr (zmatxr), tau, pdlog, eps1 (k = 1..kz+1), am, zm (zmatx)
  double *sp_zr, *sp_tau, *sp_pdlog, *sp_eps1, *sp_am, *sp_zm;
  // aten(:,:,:,pc_physic): tendencies handed in by the physical parametrizations (rcm_dyn_set_phys_tend), coupled
  // with p* like the reference's; null until first set (physics off), added at Main/mod_tendency.F90:285-314,331-335,404-411
  double *phy_u, *phy_v, *phy_t, *phy_qv, *phy_qc, *phy_pp, *phy_w;
  // diagnostics
  unsigned long long *cflbits;  // max CFL as ordered bits
  unsigned long long *negq;     // negative-q repairs
  // negative-moisture repair (Main/mod_tendency.F90:381-393; k_negq.cu): unfiltered forecast atmc%qx of qv and
  // qc, the list of negative points per (species, level) plane and its fill counts
  double *cqv, *cqc;
  int *nq_list;                 // [plane * sk + slot] = local offset jl + il*si, plane = n*kz + (k-1)
  int *nq_cnt;                  // [plane]
  int *nanflag;
  // vadv4d ind = 3 (iuwvadv = 1 with ibltyp = 2): top of the PBL per cross point, 2-D geometry; null = ind 1 / 2
  const int* kpb;
  // isladvec = 1: semi-Lagrangian horizontal advection tendencies of qv, qc left by k_sl (null otherwise)
  double *slqv, *slqc;
  // idiffu = 3: the diffusion tendencies of the one column the operator acts on (j = jdi2 / jci2), left by k_diff6:
  // [(field * kzp1 + k - 1) * (iy + 2) + i], fields RCM_D6_* (null otherwise)
  double* d6;
};
enum { RCM_D6_U = 0, RCM_D6_V, RCM_D6_T, RCM_D6_QV, RCM_D6_QC, RCM_D6_PP, RCM_D6_W, RCM_D6_TKE, RCM_D6_CHI };

// ---------------------------------------------------------------------------
// MOLOCH core (idynamic = 3; Main/mod_moloch.F90, Arakawa C grid, zita levels).
// Same padded (j,i,k) geometry and origin-shifted pointers as above: u lives on
// (jx, iy-1), v on (jx-1, iy), everything else on cross points.
#define RCM_MO_MAXF 7 /* fields advected by wafone: tetav, pai, ux, vx, wx, qv, qc */

struct MoScalars {
  double dtsec, dt, xt, dx, dz;
  double dts, zdtrdx, zdtrdy, zdtrdz, zcs2, egdts;  // sound sub-step (Main/mod_moloch.F90:720-723)
  double dta, adtrdx, adtrdy, adtrdz;               // advection step (:1109-1111)
  double anu2;
  double zz1, gzm1;                                 // surface-pressure closure (:377-383)
  int ipptls, iboudy, ifrayd, kray, nspgx, nspgd, nf;
  int rotllr;   // iproj == 'ROTLLR' branches (Main/mod_moloch.F90:773,932,1183)
};

struct MoFields {
  double *u, *v, *un, *vn, *w, *t, *pai, *tetav, *qv, *qc, *ux, *vx, *wx;
  double *s, *zdiv, *deltaw, *wwkw, *p, *tvirt;
  double *adv[RCM_MO_MAXF], *wz[RCM_MO_MAXF], *p0[RCM_MO_MAXF];
  double *fmz, *fmzf, *zru, *zrv, *zeta, *tau_c, *tau_u, *tau_v;
  double *ht, *hx, *hy, *mx2, *mx, *rmu, *rmv, *msfu, *msfv, *coriou, *coriov, *ps;
  double *gzitak, *gzitakh, *ffilt;
  double *xfc, *xgc, *xfd, *xgd;  // relaxation coefficients [ib + nspg*(k-1)]
  double *xub0, *xubt, *xvb0, *xvbt, *xtb0, *xtbt, *xqb0, *xqbt, *xpb0, *xpbt;
  int *zone_c, *zone_u, *zone_v;  // bits 0-3: south, north, west, east; bits 8..: ibnd
  double* rimbuf;
  int* nanflag;
  unsigned long long* psbits;  // [0] max of ps over the interior as ordered bits, [1] complement of the min's bits
  // physics coupling: mo_atm%tten, qxten, uten, vten handed in by the parametrizations (null until set),
  // and zetaf (kz+1 levels), the height the physics sees as atms%zq
  double *phy_t, *phy_qv, *phy_qc, *phy_u, *phy_v;
  double* zetaf;
};
