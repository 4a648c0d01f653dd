/*
 * regcm_dyn.h — C ABI of the B200-native RegCM dynamical-core step.
 *
 * This is the drop-in boundary for the reference's dyn-core hot path.  The
 * reference exposes that path as argument-less Fortran module procedures that
 * operate on module-global pointer arrays; each entry point below cites the
 * reference interface it replaces (paths relative to the reference tree).
 * A Fortran `bind(C)` interface block for these functions is shown in
 * INTEGRATION.md.
 *
 * Conventions
 *  - all reals are IEEE double (`real(rkx)`, rkx = rk8,
 *    Share/mod_realkinds.F90:42-66); integers are 32-bit (`ik4`).
 *  - arrays cross the boundary as a pointer to the first element plus explicit
 *    Fortran bounds lb[3]/ub[3] (inclusive, GLOBAL model indices), column
 *    major (j,i,k) with j fastest — exactly `c_loc(a(lb1,lb2,lb3))` of a
 *    contiguous `getmem3d` array (Share/mod_memutil.F90).  2-D arrays pass
 *    lb[2]=ub[2]=1, 1-D vertical arrays pass lb[0]=lb[1]=ub[0]=ub[1]=1.
 *  - host arrays are borrowed for the duration of the call only.
 *  - every function returns 0 on success or a negative RCM_ERR_* code;
 *    rcm_dyn_last_error() gives the message (the Fortran shim turns non-zero
 *    into `call fatal(__FILE__,__LINE__,msg)`, Share/mod_message.F90:90).
 *  - one handle per MPI rank / GPU; calls are collective across ranks.
 */
#ifndef REGCM_DYN_H
#define REGCM_DYN_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RCM_OK 0
#define RCM_ERR_ARG (-1)   /* bad argument / unsupported option            */
#define RCM_ERR_CUDA (-2)  /* CUDA runtime failure or no device            */
#define RCM_ERR_CFL (-3)   /* Main/mod_sound.F90:650-665 'CFL violation'   */
#define RCM_ERR_NAN (-4)   /* Main/mod_tendency.F90:625-703 NaN detection  */
#define RCM_ERR_NCCL (-5)  /* halo exchange / communicator failure         */
#define RCM_ERR_BDY (-6)   /* boundary record needed but not staged        */
#define RCM_ERR_STATE (-7) /* call sequence error (e.g. step before init)  */

/* prognostic fields: members of atm1/atm2 (`atmstate_a/b`,
 * Main/mpplib/mod_regcm_types.F90:65-120; Main/mod_atm_interface.F90:617-672)
 * and the boundary structures xub,xvb,xtb,xqb,xppb,xwwb
 * (Main/mod_atm_interface.F90:544-573). */
enum rcm_field {
  RCM_F_U = 0,  /* dot   (jx  ,iy  ,kz)   p*·u  */
  RCM_F_V = 1,  /* dot   (jx  ,iy  ,kz)   p*·v  */
  RCM_F_T = 2,  /* cross (jx-1,iy-1,kz)   p*·T  */
  RCM_F_QV = 3, /* cross                  p*·qv */
  RCM_F_QC = 4, /* cross                  p*·qc */
  RCM_F_PP = 5, /* cross                  p*·p' */
  RCM_F_W = 6,  /* cross (jx-1,iy-1,kz+1) p*·w  */
  RCM_F_PS = 7, /* cross (jx-1,iy-1) 2-D: p* = ps - ptop [cb], sfs%psa/psb
                 * (idynamic = 1 only; Main/mod_atm_interface.F90 `sfs`).  As a
                 * boundary field (xpsb, rcm_dyn_set_bdy) the value passed is the
                 * ICBC surface pressure in hPa; the library converts it like
                 * Main/mod_bdycod.F90:394 (`ps*0.1 - ptop`). */
  RCM_F_PAI = 8, /* cross (jx-1,iy-1,kz): Exner function Pi of the MOLOCH core
                  * (mo_atm%pai, Main/mod_atm_interface.F90:575-616); state only */
  RCM_F_COUNT = 9,
  /* Turbulent kinetic energy of the UW PBL (atm1%tke / atm2%tke, ibltyp = 2; idynamic 1, 2): cross
   * (jx-1,iy-1,kz+1), NOT coupled with p* [m2 s-2].  State and physics tendency only (no boundary record: the rim follows
   * the inflow/outflow rule of Main/mod_bdycod.F90:2412-2538).  Transport = hadv3d(ind = 1) on atm1%tke,
   * vadv3d(kz+1) on tke*psa, diffu_x3df(nuk) on atm2%tke, tketen = tkedyn*rpsa + tkephy, floor tkemin, RA filter
   * (Main/mod_tendency.F90:518-544,1414-1425,1544-1548). */
  RCM_F_TKE = 9,
  /* Tracers (atm1%chi / atm2%chi, ichem = 1; idynamic 1, 2): tracer n = 0..ntr-1 is field
   * RCM_F_CHI + n, cross (jx-1,iy-1,kz), p*·chi.  As a boundary field (chib0/chib1, che_init_bdy and chem_bdyin,
   * Main/chemlib/mod_che_bdyco.F90:118-420) the value passed is the mixing ratio of the CHBC record on model
   * levels; the library couples it with the boundary p* and forms chibt = (chib1 - chib0)/dtbdys as the reference
   * does.  The initial tracer state is the first record.  Transport = hadvtr, vadv4d(ind = 2), nudge_chi with the
   * time-dependent boundary values (ichebdy = 1), diffu_x4d, the negative fix with factor 0.1 and the RAW filter
   * (Main/mod_tendency.F90:548-586,1393-1411,1502-1511,1539-1543). */
  RCM_F_CHI = 16
};
#define RCM_MAXTR 8
/* idynamic = 3 (MOLOCH, Arakawa C grid, Main/mod_atm_interface.F90:575-616):
 * the state is single-level and uncoupled: RCM_F_U is mo_atm%u on
 * (jx, iy-1, kz), RCM_F_V is mo_atm%v on (jx-1, iy, kz), RCM_F_W is mo_atm%w on
 * (jx-1, iy-1, kz+1), RCM_F_T/QV/QC/PAI on (jx-1, iy-1, kz), RCM_F_PS is the
 * diagnosed surface pressure sfs%psa [Pa]; the `level` argument is ignored.
 * Boundary fields are U, V, T, QV as read from the ICBC file and RCM_F_PS in
 * hPa (converted like Main/mod_bdycod.F90:529); Pi on the boundary is derived
 * by `paicompute` (Main/mod_bdycod.F90:5125-5155). */

/* static domain fields: members of `mddom` read by read_domain_info
 * (Main/mod_params.F90:1733) plus the sigma vector. */
enum rcm_static {
  RCM_S_HT = 0,     /* terrain height [m], cross grid stored on (jx,iy)   */
  RCM_S_MSFX = 1,   /* map factor at cross points                         */
  RCM_S_MSFD = 2,   /* map factor at dot points                           */
  RCM_S_CORIOL = 3, /* Coriolis parameter at dot points [1/s]             */
  RCM_S_XLAT = 4,   /* latitude  of cross points [deg]                    */
  RCM_S_XLON = 5,   /* longitude of cross points [deg]                    */
  RCM_S_DLAT = 6,   /* latitude  of dot points [deg]                      */
  RCM_S_SIGMA = 7,  /* full sigma levels, kz+1 values (k = 1 is the top)  */
  /* idynamic = 3 (MOLOCH) */
  RCM_S_MSFU = 8,   /* map factor at u points (jx,iy-1)                   */
  RCM_S_MSFV = 9,   /* map factor at v points (jx-1,iy)                   */
  RCM_S_ULAT = 10,  /* latitude of u points [deg]                         */
  RCM_S_VLAT = 11,  /* latitude of v points [deg]                         */
  RCM_S_COUNT = 12
};

/* time levels of the leapfrog state (atm1 = "a" = tau, atm2 = "b" = tau-1) */
enum rcm_level { RCM_LEVEL_A = 0, RCM_LEVEL_B = 1 };

/* boundary-record slots of v3dbound (b0, b1); RCM_BDY_NEXT stages the record
 * that the state part of `bdyin` (Main/mod_bdycod.F90:669-690,772-831) will
 * move into b1 when the boundary alarm fires. */
enum rcm_bdy_slot { RCM_BDY_B0 = 0, RCM_BDY_B1 = 1, RCM_BDY_NEXT = 2 };

/* intermediate arrays that can be read back after a step for operator-level
 * parity checks (debug taps; no reference equivalent other than the module
 * variables named in the comments). */
enum rcm_work {
  RCM_W_CR = 0,    /* mdv%cr     Main/mod_tendency.F90:1179-1191          */
  RCM_W_QDOT = 1,  /* qdot       Main/mod_tendency.F90:1167-1178  (kz+1)  */
  RCM_W_XKC = 2,   /* xkc before the rdxsq*p* scaling mod_diffusion:212-230 */
  RCM_W_TH = 3,    /* th         Main/mod_tendency.F90:1350-1353          */
  RCM_W_RHO1 = 4,  /* atm1%rho   Main/mod_tendency.F90:1057-1058          */
  RCM_W_PR1 = 5,   /* atm1%pr                                             */
  RCM_W_UTEN = 6,  /* aten%u(:,:,:,pc_total) as handed to sound (x dts)   */
  RCM_W_VTEN = 7,
  RCM_W_PPTEN = 8,
  RCM_W_WTEN = 9,  /* (kz+1) */
  /* idynamic = 1 */
  RCM_W_OMEGA = 10, /* omega      Main/mod_tendency.F90:1199-1215  [cb/s]   */
  RCM_W_PHI = 11,   /* phi        Main/mod_tendency.F90:2069-2097           */
  RCM_W_PTEN = 12,  /* pten (2-D) Main/mod_tendency.F90:1131-1148, before nudging */
  /* idynamic = 3 */
  RCM_W_MO_P = 13,     /* mo_atm%p      Main/mod_moloch.F90:368-375 [Pa]      */
  RCM_W_MO_TVIRT = 14, /* mo_atm%tvirt  Main/mod_moloch.F90:321-327            */
  RCM_W_MO_FMZ = 15,   /* mo_atm%fmz    Main/mod_params.F90:2816-2824          */
  RCM_W_MO_FMZF = 16,  /* mo_atm%fmzf   (kz+1) :2827-2837                      */
  RCM_W_MO_ZETA = 17,  /* mo_atm%zeta   :2816-2824                             */
  RCM_W_MO_S = 18,     /* s (kz+1), generalized vertical velocity left by the
                        * last sound() of the step, Main/mod_moloch.F90:1022-1038 */
  /* idynamic = 1, 2: atmc%qx(:,:,:,iqv|iqc), the unfiltered leapfrog forecast AFTER the negative-value repair of
   * Main/mod_tendency.F90:381-393; defined on a step that repaired at least one point (or follows one) */
  RCM_W_CQV = 19,
  RCM_W_CQC = 20,
  RCM_W_COUNT = 21
};

/* the slices `mkslice` prepares for the physical parametrizations (type `slice`, Main/mpplib/mod_regcm_types.F90;
 * Main/mod_slice.F90:167-308, idynamic = 1, 2), cross grid, level-b (tau-1) state decoupled from p*.  3-D with kz
 * levels unless noted. */
enum rcm_slice {
  RCM_SL_TB3D = 0,   /* atms%tb3d    T [K]                                        :190 */
  RCM_SL_QVB3D = 1,  /* atms%qxb3d(iqv), floored at minqq                         :191 */
  RCM_SL_QCB3D = 2,  /* atms%qxb3d(iqc), floored at 0                             :194 */
  RCM_SL_UBX3D = 3,  /* atms%ubx3d   u on cross points [m/s]                      :180-183 */
  RCM_SL_VBX3D = 4,  /* atms%vbx3d                                                :184-187 */
  RCM_SL_PB3D = 5,   /* atms%pb3d    pressure at half levels [Pa]                 :211-217,233-235 */
  RCM_SL_PF3D = 6,   /* atms%pf3d    pressure at full levels [Pa], kz+1 levels    :221-231,239-241 */
  RCM_SL_PS2D = 7,   /* atms%ps2d    surface pressure [Pa], 2-D                   :218-220,236-238 */
  RCM_SL_RHOB3D = 8, /* atms%rhob3d  density [kg/m3]                              :253 */
  RCM_SL_TH3D = 9,   /* atms%th3d    potential temperature                        :248-251 */
  RCM_SL_TV3D = 10,  /* atms%tv3d    virtual temperature                          :202-205 */
  RCM_SL_TP3D = 11,  /* atms%tp3d    T brought to the surface pressure            :254-255 */
  RCM_SL_WPX3D = 12, /* atms%wpx3d   omega [Pa/s], interior cross points          :258-261,267-269 */
  RCM_SL_WB3D = 13,  /* atms%wb3d    w [m/s], kz+1 levels                         :262-264,270-276 */
  RCM_SL_ZQ = 14,    /* atms%zq      height of full levels [m], kz+1 levels (idynamic = 2: atm0%zf)  :283-292 */
  RCM_SL_ZA = 15,    /* atms%za      height of half levels (idynamic = 2: atm0%z) :293-295 */
  RCM_SL_DZQ = 16,   /* atms%dzq     layer thickness (idynamic = 2: atm0%dzf)     :296-298 */
  RCM_SL_QSB3D = 17, /* atms%qsb3d   saturation mixing ratio (pfwsat)             :301-303 */
  RCM_SL_RHB3D = 18, /* atms%rhb3d   relative humidity, clamped to [rhmin, rhmax] :304-307 */
  RCM_SL_COUNT = 19
};

/* vertical-mode tables of the split-explicit scheme (idynamic = 1), computed
 * once by `spinit`/`vmodes` (Main/mod_split.F90:75-175, Main/mod_vmodes.F90:86-594).
 * Column-major like the Fortran arrays. */
enum rcm_table {
  RCM_T_HBAR = 0,   /* hbar(kz)        equivalent depths, descending         */
  RCM_T_ZMATX = 1,  /* zmatx(kz,kz)    after the `fac` scaling of spinit     */
  RCM_T_ZMATXR = 2, /* zmatxr(kz,kz)                                         */
  RCM_T_TAU = 3,    /* tau(kz,kz)      as left by spinit (rows 1..nsplit)    */
  RCM_T_VARPA1 = 4, /* varpa1(kz,kz+1) as left by spinit                     */
  RCM_T_AM = 5,     /* am(kz,nsplit)                                         */
  RCM_T_AN = 6,     /* an(nsplit)                                            */
  RCM_T_TAU0 = 7,   /* tau(kz,kz) of vmodes (the matrix whose eigenproblem is solved) */
  RCM_T_COUNT = 8
};

/* Flat copy of the namelist scalars the dyn-core reads
 * (Share/mod_dynparam.F90:453-476, Main/mod_params.F90:87-118,190-300,645-665)
 * plus the decomposition (Main/mpplib/mod_mppparam.F90:1053-1401). */
typedef struct rcm_cfg {
  int32_t struct_size; /* = sizeof(rcm_cfg), ABI check                     */
  /* &dimparam */
  int32_t jx, iy, kz;
  int32_t nqx;      /* number of water species carried (2: qv,qc)          */
  /* &coreparam */
  int32_t idynamic; /* 1 hydrostatic, 2 non-hydrostatic, 3 MOLOCH          */
  /* &physicsparam (dynamics-relevant switches only) */
  int32_t iboudy;   /* 0 fixed, 1 linear relax., 5 exponential relax., 4 sponge + inflow/outflow;
                     * idynamic 1, 2 also: 2 time-dependent without relaxation, 3 = 2 + inflow/outflow
                     * (Main/mod_tendency.F90:909-983,1462-1500; Main/mod_bdycod.F90:237-250,1809-1947,2591-3122) */
  int32_t idiffu;   /* 1 = 4th order with the 2nd-order boundary ring, 2 = nine-point operator on every interior point
                     * (Main/mod_diffusion.F90:266-411,523-601,658-735,792-894); 3 = flux-limited 6th order as this snapshot
                     * of the reference has it: on the single column j = jdi2 / jci2 its loops run over (:421,611,745,902),
                     * one rank (the bounds are per tile: under MPI the reference's result depends on the decomposition),
                     * idynamic 1, 2 */
  int32_t ipptls;   /* >0: hydrometeors advected, water loading in w       */
  /* &nonhydroparam */
  int32_t ifupr;    /* upper radiative boundary condition                  */
  int32_t ifrayd;   /* Rayleigh damping at the model top                   */
  int32_t rayndamp;
  /* &boundaryparam */
  int32_t nspgx, nspgd;
  /* &globdatparam */
  int32_t ibdyfrq;  /* hours between boundary records                      */
  /* &dynparam */
  int32_t upstream_mode, stability_enhance, diffu_hgtf;
  /* decomposition: rank layout on the (jx,iy) plane, set_nproc rules      */
  int32_t nproc, myid;
  int32_t cpus_j, cpus_i; /* 0,0 = choose like set_nproc                   */
  int32_t device;         /* CUDA device ordinal, -1 = myid % device count */
  /* &hydroparam (idynamic = 1; Main/mod_params.F90:113,283-284) */
  int32_t nsplit;         /* number of split-explicit vertical modes (2)   */
  int32_t lstand;         /* 1: linearise vmodes about the standard atmosphere */
  /* idynamic = 3, &dynparam / &molochparam (Main/mod_params.F90:296-300,
   * Share/mod_dynparam.F90:189-190) */
  int32_t mo_nadv;        /* advection sub-steps per step (3)              */
  int32_t mo_nsound;      /* sound sub-steps per advection step (6)        */
  int32_t mo_nzfilt;      /* levels of the top sponge in the w solve (3)   */
  int32_t mo_rotllr;      /* 1: iproj = 'ROTLLR' (rotated lat-lon) branches of sound / wafone /
                           * compute_moloch_static (Main/mod_moloch.F90:773,932,1183; Main/mod_params.F90:2766) */
  int32_t ntr;            /* number of transported tracers, 0..RCM_MAXTR (ichem = 1 with ntr species; 0 = ichem 0) */
  /* &physicsparam / &uwparam, the part the dynamics reads (Main/mod_tendency.F90:148-156; Main/mod_params.F90:476-480) */
  int32_t ibltyp;         /* 2: the UW PBL is on, TKE (RCM_F_TKE) is carried and transported; anything else: no TKE     */
  int32_t iuwvadv;        /* 1 (with ibltyp = 2): vadv4d ind = 3 for qc and the tracers, which reshapes the vertical flux
                           * around the PBL top kpbl handed in with rcm_dyn_set_kpbl (Main/mod_advection.F90:915-957)  */
  int32_t ipgf;           /* idynamic = 1: 0 = pressure gradient force from p*.Tv (default); 1 = from the deviation of Tv from
                           * the reference profile t00pg (p/p00pg)^pgfaa1 (Main/mod_tendency.F90:1893-1964,2039-2068) */
  int32_t isladvec;       /* 1: semi-Lagrangian horizontal advection of qv, qc and the tracers (idynamic 1, 2;
                           * Main/mod_sladvection.F90).  On a decomposed domain the tiles get four ghost rows and the
                           * level-b species travel three rows wide before the operator (the departure stencil's reach) */
  int32_t iqmsl;          /* 1: quasi-monotonic limiter of the semi-Lagrangian interpolation (default 1)              */
  /* &geoparam */
  double ds;    /* grid spacing [km]                                       */
  double ptop;  /* model top pressure [cb]                                 */
  /* &timeparam */
  double dt;    /* dtsec [s]                                               */
  /* &dynparam */
  double gnu1, gnu2, ckh, adyndif, uoffc, t_extrema, q_rel_extrema;
  /* &nonhydroparam */
  double nhbet, nhxkd, rayalpha0, rayhd;
  /* &boundaryparam */
  double high_nudge, medium_nudge, low_nudge, bdy_nm, bdy_dm;
  /* &referenceatm + DOMAIN-file base state surface temperature */
  double base_state_pressure, logp_lrate, base_state_ts0;
  double mo_anu2;  /* divergence-damping coefficient of filt3d (0.6)      */
  double mo_a0, mo_b0; /* decay / stretching of the zita coordinate (0, 0.5) */
  double uw_nuk;       /* &uwparam nuk: factor of the horizontal diffusion of TKE (5.0)                              */
  double uw_tkemin;    /* floor of TKE, uwtkemin of Main/pbllib/mod_pbl_uwtcm.F90:86 (1.0e-3)                        */
  double reserved_d[3];
} rcm_cfg;

/* health metrics the reference prints (Main/mod_sound.F90:634-646) */
typedef struct rcm_diag {
  int64_t lcount;     /* rcmtimer%lcount                                   */
  double xbctime;     /* seconds since boundary record b0                  */
  double dt;          /* current leapfrog dt (dtsec, then 2*dtsec)         */
  double max_cfl;     /* max over sub-steps of last step                   */
  int32_t istep;      /* acoustic sub-steps of last step                   */
  int32_t nan_flag;
  int64_t neg_q_points; /* points repaired by the negative-q fix, last step */
  /* idynamic = 1: Bleck (1977) noise parameters of the last step, summed over
   * the interior (Main/mod_tendency.F90:1449-1459; the reference divides by
   * the number of points when it prints them, :715-716) */
  double ptntot, pt2tot;
  /* idynamic = 3: extremes of ps over the interior [Pa], the reference's
   * per-report print (Main/mod_moloch.F90:437-460) */
  double ps_max, ps_min;
  int64_t neg_chi_points; /* tracer values repaired by the negative fix of Main/mod_tendency.F90:570-582, last step,
                           * this rank */
} rcm_diag;

typedef struct rcm_dyn rcm_dyn; /* opaque */

/* Fill cfg with the reference defaults (Main/mod_params.F90:190-300,645-665;
 * Share/mod_dynparam.F90:147-154,502-508) for the given core. */
int rcm_dyn_default_cfg(rcm_cfg* cfg, int idynamic);

/* replaces: allocate_mod_atm_interface / _tend / _bdycon / _sound /
 * _diffusion, setup_model_indexes, setup_boundaries, set_nproc
 * (Main/mod_params.F90:1554-1590; Main/mod_atm_interface.F90:181-542). */
int rcm_dyn_create(const rcm_cfg* cfg, rcm_dyn** handle);

/* replaces: read_domain_info → mddom%{ht,msfx,msfd,coriol,xlat,xlon,dlat},
 * sigma (Main/mod_params.F90:1733). */
int rcm_dyn_set_static(rcm_dyn* h, int static_id, const double* a,
                       const int32_t lb[3], const int32_t ub[3]);

/* replaces: read_icbc into x?b%b0 / x?b%b1 (Main/mod_bdycod.F90:325-369,
 * 452-496,712-760).  Values are DECOUPLED (u,v [m/s], T [K], qv [kg/kg],
 * p' [Pa], w [m/s]); the library couples them with p* like
 * `couple` (Main/mod_bdycod.F90:4938-4951). */
int rcm_dyn_set_bdy(rcm_dyn* h, int field, int slot, const double* a,
                    const int32_t lb[3], const int32_t ub[3]);

/* Asynchronous variant of rcm_dyn_set_bdy(…, RCM_BDY_NEXT, …): what a driver
 * does when it reads the next ICBC record ahead of time (the reference reads it
 * synchronously inside bdyin, Main/mod_bdycod.F90:669-690).  The copy is queued on
 * a private stream and overlaps rcm_dyn_step; the record change inside
 * rcm_dyn_step waits for it on the device.  `a` must stay valid (and should be
 * page-locked) until rcm_dyn_stage_wait returns or the record has been consumed. */
int rcm_dyn_stage_bdy_async(rcm_dyn* h, int field, const double* a, const int32_t lb[3], const int32_t ub[3]);
int rcm_dyn_stage_wait(rcm_dyn* h);

/* replaces: the static part of `param` (map-factor inversion, xmsf/dmsf,
 * hsigma/dsigma/twt/qcon, init_advection, setup_bdycon,
 * make_reference_atmosphere, compute_full_coriolis_coefficients —
 * Main/mod_params.F90:1981-2074,2208-2215,2613-2741), the state part of
 * `init_bdy` (Main/mod_bdycod.F90:393-435,522-559,625-640), `init`
 * (Main/mod_init.F90:96-178,878-903), initialize_diffusion, init_sound, and
 * the first `bdyval` (Main/mod_regcm_interface.F90:146). */
int rcm_dyn_init(rcm_dyn* h);

/* replaces: nsteps x { call tend ; [call bdyin] ; call bdyval }
 * (Main/mod_regcm_interface.F90:172-228; Main/mod_tendency.F90:212;
 * Main/mod_bdycod.F90:654,1109).  Synchronous: the state is observable with
 * rcm_dyn_get_state on return. */
int rcm_dyn_step(rcm_dyn* h, int nsteps);

/* Physics coupling boundary.  idynamic = 3 (MOLOCH): `moloch` calls mkslice and physical_parametrizations after the
 * sound / advection loops and before `boundary` (Main/mod_moloch.F90:393-426); rcm_dyn_step_pre_physics runs up to
 * there, rcm_dyn_get_slice returns the members of `atms` (aliases of mo_atm, Main/mod_slice.F90:82-100, and the products
 * of mkslice for idynamic = 3, :107-146; ZQ / ZA / DZQ = mo_atm%zetaf / zeta / dz), rcm_dyn_set_phys_tend takes
 * mo_atm%tten, qxten, uten, vten (RCM_F_T, QV, QC on the cross grid, U on (jx, iy-1), V on (jx-1, iy); NOT coupled with
 * anything), rcm_dyn_step_post_physics applies t, u, v, qx += dtsec * tendency with the floors of :1548-1557 (only
 * once tendencies have been set), then boundary, Rayleigh damping, bdyin / bdyval.
 * idynamic = 1, 2:  replaces: the place where `tend` hands over to the physical
 * parametrizations and takes their tendencies back (Main/mod_tendency.F90:271 `call physical_parametrizations`,
 * :1682-1820; slices: Main/mod_slice.F90:167-308; sums: Main/mod_tendency.F90:285-314,331-335,404-411).
 *   rcm_dyn_step_pre_physics   everything of `tend` the parametrizations depend on (surface_pressures, decouple,
 *                              compute_omega, the inputs of mkslice, calc_coeff); synchronous.
 *   rcm_dyn_get_slice          one member of `atms` (enum rcm_slice), computed on request from the level-b state,
 *                              on the cross grid with the usual lb/ub window.  Valid at any time between steps; omega
 *                              (WPX3D, WB3D) of the hydrostatic core only between the pre- and post-physics calls.
 *   rcm_dyn_set_phys_tend      aten(:,:,:,pc_physic) of one prognostic (RCM_F_U, V: dot grid; T, QV, QC, PP, W, TKE: cross
 *                              grid; coupled with p* like the reference's).  Stays in effect — also for
 *                              rcm_dyn_step — until replaced or cleared; zero until first set.
 *   rcm_dyn_step_post_physics  the rest of `tend` (advection ... diffusion, the sums, forecast, time filters,
 *                              sound / splitf), then bdyin / bdyval as in rcm_dyn_step.
 * pre_physics; post_physics with no tendencies set equals rcm_dyn_step(h, 1) bit for bit. */
int rcm_dyn_step_pre_physics(rcm_dyn* h);
int rcm_dyn_get_slice(rcm_dyn* h, int slice_id, double* a, const int32_t lb[3], const int32_t ub[3]);
int rcm_dyn_set_phys_tend(rcm_dyn* h, int field, const double* a, const int32_t lb[3], const int32_t ub[3]);
int rcm_dyn_clear_phys_tend(rcm_dyn* h);
int rcm_dyn_step_post_physics(rcm_dyn* h);

/* UW-PBL coupling of the transport (ibltyp = 2).  replaces: the `kpbl` array the PBL scheme leaves for vadv4d
 * (Main/mod_atm_interface.F90:136,1135; Main/mod_advection.F90:66,92,921-956) and aten%tke(:,:,:,pc_physic)
 * (rcm_dyn_set_phys_tend with RCM_F_TKE, cross grid, kz+1 levels, uncoupled as in Main/mod_tendency.F90:530-531).
 * kpbl: int32 on the cross interior (window lb/ub as elsewhere, two indices); zero until first set, as getmem2d leaves it
 * — then ind = 3 is the plain twt interpolation.  A value > kz is an error (the reference calls `fatal`). */
int rcm_dyn_set_kpbl(rcm_dyn* h, const int32_t* kpbl, const int32_t lb[2], const int32_t ub[2]);

/* Asynchronous form of rcm_dyn_step for drivers that overlap the time loop with their own work (the reference's
 * RCM_run is synchronous: Main/mod_regcm_interface.F90:172-228; its output and boundary input block the loop).
 * rcm_dyn_step_async enqueues nsteps steps on the library's stream and returns without waiting for them (it still
 * waits once per simulated day for the tmask update of Main/mod_sound.F90:500-547); rcm_dyn_sync waits for
 * everything enqueued and returns the verdict of the batch, identical on all ranks (RCM_ERR_CFL / RCM_ERR_NAN /
 * RCM_ERR_NCCL).  rcm_dyn_step(h, n) == rcm_dyn_step_async(h, n); rcm_dyn_sync(h).  While steps are in flight
 * rcm_dyn_diag does not block: it reports the newest step whose scalars have already arrived in page-locked host
 * memory (one 32-byte D2H copy follows every step; this rank's values).  Every other accessor (get/set_state,
 * get_work, set_bdy) synchronises first. */
int rcm_dyn_step_async(rcm_dyn* h, int nsteps);
int rcm_dyn_sync(rcm_dyn* h);

/* Asynchronous state export — what `output` / write_savefile do between steps (Main/mod_regcm_interface.F90:211-222),
 * without stopping the time loop: a device-side snapshot of the field is taken in stream order (after the steps
 * already enqueued) and copied into `a` on a private stream while later steps run.  `a` should be page-locked and
 * must stay valid until rcm_dyn_get_state_wait returns.  Same field / level / bounds conventions as
 * rcm_dyn_get_state.  The first export of a field allocates its snapshot buffer (one 3-D field). */
int rcm_dyn_get_state_async(rcm_dyn* h, int field, int level, double* a,
                            const int32_t lb[3], const int32_t ub[3]);
int rcm_dyn_get_state_wait(rcm_dyn* h);

/* state export/import = what write_savefile/read_savefile carry for the
 * dynamics (Main/mod_savefile.F90:564-1264): COUPLED atm1/atm2 fields. */
int rcm_dyn_get_state(rcm_dyn* h, int field, int level, double* a,
                      const int32_t lb[3], const int32_t ub[3]);
int rcm_dyn_set_state(rcm_dyn* h, int field, int level, const double* a,
                      const int32_t lb[3], const int32_t ub[3]);

/* debug tap, see enum rcm_work */
int rcm_dyn_get_work(rcm_dyn* h, int work_id, double* a, const int32_t lb[3],
                     const int32_t ub[3]);

/* replaces: the public tables of mod_vmodes / mod_split (Main/mod_vmodes.F90:39-52,
 * Main/mod_split.F90:40-42).  Copies min(n, size) doubles, returns the table size
 * (or a negative RCM_ERR_*). */
int rcm_dyn_get_table(rcm_dyn* h, int table_id, double* out, int n);

/* Restart (the SAV path): `write_savefile` / `read_savefile` (Main/mod_savefile.F90:564-1264) keep their
 * NetCDF layer on the host and obtain / hand back the dynamical state through rcm_dyn_get_state / rcm_dyn_set_state
 * (atm1_u ... atm2_qx, atm1_w/pp, psa, psb; atm_u ... atm_pai, ps for idynamic = 3; the tracers as RCM_F_CHI + n).
 * What the SAV file holds beside the prognostics and this library owns:
 *   RCM_R_DSTOR, RCM_R_HSTOR  dstor, hstor of the split-explicit scheme (idynamic = 1; cross points, nsplit planes;
 *                             dstor_io / hstor_io, mod_savefile.F90:326-327)
 *   RCM_R_TMASK               the 13 x 13 mask of the upper radiative boundary condition (idynamic = 2, ifupr = 1;
 *                             `tmask`, mod_savefile.F90:30), passed as a (13,13,1) array
 * A restart is: rcm_dyn_create; rcm_dyn_set_static; rcm_dyn_set_bdy with the boundary records that bracket the restart
 * date (as init_bdy reads them again, Main/mod_bdycod.F90:283-330); rcm_dyn_init; rcm_dyn_set_state of every
 * prognostic at both levels; rcm_dyn_set_restart of the arrays above; rcm_dyn_set_clock with the step count and the
 * seconds since boundary record b0 that rcm_dyn_diag reported when the state was saved (ktau / xbctime of the SAV
 * file).  rcm_dyn_set_clock also refreshes the ghost rows of the restored state on a decomposed domain (collective).
 * The continued run equals the uninterrupted one bit for bit (tests/test_restart.py). */
enum rcm_restart { RCM_R_DSTOR = 0, RCM_R_HSTOR = 1, RCM_R_TMASK = 2 };
int rcm_dyn_get_restart(rcm_dyn* h, int id, double* a, const int32_t lb[3], const int32_t ub[3]);
int rcm_dyn_set_restart(rcm_dyn* h, int id, const double* a, const int32_t lb[3], const int32_t ub[3]);
int rcm_dyn_set_clock(rcm_dyn* h, int64_t lcount, double xbctime);

int rcm_dyn_diag(rcm_dyn* h, rcm_diag* out);
int rcm_dyn_last_error(rcm_dyn* h, char* buf, int len);

/* number of CUDA kernels launched by this handle so far, and the device time
 * of the last rcm_dyn_step call in milliseconds (CUDA events on the library's
 * own stream). */
int64_t rcm_dyn_launch_count(rcm_dyn* h);
double rcm_dyn_last_step_ms(rcm_dyn* h);

/* replaces: the tile extents set_nproc / setup_model_indexes give a rank
 * (global_dot_jstart..jend, istart..iend; Main/mpplib/mod_mppparam.F90:1302-1360):
 * out = {jde1, jde2, ide1, ide2} in global dot indices. */
int rcm_dyn_get_tile(rcm_dyn* h, int32_t out[4]);

/* per-kernel device timing (no reference equivalent; the reference has no
 * timers in production builds).  While enabled, every kernel
 * launch of rcm_dyn_step is bracketed by CUDA events on the library's stream;
 * rcm_dyn_profile_get returns the accumulated milliseconds and launch count
 * of kernel `kid` in [0, rcm_dyn_profile_kernels()). */
int rcm_dyn_profile(rcm_dyn* h, int enable);
int rcm_dyn_profile_get(rcm_dyn* h, int kid, double* ms, int64_t* count,
                        char* name, int len);
int rcm_dyn_profile_kernels(void);

/* replaces: the communicator set-up of set_nproc / mpi_cart_create
 * (Main/mpplib/mod_mppparam.F90:1053-1401) for nproc > 1.  Rank 0 obtains a
 * 128-byte NCCL unique id with rcm_dyn_comm_unique_id, the caller broadcasts
 * it with whatever it already has (the Fortran driver: `call bcast`, i.e.
 * MPI_Bcast; the Python harness: torch.distributed), then EVERY rank calls
 * rcm_dyn_comm_init after rcm_dyn_create and before rcm_dyn_init.  Halo
 * exchanges (the `exchange*` family, mod_mppparam.F90:209-271) then run inside
 * rcm_dyn_init / rcm_dyn_step as NCCL send/recv on packed edge buffers. */
int rcm_dyn_comm_unique_id(void* out128, int out_bytes);
int rcm_dyn_comm_init(rcm_dyn* h, const void* nccl_unique_id, int id_bytes);

/* NO COUNTERPART IN THE REFERENCE — added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * The negative-moisture repair of Main/mod_tendency.F90:381-393 (the sweep rcm_dyn_step applies to atmc%qx,
 * csrc/k_negq.cu) as a stand-alone, handle-less operator that allocates nothing: its scratch (lists of negative
 * points, fill counts) is a file-scope static __device__ array (csrc/k_negq_static.cu) of RCM_NQS_MAX_CELLS list
 * entries + RCM_NQS_MAX_PLANES counters = 4 MiB + 4 KiB, held by every GPU that loads the library until exit.
 *   q_in, q_out : DEVICE pointers (distinct), nk planes of (nj, ni) doubles, j fastest, no padding.  q_out = q_in
 *                 with every negative point of the plane interiors (j = 2..nj-1, i = 2..ni-1, 1-based) replaced by
 *                 0.01 * sum|q(j-1:j+1,i-1:i+1)| / 9 in the reference's sequential sweep order (j, then i, ascending);
 *                 the outermost ring is copied unchanged (boundary / ghost values as the sweep should see them).
 *   n_repaired  : DEVICE pointer or NULL; receives the number of repaired points.
 *   stream      : a cudaStream_t (NULL = the default stream); the call is asynchronous on it.
 * Returns RCM_ERR_ARG when nj*ni*nk > RCM_NQS_MAX_CELLS or nk > RCM_NQS_MAX_PLANES (or nj, ni < 3, nk < 1, NULL or
 * aliased arrays), RCM_ERR_CUDA on a runtime failure.  Calls no cudaMalloc / cudaFree / allocating library routine.
 * The first call on a device looks up the scratch address (cudaGetSymbolAddress): make it outside graph capture.
 * All calls share the scratch: calls that overlap on the GPU (other streams, concurrent graph launches) race. */
#define RCM_NQS_MAX_CELLS 1048576
#define RCM_NQS_MAX_PLANES 1024
int rcm_dyn_negq_repair_static(const double* q_in, double* q_out, int32_t nj, int32_t ni, int32_t nk,
                               unsigned long long* n_repaired, void* stream);

/* replaces: memory_destroy (Main/mod_regcm_interface.F90:264). */
int rcm_dyn_destroy(rcm_dyn* h);

#ifdef __cplusplus
}
#endif
#endif /* REGCM_DYN_H */
