// ORACLE — TEST INFRASTRUCTURE ONLY.
// CPU (double precision) restatement of the reference's dyn-core state and
// step for idynamic = 2 (non-hydrostatic MM5 core), 1 (hydrostatic split-explicit
// core) and 3 (MOLOCH), single rank.  See oracle/README.md.
// "parity unpinned": the reference ships no golden vectors for this path
// (SURVEY.md §8c); this restatement IS the pin.
//
// Data model follows Main/mod_atm_interface.F90:39-60 (atm1, atm2, atmc, aten,
// atmx, atm0, atms, sfs, mddom, mdv, x?b) and Main/mpplib/mod_regcm_types.F90.
#pragma once
#include <cstdint>
#include <string>
#include <functional>
#include <vector>

#include "../include/regcm_dyn.h"
#include "arr.hpp"
#include "consts.hpp"

namespace orc {

struct V3dBound {  // type v3dbound, Main/mpplib/mod_regcm_types.F90
  A3 b0, b1, bt;
};

struct V2dBound {  // type v2dbound
  A2 b0, b1, bt;
};

struct BoundArea {  // type bound_area, Main/mod_atm_interface.F90:383-542
  int nsp = 0, ns = 0, nn = 0, nw = 0, ne = 0;
  bool havebound = false;
  I2 ibnd, bsouth, bnorth, bwest, beast;
};

struct Oracle {
  rcm_cfg cfg{};
  std::string err;

  // ---- dimensions / run switches (Share/mod_dynparam.F90, mod_runparams) ----
  int jx = 0, iy = 0, kz = 0, kzp1 = 0, nqx = 2;
  int iqv = 1, iqc = 2, iqfrst = 2, iqlst = 2;
  int idif = 2;
  int iqxvadv = 1, itrvadv = 2;  // Main/mod_tendency.F90:148-154
  double dx = 0, dx2 = 0, dx4 = 0, dx8 = 0, dx16 = 0, dxsq = 0, rdx = 0,
         rdxsq = 0;
  double dtsec = 0, dt = 0, dt2 = 0;
  double ptop = 0;
  double xbctime = 0, dtbdys = 0, rdtbdy = 0;
  int64_t lcount = 0;          // rcmtimer%lcount
  int64_t model_time = 0;      // rcmtimer%model_internal_time [s]
  // alarm_day (86400 s) and alarm_in_bdy (ibdyfrq h): mod_timer.F90:277-337
  int64_t day_lastact = 0, day_now = 0;
  int64_t bdy_lastact = 0, bdy_now = -1;
  bool have_next = false;

  // ---- index families (Main/mod_atm_interface.F90:181-381), single rank ----
  int jde1, jde2, ide1, ide2, jdi1, jdi2, idi1, idi2, jdii1, jdii2, idii1,
      idii2;
  int jce1, jce2, ice1, ice2, jci1, jci2, ici1, ici2, jcii1, jcii2, icii1,
      icii2;
  int jcross1, jcross2, icross1, icross2;
  bool has_bdyleft = true, has_bdyright = true, has_bdytop = true,
       has_bdybottom = true;

  // ---- vertical coordinate (Main/mod_params.F90:2005-2010,2208-2215) ----
  A1 sigma, hsigma, dsigma, qcon, dds, xds;
  A2 twt;  // twt(k,1:2)

  // ---- mddom ----
  A2 ht, msfx, msfd, coriol, xlat, xlon, dlat, xmsf, dmsf;
  A2 ef, ddx, ddy, dmdx, dmdy, ex, crx, cry;
  A2 dpsdxm, dpsdym;

  // ---- sfs ----
  A2 psa, psb, psc, psdota, psdotb;
  A2 rpsa, rpsb, rpsda;  // Main/mod_tendency.F90:63-64

  // ---- atm0 reference atmosphere ----
  A2 a0_ps, a0_psdot;
  A3 a0_t, a0_pr, a0_rho, a0_z, a0_tf, a0_pf, a0_rhof, a0_zf, a0_dzf,
      a0_dprddx, a0_dprddy;

  // ---- atm1 (a), atm2 (b), atmc (c) ----
  A3 a1_u, a1_v, a1_t, a1_pp, a1_w, a1_pr, a1_rho;
  A4 a1_qx;
  A3 a2_u, a2_v, a2_t, a2_pp, a2_w, a2_pr;
  A4 a2_qx;
  A3 ac_u, ac_v, ac_t, ac_pp, ac_w;
  A4 ac_qx;

  // ---- atmx decoupled ----
  A3 ax_uc, ax_vc, ax_umc, ax_vmc, ax_ud, ax_vd, ax_umd, ax_vmd, ax_t, ax_tv,
      ax_pr, ax_pp, ax_w;
  A4 ax_qx;

  // ---- aten: three components pc_total=1, pc_dynamic=2, pc_physic=3 ----
  A3 uten, vten, tten, ppten, wten, udyn, vdyn, tdyn, ppdyn, wdyn;
  A4 qxten, qxdyn;

  // ---- atms slices used by the dynamics (Main/mod_slice.F90:167-263) ----
  A3 ubd3d, vbd3d, tb3d, ppb3d, wb3d;
  A4 qxb3d;

  // ---- misc work arrays ----
  A3 qdot, cr, omega;                  // mod_atm_interface qdot, mdv%cr
  A3 ucc, vcc, th, tha, thten;         // mod_tendency.F90:57-62
  A3 fg, uavg1, uavg2, vavg1, vavg2;   // mod_advection.F90:71-73
  A3 xkc, xkd, xkcf;                   // mod_diffusion
  A3 xkc_raw;  // tap: xkc before the rdxsq*p* scaling (tests only)
  A2 hgfact;
  double dydc = 0, xkhmax = 0;
  double ul = 0;                       // mod_advection.F90:106

  // ---- boundary (mod_bdycod) ----
  V3dBound xub, xvb, xtb, xqb, xppb, xwwb;
  A3 nx_u, nx_v, nx_t, nx_q, nx_pp, nx_ww;  // staged next record (decoupled)
  A2 xpsb_b0;
  A2 psdot;                            // bdycon:psdot
  A2 wue, wui, eue, eui, wve, wvi, eve, evi;  // (i,k)
  A2 sue, sui, nue, nui, sve, svi, nve, nvi;  // (j,k)
  A1 fcx, gcx, fcd, gcd;
  A2 hefc, hegc, hefd, hegd;
  A3 fg1, fg2;
  double fnudge = 0, gnudge = 0;
  BoundArea ba_cr, ba_dt;

  // ---- sound (mod_sound.F90:41-78) ----
  A3 s_aa, s_b, s_c, s_rhs, s_sigdot, s_wo, s_ca, s_g1, s_g2, s_ptend, s_pxup,
      s_pyvp, s_ucrs, s_vcrs, s_tk, s_cc, s_cdd, s_cj, s_pi, s_e, s_f;
  A2 s_astore, s_estore, s_estore_g, s_wpval, s_rpsb;
  A2 tmask;  // (-6:6,-6:6)
  double fi[13], fj[13], fk[7], fl[7];
  double xmsf_mean = 0;
  double bet = 0.4, xkd_snd = 0.1, cs = 0, bp = 0, bm = 0, bpxbm = 0,
         bpxbp = 0, dtsmax = 0, rnpts = 0;
  double xgamma = 0;
  double max_cfl_step = 0;
  int last_istep = 0;
  bool cfl_error = false;
  bool break_before_sound = false;  // test hook (oracle/taps.cpp)
  int64_t neg_q_points = 0;
  std::vector<int> rt_j, rt_i;  // first dot index of every tile along j / i (set_repair_tiles; empty = one domain)
  void repair_negative_qx();

  // ---- physics coupling boundary (oracle/physics.cpp) ----
  bool phys_on = false;
  A3 uphy, vphy, tphy, ppphy, wphy;     // aten(:,:,:,pc_physic)
  A4 qxphy;
  void pre_physics();
  void ensure_phys();
  void clear_phys();
  int set_phys(int field, const double* a, const int32_t lb[3], const int32_t ub[3]);
  int get_slice(int id, double* a, const int32_t lb[3], const int32_t ub[3]);
  inline double phy3(const A3& p, int j, int i, int k) const { return phys_on ? p(j, i, k) : d_zero; }
  inline double phy4(int j, int i, int k, int n) const { return phys_on ? qxphy(j, i, k, n) : d_zero; }

  // ---- hydrostatic core (idynamic = 1; oracle/hydro.cpp, oracle/split.cpp) ----
  // ---- UW-PBL coupling of the transport: TKE, vadv4d ind = 3 (oracle/tke.cpp) ----
  bool uw_on = false;
  double tkemin = 0;
  A3 tke_a, tke_b, tke_c, tkeps, tkedyn, tketen, tkephy;
  std::vector<int> kpb;   // (jci1:jci2, ici1:ici2)
  void allocate_tke();
  void tke_tendency();
  void bdyval_tke_copy();
  void bdyval_tke();
  void vadv_ind3(const std::function<double(int, int, int)>& f);
  int set_kpbl(const int32_t* k, const int32_t lb[2], const int32_t ub[2]);
  int tke_state(int level, double* a, const int32_t lb[3], const int32_t ub[3], bool in);
  // ---- semi-Lagrangian advection of moisture and tracers (isladvec = 1; oracle/sladvection.cpp) ----
  A3 sl_alf, sl_am2, sl_am1, sl_a0, sl_ap1, sl_bm2, sl_bm1, sl_b0, sl_bp1;
  std::vector<int> sl_xn, sl_xm1, sl_xm2, sl_xp1, sl_yn, sl_ym1, sl_ym2, sl_yp1;
  int trajcalc_x();
  int sl_rc = 0;   // RCM_ERR_CFL when a departure point lies more than one cell away (the reference calls fatal)
  void slhadv_x(const std::function<double&(int, int, int)>& ften, const std::function<double(int, int, int)>& var);
  void hdvg_x(const std::function<double&(int, int, int)>& ften, const std::function<double(int, int, int)>& var);
  // ---- tracers (oracle/tracers.cpp) ----
  int ntr = 0;
  std::vector<A3> chia, chib, chic, chix, chib3d, chidyn, chiten, nx_chi;   // atm1%chi, atm2%chi, atmc%chi, atmx%chi, ...
  std::vector<V3dBound> xchib;                                               // chib0, chib1, chibt (coupled)
  A2 cefc, cegc;                                                             // setup_che_bdycon
  int64_t neg_chi_points = 0;
  void allocate_tracers();
  void init_bdy_tracers();
  void init_state_tracers();
  void bdyin_tracers();
  void bdyval_tracers(double xt);
  void tracer_tendency();
  int tracer_state(int n, int level, double* a, const int32_t lb[3], const int32_t ub[3], bool in);
  int tracer_bdy(int n, int slot, const double* a, const int32_t lb[3], const int32_t ub[3]);

  bool hydro = false;
  A2 pten, pten_raw, rpsc, ptn2d, pt22d;          // mod_tendency.F90:63-66
  A3 phi, td, ttld, tvfac, pf3d, pb3d;        // mod_tendency.F90:57-62, atms%pf3d/pb3d
  V2dBound xpsb;                        // mod_atm_interface.F90 xpsb
  A2 nx_ps;                             // staged next ps record [hPa]
  double ptntot = 0, pt2tot = 0;
  // mod_vmodes.F90:46-52
  int nsplit = 2;
  double xps = 0, pd = 0;
  A2 vm_a0, zmatx, zmatxr, vm_tau, vm_tau0, varpa1, hydroc, hydros;
  A1 sigmah, tbarh, hbar;
  int vm_numerr = 0;
  // mod_split.F90:40-51, mod_atm_interface dstor/hstor
  A1 aam, an_, dtau;
  A2 am_, map_, uu, vv, xdelh;
  A3 uuu, vvv, ddsum, dhsum, work, dstor, hstor;
  A4 deld, delh;
  void allocate_hydro();
  void init_bdy_hydro();
  void init_state_hydro();
  void bdyin_hydro();
  void bdyval_ps();
  void surface_pressures_hydro();
  void decouple_hydro();
  void compute_omega_hydro();
  void mkslice_hydro();
  void new_pressure();
  void nudge2d(int ibdy, const A2& f, const V2dBound& bnd, A2& ften);
  void calc_coeff_hydro();
  void hadvuv_hydro(A3& ut, A3& vt, const A3& u, const A3& v);
  void vadv3d_ind1_hydro(A3& ften, const A3& f);
  void pressure_gradient_force();
  void filter_ra_2d(A2& phin, A2& phinm1, const A2& phinp1, double alpha);
  void tend_hydro();
  void vmodes();
  void spinit();
  void splitf();
  void spstep();
  int get_table(int id, double* out, int n);

  // ---- MOLOCH core (idynamic = 3; oracle/moloch.cpp) ----
  // mo_atm (Main/mod_atm_interface.F90:575-616), mod_moloch module arrays
  // (Main/mod_moloch.F90:52-91,115-157), mddom C-grid statics.
  bool moloch = false;
  int nadv = 3, nsound = 6;
  double mo_dz = 0, hzita = 0, rayzd = 0;
  double ps_max = 0, ps_min = 0;
  A1 zita, zitah, gzitak, gzitakh, ffilt, ak, bk;
  A2 msfu, msfv, ulat, vlat, coriou, coriov, hx, hy, mx2, rmu, rmv, p2d;
  A3 m_u, m_v, m_w, m_ux, m_vx, m_wx, m_t, m_pai, m_tetav, m_tvirt, m_p, m_rho,
      m_zeta, m_fmz, m_fmzf, m_deltaw, m_s, m_zdiv2, m_wwkw, m_wz, m_p0, m_ud,
      m_vd, m_zetau, m_zetav, m_zetaf, m_dz;
  // physics coupling of the MOLOCH core: tendencies of physical_parametrizations, uncoupled
  bool mo_mid = false, mo_phys_on = false;
  A3 m_tten, m_uten, m_vten;
  A4 m_qxten;
  void moloch_head();
  void moloch_tail();
  int mo_set_phys(int field, const double* a, const int32_t lb[3], const int32_t ub[3]);
  int mo_get_slice(int id, double* a, const int32_t lb[3], const int32_t ub[3]);
  A2 m_wfw, m_zpby, m_zpbw;
  A4 m_qx;
  V3dBound xpaib;
  BoundArea ba_ut, ba_vt;
  double gzita(double z) const;
  double gzitap(double z) const;
  double bzita(double z) const;
  double bzitap(double z) const;
  double md_fmz(double z, double geopot) const;
  int create_moloch();
  void compute_moloch_static();
  void init_bdy_moloch();
  void init_state_moloch();
  void init_moloch();
  void paicompute(const A2& ps, const A3& t, const A3& q, A3& pai);
  void bdyin_moloch();
  void bdyval_moloch();
  void moloch_step();
  void mo_sound(double dts);
  void mo_filt3d();
  void mo_advection(double dta);
  void wafone(A3& pp, double dta);
  void wafone4(A4& pp, int n, double dta);
  void uvstagtox();
  void xtouvstag();
  void wstagtox(const A3& w, A3& wx);
  void xtowstag(const A3& wx, A3& w);
  void xtoustag(const A3& ux, A3& u);
  void xtovstag(const A3& vx, A3& v);
  void mo_boundary();
  A1 wgtx, wgtd;   // sponge weights (Main/mod_bdycod.F90:237-250)
  void mosponge3d(A3& f, const V3dBound& bnd);
  void mosponge4d3d(A4& f, const V3dBound& bnd, int n);
  void mospongeuv(A3& fu, A3& fv, const V3dBound& bu, const V3dBound& bv);
  void monudge3d(int ibdy, A3& f, const V3dBound& bnd);
  void monudge4d3d(int ibdy, A4& f, const V3dBound& bnd, int n);
  void monudgeuv(int ibdy, A3& fu, A3& fv, const V3dBound& bu, const V3dBound& bv);
  void moraydamp(const A3& z, A3& var, const V3dBound& bnd, int j1, int j2, int i1, int i2, int k1, int k2);
  int mo_set_state(int field, const double* a, const int32_t lb[3], const int32_t ub[3]);
  int mo_get_state(int field, double* a, const int32_t lb[3], const int32_t ub[3]);
  int mo_get_work(int id, double* a, const int32_t lb[3], const int32_t ub[3]);

  // ---- life cycle (oracle/setup.cpp) ----
  int create(const rcm_cfg& c);
  int set_static(int id, const double* a, const int32_t lb[3],
                 const int32_t ub[3]);
  int set_bdy(int field, int slot, const double* a, const int32_t lb[3],
              const int32_t ub[3]);
  int init();
  int step(int n);
  int get_state(int field, int level, double* a, const int32_t lb[3],
                const int32_t ub[3]);
  int get_work(int id, double* a, const int32_t lb[3], const int32_t ub[3]);
  int set_state(int field, int level, const double* a, const int32_t lb[3],
                const int32_t ub[3]);

  // ---- setup pieces ----
  void setup_model_indexes();
  void setup_boundaries(bool ldotx, bool ldoty, BoundArea& ba);
  void allocate_all();
  void param_static();
  void make_reference_atmosphere();
  void compute_full_coriolis_coefficients();
  void init_advection();
  void setup_bdycon();
  void init_bdy();
  void init_state();
  void initialize_diffusion();
  void init_sound();
  void psc2psd(const A2& pc, A2& pd);
  void nhbase(int i1, int i2, int j1, int j2, int kx, const A1& sig,
              const A2& ter, A2& ps0, A3& pr0, A3& t0, A3& rho0, A3& z0);

  // ---- tend and its internal procedures (oracle/tendency.cpp) ----
  void tend();
  void surface_pressures();
  void decouple();
  void compute_omega();
  void mkslice();
  void init_tendencies();
  void advection();
  void curvature();
  void adiabatic();
  void boundary();
  void diffusion();

  // ---- mod_advection (oracle/advection.cpp) ----
  void start_advect();
  void hadvuv(A3& ut, A3& vt, const A3& u, const A3& v);
  void vadvuv(A3& ut, A3& vt, const A3& ud, const A3& vd);
  void hadvt(A3& ften, const A3& f);
  void hadv3d(A3& ften, const A3& f, int ind);
  void hadvqv(A4& ften, const A4& f, int iv);
  void hadvqx(A4& ften, const A4& f, int n1, int n2);
  void vadv3d(A3& ften, const A3& f, int nk, int ind);
  void vadvqv(A4& ften, const A4& f, int n);
  void vadv4d(A4& ften, const A4& f, int n1, int n2, int ind);

  // ---- mod_diffusion (oracle/diffusion.cpp) ----
  void calc_coeff();
  void calc_coeff6();
  void diffu_d(A3& ut, A3& vt, const A3& u, const A3& v);
  void diffu_x3d(A3& ften, const A3& f);
  void diffu_x3df(A3& ften, const A3& f, double fac);
  void diffu_x4d(A4& ften, const A4& f, int n1, int n2, double fac);

  // ---- mod_bdycod (oracle/bdycod.cpp) ----
  void nudge3d(int ibdy, const A3& f, const V3dBound& bnd, A3& ften);
  void nudge4d3d(int ibdy, const A4& f, const V3dBound& bnd, A4& ften, int n);
  void nudgeuv(int ibdy, const A3& fu, const A3& fv, const V3dBound& bu,
               const V3dBound& bv, A3& ftu, A3& ftv);
  void sponge3d(const V3dBound& bnd, A3& ften);
  void sponge4d(const V3dBound& bnd, A4& ften, int m);
  void spongeuv(const V3dBound& bu, const V3dBound& bv, A3& ftu, A3& ftv);
  void sponge2d(const V2dBound& bnd, A2& ften);
  void raydamp3(const A3& z, const A3& var, A3& vten, const V3dBound& bnd);
  void raydamp3f(const A3& z, const A3& var, A3& vten, double sval);
  void raydampuv(const A3& z, const A3& u, const A3& v, A3& ut, A3& vt,
                 const V3dBound& ub, const V3dBound& vb);
  void raydampqv(const A3& z, const A4& var, A4& vten, const V3dBound& bnd);
  double tau(double z, double zmax) const;
  void bdyin();
  void bdyuv(double xt);
  void bdyval();
  void timeint(const A3& a, const A3& b, A3& c, int j1, int j2, int i1, int i2,
               int k1, int k2);
  void couple(A3& a, const A2& c, int j1, int j2, int i1, int i2, int k1,
              int k2);

  // ---- mod_timefilter (oracle/timefilter.cpp) ----
  void filter_ra_3d(A3& phin, A3& phinm1, const A3& phinp1, double alpha);
  void filter_ra_uv(A3& un, A3& unm1, const A3& unp1, A3& vn, A3& vnm1,
                    const A3& vnp1, double alpha);
  void filter_raw_qv(A4& phin, A4& phinm1, const A4& phinp1, double alpha,
                     double beta, const A2& psa_, const A2& psb_);
  void filter_raw_4d(A4& phin, A4& phinm1, const A4& phinp1, double alpha,
                     double beta, int n1, int n2, double low);

  // ---- mod_sound (oracle/sound.cpp) ----
  void sound();

  // timers
  bool integrating() const { return lcount > 0; }
  bool alarm_day_act() const { return day_now == model_time; }
  bool alarm_bdy_act() const { return bdy_now == model_time; }
  void timer_advance();
};

}  // namespace orc
