// libregcm_dyn.so — host side of the B200 dyn-core: handle, static setup,
// time stepping driver and the C ABI of include/regcm_dyn.h.
//
// There is NO CPU fallback in this library: every entry point that computes
// runs CUDA kernels and fails with RCM_ERR_CUDA when no device is usable.
// The static setup (reference atmosphere, map-factor products, relaxation
// weights, Rayleigh coefficients) is evaluated once on the host in double
// precision exactly as `param` does (Main/mod_params.F90:1960-2074,2613-2741)
// and uploaded; everything that runs per step is on the device.

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/regcm_dyn.h"
#include "consts.h"
#include "dyn_types.h"
#include "halo.h"
#include "launch.h"
#include "split_host.h"

using namespace rcm;

// kernel launchers (k_*.cu)
void launch_prep(const Geom&, const Fields&, const StepScalars&, cudaStream_t);
void launch_uv(const Geom&, const Fields&, const StepScalars&, cudaStream_t);
bool tq_tma_setup();
bool ppw_tma_setup();
void launch_ppw_tma(const Geom&, const Fields&, const StepScalars&, cudaStream_t);
void launch_tq_tma(const Geom&, const Fields&, const StepScalars&, cudaStream_t);
void* uv_tma_setup(const Geom&, const Fields&, long long off3, int ni_pad);
void uv_tma_destroy(void*);
void launch_uv_tma(const Geom&, const Fields&, const StepScalars&, cudaStream_t, const void* maps);
void launch_tq(const Geom&, const Fields&, const StepScalars&, cudaStream_t);
void launch_ppw(const Geom&, const Fields&, const StepScalars&, cudaStream_t);
void launch_snd_init(const Geom&, const Fields&, const StepScalars&, cudaStream_t);
void launch_snd_uv(const Geom&, const Fields&, const StepScalars&, cudaStream_t);
void launch_xkc(const Geom&, const Fields&, const StepScalars&, cudaStream_t);
void launch_snd_wfwd(const Geom&, const Fields&, const StepScalars&, cudaStream_t);
bool wfwd_tma_setup();
void launch_snd_wfwd_tma(const Geom&, const Fields&, const StepScalars&, cudaStream_t);
void launch_snd_wback(const Geom&, const Fields&, const StepScalars&, cudaStream_t);
void launch_snd_final(const Geom&, const Fields&, const StepScalars&, cudaStream_t);
void launch_negq_repair(const Geom&, const Fields&, const StepScalars&, cudaStream_t);
void launch_slice(const Geom&, const Fields&, int id, int hydro, double ptop, double p0, double tlp, double ts0, double dx8, double* out, cudaStream_t);
void launch_window_strided(bool pack, double* buf, double* p_shifted, int j0, int i0, int k0, int n1, int n2, int nk, long long b1, long long b2, int si, long long sk, cudaStream_t st);
void launch_window(bool pack, double* buf, double* p_shifted, int j0, int i0, int k0, int n1, int n2, int nk, int si, long long sk, cudaStream_t st);
void launch_tr(const Geom&, const Fields&, const TrFields&, const StepScalars&, cudaStream_t);
void launch_tke(const Geom&, const Fields&, const UwFields&, const StepScalars&, cudaStream_t);
void launch_sl(const Geom&, const Fields&, const TrFields&, const StepScalars&, cudaStream_t);
void launch_diff6(const Geom&, const Fields&, const TrFields&, const UwFields&, const StepScalars&, int, double, cudaStream_t);
void launch_mo_phys(const Geom&, const MoFields&, double, cudaStream_t);
void launch_mo_slice(const Geom&, const MoFields&, int, double*, cudaStream_t);
void launch_tke_bdy(const Geom&, const Fields&, const UwFields&, const int2*, int, int, int, cudaStream_t);
void launch_tr_repair(const Geom&, const TrFields&, const StepScalars&, cudaStream_t);
void launch_tr_bdyval(const Geom&, const TrFields&, const int2*, int, double, int, int, cudaStream_t);
void launch_tr_timeint(const Geom&, const double*, const double*, double*, double, int, int, int, int, int, cudaStream_t);
void launch_couple(const Geom&, double*, const double*, int, int, int, int, int, cudaStream_t);
void launch_timeint(const Geom&, const double*, const double*, double*, double, int, int, int, int, int, cudaStream_t);
void launch_bdyval(const Geom&, const Fields&, const double*, const double*, const double*, const int2*, int,
                   const int2*, int, double, int, int, cudaStream_t);

// hydrostatic core (k_hydro.cu, k_split.cu)
void launch_psc2psd(const Geom&, const double*, double*, double*, int gw, cudaStream_t);
void launch_ps_convert(const Geom&, double*, double, cudaStream_t);
void launch_h_ps(const Geom&, const Fields&, cudaStream_t);
void launch_h_rimcopy(const Geom&, const Fields&, const int2*, int, const int2*, int, cudaStream_t);
void launch_h_prep(const Geom&, const Fields&, const StepScalars&, cudaStream_t);
void launch_h_uv(const Geom&, const Fields&, const StepScalars&, cudaStream_t);
void launch_h_tq(const Geom&, const Fields&, const StepScalars&, cudaStream_t);
void launch_h_psfilt(const Geom&, const Fields&, const StepScalars&, cudaStream_t);
void launch_bdyval_ps(const Geom&, const Fields&, const int2*, int, double, int, cudaStream_t);
void launch_sp_init(const Geom&, const Fields&, const StepScalars&, cudaStream_t);
void launch_sp_proj(const Geom&, const Fields&, const StepScalars&, cudaStream_t);
void launch_sp_step(const Geom&, const Fields&, const StepScalars&, int, cudaStream_t);
void launch_sp_corr(const Geom&, const Fields&, const StepScalars&, cudaStream_t);

#define CUDA_OK(call)                                                                   \
  do {                                                                                  \
    cudaError_t e_ = (call);                                                            \
    if (e_ != cudaSuccess) {                                                            \
      h->err = std::string(#call) + ": " + cudaGetErrorString(e_);                      \
      return RCM_ERR_CUDA;                                                              \
    }                                                                                   \
  } while (0)

// per-kernel device timing (CUDA events on the library's stream), enabled by
// rcm_dyn_profile(); used by bench.py for the per-kernel roofline
enum { KID_PREP, KID_XKC, KID_UV, KID_TQ, KID_PPW, KID_SND_INIT, KID_SND_UV, KID_SND_WFWD, KID_SND_WPVAL, KID_SND_WBACK,
       KID_SND_FINAL, KID_BDY, KID_HALO, KID_H_PS, KID_H_PREP, KID_H_UV, KID_H_TQ, KID_SP_PROJ, KID_SP_STEP, KID_SP_CORR,
       KID_MO_PRE, KID_MO_DIV, KID_MO_W, KID_MO_UV, KID_MO_STAG, KID_MO_WAF, KID_NEGQ, KID_TR, KID_COUNT };
static const char* const kKernelNames[KID_COUNT] = {"k_prep", "k_xkc", "k_uv", "k_tq", "k_ppw", "k_snd_init", "k_snd_uv",
                                                    "k_snd_wfwd", "k_snd_wpval", "k_snd_wback", "k_snd_final", "k_bdy", "halo",
                                                    "k_h_ps", "k_h_prep", "k_h_uv", "k_h_tq", "k_sp_proj", "k_sp_step", "k_sp_corr",
                                                    "k_mo_pre", "k_mo_div", "k_mo_w", "k_mo_uv", "k_mo_stag", "k_mo_waf", "k_negq", "k_tr"};
struct ProfRec { int kid; cudaEvent_t a, b; };

struct rcm_dyn {
  bool prof_on = false;
  bool ppw_tma = false;      // k_ppw_tma instead of k_ppw
  bool tq_tma = false;       // k_tq_tma instead of k_tq
  bool wfwd_tma = false;     // k_snd_wfwd_tma instead of k_snd_wfwd<false>
  void* uv_maps = nullptr;   // TMA tensor maps of k_uv_tma (nullptr: plain k_uv)
  bool overlap_slow = true;  // RCM_OVERLAP=0 serialises the slow-tendency kernels
  std::vector<ProfRec> prof_recs;
  std::vector<cudaEvent_t> prof_pool;
  double prof_ms[KID_COUNT] = {0};
  int64_t prof_n[KID_COUNT] = {0};
  rcm_cfg cfg{};
  Geom g{};
  Fields f{};
  StepScalars s{};
  std::string err;
  int device = 0;
  bool inited = false;
  // geometry helpers
  int nj_pad = 0, ni_pad = 0;
  long long off2 = 0, off3 = 0;  // origin shifts (elements)
  long long eoff = 0; int eni = 0;
  // host statics in the padded geometry (origin-shifted accessors below)
  std::vector<double> h_ht, h_msfx, h_msfd, h_coriol, h_xlat, h_xlon, h_dlat, h_sigma;
  // device memory
  std::vector<void*> allocs;
  double* b1[6] = {nullptr};   // boundary record b1 per bdy field (u v t qv pp w)
  double* nx[6] = {nullptr};   // staged next record (decoupled)
  // tracers
  TrFields tr{};
  UwFields uw = {};                       // TKE of the UW PBL (ibltyp = 2)
  bool uw_on = false;
  int* kpb = nullptr;                     // kpbl, origin-shifted 2-D (always allocated with ibltyp = 2; f.kpb only with iuwvadv = 1)
  std::vector<int> kpb_host;              // its host mirror (rcm_dyn_set_kpbl updates a window and uploads)
  double* tr_b1[RCM_MAXTR] = {nullptr};   // chib1
  double* tr_nx[RCM_MAXTR] = {nullptr};   // staged next record (mixing ratio)
  int64_t neg_chi = 0;
  double* psdot_bdy = nullptr; // bdycon psdot = atm0%psdot*1e-3 (Main/mod_bdycod.F90:425-431)
  double* xpsb = nullptr;      // xpsb%b0 = atm0%ps*1e-3
  // hydrostatic core
  bool hydro = false;
  double *xps_b1 = nullptr, *xps_nx = nullptr;  // xpsb%b1 and the staged next record (2-D)
  SplitTables sp;
  double ptntot = 0, pt2tot = 0;
  // MOLOCH core (idynamic = 3), mo_core.inc
  bool moloch = false;
  MoFields mf{};
  MoScalars ms{};
  double* mo_zdf = nullptr;   // full divergence (the raw one is still read by the neighbours' filter)
  int2 *mo_rim_u = nullptr, *mo_rim_v = nullptr, *mo_rim_c = nullptr, *mo_zl_c = nullptr, *mo_zl_u = nullptr, *mo_zl_v = nullptr;
  int mo_nrim_u = 0, mo_nrim_v = 0, mo_nrim_c = 0, mo_nzl_c = 0, mo_nzl_u = 0, mo_nzl_v = 0;
  std::vector<double> h_msfu, h_msfv, h_ulat, h_vlat;
  double ps_max = 0, ps_min = 0;
  int2 *rim_d = nullptr, *rim_c = nullptr;
  int nrim_d = 0, nrim_c = 0;
  std::vector<double> stage;   // host staging, one 3-D field
  // time bookkeeping (rcmtimer, alarms: Main/mpplib/mod_timer.F90)
  int64_t lcount = 0, model_time = 0;
  int64_t day_lastact = 0, day_now = 0, bdy_lastact = 0, bdy_now = 0;
  double dtsec = 0, dt = 0, dt2 = 0, xbctime = 0, dtbdys = 0, rdtbdy = 0;
  bool have_next = false;
  // sound
  double dtsmax = 0, cs = 0, rnpts = 0, xmsf_mean = 0;
  double fi[13], fj[13], fk[7], fl[7];
  int last_istep = 0;
  double max_cfl = 0;
  int64_t neg_q = 0;
  int cq_mode = 0;   // StepScalars::cq_mode of the next call: 2 after a call that repaired negative moisture, else 0
  int nan_flag = 0;
  // stream / timing
  cudaStream_t st = nullptr;
  cudaStream_t st_copy = nullptr;  // asynchronous staging of the next boundary record
  cudaEvent_t ev_stage = nullptr;
  bool stage_pending = false;
  cudaStream_t st_b = nullptr, st_c = nullptr;  // side streams: the three slow-tendency kernels are independent
  cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev_fork = nullptr, ev_jb = nullptr, ev_jc = nullptr, ev_prep = nullptr;
  bool overlap_head = true;  // RCM_OVERLAP2=0: k_xkc after k_prep and k_snd_init after the slow kernels instead of beside them
  int64_t launches = 0;
  double last_ms = 0;
  Halo halo;
  // asynchronous stepping (rcm_dyn_step_async / rcm_dyn_sync): steps enqueued and not yet finished
  bool pending = false;
  int pending_steps = 0;
  // per-step verdicts, copied D2H on the compute stream after every step into a pinned ring (read without blocking by rcm_dyn_diag)
  struct VSlot { unsigned long long cfl[2], negq, nanf, ps[2]; };
  enum { kVRing = 64 };
  VSlot* vring = nullptr;                 // pinned host memory
  cudaEvent_t vev[kVRing] = {nullptr};
  int64_t vlcount[kVRing] = {0};
  int64_t vhead = 0;                      // slots written so far
  int64_t vdone = -1;                     // newest slot known to be complete
  // asynchronous state export (rcm_dyn_get_state_async): device snapshots + D2H on a private stream
  cudaStream_t st_out = nullptr;
  enum { kSnap = 2 * (RCM_F_CHI + RCM_MAXTR) };
  double* snap[kSnap] = {nullptr};
  cudaEvent_t ev_snap[kSnap] = {nullptr}, ev_out[kSnap] = {nullptr};
  bool out_pending = false;
  // physics coupling boundary: rcm_dyn_step_pre_physics ran the head of `tend` and the tail is owed
  bool mid_step = false, head_only = false;
  double* slice_buf = nullptr;   // one 3-D field: the slice rcm_dyn_get_slice asked for
  double* stage_dev = nullptr;   // contiguous landing buffer of rcm_dyn_stage_bdy_async (one 3-D field)
  size_t stage_cap = 0;          // its capacity in doubles
};

// host accessors in the padded geometry
#define H2(v, j, i) (v)[h->off2 + (long long)(j) + (long long)(i) * h->g.si]
#define H3(v, j, i, k) (v)[h->off3 + (long long)(j) + (long long)(i) * h->g.si + (long long)(k) * h->g.sk]

static cudaEvent_t prof_event(rcm_dyn* h) {
  if (!h->prof_pool.empty()) { cudaEvent_t e = h->prof_pool.back(); h->prof_pool.pop_back(); return e; }
  cudaEvent_t e; cudaEventCreate(&e); return e;
}
#define TIMED(kid_, n_, call)                                             \
  do {                                                                    \
    if (h->prof_on) {                                                     \
      ProfRec r_; r_.kid = (kid_); r_.a = prof_event(h); r_.b = prof_event(h); \
      cudaEventRecord(r_.a, h->st); call; cudaEventRecord(r_.b, h->st);   \
      h->prof_recs.push_back(r_);                                         \
    } else { call; }                                                      \
    h->launches += (n_);                                                  \
  } while (0)
static void prof_collect(rcm_dyn* h) {
  for (ProfRec& r : h->prof_recs) {
    float ms = 0.f;
    cudaEventElapsedTime(&ms, r.a, r.b);
    h->prof_ms[r.kid] += ms; h->prof_n[r.kid] += 1;
    h->prof_pool.push_back(r.a); h->prof_pool.push_back(r.b);
  }
  h->prof_recs.clear();
}

static int mo_create(rcm_dyn* h);
static int mo_init(rcm_dyn* h);
static int mo_step_one(rcm_dyn* h);
static int mo_bdyin(rcm_dyn* h);
static int mo_bdyval(rcm_dyn* h);
static double* mo_state_ptr(rcm_dyn* h, int field);
static double* mo_work_ptr(rcm_dyn* h, int id);
static int mo_exch(rcm_dyn* h, const std::vector<HaloField>& fl, int width);

static int fail(rcm_dyn* h, int code, const std::string& msg) {
  if (h) h->err = msg;
  return code;
}
static int finish_steps(rcm_dyn* h);   // waits for steps enqueued by rcm_dyn_step_async; every accessor calls it first
static int tend(rcm_dyn* h);
static int download3(rcm_dyn* h, const double* dev_shifted, double* a, const int32_t* lb, const int32_t* ub, int nk);

// ---------------------------------------------------------------------------
extern "C" int rcm_dyn_default_cfg(rcm_cfg* c, int idynamic) {
  if (!c) return RCM_ERR_ARG;
  std::memset(c, 0, sizeof(*c));
  c->struct_size = (int32_t)sizeof(rcm_cfg);
  c->nqx = 2;
  c->idynamic = idynamic;
  c->iboudy = 5; c->idiffu = 1; c->ipptls = 1;
  c->ifupr = 1; c->ifrayd = 1; c->rayndamp = 5;
  c->nspgx = c->nspgd = 12;
  c->ibdyfrq = 6;
  c->upstream_mode = c->stability_enhance = 1;
  c->nproc = 1; c->myid = 0; c->cpus_j = c->cpus_i = 0; c->device = -1;
  c->ptop = 5.0;
  c->nsplit = 2; c->lstand = 1;  // Main/mod_params.F90:283-284
  c->mo_nadv = 3; c->mo_nsound = 6; c->mo_nzfilt = 3; c->mo_anu2 = 0.6;  // :296-300
  c->mo_a0 = 0.0; c->mo_b0 = 0.5;                                        // Share/mod_dynparam.F90:189-190
  c->iqmsl = 1;                             // Main/mod_params.F90:244
  c->uw_nuk = 5.0; c->uw_tkemin = 1.0e-3;   // Main/mod_params.F90:480; Main/pbllib/mod_pbl_uwtcm.F90:86
  if (idynamic == 2) { c->gnu1 = c->gnu2 = 0.1; c->diffu_hgtf = 0; }
  else { c->gnu1 = c->gnu2 = 0.0625; c->diffu_hgtf = 1; }
  c->ckh = c->adyndif = 1.0;
  c->uoffc = 0.25; c->t_extrema = 5.0; c->q_rel_extrema = 0.20;
  c->nhbet = 0.4; c->nhxkd = 0.1; c->rayalpha0 = 0.0003; c->rayhd = 10000.0;
  c->high_nudge = 3.0; c->medium_nudge = 2.0; c->low_nudge = 1.0;
  c->bdy_nm = c->bdy_dm = -1.0;
  c->base_state_pressure = 101325.0; c->logp_lrate = 47.70; c->base_state_ts0 = 288.15;
  return RCM_OK;
}

// ---------------------------------------------------------------------------
// decomposition: set_nproc rules (Main/mpplib/mod_mppparam.F90:1152-1172,1302-1360)
static void choose_layout(int nproc, int jx, int iy, int& cj, int& ci) {
  if (nproc == 1) { cj = ci = 1; return; }
  if (nproc == 2) { cj = 2; ci = 1; return; }  // :1152-1160 (split jx)
  cj = (int)std::lround(std::sqrt((double)nproc));
  ci = nproc / cj;
  if (jx > (int)(1.5 * iy)) { cj = cj - 1; do { cj = cj + 1; if (nproc % cj == 0) break; } while (cj < nproc); ci = nproc / cj; }
  else if (iy > (int)(1.5 * jx)) { ci = ci - 1; do { ci = ci + 1; if (nproc % ci == 0) break; } while (ci < nproc); cj = nproc / ci; }
  else { while (cj * ci != nproc && cj > 1) { cj--; ci = nproc / cj; } if (cj * ci != nproc) { cj = nproc; ci = 1; } if (cj > ci && jx <= iy) std::swap(cj, ci); }
}

static void split_range(int n, int parts, int coord, int& lo, int& hi) {
  // remainders go to the low-coordinate ranks (:1302-1328)
  const int base = n / parts, rem = n % parts;
  lo = 1 + coord * base + std::min(coord, rem);
  hi = lo + base - 1 + (coord < rem ? 1 : 0);
}

static double* dev_alloc3(rcm_dyn* h, int nk, bool shifted = true) {
  double* p = nullptr;
  const size_t n = (size_t)h->g.sk * (size_t)nk;
  if (cudaMalloc(&p, n * sizeof(double)) != cudaSuccess) return nullptr;
  cudaMemset(p, 0, n * sizeof(double));
  h->allocs.push_back(p);
  return shifted ? p + h->off3 : p;
}
static double* dev_alloc2(rcm_dyn* h) {
  double* p = nullptr;
  const size_t n = (size_t)h->g.sk;
  if (cudaMalloc(&p, n * sizeof(double)) != cudaSuccess) return nullptr;
  cudaMemset(p, 0, n * sizeof(double));
  h->allocs.push_back(p);
  return p + h->off2;
}
template <class T>
static T* dev_alloc_raw(rcm_dyn* h, size_t n) {
  T* p = nullptr;
  if (cudaMalloc(&p, std::max<size_t>(n, 1) * sizeof(T)) != cudaSuccess) return nullptr;
  cudaMemset(p, 0, std::max<size_t>(n, 1) * sizeof(T));
  h->allocs.push_back(p);
  return p;
}

extern "C" int rcm_dyn_create(const rcm_cfg* cfg, rcm_dyn** out) {
  if (!cfg || !out) return RCM_ERR_ARG;
  *out = nullptr;
  rcm_dyn* h = new rcm_dyn();
  auto bail = [&](int code, const char* msg) { std::fprintf(stderr, "rcm_dyn_create: %s\n", msg); delete h; return code; };
  h->cfg = *cfg;
  const rcm_cfg& c = h->cfg;
  if (c.struct_size != (int)sizeof(rcm_cfg)) return bail(RCM_ERR_ARG, "rcm_cfg size mismatch");
  if (c.idynamic < 1 || c.idynamic > 3) return bail(RCM_ERR_ARG, "idynamic must be 1 (hydrostatic), 2 (non-hydrostatic) or 3 (MOLOCH)");
  h->hydro = (c.idynamic == 1);
  h->moloch = (c.idynamic == 3);
  if (h->moloch && (c.mo_nadv < 1 || c.mo_nsound < 1 || c.kz < 4 || c.ipptls < 0 || c.ipptls > 1))
    return bail(RCM_ERR_ARG, "idynamic=3 needs mo_nadv >= 1, mo_nsound >= 1, kz >= 4, ipptls 0 or 1");
  if (h->hydro && (c.nsplit < 1 || c.nsplit > RCM_MAXSPLIT || c.nsplit > c.kz)) return bail(RCM_ERR_ARG, "nsplit must be 1..4");
  if (c.ipgf != 0 && c.ipgf != 1) return bail(RCM_ERR_ARG, "ipgf must be 0 or 1");
  if (c.isladvec != 0 && (c.isladvec != 1 || c.idynamic == 3))
    return bail(RCM_ERR_ARG, "isladvec must be 0 or 1; 1 runs with idynamic 1, 2");
  if (c.idiffu < 1 || c.idiffu > 3) return bail(RCM_ERR_ARG, "idiffu must be 1, 2 or 3");
  if (c.idiffu == 3 && c.idynamic != 3 && c.nproc != 1)
    return bail(RCM_ERR_ARG, "idiffu = 3 runs on one rank (the reference applies it to the last column of EVERY tile: the result would depend on the decomposition)");
  // 0 fixed, 1 / 5 relaxation, 4 sponge; sigma-coordinate cores also 2 (time-dependent, no relaxation) and 3 (2 + inflow/outflow)
  if (!(c.iboudy == 0 || c.iboudy == 1 || c.iboudy == 4 || c.iboudy == 5 || (c.idynamic != 3 && (c.iboudy == 2 || c.iboudy == 3))))
    return bail(RCM_ERR_ARG, "iboudy must be 0, 1, 4 or 5 (idynamic 1, 2: also 2, 3)");
  if (c.nqx != 2) return bail(RCM_ERR_ARG, "nqx must be 2 (qv,qc)");
  if (c.ntr < 0 || c.ntr > RCM_MAXTR || (c.ntr > 0 && c.idynamic == 3)) return bail(RCM_ERR_ARG, "ntr must be 0..8 (tracers: idynamic 1, 2)");
  if (c.jx < 2 * c.nspgx || c.iy < 2 * c.nspgx || c.kz < 3) return bail(RCM_ERR_ARG, "domain too small for the relaxation zone");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return bail(RCM_ERR_CUDA, "no CUDA device (this library has no CPU fallback)");
  h->device = (c.device >= 0) ? c.device : (c.myid % ndev);
  if (cudaSetDevice(h->device) != cudaSuccess) return bail(RCM_ERR_CUDA, "cudaSetDevice failed");

  Geom& g = h->g;
  g.jx = c.jx; g.iy = c.iy; g.kz = c.kz; g.kzp1 = c.kz + 1;
  int cj = c.cpus_j, ci = c.cpus_i;
  if (cj <= 0 || ci <= 0) choose_layout(c.nproc, c.jx, c.iy, cj, ci);
  if (cj * ci != c.nproc) return bail(RCM_ERR_ARG, "cpus_j*cpus_i != nproc");
  const int pj = c.myid % cj, pi = c.myid / cj;
  split_range(c.jx, cj, pj, g.jde1, g.jde2);
  split_range(c.iy, ci, pi, g.ide1, g.ide2);
  g.has_bdyleft = (pj == 0); g.has_bdyright = (pj == cj - 1);
  g.has_bdybottom = (pi == 0); g.has_bdytop = (pi == ci - 1);
  g.gl = g.has_bdyleft ? 0 : 1; g.gr = g.has_bdyright ? 0 : 1;
  g.gb = g.has_bdybottom ? 0 : 1; g.gt = g.has_bdytop ? 0 : 1;
  // Main/mod_atm_interface.F90:203-353
  g.jdi1 = g.jde1 + (g.has_bdyleft ? 1 : 0); g.jdii1 = g.jde1 + (g.has_bdyleft ? 2 : 0);
  g.jdi2 = g.jde2 - (g.has_bdyright ? 1 : 0); g.jdii2 = g.jde2 - (g.has_bdyright ? 2 : 0);
  g.idi1 = g.ide1 + (g.has_bdybottom ? 1 : 0); g.idii1 = g.ide1 + (g.has_bdybottom ? 2 : 0);
  g.idi2 = g.ide2 - (g.has_bdytop ? 1 : 0); g.idii2 = g.ide2 - (g.has_bdytop ? 2 : 0);
  g.jce1 = g.jde1; g.jce2 = g.has_bdyright ? g.jde2 - 1 : g.jde2;
  g.ice1 = g.ide1; g.ice2 = g.has_bdytop ? g.ide2 - 1 : g.ide2;
  g.jci1 = g.jce1 + (g.has_bdyleft ? 1 : 0); g.jcii1 = g.jce1 + (g.has_bdyleft ? 2 : 0);
  g.jci2 = g.jce2 - (g.has_bdyright ? 1 : 0); g.jcii2 = g.jce2 - (g.has_bdyright ? 2 : 0);
  g.ici1 = g.ice1 + (g.has_bdybottom ? 1 : 0); g.icii1 = g.ice1 + (g.has_bdybottom ? 2 : 0);
  g.ici2 = g.ice2 - (g.has_bdytop ? 1 : 0); g.icii2 = g.ice2 - (g.has_bdytop ? 2 : 0);
  g.jcross1 = 1; g.jcross2 = c.jx - 1; g.icross1 = 1; g.icross2 = c.iy - 1;

  const int nj = g.jde2 - g.jde1 + 1, ni = g.ide2 - g.ide1 + 1;
  if (c.nproc > 1) {
    // Every rank checks EVERY tile, so that all ranks reach the same verdict.  A tile with neighbours on both
    // sides must hold the edge slab it sends: RCM_EHALO columns of estore for the upper radiative BC (a tile on
    // a physical boundary may be narrower: the columns it cannot supply lie outside the clamped 13x13 window,
    // Main/mod_sound.F90:551-561); the reference rejects tiles under 3x3 (Main/mpplib/mod_mppparam.F90:1365).
    const int need = (c.idynamic == 2 && c.ifupr == 1) ? RCM_EHALO : RCM_HALO;
    for (int q = 0; q < cj; ++q) {
      int lo, hi; split_range(c.jx, cj, q, lo, hi);
      const int n = hi - lo + 1;
      if (n < 3 || (q > 0 && q < cj - 1 && n < need)) return bail(RCM_ERR_ARG, "tile too narrow along j for the halo width (interior tiles need >= 6 points with ifupr = 1; all tiles >= 3)");
    }
    for (int q = 0; q < ci; ++q) {
      int lo, hi; split_range(c.iy, ci, q, lo, hi);
      const int n = hi - lo + 1;
      if (n < 3 || (q > 0 && q < ci - 1 && n < need)) return bail(RCM_ERR_ARG, "tile too narrow along i for the halo width (interior tiles need >= 6 points with ifupr = 1; all tiles >= 3)");
    }
  }
  g.hw = (c.isladvec == 1 && c.nproc > 1) ? 4 : RCM_HALO;   // even, so that the 16-byte alignment of the staged tiles does not move
  h->nj_pad = ((nj + 2 * g.hw + 15) / 16) * 16;
  h->ni_pad = ni + 2 * g.hw;
  g.si = h->nj_pad;
  g.sk = (long long)h->nj_pad * h->ni_pad;
  // the kernels index with 32-bit arithmetic (dev_common.cuh): the largest element index of a field must fit
  if ((long long)(g.jde2 + RCM_EHALO + 16) + (long long)std::max(g.si, g.esi) * (g.ide2 + RCM_EHALO + 4) + g.sk * (long long)(g.kzp1 + 2) >= (1LL << 31))
    return bail(RCM_ERR_ARG, "tile too large for 32-bit element indices (more than 2^31 elements per field incl. the origin shift)");
  h->off2 = -((long long)(g.jde1 - g.hw) + (long long)g.si * (g.ide1 - g.hw));
  h->off3 = h->off2 - g.sk;
  g.esi = ((nj + 2 * RCM_EHALO + 15) / 16) * 16;
  g.pfd = 2;
  g.kc_uv = g.kc_tq = g.kc_ppw = g.kz + 1;
  if (const char* e = std::getenv("RCM_KC_UV")) g.kc_uv = std::max(1, std::atoi(e));
  if (const char* e = std::getenv("RCM_KC_TQ")) g.kc_tq = std::max(1, std::atoi(e));
  if (const char* e = std::getenv("RCM_KC_PPW")) g.kc_ppw = std::max(1, std::atoi(e));
  if (const char* e = std::getenv("RCM_OVERLAP")) h->overlap_slow = std::atoi(e) != 0;
  if (const char* e = std::getenv("RCM_OVERLAP2")) h->overlap_head = std::atoi(e) != 0;
  g.pfd_wfwd = 4; g.pfd_wback = 1;   // measured on C4: wfwd 5.50 / 5.28 / 5.05 / 4.98 ms at distance 1 / 2 / 3 / 4, wback 4.86 / 4.96 / 5.19 / 5.57
  if (const char* e = std::getenv("RCM_PFD")) g.pfd = g.pfd_wfwd = g.pfd_wback = std::max(0, std::min(8, std::atoi(e)));
  if (const char* e = std::getenv("RCM_PFD_WFWD")) g.pfd_wfwd = std::max(0, std::min(12, std::atoi(e)));
  if (const char* e = std::getenv("RCM_PFD_WBACK")) g.pfd_wback = std::max(0, std::min(12, std::atoi(e)));
  h->eni = ni + 2 * RCM_EHALO;
  h->eoff = -((long long)(g.jde1 - RCM_EHALO) + (long long)g.esi * (g.ide1 - RCM_EHALO));

  h->dtsec = c.dt; h->dt = c.dt; h->dt2 = 2.0 * c.dt;
  h->dtbdys = (double)c.ibdyfrq * 3600.0;
  h->rdtbdy = 1.0 / h->dtbdys;

  if (h->moloch) {
    if (mo_create(h) != RCM_OK) {
      for (void* p : h->allocs) cudaFree(p);
      return bail(RCM_ERR_CUDA, "device memory allocation failed");
    }
    cudaStreamCreateWithFlags(&h->st, cudaStreamNonBlocking);
    cudaStreamCreateWithFlags(&h->st_b, cudaStreamNonBlocking);
    cudaStreamCreateWithFlags(&h->st_c, cudaStreamNonBlocking);
    cudaEventCreate(&h->ev0); cudaEventCreate(&h->ev1);
    cudaEventCreate(&h->ev_fork); cudaEventCreate(&h->ev_jb); cudaEventCreate(&h->ev_jc); cudaEventCreate(&h->ev_prep);
    h->halo.init(g, c.nproc, c.myid, cj, ci);
    *out = h;
    return RCM_OK;
  }
  // ---- device memory ----
  Fields& f = h->f;
  const int kzp1 = g.kzp1;
  bool ok = true;
  auto A3 = [&](double*& p) { p = dev_alloc3(h, kzp1); ok = ok && p; };
  auto A2 = [&](double*& p) { p = dev_alloc2(h); ok = ok && p; };
  for (double** p : {&f.au, &f.av, &f.at, &f.aqv, &f.aqc, &f.app, &f.aw, &f.bu, &f.bv, &f.bt, &f.bqv, &f.bqc, &f.bpp,
                     &f.bw, &f.nat, &f.naqv, &f.naqc, &f.nbt, &f.nbqv, &f.nbqc, &f.xub0, &f.xubt, &f.xvb0, &f.xvbt,
                     &f.xtb0, &f.xtbt, &f.xqb0, &f.xqbt, &f.xpb0, &f.xpbt, &f.xwb0, &f.xwbt, &f.pr0, &f.t0, &f.rho0,
                     &f.rhof0, &f.pr1, &f.rho1, &f.th, &f.qdot, &f.cr, &f.xkc, &f.dub, &f.dvb, &f.scc, &f.scdd, &f.sca, &f.uten, &f.vten, &f.ppten, &f.wten,
                     &f.cu, &f.cv, &f.cpp, &f.cppd, &f.cw, &f.ct, &f.se, &f.sf, &f.cqv, &f.cqc})
    A3(*p);
  for (int q = 0; q < 6; ++q) { A3(h->b1[q]); A3(h->nx[q]); }
  for (double** p : {&f.psa, &f.rpsa, &f.psdota, &f.rpsda, &f.msfx, &f.msfd, &f.xmsf, &f.dmsf, &f.coriol, &f.ef, &f.ddx,
                     &f.ddy, &f.dmdx, &f.dmdy, &f.ex, &f.crx, &f.cry, &f.dpsdxm, &f.dpsdym, &f.ht, &f.ps0, &f.hgfact,
                     &f.astore, &f.wpval, &h->psdot_bdy, &h->xpsb})
    A2(*p);
  if (!h->hydro) {
    f.psb = f.psa; f.rpsb = f.rpsa; f.psdotb = f.psdota; f.rpsdb = f.rpsda;  // idynamic = 2: p* is the constant reference p*
  } else {
    for (double** p : {&f.psb, &f.rpsb, &f.psdotb, &f.rpsdb, &f.psc, &f.pten, &f.ptn2d, &f.pt22d, &f.xpsb0, &f.xpsbt, &f.map,
                       &h->xps_b1, &h->xps_nx})
      A2(*p);
    for (double** p : {&f.phi, &f.nau, &f.nav, &f.nbu, &f.nbv}) A3(*p);
    f.omega = f.th;  // tap only; th is not used by the hydrostatic core
    const int ns = c.nsplit;
    f.deld = dev_alloc3(h, 3 * ns); f.delh = dev_alloc3(h, 3 * ns);
    f.ddsum = dev_alloc3(h, ns); f.dhsum = dev_alloc3(h, ns); f.dstor = dev_alloc3(h, ns); f.hstor = dev_alloc3(h, ns);
    ok = ok && f.deld && f.delh && f.ddsum && f.dhsum && f.dstor && f.hstor;
    const size_t nt = (size_t)ns * (kzp1 + 1) + 1;
    for (double** p : {&f.sp_zr, &f.sp_tau, &f.sp_pdlog, &f.sp_eps1, &f.sp_am, &f.sp_zm}) { *p = dev_alloc_raw<double>(h, nt); ok = ok && *p; }
  }
  {
    int* z = dev_alloc_raw<int>(h, (size_t)g.sk); ok = ok && z; f.zone_c = z ? z + h->off2 : nullptr;
    z = dev_alloc_raw<int>(h, (size_t)g.sk); ok = ok && z; f.zone_d = z ? z + h->off2 : nullptr;
  }
  const int rl = std::max(1, std::min(kzp1, c.rayndamp));
  f.tau_c = dev_alloc3(h, rl); f.tau_d = dev_alloc3(h, rl); f.tau_f = dev_alloc3(h, rl);
  ok = ok && f.tau_c && f.tau_d && f.tau_f;
  const size_t nv = (size_t)g.kz + 3;
  for (double** p : {&f.sigma, &f.dsigma, &f.twt1, &f.twt2, &f.qcon, &f.xds, &f.dds}) { *p = dev_alloc_raw<double>(h, nv); ok = ok && *p; }
  f.hefc = dev_alloc_raw<double>(h, (size_t)c.nspgx * g.kz + 1); f.hegc = dev_alloc_raw<double>(h, (size_t)c.nspgx * g.kz + 1);
  f.hefd = dev_alloc_raw<double>(h, (size_t)c.nspgd * g.kz + 1); f.hegd = dev_alloc_raw<double>(h, (size_t)c.nspgd * g.kz + 1);
  {
    double* e = dev_alloc_raw<double>(h, (size_t)g.esi * h->eni); ok = ok && e; f.estore = e ? e + h->eoff : nullptr;
  }
  f.tmask = dev_alloc_raw<double>(h, 169);
  {  // one contiguous verdict block: [0..1] CFL maxima, [2] negative-q count, [3] NaN flag (one D2H copy per step)
    unsigned long long* blk = dev_alloc_raw<unsigned long long>(h, 4);
    f.cflbits = blk; f.negq = blk ? blk + 2 : nullptr; f.nanflag = blk ? (int*)(blk + 3) : nullptr;
  }
  f.nq_list = dev_alloc_raw<int>(h, (size_t)2 * g.kz * (size_t)g.sk);
  f.nq_cnt = dev_alloc_raw<int>(h, (size_t)2 * g.kz);
  ok = ok && f.hefc && f.hegc && f.hefd && f.hegd && f.tmask && f.cflbits && f.negq && f.nanflag && f.nq_list && f.nq_cnt;
  if (c.ibltyp == 2 && c.idynamic != 3) {   // UW-PBL coupling of the transport: TKE, kpbl
    UwFields& uw = h->uw;
    for (double** p : {&uw.a, &uw.b, &uw.c, &uw.phy}) A3(*p);
    uw.tkemin = c.uw_tkemin; uw.nuk = c.uw_nuk;
    int* z = dev_alloc_raw<int>(h, (size_t)g.sk); h->kpb = z ? z + h->off2 : nullptr;
    ok = ok && z;
    if (c.iuwvadv == 1) f.kpb = h->kpb;
    h->uw_on = true;
  }
  if (c.ntr > 0) {   // tracers
    TrFields& tr = h->tr;
    tr.ntr = c.ntr;
    for (int n = 0; n < c.ntr; ++n)
      for (double** p : {&tr.a[n], &tr.b[n], &tr.na[n], &tr.nb[n], &tr.c[n], &tr.b0[n], &tr.bt[n], &h->tr_b1[n], &h->tr_nx[n]}) A3(*p);
    tr.cefc = dev_alloc_raw<double>(h, (size_t)c.nspgx * g.kz + 1); tr.cegc = dev_alloc_raw<double>(h, (size_t)c.nspgx * g.kz + 1);
    int* z = dev_alloc_raw<int>(h, (size_t)g.sk); tr.ztype = z ? z + h->off2 : nullptr;
    tr.nq_list = dev_alloc_raw<int>(h, (size_t)c.ntr * g.kz * (size_t)g.sk);
    tr.nq_cnt = dev_alloc_raw<int>(h, (size_t)c.ntr * g.kz);
    tr.negcnt = dev_alloc_raw<unsigned long long>(h, 1);
    ok = ok && tr.cefc && tr.cegc && z && tr.nq_list && tr.nq_cnt && tr.negcnt;
  }
  if (c.idiffu == 3 && c.idynamic != 3) {   // the one column idiffu = 3 acts on (k_diff6)
    f.d6 = dev_alloc_raw<double>(h, (size_t)(RCM_D6_CHI + RCM_MAXTR) * g.kzp1 * (size_t)(g.iy + 2));
    ok = ok && f.d6;
  }
  if (c.isladvec == 1) {   // scratch of k_sl
    A3(f.slqv); A3(f.slqc);
    for (int n = 0; n < c.ntr; ++n) A3(h->tr.sl[n]);
  }
  if (!ok) {
    for (void* p : h->allocs) cudaFree(p);
    return bail(RCM_ERR_CUDA, "device memory allocation failed");
  }
  const size_t n2 = (size_t)g.sk;
  for (auto* v : {&h->h_ht, &h->h_msfx, &h->h_msfd, &h->h_coriol, &h->h_xlat, &h->h_xlon, &h->h_dlat}) v->assign(n2, 0.0);
  for (auto* v : {&h->h_msfx, &h->h_msfd}) v->assign(n2, 1.0);
  h->h_sigma.assign(nv, 0.0);
#ifdef RCM_EMU
  h->stage.assign((size_t)g.sk * kzp1, 0.0);
#endif
  cudaStreamCreateWithFlags(&h->st, cudaStreamNonBlocking);
  cudaStreamCreateWithFlags(&h->st_b, cudaStreamNonBlocking);
  cudaStreamCreateWithFlags(&h->st_c, cudaStreamNonBlocking);
  cudaEventCreate(&h->ev0); cudaEventCreate(&h->ev1);
  cudaEventCreate(&h->ev_fork); cudaEventCreate(&h->ev_jb); cudaEventCreate(&h->ev_jc); cudaEventCreate(&h->ev_prep);
  {
    const char* e = std::getenv("RCM_UV_TMA");
    if (!e || std::atoi(e) != 0) h->uv_maps = uv_tma_setup(g, f, h->off3, h->ni_pad);
    e = std::getenv("RCM_TQ_TMA");
    if (!e || std::atoi(e) != 0) h->tq_tma = tq_tma_setup();
    e = std::getenv("RCM_PPW_TMA");
    if (!e || std::atoi(e) != 0) h->ppw_tma = ppw_tma_setup();
    e = std::getenv("RCM_WFWD_TMA");  // opt-in: measured slower than the plain kernel
    if (e && std::atoi(e) != 0) h->wfwd_tma = wfwd_tma_setup();
  }
  h->halo.init(g, c.nproc, c.myid, cj, ci);
  *out = h;
  return RCM_OK;
}

// ---------------------------------------------------------------------------
// window copies between a caller array (Fortran bounds lb/ub) and the padded
// geometry
static void win_in(rcm_dyn* h, std::vector<double>& dst, long long off, int nk, const double* a, const int32_t* lb,
                   const int32_t* ub, int k0 /* k of plane 0 in dst */) {
  const Geom& g = h->g;
  const int jlo = std::max(lb[0], g.jde1 - RCM_HALO), jhi = std::min(ub[0], g.jde2 + RCM_HALO);
  const int ilo = std::max(lb[1], g.ide1 - RCM_HALO), ihi = std::min(ub[1], g.ide2 + RCM_HALO);
  const long long n1 = ub[0] - lb[0] + 1, n2 = ub[1] - lb[1] + 1;
  for (int k = std::max(lb[2], k0); k <= std::min(ub[2], k0 + nk - 1); ++k)
    for (int i = ilo; i <= ihi; ++i)
      for (int j = jlo; j <= jhi; ++j)
        dst[off + j + (long long)i * g.si + (long long)(k - k0) * g.sk] =
            a[(j - lb[0]) + n1 * ((i - lb[1]) + n2 * (long long)(k - lb[2]))];
}

extern "C" int rcm_dyn_set_static(rcm_dyn* h, int id, const double* a, const int32_t lb[3], const int32_t ub[3]) {
  if (!h || !a || !lb || !ub) return RCM_ERR_ARG;
  std::vector<double>* dst = nullptr;
  switch (id) {
    case RCM_S_HT: dst = &h->h_ht; break;
    case RCM_S_MSFX: dst = &h->h_msfx; break;
    case RCM_S_MSFD: dst = &h->h_msfd; break;
    case RCM_S_CORIOL: dst = &h->h_coriol; break;
    case RCM_S_XLAT: dst = &h->h_xlat; break;
    case RCM_S_XLON: dst = &h->h_xlon; break;
    case RCM_S_DLAT: dst = &h->h_dlat; break;
    case RCM_S_MSFU: dst = &h->h_msfu; break;
    case RCM_S_MSFV: dst = &h->h_msfv; break;
    case RCM_S_ULAT: dst = &h->h_ulat; break;
    case RCM_S_VLAT: dst = &h->h_vlat; break;
    case RCM_S_SIGMA:
      for (int k = lb[2]; k <= ub[2] && k <= h->g.kzp1; ++k) if (k >= 1) h->h_sigma[k] = a[k - lb[2]];
      return RCM_OK;
    default: return fail(h, RCM_ERR_ARG, "bad static id");
  }
  if (dst->empty()) return RCM_OK;   // a static field the selected core does not use
  int32_t l3[3] = {lb[0], lb[1], 1}, u3[3] = {ub[0], ub[1], 1};
  win_in(h, *dst, h->off2, 1, a, l3, u3, 1);
  return RCM_OK;
}

static int bdy_index(int field) {
  switch (field) {
    case RCM_F_U: return 0; case RCM_F_V: return 1; case RCM_F_T: return 2;
    case RCM_F_QV: return 3; case RCM_F_PP: return 4; case RCM_F_W: return 5;
  }
  return -1;
}
static double** bdy_b0(rcm_dyn* h, int q) {
  Fields& f = h->f;
  double** t[6] = {&f.xub0, &f.xvb0, &f.xtb0, &f.xqb0, &f.xpb0, &f.xwb0};
  return t[q];
}
static double** bdy_bt(rcm_dyn* h, int q) {
  Fields& f = h->f;
  double** t[6] = {&f.xubt, &f.xvbt, &f.xtbt, &f.xqbt, &f.xpbt, &f.xwbt};
  return t[q];
}

#ifndef RCM_EMU
// strided 3-D copy between a caller window (Fortran bounds lb/ub) and the
// padded device geometry: one cudaMemcpy3DAsync, no host staging.
static int copy3d(rcm_dyn* h, double* dev_shifted, const double* host, const int32_t* lb, const int32_t* ub, int nk, bool to_device,
                  cudaStream_t st = nullptr, bool sync = true) {
  if (!st) st = h->st;
  const Geom& g = h->g;
  const int jlo = std::max(lb[0], g.jde1 - RCM_HALO), jhi = std::min(ub[0], g.jde2 + RCM_HALO);
  const int ilo = std::max(lb[1], g.ide1 - RCM_HALO), ihi = std::min(ub[1], g.ide2 + RCM_HALO);
  const int klo = std::max(lb[2], 1), khi = std::min(ub[2], nk);
  if (jhi < jlo || ihi < ilo || khi < klo) return RCM_OK;
  const size_t n1 = (size_t)(ub[0] - lb[0] + 1), n2 = (size_t)(ub[1] - lb[1] + 1);
  cudaMemcpy3DParms p;
  std::memset(&p, 0, sizeof(p));
  cudaPitchedPtr hp = make_cudaPitchedPtr((void*)host, n1 * sizeof(double), n1, n2);
  cudaPitchedPtr dp = make_cudaPitchedPtr((void*)(dev_shifted - h->off3), (size_t)g.si * sizeof(double), (size_t)g.si, (size_t)h->ni_pad);
  cudaPos hpos = make_cudaPos((size_t)(jlo - lb[0]) * sizeof(double), (size_t)(ilo - lb[1]), (size_t)(klo - lb[2]));
  cudaPos dpos = make_cudaPos((size_t)(jlo - (g.jde1 - g.hw)) * sizeof(double), (size_t)(ilo - (g.ide1 - g.hw)), (size_t)(klo - 1));
  p.srcPtr = to_device ? hp : dp; p.srcPos = to_device ? hpos : dpos;
  p.dstPtr = to_device ? dp : hp; p.dstPos = to_device ? dpos : hpos;
  p.extent = make_cudaExtent((size_t)(jhi - jlo + 1) * sizeof(double), (size_t)(ihi - ilo + 1), (size_t)(khi - klo + 1));
  p.kind = to_device ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToHost;
  CUDA_OK(cudaMemcpy3DAsync(&p, st));
  if (sync) CUDA_OK(cudaStreamSynchronize(st));
  return RCM_OK;
}
#endif

static int upload3(rcm_dyn* h, double* dev_shifted, const double* a, const int32_t* lb, const int32_t* ub, int nk) {
#ifndef RCM_EMU
  return copy3d(h, dev_shifted, a, lb, ub, nk, true);
#else
  CUDA_OK(cudaMemcpy(h->stage.data(), dev_shifted - h->off3, (size_t)h->g.sk * nk * sizeof(double), cudaMemcpyDeviceToHost));
  win_in(h, h->stage, h->off2, nk, a, lb, ub, 1);  // stage plane 0 is k = 1
  CUDA_OK(cudaMemcpy(dev_shifted - h->off3, h->stage.data(), (size_t)h->g.sk * nk * sizeof(double), cudaMemcpyHostToDevice));
  return RCM_OK;
#endif
}

// 2-D p* records (idynamic = 1): a 2-D shifted pointer p is the 3-D shifted pointer p - sk with k = 1
static double* ps_slot(rcm_dyn* h, int slot) {
  if (slot == RCM_BDY_B0) return h->f.xpsb0;
  if (slot == RCM_BDY_B1) return h->xps_b1;
  if (slot == RCM_BDY_NEXT) return h->xps_nx;
  return nullptr;
}

extern "C" int rcm_dyn_set_bdy(rcm_dyn* h, int field, int slot, const double* a, const int32_t lb[3], const int32_t ub[3]) {
  if (!h || !a) return RCM_ERR_ARG;
  if (h->pending) { const int rcp_ = finish_steps(h); if (rcp_) return rcp_; }
  if (field == RCM_F_PS) {
    if (!h->hydro && !h->moloch) return fail(h, RCM_ERR_ARG, "ps boundary data exist for idynamic = 1 and 3 only");
    double* dst2 = ps_slot(h, slot);
    if (!dst2) return fail(h, RCM_ERR_ARG, "bad slot");
    cudaSetDevice(h->device);
    if (slot == RCM_BDY_NEXT) h->have_next = true;
    int32_t l3[3] = {lb[0], lb[1], 1}, u3[3] = {ub[0], ub[1], 1};
    return upload3(h, dst2 - h->g.sk, a, l3, u3, 1);
  }
  if (field >= RCM_F_CHI) {   // tracer record (mixing ratio; coupled at init / bdyin)
    const int n = field - RCM_F_CHI;
    if (n >= h->tr.ntr) return fail(h, RCM_ERR_ARG, "no such tracer");
    cudaSetDevice(h->device);
    double* dstc = slot == RCM_BDY_B0 ? h->tr.b0[n] : slot == RCM_BDY_B1 ? h->tr_b1[n] : slot == RCM_BDY_NEXT ? h->tr_nx[n] : nullptr;
    if (!dstc) return fail(h, RCM_ERR_ARG, "bad slot");
    if (slot == RCM_BDY_NEXT) h->have_next = true;
    return upload3(h, dstc, a, lb, ub, h->g.kz);
  }
  const int q = bdy_index(field);
  if (q < 0 || (h->moloch && q > 3)) return fail(h, RCM_ERR_ARG, "no boundary data for this field");
  cudaSetDevice(h->device);
  double* dst = nullptr;
  if (slot == RCM_BDY_B0) dst = *bdy_b0(h, q);
  else if (slot == RCM_BDY_B1) dst = h->b1[q];
  else if (slot == RCM_BDY_NEXT) { dst = h->nx[q]; h->have_next = true; }
  else return fail(h, RCM_ERR_ARG, "bad slot");
  const int nk = (field == RCM_F_W) ? h->g.kzp1 : h->g.kz;
  return upload3(h, dst, a, lb, ub, nk);
}

// Asynchronous staging of the NEXT boundary record: the copy runs on its own
// stream while rcm_dyn_step keeps computing; bdyin waits for it on the device.
extern "C" int rcm_dyn_stage_bdy_async(rcm_dyn* h, int field, const double* a, const int32_t lb[3], const int32_t ub[3]) {
  if (!h || !a) return RCM_ERR_ARG;
  const bool is_ps = (field == RCM_F_PS && (h->hydro || h->moloch));
  const bool is_tr = (field >= RCM_F_CHI && field - RCM_F_CHI < h->tr.ntr);
  const int q = (is_ps || is_tr) ? 0 : bdy_index(field);
  if (q < 0 || (h->moloch && q > 3)) return fail(h, RCM_ERR_ARG, "no boundary data for this field");
  cudaSetDevice(h->device);
  const int nk = is_ps ? 1 : ((field == RCM_F_W) ? h->g.kzp1 : h->g.kz);
  double* const dst_async = is_ps ? h->xps_nx - h->g.sk : is_tr ? h->tr_nx[field - RCM_F_CHI] : h->nx[q];
  int32_t lb3[3] = {lb[0], lb[1], is_ps ? 1 : lb[2]}, ub3[3] = {ub[0], ub[1], is_ps ? 1 : ub[2]};
  lb = lb3; ub = ub3;
#ifndef RCM_EMU
  if (!h->st_copy) {
    CUDA_OK(cudaStreamCreateWithFlags(&h->st_copy, cudaStreamNonBlocking));
    CUDA_OK(cudaEventCreateWithFlags(&h->ev_stage, cudaEventDisableTiming));
  }
  if (h->pending) {   // steps enqueued by rcm_dyn_step_async may still read the buffer this copy overwrites
    CUDA_OK(cudaEventRecord(h->ev_fork, h->st));
    CUDA_OK(cudaStreamWaitEvent(h->st_copy, h->ev_fork, 0));
  }
  {
    const Geom& g = h->g;
    const int jlo = std::max(lb[0], g.jde1 - RCM_HALO), jhi = std::min(ub[0], g.jde2 + RCM_HALO);
    const int ilo = std::max(lb[1], g.ide1 - RCM_HALO), ihi = std::min(ub[1], g.ide2 + RCM_HALO);
    const int klo = std::max(lb[2], 1), khi = std::min(ub[2], nk);
    if (jhi >= jlo && ihi >= ilo && khi >= klo) {
      // ONE contiguous H2D copy of the planes klo..khi of the caller's array (whatever its window) into a landing
      // buffer; the unpack kernel scatters the part that falls on this rank's padded tile
      const size_t n1 = (size_t)(ub[0] - lb[0] + 1), n2 = (size_t)(ub[1] - lb[1] + 1), nkk = (size_t)(khi - klo + 1);
      const size_t need = n1 * n2 * nkk;
      if (need > h->stage_cap) {
        if (h->stage_dev) { CUDA_OK(cudaStreamSynchronize(h->st_copy)); CUDA_OK(cudaFree(h->stage_dev)); h->stage_dev = nullptr; }
        const size_t cap = std::max(need, (size_t)g.sk * (size_t)g.kzp1);
        CUDA_OK(cudaMalloc((void**)&h->stage_dev, cap * sizeof(double)));
        h->stage_cap = cap;
      }
      CUDA_OK(cudaMemcpyAsync(h->stage_dev, a + n1 * n2 * (size_t)(klo - lb[2]), need * sizeof(double), cudaMemcpyHostToDevice, h->st_copy));
      // element (j,i,k) of the window sits at (j-lb0) + n1*((i-lb1) + n2*(k-klo)): shift the buffer so that the
      // kernel's own origin (jlo,ilo,klo) with row pitch n1 lands there
      launch_window_strided(false, h->stage_dev + (size_t)(jlo - lb[0]) + n1 * (size_t)(ilo - lb[1]), dst_async, jlo, ilo, klo,
                            jhi - jlo + 1, ihi - ilo + 1, (int)nkk, (long long)n1, (long long)(n1 * n2), g.si, g.sk, h->st_copy);
      h->launches += 1;
    } else {
      int rc = copy3d(h, dst_async, a, lb, ub, nk, true, h->st_copy, false);
      if (rc) return rc;
    }
  }
  CUDA_OK(cudaEventRecord(h->ev_stage, h->st_copy));
  h->stage_pending = true;
#else
  int rc = upload3(h, dst_async, a, lb, ub, nk);
  if (rc) return rc;
#endif
  h->have_next = true;
  return RCM_OK;
}

extern "C" int rcm_dyn_stage_wait(rcm_dyn* h) {
  if (!h) return RCM_ERR_ARG;
#ifndef RCM_EMU
  if (h->stage_pending) { cudaSetDevice(h->device); CUDA_OK(cudaStreamSynchronize(h->st_copy)); h->stage_pending = false; }
#endif
  return RCM_OK;
}

// ---------------------------------------------------------------------------
// helpers for the static setup
static int up2(rcm_dyn* h, double* dev_shifted, const std::vector<double>& v) {
  CUDA_OK(cudaMemcpy(dev_shifted - h->off2, v.data(), (size_t)h->g.sk * sizeof(double), cudaMemcpyHostToDevice));
  return RCM_OK;
}
static int up3(rcm_dyn* h, double* dev_shifted, const std::vector<double>& v, int nk) {
  CUDA_OK(cudaMemcpy(dev_shifted - h->off3, v.data(), (size_t)h->g.sk * nk * sizeof(double), cudaMemcpyHostToDevice));
  return RCM_OK;
}
static int up1(rcm_dyn* h, double* dev, const std::vector<double>& v) {
  CUDA_OK(cudaMemcpy(dev, v.data(), v.size() * sizeof(double), cudaMemcpyHostToDevice));
  return RCM_OK;
}

// relaxation-zone distance index (setup_boundaries, Main/mod_atm_interface.F90:383-542)
static void build_zone(rcm_dyn* h, bool dot, std::vector<int>& z, std::vector<int>* we = nullptr) {
  const Geom& g = h->g;
  const rcm_cfg& c = h->cfg;
  z.assign((size_t)g.sk, 0);
  if (we) we->assign((size_t)g.sk, 0);   // 1 where the point belongs to the west / east zone
  const int jx = g.jx, iy = g.iy;
  const int icx = dot ? 0 : 1, icy = dot ? 0 : 1;
  const int nsp = dot ? c.nspgd : c.nspgx;
  const int igbb1 = 2, igbb2 = nsp - 1, jgbl1 = 2, jgbl2 = nsp - 1;
  const int igbt1 = iy - icy - nsp + 2, igbt2 = iy - icy - 1;
  const int jgbr1 = jx - icx - nsp + 2, jgbr2 = jx - icx - 1;
  // the zones are computed on this rank's dot extent (jde1:jde2,ide1:ide2) and, being a function of the global
  // indices only, on its ghost rows (k_h_prep relaxes the p* tendency on the first ghost ring)
  for (int i = std::max(1, g.ide1 - RCM_HALO); i <= std::min(iy, g.ide2 + RCM_HALO); ++i)
    for (int j = std::max(1, g.jde1 - RCM_HALO); j <= std::min(jx, g.jde2 + RCM_HALO); ++j) {
      int v = 0, t = 0;
      if (i >= igbb1 && i <= igbb2 && j >= jgbl1 && j <= jgbr2 && !(j <= jgbl2 && i >= j) && !(j >= jgbr1 && i >= (jgbr2 - j + 2)))
        v = i - igbb1 + 2;  // south
      if (i >= igbt1 && i <= igbt2 && j >= jgbl1 && j <= jgbr2 && !(j <= jgbl2 && (igbt2 - i + 2) >= j) && !(j >= jgbr1 && (igbt2 - i) >= (jgbr2 - j)))
        v = igbt2 - i + 2;  // north
      if (i >= igbb1 && i <= igbt2 && j >= jgbl1 && j <= jgbl2 && !(i < igbb2 && j > i) && !(i > igbt1 && j > (igbt2 - i + 2)))
        { v = j - jgbl1 + 2; t = 1; }  // west
      if (i >= igbb1 && i <= igbt2 && j >= jgbr1 && j <= jgbr2 && !(i < igbb2 && (jgbr2 - j + 2) > i) && !(i > igbt1 && (jgbr2 - j) > (igbt2 - i)))
        { v = jgbr2 - j + 2; t = 1; }  // east
      z[h->off2 + j + (long long)i * g.si] = v;
      if (we) (*we)[h->off2 + j + (long long)i * g.si] = t;
    }
}

static double tau_fn(const rcm_cfg& c, double z, double zmax) {  // Main/mod_bdycod.F90:5115-5123
  if (z > zmax - c.rayhd) {
    const double s = std::sin(halfpi * (1.0 - (zmax - z) / c.rayhd));
    return c.rayalpha0 * (s * s);
  }
  return 0.0;
}

static int bdyval(rcm_dyn* h, const double* oat, const double* oaqv, const double* oaqc);
static int init_hydro(rcm_dyn* h);

// tracers ride along in the exchanges of the moisture species (Main/mod_tendency.F90:570,873-874,1085-1090)
static void tr_add(rcm_dyn* h, std::vector<HaloField>& fl, bool state) {
  const TrFields& tr = h->tr;
  for (int n = 0; n < tr.ntr; ++n) {
    if (state) { fl.push_back({tr.a[n], h->g.kz, 0}); fl.push_back({tr.b[n], h->g.kz, 0}); }
    else fl.push_back({tr.c[n], h->g.kz, 0});
  }
}
// exchange call sites (the reference's are listed in SURVEY.md §2.4)
static int exch_sound_pp(rcm_dyn* h) {  // atmc%t, atmc%pp  (Main/mod_sound.F90:265-266)
  if (h->cfg.nproc <= 1) return RCM_OK;
  const int kz = h->g.kz;
  if (h->halo.exchange({{h->f.cppd, kz, 0}, {h->f.ct, kz, 0}}, 1, h->st)) return fail(h, RCM_ERR_NCCL, "halo exchange (pp) failed");
  return RCM_OK;
}
// The reference exchanges atmc%u, atmc%v in every sub-step (Main/mod_sound.F90:300-301).
// Here the acoustic u,v tendencies are exchanged ONCE per step instead and k_snd_uv
// advances the first high-side ghost row itself (bit-identical to the owner's values).
static int exch_uvten(rcm_dyn* h) {
  if (h->cfg.nproc <= 1) return RCM_OK;
  const int kz = h->g.kz;
  // + the unfiltered moisture forecast atmc%qx for the negative-value repair (Main/mod_tendency.F90:381)
  std::vector<HaloField> fl = {{h->f.cppd, kz, 0}, {h->f.ct, kz, 0}, {h->f.uten, kz, 1}, {h->f.vten, kz, 1}, {h->f.cqv, kz, 0}, {h->f.cqc, kz, 0}};
  tr_add(h, fl, false);
  if (h->halo.exchange(fl, 1, h->st))
    return fail(h, RCM_ERR_NCCL, "halo exchange (uten, pp, qx forecast) failed");
  return RCM_OK;
}
static int exch_state(rcm_dyn* h) {  // atm1/atm2 all prognostics, width 2 (Main/mod_tendency.F90:866-869,1046-1083)
  if (h->cfg.nproc <= 1) return RCM_OK;
  const int kz = h->g.kz, kzp1 = h->g.kzp1;
  Fields& f = h->f;
  std::vector<HaloField> fl = {{f.au, kz, 1}, {f.av, kz, 1}, {f.at, kz, 0}, {f.aqv, kz, 0}, {f.aqc, kz, 0}, {f.app, kz, 0}, {f.aw, kzp1, 0},
                               {f.bu, kz, 1}, {f.bv, kz, 1}, {f.bt, kz, 0}, {f.bqv, kz, 0}, {f.bqc, kz, 0}, {f.bpp, kz, 0}, {f.bw, kzp1, 0}};
  tr_add(h, fl, true);
  if (h->uw_on) { fl.push_back({h->uw.a, h->g.kzp1, 0}); fl.push_back({h->uw.b, h->g.kzp1, 0}); }   // Main/mod_tendency.F90:870-872,1078-1080
  if (h->halo.exchange(fl, RCM_HALO, h->st))
    return fail(h, RCM_ERR_NCCL, "halo exchange (state) failed");
  return RCM_OK;
}
// Semi-Lagrangian advection on a decomposed domain: the departure stencil reads atm2%qx / atm2%chi three cells out
// (Main/mod_sladvection.F90:139-221,347-393; the reference allocates those two with the `sl` ghost width,
// Main/mod_atm_interface.F90:650-668, and exchanges them four wide, Main/mod_tendency.F90:1073-1077,1085-1089), so the tile is
// allocated with g.hw = 4 ghost rows and the level-b species travel once more per step, three rows wide.
static int exch_sl(rcm_dyn* h) {
  if (h->cfg.nproc <= 1) return RCM_OK;
  std::vector<HaloField> fl = {{h->f.bqv, h->g.kz, 0}, {h->f.bqc, h->g.kz, 0}};
  for (int n = 0; n < h->tr.ntr; ++n) fl.push_back({h->tr.b[n], h->g.kz, 0});
  if (h->halo.exchange(fl, 3, h->st))
    return fail(h, RCM_ERR_NCCL, "halo exchange (semi-Lagrangian species) failed");
  return RCM_OK;
}
// ---- hydrostatic core (the reference's call sites: Main/mod_tendency.F90:866-869, Main/mod_split.F90:197-198,
//      276-277,312-313,438,499,527-528,583) ----
static int exch_h_psdot(rcm_dyn* h) {   // dot-point p* tables of both time levels after psc2psd
  if (h->cfg.nproc <= 1) return RCM_OK;
  const Geom& g = h->g; Fields& f = h->f;
  if (h->halo.exchange({{f.psdota - g.sk, 1, 1}, {f.rpsda - g.sk, 1, 1}, {f.psdotb - g.sk, 1, 1}, {f.rpsdb - g.sk, 1, 1}}, RCM_HALO, h->st))
    return fail(h, RCM_ERR_NCCL, "halo exchange (psdot) failed");
  return RCM_OK;
}
static int exch_h_new(rcm_dyn* h) {   // forecast and filtered winds for splitf's divergences + the moisture forecast for the repair
  if (h->cfg.nproc <= 1) return RCM_OK;
  const int kz = h->g.kz; Fields& f = h->f;
  std::vector<HaloField> fl = {{f.nau, kz, 1}, {f.nav, kz, 1}, {f.nbu, kz, 1}, {f.nbv, kz, 1}, {f.cqv, kz, 0}, {f.cqc, kz, 0}};
  tr_add(h, fl, false);
  if (h->halo.exchange(fl, 1, h->st))
    return fail(h, RCM_ERR_NCCL, "halo exchange (new winds, qx forecast) failed");
  return RCM_OK;
}
static int exch_h_modal(rcm_dyn* h, double* p, int nplanes) {   // delh of one time slot / dhsum: nsplit planes, one row
  if (h->cfg.nproc <= 1) return RCM_OK;
  if (h->halo.exchange({{p, nplanes, 0}}, 1, h->st)) return fail(h, RCM_ERR_NCCL, "halo exchange (split-explicit modes) failed");
  return RCM_OK;
}
static int exch_h_state(rcm_dyn* h) {   // atm1/atm2 prognostics and p* of both time levels, width RCM_HALO
  if (h->cfg.nproc <= 1) return RCM_OK;
  const Geom& g = h->g; Fields& f = h->f;
  const int kz = g.kz;
  std::vector<HaloField> fl = {{f.au, kz, 1}, {f.av, kz, 1}, {f.at, kz, 0}, {f.aqv, kz, 0}, {f.aqc, kz, 0}, {f.psa - g.sk, 1, 0},
                               {f.bu, kz, 1}, {f.bv, kz, 1}, {f.bt, kz, 0}, {f.bqv, kz, 0}, {f.bqc, kz, 0}, {f.psb - g.sk, 1, 0}};
  tr_add(h, fl, true);
  if (h->uw_on) { fl.push_back({h->uw.a, h->g.kzp1, 0}); fl.push_back({h->uw.b, h->g.kzp1, 0}); }   // Main/mod_tendency.F90:870-872,1078-1080
  if (h->halo.exchange(fl, RCM_HALO, h->st))
    return fail(h, RCM_ERR_NCCL, "halo exchange (state) failed");
  return RCM_OK;
}
static int exch_estore(rcm_dyn* h) {
  if (h->cfg.nproc <= 1) return RCM_OK;
  if (h->halo.exchange_estore(h->f, h->st)) return fail(h, RCM_ERR_NCCL, "halo exchange (estore) failed");
  return RCM_OK;
}

// ---- tracers: tables, boundary records, initial state; called by rcm_dyn_init before the first bdyval ----
static void tr_extents(rcm_dyn* h, int& jclo, int& jchi, int& iclo, int& ichi) {
  const Geom& g = h->g;
  jclo = std::max(1, g.jde1 - RCM_HALO); jchi = std::min(g.jx - 1, g.jde2 + RCM_HALO);
  iclo = std::max(1, g.ide1 - RCM_HALO); ichi = std::min(g.iy - 1, g.ide2 + RCM_HALO);
}
// atm1%tke = atm2%tke = tkemin (Main/mod_init.F90:378-380)
static int init_tke(rcm_dyn* h) {
  if (!h->uw_on) return RCM_OK;
  std::vector<double> fill((size_t)h->g.sk * h->g.kzp1, h->uw.tkemin);
  CUDA_OK(cudaMemcpy(h->uw.a - h->off3, fill.data(), fill.size() * sizeof(double), cudaMemcpyHostToDevice));
  CUDA_OK(cudaMemcpy(h->uw.b - h->off3, fill.data(), fill.size() * sizeof(double), cudaMemcpyHostToDevice));
  return RCM_OK;
}
static int init_tracers(rcm_dyn* h) {
  TrFields& tr = h->tr;
  if (tr.ntr <= 0) return RCM_OK;
  const Geom& g = h->g;
  const rcm_cfg& c = h->cfg;
  const int kz = g.kz;
  // setup_che_bdycon (Main/chemlib/mod_che_bdyco.F90:1177-1217): exponential profile, dt2 = 2*dtsec
  {
    const double chdt2 = 2.0 * h->dtsec;
    const double fn = (c.bdy_nm > 0.0) ? c.bdy_nm : 0.1 / chdt2;
    const double gn = (c.bdy_dm > 0.0) ? c.bdy_dm : 1.0 / (50.0 * chdt2);
    std::vector<double> ef((size_t)c.nspgx * kz + 1, 0.0), eg((size_t)c.nspgx * kz + 1, 0.0);
    for (int k = 1; k <= kz; ++k) {
      const double hs = (h->h_sigma[k + 1] + h->h_sigma[k]) * 0.5;
      const double an = hs < 0.4 ? c.high_nudge : (hs < 0.8 ? c.medium_nudge : c.low_nudge);
      for (int n = 2; n <= c.nspgx - 1; ++n) {
        const double xfune = std::exp(-(double)(n - 2) / an);
        ef[(size_t)n + (size_t)c.nspgx * (k - 1)] = fn * xfune; eg[(size_t)n + (size_t)c.nspgx * (k - 1)] = gn * xfune;
      }
    }
    CUDA_OK(cudaMemcpy((void*)tr.cefc, ef.data(), ef.size() * sizeof(double), cudaMemcpyHostToDevice));
    CUDA_OK(cudaMemcpy((void*)tr.cegc, eg.data(), eg.size() * sizeof(double), cudaMemcpyHostToDevice));
    std::vector<int> z, we;
    build_zone(h, false, z, &we);
    CUDA_OK(cudaMemcpy((void*)(tr.ztype - h->off2), we.data(), we.size() * sizeof(int), cudaMemcpyHostToDevice));
  }
  // che_init_bdy (:257-290): couple with xpsb%b0 / xpsb%b1, chibt = (chib1 - chib0)/dtbdys
  int jclo, jchi, iclo, ichi;
  tr_extents(h, jclo, jchi, iclo, ichi);
  const double* p0 = h->hydro ? h->f.xpsb0 : h->xpsb;
  const double* p1 = h->hydro ? h->xps_b1 : h->xpsb;
  const size_t nb = (size_t)g.sk * g.kzp1 * sizeof(double);
  for (int n = 0; n < tr.ntr; ++n) {
    launch_couple(g, tr.b0[n], p0, kz, jclo, jchi, iclo, ichi, h->st);
    launch_couple(g, h->tr_b1[n], p1, kz, jclo, jchi, iclo, ichi, h->st);
    launch_tr_timeint(g, h->tr_b1[n], tr.b0[n], tr.bt[n], h->dtbdys, kz, jclo, jchi, iclo, ichi, h->st);
    CUDA_OK(cudaMemcpyAsync(tr.a[n] - h->off3, tr.b0[n] - h->off3, nb, cudaMemcpyDeviceToDevice, h->st));
    CUDA_OK(cudaMemcpyAsync(tr.b[n] - h->off3, tr.b0[n] - h->off3, nb, cudaMemcpyDeviceToDevice, h->st));
    h->launches += 3;
  }
  return RCM_OK;
}
// chem_bdyin (Main/chemlib/mod_che_bdyco.F90:326-413), after the meteorological bdyin
static int bdyin_tracers(rcm_dyn* h) {
  TrFields& tr = h->tr;
  if (tr.ntr <= 0) return RCM_OK;
  const Geom& g = h->g;
  int jclo, jchi, iclo, ichi;
  tr_extents(h, jclo, jchi, iclo, ichi);
  const double* p1 = h->hydro ? h->xps_b1 : h->xpsb;
  for (int n = 0; n < tr.ntr; ++n) {
    std::swap(tr.b0[n], h->tr_b1[n]);
    std::swap(h->tr_b1[n], h->tr_nx[n]);
    launch_couple(g, h->tr_b1[n], p1, g.kz, jclo, jchi, iclo, ichi, h->st);
    launch_tr_timeint(g, h->tr_b1[n], tr.b0[n], tr.bt[n], h->dtbdys, g.kz, jclo, jchi, iclo, ichi, h->st);
    h->launches += 2;
  }
  return RCM_OK;
}

extern "C" int rcm_dyn_init(rcm_dyn* h) {
  if (!h) return RCM_ERR_ARG;
  cudaSetDevice(h->device);
  if (h->moloch) return mo_init(h);
  const Geom& g = h->g;
  const rcm_cfg& c = h->cfg;
  Fields& f = h->f;
  const int kz = g.kz, kzp1 = g.kzp1;
  const size_t n2 = (size_t)g.sk;
  const double dx = c.ds * 1000.0, dx4 = 4.0 * dx, dx8 = 8.0 * dx, dx16 = 16.0 * dx, dxsq = dx * dx;
  const double ptop = c.ptop;
  // valid (inside the global domain) part of the padded tile
  const int jdlo = std::max(1, g.jde1 - RCM_HALO), jdhi = std::min(g.jx, g.jde2 + RCM_HALO);
  const int idlo = std::max(1, g.ide1 - RCM_HALO), idhi = std::min(g.iy, g.ide2 + RCM_HALO);
  const int jclo = jdlo, jchi = std::min(g.jx - 1, g.jde2 + RCM_HALO);
  const int iclo = idlo, ichi = std::min(g.iy - 1, g.ide2 + RCM_HALO);

  // ---- vertical tables (Main/mod_params.F90:2005-2010,2208-2215; mod_advection.F90:77-107) ----
  const size_t nv = (size_t)kz + 3;
  std::vector<double> sigma = h->h_sigma, hsigma(nv, 0.0), dsigma(nv, 0.0), twt1(nv, 0.0), twt2(nv, 0.0), qcon(nv, 0.0),
                      xds(nv, 0.0), dds(nv, 0.0);
  for (int k = 1; k <= kz; ++k) { hsigma[k] = (sigma[k + 1] + sigma[k]) * 0.5; dsigma[k] = (sigma[k + 1] - sigma[k]); }
  for (int k = 2; k <= kz; ++k) {
    twt1[k] = (sigma[k] - hsigma[k - 1]) / (hsigma[k] - hsigma[k - 1]);
    twt2[k] = 1.0 - twt1[k];
    qcon[k] = (sigma[k] - hsigma[k]) / (hsigma[k - 1] - hsigma[k]);
  }
  for (int k = 1; k <= kz; ++k) xds[k] = 1.0 / dsigma[k];
  for (int k = 2; k <= kz; ++k) dds[k] = 1.0 / (dsigma[k] + dsigma[k - 1]);
  int rc;
  if ((rc = up1(h, f.sigma, sigma))) return rc;
  if ((rc = up1(h, f.dsigma, dsigma))) return rc;
  if ((rc = up1(h, f.twt1, twt1))) return rc;
  if ((rc = up1(h, f.twt2, twt2))) return rc;
  if ((rc = up1(h, f.qcon, qcon))) return rc;
  if ((rc = up1(h, f.xds, xds))) return rc;
  if ((rc = up1(h, f.dds, dds))) return rc;

  // ---- map factors (stored inverted), xmsf, dmsf, geopotential ht (:1981-2000) ----
  std::vector<double> ht(n2, 0.0), msfx(n2, 1.0), msfd(n2, 1.0), xmsf(n2, 0.0), dmsf(n2, 0.0);
  for (int i = idlo; i <= idhi; ++i)
    for (int j = jdlo; j <= jdhi; ++j) {
      H2(ht, j, i) = H2(h->h_ht, j, i) * egrav;
      H2(msfd, j, i) = 1.0 / H2(h->h_msfd, j, i);
      H2(msfx, j, i) = 1.0 / H2(h->h_msfx, j, i);
      H2(dmsf, j, i) = 1.0 / (H2(msfd, j, i) * H2(msfd, j, i) * dx16);
      H2(xmsf, j, i) = 1.0 / (H2(msfx, j, i) * H2(msfx, j, i) * dx4);
    }

  // ---- reference atmosphere (make_reference_atmosphere :2613-2694, nhbase) ----
  std::vector<double> ps0(n2, 1.0), psdot0(n2, 1.0);
  std::vector<double> pr0((size_t)g.sk * kzp1, 1.0), t0((size_t)g.sk * kzp1, 1.0), rho0((size_t)g.sk * kzp1, 1.0),
      z0((size_t)g.sk * kzp1, 0.0), rhof0((size_t)g.sk * kzp1, 1.0), zf0((size_t)g.sk * kzp1, 0.0);
  if (!h->hydro) {  // the hydrostatic core has no reference atmosphere
    const double p0 = c.base_state_pressure, tlp = c.logp_lrate, st0 = c.base_state_ts0, ptoppa = ptop * 1000.0;
    auto nhbase = [&](int kx, const std::vector<double>& sig, std::vector<double>& pr, std::vector<double>* tt,
                      std::vector<double>& rho, std::vector<double>& zz) {
      for (int i = iclo; i <= ichi; ++i)
        for (int j = jclo; j <= jchi; ++j) {
          const double ter = (H2(ht, j, i) * regrav);
          const double ac = 0.5 * govr * ter / tlp;
          const double b = st0 / tlp;
          double alnp = -b + std::sqrt(b * b - 4.0 * ac);
          const double ps = p0 * std::exp(alnp) - ptoppa;
          H2(ps0, j, i) = ps;
          for (int k = 1; k <= kx; ++k) {
            const double p = ps * sig[k] + ptoppa;
            const double t = std::max(st0 + tlp * std::log(p / p0), tiso);
            H3(pr, j, i, k) = p;
            if (tt) H3(*tt, j, i, k) = t;
            H3(rho, j, i, k) = p / rgas / t;
            alnp = std::log(p / (ps + ptoppa));
            H3(zz, j, i, k) = std::max(-(0.5 * rovg * tlp * alnp * alnp + rovg * st0 * alnp), 0.0);
          }
        }
    };
    std::vector<double> prf((size_t)g.sk * kzp1, 1.0);
    nhbase(kz, hsigma, pr0, &t0, rho0, z0);
    nhbase(kzp1, sigma, prf, nullptr, rhof0, zf0);
  }
  // Note: the reference multiplies ht by regrav and back by egrav around
  // nhbase (:2621,2634); (ht*egrav)*regrav is what `ter` holds above, and
  // mddom%ht afterwards is ((ht*egrav)*regrav)*egrav.
  if (!h->hydro)
    for (int i = idlo; i <= idhi; ++i)
      for (int j = jdlo; j <= jdhi; ++j) H2(ht, j, i) = (H2(ht, j, i) * regrav) * egrav;

  // psc2psd (Main/mpplib/mod_mppparam.F90:13811-13862) on global indices
  auto psc2psd = [&](const std::vector<double>& pc, std::vector<double>& pd) {
    const int jx = g.jx, iy = g.iy;
    for (int i = idlo; i <= idhi; ++i)
      for (int j = jdlo; j <= jdhi; ++j) {
        const bool jl = (j == 1), jr = (j == jx), ib_ = (i == 1), it_ = (i == iy);
        // cross neighbours must lie inside the padded tile
        if ((!jl && j - 1 < jclo) || (!ib_ && i - 1 < iclo) || (!jr && j > jchi) || (!it_ && i > ichi)) continue;
        double v;
        if ((jl || jr) && (ib_ || it_)) v = H2(pc, jl ? 1 : jx - 1, ib_ ? 1 : iy - 1);
        else if (it_) v = (H2(pc, j, iy - 1) + H2(pc, j - 1, iy - 1)) * 0.5;
        else if (ib_) v = (H2(pc, j, 1) + H2(pc, j - 1, 1)) * 0.5;
        else if (jl) v = (H2(pc, 1, i) + H2(pc, 1, i - 1)) * 0.5;
        else if (jr) v = (H2(pc, jx - 1, i) + H2(pc, jx - 1, i - 1)) * 0.5;
        else v = (H2(pc, j, i) + H2(pc, j, i - 1) + H2(pc, j - 1, i) + H2(pc, j - 1, i - 1)) * 0.25;
        H2(pd, j, i) = v;
      }
  };
  psc2psd(ps0, psdot0);

  std::vector<double> dpsdxm(n2, 0.0), dpsdym(n2, 0.0);
  {
    const int gjce1 = 1, gjce2 = g.jx - 1, gice1 = 1, gice2 = g.iy - 1;
    for (int i = iclo; i <= ichi; ++i)
      for (int j = jclo; j <= jchi; ++j) {
        const double den = (H2(ps0, j, i) * dx8 * H2(msfx, j, i));
        if (j == gjce1) H2(dpsdxm, j, i) = (H2(ps0, j + 1, i) - H2(ps0, j, i)) / den;
        else if (j == gjce2) H2(dpsdxm, j, i) = (H2(ps0, j, i) - H2(ps0, j - 1, i)) / den;
        else if (j - 1 >= jclo && j + 1 <= jchi) H2(dpsdxm, j, i) = (H2(ps0, j + 1, i) - H2(ps0, j - 1, i)) / den;
        if (i == gice1) H2(dpsdym, j, i) = (H2(ps0, j, i + 1) - H2(ps0, j, i)) / den;
        else if (i == gice2) H2(dpsdym, j, i) = (H2(ps0, j, i) - H2(ps0, j, i - 1)) / den;
        else if (i - 1 >= iclo && i + 1 <= ichi) H2(dpsdym, j, i) = (H2(ps0, j, i + 1) - H2(ps0, j, i - 1)) / den;
      }
  }

  // ---- full Coriolis coefficients (:2696-2741) ----
  std::vector<double> ef(n2, 0.0), ddx(n2, 0.0), ddy(n2, 0.0), dmdx(n2, 0.0), dmdy(n2, 0.0), ex(n2, 0.0), crx(n2, 0.0), cry(n2, 0.0);
  for (int i = std::max(2, idlo + 1); i <= std::min(g.iy - 1, idhi); ++i)
    for (int j = std::max(2, jdlo + 1); j <= std::min(g.jx - 1, jdhi); ++j) {
      if (j > jchi || i > ichi) continue;
      const double dlat = H2(h->h_dlat, j, i);
      double dlatdy = 0.5 * (H2(h->h_xlat, j - 1, i) + H2(h->h_xlat, j, i) - H2(h->h_xlat, j - 1, i - 1) - H2(h->h_xlat, j, i - 1));
      if (std::fabs(dlatdy) < 1.0e-8) dlatdy = std::copysign(1.0e-8, dlatdy);
      double dlondy = 0.5 * (H2(h->h_xlon, j - 1, i) + H2(h->h_xlon, j, i) - H2(h->h_xlon, j - 1, i - 1) - H2(h->h_xlon, j, i - 1));
      if (dlondy > 180.0) dlondy = dlondy - 360.0;
      if (dlondy < -180.0) dlondy = dlondy + 360.0;
      double rotang = -std::atan(dlondy / dlatdy * std::cos(degrad * dlat));
      if (dlatdy < 0.0) rotang = rotang + mathpi;
      H2(ef, j, i) = eomeg2 * std::cos(degrad * dlat);
      H2(ddx, j, i) = std::cos(rotang);
      H2(ddy, j, i) = std::sin(rotang);
      H2(dmdx, j, i) = -0.5 * (H2(msfx, j, i) + H2(msfx, j, i - 1) - H2(msfx, j - 1, i) - H2(msfx, j - 1, i - 1)) / (dx * H2(msfd, j, i) * H2(msfd, j, i));
      H2(dmdy, j, i) = -0.5 * (H2(msfx, j, i) + H2(msfx, j - 1, i) - H2(msfx, j, i - 1) - H2(msfx, j - 1, i - 1)) / (dx * H2(msfd, j, i) * H2(msfd, j, i));
    }
  for (int i = std::max(2, iclo); i <= std::min(g.iy - 2, ichi); ++i)
    for (int j = std::max(2, jclo); j <= std::min(g.jx - 2, jchi); ++j) {
      if (j + 1 > jdhi || i + 1 > idhi) continue;
      H2(ex, j, i) = 0.25 * (H2(ef, j, i) + H2(ef, j, i + 1) + H2(ef, j + 1, i) + H2(ef, j + 1, i + 1));
      H2(crx, j, i) = 0.25 * (H2(ddx, j, i) + H2(ddx, j, i + 1) + H2(ddx, j + 1, i) + H2(ddx, j + 1, i + 1));
      H2(cry, j, i) = 0.25 * (H2(ddy, j, i) + H2(ddy, j, i + 1) + H2(ddy, j + 1, i) + H2(ddy, j + 1, i + 1));
    }

  // ---- p* (constant for idynamic = 2; Main/mod_init.F90:149-155,173-178) ----
  std::vector<double> psa(n2, 1.0), rpsa(n2, 1.0), psdota(n2, 1.0), rpsda(n2, 1.0), psdot_b(n2, 1.0), xpsb(n2, 1.0);
  for (int i = iclo; i <= ichi; ++i)
    for (int j = jclo; j <= jchi; ++j) { H2(psa, j, i) = H2(ps0, j, i) * 0.001; H2(rpsa, j, i) = 1.0 / H2(psa, j, i); H2(xpsb, j, i) = H2(ps0, j, i) * 0.001; }
  psc2psd(psa, psdota);
  for (int i = idlo; i <= idhi; ++i)
    for (int j = jdlo; j <= jdhi; ++j) { H2(rpsda, j, i) = 1.0 / H2(psdota, j, i); H2(psdot_b, j, i) = H2(psdot0, j, i) * 0.001; }

  // ---- horizontal diffusion constants (initialize_diffusion, mod_diffusion.F90:93-167) ----
  StepScalars& s = h->s;
  s.xkhmax = dxsq / (64.0 * h->dtsec);
  s.dydc = c.adyndif * vonkar * vonkar * dx * 0.25;
  double xkhz;
  if (h->hydro) {
    xkhz = c.ckh * 1.5e-3 * dxsq / h->dtsec;  // RegCM3 formulation (mod_diffusion.F90:106-107)
  } else {
    xkhz = c.ckh * dx;  // Xu et al. 2001 (:108-113)
    s.xkhmax = 2.0 * s.xkhmax;
  }
  std::vector<double> hgfact(n2, xkhz);
  if (c.diffu_hgtf == 1)
    for (int i = std::max(2, iclo + 1); i <= std::min(g.iy - 2, ichi - 1); ++i)
      for (int j = std::max(2, jclo + 1); j <= std::min(g.jx - 2, jchi - 1); ++j) {
        const double hg1 = std::fabs((H2(ht, j, i) - H2(ht, j, i - 1)) / dx), hg2 = std::fabs((H2(ht, j, i) - H2(ht, j, i + 1)) / dx);
        const double hg3 = std::fabs((H2(ht, j, i) - H2(ht, j - 1, i)) / dx), hg4 = std::fabs((H2(ht, j, i) - H2(ht, j + 1, i)) / dx);
        const double hgmax = std::max(std::max(hg1, hg2), std::max(hg3, hg4)) * regrav * 1.0e3;
        H2(hgfact, j, i) = xkhz / (1.0 + hgmax * hgmax);
      }

  // ---- relaxation weights (setup_bdycon, Main/mod_bdycod.F90:183-278) ----
  {
    double fnudge = 0.0, gnudge = 0.0;
    if (c.iboudy == 1 || c.iboudy == 5) {
      fnudge = (c.bdy_nm > 0.0) ? c.bdy_nm : 0.1 / h->dt;
      gnudge = (c.bdy_dm > 0.0) ? c.bdy_dm : 1.0 / (h->dt * 50.0);
    }
    auto table = [&](int nsp, bool dot, std::vector<double>& F, std::vector<double>& G) {
      F.assign((size_t)nsp * kz + 1, 0.0); G.assign((size_t)nsp * kz + 1, 0.0);
      for (int k = 1; k <= kz; ++k) {
        double an = (hsigma[k] < 0.4) ? c.high_nudge : ((hsigma[k] < 0.8) ? c.medium_nudge : c.low_nudge);
        for (int n = 2; n <= nsp - 1; ++n) {
          double xfun;
          if (c.iboudy == 5) xfun = std::exp(-((double)(n - 2) / an));
          else xfun = (double)(nsp - n) / (double)(nsp - 2);
          F[n + (size_t)nsp * (k - 1)] = fnudge * xfun;
          G[n + (size_t)nsp * (k - 1)] = gnudge * xfun;
          if (c.iboudy == 4) {   // sponge weights (Main/mod_bdycod.F90:237-250): F = 1 - wgt(n), G = wgt(n)
            static const double wx[3] = {0.4, 0.7, 0.9}, wd[4] = {0.20, 0.55, 0.80, 0.95};
            const double w = dot ? (n - 2 < 4 ? wd[n - 2] : 1.0) : (n - 2 < 3 ? wx[n - 2] : 1.0);
            F[n + (size_t)nsp * (k - 1)] = 1.0 - w;
            G[n + (size_t)nsp * (k - 1)] = w;
          }
        }
      }
    };
    std::vector<double> F, G;
    table(c.nspgx, false, F, G);
    if ((rc = up1(h, f.hefc, F))) return rc;
    if ((rc = up1(h, f.hegc, G))) return rc;
    table(c.nspgd, true, F, G);
    if ((rc = up1(h, f.hefd, F))) return rc;
    if ((rc = up1(h, f.hegd, G))) return rc;
  }
  {
    std::vector<int> z;
    build_zone(h, false, z);
    CUDA_OK(cudaMemcpy(f.zone_c - h->off2, z.data(), n2 * sizeof(int), cudaMemcpyHostToDevice));
    build_zone(h, true, z);
    CUDA_OK(cudaMemcpy(f.zone_d - h->off2, z.data(), n2 * sizeof(int), cudaMemcpyHostToDevice));
  }

  // ---- Rayleigh coefficients tau(z) on the static reference heights ----
  {
    const int rl = std::max(1, std::min(kzp1, c.rayndamp));
    std::vector<double> tc((size_t)g.sk * rl, 0.0), td((size_t)g.sk * rl, 0.0), tf((size_t)g.sk * rl, 0.0);
    for (int k = 1; k <= rl; ++k)
      for (int i = iclo; i <= ichi; ++i)
        for (int j = jclo; j <= jchi; ++j) {
          if (k <= kz) H3(tc, j, i, k) = tau_fn(c, H3(z0, j, i, k), H3(z0, j, i, 1));
          H3(tf, j, i, k) = tau_fn(c, H3(zf0, j, i, k), H3(zf0, j, i, 1));
          if (k <= kz && j - 1 >= jclo && i - 1 >= iclo) {
            const double zz = 0.25 * (H3(z0, j, i, k) + H3(z0, j - 1, i, k) + H3(z0, j, i - 1, k) + H3(z0, j - 1, i - 1, k));
            const double zm = 0.25 * (H3(z0, j, i, 1) + H3(z0, j - 1, i, 1) + H3(z0, j, i - 1, 1) + H3(z0, j - 1, i - 1, 1));
            H3(td, j, i, k) = tau_fn(c, zz, zm);
          }
        }
    if ((rc = up3(h, f.tau_c, tc, rl))) return rc;
    if ((rc = up3(h, f.tau_d, td, rl))) return rc;
    if ((rc = up3(h, f.tau_f, tf, rl))) return rc;
  }

  // ---- acoustic constants (init_sound, Main/mod_sound.F90:115-161) ----
  {
    h->rnpts = 1.0 / (double)((g.iy - 1 - 2) * (g.jx - 1 - 2));
    for (int q = 1; q <= 5; ++q) { h->fk[q] = 2.0; h->fl[q] = 2.0; }
    h->fk[0] = h->fl[0] = h->fk[6] = h->fl[6] = 1.0;
    for (int q = 0; q < 13; ++q) { h->fi[q] = 1.0; h->fj[q] = 1.0; }
    h->fi[0] = h->fj[0] = h->fi[12] = h->fj[12] = 0.5;
    double loc = 0.0;
    for (int i = g.ici1; i <= g.ici2; ++i)
      for (int j = g.jci1; j <= g.jci2; ++j) loc += H2(msfx, j, i);
    double maxt = -1e300;
    for (int k = 1; k <= kz; ++k)
      for (int i = g.ice1; i <= g.ice2; ++i)
        for (int j = g.jce1; j <= g.jce2; ++j) maxt = std::max(maxt, H3(t0, j, i, k));
    if (c.nproc > 1) {
      if ((rc = h->halo.allreduce_host(&loc, 1, 0)) || (rc = h->halo.allreduce_host(&maxt, 1, 1)))
        return fail(h, RCM_ERR_NCCL, "allreduce failed in init (communicator not set?)");
    }
    h->xmsf_mean = loc * h->rnpts;
    s.xgamma = 1.0 / (1.0 - rovcp);
    h->cs = std::sqrt(s.xgamma * rgas * maxt);
    s.xkd_snd = c.nhxkd;
    h->dtsmax = dx / h->cs / (1.0 + s.xkd_snd);
    s.bp = (1.0 + c.nhbet) * 0.5;
    s.bm = (1.0 - c.nhbet) * 0.5;
    s.bpxbp = s.bp * s.bp;
    s.bpxbm = s.bp * s.bm;
  }

  // ---- static scalars ----
  s.dtsec = h->dtsec;
  s.ul = c.uoffc * 0.5 * h->dt / dx;
  s.dx = dx; s.dx2 = 2.0 * dx; s.dx4 = dx4; s.dx8 = dx8; s.dx16 = dx16; s.dxsq = dxsq; s.rdx = 1.0 / dx; s.rdxsq = 1.0 / dxsq;
  s.gnu1 = c.gnu1; s.gnu2 = c.gnu2; s.t_extrema = c.t_extrema; s.q_rel_extrema = c.q_rel_extrema;
  s.ptop = ptop;
  s.upstream_mode = c.upstream_mode; s.stability_enhance = c.stability_enhance; s.ipptls = c.ipptls;
  s.iboudy = c.iboudy; s.ifrayd = c.ifrayd; s.ifupr = c.ifupr; s.idiffu = c.idiffu; s.ipgf = c.ipgf; s.isladvec = c.isladvec; s.iqmsl = c.iqmsl;
  s.kt_ray = std::min(kz, c.rayndamp); s.ktf_ray = std::min(kzp1, c.rayndamp);
  s.nspgx = c.nspgx; s.nspgd = c.nspgd;
  s.hydro = h->hydro ? 1 : 0; s.nsplit = h->hydro ? c.nsplit : 0; s.integrating = 0;

  // ---- upload statics ----
  if ((rc = up2(h, f.ht, ht)) || (rc = up2(h, f.msfx, msfx)) || (rc = up2(h, f.msfd, msfd)) || (rc = up2(h, f.xmsf, xmsf)) ||
      (rc = up2(h, f.dmsf, dmsf)) || (rc = up2(h, f.coriol, h->h_coriol)) || (rc = up2(h, f.ef, ef)) || (rc = up2(h, f.ddx, ddx)) ||
      (rc = up2(h, f.ddy, ddy)) || (rc = up2(h, f.dmdx, dmdx)) || (rc = up2(h, f.dmdy, dmdy)) || (rc = up2(h, f.ex, ex)) ||
      (rc = up2(h, f.crx, crx)) || (rc = up2(h, f.cry, cry)) || (rc = up2(h, f.dpsdxm, dpsdxm)) || (rc = up2(h, f.dpsdym, dpsdym)) ||
      (rc = up2(h, f.ps0, ps0)) || (rc = up2(h, f.hgfact, hgfact)) || (rc = up2(h, f.psa, psa)) || (rc = up2(h, f.rpsa, rpsa)) ||
      (rc = up2(h, f.psdota, psdota)) || (rc = up2(h, f.rpsda, rpsda)) || (rc = up2(h, h->psdot_bdy, psdot_b)) ||
      (rc = up2(h, h->xpsb, xpsb)))
    return rc;
  if ((rc = up3(h, f.pr0, pr0, kzp1)) || (rc = up3(h, f.t0, t0, kzp1)) || (rc = up3(h, f.rho0, rho0, kzp1)) ||
      (rc = up3(h, f.rhof0, rhof0, kzp1)))
    return rc;
  // dot-point p* on the outermost ghost columns needs cross-point p* one column
  // further out than the host window reaches: take those values from the rank
  // that owns them (the reference does the same: psc2psd then exchange)
  if (c.nproc > 1) {
    if (h->halo.exchange({{f.psdota - g.sk, 1, 1}, {f.rpsda - g.sk, 1, 1}, {h->psdot_bdy - g.sk, 1, 1}}, RCM_HALO, h->st))
      return fail(h, RCM_ERR_NCCL, "halo exchange (psdot) failed");
    CUDA_OK(cudaStreamSynchronize(h->st));
  }

  // ---- rim point lists (global physical boundary rows owned by this rank) ----
  {
    std::vector<int2> rd, rcx;
    // own rim points plus their extension into the ghost rows (so that no
    // second exchange is needed after bdyval)
    for (int i = idlo; i <= idhi; ++i)
      for (int j = jdlo; j <= jdhi; ++j)
        if (j == 1 || j == g.jx || i == 1 || i == g.iy) rd.push_back(make_int2(j, i));
    for (int i = iclo; i <= ichi; ++i)
      for (int j = jclo; j <= jchi; ++j)
        if (j == 1 || j == g.jx - 1 || i == 1 || i == g.iy - 1) rcx.push_back(make_int2(j, i));
    h->nrim_d = (int)rd.size(); h->nrim_c = (int)rcx.size();
    h->rim_d = dev_alloc_raw<int2>(h, rd.size()); h->rim_c = dev_alloc_raw<int2>(h, rcx.size());
    if (!h->rim_d || !h->rim_c) return fail(h, RCM_ERR_CUDA, "rim allocation failed");
    if (!rd.empty()) CUDA_OK(cudaMemcpy(h->rim_d, rd.data(), rd.size() * sizeof(int2), cudaMemcpyHostToDevice));
    if (!rcx.empty()) CUDA_OK(cudaMemcpy(h->rim_c, rcx.data(), rcx.size() * sizeof(int2), cudaMemcpyHostToDevice));
  }

  if (h->hydro) return init_hydro(h);
  // ---- boundary records: couple with p*, tendencies (init_bdy :393-435,522-559,625-640) ----
  h->xbctime = 0.0;
  for (int q = 0; q < 6; ++q) {
    const bool dot = (q < 2);
    const int nk = (q == 5) ? kzp1 : kz;
    const double* cpl = dot ? h->psdot_bdy : h->xpsb;
    const int j1 = dot ? jdlo : jclo, j2 = dot ? jdhi : jchi, i1 = dot ? idlo : iclo, i2 = dot ? idhi : ichi;
    launch_couple(g, *bdy_b0(h, q), cpl, nk, j1, j2, i1, i2, h->st);
    launch_couple(g, h->b1[q], cpl, nk, j1, j2, i1, i2, h->st);
    launch_timeint(g, h->b1[q], *bdy_b0(h, q), *bdy_bt(h, q), h->rdtbdy, nk, j1, j2, i1, i2, h->st);
    h->launches += 3;
  }
  // ---- initial state: atm1 = atm2 = b0 (Main/mod_init.F90:96-172) ----
  {
    const size_t nb = (size_t)g.sk * kzp1 * sizeof(double);
    double* a[6] = {f.au, f.av, f.at, f.aqv, f.app, f.aw};
    double* b[6] = {f.bu, f.bv, f.bt, f.bqv, f.bpp, f.bw};
    for (int q = 0; q < 6; ++q) {
      CUDA_OK(cudaMemcpyAsync(a[q] - h->off3, *bdy_b0(h, q) - h->off3, nb, cudaMemcpyDeviceToDevice, h->st));
      CUDA_OK(cudaMemcpyAsync(b[q] - h->off3, *bdy_b0(h, q) - h->off3, nb, cudaMemcpyDeviceToDevice, h->st));
    }
    CUDA_OK(cudaMemsetAsync(f.aqc - h->off3, 0, nb, h->st));
    CUDA_OK(cudaMemsetAsync(f.bqc - h->off3, 0, nb, h->st));
    double one[169]; for (double& x : one) x = 0.0;
    CUDA_OK(cudaMemcpyAsync(f.tmask, one, sizeof(one), cudaMemcpyHostToDevice, h->st));
  }
  h->lcount = 0; h->model_time = 0;
  h->day_lastact = 0; h->day_now = 0; h->bdy_lastact = 0; h->bdy_now = 0;
  h->inited = true;
  if ((rc = init_tracers(h))) return rc;
  if ((rc = init_tke(h))) return rc;
  // first bdyval (Main/mod_regcm_interface.F90:146)
  if ((rc = bdyval(h, f.at, f.aqv, f.aqc))) return rc;
  CUDA_OK(cudaStreamSynchronize(h->st));
  return RCM_OK;
}

// ---------------------------------------------------------------------------
// hydrostatic core: tail of rcm_dyn_init
static int upload_split_tables(rcm_dyn* h) {
  const SplitTables& T = h->sp;
  const int ns = T.nsplit, kz = T.kz, kzp1 = kz + 1;
  const size_t nt = (size_t)ns * (kzp1 + 1) + 1;
  std::vector<double> zr(nt, 0.0), tau(nt, 0.0), am(nt, 0.0), zm(nt, 0.0), pdlog(nt, 0.0), eps1(nt, 0.0);
  for (int l = 1; l <= ns; ++l)
    for (int k = 1; k <= kz; ++k) {
      const size_t q = (size_t)l + (size_t)ns * (k - 1);
      zr[q] = T.zmatxr[(size_t)(l - 1) + (size_t)kz * (k - 1)];   // zmatxr(l,k)
      tau[q] = T.tau[(size_t)(l - 1) + (size_t)kz * (k - 1)];     // tau(l,k)
      am[q] = T.am[(size_t)(k - 1) + (size_t)kz * (l - 1)];       // am(k,l)
      zm[q] = T.zmatx[(size_t)(k - 1) + (size_t)kz * (l - 1)];    // zmatx(k,l)
    }
  for (size_t q = 0; q < T.pdlog.size() && q < nt; ++q) { pdlog[q] = T.pdlog[q]; eps1[q] = T.eps1[q]; }
  int rc;
  Fields& f = h->f;
  if ((rc = up1(h, f.sp_zr, zr)) || (rc = up1(h, f.sp_tau, tau)) || (rc = up1(h, f.sp_am, am)) || (rc = up1(h, f.sp_zm, zm)) ||
      (rc = up1(h, f.sp_pdlog, pdlog)) || (rc = up1(h, f.sp_eps1, eps1)))
    return rc;
  StepScalars& s = h->s;
  s.sp_pd = T.pd;
  for (int l = 1; l <= ns; ++l) {
    s.sp_aam[l - 1] = T.aam[l]; s.sp_m2[l - 1] = (int)T.aam[l] * 2; s.sp_dtau[l - 1] = T.dtau[l];
    s.sp_hbar[l - 1] = T.hbar[l]; s.sp_an[l - 1] = T.an[l];
  }
  return RCM_OK;
}

static int download2(rcm_dyn* h, const double* dev2, std::vector<double>& v) {
  v.assign((size_t)h->g.sk, 0.0);
  CUDA_OK(cudaMemcpy(v.data(), dev2 - h->off2, (size_t)h->g.sk * sizeof(double), cudaMemcpyDeviceToHost));
  return RCM_OK;
}

// dot-point p* of a boundary record on the ghost rows (its outermost row needs cross-point p* beyond the host window)
static int exch_h_bdy_psdot(rcm_dyn* h) {
  if (h->cfg.nproc <= 1) return RCM_OK;
  if (h->halo.exchange({{h->psdot_bdy - h->g.sk, 1, 1}}, RCM_HALO, h->st)) return fail(h, RCM_ERR_NCCL, "halo exchange (boundary psdot) failed");
  return RCM_OK;
}

static int init_hydro(rcm_dyn* h) {
  const Geom& g = h->g;
  const rcm_cfg& c = h->cfg;
  Fields& f = h->f;
  const int kz = g.kz;
  int rc;
  // tile + ghost rows inside the global domain (the host window)
  const int jdlo = std::max(1, g.jde1 - RCM_HALO), jdhi = std::min(g.jx, g.jde2 + RCM_HALO);
  const int idlo = std::max(1, g.ide1 - RCM_HALO), idhi = std::min(g.iy, g.ide2 + RCM_HALO);
  const int jclo = jdlo, jchi = std::min(g.jx - 1, g.jde2 + RCM_HALO);
  const int iclo = idlo, ichi = std::min(g.iy - 1, g.ide2 + RCM_HALO);
  // ---- boundary records (init_bdy, Main/mod_bdycod.F90:393-396,408-423,522-552,625-637):
  //      p* from the ICBC surface pressure, dot-point p*, coupling, tendencies ----
  h->xbctime = 0.0;
  double* ub[4] = {f.xub0, f.xvb0, f.xtb0, f.xqb0};
  for (int rec = 0; rec < 2; ++rec) {
    double* ps = rec ? h->xps_b1 : f.xpsb0;
    launch_ps_convert(g, ps, c.ptop, h->st);
    launch_psc2psd(g, ps, h->psdot_bdy, nullptr, 0, h->st);
    if ((rc = exch_h_bdy_psdot(h))) return rc;
    for (int q = 0; q < 4; ++q) {
      const bool dot = (q < 2);
      double* a = rec ? h->b1[q] : ub[q];
      launch_couple(g, a, dot ? h->psdot_bdy : ps, kz, dot ? jdlo : jclo, dot ? jdhi : jchi, dot ? idlo : iclo, dot ? idhi : ichi, h->st);
    }
    h->launches += 6;
  }
  for (int q = 0; q < 4; ++q) {
    const bool dot = (q < 2);
    launch_timeint(g, h->b1[q], *bdy_b0(h, q), *bdy_bt(h, q), h->rdtbdy, kz, dot ? jdlo : jclo, dot ? jdhi : jchi, dot ? idlo : iclo,
                   dot ? idhi : ichi, h->st);
  }
  launch_timeint(g, h->xps_b1 - g.sk, f.xpsb0 - g.sk, f.xpsbt - g.sk, h->rdtbdy, 1, jclo, jchi, iclo, ichi, h->st);
  h->launches += 5;
  // ---- initial state (init, Main/mod_init.F90:96-147,173-178) ----
  {
    const size_t nb = (size_t)g.sk * g.kzp1 * sizeof(double), nb2 = (size_t)g.sk * sizeof(double);
    double* a[4] = {f.au, f.av, f.at, f.aqv};
    double* b[4] = {f.bu, f.bv, f.bt, f.bqv};
    for (int q = 0; q < 4; ++q) {
      CUDA_OK(cudaMemcpyAsync(a[q] - h->off3, *bdy_b0(h, q) - h->off3, nb, cudaMemcpyDeviceToDevice, h->st));
      CUDA_OK(cudaMemcpyAsync(b[q] - h->off3, *bdy_b0(h, q) - h->off3, nb, cudaMemcpyDeviceToDevice, h->st));
    }
    CUDA_OK(cudaMemsetAsync(f.aqc - h->off3, 0, nb, h->st));
    CUDA_OK(cudaMemsetAsync(f.bqc - h->off3, 0, nb, h->st));
    CUDA_OK(cudaMemcpyAsync(f.psa - h->off2, f.xpsb0 - h->off2, nb2, cudaMemcpyDeviceToDevice, h->st));
    CUDA_OK(cudaMemcpyAsync(f.psb - h->off2, f.xpsb0 - h->off2, nb2, cudaMemcpyDeviceToDevice, h->st));
  }
  launch_h_ps(g, f, h->st);
  h->launches += 1;
  if ((rc = exch_h_psdot(h))) return rc;
  h->lcount = 0; h->model_time = 0;
  h->day_lastact = 0; h->day_now = 0; h->bdy_lastact = 0; h->bdy_now = 0;
  // ---- spinit (Main/mod_split.F90:75-239): vertical modes on the host, storage on the device ----
  {
    double xps = 0.0;
    std::vector<double> tbarh((size_t)kz + 1, 0.0);
    if (!c.lstand) {  // linearisation about the horizontal mean of the data (:103-129)
      std::vector<double> psa, msfx, t;
      CUDA_OK(cudaStreamSynchronize(h->st));
      if ((rc = download2(h, f.psa, psa)) || (rc = download2(h, f.msfx, msfx))) return rc;
      t.assign((size_t)g.sk * g.kzp1, 0.0);
      CUDA_OK(cudaMemcpy(t.data(), f.at - h->off3, t.size() * sizeof(double), cudaMemcpyDeviceToHost));
      const double rnpts = 1.0 / (double)((g.iy - 1) * (g.jx - 1));
      double lxps = 0.0, lxms = 0.0;
      for (int i = g.ice1; i <= g.ice2; ++i)
        for (int j = g.jce1; j <= g.jce2; ++j) {
          const double map = 1.0 / (H2(msfx, j, i) * H2(msfx, j, i));
          lxms = lxms + rnpts * map;
          lxps = lxps + rnpts * H2(psa, j, i) * map;
        }
      std::vector<double> sums((size_t)kz + 2, 0.0);   // sumall of the reference (Main/mod_split.F90:112-127)
      sums[0] = lxms; sums[1] = lxps;
      for (int k = 1; k <= kz; ++k) {
        double ltbark = 0.0;
        for (int i = g.ice1; i <= g.ice2; ++i)
          for (int j = g.jce1; j <= g.jce2; ++j) {
            const double map = 1.0 / (H2(msfx, j, i) * H2(msfx, j, i));
            ltbark = ltbark + rnpts * map * H3(t, j, i, k) / H2(psa, j, i);
          }
        sums[(size_t)k + 1] = ltbark;
      }
      if (c.nproc > 1 && h->halo.allreduce_host(sums.data(), kz + 2, 0)) return fail(h, RCM_ERR_NCCL, "allreduce failed (spinit means)");
      lxms = sums[0]; lxps = sums[1];
      xps = lxps / lxms;
      for (int k = 1; k <= kz; ++k) tbarh[k] = sums[(size_t)k + 1] / lxms;
    }
    if (rcm_split_tables(kz, c.nsplit, c.lstand, c.ptop, h->dtsec, h->h_sigma, xps, tbarh, h->sp))
      return fail(h, RCM_ERR_ARG, "vmodes: the vertical mode problem reported errors (Main/mod_vmodes.F90:435-436)");
    if ((rc = upload_split_tables(h))) return rc;
    launch_sp_init(g, f, h->s, h->st);
    h->launches += 1;
  }
  h->inited = true;
  if ((rc = init_tracers(h))) return rc;
  if ((rc = init_tke(h))) return rc;
  // first bdyval (Main/mod_regcm_interface.F90:146)
  if ((rc = bdyval(h, f.at, f.aqv, f.aqc))) return rc;
  CUDA_OK(cudaStreamSynchronize(h->st));
  return RCM_OK;
}

// ---------------------------------------------------------------------------
static int bdyval(rcm_dyn* h, const double* oat, const double* oaqv, const double* oaqc) {
  const double xt = h->xbctime + h->dt;
  if (h->hydro) TIMED(KID_BDY, 1, launch_bdyval_ps(h->g, h->f, h->rim_c, h->nrim_c, xt, h->cfg.iboudy, h->st));
  TIMED(KID_BDY, 2, launch_bdyval(h->g, h->f, oat, oaqv, oaqc, h->rim_d, h->nrim_d, h->rim_c, h->nrim_c, xt,
                                  h->lcount > 0 ? 1 : 0, h->cfg.iboudy, h->st));
  if (h->uw_on)   // TKE rim (Main/mod_bdycod.F90:1166-1307,2412-2496), after the rim winds of this bdyval are in place
    TIMED(KID_BDY, 2, launch_tke_bdy(h->g, h->f, h->uw, h->rim_c, h->nrim_c, h->lcount > 0 ? 1 : 0, h->cfg.iboudy, h->st));
  if (h->tr.ntr > 0)   // chem_bdyval (Main/mod_bdycod.F90:2558-2564)
    TIMED(KID_BDY, 1, launch_tr_bdyval(h->g, h->tr, h->rim_c, h->nrim_c, xt, h->lcount > 0 ? 1 : 0, h->lcount > 0 ? 1 : 0, h->st));
  h->xbctime = h->xbctime + h->dtsec;
  return RCM_OK;
}

// state part of bdyin (Main/mod_bdycod.F90:669-690,772-831)
static int bdyin(rcm_dyn* h) {
  const Geom& g = h->g;
  const int kz = g.kz, kzp1 = g.kzp1;
  const int jdlo = std::max(1, g.jde1 - RCM_HALO), jdhi = std::min(g.jx, g.jde2 + RCM_HALO);
  const int idlo = std::max(1, g.ide1 - RCM_HALO), idhi = std::min(g.iy, g.ide2 + RCM_HALO);
  const int jclo = jdlo, jchi = std::min(g.jx - 1, g.jde2 + RCM_HALO);
  const int iclo = idlo, ichi = std::min(g.iy - 1, g.ide2 + RCM_HALO);
  h->xbctime = 0.0;
#ifndef RCM_EMU
  if (h->stage_pending) { CUDA_OK(cudaStreamWaitEvent(h->st, h->ev_stage, 0)); h->stage_pending = false; }
#endif
  if (h->hydro) {  // Main/mod_bdycod.F90:686,774-787,808-825
    Fields& f = h->f;
    std::swap(f.xpsb0, h->xps_b1);
    std::swap(h->xps_b1, h->xps_nx);
    launch_ps_convert(g, h->xps_b1, h->cfg.ptop, h->st);
    launch_psc2psd(g, h->xps_b1, h->psdot_bdy, nullptr, 0, h->st);
    { const int rcx = exch_h_bdy_psdot(h); if (rcx) return rcx; }
    for (int q = 0; q < 4; ++q) {
      std::swap(*bdy_b0(h, q), h->b1[q]);
      std::swap(h->b1[q], h->nx[q]);
      const bool dot = (q < 2);
      const int j1 = dot ? jdlo : jclo, j2 = dot ? jdhi : jchi, i1 = dot ? idlo : iclo, i2 = dot ? idhi : ichi;
      launch_couple(g, h->b1[q], dot ? h->psdot_bdy : h->xps_b1, kz, j1, j2, i1, i2, h->st);
      launch_timeint(g, h->b1[q], *bdy_b0(h, q), *bdy_bt(h, q), h->rdtbdy, kz, j1, j2, i1, i2, h->st);
    }
    launch_timeint(g, h->xps_b1 - g.sk, f.xpsb0 - g.sk, f.xpsbt - g.sk, h->rdtbdy, 1, jclo, jchi, iclo, ichi, h->st);
    h->launches += 11;
    h->have_next = false;
    return RCM_OK;
  }
  for (int q = 0; q < 6; ++q) {
    // b0 <- b1 (pointer swap), b1 <- staged record
    std::swap(*bdy_b0(h, q), h->b1[q]);
    std::swap(h->b1[q], h->nx[q]);
    const bool dot = (q < 2);
    const int nk = (q == 5) ? kzp1 : kz;
    const double* cpl = dot ? h->psdot_bdy : h->xpsb;
    const int j1 = dot ? jdlo : jclo, j2 = dot ? jdhi : jchi, i1 = dot ? idlo : iclo, i2 = dot ? idhi : ichi;
    launch_couple(g, h->b1[q], cpl, nk, j1, j2, i1, i2, h->st);
    launch_timeint(g, h->b1[q], *bdy_b0(h, q), *bdy_bt(h, q), h->rdtbdy, nk, j1, j2, i1, i2, h->st);
    h->launches += 2;
  }
  h->have_next = false;
  return RCM_OK;
}

// daily update of the upper-radiative-BC mask (Main/mod_sound.F90:500-547)
static int update_tmask(rcm_dyn* h) {
  const Geom& g = h->g;
  Fields& f = h->f;
  const size_t n2 = (size_t)g.sk;
  std::vector<double> astore(n2), rho1(n2), bt1(n2), psb(n2);
  CUDA_OK(cudaStreamSynchronize(h->st));
  CUDA_OK(cudaMemcpy(astore.data(), f.astore - h->off2, n2 * sizeof(double), cudaMemcpyDeviceToHost));
  CUDA_OK(cudaMemcpy(rho1.data(), f.rho1 - h->off3, n2 * sizeof(double), cudaMemcpyDeviceToHost));  // plane k = 1
  CUDA_OK(cudaMemcpy(bt1.data(), f.nbt - h->off3, n2 * sizeof(double), cudaMemcpyDeviceToHost));
  CUDA_OK(cudaMemcpy(psb.data(), f.psb - h->off2, n2 * sizeof(double), cudaMemcpyDeviceToHost));
  double atot = 0.0, rhontot = 0.0;
  for (int i = g.ici1; i <= g.ici2; ++i)
    for (int j = g.jci1; j <= g.jci2; ++j) {
      atot = atot + H2(astore, j, i);
      const double rpsb = 1.0 / H2(psb, j, i);
      const double ensq = egrav * egrav / cpd / (H2(bt1, j, i) * rpsb);
      rhontot = rhontot + H2(rho1, j, i) * std::sqrt(ensq);
    }
  if (h->cfg.nproc > 1) {
    double v[2] = {atot, rhontot};
    if (h->halo.allreduce_host(v, 2, 0)) return fail(h, RCM_ERR_NCCL, "allreduce failed (tmask)");
    atot = v[0]; rhontot = v[1];
  }
  const double abar = atot * h->rnpts, rhon = rhontot * h->rnpts;
  const double dxmsfb = 2.0 / h->s.dxsq / h->xmsf_mean;
  double tmask[169];
  for (double& x : tmask) x = 0.0;
  for (int kk = 0; kk <= 6; ++kk) {
    const double rkk = (double)kk;
    for (int ll = 0; ll <= 6; ++ll) {
      const double rll = (double)ll;
      const double xkeff = dxmsfb * std::sin(mathpi * rkk / 12.0) * std::cos(mathpi * rll / 12.0);
      const double xleff = dxmsfb * std::sin(mathpi * rll / 12.0) * std::cos(mathpi * rkk / 12.0);
      const double xkleff = std::sqrt(xkeff * xkeff + xleff * xleff);
      for (int i = -6; i <= 6; ++i) {
        const double ri = (double)i;
        for (int j = -6; j <= 6; ++j) {
          const double rj = (double)j;
          tmask[(j + 6) + 13 * (i + 6)] = tmask[(j + 6) + 13 * (i + 6)] +
              (h->fi[i + 6] * h->fj[j + 6] * h->fk[kk] * h->fl[ll]) / 144.0 * std::cos(2.0 * mathpi * rkk * ri / 12.0) *
                  std::cos(2.0 * mathpi * rll * rj / 12.0) * xkleff / (rhon - abar * xkleff);
        }
      }
    }
  }
  CUDA_OK(cudaMemcpy(f.tmask, tmask, sizeof(tmask), cudaMemcpyHostToDevice));
  return RCM_OK;
}

static void timer_advance(rcm_dyn* h) {  // Main/mpplib/mod_timer.F90:152-177,306-327
  h->model_time += (int64_t)h->dtsec;
  h->lcount += 1;
  auto check = [&](int64_t& lastact, int64_t& now, int64_t actint) {
    if (now == h->model_time) return;
    const double t1 = (double)h->model_time, t2 = (double)(lastact + actint);
    if (t1 >= t2) { lastact = (h->model_time / actint) * actint; now = h->model_time; }
  };
  check(h->day_lastact, h->day_now, 86400);
  check(h->bdy_lastact, h->bdy_now, (int64_t)h->dtbdys);
}

// one call of `tend` for the hydrostatic core (Main/mod_tendency.F90:212-616), physics off
static int tend_hydro(rcm_dyn* h) {
  const Geom& g = h->g;
  Fields& f = h->f;
  StepScalars& s = h->s;
  s.dt = h->dt;
  s.xt = h->xbctime + h->dt;
  s.integrating = h->lcount > 0 ? 1 : 0;
  h->last_istep = 0;
  int rc = RCM_OK;
  s.cq_mode = h->cq_mode;
  StepScalars s2 = s; s2.cq_mode = 1;   // second pass of k_h_tq: exits at once unless the first one found negative moisture
  CUDA_OK(cudaMemsetAsync(f.negq, 0, sizeof(unsigned long long), h->st));
  CUDA_OK(cudaMemsetAsync(f.nq_cnt, 0, (size_t)2 * g.kz * sizeof(int), h->st));
  if (h->tr.ntr > 0) {
    CUDA_OK(cudaMemsetAsync(h->tr.nq_cnt, 0, (size_t)h->tr.ntr * g.kz * sizeof(int), h->st));
    CUDA_OK(cudaMemsetAsync(h->tr.negcnt, 0, sizeof(unsigned long long), h->st));
  }
  if (!h->mid_step) {   // head of `tend`: everything physical_parametrizations depends on (Main/mod_tendency.F90:230-270)
    TIMED(KID_H_PS, 2, (launch_h_ps(g, f, h->st), launch_h_rimcopy(g, f, h->rim_d, h->nrim_d, h->rim_c, h->nrim_c, h->st)));
    TIMED(KID_HALO, 0, rc = exch_h_psdot(h));
    if (rc) return rc;
    if (!h->prof_on && h->overlap_slow && h->overlap_head) {   // as in tend(): k_xkc (level b) beside k_h_prep (level a)
      CUDA_OK(cudaEventRecord(h->ev_fork, h->st));
      CUDA_OK(cudaStreamWaitEvent(h->st_b, h->ev_fork, 0));
      launch_xkc(g, f, s, h->st_b);
      CUDA_OK(cudaEventRecord(h->ev_jb, h->st_b));
      launch_h_prep(g, f, s, h->st);
      CUDA_OK(cudaStreamWaitEvent(h->st, h->ev_jb, 0));
      h->launches += 2;
    } else {
      TIMED(KID_H_PREP, 1, launch_h_prep(g, f, s, h->st));
      TIMED(KID_XKC, 1, launch_xkc(g, f, s, h->st));
    }
  }
  if (h->head_only) { h->mid_step = true; return RCM_OK; }
  h->mid_step = false;
  if (h->cfg.idiffu == 3)   // diff_6th_coef = diff_6th_factor * 0.015625 / (2 dtsec) (Main/mod_diffusion.F90:78,154)
    TIMED(KID_XKC, 1, launch_diff6(g, f, h->tr, h->uw, s, h->uw_on ? 1 : 0, 0.12 * 0.015625 / (2.0 * h->dtsec), h->st));
  if (h->cfg.isladvec == 1) {   // semi-Lagrangian tendencies of qv, qc, tracers
    if (const int rc = exch_sl(h)) return rc;
    TIMED(KID_TR, 1, launch_sl(g, f, h->tr, s, h->st));
  }
  if (h->prof_on || !h->overlap_slow) {
    TIMED(KID_H_UV, 1, launch_h_uv(g, f, s, h->st));
    TIMED(KID_H_TQ, 1, launch_h_tq(g, f, s, h->st));
    if (s.cq_mode == 0) TIMED(KID_NEGQ, 1, launch_h_tq(g, f, s2, h->st));
  } else {
    // k_h_uv and k_h_tq read the same inputs and write disjoint outputs
    CUDA_OK(cudaEventRecord(h->ev_fork, h->st));
    CUDA_OK(cudaStreamWaitEvent(h->st_b, h->ev_fork, 0));
    launch_h_tq(g, f, s, h->st);
    if (s.cq_mode == 0) { launch_h_tq(g, f, s2, h->st); h->launches += 1; }
    launch_h_uv(g, f, s, h->st_b);
    h->launches += 2;
    CUDA_OK(cudaEventRecord(h->ev_jb, h->st_b));
    CUDA_OK(cudaStreamWaitEvent(h->st, h->ev_jb, 0));
  }
  if (h->tr.ntr > 0) TIMED(KID_TR, 1, launch_tr(g, f, h->tr, s, h->st));
  if (h->uw_on) TIMED(KID_TR, 2, launch_tke(g, f, h->uw, s, h->st));   // TKE tendency, forecast, RA filter (in place)
  // splitf reads the new winds one dot row out (Main/mod_split.F90:276-277,312-313); the repair reads the forecast moisture
  TIMED(KID_HALO, 0, rc = exch_h_new(h));
  if (rc) return rc;
  // time filter of p*, then p* on dot points as at the head of splitf (:259-260); both also on the ghost ring
  // spstep reads (the reference exchanges psdot there, :197-198)
  TIMED(KID_H_PS, 2, (launch_h_psfilt(g, f, s, h->st), launch_psc2psd(g, f.psa, f.psdota, nullptr, 1, h->st)));
  // negative-moisture repair (Main/mod_tendency.F90:381-393), after the p* filter whose result floors the qv filter (:420-425)
  TIMED(KID_NEGQ, 1, launch_negq_repair(g, f, s, h->st));
  if (h->tr.ntr > 0) TIMED(KID_NEGQ, 1, launch_tr_repair(g, h->tr, s, h->st));
  TIMED(KID_SP_PROJ, 1, launch_sp_proj(g, f, s, h->st));
  int m2max = 0;
  for (int l = 0; l < s.nsplit; ++l) m2max = std::max(m2max, s.sp_m2[l]);
  for (int n = 1; n <= m2max; ++n) {
    // the gradient of delh of the source time slot is taken on the 3x3 neighbourhood (:499,527-528,583)
    const int src = (n == 1) ? 1 : ((n & 1) ? 1 : 2);
    TIMED(KID_HALO, 0, rc = exch_h_modal(h, f.delh + (long long)(src - 1) * s.nsplit * g.sk, s.nsplit));
    if (rc) return rc;
    TIMED(KID_SP_STEP, 1, launch_sp_step(g, f, s, n, h->st));
  }
  TIMED(KID_HALO, 0, rc = exch_h_modal(h, f.dhsum, s.nsplit));   // :438
  if (rc) return rc;
  TIMED(KID_SP_CORR, 2, launch_sp_corr(g, f, s, h->st));
  std::swap(f.au, f.nau); std::swap(f.av, f.nav); std::swap(f.bu, f.nbu); std::swap(f.bv, f.nbv);
  std::swap(f.at, f.nat); std::swap(f.aqv, f.naqv); std::swap(f.aqc, f.naqc);
  std::swap(f.bt, f.nbt); std::swap(f.bqv, f.nbqv); std::swap(f.bqc, f.nbqc);
  for (int n = 0; n < h->tr.ntr; ++n) { std::swap(h->tr.a[n], h->tr.na[n]); std::swap(h->tr.b[n], h->tr.nb[n]); }
  timer_advance(h);
  if (h->lcount == 2) h->dt = h->dt2;
  return RCM_OK;
}

// one call of `tend` (Main/mod_tendency.F90:212-616), physics off
static int tend(rcm_dyn* h) {
  if (h->hydro) return tend_hydro(h);
  const Geom& g = h->g;
  Fields& f = h->f;
  StepScalars& s = h->s;
  int istep = std::max((int)(h->dt / h->dtsmax), 2);
  if (h->lcount > 0) istep = std::max(4, istep);
  s.dt = h->dt;
  s.xt = h->xbctime + h->dt;
  s.dts = h->dt / (double)istep;
  s.istep = istep; s.it = 0;
  h->last_istep = istep;
  int rc;
  s.cq_mode = h->cq_mode;
  StepScalars s2 = s; s2.cq_mode = 1;   // second pass of k_tq: exits at once unless the first one found negative moisture
  CUDA_OK(cudaMemsetAsync(f.cflbits, 0, sizeof(unsigned long long), h->st));
  CUDA_OK(cudaMemsetAsync(f.negq, 0, sizeof(unsigned long long), h->st));
  CUDA_OK(cudaMemsetAsync(f.nq_cnt, 0, (size_t)2 * g.kz * sizeof(int), h->st));
  if (h->tr.ntr > 0) {
    CUDA_OK(cudaMemsetAsync(h->tr.nq_cnt, 0, (size_t)h->tr.ntr * g.kz * sizeof(int), h->st));
    CUDA_OK(cudaMemsetAsync(h->tr.negcnt, 0, sizeof(unsigned long long), h->st));
  }
  const bool side = !h->prof_on && h->overlap_slow && h->overlap_head;
  if (!h->mid_step) {   // head of `tend`: everything physical_parametrizations depends on (Main/mod_tendency.F90:230-270)
    if (side) {
      // k_xkc reads level b only, k_prep level a only, disjoint outputs: side by side (k_xkc is memory-bound, k_prep issue-bound)
      CUDA_OK(cudaEventRecord(h->ev_fork, h->st));
      CUDA_OK(cudaStreamWaitEvent(h->st_b, h->ev_fork, 0));
      launch_xkc(g, f, s, h->st_b);
      CUDA_OK(cudaEventRecord(h->ev_jb, h->st_b));
      launch_prep(g, f, s, h->st);
      CUDA_OK(cudaStreamWaitEvent(h->st, h->ev_jb, 0));
      h->launches += 2;
    } else {
      TIMED(KID_PREP, 1, launch_prep(g, f, s, h->st));
      TIMED(KID_XKC, 1, launch_xkc(g, f, s, h->st));
    }
  }
  if (h->head_only) { h->mid_step = true; return RCM_OK; }
  h->mid_step = false;
  if (h->cfg.idiffu == 3)   // diff_6th_coef = diff_6th_factor * 0.015625 / (2 dtsec) (Main/mod_diffusion.F90:78,154)
    TIMED(KID_XKC, 1, launch_diff6(g, f, h->tr, h->uw, s, h->uw_on ? 1 : 0, 0.12 * 0.015625 / (2.0 * h->dtsec), h->st));
  if (h->cfg.isladvec == 1) {   // semi-Lagrangian tendencies of qv, qc, tracers
    if (const int rc = exch_sl(h)) return rc;
    TIMED(KID_TR, 1, launch_sl(g, f, h->tr, s, h->st));
  }
  if (h->prof_on || !h->overlap_slow) {
    TIMED(KID_UV, 1, h->uv_maps ? launch_uv_tma(g, f, s, h->st, h->uv_maps) : launch_uv(g, f, s, h->st));
    TIMED(KID_TQ, 1, h->tq_tma ? launch_tq_tma(g, f, s, h->st) : launch_tq(g, f, s, h->st));
    if (s.cq_mode == 0) TIMED(KID_NEGQ, 1, h->tq_tma ? launch_tq_tma(g, f, s2, h->st) : launch_tq(g, f, s2, h->st));
    TIMED(KID_PPW, 1, h->ppw_tma ? launch_ppw_tma(g, f, s, h->st) : launch_ppw(g, f, s, h->st));
  } else {
    // k_uv, k_tq, k_ppw read the same inputs and write disjoint outputs: run
    // them side by side so that one kernel's tail wave is filled by the others
    CUDA_OK(cudaEventRecord(h->ev_fork, h->st));
    CUDA_OK(cudaStreamWaitEvent(h->st_b, h->ev_fork, 0));
    CUDA_OK(cudaStreamWaitEvent(h->st_c, h->ev_fork, 0));
    if (h->tq_tma) launch_tq_tma(g, f, s, h->st); else launch_tq(g, f, s, h->st);
    if (s.cq_mode == 0) { if (h->tq_tma) launch_tq_tma(g, f, s2, h->st); else launch_tq(g, f, s2, h->st); h->launches += 1; }
    if (h->ppw_tma) launch_ppw_tma(g, f, s, h->st_b); else launch_ppw(g, f, s, h->st_b);
    if (h->uv_maps) launch_uv_tma(g, f, s, h->st_c, h->uv_maps); else launch_uv(g, f, s, h->st_c);
    h->launches += 3;
    // k_snd_init reads level b and the reference state only and writes the acoustic work arrays: behind k_uv (the
    // shortest of the three) it fills memory bandwidth the issue-bound slow kernels leave idle
    if (side) { launch_snd_init(g, f, s, h->st_c); h->launches += 1; }
    CUDA_OK(cudaEventRecord(h->ev_jb, h->st_b));
    CUDA_OK(cudaEventRecord(h->ev_jc, h->st_c));
    CUDA_OK(cudaStreamWaitEvent(h->st, h->ev_jb, 0));
    CUDA_OK(cudaStreamWaitEvent(h->st, h->ev_jc, 0));
  }
  if (h->tr.ntr > 0) TIMED(KID_TR, 1, launch_tr(g, f, h->tr, s, h->st));
  if (h->uw_on) TIMED(KID_TR, 2, launch_tke(g, f, h->uw, s, h->st));   // TKE tendency, forecast, RA filter (in place)
  if (!side) TIMED(KID_SND_INIT, 1, launch_snd_init(g, f, s, h->st));
  TIMED(KID_HALO, 0, rc = exch_uvten(h));  // + the first cppd, ct of the step and the moisture forecast
  if (rc) return rc;
  // negative-moisture repair (Main/mod_tendency.F90:381-393): before k_snd_wback first reads the filtered qv
  TIMED(KID_NEGQ, 1, launch_negq_repair(g, f, s, h->st));
  if (h->tr.ntr > 0) TIMED(KID_NEGQ, 1, launch_tr_repair(g, h->tr, s, h->st));
  for (int it = 1; it <= istep; ++it) {
    s.it = it;
    TIMED(KID_SND_UV, 1, launch_snd_uv(g, f, s, h->st));
    TIMED(KID_SND_WFWD, 1, (h->wfwd_tma && s.it > 1) ? launch_snd_wfwd_tma(g, f, s, h->st) : launch_snd_wfwd(g, f, s, h->st));
    if (h->cfg.ifupr == 1) {
      TIMED(KID_HALO, 0, rc = exch_estore(h));
      if (rc) return rc;
      if (h->day_now == h->model_time && it == 1)
        if ((rc = update_tmask(h))) return rc;
    }
    TIMED(KID_SND_WBACK, 1, launch_snd_wback(g, f, s, h->st));
    if (it < istep) {
      TIMED(KID_HALO, 0, rc = exch_sound_pp(h));
      if (rc) return rc;
    }
  }
  TIMED(KID_SND_FINAL, 1, launch_snd_final(g, f, s, h->st));
  // swap in the filtered t, qv, qc (the old buffers keep the previous rim)
  std::swap(f.at, f.nat); std::swap(f.aqv, f.naqv); std::swap(f.aqc, f.naqc);
  std::swap(f.bt, f.nbt); std::swap(f.bqv, f.nbqv); std::swap(f.bqc, f.nbqc);
  for (int n = 0; n < h->tr.ntr; ++n) { std::swap(h->tr.a[n], h->tr.na[n]); std::swap(h->tr.b[n], h->tr.nb[n]); }
  timer_advance(h);
  if (h->lcount == 2) h->dt = h->dt2;
  return RCM_OK;
}

// One collective for the per-rank verdicts of a call: vmax[] is max-reduced, vsum[] sum-reduced, so that every
// rank takes the same return path (the reference aborts all ranks together, Share/mod_message.F90:90-103).
// Small worlds use ONE all-gather written as a sum of one-hot slots; NaN survives both reductions.
static int reduce_verdicts(rcm_dyn* h, double* vmax, int nmax, double* vsum, int nsum) {
  const int np = h->cfg.nproc, n = nmax + nsum;
  if (np <= 1) return 0;
  if (np * n <= 64) {
    double buf[64];
    for (double& x : buf) x = 0.0;
    for (int q = 0; q < nmax; ++q) buf[h->cfg.myid * n + q] = vmax[q];
    for (int q = 0; q < nsum; ++q) buf[h->cfg.myid * n + nmax + q] = vsum[q];
    if (h->halo.allreduce_host(buf, np * n, 0)) return -5;
    for (int q = 0; q < nmax; ++q) {
      double m = buf[q];
      for (int r = 1; r < np; ++r) { const double v = buf[r * n + q]; if (v != v || (m == m && v > m)) m = v; }
      vmax[q] = m;
    }
    for (int q = 0; q < nsum; ++q) {
      double t = 0.0;
      for (int r = 0; r < np; ++r) t += buf[r * n + nmax + q];
      vsum[q] = t;
    }
    return 0;
  }
  if (nmax > 0 && h->halo.allreduce_host(vmax, nmax, 1)) return -5;
  if (nsum > 0 && h->halo.allreduce_host(vsum, nsum, 0)) return -5;
  return 0;
}

// copy the verdict scalars of the step just enqueued into the next slot of the pinned ring (stream-ordered)
static int record_verdict(rcm_dyn* h) {
  if (!h->vring) {
    CUDA_OK(cudaMallocHost((void**)&h->vring, sizeof(rcm_dyn::VSlot) * rcm_dyn::kVRing));
    std::memset(h->vring, 0, sizeof(rcm_dyn::VSlot) * rcm_dyn::kVRing);
    for (int q = 0; q < rcm_dyn::kVRing; ++q) CUDA_OK(cudaEventCreate(&h->vev[q]));
  }
  const int q = (int)(h->vhead % rcm_dyn::kVRing);
  if (h->vhead >= rcm_dyn::kVRing) CUDA_OK(cudaEventSynchronize(h->vev[q]));   // 64 steps old: complete long ago
  rcm_dyn::VSlot* v = h->vring + q;
  if (h->moloch) {
    CUDA_OK(cudaMemcpyAsync(&v->nanf, h->f.nanflag, sizeof(int), cudaMemcpyDeviceToHost, h->st));
    CUDA_OK(cudaMemcpyAsync(v->ps, h->mf.psbits, 2 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, h->st));
  } else {
    CUDA_OK(cudaMemcpyAsync(v, h->f.cflbits, 4 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, h->st));
  }
  CUDA_OK(cudaEventRecord(h->vev[q], h->st));
  h->vlcount[q] = h->lcount;
  h->vhead += 1;
  return RCM_OK;
}

// enqueue nsteps x { tend | moloch ; [bdyin] ; bdyval } on the library's stream; no host synchronisation except the
// daily tmask update (Main/mod_sound.F90:500-547) and what a record change needs
static int enqueue_steps(rcm_dyn* h, int nsteps) {
  Fields& f = h->f;
  if (!h->pending) {   // head of a batch: call-level maxima and flags start at zero
    if (h->moloch) {
      CUDA_OK(cudaMemsetAsync(f.nanflag, 0, sizeof(int), h->st));
    } else {
      CUDA_OK(cudaMemsetAsync(f.cflbits, 0, 2 * sizeof(unsigned long long), h->st));
      CUDA_OK(cudaMemsetAsync(f.negq, 0, sizeof(unsigned long long), h->st));
      CUDA_OK(cudaMemsetAsync(f.nanflag, 0, sizeof(int), h->st));
    }
    CUDA_OK(cudaEventRecord(h->ev0, h->st));
    h->pending = true;
    h->pending_steps = 0;
  }
  int rc = RCM_OK;
  for (int n = 0; n < nsteps; ++n) {
    if (h->moloch) {   // Main/mod_regcm_interface.F90:186-209: moloch; bdyin on the boundary alarm; bdyval
      if ((rc = mo_step_one(h))) break;
      if (h->bdy_now == h->model_time) {
        if (!h->have_next) { rc = fail(h, RCM_ERR_BDY, "boundary record needed but not staged"); break; }
        if ((rc = mo_bdyin(h))) break;
      }
      if ((rc = mo_bdyval(h))) break;
    } else {
      if ((rc = tend(h))) break;
      if (h->bdy_now == h->model_time) {
        if (!h->have_next) { rc = fail(h, RCM_ERR_BDY, "boundary record needed but not staged"); break; }
        if ((rc = bdyin(h))) break;
        if ((rc = bdyin_tracers(h))) break;
      }
      // after the swap in tend(): n?? hold the previous level-a buffers
      TIMED(KID_HALO, 0, rc = h->hydro ? exch_h_state(h) : exch_state(h));
      if (rc) break;
      if ((rc = bdyval(h, f.nat, f.naqv, f.naqc))) break;
    }
    h->pending_steps += 1;
    if ((rc = record_verdict(h))) break;
  }
  return rc;
}

// wait for the enqueued steps and reach the verdict of the batch (the same on every rank)
static int finish_steps(rcm_dyn* h) {
  if (!h->pending) return RCM_OK;
  Fields& f = h->f;
  const int nsteps = h->pending_steps;
  h->pending = false;
  CUDA_OK(cudaEventRecord(h->ev1, h->st));
  CUDA_OK(cudaStreamSynchronize(h->st));
  float ms = 0.f;
  cudaEventElapsedTime(&ms, h->ev0, h->ev1);
  h->last_ms = ms;
  prof_collect(h);
  h->vdone = h->vhead - 1;
  if (h->moloch) {
    int nanf = 0;
    unsigned long long pb[2] = {0, 0};
    CUDA_OK(cudaMemcpy(&nanf, f.nanflag, sizeof(nanf), cudaMemcpyDeviceToHost));
    CUDA_OK(cudaMemcpy(pb, h->mf.psbits, sizeof(pb), cudaMemcpyDeviceToHost));
    h->nan_flag = nanf;
    if (nsteps > 0 && pb[0]) { const unsigned long long mn = ~pb[1]; std::memcpy(&h->ps_max, &pb[0], 8); std::memcpy(&h->ps_min, &mn, 8); }
    int halo_err = h->halo.take_error();
    if (h->cfg.nproc > 1 && nsteps > 0) {   // maxall / minall of the reference (Main/mod_moloch.F90:440-441)
      double v[4] = {h->ps_max, -h->ps_min, (double)nanf, (double)halo_err};
      if (reduce_verdicts(h, v, 4, nullptr, 0)) return fail(h, RCM_ERR_NCCL, "allreduce (ps extremes) failed");
      h->ps_max = v[0]; h->ps_min = -v[1]; nanf = (v[2] != 0.0); h->nan_flag = nanf; halo_err = (v[3] != 0.0);
    }
    CUDA_OK(cudaGetLastError());
    if (halo_err) return fail(h, RCM_ERR_NCCL, "halo exchange: timed out waiting for a neighbour's edge data");
    if (nanf) return fail(h, RCM_ERR_CFL, "CFL violation: surface pressure is not a number (decrease dt)");   // Main/mod_moloch.F90:443-451
    return RCM_OK;
  }
  int halo_err = h->halo.take_error();
  unsigned long long blk[4] = {0, 0, 0, 0};
  CUDA_OK(cudaMemcpy(blk, f.cflbits, sizeof(blk), cudaMemcpyDeviceToHost));
  unsigned long long nq = blk[2];
  int nanf = 0; std::memcpy(&nanf, &blk[3], sizeof(nanf));
  double cfl, cfl_last; std::memcpy(&cfl, &blk[1], sizeof(cfl)); std::memcpy(&cfl_last, &blk[0], sizeof(cfl_last));
  if (h->cfg.nproc > 1 && nsteps > 0) {   // every rank must reach the same verdict (the reference's fatal() aborts all of them)
    double vm[4] = {cfl, cfl_last, (double)nanf, (double)halo_err}, vs[1] = {(double)nq};
    if (reduce_verdicts(h, vm, 4, vs, 1)) return fail(h, RCM_ERR_NCCL, "allreduce (step verdicts) failed");
    cfl = vm[0]; cfl_last = vm[1]; nanf = (vm[2] != 0.0); halo_err = (vm[3] != 0.0); nq = (unsigned long long)vs[0];
  }
  if (h->tr.ntr > 0) {   // negative tracer forecasts of the last step (this rank's)
    unsigned long long nc = 0;
    CUDA_OK(cudaMemcpy(&nc, h->tr.negcnt, sizeof(nc), cudaMemcpyDeviceToHost));
    h->neg_chi = (int64_t)nc;
  }
  h->max_cfl = cfl_last; h->neg_q = (int64_t)nq; h->nan_flag = nanf;
  h->cq_mode = nq > 0 ? 2 : 0;   // moisture went negative: the next batch stores the forecast in its first pass
  CUDA_OK(cudaGetLastError());
  if (halo_err) return fail(h, RCM_ERR_NCCL, "halo exchange: timed out waiting for a neighbour's edge data");
  if (nanf == 2) return fail(h, RCM_ERR_CFL, "SL advection: a departure point lies more than one cell away (decrease dt)");   // Main/mod_sladvection.F90:152-157
  if (nanf || cfl != cfl) return fail(h, RCM_ERR_NAN, "NaN in the acoustic solver");
  if (cfl > 1.0) {
    char buf[128]; std::snprintf(buf, sizeof(buf), "CFL violation: max sigma-dot CFL = %g", cfl);
    return fail(h, RCM_ERR_CFL, buf);
  }
  return RCM_OK;
}

extern "C" int rcm_dyn_step(rcm_dyn* h, int nsteps) {
  if (!h) return RCM_ERR_ARG;
  if (!h->inited) return fail(h, RCM_ERR_STATE, "rcm_dyn_step before rcm_dyn_init");
  cudaSetDevice(h->device);
  const int rc = enqueue_steps(h, nsteps);
  const std::string msg = h->err;
  const int rc2 = finish_steps(h);
  if (rc) { h->err = msg; return rc; }
  return rc2;
}

// ---------------------------------------------------------------------------
// Physics coupling boundary.  The reference calls physical_parametrizations in the middle of
// `tend` (Main/mod_tendency.F90:271): they read the `atms` slices built by mkslice and return aten(:,:,:,pc_physic),
// which the tendency sums add (:285-314,331-335,404-411).  Here the step splits at the same place.
static double** phys_ptr(rcm_dyn* h, int field) {
  Fields& f = h->f;
  switch (field) {
    case RCM_F_U: return &f.phy_u; case RCM_F_V: return &f.phy_v; case RCM_F_T: return &f.phy_t;
    case RCM_F_QV: return &f.phy_qv; case RCM_F_QC: return &f.phy_qc; case RCM_F_PP: return &f.phy_pp; case RCM_F_W: return &f.phy_w;
  }
  return nullptr;
}

extern "C" int rcm_dyn_set_phys_tend(rcm_dyn* h, int field, const double* a, const int32_t lb[3], const int32_t ub[3]) {
  if (!h || !a) return RCM_ERR_ARG;
  if (h->moloch) {   // mo_atm%tten, qxten, uten, vten (Main/mod_moloch.F90:1548-1557): uncoupled, u and v on their C-grid points
    MoFields& m = h->mf;
    double** pm = field == RCM_F_T ? &m.phy_t : field == RCM_F_QV ? &m.phy_qv : field == RCM_F_QC ? &m.phy_qc : field == RCM_F_U ? &m.phy_u
                  : field == RCM_F_V ? &m.phy_v : nullptr;
    if (!pm) return fail(h, RCM_ERR_ARG, "no physics tendency for this field");
    cudaSetDevice(h->device);
    if (h->pending) { const int rcp_ = finish_steps(h); if (rcp_) return rcp_; }
    if (!m.phy_t)
      for (double** q : {&m.phy_t, &m.phy_qv, &m.phy_qc, &m.phy_u, &m.phy_v}) {
        *q = dev_alloc3(h, h->g.kzp1);
        if (!*q) return fail(h, RCM_ERR_CUDA, "device memory allocation failed (physics tendencies)");
      }
    return upload3(h, *pm, a, lb, ub, h->g.kz);
  }
  double** pp = (field == RCM_F_TKE && h->uw_on) ? &h->uw.phy : phys_ptr(h, field);
  if (!pp || (h->hydro && (field == RCM_F_PP || field == RCM_F_W))) return fail(h, RCM_ERR_ARG, "no physics tendency for this field");
  cudaSetDevice(h->device);
  if (h->pending) { const int rcp_ = finish_steps(h); if (rcp_) return rcp_; }
  if (field == RCM_F_TKE) return upload3(h, *pp, a, lb, ub, h->g.kzp1);   // aten%tke(:,:,:,pc_physic), allocated with the handle
  if (!h->f.phy_t) {   // first use: all of them, zero-filled (u and v are added as a pair)
    for (int q : {RCM_F_U, RCM_F_V, RCM_F_T, RCM_F_QV, RCM_F_QC, RCM_F_PP, RCM_F_W}) {
      if (h->hydro && (q == RCM_F_PP || q == RCM_F_W)) continue;
      double* p = dev_alloc3(h, h->g.kzp1);
      if (!p) return fail(h, RCM_ERR_CUDA, "device memory allocation failed (physics tendencies)");
      *phys_ptr(h, q) = p;
    }
  }
  const int nk = (field == RCM_F_W) ? h->g.kzp1 : h->g.kz;
  return upload3(h, *pp, a, lb, ub, nk);
}

extern "C" int rcm_dyn_clear_phys_tend(rcm_dyn* h) {
  if (!h) return RCM_ERR_ARG;
  cudaSetDevice(h->device);
  if (h->pending) { const int rcp_ = finish_steps(h); if (rcp_) return rcp_; }
  if (h->moloch) {
    for (double* p : {h->mf.phy_t, h->mf.phy_qv, h->mf.phy_qc, h->mf.phy_u, h->mf.phy_v})
      if (p) CUDA_OK(cudaMemset(p - h->off3, 0, (size_t)h->g.sk * h->g.kzp1 * sizeof(double)));
    return RCM_OK;
  }
  for (int q : {RCM_F_U, RCM_F_V, RCM_F_T, RCM_F_QV, RCM_F_QC, RCM_F_PP, RCM_F_W}) {
    double** pp = phys_ptr(h, q);
    if (pp && *pp) CUDA_OK(cudaMemset(*pp - h->off3, 0, (size_t)h->g.sk * h->g.kzp1 * sizeof(double)));
  }
  if (h->uw_on) CUDA_OK(cudaMemset(h->uw.phy - h->off3, 0, (size_t)h->g.sk * h->g.kzp1 * sizeof(double)));
  return RCM_OK;
}

// kpbl of the PBL scheme for vadv4d ind = 3 (Main/mod_advection.F90:66,92,921-956)
extern "C" int rcm_dyn_set_kpbl(rcm_dyn* h, const int32_t* kpbl, const int32_t lb[2], const int32_t ub[2]) {
  if (!h || !kpbl || !lb || !ub) return RCM_ERR_ARG;
  if (!h->uw_on) return fail(h, RCM_ERR_ARG, "kpbl exists with ibltyp = 2 only (idynamic 1, 2)");
  cudaSetDevice(h->device);
  if (h->pending) { const int rcp_ = finish_steps(h); if (rcp_) return rcp_; }
  const Geom& g = h->g;
  std::vector<int>& host = h->kpb_host;   // host mirror: points outside the caller's window keep their values
  if (host.size() != (size_t)g.sk) host.assign((size_t)g.sk, 0);
  const long long n1 = ub[0] - lb[0] + 1;
  for (int i = std::max(lb[1], g.ici1); i <= std::min(ub[1], g.ici2); ++i)
    for (int j = std::max(lb[0], g.jci1); j <= std::min(ub[0], g.jci2); ++j) {
      const int v = kpbl[(j - lb[0]) + n1 * (i - lb[1])];
      if (v > g.kz) return fail(h, RCM_ERR_ARG, "kpbl is greater than kz");   // the reference calls fatal (:923-925)
      host[h->off2 + j + (long long)i * g.si] = v;
    }
  CUDA_OK(cudaMemcpy(h->kpb - h->off2, host.data(), host.size() * sizeof(int), cudaMemcpyHostToDevice));
  return RCM_OK;
}

extern "C" int rcm_dyn_step_pre_physics(rcm_dyn* h) {
  if (!h) return RCM_ERR_ARG;
  if (!h->inited) return fail(h, RCM_ERR_STATE, "rcm_dyn_step_pre_physics before rcm_dyn_init");
  cudaSetDevice(h->device);
  if (h->pending) { const int rcp_ = finish_steps(h); if (rcp_) return rcp_; }
  if (h->mid_step) return RCM_OK;   // already there
  h->head_only = true;
  const int rc = h->moloch ? mo_step_one(h) : tend(h);
  h->head_only = false;
  if (rc) return rc;
  CUDA_OK(cudaStreamSynchronize(h->st));
  return RCM_OK;
}

extern "C" int rcm_dyn_step_post_physics(rcm_dyn* h) {
  if (!h) return RCM_ERR_ARG;
  if (!h->inited) return fail(h, RCM_ERR_STATE, "rcm_dyn_step_post_physics before rcm_dyn_init");
  if (!h->mid_step) return fail(h, RCM_ERR_STATE, "rcm_dyn_step_post_physics without rcm_dyn_step_pre_physics");
  return rcm_dyn_step(h, 1);   // tend() skips the head it has already run
}

extern "C" int rcm_dyn_get_slice(rcm_dyn* h, int id, double* a, const int32_t lb[3], const int32_t ub[3]) {
  if (!h || !a) return RCM_ERR_ARG;
  if (!h->inited) return fail(h, RCM_ERR_STATE, "rcm_dyn_get_slice before rcm_dyn_init");
  if (id < 0 || id >= RCM_SL_COUNT) return fail(h, RCM_ERR_ARG, "bad slice id");
  cudaSetDevice(h->device);
  if (h->pending) { const int rcp_ = finish_steps(h); if (rcp_) return rcp_; }
  if (h->moloch) {   // aliases of mo_atm and the products of mkslice for idynamic = 3 (Main/mod_slice.F90:82-146)
    if (!h->slice_buf) { h->slice_buf = dev_alloc3(h, h->g.kzp1); if (!h->slice_buf) return fail(h, RCM_ERR_CUDA, "device memory allocation failed (slice)"); }
    launch_mo_slice(h->g, h->mf, id, h->slice_buf, h->st);
    h->launches += 1;
    int nkm = h->g.kz;
    if (id == RCM_SL_PF3D || id == RCM_SL_WB3D || id == RCM_SL_ZQ) nkm = h->g.kzp1;
    int32_t l3m[3] = {lb[0], lb[1], lb[2]}, u3m[3] = {ub[0], ub[1], ub[2]};
    if (id == RCM_SL_PS2D) { nkm = 1; l3m[2] = u3m[2] = 1; }
    return download3(h, h->slice_buf, a, l3m, u3m, nkm);
  }
  if (h->hydro && !h->mid_step && (id == RCM_SL_WPX3D || id == RCM_SL_WB3D))
    return fail(h, RCM_ERR_STATE, "omega of the hydrostatic core exists between rcm_dyn_step_pre_physics and rcm_dyn_step_post_physics");
  if (!h->slice_buf) { h->slice_buf = dev_alloc3(h, h->g.kzp1); if (!h->slice_buf) return fail(h, RCM_ERR_CUDA, "device memory allocation failed (slice)"); }
  const rcm_cfg& c = h->cfg;
  launch_slice(h->g, h->f, id, h->hydro ? 1 : 0, c.ptop, c.base_state_pressure, c.logp_lrate, c.base_state_ts0, h->s.dx8, h->slice_buf, h->st);
  h->launches += 1;
  int nk = h->g.kz;
  if (id == RCM_SL_PF3D || id == RCM_SL_WB3D || id == RCM_SL_ZQ) nk = h->g.kzp1;
  int32_t l3[3] = {lb[0], lb[1], lb[2]}, u3[3] = {ub[0], ub[1], ub[2]};
  if (id == RCM_SL_PS2D) { nk = 1; l3[2] = u3[2] = 1; }
  return download3(h, h->slice_buf, a, l3, u3, nk);
}

extern "C" int rcm_dyn_step_async(rcm_dyn* h, int nsteps) {
  if (!h) return RCM_ERR_ARG;
  if (!h->inited) return fail(h, RCM_ERR_STATE, "rcm_dyn_step_async before rcm_dyn_init");
  cudaSetDevice(h->device);
  return enqueue_steps(h, nsteps);
}

extern "C" int rcm_dyn_sync(rcm_dyn* h) {
  if (!h) return RCM_ERR_ARG;
  cudaSetDevice(h->device);
  return finish_steps(h);
}

// ---------------------------------------------------------------------------
static double* state_ptr(rcm_dyn* h, int field, int level) {
  Fields& f = h->f;
  const bool A = (level == RCM_LEVEL_A);
  switch (field) {
    case RCM_F_U: return A ? f.au : f.bu;
    case RCM_F_V: return A ? f.av : f.bv;
    case RCM_F_T: return A ? f.at : f.bt;
    case RCM_F_QV: return A ? f.aqv : f.bqv;
    case RCM_F_QC: return A ? f.aqc : f.bqc;
    case RCM_F_PP: return A ? f.app : f.bpp;
    case RCM_F_W: return A ? f.aw : f.bw;
  }
  if (field >= RCM_F_CHI && field - RCM_F_CHI < h->tr.ntr) return A ? h->tr.a[field - RCM_F_CHI] : h->tr.b[field - RCM_F_CHI];
  if (field == RCM_F_TKE && h->uw_on) return A ? h->uw.a : h->uw.b;
  return nullptr;
}

static double* ps_state(rcm_dyn* h, int level) { return (level == RCM_LEVEL_A) ? h->f.psa : h->f.psb; }

static int download3(rcm_dyn* h, const double* dev_shifted, double* a, const int32_t* lb, const int32_t* ub, int nk) {
#ifndef RCM_EMU
  return copy3d(h, const_cast<double*>(dev_shifted), a, lb, ub, nk, false);
#else
  const Geom& g = h->g;
  CUDA_OK(cudaMemcpy(h->stage.data(), dev_shifted - h->off3, (size_t)g.sk * nk * sizeof(double), cudaMemcpyDeviceToHost));
  const int jlo = std::max(lb[0], g.jde1 - RCM_HALO), jhi = std::min(ub[0], g.jde2 + RCM_HALO);
  const int ilo = std::max(lb[1], g.ide1 - RCM_HALO), ihi = std::min(ub[1], g.ide2 + RCM_HALO);
  const long long n1 = ub[0] - lb[0] + 1, n2 = ub[1] - lb[1] + 1;
  for (int k = std::max(lb[2], 1); k <= std::min(ub[2], nk); ++k)
    for (int i = ilo; i <= ihi; ++i)
      for (int j = jlo; j <= jhi; ++j)
        a[(j - lb[0]) + n1 * ((i - lb[1]) + n2 * (long long)(k - lb[2]))] =
            h->stage[h->off2 + j + (long long)i * g.si + (long long)(k - 1) * g.sk];
  return RCM_OK;
#endif
}

extern "C" int rcm_dyn_get_state(rcm_dyn* h, int field, int level, double* a, const int32_t lb[3], const int32_t ub[3]) {
  if (!h || !a) return RCM_ERR_ARG;
  cudaSetDevice(h->device);
  if (h->pending) { const int rcp_ = finish_steps(h); if (rcp_) return rcp_; }
  if (field == RCM_F_PS) {
    if (!h->hydro && !h->moloch) return fail(h, RCM_ERR_ARG, "p* is a state field for idynamic = 1 and 3 only");
    int32_t l3[3] = {lb[0], lb[1], 1}, u3[3] = {ub[0], ub[1], 1};
    return download3(h, (h->moloch ? h->mf.ps : ps_state(h, level)) - h->g.sk, a, l3, u3, 1);
  }
  double* p = h->moloch ? mo_state_ptr(h, field) : state_ptr(h, field, level);
  if (!p) return fail(h, RCM_ERR_ARG, "bad field");
  return download3(h, p, a, lb, ub, h->g.kzp1);
}

extern "C" int rcm_dyn_set_state(rcm_dyn* h, int field, int level, const double* a, const int32_t lb[3], const int32_t ub[3]) {
  if (!h || !a) return RCM_ERR_ARG;
  cudaSetDevice(h->device);
  if (h->pending) { const int rcp_ = finish_steps(h); if (rcp_) return rcp_; }
  if (field == RCM_F_PS) {
    if (!h->hydro) return fail(h, RCM_ERR_ARG, "p* is a state field for idynamic = 1 only");
    int32_t l3[3] = {lb[0], lb[1], 1}, u3[3] = {ub[0], ub[1], 1};
    return upload3(h, ps_state(h, level) - h->g.sk, a, l3, u3, 1);
  }
  double* p = h->moloch ? mo_state_ptr(h, field) : state_ptr(h, field, level);
  if (!p) return fail(h, RCM_ERR_ARG, "bad field");
  return upload3(h, p, a, lb, ub, h->g.kzp1);
}

// Asynchronous export of one prognostic field: a device-side snapshot is taken in stream order (after the steps
// already enqueued), then copied to the caller's (page-locked) array on a private stream, so that the D2H transfer
// overlaps the steps enqueued afterwards.  rcm_dyn_get_state_wait returns when every queued export has landed.
extern "C" int rcm_dyn_get_state_async(rcm_dyn* h, int field, int level, double* a, const int32_t lb[3], const int32_t ub[3]) {
  if (!h || !a) return RCM_ERR_ARG;
  cudaSetDevice(h->device);
#ifdef RCM_EMU
  return rcm_dyn_get_state(h, field, level, a, lb, ub);
#else
  if (!h->inited) return fail(h, RCM_ERR_STATE, "rcm_dyn_get_state_async before rcm_dyn_init");
  double* p = nullptr;
  int nk = h->g.kzp1;
  int32_t l3[3] = {lb[0], lb[1], lb[2]}, u3[3] = {ub[0], ub[1], ub[2]};
  if (field == RCM_F_PS) {
    if (!h->hydro && !h->moloch) return fail(h, RCM_ERR_ARG, "p* is a state field for idynamic = 1 and 3 only");
    p = (h->moloch ? h->mf.ps : ps_state(h, level)) - h->g.sk; nk = 1; l3[2] = u3[2] = 1;
  } else {
    p = h->moloch ? mo_state_ptr(h, field) : state_ptr(h, field, level);
  }
  if (!p || field < 0 || 2 * field + 1 >= rcm_dyn::kSnap) return fail(h, RCM_ERR_ARG, "bad field");
  const int q = 2 * field + (level == RCM_LEVEL_B ? 1 : 0);
  const size_t bytes = (size_t)h->g.sk * (size_t)nk * sizeof(double);
  if (!h->st_out) CUDA_OK(cudaStreamCreateWithFlags(&h->st_out, cudaStreamNonBlocking));
  if (!h->snap[q]) {   // first export of this field: the only allocation, outside the stepping path
    CUDA_OK(cudaMalloc((void**)&h->snap[q], (size_t)h->g.sk * (size_t)h->g.kzp1 * sizeof(double)));
    CUDA_OK(cudaEventCreateWithFlags(&h->ev_snap[q], cudaEventDisableTiming));
    CUDA_OK(cudaEventCreateWithFlags(&h->ev_out[q], cudaEventDisableTiming));
  } else {
    CUDA_OK(cudaStreamWaitEvent(h->st, h->ev_out[q], 0));   // the previous export of this field must have left the snapshot
  }
  const Geom& g = h->g;
  const int jlo = std::max(l3[0], g.jde1 - RCM_HALO), jhi = std::min(u3[0], g.jde2 + RCM_HALO);
  const int ilo = std::max(l3[1], g.ide1 - RCM_HALO), ihi = std::min(u3[1], g.ide2 + RCM_HALO);
  const int klo = std::max(l3[2], 1), khi = std::min(u3[2], nk);
  if (jhi < jlo || ihi < ilo || khi < klo) return RCM_OK;
  if (jlo == l3[0] && jhi == u3[0] && ilo == l3[1] && ihi == u3[1]) {
    // the window covers whole (j,i) planes of the caller's array: pack it on the device (this IS the snapshot), then
    // ONE contiguous D2H copy
    const int n1 = jhi - jlo + 1, n2 = ihi - ilo + 1, nkk = khi - klo + 1;
    launch_window(true, h->snap[q], p, jlo, ilo, klo, n1, n2, nkk, g.si, g.sk, h->st);
    h->launches += 1;
    CUDA_OK(cudaEventRecord(h->ev_snap[q], h->st));
    CUDA_OK(cudaStreamWaitEvent(h->st_out, h->ev_snap[q], 0));
    CUDA_OK(cudaMemcpyAsync(a + (size_t)n1 * n2 * (size_t)(klo - l3[2]), h->snap[q], (size_t)n1 * n2 * nkk * sizeof(double),
                            cudaMemcpyDeviceToHost, h->st_out));
  } else {
    CUDA_OK(cudaMemcpyAsync(h->snap[q], p - h->off3, bytes, cudaMemcpyDeviceToDevice, h->st));
    CUDA_OK(cudaEventRecord(h->ev_snap[q], h->st));
    CUDA_OK(cudaStreamWaitEvent(h->st_out, h->ev_snap[q], 0));
    const int rc = copy3d(h, h->snap[q] + h->off3, a, l3, u3, nk, false, h->st_out, false);
    if (rc) return rc;
  }
  CUDA_OK(cudaEventRecord(h->ev_out[q], h->st_out));
  h->out_pending = true;
  return RCM_OK;
#endif
}

extern "C" int rcm_dyn_get_state_wait(rcm_dyn* h) {
  if (!h) return RCM_ERR_ARG;
#ifndef RCM_EMU
  if (h->out_pending) { cudaSetDevice(h->device); CUDA_OK(cudaStreamSynchronize(h->st_out)); h->out_pending = false; }
#endif
  return RCM_OK;
}

extern "C" int rcm_dyn_get_work(rcm_dyn* h, int id, double* a, const int32_t lb[3], const int32_t ub[3]) {
  if (!h || !a) return RCM_ERR_ARG;
  cudaSetDevice(h->device);
  if (h->pending) { const int rcp_ = finish_steps(h); if (rcp_) return rcp_; }
  Fields& f = h->f;
  const double* p = nullptr;
  if (h->moloch) {
    p = mo_work_ptr(h, id);
    if (!p) return fail(h, RCM_ERR_ARG, "work array not available for this core");
    return download3(h, p, a, lb, ub, h->g.kzp1);
  }
  switch (id) {
    case RCM_W_CR: p = f.cr; break;
    case RCM_W_QDOT: p = f.qdot; break;
    case RCM_W_XKC: p = f.xkc; break;
    case RCM_W_TH: p = f.th; break;
    case RCM_W_RHO1: p = f.rho1; break;
    case RCM_W_PR1: p = f.pr1; break;
    case RCM_W_UTEN: p = f.uten; break;
    case RCM_W_VTEN: p = f.vten; break;
    case RCM_W_PPTEN: p = f.ppten; break;
    case RCM_W_WTEN: p = f.wten; break;
    case RCM_W_CQV: p = f.cqv; break;
    case RCM_W_CQC: p = f.cqc; break;
    case RCM_W_OMEGA: p = h->hydro ? f.omega : nullptr; break;
    case RCM_W_PHI: p = h->hydro ? f.phi : nullptr; break;
    case RCM_W_PTEN: {
      if (!h->hydro) return fail(h, RCM_ERR_ARG, "pten exists for idynamic = 1 only");
      int32_t l3[3] = {lb[0], lb[1], 1}, u3[3] = {ub[0], ub[1], 1};
      return download3(h, f.pten - h->g.sk, a, l3, u3, 1);
    }
    default: return fail(h, RCM_ERR_ARG, "bad work id");
  }
  if (!p) return fail(h, RCM_ERR_ARG, "work array not available for this core");
  return download3(h, p, a, lb, ub, h->g.kzp1);
}

extern "C" int rcm_dyn_diag(rcm_dyn* h, rcm_diag* d) {
  if (!h || !d) return RCM_ERR_ARG;
  std::memset(d, 0, sizeof(*d));
  d->lcount = h->lcount; d->xbctime = h->xbctime; d->dt = h->dt;
  d->max_cfl = h->max_cfl; d->istep = h->last_istep; d->nan_flag = h->nan_flag;
  d->neg_q_points = h->neg_q;
  d->neg_chi_points = h->neg_chi;
  if (h->pending && h->hydro) { cudaSetDevice(h->device); const int rcp_ = finish_steps(h); if (rcp_) return rcp_; }
  if (h->pending) {
    // Steps are in flight (rcm_dyn_step_async): report, WITHOUT blocking, the newest step whose verdict copy has
    // arrived in the pinned ring (this rank's values; rcm_dyn_sync gives the rank-reduced verdict of the batch).
    for (int64_t q = h->vhead - 1; q > h->vdone && q >= h->vhead - rcm_dyn::kVRing; --q)
      if (cudaEventQuery(h->vev[q % rcm_dyn::kVRing]) == cudaSuccess) { h->vdone = q; break; }
    cudaGetLastError();   // cudaErrorNotReady is not an error
    d->lcount = h->lcount - h->pending_steps;   // nothing of this batch has landed yet: the last finished step
    if (h->vdone >= 0 && h->vdone >= h->vhead - rcm_dyn::kVRing) {
      const rcm_dyn::VSlot& v = h->vring[h->vdone % rcm_dyn::kVRing];
      d->lcount = h->vlcount[h->vdone % rcm_dyn::kVRing];
      d->nan_flag = (int)(v.nanf & 0xffffffffULL);
      if (h->moloch) {
        if (v.ps[0]) { const unsigned long long mn = ~v.ps[1]; std::memcpy(&d->ps_max, &v.ps[0], 8); std::memcpy(&d->ps_min, &mn, 8); }
        return RCM_OK;
      }
      std::memcpy(&d->max_cfl, &v.cfl[0], 8);
      d->neg_q_points = (int64_t)v.negq;
    }
    return RCM_OK;
  }
  if (h->hydro && h->inited) {
    // Bleck (1977) noise parameters: per-point terms summed on the host in the
    // reference's loop order (i outer, j inner; Main/mod_tendency.F90:1452-1458)
    cudaSetDevice(h->device);
    std::vector<double> a, b;
    int rc;
    if ((rc = download2(h, h->f.ptn2d, a)) || (rc = download2(h, h->f.pt22d, b))) return rc;
    double ptn = 0.0, pt2 = 0.0;
    for (int i = h->g.ici1; i <= h->g.ici2; ++i)
      for (int j = h->g.jci1; j <= h->g.jci2; ++j) { ptn = ptn + H2(a, j, i); pt2 = pt2 + H2(b, j, i); }
    d->ptntot = ptn; d->pt2tot = pt2;
  }
  if (h->moloch) {   // max, min of ps over the interior of the last step (Main/mod_moloch.F90:437-441), reduced on the device
    d->ps_max = h->ps_max; d->ps_min = h->ps_min;
  }
  return RCM_OK;
}

// ---------------------------------------------------------------------------
// Restart support (what write_savefile / read_savefile need beside the prognostic state)
static int restart_array(rcm_dyn* h, int id, double** p, int* nk) {
  *p = nullptr; *nk = 0;
  if (id == RCM_R_DSTOR || id == RCM_R_HSTOR) {
    if (!h->hydro) return fail(h, RCM_ERR_ARG, "dstor / hstor exist for idynamic = 1 only");
    *p = (id == RCM_R_DSTOR) ? h->f.dstor : h->f.hstor; *nk = h->cfg.nsplit;
    return RCM_OK;
  }
  return fail(h, RCM_ERR_ARG, "bad restart array id");
}
extern "C" int rcm_dyn_get_restart(rcm_dyn* h, int id, double* a, const int32_t lb[3], const int32_t ub[3]) {
  if (!h || !a) return RCM_ERR_ARG;
  if (!h->inited) return fail(h, RCM_ERR_STATE, "rcm_dyn_get_restart before rcm_dyn_init");
  cudaSetDevice(h->device);
  if (h->pending) { const int rcp_ = finish_steps(h); if (rcp_) return rcp_; }
  if (id == RCM_R_TMASK) {
    if (h->hydro || h->moloch) return fail(h, RCM_ERR_ARG, "tmask exists for idynamic = 2 only");
    CUDA_OK(cudaStreamSynchronize(h->st));
    CUDA_OK(cudaMemcpy(a, h->f.tmask, 169 * sizeof(double), cudaMemcpyDeviceToHost));
    return RCM_OK;
  }
  double* p; int nk;
  const int rc = restart_array(h, id, &p, &nk);
  if (rc) return rc;
  return download3(h, p, a, lb, ub, nk);
}
extern "C" int rcm_dyn_set_restart(rcm_dyn* h, int id, const double* a, const int32_t lb[3], const int32_t ub[3]) {
  if (!h || !a) return RCM_ERR_ARG;
  if (!h->inited) return fail(h, RCM_ERR_STATE, "rcm_dyn_set_restart before rcm_dyn_init");
  cudaSetDevice(h->device);
  if (h->pending) { const int rcp_ = finish_steps(h); if (rcp_) return rcp_; }
  if (id == RCM_R_TMASK) {
    if (h->hydro || h->moloch) return fail(h, RCM_ERR_ARG, "tmask exists for idynamic = 2 only");
    CUDA_OK(cudaStreamSynchronize(h->st));
    CUDA_OK(cudaMemcpy(h->f.tmask, a, 169 * sizeof(double), cudaMemcpyHostToDevice));
    return RCM_OK;
  }
  double* p; int nk;
  const int rc = restart_array(h, id, &p, &nk);
  if (rc) return rc;
  return upload3(h, p, a, lb, ub, nk);
}
// ktau and xbctime of the saved run; alarms as the timer would have left them; halo of the restored state
extern "C" int rcm_dyn_set_clock(rcm_dyn* h, int64_t lcount, double xbctime) {
  if (!h || lcount < 0) return RCM_ERR_ARG;
  if (!h->inited) return fail(h, RCM_ERR_STATE, "rcm_dyn_set_clock before rcm_dyn_init");
  cudaSetDevice(h->device);
  if (h->pending) { const int rcp_ = finish_steps(h); if (rcp_) return rcp_; }
  h->lcount = lcount;
  h->model_time = lcount * (int64_t)h->dtsec;
  h->xbctime = xbctime;
  if (!h->moloch) h->dt = (lcount >= 2) ? h->dt2 : h->dtsec;   // leapfrog start-up, Main/mod_tendency.F90:611-616 (MOLOCH: dt = dtsec)
  h->day_lastact = (h->model_time / 86400) * 86400; h->day_now = h->day_lastact;
  const int64_t nb = (int64_t)h->dtbdys;
  h->bdy_lastact = (h->model_time / nb) * nb; h->bdy_now = h->bdy_lastact;
  int rc = RCM_OK;
  if (h->cfg.nproc > 1) {
    if (h->moloch) {
      MoFields& m = h->mf;
      const int kz = h->g.kz;
      rc = mo_exch(h, {{m.u, kz, 1}, {m.v, kz, 1}, {m.w, h->g.kzp1, 0}, {m.t, kz, 0}, {m.qv, kz, 0}, {m.qc, kz, 0}, {m.pai, kz, 0}}, RCM_HALO);
    } else {
      rc = h->hydro ? exch_h_state(h) : exch_state(h);
    }
    if (rc) return rc;
  }
  CUDA_OK(cudaStreamSynchronize(h->st));
  return RCM_OK;
}

extern "C" int rcm_dyn_last_error(rcm_dyn* h, char* buf, int len) {
  if (!buf || len <= 0) return RCM_ERR_ARG;
  const std::string m = h ? h->err : std::string("null handle");
  std::strncpy(buf, m.c_str(), (size_t)len - 1);
  buf[len - 1] = 0;
  return RCM_OK;
}

extern "C" int rcm_dyn_get_table(rcm_dyn* h, int id, double* out, int n) {
  if (!h) return RCM_ERR_ARG;
  if (!h->hydro || !h->inited) return fail(h, RCM_ERR_ARG, "vertical-mode tables exist for idynamic = 1 after rcm_dyn_init");
  const SplitTables& T = h->sp;
  std::vector<double> v;
  switch (id) {
    case RCM_T_HBAR: v.assign(T.hbar.begin() + 1, T.hbar.end()); break;
    case RCM_T_ZMATX: v = T.zmatx; break;
    case RCM_T_ZMATXR: v = T.zmatxr; break;
    case RCM_T_TAU: v = T.tau; break;
    case RCM_T_VARPA1: v = T.varpa1; break;
    case RCM_T_AM: v = T.am; break;
    case RCM_T_AN: v.assign(T.an.begin() + 1, T.an.end()); break;
    case RCM_T_TAU0: v = T.tau0; break;
    default: return fail(h, RCM_ERR_ARG, "bad table id");
  }
  const int m = std::min<int>(n, (int)v.size());
  if (out && m > 0) std::memcpy(out, v.data(), (size_t)m * sizeof(double));
  return (int)v.size();
}

extern "C" int rcm_dyn_get_tile(rcm_dyn* h, int32_t out[4]) {
  if (!h || !out) return RCM_ERR_ARG;
  out[0] = h->g.jde1; out[1] = h->g.jde2; out[2] = h->g.ide1; out[3] = h->g.ide2;
  return RCM_OK;
}

extern "C" int rcm_dyn_profile(rcm_dyn* h, int enable) {
  if (!h) return RCM_ERR_ARG;
  if (h->pending) { const int rcp_ = finish_steps(h); if (rcp_) return rcp_; }
  h->prof_on = enable != 0;
  for (int q = 0; q < KID_COUNT; ++q) { h->prof_ms[q] = 0; h->prof_n[q] = 0; }
  return RCM_OK;
}
extern "C" int rcm_dyn_profile_get(rcm_dyn* h, int kid, double* ms, int64_t* count, char* name, int len) {
  if (!h || kid < 0 || kid >= KID_COUNT) return RCM_ERR_ARG;
  if (ms) *ms = h->prof_ms[kid];
  if (count) *count = h->prof_n[kid];
  if (name && len > 0) { std::strncpy(name, kKernelNames[kid], (size_t)len - 1); name[len - 1] = 0; }
  return RCM_OK;
}
extern "C" int rcm_dyn_profile_kernels(void) { return KID_COUNT; }

extern "C" int64_t rcm_dyn_launch_count(rcm_dyn* h) { return h ? h->launches : 0; }
extern "C" double rcm_dyn_last_step_ms(rcm_dyn* h) { return h ? h->last_ms : 0.0; }

extern "C" int rcm_dyn_comm_init(rcm_dyn* h, const void* nccl_unique_id, int id_bytes) {
  if (!h) return RCM_ERR_ARG;
  cudaSetDevice(h->device);
  if (h->halo.comm_init(nccl_unique_id, id_bytes, h->err)) return RCM_ERR_NCCL;
  return RCM_OK;
}
extern "C" int rcm_dyn_comm_unique_id(void* out, int out_bytes) { return Halo::unique_id(out, out_bytes); }

extern "C" int rcm_dyn_destroy(rcm_dyn* h) {
  if (!h) return RCM_OK;
  cudaSetDevice(h->device);
  h->pending = false;
  cudaDeviceSynchronize();
  if (h->vring) { cudaFreeHost(h->vring); for (cudaEvent_t e : h->vev) if (e) cudaEventDestroy(e); }
  for (int q = 0; q < 2 * RCM_F_COUNT; ++q) {
    if (h->snap[q]) cudaFree(h->snap[q]);
    if (h->ev_snap[q]) cudaEventDestroy(h->ev_snap[q]);
    if (h->ev_out[q]) cudaEventDestroy(h->ev_out[q]);
  }
  if (h->st_out) cudaStreamDestroy(h->st_out);
  if (h->stage_dev) cudaFree(h->stage_dev);
  h->halo.destroy();
  for (cudaEvent_t e : h->prof_pool) cudaEventDestroy(e);
  for (void* p : h->allocs) cudaFree(p);
  if (h->ev0) cudaEventDestroy(h->ev0);
  if (h->ev1) cudaEventDestroy(h->ev1);
  if (h->st) cudaStreamDestroy(h->st);
  if (h->uv_maps) uv_tma_destroy(h->uv_maps);
  if (h->st_copy) { cudaStreamSynchronize(h->st_copy); cudaStreamDestroy(h->st_copy); }
  if (h->ev_stage) cudaEventDestroy(h->ev_stage);
  if (h->st_b) cudaStreamDestroy(h->st_b);
  if (h->st_c) cudaStreamDestroy(h->st_c);
  for (cudaEvent_t e : {h->ev_fork, h->ev_jb, h->ev_jc, h->ev_prep}) if (e) cudaEventDestroy(e);
  delete h;
  return RCM_OK;
}

#include "mo_core.inc"
