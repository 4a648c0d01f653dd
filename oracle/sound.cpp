// ORACLE — TEST INFRASTRUCTURE ONLY.
// Restates `sound` (Main/mod_sound.F90:163-718): acoustic sub-stepping with
// the column-implicit (Ikawa beta) tridiagonal solve for w, the upper
// radiative boundary condition, the CFL check and the closing Asselin filters.
#include <algorithm>
#include <cmath>

#include "model.hpp"

namespace orc {

void Oracle::sound() {
  const int pcnt = 0; (void)pcnt;
  A3& atm1_rho = a1_rho; A3& atm1_pr = a1_pr;
  cfl_error = false;
  int istep = std::max((int)(dt / dtsmax), 2);
  if (integrating()) istep = std::max(4, istep);
  const double dts = dt / (double)istep;
  last_istep = istep;
  max_cfl_step = 0.0;
  const double xkd = xkd_snd;
  const double gnu1 = cfg.gnu1, gnu2 = cfg.gnu2;

  for (int i = ice1; i <= ice2; ++i)
    for (int j = jce1; j <= jce2; ++j) s_rpsb(j, i) = d_one / psb(j, i);
  for (int k = 1; k <= kz; ++k)
    for (int i = ide1; i <= ide2; ++i)
      for (int j = jde1; j <= jde2; ++j) {
        ac_u(j, i, k) = a2_u(j, i, k) / psdotb(j, i);
        ac_v(j, i, k) = a2_v(j, i, k) / psdotb(j, i);
      }
  for (int k = 1; k <= kz; ++k)
    for (int i = idi1; i <= idi2; ++i)
      for (int j = jdi1; j <= jdi2; ++j) {
        uten(j, i, k) = uten(j, i, k) * dts;
        vten(j, i, k) = vten(j, i, k) * dts;
      }
  for (int k = 1; k <= kz; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) ac_qx(j, i, k, iqv) = a2_qx(j, i, k, iqv) * s_rpsb(j, i);
  for (int k = 1; k <= kz; ++k)
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j) ac_pp(j, i, k) = a2_pp(j, i, k) * s_rpsb(j, i);
  for (int k = 1; k <= kz; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) ppten(j, i, k) = ppten(j, i, k) * dts;
  for (int k = 1; k <= kzp1; ++k)
    for (int i = ice1; i <= ice2; ++i)
      for (int j = jce1; j <= jce2; ++j) ac_w(j, i, k) = a2_w(j, i, k) * s_rpsb(j, i);
  for (int k = 1; k <= kzp1; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) wten(j, i, k) = wten(j, i, k) * dts;

  for (int it = 1; it <= istep; ++it) {
    if (it > 1)
      for (int k = 1; k <= kz; ++k)
        for (int i = ici1; i <= ici2; ++i)
          for (int j = jci1; j <= jci2; ++j) ac_pp(j, i, k) = ac_pp(j, i, k) + xkd * s_pi(j, i, k);
    for (int k = 1; k <= kz; ++k) {
      int kp1 = std::min(kz, k + 1), km1 = std::max(1, k - 1);
      for (int i = ice1; i <= ice2; ++i)
        for (int j = jce1; j <= jce2; ++j)
          ac_t(j, i, k) = (ac_pp(j, i, km1) - ac_pp(j, i, kp1)) / (a0_pr(j, i, km1) - a0_pr(j, i, kp1));
    }
    // Advance u and v (:270-299)
    for (int k = 1; k <= kz; ++k)
      for (int i = idi1; i <= idi2; ++i)
        for (int j = jdi1; j <= jdi2; ++j) {
          double rho = d_rfour * (atm1_rho(j, i, k) + atm1_rho(j - 1, i, k) + atm1_rho(j, i - 1, k) + atm1_rho(j - 1, i - 1, k));
          double dppdp0 = d_rfour * (ac_t(j, i, k) + ac_t(j - 1, i, k) + ac_t(j, i - 1, k) + ac_t(j - 1, i - 1, k));
          double chh = d_half * dts / (rho * dx) / msfd(j, i);
          ac_u(j, i, k) = ac_u(j, i, k) -
              chh * (ac_pp(j, i, k) - ac_pp(j - 1, i, k) + ac_pp(j, i - 1, k) - ac_pp(j - 1, i - 1, k) -
                     a0_dprddx(j, i, k) * dppdp0);
          ac_v(j, i, k) = ac_v(j, i, k) -
              chh * (ac_pp(j, i, k) - ac_pp(j, i - 1, k) + ac_pp(j - 1, i, k) - ac_pp(j - 1, i - 1, k) -
                     a0_dprddy(j, i, k) * dppdp0);
        }
    for (int k = 1; k <= kz; ++k)
      for (int i = idi1; i <= idi2; ++i)
        for (int j = jdi1; j <= jdi2; ++j) {
          ac_u(j, i, k) = ac_u(j, i, k) + uten(j, i, k);
          ac_v(j, i, k) = ac_v(j, i, k) + vten(j, i, k);
        }
    if (it > 1)
      for (int k = 1; k <= kz; ++k)
        for (int i = ici1; i <= ici2; ++i)
          for (int j = jci1; j <= jci2; ++j) ac_pp(j, i, k) = ac_pp(j, i, k) - xkd * s_pi(j, i, k);
    // Semi-implicit solution for w and p (:317-464)
    for (int k = 1; k <= kzp1; ++k)
      for (int i = ici1; i <= ici2; ++i)
        for (int j = jci1; j <= jci2; ++j) s_wo(j, i, k) = ac_w(j, i, k);
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) {
        ac_w(j, i, kzp1) = d_half * d_rfour * regrav *
            ((ac_v(j, i + 1, kz) + ac_v(j, i, kz) + ac_v(j + 1, i + 1, kz) + ac_v(j + 1, i, kz)) *
                 (ht(j, i + 1) - ht(j, i - 1)) +
             (ac_u(j, i + 1, kz) + ac_u(j, i, kz) + ac_u(j + 1, i + 1, kz) + ac_u(j + 1, i, kz)) *
                 (ht(j + 1, i) - ht(j - 1, i))) /
            (dx * msfx(j, i));
        s_e(j, i, kz) = d_zero;
        s_f(j, i, kz) = ac_w(j, i, kzp1);
      }
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) {
        s_cc(j, i, 1) = xgamma * atm1_pr(j, i, 1) * dts / (dx * msfx(j, i));
        s_cdd(j, i, 1) = xgamma * atm1_pr(j, i, 1) * a0_rho(j, i, 1) * egrav * dts / (a0_ps(j, i) * dsigma(1));
        s_cj(j, i, 1) = d_half * a0_rho(j, i, 1) * egrav * dts;
        s_pxup(j, i, 1) = 0.0625 * (a0_pr(j + 1, i, 1) - a0_pr(j - 1, i, 1)) *
            (ac_u(j, i, 1) + ac_u(j + 1, i, 1) + ac_u(j, i + 1, 1) + ac_u(j + 1, i + 1, 1) -
             ac_u(j, i, 2) - ac_u(j + 1, i, 2) - ac_u(j, i + 1, 2) - ac_u(j + 1, i + 1, 2)) /
            (a0_pr(j, i, 1) - a0_pr(j, i, 2));
        s_pyvp(j, i, 1) = 0.0625 * (a0_pr(j, i + 1, 1) - a0_pr(j, i - 1, 1)) *
            (ac_v(j, i, 1) + ac_v(j + 1, i, 1) + ac_v(j, i + 1, 1) + ac_v(j + 1, i + 1, 1) -
             ac_v(j, i, 2) - ac_v(j + 1, i, 2) - ac_v(j, i + 1, 2) - ac_v(j + 1, i + 1, 2)) /
            (a0_pr(j, i, 1) - a0_pr(j, i, 2));
        s_ptend(j, i, 1) = ppten(j, i, 1) - d_half * s_cc(j, i, 1) *
            ((ac_v(j, i + 1, 1) * msfd(j, i + 1) - ac_v(j, i, 1) * msfd(j, i) +
              ac_v(j + 1, i + 1, 1) * msfd(j + 1, i + 1) - ac_v(j + 1, i, 1) * msfd(j + 1, i) +
              ac_u(j + 1, i, 1) * msfd(j + 1, i) - ac_u(j, i, 1) * msfd(j, i) +
              ac_u(j + 1, i + 1, 1) * msfd(j + 1, i + 1) - ac_u(j, i + 1, 1) * msfd(j, i + 1)) /
                 msfx(j, i) -
             d_two * (s_pyvp(j, i, 1) + s_pxup(j, i, 1)));
        s_tk(j, i, 1) = (d_half * a0_ps(j, i) * a0_t(j, i, 1)) /
                        (xgamma * a0_pr(j, i, 1) * a2_t(j, i, 1) * s_rpsb(j, i));
      }
    for (int k = 2; k <= kz; ++k) {
      int kp1 = std::min(k + 1, kz), km1 = k - 1;
      for (int i = ici1; i <= ici2; ++i)
        for (int j = jci1; j <= jci2; ++j) {
          s_tk(j, i, k) = (d_half * a0_ps(j, i) * a0_t(j, i, k)) /
                          (xgamma * a0_pr(j, i, k) * a2_t(j, i, k) * s_rpsb(j, i));
          double rofac = (dsigma(km1) * a0_rho(j, i, k) + dsigma(k) * a0_rho(j, i, km1)) /
                         (dsigma(km1) * atm1_rho(j, i, k) + dsigma(k) * atm1_rho(j, i, km1));
          s_cc(j, i, k) = xgamma * atm1_pr(j, i, k) * dts / (dx * msfx(j, i));
          s_cdd(j, i, k) = xgamma * atm1_pr(j, i, k) * a0_rho(j, i, k) * egrav * dts / (a0_ps(j, i) * dsigma(k));
          s_cj(j, i, k) = d_half * a0_rho(j, i, k) * egrav * dts;
          s_ca(j, i, k) = egrav * dts / (a0_pr(j, i, k) - a0_pr(j, i, km1)) * rofac;
          s_g1(j, i, k) = d_one - dsigma(km1) * s_tk(j, i, k);
          s_g2(j, i, k) = d_one + dsigma(k) * s_tk(j, i, km1);
          s_c(j, i, k) = -s_ca(j, i, k) * (s_cdd(j, i, km1) - s_cj(j, i, km1)) * s_g2(j, i, k) * bpxbp;
          s_b(j, i, k) = d_one + s_ca(j, i, k) * (s_g1(j, i, k) * (s_cdd(j, i, k) - s_cj(j, i, k)) +
                                                  s_g2(j, i, k) * (s_cdd(j, i, km1) + s_cj(j, i, km1))) * bpxbp;
          s_aa(j, i, k) = -s_ca(j, i, k) * (s_cdd(j, i, k) + s_cj(j, i, k)) * s_g1(j, i, k) * bpxbp;
          s_pyvp(j, i, k) = 0.125 * (a0_pr(j, i + 1, k) - a0_pr(j, i - 1, k)) *
              (ac_v(j, i, km1) + ac_v(j + 1, i, km1) + ac_v(j, i + 1, km1) + ac_v(j + 1, i + 1, km1) -
               ac_v(j, i, kp1) - ac_v(j + 1, i, kp1) - ac_v(j, i + 1, kp1) - ac_v(j + 1, i + 1, kp1)) /
              (a0_pr(j, i, km1) - a0_pr(j, i, kp1));
          s_pxup(j, i, k) = 0.125 * (a0_pr(j + 1, i, k) - a0_pr(j - 1, i, k)) *
              (ac_u(j, i, km1) + ac_u(j + 1, i, km1) + ac_u(j, i + 1, km1) + ac_u(j + 1, i + 1, km1) -
               ac_u(j, i, kp1) - ac_u(j + 1, i, kp1) - ac_u(j, i + 1, kp1) - ac_u(j + 1, i + 1, kp1)) /
              (a0_pr(j, i, km1) - a0_pr(j, i, kp1));
        }
    }
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) {
        s_pyvp(j, i, kz) = s_pyvp(j, i, kz) * d_half;
        s_pxup(j, i, kz) = s_pxup(j, i, kz) * d_half;
      }
    for (int k = 2; k <= kz; ++k)
      for (int i = ici1; i <= ici2; ++i)
        for (int j = jci1; j <= jci2; ++j) {
          s_ptend(j, i, k) = ppten(j, i, k) - d_half * s_cc(j, i, k) *
              ((ac_v(j, i + 1, k) * msfd(j, i + 1) - ac_v(j, i, k) * msfd(j, i) +
                ac_v(j + 1, i + 1, k) * msfd(j + 1, i + 1) - ac_v(j + 1, i, k) * msfd(j + 1, i) +
                ac_u(j + 1, i, k) * msfd(j + 1, i) - ac_u(j, i, k) * msfd(j, i) +
                ac_u(j + 1, i + 1, k) * msfd(j + 1, i + 1) - ac_u(j, i + 1, k) * msfd(j, i + 1)) /
                   msfx(j, i) -
               d_two * (s_pyvp(j, i, k) + s_pxup(j, i, k)));
          s_rhs(j, i, k) = ac_w(j, i, k) + wten(j, i, k) +
              s_ca(j, i, k) * (bpxbm * ((s_cdd(j, i, k - 1) - s_cj(j, i, k - 1)) * s_g2(j, i, k) * s_wo(j, i, k - 1) -
                                        ((s_cdd(j, i, k - 1) + s_cj(j, i, k - 1)) * s_g2(j, i, k) +
                                         (s_cdd(j, i, k) - s_cj(j, i, k)) * s_g1(j, i, k)) * s_wo(j, i, k) +
                                        (s_cdd(j, i, k) + s_cj(j, i, k)) * s_g1(j, i, k) * s_wo(j, i, k + 1)) +
                               (ac_pp(j, i, k) * s_g1(j, i, k) - ac_pp(j, i, k - 1) * s_g2(j, i, k)) +
                               (s_g1(j, i, k) * s_ptend(j, i, k) - s_g2(j, i, k) * s_ptend(j, i, k - 1)) * bp);
        }
    for (int k = 1; k <= kz; ++k)
      for (int i = ici1; i <= ici2; ++i)
        for (int j = jci1; j <= jci2; ++j) {
          s_pi(j, i, k) = ac_pp(j, i, k);
          ac_pp(j, i, k) = ac_pp(j, i, k) + s_ptend(j, i, k) +
              (s_cj(j, i, k) * (s_wo(j, i, k + 1) + s_wo(j, i, k)) +
               s_cdd(j, i, k) * (s_wo(j, i, k + 1) - s_wo(j, i, k))) * bm;
        }
    // Upward calculation of coefficients (:468-476)
    for (int k = kz; k >= 2; --k)
      for (int i = ici1; i <= ici2; ++i)
        for (int j = jci1; j <= jci2; ++j) {
          double denom = s_aa(j, i, k) * s_e(j, i, k) + s_b(j, i, k);
          s_e(j, i, k - 1) = -s_c(j, i, k) / denom;
          s_f(j, i, k - 1) = (s_rhs(j, i, k) - s_f(j, i, k) * s_aa(j, i, k)) / denom;
        }
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) s_wpval(j, i) = d_zero;
    if (cfg.ifupr == 1) {
      for (int i = ici1; i <= ici2; ++i)
        for (int j = jci1; j <= jci2; ++j) {
          double denom = (s_cdd(j, i, 1) + s_cj(j, i, 1)) * bp;
          s_estore(j, i) = ac_pp(j, i, 1) + s_f(j, i, 1) * denom;
          s_astore(j, i) = denom * s_e(j, i, 1) + (s_cj(j, i, 1) - s_cdd(j, i, 1)) * bp;
        }
      // grid_collect + bcast: single rank
      for (int i = ici1; i <= ici2; ++i)
        for (int j = jci1; j <= jci2; ++j) s_estore_g(j, i) = s_estore(j, i);
      if (alarm_day_act() && it == 1) {
        double atot = d_zero, rhontot = d_zero;
        for (int i = ici1; i <= ici2; ++i)
          for (int j = jci1; j <= jci2; ++j) {
            atot = atot + s_astore(j, i);
            double ensq = egrav * egrav / cpd / (a2_t(j, i, 1) * s_rpsb(j, i));
            rhontot = rhontot + atm1_rho(j, i, 1) * std::sqrt(ensq);
          }
        double abar = atot * rnpts, rhon = rhontot * rnpts;
        double dxmsfb = d_two / dxsq / xmsf_mean;
        tmask.fill(d_zero);
        for (int kk = 0; kk <= 6; ++kk) {
          double rkk = (double)kk;
          for (int ll = 0; ll <= 6; ++ll) {
            double rll = (double)ll;
            double xkeff = dxmsfb * std::sin(mathpi * rkk / 12.0) * std::cos(mathpi * rll / 12.0);
            double xleff = dxmsfb * std::sin(mathpi * rll / 12.0) * std::cos(mathpi * rkk / 12.0);
            double xkleff = std::sqrt(xkeff * xkeff + xleff * xleff);
            for (int i = -6; i <= 6; ++i) {
              double ri = (double)i;
              for (int j = -6; j <= 6; ++j) {
                double rj = (double)j;
                tmask(j, i) = tmask(j, i) + (fi[i + 6] * fj[j + 6] * fk[kk] * fl[ll]) / 144.0 *
                                                std::cos(2.0 * mathpi * rkk * ri / 12.0) *
                                                std::cos(2.0 * mathpi * rll * rj / 12.0) * xkleff /
                                                (rhon - abar * xkleff);
              }
            }
          }
        }
      }
      for (int i = ici1; i <= ici2; ++i)
        for (int j = jci1; j <= jci2; ++j)
          for (int nsi = -6; nsi <= 6; ++nsi) {
            int inn = std::min(std::max(icross1 + 1, i + nsi), icross2 - 1);
            for (int nsj = -6; nsj <= 6; ++nsj) {
              int jnn = std::min(std::max(jcross1 + 1, j + nsj), jcross2 - 1);
              s_wpval(j, i) = s_wpval(j, i) + s_estore_g(jnn, inn) * tmask(nsj, nsi);
            }
          }
    }
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) ac_w(j, i, 1) = s_wpval(j, i);
    for (int k = 1; k <= kz; ++k)
      for (int i = ici1; i <= ici2; ++i)
        for (int j = jci1; j <= jci2; ++j) ac_w(j, i, k + 1) = s_e(j, i, k) * ac_w(j, i, k) + s_f(j, i, k);
    // Zero-out gradient for W (:578-609)
    for (int k = 1; k <= kzp1; ++k) {
      for (int j = jci1; j <= jci2; ++j) ac_w(j, ice1, k) = ac_w(j, ici1, k);
      ac_w(jce1, ice1, k) = ac_w(jci1, ici1, k);
      ac_w(jce2, ice1, k) = ac_w(jci2, ici1, k);
      for (int j = jci1; j <= jci2; ++j) ac_w(j, ice2, k) = ac_w(j, ici2, k);
      ac_w(jce1, ice2, k) = ac_w(jci1, ici2, k);
      ac_w(jce2, ice2, k) = ac_w(jci2, ici2, k);
      for (int i = ici1; i <= ici2; ++i) ac_w(jce1, i, k) = ac_w(jci1, i, k);
      for (int i = ici1; i <= ici2; ++i) ac_w(jce2, i, k) = ac_w(jci2, i, k);
    }
    // Check CFL (:613-665)
    for (int k = 1; k <= kz; ++k)
      for (int i = ici1; i <= ici2; ++i)
        for (int j = jci1; j <= jci2; ++j) {
          s_ucrs(j, i, k) = ac_u(j, i, k) + ac_u(j, i + 1, k) + ac_u(j + 1, i, k) + ac_u(j + 1, i + 1, k);
          s_vcrs(j, i, k) = ac_v(j, i, k) + ac_v(j, i + 1, k) + ac_v(j + 1, i, k) + ac_v(j + 1, i + 1, k);
        }
    double cfl = d_zero;
    for (int k = kz; k >= 2; --k)
      for (int i = ici1; i <= ici2; ++i)
        for (int j = jci1; j <= jci2; ++j) {
          s_sigdot(j, i, k) = -a0_rhof(j, i, k) * egrav * ac_w(j, i, k) / a0_ps(j, i) -
              sigma(k) * (dpsdxm(j, i) * (twt(k, 1) * s_ucrs(j, i, k) + twt(k, 2) * s_ucrs(j, i, k - 1)) +
                          dpsdym(j, i) * (twt(k, 1) * s_vcrs(j, i, k) + twt(k, 2) * s_vcrs(j, i, k - 1)));
          double check = std::fabs(s_sigdot(j, i, k)) * dt / (dsigma(k) + dsigma(k - 1));
          cfl = std::max(check, cfl);
        }
    max_cfl_step = std::max(max_cfl_step, cfl);
    if (cfl < d_one) cfl_error = false;
    if (cfl > d_one) cfl_error = true;
    if (cfl_error) return;  // call fatal(...,'CFL violation')
    // Now compute the new pressure (:669-689)
    for (int k = 1; k <= kz; ++k)
      for (int i = ici1; i <= ici2; ++i)
        for (int j = jci1; j <= jci2; ++j) {
          double ppold = s_pi(j, i, k);
          double cddtmp = xgamma * atm1_pr(j, i, k) * a0_rho(j, i, k) * egrav * dts / (a0_ps(j, i) * dsigma(k));
          double cjtmp = a0_rho(j, i, k) * egrav * dts * d_half;
          ac_pp(j, i, k) = ac_pp(j, i, k) +
              (cjtmp * (ac_w(j, i, k + 1) + ac_w(j, i, k)) + cddtmp * (ac_w(j, i, k + 1) - ac_w(j, i, k))) * bp;
          s_pi(j, i, k) = ac_pp(j, i, k) - ppold - ppten(j, i, k);
          double cpm = cpmf(ac_qx(j, i, k, iqv));
          double dpterm = psb(j, i) * (ac_pp(j, i, k) - ppold) / (cpm * atm1_rho(j, i, k));
          a2_t(j, i, k) = a2_t(j, i, k) + gnu1 * dpterm;
          a1_t(j, i, k) = a1_t(j, i, k) + dpterm;
        }
  }
  // Apply time filtering technique (:696-712)
  for (int k = 1; k <= kz; ++k)
    for (int i = idi1; i <= idi2; ++i)
      for (int j = jdi1; j <= jdi2; ++j) {
        ac_u(j, i, k) = psdotb(j, i) * ac_u(j, i, k);
        ac_v(j, i, k) = psdotb(j, i) * ac_v(j, i, k);
      }
  filter_ra_uv(a1_u, a2_u, ac_u, a1_v, a2_v, ac_v, gnu1);
  for (int k = 1; k <= kz; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) ac_pp(j, i, k) = psb(j, i) * ac_pp(j, i, k);
  filter_ra_3d(a1_pp, a2_pp, ac_pp, gnu1);
  for (double& x : ac_w.d) if (std::fabs(x) < dlowval) x = d_zero;
  for (int k = 1; k <= kzp1; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) ac_w(j, i, k) = psb(j, i) * ac_w(j, i, k);
  filter_ra_3d(a1_w, a2_w, ac_w, gnu2);
  for (double& x : a2_w.d) if (std::fabs(x) < dlowval) x = d_zero;
  for (double& x : a1_w.d) if (std::fabs(x) < dlowval) x = d_zero;
}

}  // namespace orc
