// Physical constants of the dyn-core (Share/mod_constants.F90:20-375).
// Derived constants are built with the same operations as the reference so
// that their last bits agree.
#pragma once

#ifdef __CUDACC__
#define RCM_HD __host__ __device__
#else
#define RCM_HD
#endif

namespace rcm {
constexpr double minqq = 1.0e-8;      // mod_constants.F90:57
constexpr double mintr = 1.0e-30;     // mod_constants.F90:63 (double precision)
constexpr double dlowval = 1.0e-20;   // :68
constexpr double egrav = 9.80665;     // :85
constexpr double boltzk = 1.3806504e-23;
constexpr double navgdr = 6.02214129e23;
constexpr double amd = 28.96454;
constexpr double amw = 18.01528;
constexpr double rgasmol = navgdr * boltzk;
constexpr double c287 = rgasmol / amd;
constexpr double rgas = c287 * 1000.0;
constexpr double cpd = 3.5 * rgas;
constexpr double cvd = 2.5 * rgas;
constexpr double rdrcv = rgas / cvd;
constexpr double cpovr = cpd / rgas;
constexpr double minqc = 1.0e-10;
constexpr double tzero = 273.15;
constexpr double regrav = 1.0 / egrav;
// pressure gradient force with ipgf = 1 (Share/mod_constants.F90:359-362)
constexpr double t00pg = 287.0, p00pg = 101.325, alam = 6.5e-3, pgfaa1 = alam * rgas * regrav;
constexpr double rcpd = 1.0 / cpd;
constexpr double rovcp = rgas * rcpd;
constexpr double rovg = rgas / egrav;
constexpr double govr = egrav / rgas;
constexpr double tiso = 216.65;
constexpr double p00 = 1.0e5;
constexpr double mathpi = 3.1415926535897932384626433832795029;
constexpr double halfpi = mathpi * 0.5;
constexpr double degrad = mathpi / 180.0;
constexpr double earthrad = 6.371229e6;
constexpr double rearthrad = 1.0 / earthrad;
constexpr double eomeg = 7.2921159e-5;
constexpr double eomeg2 = 2.0 * eomeg;
constexpr double vonkar = 0.4;
constexpr double ep1 = amd / amw - 1.0;
constexpr double ep2 = amw / amd;
// Share/cpmf.inc
RCM_HD inline double cpmf(double qv) { return cpd * (1.0 + 0.80 * qv); }
}  // namespace rcm
