// ORACLE — TEST INFRASTRUCTURE ONLY.
// Physical constants restated from Share/mod_constants.F90:20-375.  Derived
// constants are computed in the same order as the reference (not typed as
// rounded literals) so that their last bits agree.
#pragma once
#include <cmath>

namespace orc {

constexpr double d_zero = 0.0, d_one = 1.0, d_two = 2.0, d_four = 4.0;
constexpr double d_half = 0.5, d_rfour = 0.25;
constexpr double d_1000 = 1000.0, d_r1000 = 0.001;

constexpr double mintr = 1.0e-30;    // mod_constants.F90:63 (double precision)
constexpr double minqq = 1.0e-8;    // mod_constants.F90:57
constexpr double dlowval = 1.0e-20; // :68

constexpr double egrav = 9.80665;             // :85
constexpr double boltzk = 1.3806504e-23;      // :92
constexpr double navgdr = 6.02214129e23;      // :94
constexpr double amd = 28.96454;              // :107
constexpr double amw = 18.01528;              // :109
constexpr double rgasmol = navgdr * boltzk;   // :130
constexpr double c287 = rgasmol / amd;        // :132
constexpr double rgas = c287 * 1000.0;        // :134
constexpr double cpd = 3.5 * rgas;            // :144
constexpr double cvd = 2.5 * rgas;            // :145
constexpr double regrav = d_one / egrav;      // :182
// pressure gradient force with ipgf = 1 (Share/mod_constants.F90:359-362)
constexpr double t00pg = 287.0, p00pg = 101.325, alam = 6.5e-3, pgfaa1 = alam * rgas * regrav;
constexpr double rcpd = d_one / cpd;          // :183
constexpr double rovcp = rgas * rcpd;         // :184
constexpr double rovg = rgas / egrav;         // :187
constexpr double rdrcv = rgas / cvd;          // :185
constexpr double cpovr = cpd / rgas;          // :186
constexpr double govr = egrav / rgas;         // :188
constexpr double tiso = 216.65;               // :198
constexpr double p00 = 1.0e5;                 // :229
constexpr double mathpi = 3.1415926535897932384626433832795029;  // :254
constexpr double halfpi = mathpi * d_half;    // :257
constexpr double degrad = mathpi / 180.0;     // :261
constexpr double earthrad = 6.371229e6;       // :282
constexpr double rearthrad = d_one / earthrad;  // :284
constexpr double eomeg = 7.2921159e-5;        // :286
constexpr double eomeg2 = d_two * eomeg;      // :287
constexpr double vonkar = 0.4;                // :296
constexpr double ep1 = amd / amw - d_one;     // :303
constexpr double ep2 = amw / amd;             // :304
constexpr double minqc = 1.0e-10;             // :58
constexpr double tzero = 273.15;

// Share/cpmf.inc: heat capacity of moist air at constant pressure
inline double cpmf(double qv) { return cpd * (d_one + 0.80 * qv); }

}  // namespace orc
