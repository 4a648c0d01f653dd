// ORACLE — TEST INFRASTRUCTURE ONLY.
// Restates Main/mod_timefilter.F90: filter_ra_3d :92-110, filter_ra_uv
// :132-156, filter_raw_4d :231-255, filter_raw_qv :257-278.
#include <algorithm>

#include "model.hpp"

namespace orc {

void Oracle::filter_ra_3d(A3& phin, A3& phinm1, const A3& phinp1, double alpha) {
  const int nk = phin.n3;
  for (int k = 1; k <= nk; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) {
        double d = alpha * (phinp1(j, i, k) + phinm1(j, i, k) - d_two * phin(j, i, k));
        phinm1(j, i, k) = phin(j, i, k) + d;
        phin(j, i, k) = phinp1(j, i, k);
      }
}

void Oracle::filter_ra_uv(A3& un, A3& unm1, const A3& unp1, A3& vn, A3& vnm1, const A3& vnp1, double alpha) {
  const int nk = un.n3;
  for (int k = 1; k <= nk; ++k)
    for (int i = idi1; i <= idi2; ++i)
      for (int j = jdi1; j <= jdi2; ++j) {
        double d = alpha * (unp1(j, i, k) + unm1(j, i, k) - d_two * un(j, i, k));
        unm1(j, i, k) = un(j, i, k) + d;
        un(j, i, k) = unp1(j, i, k);
        d = alpha * (vnp1(j, i, k) + vnm1(j, i, k) - d_two * vn(j, i, k));
        vnm1(j, i, k) = vn(j, i, k) + d;
        vn(j, i, k) = vnp1(j, i, k);
      }
}

void Oracle::filter_raw_qv(A4& phin, A4& phinm1, const A4& phinp1, double alpha, double beta, const A2& psa_, const A2& psb_) {
  const int nk = phin.n3;
  for (int k = 1; k <= nk; ++k)
    for (int i = ici1; i <= ici2; ++i)
      for (int j = jci1; j <= jci2; ++j) {
        double d = alpha * (phinp1(j, i, k, iqv) + phinm1(j, i, k, iqv) - d_two * phin(j, i, k, iqv));
        phinm1(j, i, k, iqv) = std::max(phin(j, i, k, iqv) + beta * d, minqq * psa_(j, i));
        phin(j, i, k, iqv) = std::max(phinp1(j, i, k, iqv) + (beta - d_one) * d, minqq * psb_(j, i));
      }
}

void Oracle::filter_raw_4d(A4& phin, A4& phinm1, const A4& phinp1, double alpha, double beta, int n1, int n2, double low) {
  const int nk = phin.n3;
  for (int n = n1; n <= n2; ++n)
    for (int k = 1; k <= nk; ++k)
      for (int i = ici1; i <= ici2; ++i)
        for (int j = jci1; j <= jci2; ++j) {
          double d = alpha * (phinp1(j, i, k, n) + phinm1(j, i, k, n) - d_two * phin(j, i, k, n));
          phinm1(j, i, k, n) = phin(j, i, k, n) + beta * d;
          phin(j, i, k, n) = phinp1(j, i, k, n) + (beta - d_one) * d;
          if (phinm1(j, i, k, n) < low) phinm1(j, i, k, n) = low;
          if (phin(j, i, k, n) < low) phin(j, i, k, n) = low;
        }
}

}  // namespace orc
