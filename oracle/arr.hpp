// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle/README.md).
// Fortran-style dense arrays (column-major, arbitrary lower bounds) so that the
// reference's loop nests can be restated index-for-index.
// Layout follows the reference: (j,i,k[,n]) with j fastest
// (Main/mod_atm_interface.F90:617-641, Share/mod_memutil.F90 getmemNd:
// zero-initialised on allocation, :1356-1374).
#pragma once
#include <cstddef>
#include <cstdio>
#include <cstdlib>
#include <vector>

namespace orc {

#ifdef ORACLE_CHECK
#define ORC_BND(cond, name)                                                   \
  do {                                                                        \
    if (!(cond)) {                                                            \
      std::fprintf(stderr, "oracle: index out of bounds in %s (%s:%d)\n",     \
                   name, __FILE__, __LINE__);                                 \
      std::abort();                                                           \
    }                                                                         \
  } while (0)
#else
#define ORC_BND(cond, name) ((void)0)
#endif

struct A1 {
  int l1 = 1, n1 = 0;
  std::vector<double> d;
  void alloc(int lo, int hi) {
    l1 = lo; n1 = hi - lo + 1; if (n1 < 0) n1 = 0;
    d.assign((size_t)n1, 0.0);
  }
  inline double& operator()(int a) {
    ORC_BND(a >= l1 && a < l1 + n1, "A1");
    return d[(size_t)(a - l1)];
  }
  inline const double& operator()(int a) const {
    ORC_BND(a >= l1 && a < l1 + n1, "A1");
    return d[(size_t)(a - l1)];
  }
};

struct A2 {
  int l1 = 1, l2 = 1, n1 = 0, n2 = 0;
  std::vector<double> d;
  void alloc(int lo1, int hi1, int lo2, int hi2) {
    l1 = lo1; l2 = lo2; n1 = hi1 - lo1 + 1; n2 = hi2 - lo2 + 1;
    if (n1 < 0) n1 = 0; if (n2 < 0) n2 = 0;
    d.assign((size_t)n1 * n2, 0.0);
  }
  inline size_t idx(int a, int b) const {
    ORC_BND(a >= l1 && a < l1 + n1 && b >= l2 && b < l2 + n2, "A2");
    return (size_t)(a - l1) + (size_t)n1 * (size_t)(b - l2);
  }
  inline double& operator()(int a, int b) { return d[idx(a, b)]; }
  inline const double& operator()(int a, int b) const { return d[idx(a, b)]; }
  void fill(double v) { d.assign(d.size(), v); }
};

struct A3 {
  int l1 = 1, l2 = 1, l3 = 1, n1 = 0, n2 = 0, n3 = 0;
  std::vector<double> d;
  void alloc(int lo1, int hi1, int lo2, int hi2, int lo3, int hi3) {
    l1 = lo1; l2 = lo2; l3 = lo3;
    n1 = hi1 - lo1 + 1; n2 = hi2 - lo2 + 1; n3 = hi3 - lo3 + 1;
    if (n1 < 0) n1 = 0; if (n2 < 0) n2 = 0; if (n3 < 0) n3 = 0;
    d.assign((size_t)n1 * n2 * n3, 0.0);
  }
  inline size_t idx(int a, int b, int c) const {
    ORC_BND(a >= l1 && a < l1 + n1 && b >= l2 && b < l2 + n2 && c >= l3 &&
                c < l3 + n3, "A3");
    return (size_t)(a - l1) +
           (size_t)n1 * ((size_t)(b - l2) + (size_t)n2 * (size_t)(c - l3));
  }
  inline double& operator()(int a, int b, int c) { return d[idx(a, b, c)]; }
  inline const double& operator()(int a, int b, int c) const {
    return d[idx(a, b, c)];
  }
  void fill(double v) { d.assign(d.size(), v); }
  int u3() const { return l3 + n3 - 1; }
};

struct A4 {
  int l1 = 1, l2 = 1, l3 = 1, l4 = 1, n1 = 0, n2 = 0, n3 = 0, n4 = 0;
  std::vector<double> d;
  void alloc(int lo1, int hi1, int lo2, int hi2, int lo3, int hi3, int lo4,
             int hi4) {
    l1 = lo1; l2 = lo2; l3 = lo3; l4 = lo4;
    n1 = hi1 - lo1 + 1; n2 = hi2 - lo2 + 1; n3 = hi3 - lo3 + 1;
    n4 = hi4 - lo4 + 1;
    if (n1 < 0) n1 = 0; if (n2 < 0) n2 = 0; if (n3 < 0) n3 = 0;
    if (n4 < 0) n4 = 0;
    d.assign((size_t)n1 * n2 * n3 * n4, 0.0);
  }
  inline size_t idx(int a, int b, int c, int e) const {
    ORC_BND(a >= l1 && a < l1 + n1 && b >= l2 && b < l2 + n2 && c >= l3 &&
                c < l3 + n3 && e >= l4 && e < l4 + n4, "A4");
    return (size_t)(a - l1) +
           (size_t)n1 * ((size_t)(b - l2) +
                         (size_t)n2 * ((size_t)(c - l3) +
                                       (size_t)n3 * (size_t)(e - l4)));
  }
  inline double& operator()(int a, int b, int c, int e) {
    return d[idx(a, b, c, e)];
  }
  inline const double& operator()(int a, int b, int c, int e) const {
    return d[idx(a, b, c, e)];
  }
  void fill(double v) { d.assign(d.size(), v); }
};

// logical / integer 2-D maps (boundary-zone masks)
struct I2 {
  int l1 = 1, l2 = 1, n1 = 0, n2 = 0;
  std::vector<int> d;
  void alloc(int lo1, int hi1, int lo2, int hi2, int v = 0) {
    l1 = lo1; l2 = lo2; n1 = hi1 - lo1 + 1; n2 = hi2 - lo2 + 1;
    d.assign((size_t)n1 * n2, v);
  }
  inline int& operator()(int a, int b) {
    ORC_BND(a >= l1 && a < l1 + n1 && b >= l2 && b < l2 + n2, "I2");
    return d[(size_t)(a - l1) + (size_t)n1 * (size_t)(b - l2)];
  }
  inline const int& operator()(int a, int b) const {
    ORC_BND(a >= l1 && a < l1 + n1 && b >= l2 && b < l2 + n2, "I2");
    return d[(size_t)(a - l1) + (size_t)n1 * (size_t)(b - l2)];
  }
};

}  // namespace orc
