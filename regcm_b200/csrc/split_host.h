// Tables of the split-explicit scheme computed once on the host
// (Main/mod_vmodes.F90:39-52, Main/mod_split.F90:40-42); see split_host.cu.
#pragma once
#include <vector>

struct SplitTables {
  int kz = 0, nsplit = 0, numerr = 0;
  double pd = 0, xps = 0;
  std::vector<double> hbar;                 // [1..kz]
  std::vector<double> zmatx, zmatxr, tau;   // kz x kz column-major (0-based storage); zmatx scaled by spinit's fac
  std::vector<double> tau0;                 // tau of vmodes before spinit overwrites rows 1..nsplit
  std::vector<double> varpa1;               // kz x (kz+1)
  std::vector<double> am;                   // kz x nsplit
  std::vector<double> an, aam, dtau;        // [1..nsplit]
  std::vector<double> sigmah;               // [1..kz+1]
  std::vector<double> pdlog, eps1;          // [l + nsplit*(k-1)], l = 1..nsplit, k = 1..kz+1
};

int rcm_split_tables(int kz, int nsplit, int lstand, double ptop, double dtsec, const std::vector<double>& sigma,
                     double xps_in, const std::vector<double>& tbarh_in, SplitTables& T);
