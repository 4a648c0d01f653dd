// Halo exchange of the 2-D horizontal decomposition (SURVEY.md §8 a24,
// Main/mpplib/mod_mppparam.F90 exchange*/sumall/maxall): packed multi-field
// edge buffers moved with NCCL send/recv over NVLink.  One rank = one GPU.
// With nproc == 1 every call is a no-op and NCCL is never loaded.
#pragma once
#include "launch.h"

#include <string>
#include <vector>

#include "dyn_types.h"

struct HaloField {
  double* p;   // origin-shifted field
  int nk;      // number of k planes to exchange
  int dot;     // 1: dot grid extents, 0: cross grid extents
};

struct Halo {
  Geom g{};
  int nproc = 1, myid = 0, cj = 1, ci = 1, pj = 0, pi = 0;
  int nb_w = -1, nb_e = -1, nb_s = -1, nb_n = -1;  // neighbour ranks (-1: physical boundary)
  void* comm = nullptr;       // ncclComm_t
  void* lib = nullptr;        // dlopen handle of libnccl
  double *sbuf[4] = {nullptr, nullptr, nullptr, nullptr}, *rbuf[4] = {nullptr, nullptr, nullptr, nullptr};
  size_t cap = 0;             // capacity (doubles) of each edge buffer
  double* red = nullptr;      // small device buffer for scalar reductions
  long long launches = 0;

  // ---- NVLink peer-memory transport (one GPU per rank, all on one node) ----
  // Each rank owns one device block: 2 parities x 8 directions of receive
  // buffers + 8 arrival flags.  It is exported with CUDA IPC, so that the pack
  // kernel of a neighbour stores its edge straight into this rank's receive
  // buffer over NVLink and then raises the flag; the unpack kernel waits on the
  // flags.  One phase, 8 neighbours, no NCCL call on the step path.
  bool p2p = false;
  char* pblock = nullptr;          // my block
  char* peer_block[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};  // neighbour blocks (mapped)
  int nb8[8] = {-1, -1, -1, -1, -1, -1, -1, -1};  // W, E, S, N, SW, SE, NW, NE
  size_t pcap = 0;                 // doubles per (parity, direction) buffer
  unsigned long long seq = 0;      // exchange sequence number (flag value)
  int* errflag = nullptr;          // device: set when a wait timed out
  int setup_p2p(std::string& err);
  int take_error();                // 1 if a wait on a neighbour timed out since the last call
  int exchange_p2p(const std::vector<HaloField>& fl, int width, int estore, const Fields* f, cudaStream_t st);

  void init(const Geom& g_, int nproc_, int myid_, int cj_, int ci_);
  int comm_init(const void* id, int id_bytes, std::string& err);
  static int unique_id(void* out, int out_bytes);
  void destroy();
  // generic exchange of `width` ghost rows for a list of fields, two phases
  // (west/east then south/north including the corners just received)
  int exchange(const std::vector<HaloField>& fl, int width, cudaStream_t st);
  int exchange_estore(const Fields& f, cudaStream_t st);
  int allreduce_host(double* v, int n, int op /*0 sum, 1 max*/);
};
