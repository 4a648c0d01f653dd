// Halo exchange over NCCL send/recv (see halo.h).  NCCL is bound at run time
// with dlopen so that single-GPU use needs no NCCL at all; with torch in the
// process the already-loaded libnccl.so.2 is reused.
#include "halo.h"

#include <dlfcn.h>

#include <algorithm>
#include <cstdio>
#include <cstring>

#include "launch.h"

#ifndef RCM_EMU
namespace {
typedef struct { char internal[128]; } ncclUniqueId_;
typedef int (*fn_GetUniqueId)(ncclUniqueId_*);
typedef int (*fn_CommInitRank)(void**, int, ncclUniqueId_, int);
typedef int (*fn_CommDestroy)(void*);
typedef int (*fn_Group)(void);
typedef int (*fn_SendRecv)(void*, size_t, int, int, void*, cudaStream_t);
typedef int (*fn_AllReduce)(const void*, void*, size_t, int, int, void*, cudaStream_t);
typedef const char* (*fn_ErrStr)(int);
struct Nccl {
  void* lib = nullptr;
  fn_GetUniqueId GetUniqueId = nullptr;
  fn_CommInitRank CommInitRank = nullptr;
  fn_CommDestroy CommDestroy = nullptr;
  fn_Group GroupStart = nullptr, GroupEnd = nullptr;
  fn_SendRecv Send = nullptr, Recv = nullptr;
  fn_AllReduce AllReduce = nullptr;
  fn_ErrStr ErrStr = nullptr;
  bool load() {
    if (lib) return true;
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* n : names) { lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL); if (lib) break; }
    if (!lib) return false;
    GetUniqueId = (fn_GetUniqueId)dlsym(lib, "ncclGetUniqueId");
    CommInitRank = (fn_CommInitRank)dlsym(lib, "ncclCommInitRank");
    CommDestroy = (fn_CommDestroy)dlsym(lib, "ncclCommDestroy");
    GroupStart = (fn_Group)dlsym(lib, "ncclGroupStart");
    GroupEnd = (fn_Group)dlsym(lib, "ncclGroupEnd");
    Send = (fn_SendRecv)dlsym(lib, "ncclSend");
    Recv = (fn_SendRecv)dlsym(lib, "ncclRecv");
    AllReduce = (fn_AllReduce)dlsym(lib, "ncclAllReduce");
    ErrStr = (fn_ErrStr)dlsym(lib, "ncclGetErrorString");
    return GetUniqueId && CommInitRank && CommDestroy && GroupStart && GroupEnd && Send && Recv && AllReduce;
  }
};
Nccl g_nccl;
const int kNcclDouble = 8, kNcclSum = 0, kNcclMax = 2;
}  // namespace
#else
// Host harness (tests/emu): the transport is a pair of callbacks the test
// registers (torch.distributed/gloo in tests/test_multirank_emu.py), so that the
// decomposition, the edge-box arithmetic and the two-phase ordering below are
// exercised with world_size > 1 on machines without a GPU.
extern "C" {
typedef int (*rcm_emu_sendrecv_t)(int peer, const double* sendbuf, long long nsend, double* recvbuf, long long nrecv);
typedef int (*rcm_emu_allreduce_t)(double* v, int n, int op);
}
static rcm_emu_sendrecv_t g_emu_sendrecv = nullptr;
static rcm_emu_allreduce_t g_emu_allreduce = nullptr;
extern "C" void rcm_emu_set_transport(rcm_emu_sendrecv_t sr, rcm_emu_allreduce_t ar) { g_emu_sendrecv = sr; g_emu_allreduce = ar; }
#endif

struct Box {
  double* p;      // origin-shifted field
  double* buf;    // packed edge buffer
  long long off;  // first element of this box in buf
  long long sk;
  int si, ja, ia, nj, ni, nk;
};
struct BoxList {
  Box b[36];
  int n;
  long long total;
};

__global__ void k_halo_copy(const BoxList L, const int unpack) {
  long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= L.total) return;
  int q = 0;
  long long base = 0;
  for (; q < L.n; ++q) {
    const long long sz = (long long)L.b[q].nj * L.b[q].ni * L.b[q].nk;
    if (t < base + sz) break;
    base += sz;
  }
  const Box& B = L.b[q];
  const long long r = t - base;
  const int jj = (int)(r % B.nj);
  const long long r2 = r / B.nj;
  const int ii = (int)(r2 % B.ni);
  const int kk = (int)(r2 / B.ni);
  double* fp = B.p + (long long)(B.ja + jj) + (long long)(B.ia + ii) * B.si + (long long)(kk + 1) * B.sk;
  double* bp = B.buf + B.off + r;
  if (unpack) *fp = *bp; else *bp = *fp;
}

void Halo::init(const Geom& g_, int nproc_, int myid_, int cj_, int ci_) {
  g = g_; nproc = nproc_; myid = myid_; cj = cj_; ci = ci_;
  pj = myid % cj; pi = myid / cj;
  nb_w = (pj > 0) ? myid - 1 : -1;
  nb_e = (pj < cj - 1) ? myid + 1 : -1;
  nb_s = (pi > 0) ? myid - cj : -1;
  nb_n = (pi < ci - 1) ? myid + cj : -1;
}

int Halo::unique_id(void* out, int out_bytes) {
#ifdef RCM_EMU
  (void)out; (void)out_bytes; return -5;
#else
  if (out_bytes < 128 || !g_nccl.load()) return -5;
  ncclUniqueId_ id;
  if (g_nccl.GetUniqueId(&id) != 0) return -5;
  std::memcpy(out, &id, 128);
  return 0;
#endif
}

int Halo::comm_init(const void* id, int id_bytes, std::string& err) {
  if (nproc <= 1) return 0;
#ifdef RCM_EMU
  (void)id; (void)id_bytes;
  if (!g_emu_sendrecv || !g_emu_allreduce) { err = "host harness: no transport registered"; return -5; }
  comm = (void*)1;
#else
  if (id_bytes < 128) { err = "NCCL unique id must be 128 bytes"; return -5; }
  if (!g_nccl.load()) { err = "cannot load libnccl.so.2"; return -5; }
  ncclUniqueId_ uid;
  std::memcpy(&uid, id, 128);
  int rc = g_nccl.CommInitRank(&comm, nproc, uid, myid);
  if (rc != 0) { err = std::string("ncclCommInitRank: ") + (g_nccl.ErrStr ? g_nccl.ErrStr(rc) : "error"); return -5; }
#endif
  const int nj = g.jde2 - g.jde1 + 1 + 2 * RCM_EHALO, ni = g.ide2 - g.ide1 + 1 + 2 * RCM_EHALO;
  cap = (size_t)16 * RCM_HALO * (size_t)std::max(nj, ni) * (size_t)g.kzp1;
  for (int q = 0; q < 4; ++q) {
    if (cudaMalloc(&sbuf[q], cap * sizeof(double)) != cudaSuccess || cudaMalloc(&rbuf[q], cap * sizeof(double)) != cudaSuccess) {
      err = "halo buffer allocation failed"; return -5;
    }
  }
  if (cudaMalloc(&red, 16 * sizeof(double)) != cudaSuccess) { err = "halo buffer allocation failed"; return -5; }
  return 0;
}

void Halo::destroy() {
  for (int q = 0; q < 4; ++q) { if (sbuf[q]) cudaFree(sbuf[q]); if (rbuf[q]) cudaFree(rbuf[q]); sbuf[q] = rbuf[q] = nullptr; }
  if (red) cudaFree(red);
  red = nullptr;
#ifndef RCM_EMU
  if (comm && g_nccl.CommDestroy) g_nccl.CommDestroy(comm);
#endif
  comm = nullptr;
}

int Halo::allreduce_host(double* v, int n, int op) {
  if (nproc <= 1) return 0;
  if (!comm || n > 16) return -5;
#ifdef RCM_EMU
  return g_emu_allreduce(v, n, op) ? -5 : 0;
#else
  if (cudaMemcpy(red, v, n * sizeof(double), cudaMemcpyHostToDevice) != cudaSuccess) return -5;
  if (g_nccl.AllReduce(red, red, (size_t)n, kNcclDouble, op == 1 ? kNcclMax : kNcclSum, comm, (cudaStream_t)0) != 0) return -5;
  if (cudaStreamSynchronize((cudaStream_t)0) != cudaSuccess) return -5;
  if (cudaMemcpy(v, red, n * sizeof(double), cudaMemcpyDeviceToHost) != cudaSuccess) return -5;
  return 0;
#endif
}

static int run_phase(Halo& H, BoxList& spack, BoxList& runpack, const long long scount[4], const long long rcount[4],
                     const int nb[4], int side0, cudaStream_t st) {
  if (spack.total > 0) {
    dim3 b(256), gr((unsigned)((spack.total + 255) / 256));
    RCM_LAUNCH(k_halo_copy, gr, b, st, spack, 0);
    H.launches++;
  }
#ifdef RCM_EMU
  for (int s = side0; s < side0 + 2; ++s) {  // low side first on every rank: the chain of pairwise exchanges cannot deadlock
    if (nb[s] < 0) continue;
    if (g_emu_sendrecv(nb[s], H.sbuf[s], scount[s], H.rbuf[s], rcount[s])) return -5;
  }
#else
  g_nccl.GroupStart();
  for (int s = side0; s < side0 + 2; ++s) {
    if (nb[s] < 0) continue;
    if (g_nccl.Send(H.sbuf[s], (size_t)scount[s], kNcclDouble, nb[s], H.comm, st) != 0) return -5;
    if (g_nccl.Recv(H.rbuf[s], (size_t)rcount[s], kNcclDouble, nb[s], H.comm, st) != 0) return -5;
  }
  if (g_nccl.GroupEnd() != 0) return -5;
#endif
  if (runpack.total > 0) {
    dim3 b(256), gr((unsigned)((runpack.total + 255) / 256));
    RCM_LAUNCH(k_halo_copy, gr, b, st, runpack, 1);
    H.launches++;
  }
  return 0;
}

// sides: 0 = west, 1 = east, 2 = south, 3 = north
int Halo::exchange(const std::vector<HaloField>& fl, int width, cudaStream_t st) {
  if (nproc <= 1) return 0;
  if (!comm) return -5;
  const int nb[4] = {nb_w, nb_e, nb_s, nb_n};
  const int w = width;
  // ---- phase 1: west/east, own rows only ----
  {
    BoxList sp{}, ru{};
    long long sc[4] = {0, 0, 0, 0}, rcnt[4] = {0, 0, 0, 0};
    for (const HaloField& F : fl) {
      const int j1 = F.dot ? g.jde1 : g.jce1, j2 = F.dot ? g.jde2 : g.jce2;
      const int i1 = F.dot ? g.ide1 : g.ice1, i2 = F.dot ? g.ide2 : g.ice2;
      const int ni = i2 - i1 + 1;
      for (int s = 0; s < 2; ++s) {
        if (nb[s] < 0) continue;
        Box B{}; B.p = F.p; B.si = g.si; B.sk = g.sk; B.nj = w; B.ni = ni; B.nk = F.nk; B.ia = i1;
        B.ja = (s == 0) ? j1 : j2 - w + 1; B.buf = sbuf[s]; B.off = sc[s];
        sc[s] += (long long)w * ni * F.nk; sp.b[sp.n++] = B;
        Box R = B; R.ja = (s == 0) ? j1 - w : j2 + 1; R.buf = rbuf[s]; R.off = rcnt[s];
        rcnt[s] += (long long)w * ni * F.nk; ru.b[ru.n++] = R;
      }
    }
    for (int q = 0; q < sp.n; ++q) sp.total += (long long)sp.b[q].nj * sp.b[q].ni * sp.b[q].nk;
    ru.total = sp.total;
    if ((size_t)std::max(std::max(sc[0], sc[1]), std::max(rcnt[0], rcnt[1])) > cap) return -5;
    if (run_phase(*this, sp, ru, sc, rcnt, nb, 0, st)) return -5;
  }
  // ---- phase 2: south/north, rows extended by the ghost columns just filled ----
  {
    BoxList sp{}, ru{};
    long long sc[4] = {0, 0, 0, 0}, rcnt[4] = {0, 0, 0, 0};
    for (const HaloField& F : fl) {
      const int j1 = (F.dot ? g.jde1 : g.jce1) - (nb_w >= 0 ? w : 0), j2 = (F.dot ? g.jde2 : g.jce2) + (nb_e >= 0 ? w : 0);
      const int i1 = F.dot ? g.ide1 : g.ice1, i2 = F.dot ? g.ide2 : g.ice2;
      const int nj = j2 - j1 + 1;
      for (int s = 2; s < 4; ++s) {
        if (nb[s] < 0) continue;
        Box B{}; B.p = F.p; B.si = g.si; B.sk = g.sk; B.nj = nj; B.ni = w; B.nk = F.nk; B.ja = j1;
        B.ia = (s == 2) ? i1 : i2 - w + 1; B.buf = sbuf[s]; B.off = sc[s];
        sc[s] += (long long)w * nj * F.nk; sp.b[sp.n++] = B;
        Box R = B; R.ia = (s == 2) ? i1 - w : i2 + 1; R.buf = rbuf[s]; R.off = rcnt[s];
        rcnt[s] += (long long)w * nj * F.nk; ru.b[ru.n++] = R;
      }
    }
    for (int q = 0; q < sp.n; ++q) sp.total += (long long)sp.b[q].nj * sp.b[q].ni * sp.b[q].nk;
    ru.total = sp.total;
    if ((size_t)std::max(std::max(sc[2], sc[3]), std::max(rcnt[2], rcnt[3])) > cap) return -5;
    if (run_phase(*this, sp, ru, sc, rcnt, nb, 2, st)) return -5;
  }
  return 0;
}

// width-6 halo of the 2-D estore array (replaces grid_collect + bcast of
// estore_g, Main/mod_sound.F90:494-495: only +-6 neighbours are ever read)
int Halo::exchange_estore(const Fields& f, cudaStream_t st) {
  if (nproc <= 1) return 0;
  if (!comm) return -5;
  const int nb[4] = {nb_w, nb_e, nb_s, nb_n};
  const int w = RCM_EHALO;
  // estore is addressed as p[j + esi*i]; present it to the generic copier as a
  // 1-plane field whose "k = 1" plane starts at p (sk = 0 trick: kk+1 = 1 -> + sk)
  for (int phase = 0; phase < 2; ++phase) {
    BoxList sp{}, ru{};
    long long sc[4] = {0, 0, 0, 0}, rcnt[4] = {0, 0, 0, 0};
    const int j1o = g.jci1, j2o = g.jci2, i1o = g.ici1, i2o = g.ici2;
    for (int s = 2 * phase; s < 2 * phase + 2; ++s) {
      if (nb[s] < 0) continue;
      Box B{}; B.p = f.estore; B.si = g.esi; B.sk = 0; B.nk = 1;
      Box R{};
      if (phase == 0) {
        B.nj = w; B.ni = i2o - i1o + 1; B.ia = i1o; B.ja = (s == 0) ? j1o : j2o - w + 1;
        R = B; R.ja = (s == 0) ? j1o - w : j2o + 1;
      } else {
        const int j1 = j1o - (nb_w >= 0 ? w : 0), j2 = j2o + (nb_e >= 0 ? w : 0);
        B.nj = j2 - j1 + 1; B.ni = w; B.ja = j1; B.ia = (s == 2) ? i1o : i2o - w + 1;
        R = B; R.ia = (s == 2) ? i1o - w : i2o + 1;
      }
      B.buf = sbuf[s]; B.off = 0; R.buf = rbuf[s]; R.off = 0;
      sc[s] = rcnt[s] = (long long)B.nj * B.ni;
      sp.b[sp.n++] = B; ru.b[ru.n++] = R;
    }
    for (int q = 0; q < sp.n; ++q) sp.total += (long long)sp.b[q].nj * sp.b[q].ni * sp.b[q].nk;
    ru.total = sp.total;
    if (run_phase(*this, sp, ru, sc, rcnt, nb, 2 * phase, st)) return -5;
  }
  return 0;
}
