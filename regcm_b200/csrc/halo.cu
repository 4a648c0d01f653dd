// Halo exchange over NCCL send/recv (see halo.h).  NCCL is bound at run time
// with dlopen so that single-GPU use needs no NCCL at all; with torch in the
// process the already-loaded libnccl.so.2 is reused.
#include "halo.h"

#include <dlfcn.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
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
typedef int (*fn_AllGather)(const void*, void*, size_t, int, void*, cudaStream_t);
typedef const char* (*fn_ErrStr)(int);
struct Nccl {
  void* lib = nullptr;
  fn_GetUniqueId GetUniqueId = nullptr;
  fn_CommInitRank CommInitRank = nullptr;
  fn_CommDestroy CommDestroy = nullptr;
  fn_Group GroupStart = nullptr, GroupEnd = nullptr;
  fn_SendRecv Send = nullptr, Recv = nullptr;
  fn_AllReduce AllReduce = nullptr;
  fn_AllGather AllGather = nullptr;
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
    AllGather = (fn_AllGather)dlsym(lib, "ncclAllGather");
    ErrStr = (fn_ErrStr)dlsym(lib, "ncclGetErrorString");
    return GetUniqueId && CommInitRank && CommDestroy && GroupStart && GroupEnd && Send && Recv && AllReduce;
  }
};
Nccl g_nccl;
const int kNcclDouble = 8, kNcclChar = 0, kNcclSum = 0, kNcclMax = 2;
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
  if (cudaMalloc(&red, 64 * sizeof(double)) != cudaSuccess) { err = "halo buffer allocation failed"; return -5; }
#ifndef RCM_EMU
  if (setup_p2p(err)) return -5;
#endif
  return 0;
}

void Halo::destroy() {
  for (int q = 0; q < 4; ++q) { if (sbuf[q]) cudaFree(sbuf[q]); if (rbuf[q]) cudaFree(rbuf[q]); sbuf[q] = rbuf[q] = nullptr; }
  if (red) cudaFree(red);
  red = nullptr;
#ifndef RCM_EMU
  for (int d = 0; d < 8; ++d) {
    bool first = peer_block[d] != nullptr;
    for (int q = 0; q < d; ++q) if (peer_block[q] == peer_block[d]) first = false;
    if (first) cudaIpcCloseMemHandle(peer_block[d]);
    peer_block[d] = nullptr;
  }
  if (pblock) cudaFree(pblock);
  if (errflag) cudaFree(errflag);
  pblock = nullptr; errflag = nullptr; p2p = false;
  if (comm && g_nccl.CommDestroy) g_nccl.CommDestroy(comm);
#endif
  comm = nullptr;
}

int Halo::take_error() {
#ifdef RCM_EMU
  return 0;
#else
  if (!p2p || !errflag) return 0;
  int e = 0;
  if (cudaMemcpy(&e, errflag, sizeof(int), cudaMemcpyDeviceToHost) != cudaSuccess) return 1;
  if (e) cudaMemset(errflag, 0, sizeof(int));
  return e;
#endif
}

int Halo::allreduce_host(double* v, int n, int op) {
  if (nproc <= 1) return 0;
  if (!comm || n > 64) return -5;
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


#ifndef RCM_EMU
// ===========================================================================
// NVLink peer-memory transport (see halo.h).  Directions: 0 W, 1 E, 2 S, 3 N,
// 4 SW, 5 SE, 6 NW, 7 NE; the opposite of d is kOpp[d].
namespace {
const int kOpp[8] = {1, 0, 3, 2, 7, 6, 5, 4};
const int kDj[8] = {-1, 1, 0, 0, -1, 1, -1, 1};
const int kDi[8] = {0, 0, -1, 1, -1, -1, 1, 1};

struct XField {
  double* p;     // element (j,i,kk) = p[j + si*i + sk*kk], kk from 0
  long long sk;
  int si, nk, grid;  // grid: 0 cross, 1 dot, 2 estore
};
struct XList {
  XField f[16];
  int nf;
  int s_ja[8][3], s_ia[8][3], r_ja[8][3], r_ia[8][3], nj[8][3], ni[8][3];
  double* sendbuf[8];                 // neighbour's receive buffer for what I send towards d
  double* recvbuf[8];                 // my receive buffer for what arrives from d
  unsigned long long* flag_peer[8];   // neighbour's arrival flag for my message
  unsigned long long* flag_mine;      // my 8 arrival flags
  unsigned long long seq;
  unsigned int* done;                 // block counter of the pack kernel
  int* err;
  int active[8];
};

__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}

// offset (doubles) of field fi inside the message of direction d
__device__ __forceinline__ long long x_offset(const XList& L, int d, int fi) {
  long long off = 0;
  for (int q = 0; q < fi; ++q) {
    const int gq = L.f[q].grid;
    off += (long long)L.nj[d][gq] * L.ni[d][gq] * L.f[q].nk;
  }
  return off;
}

// pack: own edge cells -> the neighbour's receive buffer (stores over NVLink);
// the last block to finish raises the neighbours' arrival flags.
__global__ void __launch_bounds__(256) k_x_pack(const XList L) {
  const int d = blockIdx.z, fi = blockIdx.y;
  if (L.active[d]) {
    const XField F = L.f[fi];
    const int nj = L.nj[d][F.grid], ni = L.ni[d][F.grid];
    const long long n = (long long)nj * ni * F.nk;
    const long long off = x_offset(L, d, fi);
    const int ja = L.s_ja[d][F.grid], ia = L.s_ia[d][F.grid];
    double* dst = L.sendbuf[d] + off;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (long long)gridDim.x * blockDim.x) {
      const int jj = (int)(t % nj);
      const long long r2 = t / nj;
      const int ii = (int)(r2 % ni), kk = (int)(r2 / ni);
      dst[t] = F.p[(long long)(ja + jj) + (long long)(ia + ii) * F.si + (long long)kk * F.sk];
    }
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int total = gridDim.x * gridDim.y * gridDim.z;
    if (atomicAdd(L.done, 1u) == total - 1) {
      *L.done = 0;
      __threadfence_system();
      for (int q = 0; q < 8; ++q)
        if (L.active[q]) st_release_sys(L.flag_peer[q], L.seq);
    }
  }
}

// unpack: wait for the message of this block's direction, then copy it into the ghost cells
__global__ void __launch_bounds__(256) k_x_unpack(const XList L) {
  const int d = blockIdx.z, fi = blockIdx.y;
  if (!L.active[d]) return;
  __shared__ int arrived;
  if (threadIdx.x == 0) {
    arrived = 1;
    const long long t0 = clock64();
    while (ld_acquire_sys(L.flag_mine + d) < L.seq) {
      if (clock64() - t0 > 20000000000LL) { atomicExch(L.err, 1); arrived = 0; break; }  // ~10 s: report instead of hanging
    }
  }
  __syncthreads();
  if (!arrived) return;  // never scatter a stale receive buffer into the ghost cells; rcm_dyn_step returns RCM_ERR_NCCL
  const XField F = L.f[fi];
  const int nj = L.nj[d][F.grid], ni = L.ni[d][F.grid];
  const long long n = (long long)nj * ni * F.nk;
  const long long off = x_offset(L, d, fi);
  const int ja = L.r_ja[d][F.grid], ia = L.r_ia[d][F.grid];
  const double* src = L.recvbuf[d] + off;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (long long)gridDim.x * blockDim.x) {
    const int jj = (int)(t % nj);
    const long long r2 = t / nj;
    const int ii = (int)(r2 % ni), kk = (int)(r2 / ni);
    F.p[(long long)(ja + jj) + (long long)(ia + ii) * F.si + (long long)kk * F.sk] = __ldcv(src + t);
  }
}
const size_t kFlagBytes = 256;
}  // namespace

int Halo::setup_p2p(std::string& err) {
  p2p = false;
  if (const char* e = std::getenv("RCM_HALO")) { if (std::strcmp(e, "nccl") == 0) return 0; }
  if (!g_nccl.AllGather) return 0;
  // neighbours in the 8 directions
  for (int d = 0; d < 8; ++d) {
    const int qj = pj + kDj[d], qi = pi + kDi[d];
    nb8[d] = (qj >= 0 && qj < cj && qi >= 0 && qi < ci) ? qj + cj * qi : -1;
  }
  struct Card { char bus[32]; cudaIpcMemHandle_t h; int ok; };
  Card mine{};
  int dev = 0;
  cudaGetDevice(&dev);
  if (cudaDeviceGetPCIBusId(mine.bus, sizeof(mine.bus), dev) != cudaSuccess) mine.bus[0] = 0;
  pcap = cap;
  const size_t bytes = kFlagBytes + (size_t)16 * pcap * sizeof(double);
  mine.ok = (cudaMalloc(&pblock, bytes) == cudaSuccess) && (cudaMemset(pblock, 0, bytes) == cudaSuccess) &&
            (cudaMalloc(&errflag, 2 * sizeof(int)) == cudaSuccess) && (cudaMemset(errflag, 0, 2 * sizeof(int)) == cudaSuccess) &&
            (cudaIpcGetMemHandle(&mine.h, pblock) == cudaSuccess);
  cudaDeviceSynchronize();
  // gather every rank's card
  Card* dcards = nullptr;
  std::vector<Card> cards(nproc);
  if (cudaMalloc(&dcards, sizeof(Card) * (nproc + 1)) != cudaSuccess) { err = "p2p: allocation failed"; return -5; }
  cudaMemcpy(dcards + nproc, &mine, sizeof(Card), cudaMemcpyHostToDevice);
  if (g_nccl.AllGather(dcards + nproc, dcards, sizeof(Card), kNcclChar, comm, (cudaStream_t)0) != 0 ||
      cudaStreamSynchronize((cudaStream_t)0) != cudaSuccess) { cudaFree(dcards); err = "p2p: allgather failed"; return -5; }
  cudaMemcpy(cards.data(), dcards, sizeof(Card) * nproc, cudaMemcpyDeviceToHost);
  cudaFree(dcards);
  bool good = true;
  for (int r = 0; r < nproc; ++r) {
    if (!cards[r].ok || cards[r].bus[0] == 0) good = false;
    for (int q = 0; q < r; ++q)
      if (std::strncmp(cards[r].bus, cards[q].bus, sizeof(mine.bus)) == 0) good = false;  // two ranks on one GPU: no spin-waits
  }
  double okv = good ? 1.0 : 0.0;
  if (good) {
    for (int d = 0; d < 8 && okv > 0; ++d) {
      if (nb8[d] < 0) continue;
      // the same neighbour can appear in several directions only on degenerate layouts; map each once
      bool mapped = false;
      for (int q = 0; q < d; ++q) if (nb8[q] == nb8[d]) { peer_block[d] = peer_block[q]; mapped = true; }
      if (mapped) continue;
      void* pp = nullptr;
      if (cudaIpcOpenMemHandle(&pp, cards[nb8[d]].h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { okv = 0.0; cudaGetLastError(); break; }
      peer_block[d] = (char*)pp;
    }
  }
  double vmin = -okv;  // min over ranks via max of the negative
  if (allreduce_host(&vmin, 1, 1)) { err = "p2p: allreduce failed"; return -5; }
  p2p = (vmin == -1.0);
  if (!p2p) {
    for (int d = 0; d < 8; ++d) {
      bool first = peer_block[d] != nullptr;
      for (int q = 0; q < d; ++q) if (peer_block[q] == peer_block[d]) first = false;
      if (first) cudaIpcCloseMemHandle(peer_block[d]);
      peer_block[d] = nullptr;
    }
  }
  return 0;
}

int Halo::exchange_p2p(const std::vector<HaloField>& fl, int width, int estore, const Fields* ff, cudaStream_t st) {
  XList L{};
  const int w = width;
  if (estore) {
    L.nf = 1;
    L.f[0].p = ff->estore; L.f[0].si = g.esi; L.f[0].sk = 0; L.f[0].nk = 1; L.f[0].grid = 2;
  } else {
    if (fl.size() > 16) return -5;
    L.nf = (int)fl.size();
    for (int q = 0; q < L.nf; ++q) {
      L.f[q].p = fl[q].p + g.sk; L.f[q].si = g.si; L.f[q].sk = g.sk; L.f[q].nk = fl[q].nk; L.f[q].grid = fl[q].dot ? 1 : 0;
    }
  }
  const int J1[3] = {g.jce1, g.jde1, g.jci1}, J2[3] = {g.jce2, g.jde2, g.jci2};
  const int I1[3] = {g.ice1, g.ide1, g.ici1}, I2[3] = {g.ice2, g.ide2, g.ici2};
  long long maxn = 0;
  ++seq;
  const int par = (int)(seq & 1ULL);
  for (int d = 0; d < 8; ++d) {
    L.active[d] = nb8[d] >= 0;
    long long msg = 0;
    for (int t = 0; t < 3; ++t) {
      const int j1 = J1[t], j2 = J2[t], i1 = I1[t], i2 = I2[t];
      const int dj = kDj[d], di = kDi[d];
      L.nj[d][t] = dj == 0 ? j2 - j1 + 1 : w;
      L.ni[d][t] = di == 0 ? i2 - i1 + 1 : w;
      L.s_ja[d][t] = dj < 0 ? j1 : (dj > 0 ? j2 - w + 1 : j1);
      L.s_ia[d][t] = di < 0 ? i1 : (di > 0 ? i2 - w + 1 : i1);
      L.r_ja[d][t] = dj < 0 ? j1 - w : (dj > 0 ? j2 + 1 : j1);
      L.r_ia[d][t] = di < 0 ? i1 - w : (di > 0 ? i2 + 1 : i1);
    }
    if (!L.active[d]) continue;
    for (int q = 0; q < L.nf; ++q) {
      const long long n = (long long)L.nj[d][L.f[q].grid] * L.ni[d][L.f[q].grid] * L.f[q].nk;
      msg += n; maxn = std::max(maxn, n);
    }
    if ((size_t)msg > pcap) return -5;
    const int od = kOpp[d];
    L.sendbuf[d] = (double*)(peer_block[d] + kFlagBytes) + ((size_t)par * 8 + od) * pcap;
    L.recvbuf[d] = (double*)(pblock + kFlagBytes) + ((size_t)par * 8 + d) * pcap;
    L.flag_peer[d] = (unsigned long long*)peer_block[d] + od;
  }
  L.flag_mine = (unsigned long long*)pblock;
  L.seq = seq;
  L.done = (unsigned int*)(errflag + 1);
  L.err = errflag;
  if (maxn == 0) return 0;
  const unsigned gx = (unsigned)std::min<long long>((maxn + 255) / 256, 64);
  dim3 b(256), gr(gx, (unsigned)L.nf, 8);
  k_x_pack<<<gr, b, 0, st>>>(L);
  k_x_unpack<<<gr, b, 0, st>>>(L);
  launches += 2;
  return cudaGetLastError() == cudaSuccess ? 0 : -5;
}
#endif

// sides: 0 = west, 1 = east, 2 = south, 3 = north
int Halo::exchange(const std::vector<HaloField>& fl, int width, cudaStream_t st) {
  if (nproc <= 1) return 0;
  if (!comm) return -5;
  // the packed message of one exchange holds at most 16 fields (XList, BoxList, buffer capacity): longer lists — the
  // state of the non-hydrostatic core with tracers and TKE — go in several messages; fields are independent of each other
  if (fl.size() > 16) {
    for (size_t q = 0; q < fl.size(); q += 16) {
      const std::vector<HaloField> part(fl.begin() + q, fl.begin() + std::min(fl.size(), q + 16));
      if (const int rc = exchange(part, width, st)) return rc;
    }
    return 0;
  }
#ifndef RCM_EMU
  if (p2p) return exchange_p2p(fl, width, 0, nullptr, st);
#endif
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
#ifndef RCM_EMU
  if (p2p) return exchange_p2p({}, RCM_EHALO, 1, &f, st);
#endif
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
