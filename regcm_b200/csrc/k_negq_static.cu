// rcm_dyn_negq_repair_static — the negative-moisture repair (Main/mod_tendency.F90:381-393, see k_negq.cu) as a
// stand-alone operator that ALLOCATES NOTHING, for callers that cannot call cudaMalloc (code under CUDA graph
// capture, tight serving loops, harnesses that forbid it).  No counterpart in the
// reference.
//
// The regular path (rcm_dyn_step -> k_tq / k_h_tq -> k_negq_repair) keeps its scratch — the list of negative
// points of every (species, level) plane and the fill counts — in arrays that rcm_dyn_create obtained with
// cudaMalloc.  Here the same scratch lives in ONE file-scope `static __device__` array of a fixed, modest size:
//
//     RCM_NQS_MAX_CELLS  = 1 048 576 list entries (4 MiB)   -> nj * ni * nk  must not exceed it
//     RCM_NQS_MAX_PLANES = 1 024 counters         (4 KiB)   -> nk            must not exceed it
//
// Every GPU that loads libregcm_dyn.so holds these 4 MiB + 4 KiB until the process exits.  All calls share the
// array and it keeps the previous call's contents: the counters are cleared at the head of every call (the
// list entries beyond a counter are never read).  Two calls that overlap on the GPU (different streams, or
// captured graphs launched concurrently) RACE on the scratch — serialise them.
//
// In host code the name of a __device__ array is a host-side stand-in, not a device address.  The device
// address is taken once per device, with that device current, by cudaGetSymbolAddress (return code checked)
// and handed to the kernels as an ordinary pointer.
#include <cuda_runtime.h>
#include <stdint.h>

#include <mutex>

#include "../../include/regcm_dyn.h"

namespace {

constexpr int kMaxCells = RCM_NQS_MAX_CELLS;
constexpr int kMaxPlanes = RCM_NQS_MAX_PLANES;
constexpr int kMaxDevices = 64;

// [0, kMaxCells): negative-point lists, plane p owns [p*nj*ni, (p+1)*nj*ni); [kMaxCells, +kMaxPlanes): fill counts
static __device__ __align__(256) int g_nqs_scratch[kMaxCells + kMaxPlanes];

std::mutex g_mu;
int* g_scratch_addr[kMaxDevices] = {nullptr};   // device address of g_nqs_scratch on each device, taken once

// q_out = q_in; append every negative INTERIOR point to the list of its plane
__global__ void __launch_bounds__(256) k_nqs_scan(const double* __restrict__ qin, double* __restrict__ qout, int* list, int* cnt,
                                                  const int nj, const int ni, const int nk) {
  const long long n2 = (long long)nj * ni, n = n2 * nk;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (long long)gridDim.x * blockDim.x) {
    const double q = qin[t];
    qout[t] = q;
    if (q < 0.0) {
      const int k = (int)(t / n2), off = (int)(t - (long long)k * n2);
      const int j = off % nj, i = off / nj;
      if (j >= 1 && j <= nj - 2 && i >= 1 && i <= ni - 2) {
        const int slot = atomicAdd(cnt + k, 1);
        list[(long long)k * n2 + slot] = off;
      }
    }
  }
}

// One block per plane.  Same rule as nq_try in k_negq.cu: a listed point is repaired once its four predecessors
// in the reference's sweep order — (j-1,i), (j-1,i-1), (j,i-1), (j+1,i-1) — are final (not negative, or not
// swept), reading the already repaired values of those and the unrepaired values of its successors.
__global__ void __launch_bounds__(256) k_nqs_repair(double* q, const int* list, const int* cnt, unsigned long long* nrep,
                                                    const int nj, const int ni) {
  const int k = blockIdx.x, c = cnt[k];
  if (c == 0) return;
  const long long n2 = (long long)nj * ni;
  double* p = q + (long long)k * n2;
  const int* l = list + (long long)k * n2;
  auto val = [&](int j, int i) { return *(const volatile double*)(p + j + (long long)i * nj); };
  auto pending = [&](int j, int i) { return j >= 1 && j <= nj - 2 && i >= 1 && i <= ni - 2 && val(j, i) < 0.0; };
  for (;;) {
    int wait = 0;
    for (int e = threadIdx.x; e < c; e += blockDim.x) {
      const int off = l[e], j = off % nj, i = off / nj;
      if (!(val(j, i) < 0.0)) continue;
      if (pending(j - 1, i) || pending(j - 1, i - 1) || pending(j, i - 1) || pending(j + 1, i - 1)) { wait = 1; continue; }
      double sum = 0.0;   // sum(abs(section)): array-element order, j fastest
      for (int ii = i - 1; ii <= i + 1; ++ii)
        for (int jj = j - 1; jj <= j + 1; ++jj) sum += fabs(val(jj, ii));
      p[j + (long long)i * nj] = 0.01 * sum / 9.0;
    }
    __threadfence_block();
    if (!__syncthreads_or(wait)) break;
  }
  if (nrep && threadIdx.x == 0) atomicAdd(nrep, (unsigned long long)c);
}

}  // namespace

extern "C" int rcm_dyn_negq_repair_static(const double* q_in, double* q_out, int32_t nj, int32_t ni, int32_t nk,
                                          unsigned long long* n_repaired, void* stream) {
  if (!q_in || !q_out || q_in == q_out) return RCM_ERR_ARG;
  if (nj < 3 || ni < 3 || nk < 1) return RCM_ERR_ARG;
  if (nk > kMaxPlanes || (long long)nj * ni * nk > (long long)kMaxCells) return RCM_ERR_ARG;  // beyond the static scratch
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) return RCM_ERR_CUDA;
  int* scratch = nullptr;
  {
    std::lock_guard<std::mutex> lk(g_mu);
    if (!g_scratch_addr[dev]) {   // set-up, once per device, with that device current
      void* a = nullptr;
      if (cudaGetSymbolAddress(&a, g_nqs_scratch) != cudaSuccess || !a) { cudaGetLastError(); return RCM_ERR_CUDA; }
      g_scratch_addr[dev] = (int*)a;
    }
    scratch = g_scratch_addr[dev];
  }
  cudaStream_t st = (cudaStream_t)stream;
  int* list = scratch;
  int* cnt = scratch + kMaxCells;
  // the scratch keeps the last call's contents: the counters (and the caller's total) must start at zero
  if (cudaMemsetAsync(cnt, 0, (size_t)nk * sizeof(int), st) != cudaSuccess) return RCM_ERR_CUDA;
  if (n_repaired && cudaMemsetAsync(n_repaired, 0, sizeof(unsigned long long), st) != cudaSuccess) return RCM_ERR_CUDA;
  const long long n = (long long)nj * ni * nk;
  const unsigned blocks = (unsigned)((n + 255) / 256 < 1184 ? (n + 255) / 256 : 1184);   // 148 SMs x 8 blocks
  k_nqs_scan<<<blocks, 256, 0, st>>>(q_in, q_out, list, cnt, nj, ni, nk);
  k_nqs_repair<<<(unsigned)nk, 256, 0, st>>>(q_out, list, cnt, n_repaired, nj, ni);
  return cudaGetLastError() == cudaSuccess ? RCM_OK : RCM_ERR_CUDA;
}
