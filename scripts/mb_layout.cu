// Micro-benchmark: HBM throughput of a column-marching kernel that streams NR read
// fields and NW written fields, for the (j,i,k) plane-major layout against the
// (j,k,i) row-major-in-k layout and a pointwise 3-D grid.  Tuning aid only.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o mb_layout mb_layout.cu
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

#define MAXF 40
struct Ptrs { double* p[MAXF]; };

template <int NR, int NW>
__global__ void __launch_bounds__(256) march(Ptrs f, int nj, int ni, int nk, long long sI, long long sK) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y * blockDim.y + threadIdx.y;
  if (j >= nj || i >= ni) return;
  const long long base = j + sI * i;
  double carry = 0.0;
  for (int k = 0; k < nk; ++k) {
    const long long o = base + sK * k;
    double v[NR];
#pragma unroll
    for (int r = 0; r < NR; ++r) v[r] = __ldg(f.p[r] + o);
    double s = carry;
#pragma unroll
    for (int r = 0; r < NR; ++r) s += v[r];
    carry = s * 0.5;
#pragma unroll
    for (int w = 0; w < NW; ++w) f.p[NR + w][o] = s + (double)w;
  }
}

template <int NR, int NW>
__global__ void __launch_bounds__(256) point(Ptrs f, int nj, int ni, int nk, long long sI, long long sK) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y * blockDim.y + threadIdx.y, k = blockIdx.z;
  if (j >= nj || i >= ni) return;
  const long long o = j + sI * i + sK * k;
  double s = 0.0;
#pragma unroll
  for (int r = 0; r < NR; ++r) s += __ldg(f.p[r] + o);
#pragma unroll
  for (int w = 0; w < NW; ++w) f.p[NR + w][o] = s + (double)w;
}

template <int NR, int NW>
void run(Ptrs f, int nj, int ni, int nk, long long pj, int bx, int by) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  const double bytes = 8.0 * (NR + NW) * nj * (double)ni * nk;
  for (int mode = 0; mode < 3; ++mode) {
    long long sI, sK;
    if (mode == 1) { sK = pj; sI = pj * nk; } else { sI = pj; sK = pj * (long long)(ni + 4); }
    dim3 b(bx, by), g((nj + bx - 1) / bx, (ni + by - 1) / by, mode == 2 ? nk : 1);
    float best = 1e30f;
    for (int it = 0; it < 4; ++it) {
      cudaEventRecord(e0);
      if (mode == 2) point<NR, NW><<<g, b>>>(f, nj, ni, nk, sI, sK);
      else march<NR, NW><<<g, b>>>(f, nj, ni, nk, sI, sK);
      cudaEventRecord(e1);
      cudaEventSynchronize(e1);
      float ms; cudaEventElapsedTime(&ms, e0, e1);
      if (it && ms < best) best = ms;
    }
    printf("NR=%2d NW=%2d block=%dx%d %-12s %8.3f ms %8.1f GB/s\n", NR, NW, bx, by,
           mode == 0 ? "march(j,i,k)" : mode == 1 ? "march(j,k,i)" : "point(j,i,k)", best, bytes / best * 1e-6);
  }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) printf("CUDA error %s\n", cudaGetErrorString(e));
}

int main(int argc, char** argv) {
  const int nj = 1000, ni = 1000, nk = 42;
  const long long pj = 1008;
  const size_t n = (size_t)pj * (ni + 4) * nk;
  Ptrs f;
  for (int a = 0; a < MAXF; ++a) { cudaMalloc(&f.p[a], n * 8); cudaMemset(f.p[a], 0, n * 8); }
  int bx = argc > 1 ? atoi(argv[1]) : 32, by = argc > 2 ? atoi(argv[2]) : 4;
  run<4, 2>(f, nj, ni, nk, pj, bx, by);
  run<8, 2>(f, nj, ni, nk, pj, bx, by);
  run<16, 6>(f, nj, ni, nk, pj, bx, by);
  run<24, 8>(f, nj, ni, nk, pj, bx, by);
  run<30, 8>(f, nj, ni, nk, pj, bx, by);
  return 0;
}
