// Bounded retry of tensor-map TMA: ONE kernel, descriptor passed as `const __grid_constant__`
// and addressed directly (no runtime-selected pointer), entry point through cudaGetDriverEntryPointByVersion, a
// rank-2 FLOAT64 tensor with a 32 x 8 box (inner extent 256 B).
//   nvcc -O2 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o tma_tensor_retry tma_tensor_retry.cu
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <vector>

constexpr int BW = 32, BH = 8;
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void k_tile(const __grid_constant__ CUtensorMap tmap, double* out, int c0, int c1) {
  __shared__ __align__(128) double tile[BW * BH];
  __shared__ __align__(8) uint64_t bar;
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&bar)), "r"(1));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar)), "r"(BW * BH * 8) : "memory");
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 ::"r"(smem_u32(tile)), "l"(reinterpret_cast<uint64_t>(&tmap)), "r"(c0), "r"(c1), "r"(smem_u32(&bar))
                 : "memory");
  }
  asm volatile("{\n.reg .pred p;\nW_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D_%=;\nbra W_%=;\nD_%=:\n}\n"
               ::"r"(smem_u32(&bar)), "r"(0) : "memory");
  for (int t = threadIdx.x; t < BW * BH; t += blockDim.x) out[t] = tile[t];
}

typedef CUresult (*fn_encode)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                              const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main() {
  const int si = 64, ni = 40;
  std::vector<double> h((size_t)si * ni);
  for (size_t q = 0; q < h.size(); ++q) h[q] = (double)q;
  double *d = nullptr, *dout = nullptr;
  cudaMalloc(&d, h.size() * 8); cudaMalloc(&dout, BW * BH * 8); cudaMemset(dout, 0, BW * BH * 8);
  cudaMemcpy(d, h.data(), h.size() * 8, cudaMemcpyHostToDevice);
  void* sym = nullptr; cudaDriverEntryPointQueryResult q;
  cudaError_t e0 = cudaGetDriverEntryPointByVersion("cuTensorMapEncodeTiled", &sym, 12000, cudaEnableDefault, &q);
  int drv = 0, rt = 0; cudaDriverGetVersion(&drv); cudaRuntimeGetVersion(&rt);
  printf("driver %d runtime %d entry %p query %d rc %d\n", drv, rt, sym, (int)q, (int)e0);
  if (!sym) return 1;
  alignas(64) CUtensorMap m;
  const cuuint64_t gdim[2] = {(cuuint64_t)si, (cuuint64_t)ni};
  const cuuint64_t gstr[1] = {(cuuint64_t)si * 8};
  const cuuint32_t box[2] = {BW, BH}, estr[2] = {1, 1};
  CUresult r = ((fn_encode)sym)(&m, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, d, gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  printf("encode rc=%d\n", (int)r);
  if (r) return 1;
  const int c0 = 5, c1 = 3;
  k_tile<<<1, 128>>>(m, dout, c0, c1);
  cudaError_t e = cudaDeviceSynchronize();
  printf("kernel: %s\n", cudaGetErrorString(e));
  if (e != cudaSuccess) return 2;
  std::vector<double> o(BW * BH);
  cudaMemcpy(o.data(), dout, o.size() * 8, cudaMemcpyDeviceToHost);
  int bad = 0;
  for (int ii = 0; ii < BH; ++ii)
    for (int jj = 0; jj < BW; ++jj)
      if (o[ii * BW + jj] != h[(size_t)(c0 + jj) + (size_t)si * (c1 + ii)]) bad++;
  printf("mismatches %d of %d\n", bad, BW * BH);
  return bad ? 3 : 0;
}
