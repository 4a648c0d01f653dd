// Minimal TMA 3-D tile load test (tuning aid): nvcc -O2 -gencode arch=compute_100a,code=sm_100a -o tma_test tma_test.cu
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <dlfcn.h>
#include <vector>

constexpr int TWMAX = 36, TH = 8;
__constant__ int dTW;
static int TW = 36;
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <int MODE>
__global__ void k(const __grid_constant__ CUtensorMap pmap, const CUtensorMap* gmap, double* out, int c0, int c1, int c2, int TW, int dims) {
  const CUtensorMap* mp = gmap ? gmap : &pmap;
  extern __shared__ __align__(128) unsigned char smraw[];
  double* tile = reinterpret_cast<double*>(smraw);
  uint64_t* bar = reinterpret_cast<uint64_t*>(smraw + TWMAX * TH * 8);
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(1));
    if (MODE >= 1) asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    if (MODE >= 2) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(TW * TH * 8) : "memory");
    if (dims == 1)
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(tile)),
                   "l"(out + 4096), "r"(TW * TH * 8), "r"(smem_u32(bar))
                   : "memory");
    else if (dims == 0) {
      // no copy at all: complete the phase by hand
      asm volatile("mbarrier.complete_tx.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(TW * TH * 8) : "memory");
    } else if (dims == 3)
      asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(
                       smem_u32(tile)),
                   "l"(mp), "r"(c0), "r"(c1), "r"(c2), "r"(smem_u32(bar))
                   : "memory");
    else
      asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
                       smem_u32(tile)),
                   "l"(mp), "r"(c0), "r"(c1), "r"(smem_u32(bar))
                   : "memory");
  }
  asm volatile(
      "{\n.reg .pred p;\nW_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D_%=;\nbra W_%=;\nD_%=:\n}\n" ::"r"(smem_u32(bar)),
      "r"(0)
      : "memory");
  for (int t = threadIdx.x; t < TW * TH; t += blockDim.x) out[t] = tile[t];
}

typedef CUresult (*fn_encode)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                              const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

template <int MODE>
int run(fn_encode enc, double* d, double* dout, int si, int ni, int nk, const std::vector<double>& h, int variant) {
  CUtensorMap m;
  TW = (variant == 2) ? 32 : 36;
  const bool f32 = (variant == 1);
  const int dims = (variant == 3 || variant == 7) ? 2 : (variant == 4) ? 1 : (variant == 5) ? 0 : 3;
  if (variant == 7) TW = 32;
  const cuuint64_t gdim[3] = {(cuuint64_t)si * (f32 ? 2 : 1), (cuuint64_t)ni, (cuuint64_t)nk};
  const cuuint64_t gstr[2] = {(cuuint64_t)si * 8, (cuuint64_t)si * ni * 8};
  const cuuint32_t box[3] = {(cuuint32_t)TW * (f32 ? 2 : 1), TH, 1}, estr[3] = {1, 1, 1};
  CUresult r = enc(&m, f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_FLOAT64, dims, d, gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (variant == 7) {  // bytes, rank 2
    const cuuint64_t gd[2] = {(cuuint64_t)si * 8, (cuuint64_t)ni * nk};
    const cuuint64_t gs[1] = {(cuuint64_t)si * 8};
    const cuuint32_t bx[2] = {256, 8}, es[2] = {1, 1};
    r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, d, gd, gs, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
            CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  }
  printf("mode %d encode rc=%d\n", MODE, (int)r);
  { const unsigned long long* w = (const unsigned long long*)&m; for (int q = 0; q < 16; ++q) printf("%016llx%c", w[q], q % 4 == 3 ? '\n' : ' '); printf("d=%p\n", (void*)d); }
  if (r) return 1;
  const int smem = TWMAX * TH * 8 + 64;
  cudaFuncSetAttribute(k<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  const int c0 = 5, c1 = 3, c2 = (dims == 3) ? 2 : 0;
  CUtensorMap* gm = nullptr;
  if (variant == 6) { cudaMalloc(&gm, sizeof(CUtensorMap)); cudaMemcpy(gm, &m, sizeof(CUtensorMap), cudaMemcpyHostToDevice); }
  printf("sizeof map %zu align %zu host addr %p\n", sizeof(CUtensorMap), alignof(CUtensorMap), (void*)&m);
  k<MODE><<<1, 128, smem>>>(m, gm, dout, f32 ? 2 * c0 : c0, c1, c2, TW, dims);
  cudaError_t e = cudaDeviceSynchronize();
  printf("mode %d kernel: %s\n", MODE, cudaGetErrorString(e));
  if (e != cudaSuccess) return 1;
  std::vector<double> o(TW * TH);
  cudaMemcpy(o.data(), dout, o.size() * 8, cudaMemcpyDeviceToHost);
  int bad = 0;
  for (int ii = 0; ii < TH; ++ii)
    for (int jj = 0; jj < TW; ++jj)
      if (o[ii * TW + jj] != h[(size_t)(c0 + jj) + (size_t)si * (c1 + ii) + (size_t)si * ni * c2]) bad++;
  printf("mode %d mismatches %d\n", MODE, bad);
  return bad;
}

int main(int argc, char** argv) {
  const int variant = argc > 1 ? atoi(argv[1]) : 0;
  const int si = 64, ni = 40, nk = 5;
  std::vector<double> h((size_t)si * ni * nk);
  for (size_t q = 0; q < h.size(); ++q) h[q] = (double)q;
  double *d, *dout;
  cudaMalloc(&d, h.size() * 8); cudaMalloc(&dout, 65536 * 8); cudaMemset(dout, 0, 65536 * 8);
  cudaMemcpy(d, h.data(), h.size() * 8, cudaMemcpyHostToDevice);
  void* sym = nullptr; cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &q);
  printf("entry %p q=%d\n", sym, (int)q);
  if (argc > 2) {
    void* lib = dlopen("libcuda.so.1", RTLD_NOW);
    sym = lib ? dlsym(lib, "cuTensorMapEncodeTiled") : nullptr;
    printf("dlsym entry %p\n", sym);
  }
  if (!sym) return 1;
  printf("variant %d\n", variant);
  run<2>((fn_encode)sym, d, dout, si, ni, nk, h, variant);
  return 0;
}
