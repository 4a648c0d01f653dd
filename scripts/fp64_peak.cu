// FP64 issue-rate microbenchmark for the second roof of the dyn-core kernels.
// Register-resident dependent chains (8 independent chains per thread), CUDA
// events, best of 5.  Prints one JSON object:
//   dfma_tinst : DFMA thread-instructions per second (x2 = FLOP/s)
//   dadd_tinst : DADD rate;  dmul_tinst : DMUL rate (what a -fmad=false build issues)
//   ddiv_tinst : IEEE double divisions per second (each ~ a dozen FP64 instructions)
// nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o fp64_peak fp64_peak.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int OP>
__global__ void __launch_bounds__(256) k(double* out, double a, double b, int iters) {
  double x[8];
#pragma unroll
  for (int q = 0; q < 8; ++q) x[q] = a + q * 1e-3 + threadIdx.x * 1e-6;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      if (OP == 0) x[q] = __fma_rn(x[q], a, b);
      else if (OP == 1) x[q] = __dadd_rn(x[q], b);
      else if (OP == 2) x[q] = __dmul_rn(x[q], a);
      else x[q] = __ddiv_rn(b, x[q] + a);
    }
  }
  double s = 0;
#pragma unroll
  for (int q = 0; q < 8; ++q) s += x[q];
  if (s == 123.456) out[0] = s;
}

template <int OP>
double rate(int iters) {
  double* d;
  cudaMalloc(&d, 8);
  cudaDeviceProp p;
  cudaGetDeviceProperties(&p, 0);
  const int blocks = p.multiProcessorCount * 8;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  double best = 0;
  for (int rep = 0; rep < 6; ++rep) {
    cudaEventRecord(e0);
    k<OP><<<blocks, 256>>>(d, 1.0000001, 1e-9, iters);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    const double r = (double)blocks * 256 * 8.0 * iters / (ms * 1e-3);
    if (rep > 0 && r > best) best = r;
  }
  cudaFree(d);
  return best;
}

int main() {
  const double fma = rate<0>(20000), add = rate<1>(20000), mul = rate<2>(20000), dv = rate<3>(2000);
  std::printf("{\"dfma_tinst\": %.4g, \"dadd_tinst\": %.4g, \"dmul_tinst\": %.4g, \"ddiv_tinst\": %.4g, \"fp64_tflops_fma\": %.3f}\n",
              fma, add, mul, dv, 2 * fma / 1e12);
  return 0;
}
