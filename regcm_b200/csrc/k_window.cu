// Window pack / unpack between the padded (j,i,k) device geometry and a CONTIGUOUS buffer laid out like the
// caller's Fortran array (n1 x n2 x nk, j fastest).  Used by the asynchronous state export and boundary staging:
// a strided cudaMemcpy3DAsync of one 3-D field costs the host ~2-3 ms to submit (tens of thousands of row
// descriptors; measured: it delayed the time loop by 13-18 ms per interval on C4), a 1-D copy of the packed window
// is one descriptor and runs at the full PCIe rate.
#include "dev_common.cuh"
#include "launch.h"

namespace {
// buf[(j-j0) + n1*((i-i0) + n2*(k-k0))]  <->  p[j + si*i + sk*k]   for j0..j0+n1-1, i0..i0+n2-1, k0..k0+nk-1
template <bool PACK>
__global__ void __launch_bounds__(256) k_window(double* __restrict__ buf, double* __restrict__ p, const int j0, const int i0, const int k0,
                                                const int n1, const int n2, const int nk, const int si, const long long sk) {
  const long long n = (long long)n1 * n2 * nk;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (long long)gridDim.x * blockDim.x) {
    const int jj = (int)(t % n1);
    const long long r = t / n1;
    const int ii = (int)(r % n2), kk = (int)(r / n2);
    double* q = p + (long long)(j0 + jj) + (long long)(i0 + ii) * si + (long long)(k0 + kk) * sk;
    if (PACK) buf[t] = *q; else *q = buf[t];
  }
}
// same with a strided buffer: buf[jj + b1*ii + b2*kk] (b1 = row pitch, b2 = plane pitch of the caller's array)
template <bool PACK>
__global__ void __launch_bounds__(256) k_window_strided(double* __restrict__ buf, double* __restrict__ p, const int j0, const int i0,
                                                        const int k0, const int n1, const int n2, const int nk, const long long b1,
                                                        const long long b2, const int si, const long long sk) {
  const long long n = (long long)n1 * n2 * nk;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (long long)gridDim.x * blockDim.x) {
    const int jj = (int)(t % n1);
    const long long r = t / n1;
    const int ii = (int)(r % n2), kk = (int)(r / n2);
    double* q = p + (long long)(j0 + jj) + (long long)(i0 + ii) * si + (long long)(k0 + kk) * sk;
    double* b = buf + jj + b1 * ii + b2 * kk;
    if (PACK) *b = *q; else *q = *b;
  }
}
}  // namespace

void launch_window_strided(bool pack, double* buf, double* p_shifted, int j0, int i0, int k0, int n1, int n2, int nk, long long b1,
                           long long b2, int si, long long sk, cudaStream_t st) {
  const long long n = (long long)n1 * n2 * nk;
  if (n <= 0) return;
  const unsigned blocks = (unsigned)std::min<long long>((n + 255) / 256, 148LL * 16);
#ifdef RCM_EMU
  if (pack) RCM_LAUNCH(k_window_strided<true>, dim3(blocks), dim3(256), st, buf, p_shifted, j0, i0, k0, n1, n2, nk, b1, b2, si, sk);
  else RCM_LAUNCH(k_window_strided<false>, dim3(blocks), dim3(256), st, buf, p_shifted, j0, i0, k0, n1, n2, nk, b1, b2, si, sk);
#else
  if (pack) k_window_strided<true><<<blocks, 256, 0, st>>>(buf, p_shifted, j0, i0, k0, n1, n2, nk, b1, b2, si, sk);
  else k_window_strided<false><<<blocks, 256, 0, st>>>(buf, p_shifted, j0, i0, k0, n1, n2, nk, b1, b2, si, sk);
#endif
}

void launch_window(bool pack, double* buf, double* p_shifted, int j0, int i0, int k0, int n1, int n2, int nk, int si, long long sk,
                   cudaStream_t st) {
  const long long n = (long long)n1 * n2 * nk;
  if (n <= 0) return;
  const unsigned blocks = (unsigned)std::min<long long>((n + 255) / 256, 148LL * 16);
#ifdef RCM_EMU
  if (pack) RCM_LAUNCH(k_window<true>, dim3(blocks), dim3(256), st, buf, p_shifted, j0, i0, k0, n1, n2, nk, si, sk);
  else RCM_LAUNCH(k_window<false>, dim3(blocks), dim3(256), st, buf, p_shifted, j0, i0, k0, n1, n2, nk, si, sk);
#else
  if (pack) k_window<true><<<blocks, 256, 0, st>>>(buf, p_shifted, j0, i0, k0, n1, n2, nk, si, sk);
  else k_window<false><<<blocks, 256, 0, st>>>(buf, p_shifted, j0, i0, k0, n1, n2, nk, si, sk);
#endif
}
