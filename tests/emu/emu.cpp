// TEST INFRASTRUCTURE ONLY — see cuda_emu.h.
#include "cuda_emu.h"
namespace rcm_emu {
thread_local dim3 g_blockIdx, g_threadIdx, g_blockDim, g_gridDim;
}
