// Kernel-launch macro.  The product build (nvcc) expands to a CUDA launch.
// The test-only host harness (tests/emu, -DRCM_EMU; never shipped, never loaded
// by the package) expands to a sequential loop over the grid so that kernel
// LOGIC can be unit-tested in CI containers that have no GPU.
#pragma once
#ifdef RCM_EMU
#include "../../tests/emu/cuda_emu.h"
#define RCM_LAUNCH(kern, grid, block, stream, ...) rcm_emu::launch(kern, grid, block, __VA_ARGS__)
#else
#include <cuda_runtime.h>
#define RCM_LAUNCH(kern, grid, block, stream, ...) kern<<<grid, block, 0, stream>>>(__VA_ARGS__)
#endif
