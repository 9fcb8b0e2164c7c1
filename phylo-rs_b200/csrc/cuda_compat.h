// cuda_compat.h — the one place that decides how kernels are launched.
//
// Product build (nvcc, sm_100a): real CUDA runtime, real <<<>>> launches.
// Test-harness build (g++ -DPHB_HOST_EMU, tests/emu/): the SAME kernel and host sources compiled
// against a fiber-based CUDA-runtime shim so the scheduler / tiling / ABI logic can be exercised
// by `pytest -m "not gpu"` in a container without a GPU.  The emulation library is test
// infrastructure only: it is built under tests/emu/, never loaded by the phylo_rs_b200 package,
// and the product library has no CPU path — it fails with PHB_ERR_CUDA when no sm_100 device exists.
#pragma once

#ifdef PHB_HOST_EMU
#include "cuda_emu.h"
#define PHB_LAUNCH(kernel, grid, block, stream, ...) \
    ::emu::launch((grid), (block), [&]() { kernel(__VA_ARGS__); })
#define PHB_LAUNCH_SMEM(kernel, grid, block, smem, stream, ...) \
    ::emu::launch((grid), (block), [&]() { kernel(__VA_ARGS__); }, (smem))
#define PHB_DYN_SMEM(name) unsigned char *name = ::emu::dyn_smem
#else
#include <cuda_runtime.h>
#define PHB_LAUNCH(kernel, grid, block, stream, ...) kernel<<<(grid), (block), 0, (stream)>>>(__VA_ARGS__)
#define PHB_LAUNCH_SMEM(kernel, grid, block, smem, stream, ...) kernel<<<(grid), (block), (smem), (stream)>>>(__VA_ARGS__)
#define PHB_DYN_SMEM(name) extern __shared__ __align__(128) unsigned char name[]
#endif
