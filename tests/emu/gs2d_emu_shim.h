// gs2d_emu_shim.h -- TEST INFRASTRUCTURE: what csrc/common.cuh gives the small next-row kernels (optim.cu, compact.cu,
// pose.cu, loss.cu), restated for a host build: they use nothing of it but the launch bookkeeping.  The CUDA side
// (vector types, intrinsics, runtime calls) is tests/emu/fake_cuda/cuda_runtime.h.
#pragma once

#include <cuda_runtime.h>

#include "gs2d.h"

namespace gs2d {
enum KernelId { K_LOSS, K_MAP_STEP, K_POSE, K_MAP_COMPACT, K_COUNT };
struct ProfScope { ProfScope(int, cudaStream_t) {} };
inline void count_launch(int = 1) {}
}  // namespace gs2d

#define GS2D_CUDA_OK(expr) do { if ((expr) != cudaSuccess) return GS2D_ERR_CUDA; } while (0)
