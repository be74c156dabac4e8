// gs2d_emu_shim.h -- TEST INFRASTRUCTURE: what csrc/common.cuh and the CUDA runtime give the small next-row kernels
// (optim.cu, compact.cu, pose.cu, loss.cu), restated for a host build under cuda_emu.h.  Only what those files use.
#pragma once

#include <atomic>
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <cstring>

#include "cuda_emu.h"
#include "gs2d.h"

using std::isfinite;
using std::isnan;

// ---- vector types -----------------------------------------------------------------------------------------------------
struct float2 { float x, y; };
struct float3 { float x, y, z; };
struct alignas(16) float4 { float x, y, z, w; };
struct uint2 { unsigned x, y; };
inline float2 make_float2(float x, float y) { return float2{x, y}; }
inline float3 make_float3(float x, float y, float z) { return float3{x, y, z}; }
inline float4 make_float4(float x, float y, float z, float w) { return float4{x, y, z, w}; }

inline float __int_as_float(int v) { float f; std::memcpy(&f, &v, 4); return f; }
inline int __float_as_int(float f) { int v; std::memcpy(&v, &f, 4); return v; }
inline float rsqrtf(float x) { return 1.0f / std::sqrt(x); }
inline float __fdividef(float a, float b) { return a / b; }
template <class T> inline T __ldg(const T* p) { return *p; }
template <class T> inline T __ldcg(const T* p) { return *p; }
template <class T> inline void __stcg(T* p, T v) { *p = v; }

// ---- runtime ----------------------------------------------------------------------------------------------------------
typedef void* cudaStream_t;
typedef int cudaError_t;
constexpr cudaError_t cudaSuccess = 0;
inline cudaError_t cudaGetLastError() { return cudaSuccess; }
inline cudaError_t cudaMemsetAsync(void* p, int v, size_t n, cudaStream_t) { std::memset(p, v, n); return cudaSuccess; }
enum cudaMemcpyKind { cudaMemcpyHostToDevice, cudaMemcpyDeviceToHost, cudaMemcpyDeviceToDevice, cudaMemcpyDefault };
inline cudaError_t cudaMemcpyAsync(void* d, const void* s, size_t n, cudaMemcpyKind, cudaStream_t) { std::memcpy(d, s, n); return cudaSuccess; }

// blocks run one after another and every CUDA thread is a host thread: a sequentially consistent RMW / fence models
// atomicAdd / __threadfence for the "last block finishes the reduction" tickets these kernels use
inline unsigned atomicAdd(unsigned* p, unsigned v) { return __atomic_fetch_add(p, v, __ATOMIC_SEQ_CST); }
inline int atomicAdd(int* p, int v) { return __atomic_fetch_add(p, v, __ATOMIC_SEQ_CST); }
inline void __threadfence() { std::atomic_thread_fence(std::memory_order_seq_cst); }

// ---- common.cuh ---------------------------------------------------------------------------------------------------------
namespace gs2d {
enum KernelId { K_LOSS, K_MAP_STEP, K_POSE, K_MAP_COMPACT, K_COUNT };
struct ProfScope { ProfScope(int, cudaStream_t) {} };
inline void count_launch(int = 1) {}
}  // namespace gs2d

#define GS2D_CUDA_OK(expr) do { if ((expr) != cudaSuccess) return GS2D_ERR_CUDA; } while (0)
