// cuda_runtime.h -- TEST INFRASTRUCTURE: a stand-in for the CUDA runtime header when the text of a .cu file is compiled
// for the HOST under tests/emu/cuda_emu.h (g++ finds this file first through -I tests/emu/fake_cuda).  Vector types,
// the device intrinsics the library's plain-CUDA kernels use (IEEE round-to-nearest forms map to the host's correctly
// rounded operations: build with -ffp-contract=off), and the handful of runtime calls their launchers make.
#pragma once

#include <atomic>
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <cstdlib>
#include <cstring>

#include "cuda_emu.h"

using std::isfinite;
using std::isnan;

// ---- vector types -----------------------------------------------------------------------------------------------------
struct float2 { float x, y; };
struct float3 { float x, y, z; };
struct alignas(16) float4 { float x, y, z, w; };
struct int2 { int x, y; };
struct uint2 { unsigned x, y; };
struct uint3 { unsigned x, y, z; };
struct alignas(16) uint4 { unsigned x, y, z, w; };
struct alignas(8) ushort4 { unsigned short x, y, z, w; };
struct dim3 { unsigned x = 1, y = 1, z = 1; dim3(unsigned x_ = 1, unsigned y_ = 1, unsigned z_ = 1) : x(x_), y(y_), z(z_) {} };
inline float2 make_float2(float x, float y) { return float2{x, y}; }
inline float3 make_float3(float x, float y, float z) { return float3{x, y, z}; }
inline float4 make_float4(float x, float y, float z, float w) { return float4{x, y, z, w}; }
inline int2 make_int2(int x, int y) { return int2{x, y}; }
inline uint2 make_uint2(unsigned x, unsigned y) { return uint2{x, y}; }
inline uint3 make_uint3(unsigned x, unsigned y, unsigned z) { return uint3{x, y, z}; }
inline uint4 make_uint4(unsigned x, unsigned y, unsigned z, unsigned w) { return uint4{x, y, z, w}; }
inline ushort4 make_ushort4(unsigned short x, unsigned short y, unsigned short z, unsigned short w) { return ushort4{x, y, z, w}; }

// CUDA's global integer / float min and max
inline int min(int a, int b) { return a < b ? a : b; }
inline int max(int a, int b) { return a > b ? a : b; }
inline unsigned min(unsigned a, unsigned b) { return a < b ? a : b; }
inline unsigned max(unsigned a, unsigned b) { return a > b ? a : b; }
inline long long min(long long a, long long b) { return a < b ? a : b; }
inline long long max(long long a, long long b) { return a > b ? a : b; }
inline float min(float a, float b) { return std::fmin(a, b); }
inline float max(float a, float b) { return std::fmax(a, b); }

// ---- intrinsics ---------------------------------------------------------------------------------------------------------
inline float __fmaf_rn(float a, float b, float c) { return std::fmaf(a, b, c); }
inline float __fmul_rn(float a, float b) { return a * b; }
inline float __fadd_rn(float a, float b) { return a + b; }
inline float __fsub_rn(float a, float b) { return a - b; }
inline float __fdiv_rn(float a, float b) { return a / b; }
inline float __frcp_rn(float a) { return 1.0f / a; }
inline float __fsqrt_rn(float a) { return std::sqrt(a); }
inline float __logf(float a) { return std::log(a); }          // approximate on the device: tolerance-only uses
inline float __expf(float a) { return std::exp(a); }
inline float __fdividef(float a, float b) { return a / b; }
inline float rsqrtf(float x) { return 1.0f / std::sqrt(x); }
inline float __int_as_float(int v) { float f; std::memcpy(&f, &v, 4); return f; }
inline int __float_as_int(float f) { int v; std::memcpy(&v, &f, 4); return v; }
inline unsigned __float_as_uint(float f) { unsigned v; std::memcpy(&v, &f, 4); return v; }
inline float __uint_as_float(unsigned v) { float f; std::memcpy(&f, &v, 4); return f; }
inline unsigned __byte_perm(unsigned x, unsigned y, unsigned s)
{
	const unsigned long long src = ((unsigned long long)y << 32) | x;
	unsigned r = 0;
	for (int i = 0; i < 4; i++) r |= (unsigned)((src >> (8 * ((s >> (4 * i)) & 7u))) & 0xffu) << (8 * i);
	return r;
}
inline unsigned __brev(unsigned x) { unsigned r = 0; for (int i = 0; i < 32; i++) r |= ((x >> i) & 1u) << (31 - i); return r; }
inline float __saturatef(float x) { return std::fmin(std::fmax(x, 0.0f), 1.0f); }
inline int __ffs(unsigned x) { return __builtin_ffs((int)x); }
[[noreturn]] inline void __trap() { std::abort(); }
template <class T> inline T __ldg(const T* p) { return *p; }
template <class T> inline T __ldcg(const T* p) { return *p; }
template <class T> inline void __stcg(T* p, T v) { *p = v; }
#ifndef __host__
#define __host__
#endif

// blocks run one after another and every CUDA thread is a host thread: sequentially consistent RMWs / fences model the
// atomics these kernels use (tickets, per-tile counters, look-back status words)
inline unsigned atomicAdd(unsigned* p, unsigned v) { return __atomic_fetch_add(p, v, __ATOMIC_SEQ_CST); }
inline int atomicAdd(int* p, int v) { return __atomic_fetch_add(p, v, __ATOMIC_SEQ_CST); }
inline unsigned long long atomicAdd(unsigned long long* p, unsigned long long v) { return __atomic_fetch_add(p, v, __ATOMIC_SEQ_CST); }
inline unsigned long long atomicExch(unsigned long long* p, unsigned long long v) { return __atomic_exchange_n(p, v, __ATOMIC_SEQ_CST); }
inline unsigned atomicExch(unsigned* p, unsigned v) { return __atomic_exchange_n(p, v, __ATOMIC_SEQ_CST); }
template <class T, class Op> inline T atomic_rmw_emu(T* p, Op op)
{
	T old = __atomic_load_n(p, __ATOMIC_SEQ_CST);
	while (!__atomic_compare_exchange_n(p, &old, op(old), false, __ATOMIC_SEQ_CST, __ATOMIC_SEQ_CST)) {}
	return old;
}
inline unsigned atomicMin(unsigned* p, unsigned v) { return atomic_rmw_emu(p, [v](unsigned o) { return o < v ? o : v; }); }
inline unsigned atomicMax(unsigned* p, unsigned v) { return atomic_rmw_emu(p, [v](unsigned o) { return o > v ? o : v; }); }
inline int atomicMin(int* p, int v) { return atomic_rmw_emu(p, [v](int o) { return o < v ? o : v; }); }
inline int atomicMax(int* p, int v) { return atomic_rmw_emu(p, [v](int o) { return o > v ? o : v; }); }
inline unsigned atomicOr(unsigned* p, unsigned v) { return __atomic_fetch_or(p, v, __ATOMIC_SEQ_CST); }
inline unsigned atomicAnd(unsigned* p, unsigned v) { return __atomic_fetch_and(p, v, __ATOMIC_SEQ_CST); }
inline float atomicAdd(float* p, float v)
{
	unsigned* u = reinterpret_cast<unsigned*>(p);
	unsigned old = __atomic_load_n(u, __ATOMIC_SEQ_CST), upd;
	float f;
	do { std::memcpy(&f, &old, 4); f += v; std::memcpy(&upd, &f, 4); } while (!__atomic_compare_exchange_n(u, &old, upd, false, __ATOMIC_SEQ_CST, __ATOMIC_SEQ_CST));
	std::memcpy(&f, &old, 4);
	return f;
}
inline void __threadfence() { std::atomic_thread_fence(std::memory_order_seq_cst); }

// ---- runtime ----------------------------------------------------------------------------------------------------------
typedef void* cudaStream_t;
typedef void* cudaEvent_t;
typedef int cudaError_t;
constexpr cudaError_t cudaSuccess = 0;
enum cudaMemcpyKind { cudaMemcpyHostToDevice, cudaMemcpyDeviceToHost, cudaMemcpyDeviceToDevice, cudaMemcpyDefault };
inline cudaError_t cudaGetLastError() { return cudaSuccess; }
inline cudaError_t cudaGetDevice(int* d) { *d = 0; return cudaSuccess; }
inline cudaError_t cudaMemsetAsync(void* p, int v, size_t n, cudaStream_t) { std::memset(p, v, n); return cudaSuccess; }
inline cudaError_t cudaMemcpyAsync(void* d, const void* s, size_t n, cudaMemcpyKind, cudaStream_t) { std::memcpy(d, s, n); return cudaSuccess; }
inline cudaError_t cudaStreamSynchronize(cudaStream_t) { return cudaSuccess; }
inline cudaError_t cudaEventCreate(cudaEvent_t* e) { *e = nullptr; return cudaSuccess; }
inline cudaError_t cudaEventRecord(cudaEvent_t, cudaStream_t) { return cudaSuccess; }
inline cudaError_t cudaEventSynchronize(cudaEvent_t) { return cudaSuccess; }
constexpr unsigned cudaStreamNonBlocking = 1, cudaEventDisableTiming = 2;
inline cudaError_t cudaStreamCreateWithFlags(cudaStream_t* s, unsigned) { static int one; *s = &one; return cudaSuccess; }   // work runs synchronously anyway
inline cudaError_t cudaEventCreateWithFlags(cudaEvent_t* e, unsigned) { static int one; *e = &one; return cudaSuccess; }
inline cudaError_t cudaStreamWaitEvent(cudaStream_t, cudaEvent_t, unsigned) { return cudaSuccess; }
inline cudaError_t cudaEventElapsedTime(float* ms, cudaEvent_t, cudaEvent_t) { *ms = 0.f; return cudaSuccess; }
inline cudaError_t cudaMemset(void* p, int v, size_t n) { std::memset(p, v, n); return cudaSuccess; }
enum cudaFuncAttribute { cudaFuncAttributeMaxDynamicSharedMemorySize };
template <class K> inline cudaError_t cudaFuncSetAttribute(K, cudaFuncAttribute, int) { return cudaSuccess; }
