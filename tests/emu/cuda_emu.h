// cuda_emu.h -- TEST INFRASTRUCTURE: runs the SOURCE TEXT of small CUDA kernels on the host, one std::thread per CUDA
// thread, so that a kernel's indexing / scan / ranking logic can be checked in a container without a GPU.
//
// What it models: a grid of blocks executed one block after another; inside a block every CUDA thread is a host
// thread; __syncthreads* are a block-wide std::barrier; warp intrinsics (__shfl_up_sync, __shfl_down_sync,
// __shfl_sync, __shfl_xor_sync, __ballot_sync, __any_sync, __all_sync) exchange through a per-warp buffer with a
// per-warp barrier (full masks only: all 32 lanes must call); `__shared__` variables become function-local statics
// (one block runs at a time), `extern __shared__` arrays a per-block buffer.  What it does NOT model: memory ordering, bank conflicts, partial warps at a
// divergent call site, timing.  It checks arithmetic and indexing -- not performance, not hardware semantics.
//
// Restrictions on the kernels: every thread of a block reaches the same barriers (a thread may return early only
// after its last barrier); blockDim.x is a multiple of 32; one-dimensional grids and blocks.
#pragma once

#include <atomic>
#include <barrier>
#include <cassert>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <memory>
#include <thread>
#include <vector>

namespace cuda_emu {

struct Dim { unsigned x = 1, y = 1, z = 1; };

struct Warp {
	std::barrier<> bar{32};
	long long slot[32];
};

struct Block {
	explicit Block(int threads, size_t dyn_smem = 0) : bar(threads), warps((threads + 31) / 32), smem_raw(dyn_smem + 256)
	{
		smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw.data()) + 255) & ~(uintptr_t)255);
	}
	std::barrier<> bar;
	std::atomic<int> acc{0};
	std::vector<Warp> warps;
	std::vector<unsigned char> smem_raw;     // `extern __shared__` storage of this block (the launch's third argument)
	unsigned char* smem;
};

inline thread_local Dim t_threadIdx, t_blockIdx;
inline Dim g_blockDim, g_gridDim;
inline thread_local Block* t_block = nullptr;

inline Warp& warp() { return t_block->warps[t_threadIdx.x >> 5]; }
inline int lane() { return (int)(t_threadIdx.x & 31u); }

inline unsigned char* dynamic_smem() { return t_block->smem; }

template <class F> void launch(unsigned grid, unsigned block, size_t dyn_smem, F&& kernel)
{
	assert(block % 32 == 0 && block <= 1024);
	g_blockDim = Dim{block, 1, 1};
	g_gridDim = Dim{grid, 1, 1};
	if (grid == 0) return;
	// one set of host threads per launch; they walk the blocks together (a barrier after every block)
	Block ctx((int)block, dyn_smem);
	std::barrier<> block_done((std::ptrdiff_t)block);
	std::vector<std::thread> th;
	th.reserve(block);
	for (unsigned t = 0; t < block; t++)
		th.emplace_back([&, t] {
			t_threadIdx = Dim{t, 0, 0};
			t_block = &ctx;
			for (unsigned b = 0; b < grid; b++) {
				t_blockIdx = Dim{b, 0, 0};
				kernel();
				block_done.arrive_and_wait();
			}
		});
	for (auto& x : th) x.join();
}

template <class F> void launch(unsigned grid, unsigned block, F&& kernel) { launch(grid, block, (size_t)0, static_cast<F&&>(kernel)); }

inline void syncthreads() { t_block->bar.arrive_and_wait(); }

inline int syncthreads_count(int pred)
{
	if (pred) t_block->acc.fetch_add(1);
	t_block->bar.arrive_and_wait();
	const int r = t_block->acc.load();
	t_block->bar.arrive_and_wait();
	if (t_threadIdx.x == 0) t_block->acc.store(0);
	t_block->bar.arrive_and_wait();
	return r;
}

template <class T, class Pick> T exchange(T v, Pick pick)
{
	static_assert(sizeof(T) <= sizeof(long long), "exchange: type too wide");
	Warp& w = warp();
	long long raw = 0;
	std::memcpy(&raw, &v, sizeof(T));
	w.slot[lane()] = raw;
	w.bar.arrive_and_wait();
	const int src = pick(lane());
	T out = v;
	if (src >= 0 && src < 32) std::memcpy(&out, &w.slot[src], sizeof(T));
	w.bar.arrive_and_wait();
	return out;
}

template <class T> T shfl_up(T v, unsigned d) { return exchange(v, [d](int l) { return l >= (int)d ? l - (int)d : -1; }); }
template <class T> T shfl_down(T v, unsigned d) { return exchange(v, [d](int l) { return l + (int)d < 32 ? l + (int)d : -1; }); }
template <class T> T shfl_xor(T v, int m) { return exchange(v, [m](int l) { return l ^ m; }); }
template <class T> T shfl_idx(T v, int s) { return exchange(v, [s](int) { return s & 31; }); }

inline unsigned ballot(int pred)
{
	Warp& w = warp();
	w.slot[lane()] = pred ? 1 : 0;
	w.bar.arrive_and_wait();
	unsigned r = 0;
	for (int l = 0; l < 32; l++) r |= (unsigned)(w.slot[l] != 0) << l;
	w.bar.arrive_and_wait();
	return r;
}

inline unsigned match_any(unsigned v)
{
	Warp& w = warp();
	w.slot[lane()] = (long long)v;
	w.bar.arrive_and_wait();
	unsigned r = 0;
	for (int l = 0; l < 32; l++) r |= (unsigned)(w.slot[l] == (long long)v) << l;
	w.bar.arrive_and_wait();
	return r;
}

}  // namespace cuda_emu

// ---- the CUDA spellings the kernel text uses ------------------------------------------------------------------------
#define __global__
#define __device__
#define __host__
#define __forceinline__ inline
#define __restrict__
#define __launch_bounds__(...)
#define __shared__ static
#define __constant__ static
#define threadIdx cuda_emu::t_threadIdx
#define blockIdx cuda_emu::t_blockIdx
#define blockDim cuda_emu::g_blockDim
#define gridDim cuda_emu::g_gridDim

inline void __syncthreads() { cuda_emu::syncthreads(); }
inline int __syncthreads_count(int p) { return cuda_emu::syncthreads_count(p); }
inline void __syncwarp(unsigned = 0xffffffffu) { cuda_emu::warp().bar.arrive_and_wait(); }
inline int __popc(unsigned x) { return __builtin_popcount(x); }
inline int __clz(int x) { return x == 0 ? 32 : __builtin_clz((unsigned)x); }
#define EMU_FULL(mask) assert((mask) == 0xffffffffu && "cuda_emu models full-warp intrinsics only")
template <class T> T __shfl_up_sync(unsigned mask, T v, unsigned d) { EMU_FULL(mask); return cuda_emu::shfl_up(v, d); }
template <class T> T __shfl_down_sync(unsigned mask, T v, unsigned d) { EMU_FULL(mask); return cuda_emu::shfl_down(v, d); }
template <class T> T __shfl_xor_sync(unsigned mask, T v, int m) { EMU_FULL(mask); return cuda_emu::shfl_xor(v, m); }
template <class T> T __shfl_sync(unsigned mask, T v, int s) { EMU_FULL(mask); return cuda_emu::shfl_idx(v, s); }
inline unsigned __ballot_sync(unsigned mask, int p) { EMU_FULL(mask); return cuda_emu::ballot(p); }
inline int __any_sync(unsigned mask, int p) { EMU_FULL(mask); return cuda_emu::ballot(p) != 0; }
inline unsigned __match_any_sync(unsigned mask, unsigned v) { EMU_FULL(mask); return cuda_emu::match_any(v); }
inline int __all_sync(unsigned mask, int p) { EMU_FULL(mask); return cuda_emu::ballot(p) == 0xffffffffu; }
