// tma_emu.cuh -- TEST INFRASTRUCTURE: host stand-in for csrc/tma.cuh (mbarrier + 1-D bulk async copy) under cuda_emu.h.
//
// An mbarrier keeps, in its 8 bytes of "shared memory": the phase bit, the arrivals still pending in the current phase,
// the arrival count it was initialised with, and the transaction bytes still expected.  A phase completes when both
// pending counts reach zero (PTX ISA, mbarrier: expect-tx / complete-tx).  The bulk copy is done synchronously by the
// issuing thread (memcpy, then complete_tx), which is one of the orders the hardware may produce.  All updates go
// through one mutex; waiters poll and yield.
#pragma once
#include <cstdint>
#include <chrono>
#include <cstring>
#include <mutex>
#include <thread>

#include <cuda_runtime.h>

namespace gs2d {

namespace tma_emu {
struct State {              // packed into the barrier's uint64_t
	uint32_t tx;            // transaction bytes still expected
	uint16_t pending;       // arrivals still expected in this phase
	uint8_t count;          // arrivals per phase
	uint8_t phase;          // parity of the CURRENT (incomplete) phase
};
static_assert(sizeof(State) == 8, "mbarrier state must fit its 8 bytes");
inline std::mutex& mu() { static std::mutex m; return m; }
inline State load(const uint64_t* bar) { State s; uint64_t v = __atomic_load_n(bar, __ATOMIC_SEQ_CST); std::memcpy(&s, &v, 8); return s; }
inline void store(uint64_t* bar, State s) { uint64_t v; std::memcpy(&v, &s, 8); __atomic_store_n(bar, v, __ATOMIC_SEQ_CST); }
inline void settle(State& s)
{
	if (s.pending == 0 && s.tx == 0) { s.phase ^= 1; s.pending = s.count; }
}
}  // namespace tma_emu

inline void mbar_init(uint64_t* bar, uint32_t count)
{
	std::lock_guard<std::mutex> lk(tma_emu::mu());
	tma_emu::store(bar, tma_emu::State{0u, (uint16_t)count, (uint8_t)count, 0});
}

inline void mbar_fence_init() { __threadfence(); }
inline void fence_proxy_async() { __threadfence(); }

inline void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes)
{
	std::lock_guard<std::mutex> lk(tma_emu::mu());
	tma_emu::State s = tma_emu::load(bar);
	s.tx += bytes;
	s.pending -= 1;
	tma_emu::settle(s);          // bytes == 0: the arrival alone may complete the phase
	tma_emu::store(bar, s);
}

// try_wait.parity(p): true once the phase of parity p has completed, i.e. the current phase has the other parity
inline void mbar_wait(uint64_t* bar, uint32_t parity)
{
	// hundreds of waiters share a few host cores with the one thread that issues the copies: back off quickly
	for (int spins = 0; tma_emu::load(bar).phase == (uint8_t)(parity & 1u); spins++) {
		if (spins < 4) std::this_thread::yield();
		else std::this_thread::sleep_for(std::chrono::microseconds(spins < 64 ? 20 : 200));
	}
}

inline void tma_load_1d(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar)
{
	std::memcpy(smem_dst, gmem_src, bytes);
	std::lock_guard<std::mutex> lk(tma_emu::mu());
	tma_emu::State s = tma_emu::load(bar);
	s.tx -= bytes;
	tma_emu::settle(s);
	tma_emu::store(bar, s);
}

}  // namespace gs2d
