// binning.cu -- key duplication (debug view of the reference's unsorted list) and the onesweep radix sort
// (stand-alone gs2d_sort_pairs + the Morton sort of distCUDA2).  The rendering path bins per tile: tile_sort.cu.
//
// Replaces duplicateWithKeys (CR/rasterizer_impl.cu:70-111), cub::DeviceRadixSort::SortPairs
// (:309-314) and cudaMemset + identifyTileRanges (:116-138, :316-323).  Keys, sorted order and
// tile ranges are bit-identical to the reference by construction: same key (tile << 32 | depth
// bits), same emission order (tile rows outer, columns inner), stable LSD sort.
//
// All kernels here are HBM-bound integer/byte movers.
#include <atomic>

#include "common.cuh"
#include "pair.cuh"

namespace gs2d {

namespace {

// ---- key duplication ---------------------------------------------------------------------------
// One thread per surfel (like the reference).  Besides the unsorted (key, surfel) pairs it
// accumulates the per-digit histograms of all radix passes, so the sort needs no extra
// read of the keys.
__global__ void __launch_bounds__(256)
duplicate_kernel(int P, unsigned gx, const float* __restrict__ depths, const ushort4* __restrict__ rect,
                 const uint32_t* __restrict__ tiles_touched, const uint32_t* __restrict__ offsets,
                 uint64_t* __restrict__ keys, uint32_t* __restrict__ vals, uint32_t* __restrict__ hist,
                 int passes)
{
	__shared__ uint32_t s_hist[SORT_MAX_PASSES * SORT_RADIX];
	for (int i = threadIdx.x; i < passes * SORT_RADIX; i += blockDim.x) s_hist[i] = 0;
	__syncthreads();
	int idx = blockIdx.x * blockDim.x + threadIdx.x;
	if (idx < P) {
		uint32_t n = tiles_touched[idx];
		if (n > 0) {
			uint32_t off = (idx == 0) ? 0u : offsets[idx - 1];
			const ushort4 r = rect[idx];
			const uint32_t dbits = __float_as_uint(depths[idx]);
			// digits of the depth half are the same for every instance of this surfel
			for (int p = 0; p < 4 && p < passes; p++)
				atomicAdd(&s_hist[p * SORT_RADIX + ((dbits >> (8 * p)) & 0xff)], n);
			for (unsigned y = r.y; y < r.w; y++) {
				for (unsigned x = r.x; x < r.z; x++) {
					uint32_t tile = y * gx + x;
					keys[off] = ((uint64_t)tile << 32) | dbits;
					vals[off] = (uint32_t)idx;
					for (int p = 4; p < passes; p++)
						atomicAdd(&s_hist[p * SORT_RADIX + ((tile >> (8 * (p - 4))) & 0xff)], 1u);
					off++;
				}
			}
		}
	}
	__syncthreads();
	for (int i = threadIdx.x; i < passes * SORT_RADIX; i += blockDim.x) {
		uint32_t c = s_hist[i];
		if (c) atomicAdd(&hist[i], c);
	}
}

// Stand-alone histogram for gs2d_sort_pairs (keys not produced by duplicate_kernel).
__global__ void __launch_bounds__(256)
histogram_kernel(long long n, const uint64_t* __restrict__ keys, uint32_t* __restrict__ hist, int passes)
{
	__shared__ uint32_t s_hist[SORT_MAX_PASSES * SORT_RADIX];
	for (int i = threadIdx.x; i < passes * SORT_RADIX; i += blockDim.x) s_hist[i] = 0;
	__syncthreads();
	for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
	     i += (long long)gridDim.x * blockDim.x) {
		uint64_t k = keys[i];
		for (int p = 0; p < passes; p++)
			atomicAdd(&s_hist[p * SORT_RADIX + (uint32_t)((k >> (8 * p)) & 0xff)], 1u);
	}
	__syncthreads();
	for (int i = threadIdx.x; i < passes * SORT_RADIX; i += blockDim.x) {
		uint32_t c = s_hist[i];
		if (c) atomicAdd(&hist[i], c);
	}
}

// Exclusive scan of each pass's 256-bin histogram (one block per pass).
__global__ void __launch_bounds__(SORT_RADIX)
digit_base_kernel(uint32_t* hist)
{
	__shared__ uint32_t s_warp[SORT_RADIX / 32];
	uint32_t* h = hist + blockIdx.x * SORT_RADIX;
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	uint32_t v = h[threadIdx.x];
	uint32_t incl = v;
#pragma unroll
	for (int o = 1; o < 32; o <<= 1) {
		uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
		if (lane >= o) incl += t;
	}
	if (lane == 31) s_warp[warp] = incl;
	__syncthreads();
	uint32_t off = 0;
	for (int w = 0; w < warp; w++) off += s_warp[w];
	h[threadIdx.x] = off + incl - v;
}

// ---- onesweep pass -------------------------------------------------------------------------------
// One launch per 8-bit digit.  Each block takes a ticket (so that a block only ever waits on
// blocks that are already running), ranks its 4096 pairs stably by digit with warp-level match,
// resolves the digit's global offset with a decoupled look-back over the earlier blocks, and
// scatters through shared memory so that the global writes are coalesced runs.
constexpr uint32_t LB_AGG = 1u << 30, LB_PREFIX = 2u << 30, LB_MASK = (1u << 30) - 1;

struct SortSmem {
	union {
		uint64_t keys[SORT_TILE];
		uint32_t vals[SORT_TILE];
	};
	uint32_t warp_cnt[SORT_THREADS / 32][SORT_RADIX];
	uint32_t block_excl[SORT_RADIX];   // start of each digit's run in the block-local order
	uint32_t global_base[SORT_RADIX];  // global position of that run's first element
	uint32_t scan_warp[SORT_RADIX / 32];
	uint32_t ticket;
};

__global__ void __launch_bounds__(SORT_THREADS)
onesweep_kernel(long long n, const uint64_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                uint64_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out, int shift,
                const uint32_t* __restrict__ digit_base, uint32_t* status, uint32_t* ticket_ctr)
{
	extern __shared__ __align__(16) unsigned char smem_raw[];
	SortSmem& sm = *reinterpret_cast<SortSmem*>(smem_raw);
	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

	if (tid == 0) sm.ticket = atomicAdd(ticket_ctr, 1u);
	for (int i = tid; i < (SORT_THREADS / 32) * SORT_RADIX; i += SORT_THREADS) (&sm.warp_cnt[0][0])[i] = 0;
	__syncthreads();
	const uint32_t bid = sm.ticket;
	const long long tile_base = (long long)bid * SORT_TILE;
	const int n_valid = (int)min((long long)SORT_TILE, n - tile_base);

	// warp-striped load: item i of lane l in warp w is element w*512 + i*32 + l of the tile
	uint64_t key[SORT_ITEMS];
	uint32_t rank[SORT_ITEMS];
	const int wbase = warp * (32 * SORT_ITEMS) + lane;
#pragma unroll
	for (int i = 0; i < SORT_ITEMS; i++) {
		int li = wbase + i * 32;
		key[i] = (li < n_valid) ? keys_in[tile_base + li] : ~0ull;
	}
	// stable rank within the warp, in (item, lane) order == memory order
	const uint32_t lt_mask = (1u << lane) - 1u;
#pragma unroll
	for (int i = 0; i < SORT_ITEMS; i++) {
		int li = wbase + i * 32;
		bool valid = li < n_valid;
		uint32_t d = (uint32_t)(key[i] >> shift) & 0xff;
		uint32_t vmask = __ballot_sync(0xffffffffu, valid);
		uint32_t peers = __match_any_sync(0xffffffffu, d) & vmask;
		uint32_t before = 0;
		if (valid) before = sm.warp_cnt[warp][d];
		__syncwarp();
		rank[i] = before + __popc(peers & lt_mask);
		if (valid && (peers & lt_mask) == 0) sm.warp_cnt[warp][d] = before + __popc(peers);
		__syncwarp();
	}
	__syncthreads();

	// per-digit: exclusive offsets of each warp, block total (thread d owns digit d)
	uint32_t total = 0;
#pragma unroll
	for (int w = 0; w < SORT_THREADS / 32; w++) {
		uint32_t c = sm.warp_cnt[w][tid];
		sm.warp_cnt[w][tid] = total;
		total += c;
	}
	// publish the block aggregate, then look back for the exclusive prefix of this digit
	uint32_t* my_status = status + (size_t)bid * SORT_RADIX + tid;
	uint32_t excl = 0;
	if (bid == 0) {
		atomicExch(my_status, LB_PREFIX | total);
	} else {
		atomicExch(my_status, LB_AGG | total);
		// look back eight predecessors at a time: the loads are independent, so their L2 latencies overlap
		int pb = (int)bid - 1;
		bool done = false;
		while (!done) {
			uint32_t sv[8];
#pragma unroll
			for (int k = 0; k < 8; k++)
			{
				sv[k] = 2u << 30;  // before block 0: an inclusive prefix of 0
				if (pb - k >= 0) sv[k] = *((volatile uint32_t*)(status + (size_t)(pb - k) * SORT_RADIX + tid));
			}
			int used = 0;
#pragma unroll
			for (int k = 0; k < 8; k++) {
				if (!done && used == k && (sv[k] >> 30) != 0) {
					excl += sv[k] & LB_MASK;
					used = k + 1;
					done = (sv[k] >> 30) == 2;
				}
			}
			pb -= used;
		}
		atomicExch(my_status, LB_PREFIX | (excl + total));
	}
	sm.global_base[tid] = digit_base[tid] + excl;
	// exclusive scan of the block totals over digits
	uint32_t incl = total;
#pragma unroll
	for (int o = 1; o < 32; o <<= 1) {
		uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
		if (lane >= o) incl += t;
	}
	if (lane == 31) sm.scan_warp[warp] = incl;
	__syncthreads();
	uint32_t woff = 0;
	for (int w = 0; w < warp; w++) woff += sm.scan_warp[w];
	sm.block_excl[tid] = woff + incl - total;
	__syncthreads();

	// block-local position of every item; stage keys in shared memory in sorted order
	uint32_t pos[SORT_ITEMS];
#pragma unroll
	for (int i = 0; i < SORT_ITEMS; i++) {
		int li = wbase + i * 32;
		uint32_t d = (uint32_t)(key[i] >> shift) & 0xff;
		pos[i] = sm.block_excl[d] + sm.warp_cnt[warp][d] + rank[i];
		if (li < n_valid) sm.keys[pos[i]] = key[i];
	}
	__syncthreads();
	uint32_t dst[SORT_ITEMS];
#pragma unroll
	for (int i = 0; i < SORT_ITEMS; i++) {
		int p = i * SORT_THREADS + tid;
		dst[i] = 0;
		if (p < n_valid) {
			uint64_t k = sm.keys[p];
			uint32_t d = (uint32_t)(k >> shift) & 0xff;
			dst[i] = sm.global_base[d] + ((uint32_t)p - sm.block_excl[d]);
			keys_out[dst[i]] = k;
		}
	}
	__syncthreads();
#pragma unroll
	for (int i = 0; i < SORT_ITEMS; i++) {
		int li = wbase + i * 32;
		if (li < n_valid) sm.vals[pos[i]] = vals_in[tile_base + li];
	}
	__syncthreads();
#pragma unroll
	for (int i = 0; i < SORT_ITEMS; i++) {
		int p = i * SORT_THREADS + tid;
		if (p < n_valid) vals_out[dst[i]] = sm.vals[p];
	}
}

}  // namespace

uint32_t higher_msb(uint32_t n)
{
	// number of bits needed to represent n (0 -> 0): the next-highest bit of the MSB
	uint32_t b = 0;
	while (b < 32 && (n >> b)) b++;
	return b;
}

int launch_duplicate(const gs2d_settings& s, const GeomState& g, const int32_t*, const BinningState& b,
                     int64_t I, cudaStream_t st)
{
	(void)I;
	const unsigned gx = (s.image_width + TILE - 1) / TILE, gy = (s.image_height + TILE - 1) / TILE;
	const int passes = sort_passes(32 + (int)higher_msb(gx * gy));
	{ ProfScope ps(K_MEMSET, st); GS2D_CUDA_OK(cudaMemsetAsync(b.sort.hist, 0, b.sort.zero_bytes, st)); }
	{
		ProfScope ps(K_DUPLICATE, st);
		duplicate_kernel<<<(s.P + 255) / 256, 256, 0, st>>>(s.P, gx, g.depths, (const ushort4*)g.rect,
		                                                    g.tiles_touched, g.point_offsets, b.keys_dbg, b.vals_dbg,
		                                                    b.sort.hist, passes);
		count_launch();
	}
	GS2D_CUDA_OK(cudaGetLastError());
	return GS2D_OK;
}

// Sorts (keys_a, vals_a) on bits [0, end_bit).  Histograms in ss.hist must already hold the digit
// counts when `have_hist` (they are turned into exclusive bases here).  The passes ping-pong
// a -> b -> a ...; *result_in_b tells where the sorted data ends up.
static int sort_passes_run(int64_t n, uint64_t* keys_a, uint32_t* vals_a, uint64_t* keys_b, uint32_t* vals_b,
                           int end_bit, const SortScratch& ss, bool* result_in_b, cudaStream_t st)
{
	const int passes = sort_passes(end_bit);
	if (passes > SORT_MAX_PASSES) return GS2D_ERR_INVALID;
	{
		// the dynamic shared-memory opt-in is per device: made once for every device the sort runs on
		static std::atomic<unsigned long long> opted_in{0};
		int dev = 0;
		GS2D_CUDA_OK(cudaGetDevice(&dev));
		if (dev >= 64 || !((opted_in.load(std::memory_order_relaxed) >> dev) & 1ull)) {
			GS2D_CUDA_OK(cudaFuncSetAttribute(onesweep_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
			                                  (int)sizeof(SortSmem)));
			if (dev < 64) opted_in.fetch_or(1ull << dev, std::memory_order_relaxed);
		}
	}
	{
		ProfScope ps(K_DIGIT_BASE, st);
		digit_base_kernel<<<passes, SORT_RADIX, 0, st>>>(ss.hist);
		count_launch();
	}
	const int blocks = (int)((n + SORT_TILE - 1) / SORT_TILE);
	bool in_a = true;
	for (int p = 0; p < passes; p++) {
		const uint64_t* ki = in_a ? keys_a : keys_b;
		const uint32_t* vi = in_a ? vals_a : vals_b;
		uint64_t* ko = in_a ? keys_b : keys_a;
		uint32_t* vo = in_a ? vals_b : vals_a;
		{
			ProfScope ps(K_ONESWEEP, st);
			onesweep_kernel<<<blocks, SORT_THREADS, sizeof(SortSmem), st>>>(
				n, ki, vi, ko, vo, p * SORT_RADIX_BITS, ss.hist + p * SORT_RADIX,
				ss.status + (size_t)p * ss.blocks * SORT_RADIX, ss.tickets + p);
			count_launch();
		}
		in_a = !in_a;
	}
	GS2D_CUDA_OK(cudaGetLastError());
	*result_in_b = !in_a;
	return GS2D_OK;
}

int launch_sort_pairs(int64_t n, const uint64_t* keys_in, const uint32_t* vals_in, uint64_t* keys_b,
                      uint32_t* vals_b, uint64_t* keys_a_rw, uint32_t* vals_a_rw, int end_bit,
                      const SortScratch& ss, bool* result_in_b, cudaStream_t st)
{
	// keys_in/vals_in == keys_a_rw/vals_a_rw when the histograms were built by duplicate_kernel
	// (ss.hist already populated); otherwise copy in and histogram here.
	if (n <= 0) { *result_in_b = false; return GS2D_OK; }
	if (keys_in != keys_a_rw) {
		const int passes = sort_passes(end_bit);
		{ ProfScope ps(K_MEMSET, st); GS2D_CUDA_OK(cudaMemsetAsync(ss.hist, 0, ss.zero_bytes, st)); }
		GS2D_CUDA_OK(cudaMemcpyAsync(keys_a_rw, keys_in, sizeof(uint64_t) * (size_t)n, cudaMemcpyDeviceToDevice, st));
		GS2D_CUDA_OK(cudaMemcpyAsync(vals_a_rw, vals_in, sizeof(uint32_t) * (size_t)n, cudaMemcpyDeviceToDevice, st));
		int64_t want = (n + 256 * 16 - 1) / (256 * 16);
		int blocks = (int)(want < 148 * 8 ? want : 148 * 8);
		{
			ProfScope ps(K_HISTOGRAM, st);
			histogram_kernel<<<blocks, 256, 0, st>>>(n, keys_a_rw, ss.hist, passes);
			count_launch();
		}
	}
	return sort_passes_run(n, keys_a_rw, vals_a_rw, keys_b, vals_b, end_bit, ss, result_in_b, st);
}

}  // namespace gs2d
