// tile_sort.cu -- tile binning without a global sort: count, scatter, sort inside each tile.
//
// Replaces duplicateWithKeys + cub::DeviceRadixSort::SortPairs + identifyTileRanges
// (CR/rasterizer_impl.cu:70-138, :296-323) for the rendering path.  The reference sorts all
// (tile << 32 | depth bits, surfel) pairs of the frame with a stable global radix sort.  The order that sort
// produces is a total order that can be stated without the sort: tiles ascending; inside a tile depth bits
// ascending; equal depths by surfel index ascending (duplicateWithKeys emits surfels in index order and a
// surfel touches a tile once, so "stable" means exactly that).  Hence:
//   1. preprocess_forward counts the instances of every tile (one RED per instance, preprocess.cu; every tile's
//      counter and cursor sit in a 128-byte line of their own, TC_STRIDE: L2 serialises atomics on a line);
//   2. one extra block of the tiles_touched scan (preprocess.cu) scans the 3 225 counts -> tile ranges
//      (identifyTileRanges's output) + write cursors;
//   3. scatter_kernel drops every instance's (depth bits << 32 | surfel) word into its tile's segment, in
//      whatever order the atomics resolve;
//   4. tile_sort_kernel (one CTA per tile) sorts the segment by that 64-bit word in shared memory -- unique
//      keys, so the result does not depend on the scatter order and equals the reference's sorted list bit for
//      bit -- then, while the tile's list is on chip, writes keys / point list, gathers the 80-byte instance
//      records (binning.cu's gather) and computes the footprint masks (footprint.cuh).
// Six onesweep passes over 0.8 M pairs, the gather and the footprint launch become one kernel whose sort part
// works on ~250 keys in shared memory.  Tiles with more than TS_CAP instances sort in place in global memory
// with the same network (slow, correct).
#include "common.cuh"
#include "footprint.cuh"

namespace gs2d {

namespace {

#ifndef GS2D_SCATTER_UNROLL
#define GS2D_SCATTER_UNROLL 1
#endif
constexpr int TS_CAP = 2048;   // instances of a tile sorted in shared memory (16 KB of keys)

// One thread per surfel, like duplicateWithKeys -- but the destination is the tile's own segment.
__global__ void __launch_bounds__(256)
scatter_kernel(int P, unsigned gx, const float* __restrict__ depths, const ushort4* __restrict__ rect,
               const uint32_t* __restrict__ tiles_touched, uint32_t* __restrict__ cursor,
               uint64_t* __restrict__ pairs, const uint32_t* __restrict__ total, uint32_t capacity)
{
	// early launch (before the host knows num_rendered): only when the caller's arena is large enough
	if (total && *total > capacity) return;
	const int idx = blockIdx.x * blockDim.x + threadIdx.x;
	if (idx >= P) return;
	// the rectangle and the depth are requested together with the visibility test (one memory round trip, not two;
	// for an invisible surfel they are stale words that nobody uses)
	const uint32_t touched = tiles_touched[idx];
	const ushort4 r = rect[idx];
	const float depth = depths[idx];
	if (touched == 0) return;
	const uint64_t word = ((uint64_t)__float_as_uint(depth) << 32) | (uint32_t)idx;
#if GS2D_SCATTER_UNROLL
	// four cursor atomics in flight per thread: the returned slot is only needed by the store that follows
	const unsigned w = r.z - r.x, cnt = w * (r.w - r.y);
	for (unsigned e0 = 0; e0 < cnt; e0 += 4) {
		uint32_t slot[4];
#pragma unroll
		for (int k = 0; k < 4; k++) {
			const unsigned e = e0 + k;
			if (e < cnt) {
				const unsigned yy = e / w;
				slot[k] = atomicAdd(&cursor[(size_t)((r.y + yy) * gx + r.x + (e - yy * w)) * TC_STRIDE], 1u);
			}
		}
#pragma unroll
		for (int k = 0; k < 4; k++)
			if (e0 + k < cnt) pairs[slot[k]] = word;
	}
#else
	for (unsigned y = r.y; y < r.w; y++)
		for (unsigned x = r.x; x < r.z; x++) pairs[atomicAdd(&cursor[(size_t)(y * gx + x) * TC_STRIDE], 1u)] = word;
#endif
}

// Ascending sort of keys[0, n) by 256 threads in memory (the fallback for tiles beyond TS_CAP): bitonic network in its "every exchange puts the smaller key at
// the lower index" form (first step of a merge mirrors the partner), so the virtual +inf padding up to the next
// power of two never moves and pairs that reach it are simply skipped.
__device__ __forceinline__ void cta_sort(uint64_t* keys, int n, int tid)
{
	int N = 1;
	while (N < n) N <<= 1;
	const int half = N >> 1;
	for (int k = 2; k <= N; k <<= 1) {
		const int hk = k >> 1;
		for (int t = tid; t < half; t += 256) {
			const int blk = t / hk, off = t - blk * hk;
			const int i = blk * k + off, j = blk * k + (k - 1 - off);
			if (j < n) {
				const uint64_t a = keys[i], b = keys[j];
				if (a > b) { keys[i] = b; keys[j] = a; }
			}
		}
		__syncthreads();
		for (int s = hk >> 1; s >= 1; s >>= 1) {
			for (int t = tid; t < half; t += 256) {
				const int i = ((t / s) * 2 * s) + (t % s), j = i + s;
				if (j < n) {
					const uint64_t a = keys[i], b = keys[j];
					if (a > b) { keys[i] = b; keys[j] = a; }
				}
			}
			__syncthreads();
		}
	}
}

// Register version for n <= 256 * E: thread t holds the keys at positions r * 256 + t (r < E), padded with
// +inf.  One exchange step pairs position e with e ^ mask; the position whose `dirbit` is clear keeps the
// smaller key.  The partner sits in the same thread (mask is a multiple of 256), in the same warp (shuffle) or
// elsewhere (through shared memory); the network is unrolled at compile time, so every mask is a constant.
template <int E>
__device__ __forceinline__ void exchange_step(uint64_t (&key)[E], const int mask, const int dirbit, uint64_t* smem,
                                              int tid)
{
	const int mr = mask >> 8, mt = mask & 255;
	if (mt == 0) {
#pragma unroll
		for (int r = 0; r < E; r++) {
			if ((r ^ mr) > r) {
				const uint64_t a = key[r], b = key[r ^ mr];
				key[r] = a < b ? a : b;
				key[r ^ mr] = a < b ? b : a;
			}
		}
		return;
	}
	uint64_t got[E];
	if (mt < 32) {
#pragma unroll
		for (int r = 0; r < E; r++) got[r] = __shfl_xor_sync(0xffffffffu, key[r ^ mr], mt);
	} else {
#pragma unroll
		for (int r = 0; r < E; r++) smem[r * 256 + tid] = key[r];
		__syncthreads();
#pragma unroll
		for (int r = 0; r < E; r++) got[r] = smem[(r ^ mr) * 256 + (tid ^ mt)];
		__syncthreads();
	}
#pragma unroll
	for (int r = 0; r < E; r++) {
		const bool lower = (((r << 8) | tid) & dirbit) == 0;
		const uint64_t a = key[r], b = got[r];
		key[r] = (lower == (a < b)) ? a : b;
	}
}

// Loads the segment, sorts it, leaves the sorted keys in smem[0, n).
template <int E>
__device__ __forceinline__ void cta_sort_regs(const uint64_t* seg, int n, uint64_t* smem, int tid)
{
	uint64_t key[E];
#pragma unroll
	for (int r = 0; r < E; r++) key[r] = (r * 256 + tid) < n ? seg[r * 256 + tid] : ~0ull;
#pragma unroll
	for (int k = 2; k <= 256 * E; k <<= 1) {
		exchange_step<E>(key, k - 1, k >> 1, smem, tid);
#pragma unroll
		for (int s = k >> 2; s >= 1; s >>= 1) exchange_step<E>(key, s, s, smem, tid);
	}
#pragma unroll
	for (int r = 0; r < E; r++) smem[r * 256 + tid] = key[r];
	__syncthreads();
}

#ifndef GS2D_TS_GATHER_ILP
#define GS2D_TS_GATHER_ILP 4      // record parts in flight per thread in the gather (0 = one at a time)
#endif
#ifndef GS2D_TS_MIN_BLOCKS
#define GS2D_TS_MIN_BLOCKS 5
#endif
__global__ void __launch_bounds__(256, GS2D_TS_MIN_BLOCKS)
tile_sort_kernel(int W, int H, unsigned gx, const uint2* __restrict__ ranges, uint64_t* pairs,
                 const float4* __restrict__ splat, const ushort4* __restrict__ rect,
                 const float4* __restrict__ cull, const uint32_t* __restrict__ offsets,
                 uint64_t* __restrict__ keys_sorted, uint32_t* __restrict__ point_list, float4* instances,
                 uint32_t* __restrict__ hit_words, uint32_t* __restrict__ hit_count,
                 uint32_t* __restrict__ zero_words, int n_zero_words, unsigned tile0)
{
	__shared__ uint64_t s_keys[TS_CAP];
	const int tid = threadIdx.x;
	const unsigned tile = blockIdx.x + tile0;      // the launch covers tiles [tile0, tile0 + gridDim.x)
	// status words + counters of the pair-count scan that follows (instead of a memset launch)
	if (tile == 0)
		for (int i = tid; i < n_zero_words; i += 256) zero_words[i] = 0u;
	const uint2 range = ranges[tile];
	const int n = (int)(range.y - range.x);
	if (n == 0) return;
	uint64_t* seg = pairs + range.x;
	uint64_t* kp = seg;
	if (n <= TS_CAP) {
		kp = s_keys;
		if (n <= 256) cta_sort_regs<1>(seg, n, s_keys, tid);
		else if (n <= 512) cta_sort_regs<2>(seg, n, s_keys, tid);
		else if (n <= 1024) cta_sort_regs<4>(seg, n, s_keys, tid);
		else cta_sort_regs<8>(seg, n, s_keys, tid);
	} else {
		__syncthreads();
		cta_sort(seg, n, tid);
	}

	const int tx = (int)(tile % gx), ty = (int)(tile / gx);
	// the reference's sorted key / value lists (gs2d_state_view, footprint audit)
	for (int k = tid; k < n; k += 256) {
		const uint64_t w = kp[k];
		keys_sorted[range.x + k] = ((uint64_t)tile << 32) | (w >> 32);
		point_list[range.x + k] = (uint32_t)w;
	}
	// gather: 5 consecutive threads copy the 5 float4 of a record -> the tile's list is written contiguously (a pure
	// copy: the two record words that depend on the tile are patched in below by the instance's own thread, so no lane
	// of this loop takes a special path).
	float4* tile_rec = instances + 5 * (size_t)range.x;
#if GS2D_TS_GATHER_ILP
	// GU record parts per thread and round, all loads requested before the first store (the records sit at
	// random places of a 40 MB array: the loop waits on DRAM, not on anything it computes)
	constexpr int GU = GS2D_TS_GATHER_ILP;
	for (int t0 = tid; t0 < 5 * n; t0 += GU * 256) {
		float4 v[GU];
#pragma unroll
		for (int u = 0; u < GU; u++) {
			const int t = t0 + u * 256;
			if (t < 5 * n) {
				const int k = t / 5, part = t - 5 * k;
				v[u] = splat[5 * (size_t)(uint32_t)kp[k] + part];
			}
		}
#pragma unroll
		for (int u = 0; u < GU; u++) {
			const int t = t0 + u * 256;
			if (t < 5 * n) tile_rec[t] = v[u];
		}
	}
#else
	for (int t = tid; t < 5 * n; t += 256) {
		const int k = t / 5, part = t - 5 * k;
		const uint32_t idx = (uint32_t)kp[k];
		tile_rec[t] = splat[5 * (size_t)idx + part];
	}
#endif
	__syncthreads();   // the records are complete copies: they are patched and read back below by other threads
	// Footprints: a thread per instance.  It first finishes its record -- word 15 = in-tile footprint box, word 19 = the
	// instance's gradient row (surfel-major, see common.cuh) -- then walks the rows of the box (the row-independent half
	// of the closed form is evaluated once per instance, not once per lane of a quarter-warp).  A thread ORs its rows
	// into its own 8 words in shared memory (stride 9: conflict-free for the owners and for the coalesced copy-out that
	// follows).  The sorted keys are still needed (surfel index), so the words have their own array.
	__shared__ uint32_t s_words[256 * 9];
	const float ox = (float)(tx * TILE), oy = (float)(ty * TILE);
	for (int q0 = 0; q0 < n; q0 += 256) {
		const int k = q0 + tid;
		uint32_t* mine = s_words + 9 * tid;
#pragma unroll
		for (int w = 0; w < 8; w++) mine[w] = 0u;
		uint32_t bw = 0, cnt = 0;
		if (k < n) {
			const uint32_t idx = (uint32_t)kp[k];
			bw = tile_box_word(cull[idx], tx, ty, W, H);
			const ushort4 r = rect[idx];
			const uint32_t first = idx == 0 ? 0u : offsets[idx - 1];
			float* rec_words = reinterpret_cast<float*>(tile_rec + 5 * (size_t)k);
			rec_words[REC_MASK] = __uint_as_float(bw);
			rec_words[REC_ROW] = __uint_as_float(first + (uint32_t)(ty - r.y) * (uint32_t)(r.z - r.x) + (uint32_t)(tx - r.x));
			footprint_rows(tile_rec + 5 * (size_t)k, bw, ox, oy, [&](int row, uint32_t bits) {
				mine[row >> 1] |= bits << (16 * (row & 1));
				cnt += __popc(bits);
			});
			hit_count[range.x + k] = cnt;
		}
		__syncthreads();
		const int words = 8 * min(n - q0, 256);
		uint32_t* dst = hit_words + 8 * ((size_t)range.x + q0);
		for (int e = tid; e < words; e += 256) dst[e] = s_words[9 * (e >> 3) + (e & 7)];
		__syncthreads();
	}
}

}  // namespace

// Scatter into `pairs` (capacity words).  With `early`, num_rendered is read on the device (the host has not
// seen it yet) and the launch is a no-op when it exceeds the capacity.
int launch_scatter(const gs2d_settings& s, const GeomState& g, const ImageState& im, uint64_t* pairs, bool early,
                   int64_t capacity, cudaStream_t st)
{
	const unsigned gx = (s.image_width + TILE - 1) / TILE;
	if (s.P <= 0) return GS2D_OK;
	ProfScope ps(K_SCATTER, st);
	const uint32_t cap = (uint32_t)(capacity < 0 ? 0 : (capacity > 0xffffffffll ? 0xffffffffll : capacity));
	scatter_kernel<<<(s.P + 255) / 256, 256, 0, st>>>(s.P, gx, g.depths, (const ushort4*)g.rect, g.tiles_touched,
	                                                  im.tile_cursor, pairs, early ? g.counters + 1 : nullptr, cap);
	count_launch();
	GS2D_CUDA_OK(cudaGetLastError());
	return GS2D_OK;
}

// `pairs`: the scattered words (b.keys_a, or the caller's arena when the scatter was launched early).
int launch_tile_sort(const gs2d_settings& s, const GeomState& g, const ImageState& im, const BinningState& b,
                     uint64_t* pairs, int64_t I, cudaStream_t st, int tile0, int ntiles)
{
	const int W = s.image_width, H = s.image_height;
	const unsigned gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
	const int tiles = (int)(gx * gy);
	if (tiles == 0 || I <= 0) return GS2D_OK;
	if (ntiles < 0) { tile0 = 0; ntiles = tiles; }     // the whole frame
	if (ntiles == 0) return GS2D_OK;
	// (tile ranges and cursors come from the extra block of the tiles_touched scan, preprocess.cu)
	ProfScope ps(K_TILE_SORT, st);
	tile_sort_kernel<<<ntiles, 256, 0, st>>>(W, H, gx, (const uint2*)im.ranges, pairs, (const float4*)g.splat,
	                                        (const ushort4*)g.rect, (const float4*)g.cull, g.point_offsets, b.keys_b,
	                                        b.vals_b, (float4*)b.instances, b.hit_inst, b.hit_count,
	                                        (uint32_t*)b.pair_status,
	                                        (int)(((char*)(b.pair_counters + 8) - (char*)b.pair_status) / sizeof(uint32_t)),
	                                        (unsigned)tile0);
	count_launch();
	GS2D_CUDA_OK(cudaGetLastError());
	return GS2D_OK;
}

}  // namespace gs2d
