// composite_fwd.cu -- per-tile front-to-back compositing (forward).
//
// Replaces FORWARD::render / renderCUDA<3> (CR/forward.cu:258-467): ray-splat intersection,
// low-pass filter, alpha, early ray termination, surface-aware (SA) depth or 2DGS distortion,
// normals, median depth.  The per-(pixel, instance) arithmetic keeps the reference's operation
// order, so a pixel's result is the same sequence of fp32 operations.  What changes is which pairs
// are visited and how the data is fed:
//   * one CTA per 16x16 tile, one thread per pixel (warp k owns an 8 x 4 pixel block, pair.cuh);
//   * the tile's instance records (80 B each, sorted, contiguous in HBM) and their footprint masks
//     (32 B each, footprint.cu) arrive by TMA bulk copies in chunks of 128 instances; FWD_CHUNKS chunks
//     (384 instances: all of a tile for 99.9 % / 83 % of the tiles of the bench views) are requested AT ONCE
//     and stay resident together ("super-batch");
//   * the instance-major footprint masks are transposed into per-PIXEL hit words with warp shuffles,
//     and every pixel thread then walks ITS OWN set bits of the WHOLE super-batch front to back in ONE
//     loop.  A pixel visits ~18 instances per tile instead of ~300.  Lanes of a warp run independent
//     lists and the warp iterates as often as its longest list -- measured (scripts/pair_stats.py) 0.60 - 0.63
//     of the lanes busy when the lanes re-synchronise after every 128 instances (round 1), 0.75 when they
//     only do so per tile: short and long stretches of a pixel's list average out over the tile;
//   * a pixel stops at T < 1e-4 like the reference; the CTA stops staging when all 256 have;
//   * the pairs that really contributed (what the backward pass needs to know about visibility and
//     termination) are the footprint bits minus the rejected ones, in front of the termination point:
//     rejected bits are cleared in the pixel's own hit words, no second set of words is kept.
// The arithmetic stays IEEE throughout (division, expf): the SA variance D2 - 2 md D + md^2 (1 - T) is a catastrophic
// cancellation (~1e-6 from terms ~10) that the backward pass divides by, so even a 1e-6 relative change of one blend
// weight (measured: MUFU ex2 for the confidence exponent) moves gradients by more than the 1e-3 bar.  Images are
// bit-identical to the reference build.
#include <atomic>

#include "common.cuh"
#include "tma.cuh"
#include "pair.cuh"

namespace gs2d {

namespace {

constexpr int FWD_BATCH = 128;              // instances per TMA chunk
#ifndef GS2D_FWD_CHUNKS
#define GS2D_FWD_CHUNKS 3
#endif
constexpr int FWD_CHUNKS = GS2D_FWD_CHUNKS;
constexpr int FWD_SUPER = FWD_BATCH * FWD_CHUNKS;   // instances resident at once
constexpr int FWD_WORDS = FWD_SUPER / 32;           // hit words per pixel per super-batch
static_assert(FWD_BATCH / 32 <= HIT_PIX_PAD, "hit_pix padding");
constexpr int FWD_SORT_BINS = 64;                   // list lengths 0 .. 62, 63 = longer

struct FwdSmem {
	float4 rec[FWD_CHUNKS][FWD_BATCH * 5];    // contiguous: instance i of the super-batch is rec[0][5 * i]
	uint32_t mask[FWD_CHUNKS][FWD_BATCH * 8];
	uint32_t hits[FWD_WORDS][TILE_PIXELS];    // per-pixel hit words (live pixels; rejected pairs get cleared)
	uint64_t bar[FWD_CHUNKS];
	uint32_t bins[FWD_SORT_BINS];             // counting sort of the pixels by list length
	uint8_t order[TILE_PIXELS];               // thread -> natural owner of the pixel it walks
};

struct PixelFwd {
	float T = 1.0f;
	float C0 = 0.f, C1 = 0.f, C2 = 0.f;
	float N0 = 0.f, N1 = 0.f, N2 = 0.f;
	float D = 0.f, M1 = 0.f, M2 = 0.f, D2 = 0.f, distortion = 0.f;
	float median_depth = 0.f;
	uint32_t median_contributor = 0;   // 0 = none (the reference stores float(-1) -> 0u)
	uint32_t last_contributor = 0;
};

// 32x32 bit-matrix transpose across a warp: lane r holds row r (bit c = column c); on return lane c holds
// column c (bit r = row r).  Five block-swap steps of one shuffle each.
__device__ __forceinline__ uint32_t transpose32(uint32_t x, int lane)
{
#pragma unroll
	for (int d = 16; d >= 1; d >>= 1) {
		const uint32_t m = d == 16 ? 0x0000ffffu : d == 8 ? 0x00ff00ffu : d == 4 ? 0x0f0f0f0fu : d == 2 ? 0x33333333u : 0x55555555u;
		const uint32_t y = __shfl_xor_sync(0xffffffffu, x, d);
		const bool up = lane & d;
		const uint32_t keep = up ? ~m : m;
		const uint32_t moved = up ? (y >> d) : (y << d);
		x = (x & keep) | (moved & ~keep);
	}
	return x;
}

// The order-dependent half of the reference's loop body (CR/forward.cu:388-441) for one accepted pair:
// termination test, median depth, SA confidence blend or 2DGS distortion, the weighted sums.  Sets `done`
// (and adds nothing) when the pair would push the transmittance below 1e-4.
template <bool USE_SA>
__device__ __forceinline__ void blend_pair(const float alpha, float depth, const float4 r3, const float4 r4,
                                           const uint32_t contributor, PixelFwd& px, bool& done)
{
	const float test_T = px.T * (1 - alpha);
	if (test_T < 0.0001f) {
		done = true;
		return;
	}
	const float w = alpha * px.T;
	const bool is_median = px.T > 0.5f;   // the reference compares in double; 0.5 is exact, so the float compare decides the same
	if (is_median) {
		px.median_depth = depth;
		px.median_contributor = contributor;
	}
	if (USE_SA) {
		// While the pixel is still in front of its median the expected depth IS this depth: error = 0, conf =
		// expf(-0) = 1 and the blend returns `depth` unchanged, so the block is skipped (it would send 0 / x
		// through the division's slow path for every such pair).
		if (px.D > 0 && !is_median) {
			const float exp_depth = px.median_depth;
			const float error = (exp_depth - depth) * (exp_depth - depth);
			float exp_std = (px.D2 - 2 * px.D * exp_depth) / (1 - px.T) + exp_depth * exp_depth;
			exp_std = fmaxf(exp_std, 1e-7f);
			const float conf = expf(-error / (4 * exp_std));
			depth = conf * depth + (1 - conf) * exp_depth;
		}
		px.D += depth * w;
		px.D2 += depth * depth * w;
	} else {
		const float A = 1 - px.T;
		const float m = GS2D_FAR / (GS2D_FAR - GS2D_NEAR) * (1 - GS2D_NEAR / depth);
		px.distortion += (m * m * A + px.M2 - 2 * m * px.M1) * w;
		px.D += depth * w;
		px.M1 += m * w;
		px.M2 += m * m * w;
	}
	px.N0 += r3.x * w; px.N1 += r3.y * w; px.N2 += r3.z * w;
	px.C0 += r4.x * w; px.C1 += r4.y * w; px.C2 += r4.z * w;
	px.T = test_T;
	px.last_contributor = contributor;
}

__device__ __forceinline__ uint32_t bits_below(int k) { return k <= 0 ? 0u : (k >= 32 ? 0xffffffffu : ((1u << k) - 1u)); }


#ifndef GS2D_FWD_MIN_BLOCKS
#define GS2D_FWD_MIN_BLOCKS 4
#endif
template <bool USE_SA>
__global__ void __launch_bounds__(256, GS2D_FWD_MIN_BLOCKS)
composite_forward_kernel(int W, int H, unsigned gx, const uint2* __restrict__ ranges,
                         const float4* __restrict__ instances, const uint32_t* __restrict__ hit_inst,
                         uint32_t* __restrict__ hit_pix, const float* __restrict__ bg, float* __restrict__ final_T,
                         uint32_t* __restrict__ n_contrib, float* __restrict__ median_depth_out,
                         float* __restrict__ depth_std_out, float* __restrict__ out_color,
                         float* __restrict__ out_others, unsigned tile0)
{
	extern __shared__ __align__(128) unsigned char fwd_smem_raw[];
	FwdSmem& sm = *reinterpret_cast<FwdSmem*>(fwd_smem_raw);
	uint64_t* s_bar = sm.bar;

	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
	const unsigned tile = blockIdx.x + tile0;      // the launch covers tiles [tile0, tile0 + gridDim.x)
	const unsigned tx = tile % gx, ty = tile / gx;
	// The transposition below hands thread t the hit words of its NATURAL pixel (warp k = 8 x 4 block k, pair.cuh).
	// The walk is done by another thread: pixels are sorted by list length once per tile, so that the 32 lists of a
	// warp are alike (lanes busy: 0.75 natural -> 0.93 sorted, scripts/pair_stats.py).
	const int nat_ptid = pixel_thread(thread_lx<1>(tid), thread_ly<1>(tid));   // row-major index of the natural pixel
	const uint2 range = ranges[tile];
	const int n = (int)(range.y - range.x);
	const int nsb = (n + FWD_SUPER - 1) / FWD_SUPER;

	if (tid == 0) {
		for (int s = 0; s < FWD_CHUNKS; s++) mbar_init(&s_bar[s], 1);
		mbar_fence_init();
	}
	__syncthreads();
	const float4* tile_rec = instances + 5 * (size_t)range.x;
	const uint32_t* tile_mask = hit_inst + 8 * (size_t)range.x;
	uint32_t* pix_words = hit_pix + ((size_t)(range.x >> 5) + HIT_PIX_PAD * (size_t)tile) * 2 * TILE_PIXELS;
	const float4* recs = &sm.rec[0][0];

	PixelFwd px;
	int o = tid;                   // natural owner of the pixel this thread walks (sorted at the first super-batch)
	int ptid = nat_ptid;
	unsigned pix_x = tx * TILE + thread_lx<1>(tid), pix_y = ty * TILE + thread_ly<1>(tid);
	bool inside = pix_x < (unsigned)W && pix_y < (unsigned)H;
	float pxf = (float)pix_x, pyf = (float)pix_y;
	bool done = !inside;

	for (int sb = 0; sb < nsb; sb++) {
		if (sb > 0) {
			// the previous super-batch is consumed by everyone -> its memory takes the next one; CTA-wide exit vote
			const int num_done = __syncthreads_count(done);
			if (num_done == TILE_PIXELS) break;
		}
		const int first = sb * FWD_SUPER;                       // first instance of the super-batch (in the tile)
		const int cnt_sb = min(n - first, FWD_SUPER);
		const int nch = (cnt_sb + FWD_BATCH - 1) / FWD_BATCH;   // chunks in flight: all requested at once
		if (tid == 0) {
			for (int c = 0; c < nch; c++) {
				const uint32_t cnt = (uint32_t)min(cnt_sb - c * FWD_BATCH, FWD_BATCH);
				const size_t at = (size_t)first + (size_t)c * FWD_BATCH;
				mbar_arrive_expect_tx(&s_bar[c], cnt * (80u + 32u));
				tma_load_1d(&sm.rec[c][0], tile_rec + 5 * at, cnt * 80u, &s_bar[c]);
				tma_load_1d(&sm.mask[c][0], tile_mask + 8 * at, cnt * 32u, &s_bar[c]);
			}
		}
		const uint32_t parity = (uint32_t)sb & 1u;              // every chunk barrier completes one phase per super-batch
		const int nw = nch * (FWD_BATCH / 32);                  // hit words per pixel in this super-batch
		// Transpose in registers, chunk by chunk as the chunks land: this warp's byte-gathered word of instance j
		// holds its 32 pixels; a 32x32 bit transpose turns the words of 32 instances into one hit word per pixel
		// (bit j <-> instance 32 g + j).  The raw words are the footprint (a gradient-pair slot each in the backward
		// pass) and go out at once; the copy in shared memory is what the walk consumes and edits.
		for (int c = 0; c < nch; c++) {
			mbar_wait(&s_bar[c], parity);
			const int cnt = min(cnt_sb - c * FWD_BATCH, FWD_BATCH);
			const uint32_t* masks = &sm.mask[c][0];
#pragma unroll
			for (int g = 0; g < FWD_BATCH / 32; g++) {
				const int j = g * 32 + lane;
				const uint32_t mine = j < cnt ? warp_word<1>(masks + j * 8, warp) : 0u;
				const uint32_t raw = __any_sync(0xffffffffu, mine != 0u) ? transpose32(mine, lane) : 0u;
				const int w = c * (FWD_BATCH / 32) + g;
				sm.hits[w][tid] = raw;
				pix_words[(size_t)((sb * FWD_WORDS + w) * 2 + 1) * TILE_PIXELS + nat_ptid] = raw;
			}
		}

		if (sb == 0) {
			// counting sort of the 256 pixels by the length of their list, longest first
			int c = 0;
			for (int w = 0; w < nw; w++) c += __popc(sm.hits[w][tid]);
			c = (FWD_SORT_BINS - 1) - min(c, FWD_SORT_BINS - 1);
			if (tid < FWD_SORT_BINS) sm.bins[tid] = 0u;
			__syncthreads();
			const uint32_t rank = atomicAdd(&sm.bins[c], 1u);
			__syncthreads();
			if (warp == 0) {
				const uint32_t a = sm.bins[2 * lane], b = sm.bins[2 * lane + 1];
				uint32_t incl = a + b;
#pragma unroll
				for (int d = 1; d < 32; d <<= 1) {
					const uint32_t t = __shfl_up_sync(0xffffffffu, incl, d);
					if (lane >= d) incl += t;
				}
				sm.bins[2 * lane] = incl - a - b;
				sm.bins[2 * lane + 1] = incl - b;
			}
			__syncthreads();
			sm.order[sm.bins[c] + rank] = (uint8_t)tid;
			__syncthreads();
			o = sm.order[tid];
			const int lx = thread_lx<1>(o), ly = thread_ly<1>(o);
			ptid = pixel_thread(lx, ly);
			pix_x = tx * TILE + lx;
			pix_y = ty * TILE + ly;
			inside = pix_x < (unsigned)W && pix_y < (unsigned)H;
			pxf = (float)pix_x;
			pyf = (float)pix_y;
			done = !inside;
		} else {
			__syncthreads();           // the hit words were written by the pixels' natural owners
		}

		// Every pixel walks its own hits of the WHOLE super-batch front to back, in one loop: a warp iterates as
		// often as its longest pixel list.
		const bool done_before = done;
		int stop = FWD_SUPER;          // instance (in the super-batch) at which this pixel terminated
		{
			int g = 0;
			uint32_t m = done ? 0u : sm.hits[0][o];
			for (;;) {
				// advance to the next non-empty word (finished pixels sit at the last word with m = 0)
				while (m == 0u && g < nw - 1) {
					g++;
					m = done ? 0u : sm.hits[g][o];
				}
				if (!__any_sync(0xffffffffu, m != 0u)) break;
				if (m != 0u) {
					const int jl = __ffs(m) - 1;
					m &= m - 1;
					const int j = g * 32 + jl;
					const float4* rec = recs + j * 5;
					PairFront f;
					if (pair_front(rec[0], rec[1], rec[2], pxf, pyf, f)) {
						blend_pair<USE_SA>(f.alpha, f.depth, rec[3], rec[4], (uint32_t)(first + j + 1), px, done);
						if (done) {   // T < 1e-4: nothing behind this pair counts; park the lane at the last word
							stop = j;
							m = 0u;
							g = nw - 1;
						}
					} else {
						sm.hits[g][o] &= ~(1u << jl);   // in the footprint, not a contributor (1.4 % of the pairs)
					}
				}
			}
		}
		// for the backward pass (coalesced, one word per pixel thread): the pairs that really contributed
		for (int w = 0; w < nw; w++) {
			const uint32_t used = done_before ? 0u : (sm.hits[w][o] & bits_below(stop - 32 * w));
			pix_words[(size_t)((sb * FWD_WORDS + w) * 2 + 0) * TILE_PIXELS + ptid] = used;
		}
	}

	if (inside) {
		const size_t HW = (size_t)H * W;
		const size_t pid = (size_t)W * pix_y + pix_x;
		const float T = px.T;
		final_T[pid] = T;
		final_T[pid + HW] = px.M1;
		final_T[pid + 2 * HW] = px.M2;
		n_contrib[pid] = px.last_contributor;
		n_contrib[pid + HW] = px.median_contributor;
		out_color[pid] = px.C0 + T * bg[0];
		out_color[pid + HW] = px.C1 + T * bg[1];
		out_color[pid + 2 * HW] = px.C2 + T * bg[2];
		const float md = px.median_depth;
		out_others[pid + 0 * HW] = px.D;
		out_others[pid + 1 * HW] = 1 - T;
		out_others[pid + 2 * HW] = px.N0;
		out_others[pid + 3 * HW] = px.N1;
		out_others[pid + 4 * HW] = px.N2;
		out_others[pid + 5 * HW] = md;
		median_depth_out[pid] = md;
		depth_std_out[pid] = px.D2 - 2 * md * px.D + md * md * (1 - T);
		if (USE_SA)
			out_others[pid + 6 * HW] = px.D2 - 2 * md * px.D + (1 - T) * md * md;
		else
			out_others[pid + 6 * HW] = px.distortion;
	}
}

template <bool USE_SA>
int launch_variant(const gs2d_settings& s, const gs2d_inputs& in, const ImageState& im, const float* instances,
                   const uint32_t* hit_inst, uint32_t* hit_pix, float* out_color, float* out_others, unsigned gx,
                   unsigned tile0, unsigned ntiles, cudaStream_t st)
{
	// the opt-in to > 48 KB of dynamic shared memory is per device: made once for every device this variant runs on
	static std::atomic<unsigned long long> opted_in{0};
	int dev = 0;
	GS2D_CUDA_OK(cudaGetDevice(&dev));
	if (dev >= 64 || !((opted_in.load(std::memory_order_relaxed) >> dev) & 1ull)) {
		GS2D_CUDA_OK(cudaFuncSetAttribute(composite_forward_kernel<USE_SA>,
		                                  cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(FwdSmem)));
		if (dev < 64) opted_in.fetch_or(1ull << dev, std::memory_order_relaxed);
	}
	composite_forward_kernel<USE_SA><<<ntiles, 256, sizeof(FwdSmem), st>>>(
		s.image_width, s.image_height, gx, (const uint2*)im.ranges, (const float4*)instances, hit_inst, hit_pix,
		in.background, im.final_T, im.n_contrib, im.median_depth, im.depth_std, out_color, out_others, tile0);
	return GS2D_OK;
}

}  // namespace

int launch_composite_forward(const gs2d_settings& s, const gs2d_inputs& in, const ImageState& im,
                             const float* instances, const uint32_t* hit_inst, uint32_t* hit_pix, float* out_color,
                             float* out_others, cudaStream_t st, int tile0, int ntiles)
{
	const int W = s.image_width, H = s.image_height;
	const unsigned gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
	if (gx * gy == 0) return GS2D_OK;
	if (ntiles < 0) { tile0 = 0; ntiles = (int)(gx * gy); }     // the whole frame
	if (ntiles == 0) return GS2D_OK;
	ProfScope ps(K_COMPOSITE_FWD, st);
	const int rc = s.use_sa ? launch_variant<true>(s, in, im, instances, hit_inst, hit_pix, out_color, out_others, gx, (unsigned)tile0, (unsigned)ntiles, st)
	                        : launch_variant<false>(s, in, im, instances, hit_inst, hit_pix, out_color, out_others, gx, (unsigned)tile0, (unsigned)ntiles, st);
	if (rc != GS2D_OK) return rc;
	count_launch();
	GS2D_CUDA_OK(cudaGetLastError());
	return GS2D_OK;
}

}  // namespace gs2d
