// composite_fwd.cu -- per-tile front-to-back compositing (forward).
//
// Replaces FORWARD::render / renderCUDA<3> (CR/forward.cu:258-467): ray-splat intersection,
// low-pass filter, alpha, early ray termination, surface-aware (SA) depth or 2DGS distortion,
// normals, median depth.  The per-(pixel, instance) arithmetic keeps the reference's operation
// order, so a pixel's result is the same sequence of fp32 operations (images are bit-identical to
// the reference build).  What changes is which pairs are visited and how the data is fed:
//   * one CTA per 16x16 tile, one thread per pixel, warp k owns tile rows 2k and 2k+1;
//   * the tile's instance records (80 B each, sorted, contiguous in HBM) and their footprint masks
//     (32 B each, footprint.cu) are staged by TMA bulk copies, 128 instances per batch, into a
//     three-stage shared-memory ring guarded by mbarriers; the first three batches are requested at once and every
//     thread waits on the mbarrier itself, so lists of up to 384 instances need no CTA barrier at all;
//   * per batch the instance-major footprint masks are transposed into per-PIXEL hit words with warp
//     ballots (registers only: word k of a mask is exactly warp k's 32 pixels), and every pixel thread
//     then walks only ITS OWN set bits in depth order.  A pixel visits ~17 instances per tile instead of ~250; lanes of a
//     warp run independent lists (the trip count of a warp is the longest list among its 32 pixels);
//   * a pixel stops at T < 1e-4 like the reference; the CTA stops staging when all 256 have.
// Bound by FP32 issue + MUFU on the visited pairs; HBM traffic is 112 B per instance + 68 B per pixel.
#include "common.cuh"
#include "tma.cuh"
#include "pair.cuh"

namespace gs2d {

namespace {

#ifndef GS2D_FWD_BATCH
#define GS2D_FWD_BATCH 128
#endif
constexpr int FWD_BATCH = GS2D_FWD_BATCH;   // instances per staged batch (a multiple of 32)
constexpr int FWD_GROUPS = FWD_BATCH / 32;  // hit words per pixel per batch
#ifndef GS2D_FWD_STAGES
#define GS2D_FWD_STAGES 3
#endif
constexpr int FWD_STAGES = GS2D_FWD_STAGES;
static_assert(FWD_BATCH % 32 == 0 && FWD_BATCH / 32 <= HIT_PIX_PAD, "hit_pix padding");

struct FwdSmem {
	float4 rec[FWD_STAGES][FWD_BATCH * 5];
	uint32_t mask[FWD_STAGES][FWD_BATCH * 8];
	uint32_t hits[FWD_GROUPS][TILE_PIXELS];   // per-pixel hit words of the current batch
	uint32_t used[FWD_GROUPS][TILE_PIXELS];   // the pairs that really contributed
	uint64_t bar[FWD_STAGES];
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

template <bool USE_SA>
__device__ __forceinline__ bool eval_pair(const float4* __restrict__ rec, float pxf, float pyf,
                                          uint32_t contributor, PixelFwd& px, bool& done)
{
	PairFront f;
	if (!pair_front(rec[0], rec[1], rec[2], pxf, pyf, f)) return false;
	const float alpha = f.alpha;
	float depth = f.depth;
	const float test_T = px.T * (1 - alpha);
	if (test_T < 0.0001f) {
		done = true;
		return false;
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
			float exp_std = (px.D2 - 2 * px.D * exp_depth) / (1 - px.T) + exp_depth * exp_depth;
			exp_std = fmaxf(exp_std, 1e-7f);
			const float error = (exp_depth - depth) * (exp_depth - depth);
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
	const float4 r3 = rec[3], r4 = rec[4];
	px.N0 += r3.x * w; px.N1 += r3.y * w; px.N2 += r3.z * w;
	px.C0 += r4.x * w; px.C1 += r4.y * w; px.C2 += r4.z * w;
	px.T = test_T;
	px.last_contributor = contributor;
	return true;
}

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
                         float* __restrict__ out_others)
{
	extern __shared__ __align__(128) unsigned char fwd_smem_raw[];
	FwdSmem& sm = *reinterpret_cast<FwdSmem*>(fwd_smem_raw);
	float4 (*s_rec)[FWD_BATCH * 5] = sm.rec;
	uint32_t (*s_mask)[FWD_BATCH * 8] = sm.mask;
	uint64_t* s_bar = sm.bar;

	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
	const unsigned tile = blockIdx.x;
	const unsigned tx = tile % gx, ty = tile / gx;
	const int lx = thread_lx<GS2D_FWD_BLOCK>(tid), ly = thread_ly<GS2D_FWD_BLOCK>(tid);   // thread <-> tile pixel: pair.cuh
	const int ptid = pixel_thread(lx, ly);               // row-major pixel index: where this pixel's words go in hit_pix
	const unsigned pix_x = tx * TILE + lx;
	const unsigned pix_y = ty * TILE + ly;
	const bool inside = pix_x < (unsigned)W && pix_y < (unsigned)H;
	const float pxf = (float)pix_x, pyf = (float)pix_y;

	const uint2 range = ranges[tile];
	const int n = (int)(range.y - range.x);
	const int nb = (n + FWD_BATCH - 1) / FWD_BATCH;

	if (tid == 0) {
		for (int s = 0; s < FWD_STAGES; s++) mbar_init(&s_bar[s], 1);
		mbar_fence_init();
	}
	__syncthreads();
	const float4* tile_rec = instances + 5 * (size_t)range.x;
	const uint32_t* tile_mask = hit_inst + 8 * (size_t)range.x;
	auto issue = [&](int b, int stage) {
		const uint32_t cnt = (uint32_t)min(n - b * FWD_BATCH, FWD_BATCH);
		mbar_arrive_expect_tx(&s_bar[stage], cnt * (80u + 32u));
		tma_load_1d(&s_rec[stage][0], tile_rec + 5 * (size_t)b * FWD_BATCH, cnt * 80u, &s_bar[stage]);
		tma_load_1d(&s_mask[stage][0], tile_mask + 8 * (size_t)b * FWD_BATCH, cnt * 32u, &s_bar[stage]);
	};
	// The first FWD_STAGES batches are requested at once: tiles of up to FWD_STAGES * FWD_BATCH instances (most of
	// them) run without a single CTA barrier -- every thread waits on the batch's mbarrier itself, so the 8 warps of
	// a tile drift apart freely instead of waiting for the slowest one at every batch (the barrier was the top stall
	// of this kernel).  Longer lists reuse the stages: one barrier per batch, which is also the CTA-wide exit vote.
	if (tid == 0)
		for (int b = 0; b < min(nb, FWD_STAGES); b++) issue(b, b);

	uint32_t* pix_words = hit_pix + ((size_t)(range.x >> 5) + HIT_PIX_PAD * (size_t)tile) * 2 * TILE_PIXELS;

	PixelFwd px;
	bool done = !inside;
	int first_unwaited = 0;       // batches below this index have been waited for by every thread

	for (int b = 0; b < nb; b++) {
		const int stage = b % FWD_STAGES;
		const uint32_t parity = (uint32_t)(b / FWD_STAGES) & 1u;
		if (b >= 1 && b + FWD_STAGES - 1 < nb) {
			// batch b-1 is consumed by everyone -> its stage takes batch b + FWD_STAGES - 1; CTA-wide exit vote
			const int num_done = __syncthreads_count(done);
			if (num_done == TILE_PIXELS) break;
			if (tid == 0) issue(b + FWD_STAGES - 1, (b + FWD_STAGES - 1) % FWD_STAGES);
		}
		mbar_wait(&s_bar[stage], parity);
		first_unwaited = b + 1;

		const int cnt = min(n - b * FWD_BATCH, FWD_BATCH);
		const float4* recs = &s_rec[stage][0];
		const uint32_t* masks = &s_mask[stage][0];
		// Transpose in registers: word `warp` of instance j holds this warp's 32 pixels; a 32x32 bit transpose turns
		// the words of 32 instances into one hit word per pixel (bit j <-> instance g*32 + j).
#pragma unroll
		for (int g = 0; g < FWD_GROUPS; g++) {
			const int j = g * 32 + lane;
			const uint32_t mine = j < cnt ? warp_word<GS2D_FWD_BLOCK>(masks + j * 8, warp) : 0u;
			sm.hits[g][tid] = __any_sync(0xffffffffu, mine != 0u) ? transpose32(mine, lane) : 0u;
			sm.used[g][tid] = 0u;
		}

		// Every pixel walks its own hits of the WHOLE batch front to back, in one loop: a warp iterates as often as
		// its longest pixel list (lists of 128 instances balance better between the lanes than lists of 32).
		{
			int g = done ? FWD_GROUPS - 1 : 0;
			uint32_t m = done ? 0u : sm.hits[0][tid], used = 0u;
			for (;;) {
				// advance to the next non-empty word (finished pixels sit at the last word with m = 0)
				while (m == 0u && g < FWD_GROUPS - 1) {
					sm.used[g][tid] = used;
					used = 0u;
					g++;
					m = sm.hits[g][tid];
				}
				if (!__any_sync(0xffffffffu, m != 0u)) break;
				if (m != 0u) {
					const int jl = __ffs(m) - 1;
					m &= m - 1;
					const int j = g * 32 + jl;
					if (eval_pair<USE_SA>(recs + j * 5, pxf, pyf, (uint32_t)(b * FWD_BATCH + j + 1), px, done))
						used |= 1u << jl;
					if (done) {   // T < 1e-4: nothing behind this pair counts; park the lane at the last word
						m = 0u;
						if (g < FWD_GROUPS - 1) {
							sm.used[g][tid] = used;
							used = 0u;
							g = FWD_GROUPS - 1;
						}
					}
				}
			}
			sm.used[g][tid] = used;
		}
		// for the backward pass (coalesced, one word per pixel thread): the pairs that really contributed,
		// and the pairs of the footprint masks (they own a gradient-pair slot each)
#pragma unroll
		for (int g = 0; g < FWD_GROUPS; g++) {
			pix_words[(size_t)((b * FWD_GROUPS + g) * 2 + 0) * TILE_PIXELS + ptid] = sm.used[g][tid];
			pix_words[(size_t)((b * FWD_GROUPS + g) * 2 + 1) * TILE_PIXELS + ptid] = sm.hits[g][tid];
		}
	}

	// early exit: no bulk copy may still be in flight when the CTA retires (requested but never consumed batches)
	if (tid == 0)
		for (int bb = first_unwaited; bb < min(nb, first_unwaited + FWD_STAGES - 1); bb++)
			mbar_wait(&s_bar[bb % FWD_STAGES], (uint32_t)(bb / FWD_STAGES) & 1u);

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

}  // namespace

int launch_composite_forward(const gs2d_settings& s, const gs2d_inputs& in, const ImageState& im,
                             const float* instances, const uint32_t* hit_inst, uint32_t* hit_pix, float* out_color,
                             float* out_others, cudaStream_t st)
{
	const int W = s.image_width, H = s.image_height;
	const unsigned gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
	if (gx * gy == 0) return GS2D_OK;
	static bool attr_set = false;
	if (!attr_set) {
		GS2D_CUDA_OK(cudaFuncSetAttribute(composite_forward_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(FwdSmem)));
		GS2D_CUDA_OK(cudaFuncSetAttribute(composite_forward_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(FwdSmem)));
		attr_set = true;
	}
	ProfScope ps(K_COMPOSITE_FWD, st);
	if (s.use_sa)
		composite_forward_kernel<true><<<gx * gy, 256, sizeof(FwdSmem), st>>>(
			W, H, gx, (const uint2*)im.ranges, (const float4*)instances, hit_inst, hit_pix, in.background, im.final_T,
			im.n_contrib, im.median_depth, im.depth_std, out_color, out_others);
	else
		composite_forward_kernel<false><<<gx * gy, 256, sizeof(FwdSmem), st>>>(
			W, H, gx, (const uint2*)im.ranges, (const float4*)instances, hit_inst, hit_pix, in.background, im.final_T,
			im.n_contrib, im.median_depth, im.depth_std, out_color, out_others);
	count_launch();
	GS2D_CUDA_OK(cudaGetLastError());
	return GS2D_OK;
}

}  // namespace gs2d
