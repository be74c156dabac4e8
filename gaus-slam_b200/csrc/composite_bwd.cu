// composite_bwd.cu -- per-tile back-to-front re-traversal (backward of the compositing).
//
// Replaces BACKWARD::render / renderCUDA<3> (CR/backward.cu:143-463).  Same per-(pixel, instance)
// gradient algebra (including the reference's deliberate non-analytic parts: SA confidence from the
// *final* per-pixel statistics, opacity gradient through the clamped alpha), but the work is split by
// what it depends on, so that every lane does useful work and nothing is reduced with atomics:
//
//   phase A (per PIXEL, sequential back to front; one CTA per tile):
//       the order-dependent part only -- transmittance T = T / (1 - alpha), the suffix recurrences for
//       colour / depth / alpha / normal / dL_dT (CR/backward.cu:323-410).  Every pixel walks ITS OWN list of
//       the pairs that contributed in the forward pass (recorded there, one word per pixel and 32
//       instances), i.e. ~18 instances per tile instead of ~300 and no skip tests, and leaves three scalars
//       per pair -- dL/dalpha, dL/dz, w = alpha * T -- plus its pixel id in the pair's 16-byte slot.  Slots
//       live in global memory (L2), laid out per instance: slot = first slot of the instance (exclusive
//       scan of the footprint-mask popcounts, computed in the forward pass) + rank of the pixel inside the
//       instance's mask.  Pairs of the mask that did not contribute get an "empty" marker.
//       Up to BWD_SUPER instances of the tile (all of it for the BASELINE views) are resident at once, and the
//       walk of a pixel is done by a thread chosen by a counting sort of the 256 pixels by list length: the 32
//       lists a warp walks in lockstep are alike (27 of 32 lanes busy per instruction, 16 before).
//   phase B (per INSTANCE, same CTA, four lanes per instance, instances in the order of their slot counts):
//       the geometric chain rule dL/d{Tu,Tv,Tw}, dL/dmean2D, dL/dopacity, dL/dnormal, dL/dcolour
//       (CR/backward.cu:343, :396, :412-460) from the instance's slots (one coalesced read) + its record,
//       reduce-scattered over the four lanes with 15 shuffles and written as one 16-byte + one 4-byte store per
//       lane into the instance's own gradient row: the reference issues 16 global atomicAdd per contributing
//       (pixel, instance) pair; here a row is written exactly once, so no atomics and no zero-fill.
//
// Records + masks of a tile arrive by TMA bulk copies in chunks of 128 instances, all chunks of a super-batch
// requested at once, one mbarrier each.  Rows are indexed by the instance's position in surfel-major (unsorted)
// order, so preprocess_bwd.cu reads each surfel's rows contiguously.  What bounds the kernel now is the L1 /
// shared-memory data pipe (an 80-byte record from a random row per pair), not the schedulers.
#include <atomic>

#include "common.cuh"
#include "tma.cuh"
#include "pair.cuh"

namespace gs2d {

namespace {

#ifndef GS2D_BWD_DISCARD
#define GS2D_BWD_DISCARD 1
#endif
#ifndef GS2D_BWD_PREFETCH
#define GS2D_BWD_PREFETCH 2
#endif
#ifndef GS2D_BWD_SLOT_CG
#define GS2D_BWD_SLOT_CG 0
#endif
#if GS2D_BWD_SLOT_CG
#define SLOT_LD(p) __ldcg(p)
#else
#define SLOT_LD(p) (*(p))
#endif
#ifndef GS2D_BWD_PREFETCH_PIX
#define GS2D_BWD_PREFETCH_PIX 1
#endif
#ifndef GS2D_BWD_ROW128
#define GS2D_BWD_ROW128 0
#endif
constexpr int BWD_BATCH = 128;                        // instances per TMA chunk
// Instances resident at once ("super-batch").  352 keeps the kernel's shared memory under 56 KB (four CTAs per SM);
// the tiles of the BASELINE mapping views hold 307 instances on average (p99 360): a denser tile takes a second round.
#ifndef GS2D_BWD_SUPER
#define GS2D_BWD_SUPER 352
#endif
constexpr int BWD_SUPER = GS2D_BWD_SUPER;
static_assert(BWD_SUPER % 32 == 0, "super-batches start at a word boundary of the per-pixel words");
constexpr int BWD_CHUNKS = (BWD_SUPER + BWD_BATCH - 1) / BWD_BATCH;
constexpr int BWD_WORDS = BWD_SUPER / 32;             // contribution words per pixel per super-batch
constexpr int BWD_PIX_BINS = 64;                      // pixel list lengths 0 .. 62, 63 = longer
constexpr int BWD_INST_BINS = 32;                     // phase-B iterations of an instance 0 .. 30, 31 = more
static_assert(BWD_BATCH / 32 <= HIT_PIX_PAD, "hit_pix padding");
constexpr int BWD_PIX_SLOTS = TILE_PIXELS + 3 * TILE;
__device__ __forceinline__ int pix_slot(int p) { return p + 3 * (p >> 4); }

__device__ __forceinline__ void prefetch_l1(const void* p)
{
#ifdef __CUDA_ARCH__
	asm volatile("prefetch.global.L1 [%0];" ::"l"(p));
#else
	(void)p;
#endif
}

struct PixelConst {
	float T_final, final_A, final_D, final_D2;
	float dpix0, dpix1, dpix2;
	float dL_ddepth, dL_daccum, dL_dreg, dL_dmedian;
	float dn0, dn1, dn2;
	float mm, inv_4std, bg_dot;      // inv_4std = 1 / (4 max(mstd / (1 - T_final), 1e-7)), CR/backward.cu:349
	uint32_t median_contributor;
};

// Back-to-front state of a pixel.  The reference keeps, for each of its 8 blended channels x (rgb, depth,
// alpha, normal) the running "what lies behind" value acc_x and adds (x - acc_x) * dL/dx to dL/dalpha
// (CR/backward.cu:336-345, :384-397), plus the same construction for the regulariser weight (last_dL_dT,
// :370-376).  All nine recurrences are linear with the same coefficients, acc <- alpha * x + (1 - alpha) * acc,
// so their weighted sum R = sum_x acc_x * dL/dx + last_dL_dT obeys the same recurrence with
// Y = sum_x x * dL/dx + dL_dweight in the place of x: one scalar instead of 17 (same value up to fp32 summation
// order; the gradient bar is 1e-3 relative).
struct PixelBwd {
	float T;
	float behind = 0.f;   // R
};

struct BwdSmem {
	float4 rec[BWD_SUPER * 5];                // contiguous: instance i of the super-batch is rec[5 * i]
	uint32_t mask[BWD_SUPER * 8];
	union {
		uint32_t used[BWD_WORDS][TILE_PIXELS];    // phase A: per-pixel contribution words, column = row-major pixel id
		struct {
			// phase B: the pixel's image gradients, dL_dpix (3) | dL_dnormal.x and dL_dnormal.yz, at index pix_slot(pixel):
			// two vector loads per pair, and the padding keeps the pixels of one instance (neighbours in x and y, read by
			// the lanes of one quarter-warp) in different banks
			float4 pixa[BWD_PIX_SLOTS];
			float2 pixb[BWD_PIX_SLOTS];
			uint16_t ord[BWD_SUPER];              // phase B: instances of the super-batch, longest slot list first
		} b;
	};
	uint2 pref[BWD_SUPER];                    // 8 packed bytes: hits of the instance in mask words 0..k-1
	uint32_t base[BWD_SUPER + 1];             // first pair slot of every instance of the super-batch (global index)
	uint32_t bins[BWD_PIX_BINS];
	uint8_t order[TILE_PIXELS];               // phase A: thread -> row-major id of the pixel it walks
	uint32_t max_last[8];
	uint64_t bar[BWD_CHUNKS];
};

// Phase A for one (pixel, instance) pair: CR/backward.cu:283-410 minus the geometric chain.
// The forward pass recorded which pairs contributed (hit_pix), so there are no tests here.
template <bool USE_SA>
__device__ __forceinline__ void pixel_step(const float4* __restrict__ rec, float pxf, float pyf, uint32_t q,
                                           const PixelConst& pc, PixelBwd& st, float& dL_dalpha_out,
                                           float& dL_dz_out, float& w_out)
{
	PairFront f;
	float inv_pz;
	pair_geom_fast(rec[0], rec[1], rec[2], pxf, pyf, f, inv_pz);
	const float alpha = f.alpha;
	float c_d = f.depth;
	const float4 r3 = rec[3], r4 = rec[4];
	const float inv_1ma = __fdividef(1.0f, 1.f - alpha);
	st.T = st.T * inv_1ma;
	const float T = st.T;
	const float w = alpha * T;

	float conf = 1;
	if (USE_SA) {
		// confidence from the FINAL per-pixel statistics, treated as a constant (CR/backward.cu:349)
		conf = T < 0.5f ? __expf(-(c_d - pc.mm) * (c_d - pc.mm) * pc.inv_4std) : 1.f;
		c_d = c_d * conf + pc.mm * (1 - conf);
	}

	float dL_dz = 0.0f;
	float dL_dweight = 0;
	if (q + 1 == pc.median_contributor) dL_dz += pc.dL_dmedian;
	if (USE_SA) {
		dL_dweight = ((c_d - pc.mm) * (c_d - pc.mm)) * pc.dL_dreg;
		dL_dz += conf * 2.0f * w * (c_d - pc.mm) * pc.dL_dreg;
	} else {
		const float m_d = GS2D_FAR / (GS2D_FAR - GS2D_NEAR) * (1 - GS2D_NEAR / c_d);
		const float dmd_dd = (GS2D_FAR * GS2D_NEAR) / ((GS2D_FAR - GS2D_NEAR) * c_d * c_d);
		dL_dweight = (pc.final_D2 + m_d * m_d * pc.final_A - 2 * m_d * pc.final_D) * pc.dL_dreg;
		const float dL_dmd = 2.0f * (T * alpha) * (m_d * pc.final_A - pc.final_D) * pc.dL_dreg;
		dL_dz += dL_dmd * dmd_dd;
	}
	// Y: this instance's blended values weighted by the pixel's gradients (+ the regulariser weight)
	float Y = dL_dweight + pc.dL_daccum;
	Y += r4.x * pc.dpix0; Y += r4.y * pc.dpix1; Y += r4.z * pc.dpix2;
	Y += r3.x * pc.dn0;   Y += r3.y * pc.dn1;   Y += r3.z * pc.dn2;
	Y += c_d * pc.dL_ddepth;
	const float ahead = Y - st.behind;
	st.behind += alpha * ahead;
	float dL_dalpha = ahead * T;
	dL_dalpha += (-pc.T_final * inv_1ma) * pc.bg_dot;
	dL_dz += conf * alpha * T * pc.dL_ddepth;

	dL_dalpha_out = dL_dalpha;
	dL_dz_out = dL_dz;
	w_out = w;
}

// Lanes per instance in phase B (8 or 4).  With 4 lanes the 20 row words are reduce-scattered to 5 per lane; GP maps row
// word w to its position in the accumulator array so that lane L ends up with words 4L .. 4L+3 (one 16-byte store) and
// word 16 + L.  With 8 lanes GP is the identity.
#ifndef GS2D_BWD_GRP
#define GS2D_BWD_GRP 4
#endif
constexpr int BWD_GRP = GS2D_BWD_GRP;
static_assert(BWD_GRP == 8 || BWD_GRP == 4, "phase B: a quarter or an eighth of a warp per instance");
__host__ __device__ constexpr int GP(int w) { return BWD_GRP == 8 ? w : (w < 16 ? 5 * (w / 4) + (w % 4) : 5 * (w - 16) + 4); }

// Phase B for one (instance, pixel) pair: the geometric chain (CR/backward.cu:343, :396, :412-460).
// A = dL/dalpha, dz = dL/dz, w = alpha * T from phase A; adds the 18 per-instance terms into g.
__device__ __forceinline__ void instance_step(const float4 r0, const float4 r1, const float4 r2, float pxf, float pyf,
                                              float A, float dz, float w, const float4 pa, const float2 pb,
                                              float (&g)[24])
{
	PairFront f;
	float inv_pz;
	pair_geom_fast(r0, r1, r2, pxf, pyf, f, inv_pz);
	const float3 Tw = make_float3(r2.x, r2.y, r2.z);
	const float opa = r2.w, G = f.G;
	g[GP(GR_COL + 0)] += w * pa.x;
	g[GP(GR_COL + 1)] += w * pa.y;
	g[GP(GR_COL + 2)] += w * pa.z;
	g[GP(GR_NRM + 0)] += w * pa.w;
	g[GP(GR_NRM + 1)] += w * pb.x;
	g[GP(GR_NRM + 2)] += w * pb.y;
	const float dL_dG = opa * A;
	if (f.rho3d <= f.rho2d) {
		const float dL_dsx = dL_dG * -G * f.s.x + dz * Tw.x;
		const float dL_dsy = dL_dG * -G * f.s.y + dz * Tw.y;
		const float dsx_pz = dL_dsx * inv_pz;
		const float dsy_pz = dL_dsy * inv_pz;
		const float3 dL_dp = make_float3(dsx_pz, dsy_pz, -(dsx_pz * f.s.x + dsy_pz * f.s.y));
		const float3 k = f.k, l = f.l;
		const float3 dL_dk = make_float3(l.y * dL_dp.z - l.z * dL_dp.y, l.z * dL_dp.x - l.x * dL_dp.z,
		                                 l.x * dL_dp.y - l.y * dL_dp.x);
		const float3 dL_dl = make_float3(dL_dp.y * k.z - dL_dp.z * k.y, dL_dp.z * k.x - dL_dp.x * k.z,
		                                 dL_dp.x * k.y - dL_dp.y * k.x);
		g[GP(0)] += -dL_dk.x; g[GP(1)] += -dL_dk.y; g[GP(2)] += -dL_dk.z;
		g[GP(3)] += -dL_dl.x; g[GP(4)] += -dL_dl.y; g[GP(5)] += -dL_dl.z;
		g[GP(6)] += pxf * dL_dk.x + pyf * dL_dl.x + dz * f.s.x;
		g[GP(7)] += pxf * dL_dk.y + pyf * dL_dl.y + dz * f.s.y;
		g[GP(8)] += pxf * dL_dk.z + pyf * dL_dl.z + dz * 1.0f;
	} else {
		g[GP(8)] += dz;
		g[GP(GR_M2D + 0)] += dL_dG * (-G * GS2D_FILTER_INV_SQUARE * f.d.x);
		g[GP(GR_M2D + 1)] += dL_dG * (-G * GS2D_FILTER_INV_SQUARE * f.d.y);
	}
	g[GP(GR_OPA)] += G * A;
}

// Sum g[0..17] over the 8 lanes of each quarter-warp (reduce-scatter: 12 + 6 + 3 shuffles for four instances
// at once).  g is padded to 24 values; on return lane L of a group (L = lane & 7) holds the totals of
// g[3L], g[3L+1], g[3L+2] in g[0..2] (lanes 6, 7 hold padding).
__device__ __forceinline__ void oct_reduce24(float (&g)[24], int lane)
{
	{
		const bool up = lane & 4;
#pragma unroll
		for (int i = 0; i < 12; i++) {
			const float send = up ? g[i] : g[i + 12];
			const float keep = up ? g[i + 12] : g[i];
			g[i] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
		}
	}
	{
		const bool up = lane & 2;
#pragma unroll
		for (int i = 0; i < 6; i++) {
			const float send = up ? g[i] : g[i + 6];
			const float keep = up ? g[i + 6] : g[i];
			g[i] = keep + __shfl_xor_sync(0xffffffffu, send, 2);
		}
	}
	{
		const bool up = lane & 1;
#pragma unroll
		for (int i = 0; i < 3; i++) {
			const float send = up ? g[i] : g[i + 3];
			const float keep = up ? g[i + 3] : g[i];
			g[i] = keep + __shfl_xor_sync(0xffffffffu, send, 1);
		}
	}
}

// Four lanes per instance: sum g[0..19] over the 4 lanes of each group (reduce-scatter: 10 + 5 shuffles for eight
// instances at once); on return lane L of a group (L = lane & 3) holds the totals of positions 5L .. 5L+4 in g[0..4].
__device__ __forceinline__ void quad_reduce20(float (&g)[24], int lane)
{
	{
		const bool up = lane & 2;
#pragma unroll
		for (int i = 0; i < 10; i++) {
			const float send = up ? g[i] : g[i + 10];
			const float keep = up ? g[i + 10] : g[i];
			g[i] = keep + __shfl_xor_sync(0xffffffffu, send, 2);
		}
	}
	{
		const bool up = lane & 1;
#pragma unroll
		for (int i = 0; i < 5; i++) {
			const float send = up ? g[i] : g[i + 5];
			const float keep = up ? g[i + 5] : g[i];
			g[i] = keep + __shfl_xor_sync(0xffffffffu, send, 1);
		}
	}
}

#ifndef GS2D_BWD_MIN_BLOCKS
#define GS2D_BWD_MIN_BLOCKS 4
#endif
template <bool USE_SA>
__global__ void __launch_bounds__(256, GS2D_BWD_MIN_BLOCKS)
composite_backward_pixels_kernel(int W, int H, unsigned gx, const uint2* __restrict__ ranges,
                                 const float4* __restrict__ instances, const uint32_t* __restrict__ hit_inst,
                                 const uint32_t* __restrict__ hit_pix, const uint32_t* __restrict__ pair_incl,
                                 const float* __restrict__ bg, const float* __restrict__ final_T,
                                 const uint32_t* __restrict__ n_contrib, const float* __restrict__ median_depth,
                                 const float* __restrict__ depth_std, const float* __restrict__ dL_dpix,
                                 const float* __restrict__ dL_dothers, float4* __restrict__ slots,
                                 float* __restrict__ grad_rows, uint32_t slot_cap, uint32_t* __restrict__ overflow)
{
	extern __shared__ __align__(128) unsigned char bwd_smem_raw[];
	BwdSmem& sm = *reinterpret_cast<BwdSmem*>(bwd_smem_raw);

	// Thread t is the NATURAL owner of row-major tile pixel t (warp k = tile rows 2k, 2k+1 = footprint-mask word k,
	// lane = bit): it moves that pixel's words.  The back-to-front WALK of a pixel is done by another thread: pixels
	// are sorted by list length once per tile so that the 32 lists a warp walks in lockstep are alike.
	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
	const unsigned tile = blockIdx.x;
	const unsigned tx = tile % gx, ty = tile / gx;
	const size_t HW = (size_t)H * W;

	// The loads that do not depend on the tile's range go out first (the pixel's last contributor, the lines of the
	// per-pixel state), so that they travel beside the range and the overflow check, not behind them.
	const unsigned nat_x = tx * TILE + (tid & 15), nat_y = ty * TILE + (tid >> 4);
	const bool nin = nat_x < (unsigned)W && nat_y < (unsigned)H;
	const size_t npid = (size_t)W * nat_y + nat_x;
	uint32_t wmax = nin ? n_contrib[npid] : 0u;
#if GS2D_BWD_PREFETCH_PIX
	// The per-pixel state is loaded further down by the thread that WALKS the pixel (known only after the sort of
	// the pixels by list length): request the tile's lines now, in natural order, so that those loads hit L1 / L2.
	if (nin) {
		prefetch_l1(final_T + npid); prefetch_l1(final_T + npid + HW); prefetch_l1(final_T + npid + 2 * HW);
		prefetch_l1(n_contrib + npid + HW); prefetch_l1(median_depth + npid); prefetch_l1(depth_std + npid);
		prefetch_l1(dL_dpix + npid); prefetch_l1(dL_dpix + npid + HW); prefetch_l1(dL_dpix + npid + 2 * HW);
#pragma unroll
		for (int c = 0; c < 7; c++) prefetch_l1(dL_dothers + npid + c * HW);
	}
#endif
	const uint2 range = ranges[tile];
	const int n = (int)(range.y - range.x);
	// Memory safety against a wrong pair count from the caller (it sized `slots`): a tile whose slots would end beyond
	// the scratch writes nothing and raises the flag (pair_counters[3] of the binning state) instead.
	if (n > 0 && pair_incl[(size_t)range.x + n - 1] > slot_cap) {
		if (threadIdx.x == 0) *overflow = 1u;
		return;
	}

	// deepest last contributor of the tile: nothing behind it is ever visited
	{
#pragma unroll
		for (int o = 16; o >= 1; o >>= 1) wmax = max(wmax, __shfl_xor_sync(0xffffffffu, wmax, o));
		if (lane == 0) sm.max_last[warp] = wmax;
	}
	if (tid == 0) {
		for (int c = 0; c < BWD_CHUNKS; c++) mbar_init(&sm.bar[c], 1);
		mbar_fence_init();
	}
	__syncthreads();
	uint32_t bmax = 0;
#pragma unroll
	for (int i = 0; i < 8; i++) bmax = max(bmax, sm.max_last[i]);
	const int n_used = min((int)bmax, n);            // instances [0, n_used) can matter
	const int n_done = min(((n_used + BWD_BATCH - 1) / BWD_BATCH) * BWD_BATCH, n);   // whole chunks: all their slots get written
	const float4* tile_rec = instances + 5 * (size_t)range.x;
	// rows of the instances nobody reaches (behind every pixel's last contributor): zeros
	for (int q = n_done + (tid >> 2); q < n; q += 64) {
		const uint32_t row = __float_as_uint(tile_rec[5 * (size_t)q + 4].w);
		float4* dst = reinterpret_cast<float4*>(grad_rows + (size_t)row * GRAD_WORDS);
		dst[tid & 3] = make_float4(0.f, 0.f, 0.f, 0.f);
		if ((tid & 3) == 0) dst[4] = make_float4(0.f, 0.f, 0.f, 0.f);
	}
	if (n_done == 0) return;
	const int nsb = (n_done + BWD_SUPER - 1) / BWD_SUPER;

	const uint32_t* tile_mask = hit_inst + 8 * (size_t)range.x;
	const uint32_t* pix_words = hit_pix + ((size_t)(range.x >> 5) + HIT_PIX_PAD * (size_t)tile) * 2 * TILE_PIXELS;
	const float4 empty = make_float4(0.f, 0.f, 0.f, __int_as_float(-1));
	const float4* recs = &sm.rec[0];
	const uint32_t* masks = &sm.mask[0];

	// pair slot of (instance j of the super-batch, pixel with mask word `word`, bits below it `lt`)
	auto slot_of = [&](int j, int word, uint32_t lt) -> float4* {
		const uint2 pf = sm.pref[j];
		const uint32_t before = (uint32_t)((((unsigned long long)pf.y << 32) | pf.x) >> (8 * word)) & 0xffu;
		return slots + (sm.base[j] + before + __popc(masks[j * 8 + word] & lt));
	};

	PixelConst pc;
	PixelBwd st;
	int o = tid;                                      // row-major id of the pixel this thread walks
	float pxf = 0.f, pyf = 0.f;
	uint32_t o_lt = 0;
	uint32_t bar_phase = 0;

	for (int it = 0; it < nsb; it++) {
		const int sb = nsb - 1 - it;
		const int first = sb * BWD_SUPER;                       // first instance of the super-batch (in the tile)
		const int cnt_sb = min(n_done - first, BWD_SUPER);
		const int nch = (cnt_sb + BWD_BATCH - 1) / BWD_BATCH;
		const int nw = (cnt_sb + 31) >> 5;
		if (it > 0) __syncthreads();   // the previous super-batch is fully consumed
		if (tid == 0) {
			for (int c = 0; c < nch; c++) {
				const uint32_t cnt = (uint32_t)min(cnt_sb - c * BWD_BATCH, BWD_BATCH);
				const size_t at = (size_t)first + (size_t)c * BWD_BATCH;
				mbar_arrive_expect_tx(&sm.bar[c], cnt * (80u + 32u));
				tma_load_1d(&sm.rec[5 * c * BWD_BATCH], tile_rec + 5 * at, cnt * 80u, &sm.bar[c]);
				tma_load_1d(&sm.mask[8 * c * BWD_BATCH], tile_mask + 8 * at, cnt * 32u, &sm.bar[c]);
			}
		}
		// natural owners: the pixel's contribution words of this super-batch (coalesced) -> shared memory
		// (all words are requested before the first is used: the loop would otherwise pay one L2 round trip per word)
		int len = 0;
		{
			uint32_t u[BWD_WORDS];
#pragma unroll
			for (int w = 0; w < BWD_WORDS; w++)
				u[w] = w < nw ? pix_words[(size_t)(((first >> 5) + w) * 2 + 0) * TILE_PIXELS + tid] : 0u;
#pragma unroll
			for (int w = 0; w < BWD_WORDS; w++) {
				if (w < nw) sm.used[w][tid] = u[w];
				len += __popc(u[w]);
			}
		}
		{
			constexpr int PERB = (BWD_SUPER + 256) / 256;       // both loads of a thread are requested before either is stored
			uint32_t bv[PERB];
#pragma unroll
			for (int r = 0; r < PERB; r++) {
				const int i = r * 256 + tid;
				const size_t gi = (size_t)range.x + first + i;
				bv[r] = (i <= cnt_sb && gi) ? pair_incl[gi - 1] : 0u;
			}
#pragma unroll
			for (int r = 0; r < PERB; r++)
				if (r * 256 + tid <= cnt_sb) sm.base[r * 256 + tid] = bv[r];
		}
		if (it == 0) {
			// counting sort of the 256 pixels by the length of their list (deepest super-batch), longest first
			const int c = (BWD_PIX_BINS - 1) - min(len, BWD_PIX_BINS - 1);
			if (tid < BWD_PIX_BINS) sm.bins[tid] = 0u;
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
			o_lt = (1u << (o & 31)) - 1u;
			const unsigned pix_x = tx * TILE + (o & 15), pix_y = ty * TILE + (o >> 4);
			const bool inside = pix_x < (unsigned)W && pix_y < (unsigned)H;
			pxf = (float)pix_x;
			pyf = (float)pix_y;
			const size_t pid = (size_t)W * pix_y + pix_x;
			pc.T_final = inside ? final_T[pid] : 0.f;
			pc.median_contributor = inside ? n_contrib[pid + HW] : 0u;
			pc.final_A = 1 - pc.T_final;
			pc.final_D = inside ? final_T[pid + HW] : 0.f;
			pc.final_D2 = inside ? final_T[pid + 2 * HW] : 0.f;
			pc.dpix0 = pc.dpix1 = pc.dpix2 = 0.f;
			pc.dL_ddepth = pc.dL_daccum = pc.dL_dreg = pc.dL_dmedian = 0.f;
			pc.dn0 = pc.dn1 = pc.dn2 = 0.f;
			pc.mm = 0.f;
			pc.inv_4std = 0.f;
			if (inside) {
				pc.dpix0 = dL_dpix[pid];
				pc.dpix1 = dL_dpix[pid + HW];
				pc.dpix2 = dL_dpix[pid + 2 * HW];
				pc.dL_ddepth = dL_dothers[pid + 0 * HW];
				pc.dL_daccum = dL_dothers[pid + 1 * HW];
				pc.dn0 = dL_dothers[pid + 2 * HW];
				pc.dn1 = dL_dothers[pid + 3 * HW];
				pc.dn2 = dL_dothers[pid + 4 * HW];
				pc.dL_dmedian = dL_dothers[pid + 5 * HW];
				pc.dL_dreg = dL_dothers[pid + 6 * HW];
				pc.mm = median_depth[pid];
				pc.inv_4std = 1.0f / (4 * fmaxf(depth_std[pid] / (1 - pc.T_final), 1e-7f));
			}
			{
				float acc = 0;
				acc += bg[0] * pc.dpix0;
				acc += bg[1] * pc.dpix1;
				acc += bg[2] * pc.dpix2;
				pc.bg_dot = acc;
			}
			st.T = pc.T_final;
		}
		// a chunk barrier completes one phase per super-batch that USES it (the deepest super-batch, taken first, may
		// use fewer chunks than the others): bar_phase holds the parity of every barrier's next phase
		if (warp == 0)
			for (int c = 0; c < nch; c++) mbar_wait(&sm.bar[c], (bar_phase >> c) & 1u);
		bar_phase ^= (1u << nch) - 1u;
		__syncthreads();
		for (int i = tid; i < cnt_sb; i += 256) {
			uint32_t run = 0, lo = 0, hi = 0;
#pragma unroll
			for (int k = 0; k < 8; k++) {
				if (k < 4) lo |= run << (8 * k); else hi |= run << (8 * (k - 4));
				run += __popc(masks[i * 8 + k]);
			}
			sm.pref[i] = make_uint2(lo, hi);
		}
		__syncthreads();

		// ---- phase A.  Natural owners first mark the footprint pairs that did not contribute (1.4 % of the slots:
		// rejected by the exact tests, or behind the pixel's termination point) ...
		{
			const uint32_t lane_lt = (1u << lane) - 1u;
			uint32_t ew[BWD_WORDS];
#pragma unroll
			for (int w = 0; w < BWD_WORDS; w++)
				ew[w] = w < nw ? (pix_words[(size_t)(((first >> 5) + w) * 2 + 1) * TILE_PIXELS + tid] & ~sm.used[w][tid]) : 0u;
#pragma unroll
			for (int w = 0; w < BWD_WORDS; w++) {
				uint32_t e = ew[w];
				while (e) {
					const int jl = __ffs(e) - 1;
					e &= e - 1;
					*slot_of(w * 32 + jl, warp, lane_lt) = empty;
				}
			}
		}
		// ... then every (sorted) pixel walks ITS contributing pairs of the WHOLE super-batch back to front in one loop
		// (a warp iterates as often as its longest list; lists of a whole tile, sorted, are alike).
		{
			const int oword = o >> 5;
			int g = nw - 1;
			uint32_t m = sm.used[g][o];
			for (;;) {
				while (m == 0u && g > 0) {     // advance to the next non-empty word (towards the front)
					g--;
					m = sm.used[g][o];
				}
				if (!__any_sync(0xffffffffu, m != 0u)) break;
				if (m != 0u) {
					const int jl = 31 - __clz(m);
					m &= ~(1u << jl);
					const int j = g * 32 + jl;
					float4* dst = slot_of(j, oword, o_lt);
					float A, dz, w;
					pixel_step<USE_SA>(recs + j * 5, pxf, pyf, (uint32_t)(first + j), pc, st, A, dz, w);
#if GS2D_BWD_SLOT_CG
					__stcg(dst, make_float4(A, dz, w, __int_as_float(o)));   // L2 only: phase B of this CTA reads it back once
#else
					*dst = make_float4(A, dz, w, __int_as_float(o));
#endif
				}
			}
		}
	}
	__syncthreads();   // every slot of the tile's instances is written; the words are dead
	// the pixel's gradient constants for phase B take the place of the words
	sm.b.pixa[pix_slot(o)] = make_float4(pc.dpix0, pc.dpix1, pc.dpix2, pc.dn0);
	sm.b.pixb[pix_slot(o)] = make_float2(pc.dn1, pc.dn2);

	// ---- phase B: a quarter-warp per instance, lanes over the instance's pair slots (just written by this CTA:
	// L2-hot); pixel constants from shared memory; the record from shared memory when the whole tile is resident (one
	// super-batch), else from global memory.  Instances are taken, window by window, in the order of their slot counts
	// (counting sort by ceil(slots / 8)) so that the four of a warp iterate equally often.
	const bool resident = nsb == 1;
	for (int first = 0; first < n_done; first += BWD_SUPER) {
		const int cnt_sb = min(n_done - first, BWD_SUPER);
		if (!resident) {
			__syncthreads();   // the previous window's order / bases are consumed
			for (int i = tid; i <= cnt_sb; i += 256) {
				const size_t gi = (size_t)range.x + first + i;
				sm.base[i] = gi ? pair_incl[gi - 1] : 0u;
			}
		}
		{
			constexpr int PER = (BWD_SUPER + 255) / 256;
			int cls[PER];
			uint32_t rnk[PER];
			if (tid < BWD_INST_BINS) sm.bins[tid] = 0u;
			__syncthreads();
#pragma unroll
			for (int r = 0; r < PER; r++) {
				const int i = r * 256 + tid;
				int c = BWD_INST_BINS;      // beyond the list: a class of its own
				if (i < cnt_sb) {
					const uint32_t cnt = sm.base[i + 1] - sm.base[i];
					c = (BWD_INST_BINS - 1) - (int)min((cnt + (uint32_t)(BWD_GRP - 1)) / (uint32_t)BWD_GRP, (uint32_t)(BWD_INST_BINS - 1));
				}
				const uint32_t peers = __match_any_sync(0xffffffffu, c);
				const int leader = __ffs(peers) - 1;
				uint32_t at = 0;
				if (lane == leader && c < BWD_INST_BINS) at = atomicAdd(&sm.bins[c], (uint32_t)__popc(peers));
				at = __shfl_sync(0xffffffffu, at, leader);
				cls[r] = c;
				rnk[r] = at + __popc(peers & ((1u << lane) - 1u));
			}
			__syncthreads();
			if (warp == 0) {
				const uint32_t v = lane < BWD_INST_BINS ? sm.bins[lane] : 0u;
				uint32_t incl = v;
#pragma unroll
				for (int d = 1; d < 32; d <<= 1) {
					const uint32_t t = __shfl_up_sync(0xffffffffu, incl, d);
					if (lane >= d) incl += t;
				}
				if (lane < BWD_INST_BINS) sm.bins[lane] = incl - v;
			}
			__syncthreads();
#pragma unroll
			for (int r = 0; r < PER; r++)
				if (cls[r] < BWD_INST_BINS) sm.b.ord[sm.bins[cls[r]] + rnk[r]] = (uint16_t)(r * 256 + tid);
			__syncthreads();
		}
		const int grp = tid / BWD_GRP, gl = tid % BWD_GRP;
		for (int q0 = 0; q0 < cnt_sb; q0 += 256 / BWD_GRP) {
			const int qi = q0 + grp;
			float g[24];
#pragma unroll
			for (int k = 0; k < 24; k++) g[k] = 0.f;
			uint32_t row = 0;
			if (qi < cnt_sb) {
				const int q = sm.b.ord[qi];
				float4 r0, r1, r2;
				if (resident) {
					const float4* rec = recs + 5 * q;
					r0 = rec[0]; r1 = rec[1]; r2 = rec[2];
					row = __float_as_uint(rec[4].w);
				} else {
					const float4* rec = tile_rec + 5 * (size_t)(first + q);
					r0 = rec[0]; r1 = rec[1]; r2 = rec[2];
					row = __float_as_uint(rec[4].w);
				}
				const uint32_t s0 = sm.base[q], s1 = sm.base[q + 1];
#if GS2D_BWD_PREFETCH
				// GS2D_BWD_PREFETCH slots of the lane are in flight while one is worked on (a slot takes ~100 instructions,
				// an L2 round trip is longer than that)
				constexpr int PF = GS2D_BWD_PREFETCH;
				float4 nx[PF];
#pragma unroll
				for (int d = 0; d < PF; d++)
					nx[d] = (s0 + gl + d * BWD_GRP < s1) ? SLOT_LD(slots + s0 + gl + d * BWD_GRP)
					                                      : make_float4(0.f, 0.f, 0.f, __int_as_float(-1));
#endif
				for (uint32_t k = s0 + gl; k < s1; k += BWD_GRP) {
#if GS2D_BWD_PREFETCH
					const float4 sl = nx[0];
#pragma unroll
					for (int d = 0; d + 1 < PF; d++) nx[d] = nx[d + 1];
					if (k + PF * BWD_GRP < s1) nx[PF - 1] = SLOT_LD(slots + k + PF * BWD_GRP);
#else
					const float4 sl = SLOT_LD(slots + k);
#endif
					const int p = __float_as_int(sl.w);
					if (p >= 0)
						instance_step(r0, r1, r2, (float)(tx * TILE + (p & 15)), (float)(ty * TILE + (p >> 4)), sl.x, sl.y,
						              sl.z, sm.b.pixa[pix_slot(p)], sm.b.pixb[pix_slot(p)], g);
				}
			}
			__syncwarp();
#if GS2D_BWD_GRP == 4
			quad_reduce20(g, lane);
			if (qi < cnt_sb) {
				float* dst = grad_rows + (size_t)row * GRAD_WORDS;
				reinterpret_cast<float4*>(dst)[gl] = make_float4(g[0], g[1], g[2], g[3]);
				dst[16 + gl] = g[4];
			}
#else
			oct_reduce24(g, lane);
			// lane L of a group holds row words 3L .. 3L+2; regroup so that lane M < 5 holds words 4M .. 4M+3 (word e
			// sits in lane e / 3, slot e % 3; words 18, 19 are the zero padding of lane 6) and the row goes out as five
			// 16-byte stores instead of eighteen 4-byte ones
#if GS2D_BWD_ROW128
			float4 out;
			{
				const int gbase = lane & ~7;
				float v[4];
#pragma unroll
				for (int i = 0; i < 4; i++) {
					const int slot = (i + gl) & 3;
					const float send = slot == 0 ? g[0] : (slot == 1 ? g[1] : g[2]);
					v[i] = __shfl_sync(0xffffffffu, send, gbase + min((4 * gl + i) / 3, 7));
				}
				out = make_float4(v[0], v[1], v[2], v[3]);
			}
			if (qi < cnt_sb && gl < 5)
				reinterpret_cast<float4*>(grad_rows + (size_t)row * GRAD_WORDS)[gl] = out;
#else
			if (qi < cnt_sb && gl < 6) {
				float* dst = grad_rows + (size_t)row * GRAD_WORDS + 3 * gl;
				dst[0] = g[0]; dst[1] = g[1]; dst[2] = g[2];
			}
#endif
#endif
		}
	}
#if GS2D_BWD_DISCARD
	// The tile's pair slots are dead now (written in phase A, read back once in phase B): tell L2 to drop the lines
	// that lie wholly inside this tile's slot range instead of writing them back to HBM (16 B x ~15 M pairs a frame).
	__syncthreads();
	{
		// (with the whole tile resident the two ends of its slot range are still in shared memory: no global round trip
		// at the very end of the CTA, where nothing is left to hide it)
		const size_t gi0 = (size_t)range.x;
		const uint32_t t0 = resident ? sm.base[0] : (gi0 ? pair_incl[gi0 - 1] : 0u);
		const uint32_t t1 = resident ? sm.base[n_done] : pair_incl[gi0 + n_done - 1];
		const uintptr_t a0 = (reinterpret_cast<uintptr_t>(slots + t0) + 127) & ~(uintptr_t)127;
		const uintptr_t a1 = reinterpret_cast<uintptr_t>(slots + t1) & ~(uintptr_t)127;
		for (uintptr_t a = a0 + (uintptr_t)tid * 128; a < a1; a += 256 * 128)
			asm volatile("discard.global.L2 [%0], 128;" ::"l"(a) : "memory");
	}
#endif
}

template <bool USE_SA, typename... Args>
int launch_bwd_variant(unsigned blocks, cudaStream_t st, Args... args)
{
	// the opt-in to > 48 KB of dynamic shared memory is per device: made once for every device this variant runs on
	static std::atomic<unsigned long long> opted_in{0};
	int dev = 0;
	GS2D_CUDA_OK(cudaGetDevice(&dev));
	if (dev >= 64 || !((opted_in.load(std::memory_order_relaxed) >> dev) & 1ull)) {
		GS2D_CUDA_OK(cudaFuncSetAttribute(composite_backward_pixels_kernel<USE_SA>,
		                                  cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(BwdSmem)));
		if (dev < 64) opted_in.fetch_or(1ull << dev, std::memory_order_relaxed);
	}
	composite_backward_pixels_kernel<USE_SA><<<blocks, 256, sizeof(BwdSmem), st>>>(args...);
	return GS2D_OK;
}

}  // namespace

size_t composite_backward_scratch_bytes(int64_t I, int64_t pairs, int64_t tiles)
{
	Carver c(nullptr);
	c.take<float>((size_t)(I > 0 ? I : 0) * GRAD_WORDS);
	c.take<float4>((size_t)(pairs > 0 ? pairs : 0) + 16);
	(void)tiles;
	return c.total();
}

int launch_composite_backward(const gs2d_settings& s, const gs2d_inputs& in, const ImageState& im,
                              const BinningState& b, int64_t I, int64_t pairs, const float* dL_dcolor,
                              const float* dL_dothers, void* scratch, float** grad_rows_out, cudaStream_t st)
{
	const int W = s.image_width, H = s.image_height;
	const unsigned gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
	Carver c(scratch);
	float* rows = c.take<float>((size_t)I * GRAD_WORDS);
	float4* slots = c.take<float4>((size_t)(pairs > 0 ? pairs : 0) + 16);
	*grad_rows_out = rows;
	if (gx * gy == 0 || I <= 0) return GS2D_OK;
	{
		ProfScope ps(K_COMPOSITE_BWD, st);
		const uint32_t cap = (uint32_t)(pairs > 0 ? pairs : 0);
		const int rc = s.use_sa
			? launch_bwd_variant<true>(gx * gy, st, W, H, gx, (const uint2*)im.ranges, (const float4*)b.instances,
			                           (const uint32_t*)b.hit_inst, (const uint32_t*)b.hit_pix, (const uint32_t*)b.pair_incl,
			                           (const float*)in.background, (const float*)im.final_T, (const uint32_t*)im.n_contrib,
			                           (const float*)im.median_depth, (const float*)im.depth_std, dL_dcolor, dL_dothers,
			                           slots, rows, cap, b.pair_counters + 3)
			: launch_bwd_variant<false>(gx * gy, st, W, H, gx, (const uint2*)im.ranges, (const float4*)b.instances,
			                            (const uint32_t*)b.hit_inst, (const uint32_t*)b.hit_pix, (const uint32_t*)b.pair_incl,
			                            (const float*)in.background, (const float*)im.final_T, (const uint32_t*)im.n_contrib,
			                            (const float*)im.median_depth, (const float*)im.depth_std, dL_dcolor, dL_dothers,
			                            slots, rows, cap, b.pair_counters + 3);
		if (rc != GS2D_OK) return rc;
		count_launch();
	}
	GS2D_CUDA_OK(cudaGetLastError());
	GS2D_CUDA_OK(cudaGetLastError());
	return GS2D_OK;
}

}  // namespace gs2d
