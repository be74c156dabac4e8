// composite_bwd.cu -- per-tile back-to-front re-traversal (backward of the compositing).
//
// Replaces BACKWARD::render / renderCUDA<3> (CR/backward.cu:143-463).  Same per-(pixel, instance)
// gradient algebra (including the reference's deliberate non-analytic parts: SA confidence from the
// *final* per-pixel statistics, opacity gradient through the clamped alpha), but the work is split by
// what it depends on, so that every lane does useful work and nothing is reduced with atomics:
//
//   kernel A (per PIXEL, sequential back to front; one CTA per tile, lanes = 32 pixels of a warp):
//       the order-dependent part only -- transmittance T = T / (1 - alpha), the suffix recurrences for
//       colour / depth / alpha / normal / dL_dT (CR/backward.cu:323-410).  Every pixel walks ITS OWN list of
//       the pairs that contributed in the forward pass (recorded there, one word per pixel and 32
//       instances), i.e. ~17 instances per tile instead of ~250 and no skip tests, and leaves three scalars
//       per pair -- dL/dalpha, dL/dz, w = alpha * T -- plus its pixel id in the pair's 16-byte slot.  Slots
//       live in global memory (L2), laid out per instance: slot = first slot of the instance (exclusive
//       scan of the footprint-mask popcounts, computed in the forward pass) + rank of the pixel inside the
//       instance's mask.  Pairs of the mask that did not contribute get an "empty" marker.
//   kernel B (per INSTANCE, fully parallel over all instances of the frame, half a warp per instance):
//       the geometric chain rule dL/d{Tu,Tv,Tw}, dL/dmean2D, dL/dopacity, dL/dnormal, dL/dcolour
//       (CR/backward.cu:343, :396, :412-460) from the instance's slots (one coalesced read) + its record,
//       summed over the instance's pixels with 19 shuffles, and written as ONE plain 72-byte store into the
//       instance's own gradient row: the reference issues 16 global atomicAdd per contributing
//       (pixel, instance) pair; here a row is written exactly once, so no atomics and no zero-fill.
//
// Records + masks of a tile arrive in batches of 128 by TMA (two-stage mbarrier ring).  Rows are indexed by
// the instance's position in surfel-major (unsorted) order, so preprocess_bwd.cu reads each surfel's rows
// contiguously.
#include "common.cuh"
#include "tma.cuh"
#include "pair.cuh"

namespace gs2d {

namespace {

#ifndef GS2D_BWD_DISCARD
#define GS2D_BWD_DISCARD 1
#endif
#ifndef GS2D_BWD_PREFETCH
#define GS2D_BWD_PREFETCH 1
#endif
constexpr int BWD_BATCH = 128;
constexpr int BWD_STAGES = 2;

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
	float4 rec[BWD_STAGES][BWD_BATCH * 5];
	uint32_t mask[BWD_STAGES][BWD_BATCH * 8];
	uint2 pref[BWD_BATCH];              // 8 packed bytes: hits of the instance in mask words 0..k-1
	uint32_t base[BWD_BATCH];           // first pair slot of every instance of the batch (global index)
	float pix[6][TILE_PIXELS];          // dL_dpix (3), dL_dnormal (3) per pixel, for phase B
	uint32_t max_last[8];
	uint64_t bar[BWD_STAGES];
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

// Phase B for one (instance, pixel) pair: the geometric chain (CR/backward.cu:343, :396, :412-460).
// A = dL/dalpha, dz = dL/dz, w = alpha * T from phase A; adds the 18 per-instance terms into g.
__device__ __forceinline__ void instance_step(const float4 r0, const float4 r1, const float4 r2, float pxf, float pyf,
                                              float A, float dz, float w, const float* __restrict__ pixc, int pid,
                                              float (&g)[24])
{
	PairFront f;
	float inv_pz;
	pair_geom_fast(r0, r1, r2, pxf, pyf, f, inv_pz);
	const float3 Tw = make_float3(r2.x, r2.y, r2.z);
	const float opa = r2.w, G = f.G;
	g[GR_COL + 0] += w * pixc[0 * TILE_PIXELS + pid];
	g[GR_COL + 1] += w * pixc[1 * TILE_PIXELS + pid];
	g[GR_COL + 2] += w * pixc[2 * TILE_PIXELS + pid];
	g[GR_NRM + 0] += w * pixc[3 * TILE_PIXELS + pid];
	g[GR_NRM + 1] += w * pixc[4 * TILE_PIXELS + pid];
	g[GR_NRM + 2] += w * pixc[5 * TILE_PIXELS + pid];
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
		g[0] += -dL_dk.x; g[1] += -dL_dk.y; g[2] += -dL_dk.z;
		g[3] += -dL_dl.x; g[4] += -dL_dl.y; g[5] += -dL_dl.z;
		g[6] += pxf * dL_dk.x + pyf * dL_dl.x + dz * f.s.x;
		g[7] += pxf * dL_dk.y + pyf * dL_dl.y + dz * f.s.y;
		g[8] += pxf * dL_dk.z + pyf * dL_dl.z + dz * 1.0f;
	} else {
		g[8] += dz;
		g[GR_M2D + 0] += dL_dG * (-G * GS2D_FILTER_INV_SQUARE * f.d.x);
		g[GR_M2D + 1] += dL_dG * (-G * GS2D_FILTER_INV_SQUARE * f.d.y);
	}
	g[GR_OPA] += G * A;
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
	__shared__ BwdSmem sm;

	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
	const unsigned tile = blockIdx.x;
	const unsigned tx = tile % gx, ty = tile / gx;
	const int lx = thread_lx<GS2D_BWD_BLOCK>(tid), ly = thread_ly<GS2D_BWD_BLOCK>(tid);   // thread <-> tile pixel: pair.cuh
	const int ptid = pixel_thread(lx, ly);               // row-major pixel index: this pixel's words in hit_pix
	const unsigned pix_x = tx * TILE + lx;
	const unsigned pix_y = ty * TILE + ly;
	const bool inside = pix_x < (unsigned)W && pix_y < (unsigned)H;
	const float pxf = (float)pix_x, pyf = (float)pix_y;
	const size_t HW = (size_t)H * W;
	const size_t pid = (size_t)W * pix_y + pix_x;

	const uint2 range = ranges[tile];
	const int n = (int)(range.y - range.x);
	// Memory safety against a wrong pair count from the caller (it sized `slots`): a tile whose slots would end beyond
	// the scratch writes nothing and raises the flag (pair_counters[3] of the binning state) instead.
	if (n > 0 && pair_incl[(size_t)range.x + n - 1] > slot_cap) {
		if (threadIdx.x == 0) *overflow = 1u;
		return;
	}

	PixelConst pc;
	pc.T_final = inside ? final_T[pid] : 0.f;
	const uint32_t last_contributor = inside ? n_contrib[pid] : 0u;
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
	sm.pix[0][tid] = pc.dpix0; sm.pix[1][tid] = pc.dpix1; sm.pix[2][tid] = pc.dpix2;
	sm.pix[3][tid] = pc.dn0;   sm.pix[4][tid] = pc.dn1;   sm.pix[5][tid] = pc.dn2;
	PixelBwd st;
	st.T = pc.T_final;

	// deepest last contributor of the tile: nothing behind it is ever visited
	uint32_t wmax = last_contributor;
#pragma unroll
	for (int o = 16; o >= 1; o >>= 1) wmax = max(wmax, __shfl_xor_sync(0xffffffffu, wmax, o));
	if (lane == 0) sm.max_last[warp] = wmax;
	if (tid == 0) {
		for (int s = 0; s < BWD_STAGES; s++) mbar_init(&sm.bar[s], 1);
		mbar_fence_init();
	}
	__syncthreads();
	uint32_t bmax = 0;
#pragma unroll
	for (int i = 0; i < 8; i++) bmax = max(bmax, sm.max_last[i]);
	const int n_used = min((int)bmax, n);            // instances [0, n_used) can matter
	const int nb = (n_used + BWD_BATCH - 1) / BWD_BATCH;
	const int n_done = min(nb * BWD_BATCH, n);       // instances of whole batches: all their slots get written
	const float4* tile_rec = instances + 5 * (size_t)range.x;
	// rows of the instances nobody reaches (behind every pixel's last contributor): zeros
	for (int q = n_done + (tid >> 2); q < n; q += 64) {
		const uint32_t row = __float_as_uint(tile_rec[5 * (size_t)q + 4].w);
		float4* dst = reinterpret_cast<float4*>(grad_rows + (size_t)row * GRAD_WORDS);
		dst[tid & 3] = make_float4(0.f, 0.f, 0.f, 0.f);
		if ((tid & 3) == 0) dst[4] = make_float4(0.f, 0.f, 0.f, 0.f);
	}
	if (nb == 0) return;

	const uint32_t* tile_mask = hit_inst + 8 * (size_t)range.x;
	auto issue = [&](int b, int stage) {
		const uint32_t cnt = (uint32_t)min(n - b * BWD_BATCH, BWD_BATCH);
		mbar_arrive_expect_tx(&sm.bar[stage], cnt * (80u + 32u));
		tma_load_1d(&sm.rec[stage][0], tile_rec + 5 * (size_t)b * BWD_BATCH, cnt * 80u, &sm.bar[stage]);
		tma_load_1d(&sm.mask[stage][0], tile_mask + 8 * (size_t)b * BWD_BATCH, cnt * 32u, &sm.bar[stage]);
	};
	if (tid == 0) issue(nb - 1, 0);

	const uint32_t lane_lt = (1u << lane) - 1u;
	const uint32_t* pix_words = hit_pix + ((size_t)(range.x >> 5) + HIT_PIX_PAD * (size_t)tile) * 2 * TILE_PIXELS;
	const float4 empty = make_float4(0.f, 0.f, 0.f, __int_as_float(-1));

	for (int it = 0; it < nb; it++) {
		const int b = nb - 1 - it;
		const int stage = it & 1;
		const uint32_t parity = (uint32_t)(it >> 1) & 1u;
		__syncthreads();   // batch it-1 fully consumed -> its stage and pref/base may be overwritten
		if (tid == 0 && it + 1 < nb) issue(b - 1, stage ^ 1);
		const int cnt = min(n - b * BWD_BATCH, BWD_BATCH);
		// words of this pixel: pairs that contributed / pairs that own a slot
		uint32_t used[BWD_BATCH / 32], foot[BWD_BATCH / 32];
#pragma unroll
		for (int g = 0; g < BWD_BATCH / 32; g++) {
			used[g] = pix_words[(size_t)((b * (BWD_BATCH / 32) + g) * 2 + 0) * TILE_PIXELS + ptid];
			foot[g] = pix_words[(size_t)((b * (BWD_BATCH / 32) + g) * 2 + 1) * TILE_PIXELS + ptid];
		}
		if (tid < cnt) {
			const size_t gi = (size_t)range.x + (size_t)b * BWD_BATCH + tid;
			sm.base[tid] = gi ? pair_incl[gi - 1] : 0u;
		}
		if (warp == 0) mbar_wait(&sm.bar[stage], parity);
		__syncthreads();

		const float4* recs = &sm.rec[stage][0];
		const uint32_t* masks = &sm.mask[stage][0];
#if GS2D_BWD_BLOCK
		// re-order every mask in place from row-pair words to warp-block words (pair.cuh): word pair (2m, 2m+1) holds
		// tile rows 4m..4m+3, i.e. exactly the two 8 x 4 blocks of block row m
		for (int e = tid; e < cnt * 4; e += 256) {
			uint2* w2 = reinterpret_cast<uint2*>(&sm.mask[stage][0]) + e;
			const uint2 rows = *w2;
			*w2 = make_uint2(__byte_perm(rows.x, rows.y, 0x6420), __byte_perm(rows.x, rows.y, 0x7531));
		}
		__syncthreads();
#endif
		if (tid < cnt) {
			uint32_t run = 0, lo = 0, hi = 0;
#pragma unroll
			for (int k = 0; k < 8; k++) {
				if (k < 4) lo |= run << (8 * k); else hi |= run << (8 * (k - 4));
				run += __popc(masks[tid * 8 + k]);
			}
			sm.pref[tid] = make_uint2(lo, hi);
		}
		__syncthreads();

		// every pixel walks the pairs of its footprint words of the WHOLE batch, back to front, in one loop (a warp
		// iterates as often as its longest pixel list; lists of 128 instances balance better than lists of 32)
		{
			constexpr int NG = BWD_BATCH / 32;
			int g = NG - 1;
			uint32_t m = foot[NG - 1], u = used[NG - 1];
			for (;;) {
				while (m == 0u && g > 0) {     // advance to the next non-empty word (towards the front)
					g--;
#pragma unroll
					for (int k = 0; k < NG - 1; k++)
						if (g == k) { m = foot[k]; u = used[k]; }
				}
				if (!__any_sync(0xffffffffu, m != 0u)) break;
				if (m != 0u) {
					const int jl = 31 - __clz(m);
					m &= ~(1u << jl);
					const int j = g * 32 + jl;
					const uint2 pf = sm.pref[j];
					const uint32_t before = ((warp < 4 ? pf.x >> (8 * warp) : pf.y >> (8 * (warp - 4))) & 0xffu);
					float4* dst = slots + (sm.base[j] + before + __popc(masks[j * 8 + warp] & lane_lt));
					if ((u >> jl) & 1u) {
						float A, dz, w;
						pixel_step<USE_SA>(recs + j * 5, pxf, pyf, (uint32_t)(b * BWD_BATCH + j), pc, st, A, dz, w);
						*dst = make_float4(A, dz, w, __int_as_float(tid));
					} else {
						*dst = empty;
					}
				}
			}
		}
	}

	// ---- phase B: a quarter-warp per instance of the tile, lanes over the instance's pair slots (just written by
	// this CTA: L2-hot); pixel constants from shared memory, the record straight from global memory.
	__syncthreads();
	const int grp = tid >> 3, gl = tid & 7;
	for (int q0 = 0; q0 < n_done; q0 += 32) {
		const int q = q0 + grp;
		float g[24];
#pragma unroll
		for (int k = 0; k < 24; k++) g[k] = 0.f;
		uint32_t row = 0;
		if (q < n_done) {
			const float4* rec = tile_rec + 5 * (size_t)q;
			const float4 r0 = rec[0], r1 = rec[1], r2 = rec[2];
			row = __float_as_uint(rec[4].w);
			const size_t gi = (size_t)range.x + q;
			const uint32_t s0 = gi ? pair_incl[gi - 1] : 0u, s1 = pair_incl[gi];
#if GS2D_BWD_PREFETCH
			// the next slot is requested before the current one is worked on (phase B waits on L2 otherwise)
			float4 nx = (s0 + gl < s1) ? slots[s0 + gl] : make_float4(0.f, 0.f, 0.f, __int_as_float(-1));
#endif
			for (uint32_t k = s0 + gl; k < s1; k += 8) {
#if GS2D_BWD_PREFETCH
				const float4 sl = nx;
				if (k + 8 < s1) nx = slots[k + 8];
#else
				const float4 sl = slots[k];
#endif
				const int p = __float_as_int(sl.w);
				if (p >= 0)
					instance_step(r0, r1, r2, (float)(tx * TILE + thread_lx<GS2D_BWD_BLOCK>(p)), (float)(ty * TILE + thread_ly<GS2D_BWD_BLOCK>(p)), sl.x, sl.y,
					              sl.z, &sm.pix[0][0], p, g);
			}
		}
		__syncwarp();
		oct_reduce24(g, lane);
		if (q < n_done && gl < 6) {
			float* dst = grad_rows + (size_t)row * GRAD_WORDS + 3 * gl;
			dst[0] = g[0]; dst[1] = g[1]; dst[2] = g[2];
		}
	}
#if GS2D_BWD_DISCARD
	// The tile's pair slots are dead now (written in phase A, read back once in phase B): tell L2 to drop the lines
	// that lie wholly inside this tile's slot range instead of writing them back to HBM (16 B x ~15 M pairs a frame).
	__syncthreads();
	{
		const size_t gi0 = (size_t)range.x;
		const uint32_t t0 = gi0 ? pair_incl[gi0 - 1] : 0u, t1 = pair_incl[gi0 + n_done - 1];
		const uintptr_t a0 = (reinterpret_cast<uintptr_t>(slots + t0) + 127) & ~(uintptr_t)127;
		const uintptr_t a1 = reinterpret_cast<uintptr_t>(slots + t1) & ~(uintptr_t)127;
		for (uintptr_t a = a0 + (uintptr_t)tid * 128; a < a1; a += 256 * 128)
			asm volatile("discard.global.L2 [%0], 128;" ::"l"(a) : "memory");
	}
#endif
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
		if (s.use_sa)
			composite_backward_pixels_kernel<true><<<gx * gy, 256, 0, st>>>(
				W, H, gx, (const uint2*)im.ranges, (const float4*)b.instances, b.hit_inst, b.hit_pix, b.pair_incl,
				in.background, im.final_T, im.n_contrib, im.median_depth, im.depth_std, dL_dcolor, dL_dothers, slots,
				rows, (uint32_t)(pairs > 0 ? pairs : 0), b.pair_counters + 3);
		else
			composite_backward_pixels_kernel<false><<<gx * gy, 256, 0, st>>>(
				W, H, gx, (const uint2*)im.ranges, (const float4*)b.instances, b.hit_inst, b.hit_pix, b.pair_incl,
				in.background, im.final_T, im.n_contrib, im.median_depth, im.depth_std, dL_dcolor, dL_dothers, slots,
				rows, (uint32_t)(pairs > 0 ? pairs : 0), b.pair_counters + 3);
		count_launch();
	}
	GS2D_CUDA_OK(cudaGetLastError());
	GS2D_CUDA_OK(cudaGetLastError());
	return GS2D_OK;
}

}  // namespace gs2d
