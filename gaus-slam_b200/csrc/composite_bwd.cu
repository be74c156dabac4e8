// composite_bwd.cu -- per-tile back-to-front re-traversal (backward of the compositing).
//
// Replaces BACKWARD::render / renderCUDA<3> (CR/backward.cu:143-463).  Same per-(pixel, instance)
// gradient algebra (including the reference's deliberate non-analytic parts: SA confidence from the
// *final* per-pixel statistics, opacity gradient through the clamped alpha), but the work is split by
// what it depends on, so that every lane does useful work and nothing is reduced with atomics:
//
//   phase A (per PIXEL, sequential back to front, lanes = 32 pixels of a warp):
//       the order-dependent part only -- transmittance T = T / (1 - alpha), the suffix recurrences for
//       colour / depth / alpha / normal / dL_dT (CR/backward.cu:323-410).  Every pixel walks ITS OWN list of
//       the pairs that contributed in the forward pass (recorded there, one word per pixel and 32
//       instances), i.e. ~17 instances per tile instead of ~250 and no skip tests, and leaves three scalars per pair -- dL/dalpha, dL/dz, w = alpha * T -- plus its
//       pixel id in the pair's 16-byte slot in shared memory.  Slots are laid out per instance
//       (slot = instance base + rank of the pixel inside the instance's mask).
//   phase B (per INSTANCE, lanes = that instance's contributing pixels, half a warp per instance):
//       the geometric chain rule dL/d{Tu,Tv,Tw}, dL/dmean2D, dL/dopacity, dL/dnormal, dL/dcolour
//       (CR/backward.cu:343, :396, :412-460) from the slot + the instance record, summed over the
//       instance's pixels with 19 shuffles, and written as ONE plain 72-byte store into the instance's own
//       gradient row: the reference issues 16 global atomicAdd per contributing (pixel, instance) pair;
//       here a row is written exactly once, so no atomics and no zero-fill of the rows are needed.
//
// Instances come in batches of 128 (records + masks staged by TMA, two-stage mbarrier ring); a batch is
// cut into sub-ranges whose pair count fits the slot buffer.  Rows are indexed by the instance's position
// in surfel-major (unsorted) order, so preprocess_bwd.cu reads each surfel's rows contiguously.
#include "common.cuh"
#include "tma.cuh"
#include "pair.cuh"

namespace gs2d {

namespace {

constexpr int BWD_BATCH = 128;
constexpr int BWD_STAGES = 2;
constexpr int BWD_CAP = 2304;     // pair slots (16 B each) per sub-range; >= 256 = one full-tile instance

struct PixelConst {
	float T_final, final_A, final_D, final_D2;
	float dpix0, dpix1, dpix2;
	float dL_ddepth, dL_daccum, dL_dreg, dL_dmedian;
	float dn0, dn1, dn2;
	float mm, inv_4std, bg_dot;      // inv_4std = 1 / (4 max(mstd / (1 - T_final), 1e-7)), CR/backward.cu:349
	uint32_t median_contributor;
};

struct PixelBwd {
	float T;
	float acc_c0 = 0.f, acc_c1 = 0.f, acc_c2 = 0.f;
	float last_c0 = 0.f, last_c1 = 0.f, last_c2 = 0.f;
	float last_alpha = 0.f;
	float last_depth = 0.f, acc_depth = 0.f, acc_alpha = 0.f;
	float acc_n0 = 0.f, acc_n1 = 0.f, acc_n2 = 0.f;
	float last_n0 = 0.f, last_n1 = 0.f, last_n2 = 0.f;
	float last_dL_dT = 0.f;
};

struct BwdSmem {
	float4 rec[BWD_STAGES][BWD_BATCH * 5];
	uint32_t mask[BWD_STAGES][BWD_BATCH * 8];
	float4 slot[BWD_CAP];
	float pix[6][TILE_PIXELS];          // dL_dpix (3), dL_dnormal (3) per pixel, for phase B
	uint2 pref[BWD_BATCH];              // 8 packed bytes: hits of the instance in mask words 0..k-1
	uint32_t base[BWD_BATCH + 1];       // exclusive prefix of the instances' hit counts within the batch
	uint32_t wsum[8];
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

	float dL_dalpha = 0.0f;
	st.acc_c0 = st.last_alpha * st.last_c0 + (1.f - st.last_alpha) * st.acc_c0;
	st.last_c0 = r4.x;
	dL_dalpha += (r4.x - st.acc_c0) * pc.dpix0;
	st.acc_c1 = st.last_alpha * st.last_c1 + (1.f - st.last_alpha) * st.acc_c1;
	st.last_c1 = r4.y;
	dL_dalpha += (r4.y - st.acc_c1) * pc.dpix1;
	st.acc_c2 = st.last_alpha * st.last_c2 + (1.f - st.last_alpha) * st.acc_c2;
	st.last_c2 = r4.z;
	dL_dalpha += (r4.z - st.acc_c2) * pc.dpix2;

	float conf = 1;
	if (USE_SA) {
		// confidence from the FINAL per-pixel statistics, treated as a constant (CR/backward.cu:349)
		conf = T < 0.5f ? __expf(-(c_d - pc.mm) * (c_d - pc.mm) * pc.inv_4std) : 1.f;
		c_d = c_d * conf + pc.mm * (1 - conf);
	}

	float dL_dz = 0.0f;
	float dL_dweight = 0;
	if (q + 1 == pc.median_contributor) dL_dz += pc.dL_dmedian;
	float m_d = 0.f, dmd_dd = 0.f;
	if (USE_SA) {
		dL_dweight += ((c_d - pc.mm) * (c_d - pc.mm)) * pc.dL_dreg;
	} else {
		m_d = GS2D_FAR / (GS2D_FAR - GS2D_NEAR) * (1 - GS2D_NEAR / c_d);
		dmd_dd = (GS2D_FAR * GS2D_NEAR) / ((GS2D_FAR - GS2D_NEAR) * c_d * c_d);
		dL_dweight += (pc.final_D2 + m_d * m_d * pc.final_A - 2 * m_d * pc.final_D) * pc.dL_dreg;
	}
	dL_dalpha += dL_dweight - st.last_dL_dT;
	st.last_dL_dT = dL_dweight * alpha + (1 - alpha) * st.last_dL_dT;
	if (USE_SA) {
		dL_dz += conf * 2.0f * w * (c_d - pc.mm) * pc.dL_dreg;
	} else {
		const float dL_dmd = 2.0f * (T * alpha) * (m_d * pc.final_A - pc.final_D) * pc.dL_dreg;
		dL_dz += dL_dmd * dmd_dd;
	}
	st.acc_depth = st.last_alpha * st.last_depth + (1.f - st.last_alpha) * st.acc_depth;
	st.last_depth = c_d;
	dL_dalpha += (c_d - st.acc_depth) * pc.dL_ddepth;
	st.acc_alpha = st.last_alpha * 1.0f + (1.f - st.last_alpha) * st.acc_alpha;
	dL_dalpha += (1 - st.acc_alpha) * pc.dL_daccum;

	st.acc_n0 = st.last_alpha * st.last_n0 + (1.f - st.last_alpha) * st.acc_n0;
	st.last_n0 = r3.x;
	dL_dalpha += (r3.x - st.acc_n0) * pc.dn0;
	st.acc_n1 = st.last_alpha * st.last_n1 + (1.f - st.last_alpha) * st.acc_n1;
	st.last_n1 = r3.y;
	dL_dalpha += (r3.y - st.acc_n1) * pc.dn1;
	st.acc_n2 = st.last_alpha * st.last_n2 + (1.f - st.last_alpha) * st.acc_n2;
	st.last_n2 = r3.z;
	dL_dalpha += (r3.z - st.acc_n2) * pc.dn2;

	dL_dalpha *= T;
	st.last_alpha = alpha;
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
                                              float (&g)[GR_USED])
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

// Sum g[0..17] over the 16 lanes of each half-warp.  On return lane L of a half holds the total of
// g[L] in `a`; lanes with bit 3 clear / set hold the totals of g[16] / g[17] in `b`.
__device__ __forceinline__ void half_reduce18(float (&g)[GR_USED], int lane, float& a, float& b)
{
	{
		const bool up = lane & 8;
#pragma unroll
		for (int i = 0; i < 8; i++) {
			const float send = up ? g[i] : g[i + 8];
			const float keep = up ? g[i + 8] : g[i];
			g[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
		}
		const float send = up ? g[16] : g[17];
		const float keep = up ? g[17] : g[16];
		g[16] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
	}
	{
		const bool up = lane & 4;
#pragma unroll
		for (int i = 0; i < 4; i++) {
			const float send = up ? g[i] : g[i + 4];
			const float keep = up ? g[i + 4] : g[i];
			g[i] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
		}
		g[16] += __shfl_xor_sync(0xffffffffu, g[16], 4);
	}
	{
		const bool up = lane & 2;
#pragma unroll
		for (int i = 0; i < 2; i++) {
			const float send = up ? g[i] : g[i + 2];
			const float keep = up ? g[i + 2] : g[i];
			g[i] = keep + __shfl_xor_sync(0xffffffffu, send, 2);
		}
		g[16] += __shfl_xor_sync(0xffffffffu, g[16], 2);
	}
	{
		const bool up = lane & 1;
		const float send = up ? g[0] : g[1];
		const float keep = up ? g[1] : g[0];
		g[0] = keep + __shfl_xor_sync(0xffffffffu, send, 1);
		g[16] += __shfl_xor_sync(0xffffffffu, g[16], 1);
	}
	a = g[0];
	b = g[16];
}

template <bool USE_SA>
__global__ void __launch_bounds__(256, 3)
composite_backward_kernel(int W, int H, unsigned gx, const uint2* __restrict__ ranges,
                          const float4* __restrict__ instances, const uint32_t* __restrict__ hit_inst,
                          const uint32_t* __restrict__ hit_pix, const float* __restrict__ bg,
                          const float* __restrict__ final_T,
                          const uint32_t* __restrict__ n_contrib, const float* __restrict__ median_depth,
                          const float* __restrict__ depth_std, const float* __restrict__ dL_dpix,
                          const float* __restrict__ dL_dothers, float* __restrict__ grad_rows)
{
	extern __shared__ __align__(128) unsigned char smem_raw[];
	BwdSmem& sm = *reinterpret_cast<BwdSmem*>(smem_raw);

	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
	const unsigned tile = blockIdx.x;
	const unsigned tx = tile % gx, ty = tile / gx;
	const unsigned pix_x = tx * TILE + (tid & 15);       // thread t <-> tile pixel (t & 15, t >> 4), pair.cuh
	const unsigned pix_y = ty * TILE + (tid >> 4);
	const bool inside = pix_x < (unsigned)W && pix_y < (unsigned)H;
	const float pxf = (float)pix_x, pyf = (float)pix_y;
	const size_t HW = (size_t)H * W;
	const size_t pid = (size_t)W * pix_y + pix_x;

	const uint2 range = ranges[tile];
	const int n = (int)(range.y - range.x);

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

	const float4* tile_rec = instances + 5 * (size_t)range.x;
	const uint32_t* tile_mask = hit_inst + 8 * (size_t)range.x;
	auto issue = [&](int b, int stage) {
		const uint32_t cnt = (uint32_t)min(n_used - b * BWD_BATCH, BWD_BATCH);
		mbar_arrive_expect_tx(&sm.bar[stage], cnt * (80u + 32u));
		tma_load_1d(&sm.rec[stage][0], tile_rec + 5 * (size_t)b * BWD_BATCH, cnt * 80u, &sm.bar[stage]);
		tma_load_1d(&sm.mask[stage][0], tile_mask + 8 * (size_t)b * BWD_BATCH, cnt * 32u, &sm.bar[stage]);
	};
	if (tid == 0 && nb > 0) issue(nb - 1, 0);

	// rows of the instances nobody reaches (behind every pixel's last contributor): zeros
	for (int q = n_used + (tid >> 2); q < n; q += 64) {
		const uint32_t row = __float_as_uint(tile_rec[5 * (size_t)q + 4].w);
		float4* dst = reinterpret_cast<float4*>(grad_rows + (size_t)row * GRAD_WORDS);
		dst[tid & 3] = make_float4(0.f, 0.f, 0.f, 0.f);
		if ((tid & 3) == 0) dst[4] = make_float4(0.f, 0.f, 0.f, 0.f);
	}

	const uint32_t lane_lt = (1u << lane) - 1u;
	const uint32_t* pix_words = hit_pix + ((size_t)(range.x >> 5) + 4 * (size_t)tile) * TILE_PIXELS;
	const int hw = tid >> 4, hl = tid & 15;          // half-warp id / lane inside it (phase B)

	for (int it = 0; it < nb; it++) {
		const int b = nb - 1 - it;
		const int stage = it & 1;
		const uint32_t parity = (uint32_t)(it >> 1) & 1u;
		__syncthreads();   // batch it-1 fully consumed -> its stage may be overwritten
		if (tid == 0 && it + 1 < nb) issue(b - 1, stage ^ 1);
		if (warp == 0) mbar_wait(&sm.bar[stage], parity);
		__syncthreads();

		const int cnt = min(n_used - b * BWD_BATCH, BWD_BATCH);
		const float4* recs = &sm.rec[stage][0];
		const uint32_t* masks = &sm.mask[stage][0];

		// per-instance hit counts: packed prefix over the 8 mask words, block-wide exclusive scan
		if (tid < BWD_BATCH) {
			uint32_t run = 0, lo = 0, hi = 0;
			if (tid < cnt) {
#pragma unroll
				for (int k = 0; k < 8; k++) {
					if (k < 4) lo |= run << (8 * k); else hi |= run << (8 * (k - 4));
					run += __popc(masks[tid * 8 + k]);
				}
			}
			sm.pref[tid] = make_uint2(lo, hi);
			uint32_t incl = run;
#pragma unroll
			for (int o = 1; o < 32; o <<= 1) {
				const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
				if (lane >= o) incl += t;
			}
			if (lane == 31) sm.wsum[warp] = incl;
			sm.base[tid + 1] = incl;   // warp-local inclusive value; the other warps' totals are added below
		}
		__syncthreads();
		if (tid < BWD_BATCH) {
			uint32_t off = 0;
			for (int w = 0; w < warp; w++) off += sm.wsum[w];
			if (off) sm.base[tid + 1] += off;
		}
		if (tid == 0) sm.base[0] = 0;

		// the pairs of this batch that contributed in the forward pass, one word per pixel and 32 instances
		uint32_t hits[BWD_BATCH / 32];
#pragma unroll
		for (int g = 0; g < BWD_BATCH / 32; g++)
			hits[g] = pix_words[(size_t)(b * (BWD_BATCH / 32) + g) * TILE_PIXELS + tid];
		__syncthreads();

		// sub-ranges of roughly equal pair counts, each fitting the slot buffer
		const uint32_t total = sm.base[cnt];
		const uint32_t parts = (total + BWD_CAP - 1) / BWD_CAP;
		const uint32_t target = parts ? (total + parts - 1) / parts + 128 : 0;
		const uint32_t cap = min((uint32_t)BWD_CAP, max(target, 256u));
		int hi = cnt;
		while (hi > 0) {
			// largest sub-range [lo, hi) whose pairs fit the slot buffer
			const uint32_t total_hi = sm.base[hi];
			const bool fits = tid < hi && (total_hi - sm.base[tid] <= cap);
			const int lo = hi - __syncthreads_count(fits);   // also: phase B of the previous sub-range is done
			const uint32_t base_lo = sm.base[lo];
			// slots of pairs that did not contribute (mask superset, terminated pixels) stay marked empty
			for (uint32_t k = tid; k < total_hi - base_lo; k += 256) sm.slot[k].w = __int_as_float(-1);
			__syncthreads();

			// ---- phase A: every pixel walks its hits in [lo, hi), back to front
#pragma unroll
			for (int g = BWD_BATCH / 32 - 1; g >= 0; g--) {
				const int a = max(lo - 32 * g, 0), e = min(hi - 32 * g, 32);
				if (e <= a) continue;
				uint32_t m = hits[g] & (e == 32 ? 0xffffffffu : ((1u << e) - 1u)) & ~((1u << a) - 1u);
				while (m) {
					const int jl = 31 - __clz(m);
					m &= ~(1u << jl);
					const int j = g * 32 + jl;
					const uint32_t q = (uint32_t)(b * BWD_BATCH + j);
					const uint2 pf = sm.pref[j];
					const uint32_t before = ((warp < 4 ? pf.x >> (8 * warp) : pf.y >> (8 * (warp - 4))) & 0xffu);
					const uint32_t si = sm.base[j] - base_lo + before + __popc(masks[j * 8 + warp] & lane_lt);
					float A, dz, w;
					pixel_step<USE_SA>(recs + j * 5, pxf, pyf, q, pc, st, A, dz, w);
					sm.slot[si] = make_float4(A, dz, w, __int_as_float(tid));
				}
			}
			__syncthreads();

			// ---- phase B: half a warp per instance, lanes over the instance's pairs
			const int rounds = (hi - lo + 15) >> 4;
			for (int r = 0; r < rounds; r++) {
				const int j = hi - 1 - (r * 16 + hw);
				float g[GR_USED];
#pragma unroll
				for (int i = 0; i < GR_USED; i++) g[i] = 0.f;
				uint32_t row = 0;
				if (j >= lo) {
					const float4 r0 = recs[j * 5], r1 = recs[j * 5 + 1], r2 = recs[j * 5 + 2];
					row = __float_as_uint(recs[j * 5 + 4].w);
					const uint32_t s0 = sm.base[j] - base_lo, c = sm.base[j + 1] - sm.base[j];
					for (uint32_t k = hl; k < c; k += 16) {
						const float4 sl = sm.slot[s0 + k];
						const int p = __float_as_int(sl.w);
						if (p >= 0)
							instance_step(r0, r1, r2, (float)(tx * TILE + (p & 15)), (float)(ty * TILE + (p >> 4)),
							              sl.x, sl.y, sl.z, &sm.pix[0][0], p, g);
					}
				}
				__syncwarp();
				float a, bsum;
				half_reduce18(g, lane, a, bsum);
				if (j >= lo) {
					float* dst = grad_rows + (size_t)row * GRAD_WORDS;
					dst[hl] = a;
					if ((hl & 7) == 0) dst[16 + (hl >> 3)] = bsum;
				}
			}
			hi = lo;
		}
	}
}

}  // namespace

int launch_composite_backward(const gs2d_settings& s, const gs2d_inputs& in, const ImageState& im,
                              const float* instances, const uint32_t* hit_inst, const uint32_t* hit_pix,
                              const float* dL_dcolor, const float* dL_dothers, float* grad_rows, cudaStream_t st)
{
	const int W = s.image_width, H = s.image_height;
	const unsigned gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
	if (gx * gy == 0) return GS2D_OK;
	static bool attr_set = false;
	if (!attr_set) {
		GS2D_CUDA_OK(cudaFuncSetAttribute(composite_backward_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
		                                  (int)sizeof(BwdSmem)));
		GS2D_CUDA_OK(cudaFuncSetAttribute(composite_backward_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
		                                  (int)sizeof(BwdSmem)));
		attr_set = true;
	}
	ProfScope ps(K_COMPOSITE_BWD, st);
	if (s.use_sa)
		composite_backward_kernel<true><<<gx * gy, 256, sizeof(BwdSmem), st>>>(
			W, H, gx, (const uint2*)im.ranges, (const float4*)instances, hit_inst, hit_pix, in.background, im.final_T,
			im.n_contrib, im.median_depth, im.depth_std, dL_dcolor, dL_dothers, grad_rows);
	else
		composite_backward_kernel<false><<<gx * gy, 256, sizeof(BwdSmem), st>>>(
			W, H, gx, (const uint2*)im.ranges, (const float4*)instances, hit_inst, hit_pix, in.background, im.final_T,
			im.n_contrib, im.median_depth, im.depth_std, dL_dcolor, dL_dothers, grad_rows);
	count_launch();
	GS2D_CUDA_OK(cudaGetLastError());
	return GS2D_OK;
}

}  // namespace gs2d
