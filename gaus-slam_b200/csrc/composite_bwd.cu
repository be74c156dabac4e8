// composite_bwd.cu -- per-tile back-to-front re-traversal (backward of the compositing).
//
// Replaces BACKWARD::render / renderCUDA<3> (CR/backward.cu:143-463).  Same per-(pixel, instance)
// gradient algebra (including the reference's deliberate non-analytic parts: SA confidence from
// the *final* per-pixel statistics, opacity gradient through the clamped alpha), different
// data flow:
//   * instance records staged by TMA bulk copies, newest-first, two-stage mbarrier ring;
//   * per-warp ballot over the records' 8-bit block masks, so a warp only re-evaluates
//     instances that can touch its 8x4 pixels, and only those not behind the warp's deepest
//     last contributor;
//   * the 18 per-pair gradient terms are summed across the warp with a shuffle reduce-scatter
//     (21 SHFL) and leave the SM as ONE 18-lane RED per (warp, instance) into the instance's own
//     80-byte gradient row -- the reference issues 16 scattered atomicAdd per contributing
//     *pixel* (CR/backward.cu:343, :396, :441-449, :454-456, :460).  Rows are indexed by the
//     instance's position in surfel-major (unsorted) order, so the per-surfel reduction in
//     preprocess_bwd.cu reads contiguous rows and needs no atomics at all.
#include "common.cuh"
#include "tma.cuh"

namespace gs2d {

namespace {

constexpr int BWD_BATCH = 128;
constexpr int BWD_STAGES = 2;

struct PixelConst {
	float T_final, final_A, final_D, final_D2;
	float dpix0, dpix1, dpix2;
	float dL_ddepth, dL_daccum, dL_dreg, dL_dmedian;
	float dn0, dn1, dn2;
	float mm, mstd, bg_dot;
	uint32_t last_contributor, median_contributor;
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

// Evaluates one (pixel, instance) pair; on contribution fills g[0..17] and returns true.
template <bool USE_SA>
__device__ __forceinline__ bool grad_pair(const float4* __restrict__ rec, float pxf, float pyf, uint32_t q,
                                          const PixelConst& pc, PixelBwd& st, float (&g)[GR_USED])
{
	const float4 r0 = rec[0], r1 = rec[1], r2 = rec[2];
	const float3 Tu = make_float3(r0.x, r0.y, r0.z);
	const float3 Tv = make_float3(r1.x, r1.y, r1.z);
	const float3 Tw = make_float3(r2.x, r2.y, r2.z);
	const float2 xy = make_float2(r0.w, r1.w);
	const float opa = r2.w;

	const float3 k = make_float3(pxf * Tw.x - Tu.x, pxf * Tw.y - Tu.y, pxf * Tw.z - Tu.z);
	const float3 l = make_float3(pyf * Tw.x - Tv.x, pyf * Tw.y - Tv.y, pyf * Tw.z - Tv.z);
	const float3 p = make_float3(k.y * l.z - k.z * l.y, k.z * l.x - k.x * l.z, k.x * l.y - k.y * l.x);
	if (p.z == 0.0) return false;
	const float2 s = make_float2(p.x / p.z, p.y / p.z);
	const float rho3d = (s.x * s.x + s.y * s.y);
	const float2 d = make_float2(xy.x - pxf, xy.y - pyf);
	const float rho2d = GS2D_FILTER_INV_SQUARE * (d.x * d.x + d.y * d.y);
	const float rho = fminf(rho3d, rho2d);
	float c_d = (rho3d <= rho2d) ? (s.x * Tw.x + s.y * Tw.y) + Tw.z : Tw.z;
	if (c_d < GS2D_NEAR) return false;
	const float power = -0.5f * rho;
	if (power > 0.0f) return false;
	const float G = expf(power);
	const float alpha = fminf(0.99f, opa * G);
	if (alpha < 1.0f / 255.0f) return false;

	const float4 r3 = rec[3], r4 = rec[4];
	st.T = st.T / (1.f - alpha);
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
	g[GR_COL + 0] = w * pc.dpix0;
	g[GR_COL + 1] = w * pc.dpix1;
	g[GR_COL + 2] = w * pc.dpix2;

	float conf = 1;
	if (USE_SA) {
		// confidence from the FINAL per-pixel statistics, treated as a constant (CR/backward.cu:349)
		conf = T < 0.5f ? expf(-(c_d - pc.mm) * (c_d - pc.mm) / (4 * fmaxf(pc.mstd / (1 - pc.T_final), 1e-7f))) : 1.f;
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
	g[GR_NRM + 0] = alpha * T * pc.dn0;
	g[GR_NRM + 1] = alpha * T * pc.dn1;
	g[GR_NRM + 2] = alpha * T * pc.dn2;

	dL_dalpha *= T;
	st.last_alpha = alpha;
	dL_dalpha += (-pc.T_final / (1.f - alpha)) * pc.bg_dot;

	const float dL_dG = opa * dL_dalpha;
	dL_dz += conf * alpha * T * pc.dL_ddepth;

	if (rho3d <= rho2d) {
		const float dL_dsx = dL_dG * -G * s.x + dL_dz * Tw.x;
		const float dL_dsy = dL_dG * -G * s.y + dL_dz * Tw.y;
		const float dsx_pz = dL_dsx / p.z;
		const float dsy_pz = dL_dsy / p.z;
		const float3 dL_dp = make_float3(dsx_pz, dsy_pz, -(dsx_pz * s.x + dsy_pz * s.y));
		const float3 dL_dk = make_float3(l.y * dL_dp.z - l.z * dL_dp.y, l.z * dL_dp.x - l.x * dL_dp.z,
		                                 l.x * dL_dp.y - l.y * dL_dp.x);
		const float3 dL_dl = make_float3(dL_dp.y * k.z - dL_dp.z * k.y, dL_dp.z * k.x - dL_dp.x * k.z,
		                                 dL_dp.x * k.y - dL_dp.y * k.x);
		g[0] = -dL_dk.x; g[1] = -dL_dk.y; g[2] = -dL_dk.z;
		g[3] = -dL_dl.x; g[4] = -dL_dl.y; g[5] = -dL_dl.z;
		g[6] = pxf * dL_dk.x + pyf * dL_dl.x + dL_dz * s.x;
		g[7] = pxf * dL_dk.y + pyf * dL_dl.y + dL_dz * s.y;
		g[8] = pxf * dL_dk.z + pyf * dL_dl.z + dL_dz * 1.0f;
		g[GR_M2D + 0] = 0.f;
		g[GR_M2D + 1] = 0.f;
	} else {
		const float dG_ddelx = -G * GS2D_FILTER_INV_SQUARE * d.x;
		const float dG_ddely = -G * GS2D_FILTER_INV_SQUARE * d.y;
#pragma unroll
		for (int i = 0; i < 8; i++) g[i] = 0.f;
		g[8] = dL_dz;
		g[GR_M2D + 0] = dL_dG * dG_ddelx;
		g[GR_M2D + 1] = dL_dG * dG_ddely;
	}
	g[GR_OPA] = G * dL_dalpha;
	return true;
}

// Sum g[0..17] over the 32 lanes.  On return lane L (L < 16) holds the total of g[L] in `a`,
// and lanes with bit 4 clear / set hold the totals of g[16] / g[17] in `b`.
__device__ __forceinline__ void warp_reduce18(float (&g)[GR_USED], int lane, float& a, float& b)
{
	// values 0..15: recursive halving over lane bits 3..0, then fold the two half-warps
	{
		const bool up = lane & 8;
#pragma unroll
		for (int i = 0; i < 8; i++) {
			float send = up ? g[i] : g[i + 8];
			float keep = up ? g[i + 8] : g[i];
			g[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
		}
	}
	{
		const bool up = lane & 4;
#pragma unroll
		for (int i = 0; i < 4; i++) {
			float send = up ? g[i] : g[i + 4];
			float keep = up ? g[i + 4] : g[i];
			g[i] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
		}
	}
	{
		const bool up = lane & 2;
#pragma unroll
		for (int i = 0; i < 2; i++) {
			float send = up ? g[i] : g[i + 2];
			float keep = up ? g[i + 2] : g[i];
			g[i] = keep + __shfl_xor_sync(0xffffffffu, send, 2);
		}
	}
	{
		const bool up = lane & 1;
		float send = up ? g[0] : g[1];
		float keep = up ? g[1] : g[0];
		g[0] = keep + __shfl_xor_sync(0xffffffffu, send, 1);
	}
	a = g[0] + __shfl_xor_sync(0xffffffffu, g[0], 16);
	// values 16, 17
	{
		const bool up = lane & 16;
		float send = up ? g[16] : g[17];
		float keep = up ? g[17] : g[16];
		float v = keep + __shfl_xor_sync(0xffffffffu, send, 16);
#pragma unroll
		for (int o = 8; o >= 1; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
		b = v;
	}
}

template <bool USE_SA>
__global__ void __launch_bounds__(256)
composite_backward_kernel(int W, int H, unsigned gx, const uint2* __restrict__ ranges,
                          const float4* __restrict__ instances, const float* __restrict__ bg,
                          const float* __restrict__ final_T, const uint32_t* __restrict__ n_contrib,
                          const float* __restrict__ median_depth, const float* __restrict__ depth_std,
                          const float* __restrict__ dL_dpix, const float* __restrict__ dL_dothers,
                          float* __restrict__ grad_rows)
{
	__shared__ __align__(128) float4 s_rec[BWD_STAGES][BWD_BATCH * 5];
	__shared__ __align__(8) uint64_t s_bar[BWD_STAGES];
	__shared__ uint32_t s_max_last[8];

	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
	const unsigned tile = blockIdx.x;
	const unsigned tx = tile % gx, ty = tile / gx;
	const unsigned pix_x = tx * TILE + (warp & 1) * 8 + (lane & 7);
	const unsigned pix_y = ty * TILE + (warp >> 1) * 4 + (lane >> 3);
	const bool inside = pix_x < (unsigned)W && pix_y < (unsigned)H;
	const float pxf = (float)pix_x, pyf = (float)pix_y;
	const size_t HW = (size_t)H * W;
	const size_t pid = (size_t)W * pix_y + pix_x;

	const uint2 range = ranges[tile];
	const int n = (int)(range.y - range.x);

	PixelConst pc;
	pc.T_final = inside ? final_T[pid] : 0.f;
	pc.last_contributor = inside ? n_contrib[pid] : 0u;
	pc.median_contributor = inside ? n_contrib[pid + HW] : 0u;
	pc.final_A = 1 - pc.T_final;
	pc.final_D = inside ? final_T[pid + HW] : 0.f;
	pc.final_D2 = inside ? final_T[pid + 2 * HW] : 0.f;
	pc.dpix0 = pc.dpix1 = pc.dpix2 = 0.f;
	pc.dL_ddepth = pc.dL_daccum = pc.dL_dreg = pc.dL_dmedian = 0.f;
	pc.dn0 = pc.dn1 = pc.dn2 = 0.f;
	pc.mm = pc.mstd = 0.f;
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
		pc.mstd = depth_std[pid];
	}
	{
		float acc = 0;
		acc += bg[0] * pc.dpix0;
		acc += bg[1] * pc.dpix1;
		acc += bg[2] * pc.dpix2;
		pc.bg_dot = acc;
	}
	PixelBwd st;
	st.T = pc.T_final;

	// deepest last contributor of the warp / of the tile: nothing behind it is ever visited
	uint32_t wmax = pc.last_contributor;
#pragma unroll
	for (int o = 16; o >= 1; o >>= 1) wmax = max(wmax, __shfl_xor_sync(0xffffffffu, wmax, o));
	if (lane == 0) s_max_last[warp] = wmax;
	if (tid == 0) {
		for (int s = 0; s < BWD_STAGES; s++) mbar_init(&s_bar[s], 1);
		mbar_fence_init();
	}
	__syncthreads();
	uint32_t bmax = 0;
#pragma unroll
	for (int i = 0; i < 8; i++) bmax = max(bmax, s_max_last[i]);
	const int n_used = min((int)bmax, n);            // instances [0, n_used) can matter
	const int nb = (n_used + BWD_BATCH - 1) / BWD_BATCH;
	if (nb == 0) return;

	const float4* tile_src = instances + 5 * (size_t)range.x;
	auto issue = [&](int b, int stage) {
		const int cnt = min(n_used - b * BWD_BATCH, BWD_BATCH);
		const uint32_t bytes = (uint32_t)cnt * 80u;
		mbar_arrive_expect_tx(&s_bar[stage], bytes);
		tma_load_1d(&s_rec[stage][0], tile_src + 5 * (size_t)b * BWD_BATCH, bytes, &s_bar[stage]);
	};
	if (tid == 0) issue(nb - 1, 0);

	for (int it = 0; it < nb; it++) {
		const int b = nb - 1 - it;
		const int stage = it & 1;
		const uint32_t parity = (uint32_t)(it >> 1) & 1u;
		__syncthreads();   // batch it-1 fully consumed -> its stage may be overwritten
		if (tid == 0 && it + 1 < nb) issue(b - 1, stage ^ 1);
		mbar_wait(&s_bar[stage], parity);

		const int cnt = min(n_used - b * BWD_BATCH, BWD_BATCH);
		const float4* recs = &s_rec[stage][0];
		const int q0 = b * BWD_BATCH;
		if ((uint32_t)q0 >= wmax) continue;
		for (int g0 = ((cnt - 1) / 32) * 32; g0 >= 0; g0 -= 32) {
			const int mine = g0 + lane;
			bool want = false;
			if (mine < cnt && (uint32_t)(q0 + mine) < wmax)
			{
				const uint32_t bw = __float_as_uint(recs[mine * 5 + 3].w);
				const int bx0 = (warp & 1) * 8, by0 = (warp >> 1) * 4;
				want = ((bw >> 16) & 1u) && (int)(bw & 15) <= bx0 + 7 && (int)((bw >> 8) & 15) >= bx0 &&
				       (int)((bw >> 4) & 15) <= by0 + 3 && (int)((bw >> 12) & 15) >= by0;
			}
			uint32_t bal = __ballot_sync(0xffffffffu, want);
			while (bal) {
				const int jl = 31 - __clz(bal);
				bal &= ~(1u << jl);
				const int j = g0 + jl;
				const uint32_t q = (uint32_t)(q0 + j);
				float g[GR_USED];
				bool hit = false;
				if (q < pc.last_contributor) hit = grad_pair<USE_SA>(recs + j * 5, pxf, pyf, q, pc, st, g);
				if (!__any_sync(0xffffffffu, hit)) continue;
				if (!hit) {
#pragma unroll
					for (int i = 0; i < GR_USED; i++) g[i] = 0.f;
				}
				float a, bsum;
				warp_reduce18(g, lane, a, bsum);
				const uint32_t row = __float_as_uint(recs[j * 5 + 4].w);
				float* dst = grad_rows + (size_t)row * GRAD_WORDS;
				if (lane < 16) atomicAdd(dst + lane, a);
				if ((lane & 15) == 0) atomicAdd(dst + 16 + (lane >> 4), bsum);
			}
		}
	}
}

}  // namespace

int launch_composite_backward(const gs2d_settings& s, const gs2d_inputs& in, const ImageState& im,
                              const float* instances, const uint32_t* hit_inst, const float* dL_dcolor,
                              const float* dL_dothers, float* grad_rows, cudaStream_t st)
{
	const int W = s.image_width, H = s.image_height;
	const unsigned gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
	if (gx * gy == 0) return GS2D_OK;
	ProfScope ps(K_COMPOSITE_BWD, st);
	if (s.use_sa)
		composite_backward_kernel<true><<<gx * gy, 256, 0, st>>>(
			W, H, gx, (const uint2*)im.ranges, (const float4*)instances, in.background, im.final_T,
			im.n_contrib, im.median_depth, im.depth_std, dL_dcolor, dL_dothers, grad_rows);
	else
		composite_backward_kernel<false><<<gx * gy, 256, 0, st>>>(
			W, H, gx, (const uint2*)im.ranges, (const float4*)instances, in.background, im.final_T,
			im.n_contrib, im.median_depth, im.depth_std, dL_dcolor, dL_dothers, grad_rows);
	count_launch();
	GS2D_CUDA_OK(cudaGetLastError());
	return GS2D_OK;
}

}  // namespace gs2d
