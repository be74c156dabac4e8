// composite_fwd.cu -- per-tile front-to-back compositing (forward).
//
// Replaces FORWARD::render / renderCUDA<3> (CR/forward.cu:258-467): ray-splat intersection,
// low-pass filter, alpha, early ray termination, surface-aware (SA) depth or 2DGS distortion,
// normals, median depth.  The per-(pixel, instance) arithmetic keeps the reference's operation
// order, so a pixel's result is the same sequence of fp32 operations; what changes is how the
// work is fed:
//   * one CTA per 16x16 tile, each warp owns an 8x4 pixel block;
//   * the tile's instance records (80 B each, already in sorted order and contiguous in HBM)
//     are staged with one TMA bulk copy per 128-instance batch into a two-stage shared-memory
//     ring guarded by mbarriers, so the copy of batch b+1 overlaps the math of batch b;
//   * every record carries an 8-bit block mask (which 8x4 blocks its >= 1/255-alpha footprint can
//     touch); each warp ballots 32 masks at a time and only visits the instances whose bit is
//     set -- a skipped instance provably changes no pixel state, so results are unchanged;
//   * a warp whose 32 pixels have all terminated stops evaluating; the CTA stops staging when
//     all 256 have.
// Bound by FP32 issue + MUFU (exp, reciprocal) on the surviving pairs; HBM traffic is
// 80 B per instance + 68 B per pixel.
#include "common.cuh"
#include "tma.cuh"

namespace gs2d {

namespace {

constexpr int FWD_BATCH = 128;
constexpr int FWD_STAGES = 2;

struct PixelFwd {
	float T = 1.0f;
	float C0 = 0.f, C1 = 0.f, C2 = 0.f;
	float N0 = 0.f, N1 = 0.f, N2 = 0.f;
	float D = 0.f, M1 = 0.f, M2 = 0.f, D2 = 0.f, distortion = 0.f;
	float median_depth = 0.f;
	uint32_t median_contributor = 0;   // 0 = none (the reference stores float(-1) -> 0u)
	uint32_t last_contributor = 0;
};

template <bool USE_SA>
__device__ __forceinline__ void eval_pair(const float4* __restrict__ rec, float pxf, float pyf,
                                          uint32_t contributor, PixelFwd& px, bool& done)
{
	const float4 r0 = rec[0], r1 = rec[1], r2 = rec[2];
	const float3 Tu = make_float3(r0.x, r0.y, r0.z);
	const float3 Tv = make_float3(r1.x, r1.y, r1.z);
	const float3 Tw = make_float3(r2.x, r2.y, r2.z);
	const float2 xy = make_float2(r0.w, r1.w);
	const float opa = r2.w;

	// two homogeneous planes through the pixel, intersected with the splat plane
	const float3 k = make_float3(pxf * Tw.x - Tu.x, pxf * Tw.y - Tu.y, pxf * Tw.z - Tu.z);
	const float3 l = make_float3(pyf * Tw.x - Tv.x, pyf * Tw.y - Tv.y, pyf * Tw.z - Tv.z);
	const float3 p = make_float3(k.y * l.z - k.z * l.y, k.z * l.x - k.x * l.z, k.x * l.y - k.y * l.x);
	if (p.z == 0.0) return;
	const float2 s = make_float2(p.x / p.z, p.y / p.z);
	const float rho3d = (s.x * s.x + s.y * s.y);
	const float2 d = make_float2(xy.x - pxf, xy.y - pyf);
	const float rho2d = GS2D_FILTER_INV_SQUARE * (d.x * d.x + d.y * d.y);

	const float rho = fminf(rho3d, rho2d);
	float depth = (rho3d <= rho2d) ? (s.x * Tw.x + s.y * Tw.y) + Tw.z : Tw.z;
	if (depth < GS2D_NEAR) return;

	const float power = -0.5f * rho;
	if (power > 0.0f) return;
	const float alpha = fminf(0.99f, opa * expf(power));
	if (alpha < 1.0f / 255.0f) return;
	const float test_T = px.T * (1 - alpha);
	if (test_T < 0.0001f) {
		done = true;
		return;
	}
	const float w = alpha * px.T;
	if (px.T > 0.5) {
		px.median_depth = depth;
		px.median_contributor = contributor;
	}

	if (USE_SA) {
		if (px.D > 0) {
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
}

template <bool USE_SA>
__global__ void __launch_bounds__(256)
composite_forward_kernel(int W, int H, unsigned gx, const uint2* __restrict__ ranges,
                         const float4* __restrict__ instances, const float* __restrict__ bg,
                         float* __restrict__ final_T, uint32_t* __restrict__ n_contrib,
                         float* __restrict__ median_depth_out, float* __restrict__ depth_std_out,
                         float* __restrict__ out_color, float* __restrict__ out_others)
{
	__shared__ __align__(128) float4 s_rec[FWD_STAGES][FWD_BATCH * 5];
	__shared__ __align__(8) uint64_t s_bar[FWD_STAGES];

	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
	const unsigned tile = blockIdx.x;
	const unsigned tx = tile % gx, ty = tile / gx;
	const unsigned pix_x = tx * TILE + (warp & 1) * 8 + (lane & 7);
	const unsigned pix_y = ty * TILE + (warp >> 1) * 4 + (lane >> 3);
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
	const float4* tile_src = instances + 5 * (size_t)range.x;
	if (tid == 0 && nb > 0) {
		const uint32_t bytes = (uint32_t)min(n, FWD_BATCH) * 80u;
		mbar_arrive_expect_tx(&s_bar[0], bytes);
		tma_load_1d(&s_rec[0][0], tile_src, bytes, &s_bar[0]);
	}

	PixelFwd px;
	bool done = !inside;

	for (int b = 0; b < nb; b++) {
		const int stage = b & 1;
		const uint32_t parity = (uint32_t)(b >> 1) & 1u;
		// everyone is finished with the other stage (batch b-1); also the CTA-wide exit vote
		const int num_done = __syncthreads_count(done);
		if (num_done == TILE_PIXELS) {
			mbar_wait(&s_bar[stage], parity);   // do not exit with a bulk copy in flight
			break;
		}
		if (tid == 0 && b + 1 < nb) {
			const int cnt = min(n - (b + 1) * FWD_BATCH, FWD_BATCH);
			const uint32_t bytes = (uint32_t)cnt * 80u;
			mbar_arrive_expect_tx(&s_bar[stage ^ 1], bytes);
			tma_load_1d(&s_rec[stage ^ 1][0], tile_src + 5 * (size_t)(b + 1) * FWD_BATCH, bytes, &s_bar[stage ^ 1]);
		}
		mbar_wait(&s_bar[stage], parity);

		if (__all_sync(0xffffffffu, done)) continue;
		const int cnt = min(n - b * FWD_BATCH, FWD_BATCH);
		const float4* recs = &s_rec[stage][0];
		for (int g = 0; g < cnt; g += 32) {
			const int mine = g + lane;
			bool want = false;
			if (mine < cnt) want = (__float_as_uint(recs[mine * 5 + 3].w) >> warp) & 1u;
			uint32_t bal = __ballot_sync(0xffffffffu, want);
			while (bal) {
				const int j = g + (__ffs(bal) - 1);
				bal &= bal - 1;
				if (!done) eval_pair<USE_SA>(recs + j * 5, pxf, pyf, (uint32_t)(b * FWD_BATCH + j + 1), px, done);
			}
			if (__all_sync(0xffffffffu, done)) break;
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

}  // namespace

int launch_composite_forward(const gs2d_settings& s, const gs2d_inputs& in, const ImageState& im,
                             const float* instances, float* out_color, float* out_others, cudaStream_t st)
{
	const int W = s.image_width, H = s.image_height;
	const unsigned gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
	if (gx * gy == 0) return GS2D_OK;
	ProfScope ps(K_COMPOSITE_FWD, st);
	if (s.use_sa)
		composite_forward_kernel<true><<<gx * gy, 256, 0, st>>>(
			W, H, gx, (const uint2*)im.ranges, (const float4*)instances, in.background, im.final_T,
			im.n_contrib, im.median_depth, im.depth_std, out_color, out_others);
	else
		composite_forward_kernel<false><<<gx * gy, 256, 0, st>>>(
			W, H, gx, (const uint2*)im.ranges, (const float4*)instances, in.background, im.final_T,
			im.n_contrib, im.median_depth, im.depth_std, out_color, out_others);
	count_launch();
	GS2D_CUDA_OK(cudaGetLastError());
	return GS2D_OK;
}

}  // namespace gs2d
