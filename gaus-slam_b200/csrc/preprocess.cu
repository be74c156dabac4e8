// preprocess.cu -- per-surfel forward preprocess + tile-count scan + mark_visible.
//
// Replaces FORWARD::preprocess / preprocessCUDA<3> (CR/forward.cu:150-253, with
// compute_transmat :75-115, compute_aabb :119-147, in_frustum CR/auxiliary.h:184-209,
// getRect :66-76), cub::DeviceScan::InclusiveSum (CR/rasterizer_impl.cu:283) and
// checkFrustum (CR/rasterizer_impl.cu:54-66).
//
// One thread per surfel, HBM-bound: 40 B read + 8 B written per surfel, +100 B per visible
// surfel (80 B splat record, depth, tile rect, contribution bounds).  The arithmetic that
// decides radii / tile rects / depth bits keeps the reference's operation order so that the
// binning stays bit-exact; everything else (record packing, contribution bounds for warp-level
// culling in the compositing kernels) is new.
#include "common.cuh"
#include "math.cuh"
#include "exact.cuh"

namespace gs2d {

namespace {

__device__ __forceinline__ float3 sh_to_rgb(int idx, int deg, int max_coeffs, float3 pos, float3 campos,
                                            const float* shs, uint8_t* clamped)
{
	float3 dir = make_float3(pos.x - campos.x, pos.y - campos.y, pos.z - campos.z);
	float len = sqrtf(dir.x * dir.x + dir.y * dir.y + dir.z * dir.z);
	dir = make_float3(dir.x / len, dir.y / len, dir.z / len);
	const float3* sh = ((const float3*)shs) + (size_t)idx * max_coeffs;
	float3 res = scale3(kSH0, sh[0]);
	if (deg > 0) {
		float x = dir.x, y = dir.y, z = dir.z;
		res.x = res.x - kSH1 * y * sh[1].x + kSH1 * z * sh[2].x - kSH1 * x * sh[3].x;
		res.y = res.y - kSH1 * y * sh[1].y + kSH1 * z * sh[2].y - kSH1 * x * sh[3].y;
		res.z = res.z - kSH1 * y * sh[1].z + kSH1 * z * sh[2].z - kSH1 * x * sh[3].z;
		if (deg > 1) {
			float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
			float b4 = kSH2[0] * xy, b5 = kSH2[1] * yz, b6 = kSH2[2] * (2.0f * zz - xx - yy);
			float b7 = kSH2[3] * xz, b8 = kSH2[4] * (xx - yy);
			res.x = res.x + b4 * sh[4].x + b5 * sh[5].x + b6 * sh[6].x + b7 * sh[7].x + b8 * sh[8].x;
			res.y = res.y + b4 * sh[4].y + b5 * sh[5].y + b6 * sh[6].y + b7 * sh[7].y + b8 * sh[8].y;
			res.z = res.z + b4 * sh[4].z + b5 * sh[5].z + b6 * sh[6].z + b7 * sh[7].z + b8 * sh[8].z;
			if (deg > 2) {
				float c9 = kSH3[0] * y * (3.0f * xx - yy), c10 = kSH3[1] * xy * z;
				float c11 = kSH3[2] * y * (4.0f * zz - xx - yy);
				float c12 = kSH3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy);
				float c13 = kSH3[4] * x * (4.0f * zz - xx - yy), c14 = kSH3[5] * z * (xx - yy);
				float c15 = kSH3[6] * x * (xx - 3.0f * yy);
				res.x = res.x + c9 * sh[9].x + c10 * sh[10].x + c11 * sh[11].x + c12 * sh[12].x +
				        c13 * sh[13].x + c14 * sh[14].x + c15 * sh[15].x;
				res.y = res.y + c9 * sh[9].y + c10 * sh[10].y + c11 * sh[11].y + c12 * sh[12].y +
				        c13 * sh[13].y + c14 * sh[14].y + c15 * sh[15].y;
				res.z = res.z + c9 * sh[9].z + c10 * sh[10].z + c11 * sh[11].z + c12 * sh[12].z +
				        c13 * sh[13].z + c14 * sh[14].z + c15 * sh[15].z;
			}
		}
	}
	res.x += 0.5f; res.y += 0.5f; res.z += 0.5f;
	clamped[3 * idx + 0] = (res.x < 0);
	clamped[3 * idx + 1] = (res.y < 0);
	clamped[3 * idx + 2] = (res.z < 0);
	return make_float3(fmaxf(res.x, 0.0f), fmaxf(res.y, 0.0f), fmaxf(res.z, 0.0f));
}

// Screen-space bounds (pixels) outside which this surfel cannot reach alpha >= 1/255.
// A pixel contributes only if min(rho3d, rho2d) <= tau with tau = 2 ln(255 * opacity)
// (CR/forward.cu:368-386).  {rho3d <= c^2} is the projection of the disc |uv| <= c, whose
// axis-aligned bounds follow from the same dual-conic formula the reference uses at c = 3
// (CR/forward.cu:119-147), valid while the whole disc stays in front of the eye plane;
// {rho2d <= tau} is a disc of radius < 0.34 px around the stored centre.  The result is
// inflated (2 % + 0.05 on c^2, 1 px on the box) so fp32 rounding in the per-pixel evaluation
// can never put a contributing pixel outside it; when the closed form is not safely valid the
// bounds are infinite (no culling).
__device__ __forceinline__ float4 contribution_bounds(const float3& Tu, const float3& Tv, const float3& Tw,
                                                      float2 centre, float opacity)
{
	const float INF = __int_as_float(0x7f800000);
	float4 all = make_float4(-INF, -INF, INF, INF);
	float a = 255.0f * opacity;
	if (!(a >= 1.0f)) {
		// alpha <= opacity < 1/255 for every pixel (NaN opacity falls through to "all").
		if (opacity == opacity) return make_float4(INF, INF, -INF, -INF);
		return all;
	}
	float c2 = 2.0f * __logf(a) * 1.02f + 0.05f;
	float d = c2 * (Tw.x * Tw.x + Tw.y * Tw.y) - Tw.z * Tw.z;
	if (!(d < -1e-3f * Tw.z * Tw.z) || !(Tw.z > 0.0f)) return all;
	float fx = c2 / d, fz = -1.0f / d;
	float cx = fx * (Tu.x * Tw.x + Tu.y * Tw.y) + fz * Tu.z * Tw.z;
	float cy = fx * (Tv.x * Tw.x + Tv.y * Tw.y) + fz * Tv.z * Tw.z;
	float ex2 = cx * cx - (fx * (Tu.x * Tu.x + Tu.y * Tu.y) + fz * Tu.z * Tu.z);
	float ey2 = cy * cy - (fx * (Tv.x * Tv.x + Tv.y * Tv.y) + fz * Tv.z * Tv.z);
	if (!(ex2 == ex2) || !(ey2 == ey2)) return all;
	float ex = sqrtf(fmaxf(ex2, 0.0f)) * 1.001f + 1.0f + 1e-4f * fabsf(cx);
	float ey = sqrtf(fmaxf(ey2, 0.0f)) * 1.001f + 1.0f + 1e-4f * fabsf(cy);
	float4 r = make_float4(cx - ex, cy - ey, cx + ex, cy + ey);
	// low-pass filter footprint around the stored centre
	r.x = fminf(r.x, centre.x - 0.5f); r.y = fminf(r.y, centre.y - 0.5f);
	r.z = fmaxf(r.z, centre.x + 0.5f); r.w = fmaxf(r.w, centre.y + 0.5f);
	if (!(r.x == r.x) || !(r.y == r.y) || !(r.z == r.z) || !(r.w == r.w)) return all;
	return r;
}

__global__ void __launch_bounds__(256)
preprocess_forward_kernel(int P, int D, int M,
	const float* __restrict__ means3D, const float2* __restrict__ scales, float scale_modifier,
	const float4* __restrict__ rotations, const float* __restrict__ opacities,
	const float* __restrict__ shs, uint8_t* __restrict__ clamped,
	const float* __restrict__ transMat_precomp, const float* __restrict__ colors_precomp,
	const float* __restrict__ viewmatrix, const float* __restrict__ projmatrix,
	const float* __restrict__ campos, int W, int H, unsigned gx, unsigned gy, int prefiltered,
	int* __restrict__ radii, float* __restrict__ depths, float4* __restrict__ splat,
	ushort4* __restrict__ rect, float4* __restrict__ cull, uint32_t* __restrict__ tiles_touched,
	uint32_t* __restrict__ tile_count, uint32_t* __restrict__ zero_words, int n_zero_words)
{
	int idx = blockIdx.x * blockDim.x + threadIdx.x;
	// status words + counters of the tiles_touched scan that follows (instead of a memset launch)
	for (int i = idx; i < n_zero_words; i += gridDim.x * blockDim.x) zero_words[i] = 0u;
	if (idx >= P) return;

	radii[idx] = 0;
	tiles_touched[idx] = 0;

	const float3 p_orig = make_float3(means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2]);
	// all per-surfel inputs are requested up front (one memory round trip; the kernel is latency-bound)
	const bool own_transform = transMat_precomp == nullptr;
	const float2 scale_in = own_transform ? scales[idx] : make_float2(1.f, 1.f);
	const float4 rot_in = own_transform ? rotations[idx] : make_float4(1.f, 0.f, 0.f, 0.f);
	const float opacity = opacities[idx];
	float3 rgb = make_float3(0.f, 0.f, 0.f);
	if (colors_precomp != nullptr)
		rgb = make_float3(colors_precomp[3 * idx], colors_precomp[3 * idx + 1], colors_precomp[3 * idx + 2]);
	const float3 p_view = xform_point_exact(p_orig, viewmatrix);
	if (p_view.z <= 0.2f) {
		if (prefiltered) __trap();
		return;
	}

	float3 Tu, Tv, Tw, normal;
	if (transMat_precomp == nullptr) {
		const SplatFrame fr = splat_frame_exact(p_orig, scale_in, scale_modifier, rot_in, projmatrix, W, H);
		Tu = fr.Tu; Tv = fr.Tv; Tw = fr.Tw;
		normal = xform_vec_exact(fr.L2, viewmatrix);
	} else {
		const float* t = transMat_precomp + 9 * (size_t)idx;
		Tu = make_float3(t[0], t[1], t[2]);
		Tv = make_float3(t[3], t[4], t[5]);
		Tw = make_float3(t[6], t[7], t[8]);
		normal = make_float3(0.0f, 0.0f, 1.0f);
	}

	// two-sided surfels: make the normal face the camera (DUAL_VISIABLE, CR/forward.cu:211-216)
	const float pdotn = dot3_ref(p_view.x, normal.x, p_view.y, normal.y, p_view.z, normal.z);
	if (pdotn == 0.0f) return;
	const float multiplier = pdotn < 0.0f ? 1.0f : -1.0f;
	normal = make_float3(__fmul_rn(normal.x, multiplier), __fmul_rn(normal.y, multiplier), __fmul_rn(normal.z, multiplier));

	// centre + half extent of the 3-sigma bounding box (CR/forward.cu:119-147, :221-232)
	float2 centre, extent;
	if (!aabb_exact(Tu, Tv, Tw, centre, extent)) return;
	const float radius = ceilf(fmaxf(extent.x, extent.y));

	uint2 rmin, rmax;
	tile_rect_exact(centre.x, centre.y, (int)radius, gx, gy, rmin, rmax);
	if ((rmax.x - rmin.x) * (rmax.y - rmin.y) == 0) return;

	if (colors_precomp == nullptr) {
		float3 cp = make_float3(campos[0], campos[1], campos[2]);
		rgb = sh_to_rgb(idx, D, M, p_orig, cp, shs, clamped);
	}

	depths[idx] = p_view.z;
	radii[idx] = (int)radius;
	float4* rec = splat + 5 * (size_t)idx;
	rec[0] = make_float4(Tu.x, Tu.y, Tu.z, centre.x);
	rec[1] = make_float4(Tv.x, Tv.y, Tv.z, centre.y);
	rec[2] = make_float4(Tw.x, Tw.y, Tw.z, opacity);
	rec[3] = make_float4(normal.x, normal.y, normal.z, 0.0f);
	rec[4] = make_float4(rgb.x, rgb.y, rgb.z, 0.0f);
	rect[idx] = make_ushort4((unsigned short)rmin.x, (unsigned short)rmin.y, (unsigned short)rmax.x,
	                         (unsigned short)rmax.y);
	cull[idx] = contribution_bounds(Tu, Tv, Tw, centre, opacity);
	tiles_touched[idx] = (rmax.y - rmin.y) * (rmax.x - rmin.x);
	// instances per tile, for the tile-local sort (tile_sort.cu): one RED per instance
	if (tile_count)
		for (unsigned y = rmin.y; y < rmax.y; y++)
			for (unsigned x = rmin.x; x < rmax.x; x++) atomicAdd(&tile_count[(size_t)(y * gx + x) * TC_STRIDE], 1u);
}

// ---- inclusive scan of tiles_touched: single pass, decoupled look-back -----------------------
// status word: bits 63..62 flag (1 = block aggregate, 2 = inclusive prefix), low 62 bits value.
constexpr unsigned long long ST_AGG = 1ull << 62, ST_PREFIX = 2ull << 62, ST_MASK = (1ull << 62) - 1;

// Exclusive scan of side_n counts by one CTA of SCAN_THREADS threads: ranges[t] = (start, end) -- (0, 0) for an
// empty tile, like the reference's memset -- and cursor[t] = start.  This CTA is the tail of the launch, so it is
// written for latency: a thread owns SIDE_ITEMS consecutive tiles and requests all its counts (one line each, see
// TC_STRIDE) before it uses any; one block scan per SCAN_THREADS * SIDE_ITEMS tiles.
constexpr int SIDE_ITEMS = 16;
__device__ void side_scan_block(int side_n, const uint32_t* __restrict__ cnt, uint2* __restrict__ ranges,
                                uint32_t* __restrict__ cursor, uint32_t* s_warp)
{
	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
	uint32_t carry = 0;
	for (int base = 0; base < side_n; base += SCAN_THREADS * SIDE_ITEMS) {
		const int i0 = base + tid * SIDE_ITEMS;
		uint32_t v[SIDE_ITEMS];
#pragma unroll
		for (int k = 0; k < SIDE_ITEMS; k++) v[k] = (i0 + k) < side_n ? cnt[(size_t)(i0 + k) * TC_STRIDE] : 0u;
		uint32_t tsum = 0;
#pragma unroll
		for (int k = 0; k < SIDE_ITEMS; k++) tsum += v[k];
		uint32_t incl = tsum;
#pragma unroll
		for (int o = 1; o < 32; o <<= 1) {
			const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
			if (lane >= o) incl += t;
		}
		__syncthreads();   // s_warp of the previous round is consumed
		if (lane == 31) s_warp[warp] = incl;
		__syncthreads();
		uint32_t woff = 0, total = 0;
#pragma unroll
		for (int w = 0; w < SCAN_THREADS / 32; w++) {
			const uint32_t t = s_warp[w];
			if (w < warp) woff += t;
			total += t;
		}
		uint32_t start = carry + woff + incl - tsum;
#pragma unroll
		for (int k = 0; k < SIDE_ITEMS; k++) {
			if (i0 + k < side_n) {
				ranges[i0 + k] = v[k] ? make_uint2(start, start + v[k]) : make_uint2(0u, 0u);
				cursor[(size_t)(i0 + k) * TC_STRIDE] = start;
			}
			start += v[k];
		}
		carry += total;
	}
}

__global__ void __launch_bounds__(SCAN_THREADS)
scan_tiles_kernel(const uint32_t* __restrict__ in, uint32_t* __restrict__ out, int n,
                  unsigned long long* status, uint32_t* counters, int side_n, const uint32_t* __restrict__ side_in,
                  uint2* __restrict__ side_ranges, uint32_t* __restrict__ side_cursor)
{
	__shared__ uint32_t s_block;
	__shared__ uint32_t s_warp[SCAN_THREADS / 32];
	__shared__ unsigned long long s_prefix;
	if (threadIdx.x == 0) s_block = atomicAdd(&counters[0], 1u);
	__syncthreads();
	const uint32_t bid = s_block;
	if ((long long)bid * SCAN_TILE >= n) {
		// The one extra block of the launch: exclusive scan of the (few thousand) per-tile instance counts ->
		// tile ranges + scatter cursors (tile_sort.cu), done beside the big scan instead of in its own launch.
		side_scan_block(side_n, side_in, side_ranges, side_cursor, s_warp);
		return;
	}
	const int base = bid * SCAN_TILE + threadIdx.x * SCAN_ITEMS;

	uint32_t v[SCAN_ITEMS];
	uint32_t tsum = 0;
#pragma unroll
	for (int i = 0; i < SCAN_ITEMS; i++) v[i] = (base + i < n) ? in[base + i] : 0u;
#pragma unroll
	for (int i = 0; i < SCAN_ITEMS; i++) tsum += v[i];
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	uint32_t incl = tsum;
#pragma unroll
	for (int o = 1; o < 32; o <<= 1) {
		uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
		if (lane >= o) incl += t;
	}
	if (lane == 31) s_warp[warp] = incl;
	__syncthreads();
	uint32_t warp_off = 0, block_total = 0;
#pragma unroll
	for (int w = 0; w < SCAN_THREADS / 32; w++) {
		uint32_t t = s_warp[w];
		if (w < warp) warp_off += t;
		block_total += t;
	}
	if (warp == 0) {
		// warp-wide look-back: lane l inspects predecessor pb - l; a window is consumed up to its first inclusive prefix
		unsigned long long excl = 0;
		if (bid == 0) {
			if (lane == 0) atomicExch(&status[0], ST_PREFIX | (unsigned long long)block_total);
		} else {
			if (lane == 0) atomicExch(&status[bid], ST_AGG | (unsigned long long)block_total);
			int pb = (int)bid - 1;
			while (true) {
				const int idx = pb - lane;
				unsigned long long sv = 2ull << 62;  // before block 0: an inclusive prefix of 0
				if (idx >= 0) sv = *((volatile unsigned long long*)&status[idx]);
				const uint32_t flag = (uint32_t)(sv >> 62);
				const uint32_t ready = __ballot_sync(0xffffffffu, flag != 0);
				const uint32_t pref = __ballot_sync(0xffffffffu, flag == 2);
				const int last = pref ? __ffs(pref) - 1 : 31;
				const uint32_t need = last == 31 ? 0xffffffffu : ((2u << last) - 1u);
				if ((ready & need) != need) continue;
				unsigned long long part = lane <= last ? (sv & ST_MASK) : 0ull;
#pragma unroll
				for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
				excl += part;
				if (pref) break;
				pb -= 32;
			}
			if (lane == 0) atomicExch(&status[bid], ST_PREFIX | (excl + block_total));
		}
		if (lane == 0) {
			s_prefix = excl;
			if ((long long)(bid + 1) * SCAN_TILE >= n) counters[1] = (uint32_t)(excl + block_total);
		}
	}
	__syncthreads();
	uint32_t run = (uint32_t)s_prefix + warp_off + (incl - tsum);
#pragma unroll
	for (int i = 0; i < SCAN_ITEMS; i++) {
		run += v[i];
		if (base + i < n) out[base + i] = run;
	}
}

__global__ void mark_visible_kernel(int P, const float* __restrict__ means3D,
                                    const float* __restrict__ viewmatrix, uint8_t* __restrict__ present)
{
	int idx = blockIdx.x * blockDim.x + threadIdx.x;
	if (idx >= P) return;
	float3 p = make_float3(means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2]);
	float3 pv = xform_point_exact(p, viewmatrix);
	present[idx] = pv.z <= 0.2f ? 0 : 1;
}

}  // namespace

int launch_preprocess_forward(const gs2d_settings& s, const gs2d_inputs& in, const GeomState& g,
                              uint32_t* tile_count, int32_t* radii, cudaStream_t st)
{
	const int P = s.P;
	const unsigned gx = (s.image_width + TILE - 1) / TILE, gy = (s.image_height + TILE - 1) / TILE;
	{
		ProfScope ps(K_PREPROCESS_FWD, st);
		preprocess_forward_kernel<<<(P + 255) / 256, 256, 0, st>>>(
			P, s.sh_degree, s.sh_coeffs, in.means3D, (const float2*)in.scales, s.scale_modifier,
			(const float4*)in.rotations, in.opacities, in.shs, g.clamped, in.transMat_precomp,
			in.colors_precomp, in.viewmatrix, in.projmatrix, in.campos, s.image_width, s.image_height,
			gx, gy, s.prefiltered, radii, g.depths, (float4*)g.splat, (ushort4*)g.rect, (float4*)g.cull,
			g.tiles_touched, tile_count, (uint32_t*)g.scan_status,
			(int)(((char*)(g.counters + 8) - (char*)g.scan_status) / sizeof(uint32_t)));
		count_launch();
	}
	GS2D_CUDA_OK(cudaGetLastError());
	return GS2D_OK;
}

// Inclusive scan of n u32 values (single pass, decoupled look-back).  status: n / SCAN_TILE + 1 u64 words,
// counters: 8 u32 words (both zeroed here); counters[1] receives the total.
int launch_scan_u32(const uint32_t* in, uint32_t* out, long long n, uint64_t* status, uint32_t* counters,
                    cudaStream_t st, bool clear, int side_n, const uint32_t* side_in, uint32_t* side_ranges,
                    uint32_t* side_cursor)
{
	const int blocks = (int)((n + SCAN_TILE - 1) / SCAN_TILE);
	if (clear) {
		{ ProfScope ps(K_MEMSET, st); GS2D_CUDA_OK(cudaMemsetAsync(status, 0, sizeof(uint64_t) * (size_t)(blocks + 1), st)); }
		{ ProfScope ps(K_MEMSET, st); GS2D_CUDA_OK(cudaMemsetAsync(counters, 0, sizeof(uint32_t) * 8, st)); }
	}
	const int extra = side_n > 0 ? 1 : 0;
	if (blocks + extra > 0) {
		ProfScope ps(K_SCAN, st);
		scan_tiles_kernel<<<blocks + extra, SCAN_THREADS, 0, st>>>(in, out, (int)n, (unsigned long long*)status, counters,
		                                                          side_n, side_in, (uint2*)side_ranges, side_cursor);
		count_launch();
	}
	GS2D_CUDA_OK(cudaGetLastError());
	return GS2D_OK;
}

int launch_scan_tiles(const GeomState& g, int P, cudaStream_t st, int tiles, const uint32_t* tile_count,
                      uint32_t* ranges, uint32_t* tile_cursor)
{
	// status words and counters (adjacent in the geom buffer) were zeroed by preprocess_forward_kernel
	return launch_scan_u32(g.tiles_touched, g.point_offsets, P, g.scan_status, g.counters, st, false, tiles, tile_count,
	                       ranges, tile_cursor);
}

int launch_mark_visible(int P, const float* means3D, const float* view, const float* proj,
                        uint8_t* present, cudaStream_t st)
{
	(void)proj;
	{
		ProfScope ps(K_MARK_VISIBLE, st);
		mark_visible_kernel<<<(P + 255) / 256, 256, 0, st>>>(P, means3D, view, present);
		count_launch();
	}
	GS2D_CUDA_OK(cudaGetLastError());
	return GS2D_OK;
}

}  // namespace gs2d
