// loss.cu -- fused mapping loss + image gradients (the step right after the rasterizer).
//
// Replaces, for the mapping branch of get_loss (slam/Loss.py:22-25, :36, :51-63) together with the weight
// normalisation of the rendered depth that precedes it (render/__init__.py:46-49), about 25 elementwise /
// reduction launches of torch (+ as many in autograd) by two kernels:
//
//   depth  = allmap[0] / (allmap[1] + eps);  depth = 0 where depth > far or depth < near       (render/__init__.py)
//   x      = nan_to_num(x, 0, 0) for colour, depth, dist (NaN -> 0, +inf -> 0, -inf -> -FLT_MAX)   (Loss.py:22-25)
//   mask   = gt_depth > 1e-5  and  depth > 1e-5                                                  (Loss.py:36)
//   loss   = w_color * mean_{mask, 3 channels} |colour - gt_colour|
//          + w_depth * mean_{mask} |depth - gt_depth|  +  w_dist * mean_{mask} allmap[6]         (Loss.py:52-63)
//
// (use_edge_growth = False, use_normal_loss = False, enable_exposure = False: the defaults of every shipped
// config; the other branches are not built.)  Forward: one pass, per-block partial sums, the last block adds
// them in a fixed order (deterministic) and leaves {loss, count, the three sums} in device memory.  Backward: one
// pass that recomputes the mask and writes dL/dcolour [3,H,W] and dL/dallmap [7,H,W] (channels 2..5 get zeros),
// scaled by the upstream gradient read from device memory -- nothing goes through the host.
// Both kernels are pure HBM streams: 40 B read per pixel forward, 40 B read + 40 B written backward.
#include "common.cuh"

#include <cfloat>
#include <initializer_list>

namespace gs2d {

namespace {

struct LossPixel {
	float c[3], depth, dist, gc[3], gd;
	bool mask, mask_c, c_ok[3], dist_ok;   // mask: depth term; mask_c: colour and distortion terms
	float inv_a;     // 1 / (alpha + eps)
};

__device__ __forceinline__ float nan_to_num00(float v, bool& passes)
{
	// torch.nan_to_num(v, 0, 0): nan = 0, posinf = 0, neginf = default (lowest finite); gradient passes only
	// where the value was finite
	passes = isfinite(v);
	if (passes) return v;
	return (v < 0.f) ? -FLT_MAX : 0.f;
}

__device__ __forceinline__ LossPixel load_pixel(size_t p, size_t HW, const float* __restrict__ color,
                                                const float* __restrict__ allmap, const float* __restrict__ gt_color,
                                                const float* __restrict__ gt_depth, int gt_hwc, float eps, float near_,
                                                float far_, float silmask_th, float edge_thres)
{
	LossPixel px;
#pragma unroll
	for (int k = 0; k < 3; k++) {
		px.c[k] = nan_to_num00(color[k * HW + p], px.c_ok[k]);
		px.gc[k] = gt_hwc ? gt_color[3 * p + k] : gt_color[k * HW + p];
	}
	px.gd = gt_depth[p];
	const float D = allmap[p], A = allmap[HW + p];
	px.inv_a = 1.0f / (A + eps);
	float d = D * px.inv_a;
	const bool outlier = (d > far_) || (d < near_);
	if (outlier) d = 0.f;
	bool d_ok;
	d = nan_to_num00(d, d_ok);
	px.depth = d;
	px.dist = nan_to_num00(allmap[6 * HW + p], px.dist_ok);
	px.mask = (px.gd > 1e-5f) && (d > 1e-5f);   // implies: finite, not an outlier
	// tracking branch (slam/Loss.py:42-43): nan_to_num(alpha, 0, 0) > silmask_th as well
	if (silmask_th >= 0.f) px.mask = px.mask && isfinite(A) && (A > silmask_th);
	// mapping with densify.use_edge_growth (slam/Loss.py:53): colour and distortion are averaged over
	// nan_to_num(alpha, 0, 0) > edge_thres instead of over the depth mask
	px.mask_c = edge_thres >= 0.f ? (isfinite(A) && (A > edge_thres)) : px.mask;
	return px;
}

// CTAs per SM of the two kernels (the scratch holds partials of up to 148 * 8 CTAs).  Every forward CTA ends with one
// atomic on the same ticket word and L2 serialises those, so its grid is small: 12.0 us at 2 CTAs / SM, 12.9 at 4,
// 13.7 at 8 (1200 x 680).  The backward kernel has no such tail: 11.5 us at 4, 12.2 at 8, 14.0 at 2.
constexpr int LOSS_FWD_CTAS_PER_SM = 2, LOSS_BWD_CTAS_PER_SM = 4;
constexpr int LOSS_THREADS = 256;
constexpr int LOSS_COLS = 8;        // per-block partials: count, colour sum, depth sum, distortion sum, colour count, 3 spare

// Four consecutive pixels at once (128-bit loads of every plane): HW % 4 == 0 and 16-byte aligned planes.
struct Quad {
	float v[4];
};
__device__ __forceinline__ Quad ldq(const float* __restrict__ p)
{
	const float4 t = *reinterpret_cast<const float4*>(p);
	Quad q;
	q.v[0] = t.x; q.v[1] = t.y; q.v[2] = t.z; q.v[3] = t.w;
	return q;
}
__device__ __forceinline__ void stq(float* __restrict__ p, const float (&v)[4])
{
	*reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
}
// the same pixel record as load_pixel(), for pixel i of a quad whose planes were loaded with ldq()
__device__ __forceinline__ LossPixel quad_pixel(int i, const Quad (&c)[3], const Quad (&gcq)[3], const Quad& gdq,
                                                const Quad& Dq, const Quad& Aq, const Quad& Rq, int gt_hwc, float eps,
                                                float near_, float far_, float silmask_th, float edge_thres)
{
	LossPixel px;
#pragma unroll
	for (int k = 0; k < 3; k++) {
		px.c[k] = nan_to_num00(c[k].v[i], px.c_ok[k]);
		// HWC: the quad's 12 colour values are consecutive: pixel i, channel k sits at 3 i + k
		px.gc[k] = gt_hwc ? gcq[(3 * i + k) >> 2].v[(3 * i + k) & 3] : gcq[k].v[i];
	}
	px.gd = gdq.v[i];
	const float D = Dq.v[i], A = Aq.v[i];
	px.inv_a = 1.0f / (A + eps);
	float d = D * px.inv_a;
	const bool outlier = (d > far_) || (d < near_);
	if (outlier) d = 0.f;
	bool d_ok;
	d = nan_to_num00(d, d_ok);
	px.depth = d;
	px.dist = nan_to_num00(Rq.v[i], px.dist_ok);
	px.mask = (px.gd > 1e-5f) && (d > 1e-5f);
	if (silmask_th >= 0.f) px.mask = px.mask && isfinite(A) && (A > silmask_th);
	px.mask_c = edge_thres >= 0.f ? (isfinite(A) && (A > edge_thres)) : px.mask;
	return px;
}
struct QuadIn {
	Quad c[3], gc[3], gd, D, A, R;
};
__device__ __forceinline__ QuadIn load_quad(size_t p, size_t HW, const float* __restrict__ color,
                                            const float* __restrict__ allmap, const float* __restrict__ gt_color,
                                            const float* __restrict__ gt_depth, int gt_hwc)
{
	QuadIn q;
#pragma unroll
	for (int k = 0; k < 3; k++) {
		q.c[k] = ldq(color + k * HW + p);
		q.gc[k] = gt_hwc ? ldq(gt_color + 3 * p + 4 * k) : ldq(gt_color + k * HW + p);
	}
	q.gd = ldq(gt_depth + p);
	q.D = ldq(allmap + p);
	q.A = ldq(allmap + HW + p);
	q.R = ldq(allmap + 6 * HW + p);
	return q;
}

template <bool VEC>
__global__ void __launch_bounds__(LOSS_THREADS)
mapping_loss_forward_kernel(size_t HW, const float* __restrict__ color, const float* __restrict__ allmap,
                            const float* __restrict__ gt_color, const float* __restrict__ gt_depth, int gt_hwc,
                            float w_color, float w_depth, float w_dist, float eps, float near_, float far_,
                            int tracking, float silmask_th, float edge_thres,
                            float* __restrict__ partials, uint32_t* __restrict__ ticket, float* __restrict__ state)
{
	float cnt = 0.f, cntc = 0.f, sc = 0.f, sd = 0.f, sr = 0.f;
	const float sil = tracking ? silmask_th : -1.f;
	const float edge = tracking ? -1.f : edge_thres;
	auto add_pixel = [&](const LossPixel& px) {
		if (px.mask) {
			cnt += 1.f;
			sd += fabsf(px.depth - px.gd);
		}
		if (px.mask_c) {
			cntc += 1.f;
			sc += fabsf(px.c[0] - px.gc[0]) + fabsf(px.c[1] - px.gc[1]) + fabsf(px.c[2] - px.gc[2]);
			sr += px.dist;
		}
	};
	if (VEC) {
		for (size_t p = 4 * ((size_t)blockIdx.x * LOSS_THREADS + threadIdx.x); p < HW; p += 4 * (size_t)gridDim.x * LOSS_THREADS) {
			const QuadIn q = load_quad(p, HW, color, allmap, gt_color, gt_depth, gt_hwc);
#pragma unroll
			for (int i = 0; i < 4; i++) {
				add_pixel(quad_pixel(i, q.c, q.gc, q.gd, q.D, q.A, q.R, gt_hwc, eps, near_, far_, sil, edge));
			}
		}
	} else
	for (size_t p = (size_t)blockIdx.x * LOSS_THREADS + threadIdx.x; p < HW; p += (size_t)gridDim.x * LOSS_THREADS) {
		add_pixel(load_pixel(p, HW, color, allmap, gt_color, gt_depth, gt_hwc, eps, near_, far_, sil, edge));
	}
	__shared__ float s_part[LOSS_THREADS / 32][LOSS_COLS];
	__shared__ bool s_last;
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
	for (int o = 16; o >= 1; o >>= 1) {
		cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
		sc += __shfl_xor_sync(0xffffffffu, sc, o);
		sd += __shfl_xor_sync(0xffffffffu, sd, o);
		sr += __shfl_xor_sync(0xffffffffu, sr, o);
		cntc += __shfl_xor_sync(0xffffffffu, cntc, o);
	}
	if (lane == 0) {
		s_part[warp][0] = cnt; s_part[warp][1] = sc; s_part[warp][2] = sd; s_part[warp][3] = sr; s_part[warp][4] = cntc;
		s_part[warp][5] = s_part[warp][6] = s_part[warp][7] = 0.f;
	}
	__syncthreads();
	if (threadIdx.x < LOSS_COLS) {
		float t = 0.f;
#pragma unroll
		for (int w = 0; w < LOSS_THREADS / 32; w++) t += s_part[w][threadIdx.x];
		partials[LOSS_COLS * blockIdx.x + threadIdx.x] = t;
	}
	__threadfence();
	__syncthreads();
	if (threadIdx.x == 0) s_last = atomicAdd(ticket, 1u) == gridDim.x - 1;
	__syncthreads();
	if (!s_last) return;
	// last block: add the per-block partials in block order (deterministic), in double
	// (8 columns x 32 chunks in parallel, then the chunks in order: the order of additions is fixed)
	__shared__ double s_sum[LOSS_COLS];
	__shared__ double s_chunk[LOSS_THREADS / LOSS_COLS][LOSS_COLS];
	{
		const unsigned col = threadIdx.x % LOSS_COLS, chunk = threadIdx.x / LOSS_COLS;
		double t = 0.0;
		for (unsigned b = chunk; b < gridDim.x; b += LOSS_THREADS / LOSS_COLS) t += (double)__ldcg(&partials[LOSS_COLS * b + col]);
		s_chunk[chunk][col] = t;
	}
	__syncthreads();
	if (threadIdx.x < LOSS_COLS) {
		double t = 0.0;
		for (int c = 0; c < LOSS_THREADS / LOSS_COLS; c++) t += s_chunk[c][threadIdx.x];
		s_sum[threadIdx.x] = t;
	}
	__syncthreads();
	if (threadIdx.x == 0) {
		const double n = s_sum[0], nc = s_sum[4];
		// mean over an empty selection is NaN in torch; kept
		// mapping: masked means (slam/Loss.py:52-58; colour and distortion over their own mask, which is the depth
		// mask unless use_edge_growth); tracking: masked sums, no distortion term (:46-47)
		const double loss = tracking ? (double)w_color * s_sum[1] + (double)w_depth * s_sum[2]
		                             : (double)w_color * (s_sum[1] / (3.0 * nc)) + (double)w_depth * (s_sum[2] / n) +
		                                   (double)w_dist * (s_sum[3] / nc);
		state[0] = (float)loss;
		state[1] = (float)n;
		state[2] = (float)s_sum[1];
		state[3] = (float)s_sum[2];
		state[4] = (float)s_sum[3];
		state[5] = (float)nc;
		*ticket = 0u;   // ready for the next call
	}
}

template <bool VEC>
__global__ void __launch_bounds__(LOSS_THREADS)
mapping_loss_backward_kernel(size_t HW, const float* __restrict__ color, const float* __restrict__ allmap,
                             const float* __restrict__ gt_color, const float* __restrict__ gt_depth, int gt_hwc,
                             float w_color, float w_depth, float w_dist, float eps, float near_, float far_,
                             int tracking, float silmask_th, float edge_thres,
                             const float* __restrict__ state, const float* __restrict__ grad_out,
                             float* __restrict__ dL_dcolor, float* __restrict__ dL_dallmap)
{
	const float n = state[1], nc = state[5];
	const float g = grad_out ? grad_out[0] : 1.0f;
	const float gc = tracking ? g * w_color : (nc > 0.f ? g * w_color / (3.0f * nc) : 0.f);
	const float gd = tracking ? g * w_depth : (n > 0.f ? g * w_depth / n : 0.f);
	const float gr = tracking ? 0.f : (nc > 0.f ? g * w_dist / nc : 0.f);
	const float edge = tracking ? -1.f : edge_thres;
	const float sil = tracking ? silmask_th : -1.f;
	// per-pixel gradients of one pixel record: (dc[3], dD, dA, dR)
	auto pixel_grads = [&](const LossPixel& px, float (&dc)[3], float& dD, float& dA, float& dR) {
		dc[0] = dc[1] = dc[2] = 0.f;
		dD = dA = dR = 0.f;
		if (px.mask_c) {
#pragma unroll
			for (int k = 0; k < 3; k++) {
				const float e = px.c[k] - px.gc[k];
				dc[k] = px.c_ok[k] ? (e > 0.f ? gc : (e < 0.f ? -gc : 0.f)) : 0.f;   // sign(0) = 0 like torch.abs
			}
			dR = px.dist_ok ? gr : 0.f;
		}
		if (px.mask) {
			const float e = px.depth - px.gd;
			const float dd = e > 0.f ? gd : (e < 0.f ? -gd : 0.f);
			dD = dd * px.inv_a;                       // depth = D / (A + eps)
			dA = -dd * px.depth * px.inv_a;
		}
	};
	if (VEC) {
		const float zero[4] = {0.f, 0.f, 0.f, 0.f};
		for (size_t p = 4 * ((size_t)blockIdx.x * LOSS_THREADS + threadIdx.x); p < HW; p += 4 * (size_t)gridDim.x * LOSS_THREADS) {
			const QuadIn q = load_quad(p, HW, color, allmap, gt_color, gt_depth, gt_hwc);
			float o_c[3][4], o_D[4], o_A[4], o_R[4];
#pragma unroll
			for (int i = 0; i < 4; i++) {
				const LossPixel px = quad_pixel(i, q.c, q.gc, q.gd, q.D, q.A, q.R, gt_hwc, eps, near_, far_, sil, edge);
				float dc[3];
				pixel_grads(px, dc, o_D[i], o_A[i], o_R[i]);
				o_c[0][i] = dc[0]; o_c[1][i] = dc[1]; o_c[2][i] = dc[2];
			}
			stq(dL_dcolor + p, o_c[0]); stq(dL_dcolor + HW + p, o_c[1]); stq(dL_dcolor + 2 * HW + p, o_c[2]);
			stq(dL_dallmap + p, o_D); stq(dL_dallmap + HW + p, o_A);
			stq(dL_dallmap + 2 * HW + p, zero); stq(dL_dallmap + 3 * HW + p, zero); stq(dL_dallmap + 4 * HW + p, zero);
			stq(dL_dallmap + 5 * HW + p, zero); stq(dL_dallmap + 6 * HW + p, o_R);
		}
		return;
	}
	for (size_t p = (size_t)blockIdx.x * LOSS_THREADS + threadIdx.x; p < HW; p += (size_t)gridDim.x * LOSS_THREADS) {
		const LossPixel px = load_pixel(p, HW, color, allmap, gt_color, gt_depth, gt_hwc, eps, near_, far_, sil, edge);
		float dc[3], dD, dA, dR;
		pixel_grads(px, dc, dD, dA, dR);
		dL_dcolor[p] = dc[0]; dL_dcolor[HW + p] = dc[1]; dL_dcolor[2 * HW + p] = dc[2];
		dL_dallmap[p] = dD; dL_dallmap[HW + p] = dA;
		dL_dallmap[2 * HW + p] = 0.f; dL_dallmap[3 * HW + p] = 0.f; dL_dallmap[4 * HW + p] = 0.f;
		dL_dallmap[5 * HW + p] = 0.f; dL_dallmap[6 * HW + p] = dR;
	}
}

inline bool loss_vec_ok(size_t HW, std::initializer_list<const void*> ptrs)
{
	if (HW & 3) return false;
	for (const void* q : ptrs)
		if ((uintptr_t)q & 15) return false;
	return true;
}

inline int loss_blocks(size_t HW, bool vec, int ctas_per_sm)
{
	const size_t items = vec ? HW / 4 : HW;
	const size_t want = (items + LOSS_THREADS - 1) / LOSS_THREADS;
	// grid-stride loops over a grid of `ctas_per_sm` CTAs per SM (measured per kernel, see the callers)
	const size_t cap = 148 * (size_t)ctas_per_sm;
	return (int)(want < cap ? (want ? want : 1) : cap);
}

}  // namespace

}  // namespace gs2d

using namespace gs2d;

extern "C" {

size_t gs2d_mapping_loss_scratch_bytes(int32_t, int32_t)
{
	// per-block partials (LOSS_COLS floats x 148*8 blocks), the ticket, the 8-float state
	return sizeof(float) * LOSS_COLS * 148 * 8 + 128 + sizeof(float) * 8;
}

static inline void loss_carve(void* scratch, float** partials, uint32_t** ticket, float** state)
{
	char* b = (char*)scratch;
	*partials = (float*)b;
	*ticket = (uint32_t*)(b + sizeof(float) * LOSS_COLS * 148 * 8);
	*state = (float*)(b + sizeof(float) * LOSS_COLS * 148 * 8 + 128);
}

int gs2d_mapping_loss_forward(int32_t W, int32_t H, const float* color, const float* allmap, const float* gt_color,
                              const float* gt_depth, int32_t gt_color_hwc, const gs2d_loss_params* lp, void* scratch,
                              float* loss_out, void* stream)
{
	if (W <= 0 || H <= 0 || !color || !allmap || !gt_color || !gt_depth || !lp || !scratch) return GS2D_ERR_INVALID;
	cudaStream_t st = (cudaStream_t)stream;
	float *partials, *state;
	uint32_t* ticket;
	loss_carve(scratch, &partials, &ticket, &state);
	const size_t HW = (size_t)W * H;
	{
		ProfScope ps(K_LOSS, st);
		const bool vec = loss_vec_ok(HW, {color, allmap, gt_color, gt_depth});
		const int blocks = loss_blocks(HW, vec, LOSS_FWD_CTAS_PER_SM);
		if (vec)
			mapping_loss_forward_kernel<true><<<blocks, LOSS_THREADS, 0, st>>>(
				HW, color, allmap, gt_color, gt_depth, gt_color_hwc, lp->w_color, lp->w_depth, lp->w_dist, lp->eps,
				lp->depth_near, lp->depth_far, lp->tracking, lp->silmask_th, lp->edge_thres, partials, ticket, state);
		else
			mapping_loss_forward_kernel<false><<<blocks, LOSS_THREADS, 0, st>>>(
				HW, color, allmap, gt_color, gt_depth, gt_color_hwc, lp->w_color, lp->w_depth, lp->w_dist, lp->eps,
				lp->depth_near, lp->depth_far, lp->tracking, lp->silmask_th, lp->edge_thres, partials, ticket, state);
		count_launch();
	}
	GS2D_CUDA_OK(cudaGetLastError());
	if (loss_out) GS2D_CUDA_OK(cudaMemcpyAsync(loss_out, state, sizeof(float), cudaMemcpyDeviceToDevice, st));
	return GS2D_OK;
}

int gs2d_mapping_loss_backward(int32_t W, int32_t H, const float* color, const float* allmap, const float* gt_color,
                               const float* gt_depth, int32_t gt_color_hwc, const gs2d_loss_params* lp,
                               const void* scratch, const float* grad_loss, float* dL_dcolor, float* dL_dallmap,
                               void* stream)
{
	if (W <= 0 || H <= 0 || !color || !allmap || !gt_color || !gt_depth || !lp || !scratch || !dL_dcolor || !dL_dallmap)
		return GS2D_ERR_INVALID;
	cudaStream_t st = (cudaStream_t)stream;
	float *partials, *state;
	uint32_t* ticket;
	loss_carve(const_cast<void*>(scratch), &partials, &ticket, &state);
	const size_t HW = (size_t)W * H;
	{
		ProfScope ps(K_LOSS, st);
		const bool vec = loss_vec_ok(HW, {color, allmap, gt_color, gt_depth, dL_dcolor, dL_dallmap});
		const int blocks = loss_blocks(HW, vec, LOSS_BWD_CTAS_PER_SM);
		if (vec)
			mapping_loss_backward_kernel<true><<<blocks, LOSS_THREADS, 0, st>>>(
				HW, color, allmap, gt_color, gt_depth, gt_color_hwc, lp->w_color, lp->w_depth, lp->w_dist, lp->eps,
				lp->depth_near, lp->depth_far, lp->tracking, lp->silmask_th, lp->edge_thres, state, grad_loss, dL_dcolor, dL_dallmap);
		else
			mapping_loss_backward_kernel<false><<<blocks, LOSS_THREADS, 0, st>>>(
				HW, color, allmap, gt_color, gt_depth, gt_color_hwc, lp->w_color, lp->w_depth, lp->w_dist, lp->eps,
				lp->depth_near, lp->depth_far, lp->tracking, lp->silmask_th, lp->edge_thres, state, grad_loss, dL_dcolor, dL_dallmap);
		count_launch();
	}
	GS2D_CUDA_OK(cudaGetLastError());
	return GS2D_OK;
}

}  // extern "C"
