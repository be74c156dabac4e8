// optim.cu -- fused map optimiser step (the step right after the gradient all-reduce).
//
// The reference keeps five raw parameter tensors (scene/Gaussians.py:106-114: _xyz [P,3], _opacity [P,1] logit,
// _scaling [P,S] log, _rotation [P,4] unnormalised, _rgb [P,3]), renders their activations (get_render_params,
// :336-358: sigmoid / exp (tiled to two columns when isotropic) / F.normalize) and steps them with
// torch.optim.Adam(lr=0, eps=1e-15) with one learning rate per group (:124-137).  Per iteration that is the
// autograd backward of three activations plus a foreach Adam over five groups (~20 launches, every parameter and
// moment read and written several times) and, at the next iteration, three activation launches.
//
// Here: ONE kernel.  It reads the gradient w.r.t. the RENDER parameters straight from the flat [13 * P] buffer
// the backward pass wrote / the all-reduce summed (gs2d_b200/mapping.py: means3D 3P | opacities P | scales 2P |
// rotations 4P | colors 3P), applies the chain rule of the activation, does the Adam update of
// torch/optim/adam.py (_single_tensor_adam, no amsgrad / weight decay / maximize):
//     m <- m + (g - m) (1 - b1);  v <- v b2 + (1 - b2) g g;
//     p <- p - (lr / bc1) * m / (sqrt(v) / sqrt(bc2) + eps),     bc_k = 1 - b_k^t
// and writes the activations of the UPDATED parameters for the next forward pass.  Pure HBM stream: per surfel
// 13 (param) + 13 (grad) + 26 (moments) floats in, 13 + 26 + 7 (opacity, 2 scales, 4 rotation) out = 392 B.
// Work item = 4 consecutive floats of one group (for the rotation group: one quaternion); 128-bit accesses
// whenever the group's pointers are 16-byte aligned.
#include "common.cuh"

#include <cmath>

namespace gs2d {

namespace {

constexpr int OPT_THREADS = 256;

enum Group { G_XYZ = 0, G_OPACITY, G_SCALING, G_ROTATION, G_RGB, G_COUNT };

struct GroupArgs {
	float* param;            // raw parameter, n floats
	float* exp_avg;          // Adam first moment   (NULL in activate-only launches)
	float* exp_avg_sq;       // Adam second moment
	const float* grad;       // gradient w.r.t. the render parameter (section of the flat buffer)
	float* activated;        // render parameter out (NULL: not written; xyz / rgb are their own activation)
	long long n;             // number of raw floats
	float step_size;         // lr / (1 - beta1^t), computed in double like torch/optim/adam.py
	float bc2_sqrt;          // sqrt(1 - beta2^t); t is per group (torch keeps one step counter per parameter)
	int vec;                 // all of the above 16-byte aligned
	int block0;              // first block of this group
};

struct StepArgs {
	GroupArgs g[G_COUNT];
	float beta1, beta2, w1, w2, eps;             // w_k = 1 - beta_k (rounded from double, as torch's scalars are)
	int scale_cols;          // S: 2 (anisotropic) or 1 (isotropic: exp(s) tiled to both render columns)
	int update;              // 0: activation only
};

__device__ __forceinline__ void ld4(const float* __restrict__ p, long long i, long long n, bool vec, float (&o)[4])
{
	if (vec && i + 4 <= n) {
		const float4 t = *reinterpret_cast<const float4*>(p + i);
		o[0] = t.x; o[1] = t.y; o[2] = t.z; o[3] = t.w;
	} else {
#pragma unroll
		for (int k = 0; k < 4; k++) o[k] = (i + k < n) ? p[i + k] : 0.f;
	}
}

__device__ __forceinline__ void st4(float* __restrict__ p, long long i, long long n, bool vec, const float (&o)[4])
{
	if (vec && i + 4 <= n) {
		*reinterpret_cast<float4*>(p + i) = make_float4(o[0], o[1], o[2], o[3]);
	} else {
#pragma unroll
		for (int k = 0; k < 4; k++)
			if (i + k < n) p[i + k] = o[k];
	}
}

__device__ __forceinline__ float sigmoidf(float x) { return 1.0f / (1.0f + expf(-x)); }

// F.normalize(q, dim=-1, eps=1e-12): q / max(|q|, eps)
__device__ __forceinline__ void normalize4(const float (&q)[4], float (&o)[4], float& n)
{
	n = fmaxf(sqrtf(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]), 1e-12f);
#pragma unroll
	for (int k = 0; k < 4; k++) o[k] = q[k] / n;
}

__global__ void __launch_bounds__(OPT_THREADS) map_step_kernel(const StepArgs a)
{
	int gi = 0;
#pragma unroll
	for (int k = 1; k < G_COUNT; k++)
		if ((int)blockIdx.x >= a.g[k].block0) gi = k;
	const GroupArgs& G = a.g[gi];
	const long long i = 4ll * ((long long)(blockIdx.x - G.block0) * OPT_THREADS + threadIdx.x);
	if (i >= G.n) return;
	const bool vec = G.vec != 0;

	float p[4];
	ld4(G.param, i, G.n, vec, p);

	if (a.update) {
		// gradient w.r.t. the raw parameter: chain rule of the activation ---------------------------------
		float g[4];
		if (gi == G_SCALING && a.scale_cols == 1) {
			// scales = tile(exp(s), (1, 2)): the raw gradient is the sum of the two render columns
#pragma unroll
			for (int k = 0; k < 4; k++)
				g[k] = (i + k < G.n) ? G.grad[2 * (i + k)] + G.grad[2 * (i + k) + 1] : 0.f;
		} else {
			ld4(G.grad, i, G.n, vec, g);
		}
		if (gi == G_OPACITY) {
#pragma unroll
			for (int k = 0; k < 4; k++) {       // sigmoid_backward: grad * (1 - y) * y
				const float y = sigmoidf(p[k]);
				g[k] = g[k] * (1.0f - y) * y;
			}
		} else if (gi == G_SCALING) {
#pragma unroll
			for (int k = 0; k < 4; k++) g[k] = g[k] * expf(p[k]);
		} else if (gi == G_ROTATION) {
			float qh[4], n;
			normalize4(p, qh, n);
			const float d = qh[0] * g[0] + qh[1] * g[1] + qh[2] * g[2] + qh[3] * g[3];
#pragma unroll
			for (int k = 0; k < 4; k++) g[k] = (g[k] - qh[k] * d) / n;
		}
		// Adam ---------------------------------------------------------------------------------------------
		float m[4], v[4];
		ld4(G.exp_avg, i, G.n, vec, m);
		ld4(G.exp_avg_sq, i, G.n, vec, v);
#pragma unroll
		for (int k = 0; k < 4; k++) {
			// exp_avg.lerp_(grad, 1 - beta1): torch's lerp formula (ATen/native/Lerp.h)
			const float diff = g[k] - m[k];
			m[k] = (a.w1 < 0.5f) ? m[k] + a.w1 * diff : g[k] - diff * (1.0f - a.w1);
			// exp_avg_sq.mul_(beta2).addcmul_(grad, grad, value=1 - beta2)
			v[k] = v[k] * a.beta2 + a.w2 * g[k] * g[k];
			// denom = exp_avg_sq.sqrt() / sqrt(bc2) + eps;  param.addcdiv_(exp_avg, denom, value=-step_size)
			const float denom = sqrtf(v[k]) / G.bc2_sqrt + a.eps;
			p[k] = p[k] - G.step_size * (m[k] / denom);
		}
		st4(G.exp_avg, i, G.n, vec, m);
		st4(G.exp_avg_sq, i, G.n, vec, v);
		st4(G.param, i, G.n, vec, p);
	}

	// activation of the (updated) parameter: the next forward pass reads these ------------------------------
	if (G.activated == nullptr) return;
	float o[4];
	if (gi == G_OPACITY) {
#pragma unroll
		for (int k = 0; k < 4; k++) o[k] = sigmoidf(p[k]);
		st4(G.activated, i, G.n, vec, o);
	} else if (gi == G_SCALING) {
#pragma unroll
		for (int k = 0; k < 4; k++) o[k] = expf(p[k]);
		if (a.scale_cols == 1) {
#pragma unroll
			for (int k = 0; k < 4; k++)
				if (i + k < G.n) { G.activated[2 * (i + k)] = o[k]; G.activated[2 * (i + k) + 1] = o[k]; }
		} else {
			st4(G.activated, i, G.n, vec, o);
		}
	} else if (gi == G_ROTATION) {
		float n;
		normalize4(p, o, n);
		st4(G.activated, i, G.n, vec, o);
	} else {
		st4(G.activated, i, G.n, vec, p);
	}
}

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

int launch_map_step(int P, int scale_cols, const gs2d_map_state* ms, const float* flat_grad, const gs2d_adam_params* ap,
                    float* opacities, float* scales, float* rotations, bool update, cudaStream_t st)
{
	StepArgs a{};
	float* param[G_COUNT] = {ms->xyz, ms->opacity, ms->scaling, ms->rotation, ms->rgb};
	const long long width[G_COUNT] = {3, 1, scale_cols, 4, 3};
	const long long gsec[G_COUNT] = {0, 3, 4, 6, 10};      // flat-buffer section starts, in units of P floats
	float* act[G_COUNT] = {nullptr, opacities, scales, rotations, nullptr};
	int blocks = 0;
	for (int k = 0; k < G_COUNT; k++) {
		GroupArgs& G = a.g[k];
		G.param = param[k];
		G.exp_avg = update ? ms->exp_avg[k] : nullptr;
		G.exp_avg_sq = update ? ms->exp_avg_sq[k] : nullptr;
		G.grad = update ? flat_grad + gsec[k] * (long long)P : nullptr;
		G.activated = act[k];
		G.n = width[k] * (long long)P;
		// torch counts steps per parameter: a group that was skipped before (no gradient) is behind the others
		const double t = update ? (double)(ap->group_step[k] > 0 ? ap->group_step[k] : ap->step) : 1.0;
		const double bc1 = update ? 1.0 - pow(ap->beta1, t) : 1.0;
		const double bc2 = update ? 1.0 - pow(ap->beta2, t) : 1.0;
		G.step_size = update ? (float)(ap->lr[k] / bc1) : 0.f;
		G.bc2_sqrt = (float)sqrt(bc2);
		G.vec = aligned16(G.param) && aligned16(G.exp_avg) && aligned16(G.exp_avg_sq) && aligned16(G.grad) &&
		        aligned16(G.activated);
		G.block0 = blocks;
		// activation-only launches skip the groups that are their own activation; an update skips the groups the
		// caller marks with a negative learning rate: "this parameter has no gradient" -- torch.optim.Adam leaves
		// parameters whose .grad is None untouched, moments included (Renderer_BA / Renderer_tracking detach the
		// rotations, render/__init__.py:36, :98, so _rotation.grad stays None in those steps)
		const bool skip = (!update && G.activated == nullptr) || (update && ap->lr[k] < 0.0);
		const long long items = skip ? 0 : (G.n + 3) / 4;
		blocks += (int)((items + OPT_THREADS - 1) / OPT_THREADS);
	}
	if (update) {
		a.beta1 = (float)ap->beta1; a.beta2 = (float)ap->beta2; a.eps = (float)ap->eps;
		a.w1 = (float)(1.0 - ap->beta1); a.w2 = (float)(1.0 - ap->beta2);
	}
	a.scale_cols = scale_cols;
	a.update = update ? 1 : 0;
	if (blocks == 0) return GS2D_OK;
	{
		ProfScope ps(K_MAP_STEP, st);
		map_step_kernel<<<blocks, OPT_THREADS, 0, st>>>(a);
		count_launch();
	}
	GS2D_CUDA_OK(cudaGetLastError());
	return GS2D_OK;
}

}  // namespace

}  // namespace gs2d

using namespace gs2d;

extern "C" {

int gs2d_map_activate(int32_t P, int32_t scale_cols, const gs2d_map_state* ms, float* opacities, float* scales,
                      float* rotations, void* stream)
{
	if (P < 0 || (scale_cols != 1 && scale_cols != 2) || !ms) return GS2D_ERR_INVALID;
	if (P == 0) return GS2D_OK;
	if (!ms->opacity || !ms->scaling || !ms->rotation || !opacities || !scales || !rotations) return GS2D_ERR_INVALID;
	return launch_map_step(P, scale_cols, ms, nullptr, nullptr, opacities, scales, rotations, false, (cudaStream_t)stream);
}

int gs2d_map_adam_step(int32_t P, int32_t scale_cols, const gs2d_map_state* ms, const float* flat_grad,
                       const gs2d_adam_params* ap, float* opacities, float* scales, float* rotations, void* stream)
{
	if (P < 0 || (scale_cols != 1 && scale_cols != 2) || !ms || !ap) return GS2D_ERR_INVALID;
	if (P == 0) return GS2D_OK;
	if (!flat_grad || !ms->xyz || !ms->opacity || !ms->scaling || !ms->rotation || !ms->rgb) return GS2D_ERR_INVALID;
	for (int k = 0; k < 5; k++)
		if (!ms->exp_avg[k] || !ms->exp_avg_sq[k]) return GS2D_ERR_INVALID;
	for (int k = 0; k < 5; k++)
		if (ap->group_step[k] < 0) return GS2D_ERR_INVALID;
	if (ap->step < 1 || !(ap->beta1 >= 0.0 && ap->beta1 < 1.0) || !(ap->beta2 >= 0.0 && ap->beta2 < 1.0)) return GS2D_ERR_INVALID;
	// the three activation outputs are optional, but all or none
	const int n_act = (opacities != nullptr) + (scales != nullptr) + (rotations != nullptr);
	if (n_act != 0 && n_act != 3) return GS2D_ERR_INVALID;
	return launch_map_step(P, scale_cols, ms, flat_grad, ap, opacities, scales, rotations, true, (cudaStream_t)stream);
}

}  // extern "C"
