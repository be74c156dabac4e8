// pose.cu -- fused pose transform + pose-gradient reduction (the step either side of the
// rasterizer in tracking and bundle adjustment).
//
// Renderer_tracking / Renderer_BA (render/__init__.py:35-36, :97-98) render with an identity camera and move the
// map into the camera frame first:
//     means3D'   = (w2c[:3,:3] @ means3D.T + w2c[:3,3:]).T
//     rotations' = quaternion_multiply(matrix_to_quaternion(w2c[None,:3,:3]), rotations).detach()
// with w2c = Transform.get_transform_matrix (scene/Frame.py:87-96): build_rotation(F.normalize(_cam_rot)) and
// _cam_trans.  In torch that is a GEMM, a transpose, ~15 small quaternion launches forward and, backward, a
// [P,3] x [3,P] GEMM + reductions + the autograd chain of the 4x4 assembly -- per tracking iteration.
//
// Forward here: one kernel, 28 B in + 28 B out per surfel.  Backward: one kernel that reduces
//     dL/dR = sum_i g_i x_i^T,  dL/dt = sum_i g_i            (g = dL/dmeans3D' from the rasterizer's backward)
// over the map (24 B per surfel; per-block partials, the last block adds them in a fixed order: deterministic),
// optionally writes dL/dmeans3D = R^T g (bundle adjustment), and finishes the chain rule down to the seven pose
// scalars (dL/d_cam_rot through quaternion_to_matrix and F.normalize, dL/d_cam_trans) in the same launch.
//
// pytorch3d is not part of the reference checkout (requirements: pytorch3d, version unpinned); its published
// algorithms are restated: quaternion_to_matrix (two_s = 2 / |q|^2 form), matrix_to_quaternion (four candidates,
// the best conditioned one wins, floor 0.1), quaternion_raw_multiply, standardize_quaternion (real part >= 0).
#include "common.cuh"

namespace gs2d {

namespace {

constexpr int POSE_THREADS = 256;
constexpr int POSE_MAX_BLOCKS = 148 * 4;
constexpr int POSE_CHUNKS = 20;           // the last block sums the per-block partials as 12 columns x 20 chunks

struct Pose {
	float R[9];      // row-major
	float t[3];
	float q[4];      // quaternion of R (real first), standardised
};

// pytorch3d.transforms.quaternion_to_matrix
__device__ __forceinline__ void quat_to_matrix(const float (&q)[4], float (&R)[9])
{
	const float r = q[0], i = q[1], j = q[2], k = q[3];
	const float two_s = 2.0f / (r * r + i * i + j * j + k * k);
	R[0] = 1.f - two_s * (j * j + k * k); R[1] = two_s * (i * j - k * r);       R[2] = two_s * (i * k + j * r);
	R[3] = two_s * (i * j + k * r);       R[4] = 1.f - two_s * (i * i + k * k); R[5] = two_s * (j * k - i * r);
	R[6] = two_s * (i * k - j * r);       R[7] = two_s * (j * k + i * r);       R[8] = 1.f - two_s * (i * i + j * j);
}

__device__ __forceinline__ float sqrt_positive_part(float x) { return x > 0.f ? sqrtf(x) : 0.f; }

// pytorch3d.transforms.matrix_to_quaternion (+ standardize_quaternion)
__device__ __forceinline__ void matrix_to_quat(const float (&m)[9], float (&q)[4])
{
	const float m00 = m[0], m01 = m[1], m02 = m[2], m10 = m[3], m11 = m[4], m12 = m[5], m20 = m[6], m21 = m[7], m22 = m[8];
	const float qa[4] = {sqrt_positive_part(1.f + m00 + m11 + m22), sqrt_positive_part(1.f + m00 - m11 - m22),
	                     sqrt_positive_part(1.f - m00 + m11 - m22), sqrt_positive_part(1.f - m00 - m11 + m22)};
	const float cand[4][4] = {{qa[0] * qa[0], m21 - m12, m02 - m20, m10 - m01},
	                          {m21 - m12, qa[1] * qa[1], m10 + m01, m02 + m20},
	                          {m02 - m20, m10 + m01, qa[2] * qa[2], m12 + m21},
	                          {m10 - m01, m20 + m02, m21 + m12, qa[3] * qa[3]}};
	int best = 0;                      // argmax: first maximum
#pragma unroll
	for (int c = 1; c < 4; c++)
		if (qa[c] > qa[best]) best = c;
#pragma unroll
	for (int c = 0; c < 4; c++) {
		float v = cand[0][c], d = qa[0];
		if (best == 1) { v = cand[1][c]; d = qa[1]; }
		if (best == 2) { v = cand[2][c]; d = qa[2]; }
		if (best == 3) { v = cand[3][c]; d = qa[3]; }
		q[c] = v / (2.0f * fmaxf(d, 0.1f));
	}
	if (q[0] < 0.f) {
#pragma unroll
		for (int c = 0; c < 4; c++) q[c] = -q[c];
	}
}

// F.normalize(q[None])[0]
__device__ __forceinline__ float normalize_quat(const float* __restrict__ q_in, float (&qh)[4])
{
	const float a = q_in[0], b = q_in[1], c = q_in[2], d = q_in[3];
	const float n = fmaxf(sqrtf(a * a + b * b + c * c + d * d), 1e-12f);
	qh[0] = a / n; qh[1] = b / n; qh[2] = c / n; qh[3] = d / n;
	return n;
}

// exactly one of {(cam_rot, cam_trans), w2c}
__device__ __forceinline__ void load_pose(const float* __restrict__ cam_rot, const float* __restrict__ cam_trans,
                                          const float* __restrict__ w2c, Pose& p)
{
	if (w2c != nullptr) {
#pragma unroll
		for (int r = 0; r < 3; r++) {
#pragma unroll
			for (int c = 0; c < 3; c++) p.R[3 * r + c] = w2c[4 * r + c];
			p.t[r] = w2c[4 * r + 3];
		}
	} else {
		float qh[4];
		normalize_quat(cam_rot, qh);
		quat_to_matrix(qh, p.R);
		p.t[0] = cam_trans[0]; p.t[1] = cam_trans[1]; p.t[2] = cam_trans[2];
	}
	matrix_to_quat(p.R, p.q);
}

__global__ void __launch_bounds__(POSE_THREADS) pose_transform_forward_kernel(
	int P, const float* __restrict__ xyz, const float* __restrict__ rotations, const float* __restrict__ cam_rot,
	const float* __restrict__ cam_trans, const float* __restrict__ w2c, float* __restrict__ means_out,
	float* __restrict__ rot_out)
{
	Pose ps;
	load_pose(cam_rot, cam_trans, w2c, ps);
	for (int i = blockIdx.x * POSE_THREADS + threadIdx.x; i < P; i += gridDim.x * POSE_THREADS) {
		const float x = xyz[3 * i], y = xyz[3 * i + 1], z = xyz[3 * i + 2];
		means_out[3 * i]     = ps.R[0] * x + ps.R[1] * y + ps.R[2] * z + ps.t[0];
		means_out[3 * i + 1] = ps.R[3] * x + ps.R[4] * y + ps.R[5] * z + ps.t[1];
		means_out[3 * i + 2] = ps.R[6] * x + ps.R[7] * y + ps.R[8] * z + ps.t[2];
		if (rot_out != nullptr) {
			const float4 b = reinterpret_cast<const float4*>(rotations)[i];     // (w, x, y, z)
			const float aw = ps.q[0], ax = ps.q[1], ay = ps.q[2], az = ps.q[3];
			float4 o;                                                            // quaternion_raw_multiply(a, b)
			o.x = aw * b.x - ax * b.y - ay * b.z - az * b.w;
			o.y = aw * b.y + ax * b.x + ay * b.w - az * b.z;
			o.z = aw * b.z - ax * b.w + ay * b.x + az * b.y;
			o.w = aw * b.w + ax * b.z - ay * b.y + az * b.x;
			if (o.x < 0.f) { o.x = -o.x; o.y = -o.y; o.z = -o.z; o.w = -o.w; }   // standardize_quaternion
			reinterpret_cast<float4*>(rot_out)[i] = o;
		}
	}
}

// partial sums: [0..8] = dL/dR row-major, [9..11] = dL/dt
__global__ void __launch_bounds__(POSE_THREADS) pose_transform_backward_kernel(
	int P, const float* __restrict__ xyz, const float* __restrict__ cam_rot, const float* __restrict__ cam_trans,
	const float* __restrict__ w2c, const float* __restrict__ dL_dmeans, float* __restrict__ partials,
	unsigned int* __restrict__ ticket, float* __restrict__ dL_dcam_rot, float* __restrict__ dL_dcam_trans,
	float* __restrict__ dL_dw2c, float* __restrict__ dL_dxyz)
{
	__shared__ float sm[POSE_THREADS / 32][12];
	__shared__ bool last;
	Pose ps;
	if (dL_dxyz != nullptr) load_pose(cam_rot, cam_trans, w2c, ps);
	float acc[12];
#pragma unroll
	for (int k = 0; k < 12; k++) acc[k] = 0.f;
	for (int i = blockIdx.x * POSE_THREADS + threadIdx.x; i < P; i += gridDim.x * POSE_THREADS) {
		const float x[3] = {xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]};
		const float g[3] = {dL_dmeans[3 * i], dL_dmeans[3 * i + 1], dL_dmeans[3 * i + 2]};
#pragma unroll
		for (int a = 0; a < 3; a++) {
#pragma unroll
			for (int b = 0; b < 3; b++) acc[3 * a + b] += g[a] * x[b];
			acc[9 + a] += g[a];
		}
		if (dL_dxyz != nullptr) {     // R^T g
			dL_dxyz[3 * i]     = ps.R[0] * g[0] + ps.R[3] * g[1] + ps.R[6] * g[2];
			dL_dxyz[3 * i + 1] = ps.R[1] * g[0] + ps.R[4] * g[1] + ps.R[7] * g[2];
			dL_dxyz[3 * i + 2] = ps.R[2] * g[0] + ps.R[5] * g[1] + ps.R[8] * g[2];
		}
	}
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
	for (int k = 0; k < 12; k++) {
		float v = acc[k];
#pragma unroll
		for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
		if (lane == 0) sm[warp][k] = v;
	}
	__syncthreads();
	if (threadIdx.x < 12) {
		float v = 0.f;
#pragma unroll
		for (int w = 0; w < POSE_THREADS / 32; w++) v += sm[w][threadIdx.x];
		partials[12 * blockIdx.x + threadIdx.x] = v;
	}
	__threadfence();
	__syncthreads();
	if (threadIdx.x == 0) last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
	__syncthreads();
	if (!last) return;
	__threadfence();
	// the last block: fixed-order sum of the per-block partials, then the chain rule of the pose parametrisation
	__shared__ float G[12];
	__shared__ float red[POSE_CHUNKS][12];
	if (threadIdx.x < 12 * POSE_CHUNKS) {
		const unsigned col = threadIdx.x % 12, chunk = threadIdx.x / 12;
		float v = 0.f;
		for (unsigned b = chunk; b < gridDim.x; b += POSE_CHUNKS) v += __ldcg(&partials[12 * b + col]);
		red[chunk][col] = v;
	}
	__syncthreads();
	if (threadIdx.x < 12) {
		float v = 0.f;
#pragma unroll
		for (int c = 0; c < POSE_CHUNKS; c++) v += red[c][threadIdx.x];
		G[threadIdx.x] = v;
	}
	__syncthreads();
	if (threadIdx.x != 0) return;
	*ticket = 0u;                       // ready for the next launch
	if (dL_dw2c != nullptr) {           // matrix form: gradient of the 4x4 (last row does not take part)
#pragma unroll
		for (int r = 0; r < 3; r++) {
#pragma unroll
			for (int c = 0; c < 3; c++) dL_dw2c[4 * r + c] = G[3 * r + c];
			dL_dw2c[4 * r + 3] = G[9 + r];
		}
#pragma unroll
		for (int c = 0; c < 4; c++) dL_dw2c[12 + c] = 0.f;
		return;
	}
	dL_dcam_trans[0] = G[9]; dL_dcam_trans[1] = G[10]; dL_dcam_trans[2] = G[11];
	// R = I + two_s * N(qh), two_s = 2 / (qh . qh), qh = q / |q|
	float qh[4];
	const float n = normalize_quat(cam_rot, qh);
	const float r = qh[0], i = qh[1], j = qh[2], k = qh[3];
	const float s = r * r + i * i + j * j + k * k;
	const float two_s = 2.0f / s;
	const float N[9] = {-(j * j + k * k), i * j - k * r, i * k + j * r,
	                    i * j + k * r, -(i * i + k * k), j * k - i * r,
	                    i * k - j * r, j * k + i * r, -(i * i + j * j)};
	float d_two_s = 0.f;
#pragma unroll
	for (int e = 0; e < 9; e++) d_two_s += G[e] * N[e];
	float dq[4];
	dq[0] = two_s * (-k * G[1] + j * G[2] + k * G[3] - i * G[5] - j * G[6] + i * G[7]);
	dq[1] = two_s * (j * G[1] + k * G[2] + j * G[3] - 2.f * i * G[4] - r * G[5] + k * G[6] + r * G[7] - 2.f * i * G[8]);
	dq[2] = two_s * (-2.f * j * G[0] + i * G[1] + r * G[2] + i * G[3] + k * G[5] - r * G[6] + k * G[7] - 2.f * j * G[8]);
	dq[3] = two_s * (-2.f * k * G[0] - r * G[1] + i * G[2] + r * G[3] - 2.f * k * G[4] + j * G[5] + i * G[6] + j * G[7]);
	const float c2 = -4.0f * d_two_s / (s * s);       // d two_s / d qh = -4 qh / s^2
#pragma unroll
	for (int c = 0; c < 4; c++) dq[c] += c2 * qh[c];
	// F.normalize backward: (dqh - qh (qh . dqh)) / |q|
	const float dot = qh[0] * dq[0] + qh[1] * dq[1] + qh[2] * dq[2] + qh[3] * dq[3];
#pragma unroll
	for (int c = 0; c < 4; c++) dL_dcam_rot[c] = (dq[c] - qh[c] * dot) / n;
}

inline int pose_blocks(int P)
{
	const int want = (P + POSE_THREADS - 1) / POSE_THREADS;
	return want < 1 ? 1 : (want < POSE_MAX_BLOCKS ? want : POSE_MAX_BLOCKS);
}

}  // namespace

}  // namespace gs2d

using namespace gs2d;

namespace gs2d {
namespace {

// Adam on the tracker's seven pose scalars (scene/Frame.py:66-73: torch.optim.Adam over cam_rot [4] and cam_trans [3],
// one learning rate each) in torch's operation order (torch/optim/adam.py: lerp for the first moment,
// sqrt(v) / sqrt(bc2) + eps, scalars computed in double on the host): one 7-thread launch instead of the ~25 foreach
// launches torch issues for two tiny parameters.
struct PoseAdamArgs {
	float* param[2];
	const float* grad[2];
	float* state;            // exp_avg[7] | exp_avg_sq[7], cam_rot first
	float step_size[2];      // lr / (1 - beta1^t), per parameter (torch keeps one step counter per parameter)
	float bc2_sqrt[2];
	float w1, beta2, w2, eps;
};

__global__ void pose_adam_kernel(const PoseAdamArgs a)
{
	const int i = threadIdx.x;
	if (i >= 7) return;
	const int grp = i < 4 ? 0 : 1, k = i < 4 ? i : i - 4;
	if (a.grad[grp] == nullptr) return;      // no gradient: parameter and moments untouched, like torch
	const float g = a.grad[grp][k];
	float m = a.state[i], v = a.state[7 + i], p = a.param[grp][k];
	const float diff = g - m;
	m = (a.w1 < 0.5f) ? m + a.w1 * diff : g - diff * (1.0f - a.w1);      // exp_avg.lerp_(grad, 1 - beta1)
	v = v * a.beta2 + a.w2 * g * g;                                        // mul_(beta2).addcmul_(grad, grad, 1 - beta2)
	const float denom = sqrtf(v) / a.bc2_sqrt[grp] + a.eps;
	p = p - a.step_size[grp] * (m / denom);                                // addcdiv_(exp_avg, denom, value=-step_size)
	a.state[i] = m;
	a.state[7 + i] = v;
	a.param[grp][k] = p;
}

}  // namespace
}  // namespace gs2d

extern "C" {

int gs2d_pose_adam_step(float* cam_rot, float* cam_trans, const float* dL_dcam_rot, const float* dL_dcam_trans,
                        float* state, const gs2d_pose_adam_params* ap, void* stream)
{
	if (!cam_rot || !cam_trans || !state || !ap) return GS2D_ERR_INVALID;
	if (!dL_dcam_rot && !dL_dcam_trans) return GS2D_OK;
	PoseAdamArgs a;
	a.param[0] = cam_rot; a.param[1] = cam_trans;
	a.grad[0] = dL_dcam_rot; a.grad[1] = dL_dcam_trans;
	a.state = state;
	for (int k = 0; k < 2; k++) {
		const long long t = ap->step[k];
		if (a.grad[k] && t < 1) return GS2D_ERR_INVALID;
		const double bc1 = 1.0 - pow(ap->beta1, (double)(t > 0 ? t : 1)), bc2 = 1.0 - pow(ap->beta2, (double)(t > 0 ? t : 1));
		a.step_size[k] = (float)(ap->lr[k] / bc1);
		a.bc2_sqrt[k] = (float)sqrt(bc2);
	}
	a.w1 = (float)(1.0 - ap->beta1);
	a.beta2 = (float)ap->beta2;
	a.w2 = (float)(1.0 - ap->beta2);
	a.eps = (float)ap->eps;
	cudaStream_t st = (cudaStream_t)stream;
	{
		ProfScope ps(K_POSE, st);
		pose_adam_kernel<<<1, 32, 0, st>>>(a);
		count_launch();
	}
	GS2D_CUDA_OK(cudaGetLastError());
	return GS2D_OK;
}

int gs2d_pose_transform_forward(int32_t P, const float* xyz, const float* rotations, const float* cam_rot,
                                const float* cam_trans, const float* w2c, float* means3D_out, float* rotations_out,
                                void* stream)
{
	if (P < 0) return GS2D_ERR_INVALID;
	if (P == 0) return GS2D_OK;
	const bool quat = cam_rot != nullptr && cam_trans != nullptr;
	if (!xyz || !means3D_out || (quat == (w2c != nullptr)) || (!quat && (cam_rot || cam_trans))) return GS2D_ERR_INVALID;
	if ((rotations == nullptr) != (rotations_out == nullptr)) return GS2D_ERR_INVALID;
	if (rotations && ((reinterpret_cast<uintptr_t>(rotations) | reinterpret_cast<uintptr_t>(rotations_out)) & 15u))
		return GS2D_ERR_INVALID;
	cudaStream_t st = (cudaStream_t)stream;
	{
		ProfScope ps(K_POSE, st);
		pose_transform_forward_kernel<<<pose_blocks(P), POSE_THREADS, 0, st>>>(P, xyz, rotations, cam_rot, cam_trans, w2c,
		                                                                      means3D_out, rotations_out);
		count_launch();
	}
	GS2D_CUDA_OK(cudaGetLastError());
	return GS2D_OK;
}

size_t gs2d_pose_transform_scratch_bytes(void)
{
	return sizeof(float) * 12 * POSE_MAX_BLOCKS + 128;     // per-block partials + the ticket
}

int gs2d_pose_transform_backward(int32_t P, const float* xyz, const float* cam_rot, const float* cam_trans,
                                 const float* w2c, const float* dL_dmeans3D, void* scratch, float* dL_dcam_rot,
                                 float* dL_dcam_trans, float* dL_dw2c, float* dL_dxyz, void* stream)
{
	if (P < 0 || !scratch) return GS2D_ERR_INVALID;
	const bool quat = cam_rot != nullptr && cam_trans != nullptr;
	if ((quat == (w2c != nullptr)) || (!quat && (cam_rot || cam_trans))) return GS2D_ERR_INVALID;
	if (quat ? (!dL_dcam_rot || !dL_dcam_trans || dL_dw2c) : (!dL_dw2c || dL_dcam_rot || dL_dcam_trans)) return GS2D_ERR_INVALID;
	if (P > 0 && (!xyz || !dL_dmeans3D)) return GS2D_ERR_INVALID;
	cudaStream_t st = (cudaStream_t)stream;
	float* partials = (float*)scratch;
	unsigned int* ticket = (unsigned int*)((char*)scratch + sizeof(float) * 12 * POSE_MAX_BLOCKS);
	{
		ProfScope ps(K_POSE, st);
		pose_transform_backward_kernel<<<pose_blocks(P), POSE_THREADS, 0, st>>>(
			P, xyz, cam_rot, cam_trans, w2c, dL_dmeans3D, partials, ticket, dL_dcam_rot, dL_dcam_trans, dL_dw2c, dL_dxyz);
		count_launch();
	}
	GS2D_CUDA_OK(cudaGetLastError());
	return GS2D_OK;
}

}  // extern "C"
