// math.cuh -- small fixed-size matrix helpers for the per-surfel kernels.
//
// Storage is column-major, M.c[col][row], and every product is the textbook
// Result[c][r] = sum_k A[k][r] * B[c][k] accumulated left to right in k.  The summation order
// matters: tile keys, radii and tile counts have to be bit-identical to the reference
// (BASELINE.json), whose GLM products have this same order (CR/forward.cu:113,
// CR/backward.cu:526-527, :591), and nvcc's FMA contraction is a function of the expression
// tree.  Nothing here is approximated (no fast-math, IEEE division and sqrt).
#pragma once
#include <cuda_runtime.h>

namespace gs2d {

template <int C, int R>
struct Mat {
	float c[C][R];
};

// (R x K) * (K x C): A has K columns of R rows, B has C columns of K rows.
template <int K, int R, int C>
__device__ __forceinline__ Mat<C, R> matmul(const Mat<K, R>& A, const Mat<C, K>& B)
{
	Mat<C, R> out;
#pragma unroll
	for (int c = 0; c < C; c++) {
#pragma unroll
		for (int r = 0; r < R; r++) {
			float acc = A.c[0][r] * B.c[c][0];
#pragma unroll
			for (int k = 1; k < K; k++) acc = acc + A.c[k][r] * B.c[c][k];
			out.c[c][r] = acc;
		}
	}
	return out;
}

template <int C, int R>
__device__ __forceinline__ Mat<R, C> transpose(const Mat<C, R>& A)
{
	Mat<R, C> out;
#pragma unroll
	for (int c = 0; c < C; c++)
#pragma unroll
		for (int r = 0; r < R; r++) out.c[r][c] = A.c[c][r];
	return out;
}

__device__ __forceinline__ float3 mul3(float3 a, float3 b) { return make_float3(a.x * b.x, a.y * b.y, a.z * b.z); }
__device__ __forceinline__ float3 scale3(float f, float3 a) { return make_float3(f * a.x, f * a.y, f * a.z); }
__device__ __forceinline__ float sum3(float3 a) { return a.x + a.y + a.z; }

// p' = M * (p, 1) for a 4x4 stored as 16 floats with element (row r, col c) at m[4c + r]
// (CR/auxiliary.h:78-97: the torch tensors handed in are the transposed matrices).
__device__ __forceinline__ float3 xform_point_4x3(const float3& p, const float* m)
{
	return make_float3(m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12],
	                   m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13],
	                   m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14]);
}
__device__ __forceinline__ float3 xform_vec_4x3(const float3& p, const float* m)
{
	return make_float3(m[0] * p.x + m[4] * p.y + m[8] * p.z,
	                   m[1] * p.x + m[5] * p.y + m[9] * p.z,
	                   m[2] * p.x + m[6] * p.y + m[10] * p.z);
}
__device__ __forceinline__ float3 xform_vec_4x3_T(const float3& p, const float* m)
{
	return make_float3(m[0] * p.x + m[1] * p.y + m[2] * p.z,
	                   m[4] * p.x + m[5] * p.y + m[6] * p.z,
	                   m[8] * p.x + m[9] * p.y + m[10] * p.z);
}

// Unit quaternion (w,x,y,z) -> rotation matrix; q is normalised here with rsqrtf like the
// reference does (CR/auxiliary.h:212-234).
__device__ __forceinline__ Mat<3, 3> quat_to_rot(float q0, float q1, float q2, float q3)
{
	float s = rsqrtf(q3 * q3 + q0 * q0 + q1 * q1 + q2 * q2);
	float w = q0 * s, x = q1 * s, y = q2 * s, z = q3 * s;
	Mat<3, 3> R;
	R.c[0][0] = 1.f - 2.f * (y * y + z * z);
	R.c[0][1] = 2.f * (x * y + w * z);
	R.c[0][2] = 2.f * (x * z - w * y);
	R.c[1][0] = 2.f * (x * y - w * z);
	R.c[1][1] = 1.f - 2.f * (x * x + z * z);
	R.c[1][2] = 2.f * (y * z + w * x);
	R.c[2][0] = 2.f * (x * z + w * y);
	R.c[2][1] = 2.f * (y * z - w * x);
	R.c[2][2] = 1.f - 2.f * (x * x + y * y);
	return R;
}

__device__ __forceinline__ Mat<3, 3> scale_mat(float sx, float sy, float mod)
{
	Mat<3, 3> S;
#pragma unroll
	for (int c = 0; c < 3; c++)
#pragma unroll
		for (int r = 0; r < 3; r++) S.c[c][r] = (c == r) ? 1.f : 0.f;
	S.c[0][0] = mod * sx;
	S.c[1][1] = mod * sy;
	return S;
}

// world -> NDC as a column-major 4x4 built from the 16 floats handed in
// (CR/forward.cu:99-104: column c = (pm[c], pm[4+c], pm[8+c], pm[12+c])).
__device__ __forceinline__ Mat<4, 4> world2ndc_mat(const float* pm)
{
	Mat<4, 4> B;
#pragma unroll
	for (int c = 0; c < 4; c++)
#pragma unroll
		for (int r = 0; r < 4; r++) B.c[c][r] = pm[4 * r + c];
	return B;
}

// NDC -> pixel, 3 columns of 4 rows (CR/forward.cu:106-110).
__device__ __forceinline__ Mat<3, 4> ndc2pix_mat(int W, int H)
{
	Mat<3, 4> N;
	N.c[0][0] = float(W) / 2.0; N.c[0][1] = 0.0f; N.c[0][2] = 0.0f; N.c[0][3] = float(W - 1) / 2.0;
	N.c[1][0] = 0.0f; N.c[1][1] = float(H) / 2.0; N.c[1][2] = 0.0f; N.c[1][3] = float(H - 1) / 2.0;
	N.c[2][0] = 0.0f; N.c[2][1] = 0.0f; N.c[2][2] = 0.0f; N.c[2][3] = 1.0f;
	return N;
}

// splat -> world, 3 columns of 4 rows: (L0,0), (L1,0), (p,1)  (CR/forward.cu:93-97).
__device__ __forceinline__ Mat<3, 4> splat2world_mat(const Mat<3, 3>& L, const float3& p)
{
	Mat<3, 4> M;
	M.c[0][0] = L.c[0][0]; M.c[0][1] = L.c[0][1]; M.c[0][2] = L.c[0][2]; M.c[0][3] = 0.0f;
	M.c[1][0] = L.c[1][0]; M.c[1][1] = L.c[1][1]; M.c[1][2] = L.c[1][2]; M.c[1][3] = 0.0f;
	M.c[2][0] = p.x; M.c[2][1] = p.y; M.c[2][2] = p.z; M.c[2][3] = 1.0f;
	return M;
}

// Tile rectangle of a screen-space disc (CR/auxiliary.h:66-76): note the int radius and the
// truncation toward zero before the clamp.
__device__ __forceinline__ void tile_rect(float px, float py, int max_radius, unsigned gx, unsigned gy,
                                          uint2& rmin, uint2& rmax)
{
	rmin.x = min(gx, (unsigned)max((int)0, (int)((px - max_radius) / TILE)));
	rmin.y = min(gy, (unsigned)max((int)0, (int)((py - max_radius) / TILE)));
	rmax.x = min(gx, (unsigned)max((int)0, (int)((px + max_radius + TILE - 1) / TILE)));
	rmax.y = min(gy, (unsigned)max((int)0, (int)((py + max_radius + TILE - 1) / TILE)));
}

// Spherical-harmonics basis constants (the standard real SH normalisation used by 3DGS).
__device__ const float kSH0 = 0.28209479177387814f;
__device__ const float kSH1 = 0.4886025119029199f;
__device__ const float kSH2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                                  -1.0925484305920792f, 0.5462742152960396f};
__device__ const float kSH3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                                  0.3731763325901154f, -0.4570457994644658f, 1.445305721320277f,
                                  -0.5900435899266435f};

}  // namespace gs2d
