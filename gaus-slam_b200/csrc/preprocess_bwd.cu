// preprocess_bwd.cu -- per-surfel backward: gradient-row reduction + transform chain rule.
//
// Replaces BACKWARD::preprocess / preprocessCUDA<3> + compute_transmat_aabb (CR/backward.cu:466-664),
// quat_to_rotmat_vjp (CR/auxiliary.h:237-281), the SH backward (CR/backward.cu:20-139) and the nine
// torch::zeros fills of RAST/rasterize_points.cu:192-200: every output element is written here
// exactly once (zeros for culled surfels), so no memset of P x 112 B is needed.
//
// One thread per surfel.  A visible surfel first sums its tiles_touched gradient rows -- they are
// contiguous because rows are laid out in surfel-major instance order -- which replaces the
// reference's global atomics with a deterministic, atomic-free reduction.  HBM-bound:
// 80 B per instance + ~80 B in / 112 B out per surfel.
#include "common.cuh"
#include "math.cuh"

namespace gs2d {

namespace {

__device__ __forceinline__ float3 normalize_vjp(float3 v, float3 dv)
{
	// d/dv of v/|v| applied to dv
	float s2 = v.x * v.x + v.y * v.y + v.z * v.z;
	float inv = 1.0f / sqrtf(s2 * s2 * s2);
	float3 o;
	o.x = ((+s2 - v.x * v.x) * dv.x - v.y * v.x * dv.y - v.z * v.x * dv.z) * inv;
	o.y = (-v.x * v.y * dv.x + (s2 - v.y * v.y) * dv.y - v.z * v.y * dv.z) * inv;
	o.z = (-v.x * v.z * dv.x - v.y * v.z * dv.y + (s2 - v.z * v.z) * dv.z) * inv;
	return o;
}

// SH colour backward: dL/dsh[k] = basis_k * dL/dRGB and the part of dL/dmean that flows through
// the view direction.  Returns that mean gradient.
__device__ float3 sh_backward(int idx, int deg, int max_coeffs, float3 pos, float3 campos,
                              const float* __restrict__ shs, const uint8_t* __restrict__ clamped,
                              float3 dL_dRGB, float* __restrict__ dL_dshs)
{
	float3 dir_orig = make_float3(pos.x - campos.x, pos.y - campos.y, pos.z - campos.z);
	float len = sqrtf(dir_orig.x * dir_orig.x + dir_orig.y * dir_orig.y + dir_orig.z * dir_orig.z);
	const float x = dir_orig.x / len, y = dir_orig.y / len, z = dir_orig.z / len;
	const float3* sh = ((const float3*)shs) + (size_t)idx * max_coeffs;
	float3* out = ((float3*)dL_dshs) + (size_t)idx * max_coeffs;
	if (clamped[3 * idx + 0]) dL_dRGB.x = 0.f;
	if (clamped[3 * idx + 1]) dL_dRGB.y = 0.f;
	if (clamped[3 * idx + 2]) dL_dRGB.z = 0.f;

	const int ncoef = (deg + 1) * (deg + 1);
	float b[16], gx[16], gy[16], gz[16];
#pragma unroll
	for (int i = 0; i < 16; i++) { b[i] = 0.f; gx[i] = 0.f; gy[i] = 0.f; gz[i] = 0.f; }
	const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
	b[0] = kSH0;
	b[1] = -kSH1 * y;  gy[1] = -kSH1;
	b[2] = kSH1 * z;   gz[2] = kSH1;
	b[3] = -kSH1 * x;  gx[3] = -kSH1;
	b[4] = kSH2[0] * xy;  gx[4] = kSH2[0] * y;  gy[4] = kSH2[0] * x;
	b[5] = kSH2[1] * yz;  gy[5] = kSH2[1] * z;  gz[5] = kSH2[1] * y;
	b[6] = kSH2[2] * (2.f * zz - xx - yy);  gx[6] = kSH2[2] * -2.f * x;  gy[6] = kSH2[2] * -2.f * y;  gz[6] = kSH2[2] * 4.f * z;
	b[7] = kSH2[3] * xz;  gx[7] = kSH2[3] * z;  gz[7] = kSH2[3] * x;
	b[8] = kSH2[4] * (xx - yy);  gx[8] = kSH2[4] * 2.f * x;  gy[8] = kSH2[4] * -2.f * y;
	b[9] = kSH3[0] * y * (3.f * xx - yy);  gx[9] = kSH3[0] * 6.f * xy;  gy[9] = kSH3[0] * 3.f * (xx - yy);
	b[10] = kSH3[1] * xy * z;  gx[10] = kSH3[1] * yz;  gy[10] = kSH3[1] * xz;  gz[10] = kSH3[1] * xy;
	b[11] = kSH3[2] * y * (4.f * zz - xx - yy);  gx[11] = kSH3[2] * -2.f * xy;
	gy[11] = kSH3[2] * (-3.f * yy + 4.f * zz - xx);  gz[11] = kSH3[2] * 8.f * yz;
	b[12] = kSH3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy);  gx[12] = kSH3[3] * -6.f * xz;
	gy[12] = kSH3[3] * -6.f * yz;  gz[12] = kSH3[3] * 3.f * (2.f * zz - xx - yy);
	b[13] = kSH3[4] * x * (4.f * zz - xx - yy);  gx[13] = kSH3[4] * (-3.f * xx + 4.f * zz - yy);
	gy[13] = kSH3[4] * -2.f * xy;  gz[13] = kSH3[4] * 8.f * xz;
	b[14] = kSH3[5] * z * (xx - yy);  gx[14] = kSH3[5] * 2.f * xz;  gy[14] = kSH3[5] * -2.f * yz;
	gz[14] = kSH3[5] * (xx - yy);
	b[15] = kSH3[6] * x * (xx - 3.f * yy);  gx[15] = kSH3[6] * 3.f * (xx - yy);  gy[15] = kSH3[6] * -6.f * xy;

	float3 dL_ddir = make_float3(0.f, 0.f, 0.f);
#pragma unroll
	for (int kk = 0; kk < 16; kk++) {
		if (kk < ncoef && kk < max_coeffs) {
			out[kk] = make_float3(b[kk] * dL_dRGB.x, b[kk] * dL_dRGB.y, b[kk] * dL_dRGB.z);
			float dotv = sh[kk].x * dL_dRGB.x + sh[kk].y * dL_dRGB.y + sh[kk].z * dL_dRGB.z;
			dL_ddir.x += gx[kk] * dotv;
			dL_ddir.y += gy[kk] * dotv;
			dL_ddir.z += gz[kk] * dotv;
		}
	}
	return normalize_vjp(dir_orig, dL_ddir);
}

// multimem.red: the NVSwitch adds the value into EVERY rank's copy of the address (an NVLS multicast mapping).
#ifdef __CUDACC__
__device__ __forceinline__ void mc_red(float* mc, float v)
{
	asm volatile("multimem.red.relaxed.sys.global.add.f32 [%0], %1;" ::"l"(mc), "f"(v) : "memory");
}
__device__ __forceinline__ void mc_red2(float* mc, float a, float b)
{
	asm volatile("multimem.red.relaxed.sys.global.add.v2.f32 [%0], {%1, %2};" ::"l"(mc), "f"(a), "f"(b) : "memory");
}
__device__ __forceinline__ void mc_red4(float* mc, float4 v)
{
	asm volatile("multimem.red.relaxed.sys.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "f"(v.x), "f"(v.y),
	             "f"(v.z), "f"(v.w)
	             : "memory");
}
#else   // host compilation of this file (no PTX): one "rank", the multicast address is the buffer itself
inline void mc_red(float* mc, float v) { mc[0] += v; }
inline void mc_red2(float* mc, float a, float b) { mc[0] += a; mc[1] += b; }
inline void mc_red4(float* mc, float4 v) { mc[0] += v.x; mc[1] += v.y; mc[2] += v.z; mc[3] += v.w; }
#endif

// REDUCE: the five map-parameter gradients (xyz | opacity | scale | rotation | rgb, the flat [13 * P] layout of
// gs2d_b200/mapping.py) are not stored but ADDED, through the switch, into every rank's copy of a symmetric buffer
// that was zeroed beforehand (mc_flat = its multicast address): the all-reduce of the view-sharded mapping step
// happens in this kernel's epilogue, tile of surfels by tile of surfels, instead of in a pass of its own.  Surfels
// that are not visible in this rank's view add nothing.
#ifndef GS2D_PBWD_ROWS2
#define GS2D_PBWD_ROWS2 1
#endif
#ifndef GS2D_PBWD_MIN_BLOCKS
#define GS2D_PBWD_MIN_BLOCKS 4
#endif
template <bool REDUCE>
__global__ void __launch_bounds__(256, GS2D_PBWD_MIN_BLOCKS)
preprocess_backward_kernel(int P, int D, int M,
	const float3* __restrict__ means3D, const int* __restrict__ radii, const float* __restrict__ shs,
	const uint8_t* __restrict__ clamped, const float2* __restrict__ scales,
	const float4* __restrict__ rotations, const float* __restrict__ transMat_precomp,
	const float* __restrict__ splat, const float* __restrict__ viewmatrix,
	const float* __restrict__ projmatrix, float focal_x, float focal_y, float tan_fovx, float tan_fovy,
	const float* __restrict__ campos, const uint32_t* __restrict__ tiles_touched,
	const uint32_t* __restrict__ offsets, const float* __restrict__ grad_rows,
	float* __restrict__ dL_dtransMat, float* __restrict__ dL_dcolors, float* __restrict__ dL_dopacity,
	float* __restrict__ dL_dsh, float* __restrict__ dL_dmean2D, float* __restrict__ dL_dmean3D,
	float* __restrict__ dL_dscale, float* __restrict__ dL_drot, float* mc_flat)
{
	const int idx = blockIdx.x * blockDim.x + threadIdx.x;
	if (idx >= P) return;

	float dT[9];
#pragma unroll
	for (int i = 0; i < 9; i++) dT[i] = 0.f;
	float dm2x = 0.f, dm2y = 0.f, dopa = 0.f;
	float3 dnrm = make_float3(0.f, 0.f, 0.f), dcol = make_float3(0.f, 0.f, 0.f);
	float3 dmean = make_float3(0.f, 0.f, 0.f);
	float2 dscale = make_float2(0.f, 0.f);
	float4 drot = make_float4(0.f, 0.f, 0.f, 0.f);
	float dmean2D_x = 0.f, dmean2D_y = 0.f;
	// everything a visible surfel needs besides its rows is requested up front, next to the visibility flag itself: the
	// kernel is a chain of dependent global loads (flag -> row range -> rows), and this takes one link out of it
	const int radius = radii[idx];
	const uint32_t cnt_all = tiles_touched[idx];
	const uint32_t start_all = idx == 0 ? 0u : offsets[idx - 1];
	const bool visible = radius > 0;

	if (visible) {
		const uint32_t cnt = cnt_all;
		const uint32_t start = start_all;
		const float4* rows = (const float4*)(grad_rows + (size_t)start * GRAD_WORDS);
		auto add_row = [&](const float4 a, const float4 b, const float4 c, const float4 d, const float4 e) {
			dT[0] += a.x; dT[1] += a.y; dT[2] += a.z; dT[3] += a.w;
			dT[4] += b.x; dT[5] += b.y; dT[6] += b.z; dT[7] += b.w;
			dT[8] += c.x; dm2x += c.y; dm2y += c.z; dopa += c.w;
			dnrm.x += d.x; dnrm.y += d.y; dnrm.z += d.z; dcol.x += d.w;
			dcol.y += e.x; dcol.z += e.y;
		};
#if GS2D_PBWD_ROWS2
		// two rows per round, all ten loads requested before the first is used (a surfel has 2.1 rows on average; the
		// kernel waits on memory, not on arithmetic); the additions stay in row order
		uint32_t r = 0;
		for (; r + 2 <= cnt; r += 2) {
			const float4 a0 = rows[5 * r + 0], b0 = rows[5 * r + 1], c0 = rows[5 * r + 2], d0 = rows[5 * r + 3],
			             e0 = rows[5 * r + 4];
			const float4 a1 = rows[5 * r + 5], b1 = rows[5 * r + 6], c1 = rows[5 * r + 7], d1 = rows[5 * r + 8],
			             e1 = rows[5 * r + 9];
			add_row(a0, b0, c0, d0, e0);
			add_row(a1, b1, c1, d1, e1);
		}
		if (r < cnt) add_row(rows[5 * r + 0], rows[5 * r + 1], rows[5 * r + 2], rows[5 * r + 3], rows[5 * r + 4]);
#else
		for (uint32_t r = 0; r < cnt; r++)
			add_row(rows[5 * r + 0], rows[5 * r + 1], rows[5 * r + 2], rows[5 * r + 3], rows[5 * r + 4]);
#endif

		const int W = int(focal_x * tan_fovx * 2);
		const int H = int(focal_y * tan_fovy * 2);
		const float3 p_orig = means3D[idx];
		float Tw8;   // T[2][2] as stored by the forward (view depth of the centre)

		if (scales == nullptr) {
			// precomputed transform: only the centre term is folded back into dL/dT
			const float* t = transMat_precomp + 9 * (size_t)idx;
			Tw8 = t[8];
			if (dm2x != 0 || dm2y != 0) {
				const float3 Tu = make_float3(t[0], t[1], t[2]), Tv = make_float3(t[3], t[4], t[5]),
				             Tw = make_float3(t[6], t[7], t[8]);
				const float distance = Tw.x * Tw.x + Tw.y * Tw.y - Tw.z * Tw.z;
				const float f = 1 / (distance);
				dT[0] += dm2x * (f * Tw.x); dT[1] += dm2x * (f * Tw.y); dT[2] += dm2x * (-f * Tw.z);
				dT[3] += dm2y * (f * Tw.x); dT[4] += dm2y * (f * Tw.y); dT[5] += dm2y * (-f * Tw.z);
				dT[6] += dm2x * (Tu.x * (f - 2 * f * f * Tw.x * Tw.x)) + dm2y * (Tv.x * (f - 2 * f * f * Tw.x * Tw.x));
				dT[7] += dm2x * (Tu.y * (f - 2 * f * f * Tw.y * Tw.y)) + dm2y * (Tv.y * (f - 2 * f * f * Tw.y * Tw.y));
				dT[8] += dm2x * (-Tu.z * (f + 2 * f * f * Tw.z * Tw.z)) + dm2y * (-Tv.z * (f + 2 * f * f * Tw.z * Tw.z));
			}
			dmean2D_x = dT[2] * Tw8 * 0.5f * float(W);
			dmean2D_y = dT[5] * Tw8 * 0.5f * float(H);
		} else {
			Tw8 = splat[(size_t)idx * REC_WORDS + 10];
			// densification proxy uses the raw accumulated dL/dT (CR/backward.cu:660-663)
			dmean2D_x = dT[2] * Tw8 * 0.5f * float(W);
			dmean2D_y = dT[5] * Tw8 * 0.5f * float(H);

			const float4 q = rotations[idx];
			const float2 sc = scales[idx];
			const Mat<3, 3> R = quat_to_rot(q.x, q.y, q.z, q.w);
			const Mat<3, 3> S = scale_mat(sc.x, sc.y, 1.0f);   // scale_modifier ignored, like the reference
			const Mat<3, 3> L = matmul<3, 3, 3>(R, S);
			const Mat<3, 4> Mm = splat2world_mat(L, p_orig);
			const Mat<4, 4> w2n = world2ndc_mat(projmatrix);
			const Mat<3, 4> n2p = ndc2pix_mat(W, H);
			const Mat<3, 4> Pm = matmul<4, 4, 3>(w2n, n2p);
			const Mat<3, 3> T = matmul<4, 3, 3>(transpose(Mm), Pm);
			const float3 normal = xform_vec_4x3(make_float3(L.c[2][0], L.c[2][1], L.c[2][2]), viewmatrix);

			Mat<3, 3> dLdT;
#pragma unroll
			for (int c = 0; c < 3; c++)
#pragma unroll
				for (int r = 0; r < 3; r++) dLdT.c[c][r] = dT[3 * c + r];
			if (dm2x != 0 || dm2y != 0) {
				const float3 Tu = make_float3(T.c[0][0], T.c[0][1], T.c[0][2]);
				const float3 Tv = make_float3(T.c[1][0], T.c[1][1], T.c[1][2]);
				const float3 Tw = make_float3(T.c[2][0], T.c[2][1], T.c[2][2]);
				const float distance = Tw.x * Tw.x + Tw.y * Tw.y - Tw.z * Tw.z;
				const float f = 1 / (distance);
				const float dpx_dT00 = f * Tw.x, dpx_dT01 = f * Tw.y, dpx_dT02 = -f * Tw.z;
				const float dpy_dT10 = f * Tw.x, dpy_dT11 = f * Tw.y, dpy_dT12 = -f * Tw.z;
				const float dpx_dT30 = Tu.x * (f - 2 * f * f * Tw.x * Tw.x);
				const float dpx_dT31 = Tu.y * (f - 2 * f * f * Tw.y * Tw.y);
				const float dpx_dT32 = -Tu.z * (f + 2 * f * f * Tw.z * Tw.z);
				const float dpy_dT30 = Tv.x * (f - 2 * f * f * Tw.x * Tw.x);
				const float dpy_dT31 = Tv.y * (f - 2 * f * f * Tw.y * Tw.y);
				const float dpy_dT32 = -Tv.z * (f + 2 * f * f * Tw.z * Tw.z);
				dLdT.c[0][0] += dm2x * dpx_dT00; dLdT.c[0][1] += dm2x * dpx_dT01; dLdT.c[0][2] += dm2x * dpx_dT02;
				dLdT.c[1][0] += dm2y * dpy_dT10; dLdT.c[1][1] += dm2y * dpy_dT11; dLdT.c[1][2] += dm2y * dpy_dT12;
				dLdT.c[2][0] += dm2x * dpx_dT30 + dm2y * dpy_dT30;
				dLdT.c[2][1] += dm2x * dpx_dT31 + dm2y * dpy_dT31;
				dLdT.c[2][2] += dm2x * dpx_dT32 + dm2y * dpy_dT32;
			}
			const Mat<3, 4> dLdM = matmul<3, 4, 3>(Pm, transpose(dLdT));
			float3 dL_dtn = xform_vec_4x3_T(dnrm, viewmatrix);
			const float3 p_view = xform_point_4x3(p_orig, viewmatrix);
			const float cosv = -sum3(mul3(p_view, normal));
			const float multiplier = cosv > 0 ? 1 : -1;
			dL_dtn = scale3(multiplier, dL_dtn);

			Mat<3, 3> dRS;
			dRS.c[0][0] = dLdM.c[0][0]; dRS.c[0][1] = dLdM.c[0][1]; dRS.c[0][2] = dLdM.c[0][2];
			dRS.c[1][0] = dLdM.c[1][0]; dRS.c[1][1] = dLdM.c[1][1]; dRS.c[1][2] = dLdM.c[1][2];
			dRS.c[2][0] = dL_dtn.x; dRS.c[2][1] = dL_dtn.y; dRS.c[2][2] = dL_dtn.z;
			Mat<3, 3> vR;
#pragma unroll
			for (int r = 0; r < 3; r++) {
				vR.c[0][r] = dRS.c[0][r] * sc.x;
				vR.c[1][r] = dRS.c[1][r] * sc.y;
				vR.c[2][r] = dRS.c[2][r];
			}
			{
				// VJP of quat_to_rot w.r.t. the normalised quaternion (no normalisation Jacobian)
				const float s = rsqrtf(q.w * q.w + q.x * q.x + q.y * q.y + q.z * q.z);
				const float w = q.x * s, x = q.y * s, y = q.z * s, z = q.w * s;
				drot.x = 2.f * (x * (vR.c[1][2] - vR.c[2][1]) + y * (vR.c[2][0] - vR.c[0][2]) +
				                z * (vR.c[0][1] - vR.c[1][0]));
				drot.y = 2.f * (-2.f * x * (vR.c[1][1] + vR.c[2][2]) + y * (vR.c[0][1] + vR.c[1][0]) +
				                z * (vR.c[0][2] + vR.c[2][0]) + w * (vR.c[1][2] - vR.c[2][1]));
				drot.z = 2.f * (x * (vR.c[0][1] + vR.c[1][0]) - 2.f * y * (vR.c[0][0] + vR.c[2][2]) +
				                z * (vR.c[1][2] + vR.c[2][1]) + w * (vR.c[2][0] - vR.c[0][2]));
				drot.w = 2.f * (x * (vR.c[0][2] + vR.c[2][0]) + y * (vR.c[1][2] + vR.c[2][1]) -
				                2.f * z * (vR.c[0][0] + vR.c[1][1]) + w * (vR.c[0][1] - vR.c[1][0]));
			}
			dscale.x = dRS.c[0][0] * R.c[0][0] + dRS.c[0][1] * R.c[0][1] + dRS.c[0][2] * R.c[0][2];
			dscale.y = dRS.c[1][0] * R.c[1][0] + dRS.c[1][1] * R.c[1][1] + dRS.c[1][2] * R.c[1][2];
			dmean = make_float3(dLdM.c[2][0], dLdM.c[2][1], dLdM.c[2][2]);
		}

		if (shs != nullptr) {
			const float3 cp = make_float3(campos[0], campos[1], campos[2]);
			const float3 add = sh_backward(idx, D, M, p_orig, cp, shs, clamped, dcol, dL_dsh);
			dmean.x += add.x; dmean.y += add.y; dmean.z += add.z;
		}
	} else if (shs != nullptr && M > 0) {
		float3* out = ((float3*)dL_dsh) + (size_t)idx * M;
		for (int kk = 0; kk < M; kk++) out[kk] = make_float3(0.f, 0.f, 0.f);
	}
	if (visible && shs != nullptr) {
		// coefficients above the active degree receive no gradient
		const int ncoef = (D + 1) * (D + 1);
		float3* out = ((float3*)dL_dsh) + (size_t)idx * M;
		for (int kk = ncoef; kk < M; kk++) out[kk] = make_float3(0.f, 0.f, 0.f);
	}

#pragma unroll
	for (int i = 0; i < 9; i++) dL_dtransMat[9 * (size_t)idx + i] = dT[i];
	dL_dmean2D[3 * (size_t)idx + 0] = dmean2D_x;
	dL_dmean2D[3 * (size_t)idx + 1] = dmean2D_y;
	dL_dmean2D[3 * (size_t)idx + 2] = 0.f;
	if (REDUCE) {
		if (visible) {
			const size_t Ps = (size_t)P, i = (size_t)idx;
			mc_red(mc_flat + 3 * i + 0, dmean.x);
			mc_red(mc_flat + 3 * i + 1, dmean.y);
			mc_red(mc_flat + 3 * i + 2, dmean.z);
			mc_red(mc_flat + 3 * Ps + i, dopa);
			mc_red2(mc_flat + 4 * Ps + 2 * i, dscale.x, dscale.y);
			mc_red4(mc_flat + 6 * Ps + 4 * i, drot);
			mc_red(mc_flat + 10 * Ps + 3 * i + 0, dcol.x);
			mc_red(mc_flat + 10 * Ps + 3 * i + 1, dcol.y);
			mc_red(mc_flat + 10 * Ps + 3 * i + 2, dcol.z);
		}
		return;
	}
	dL_dcolors[3 * (size_t)idx + 0] = dcol.x;
	dL_dcolors[3 * (size_t)idx + 1] = dcol.y;
	dL_dcolors[3 * (size_t)idx + 2] = dcol.z;
	dL_dopacity[idx] = dopa;
	dL_dmean3D[3 * (size_t)idx + 0] = dmean.x;
	dL_dmean3D[3 * (size_t)idx + 1] = dmean.y;
	dL_dmean3D[3 * (size_t)idx + 2] = dmean.z;
	dL_dscale[2 * (size_t)idx + 0] = dscale.x;
	dL_dscale[2 * (size_t)idx + 1] = dscale.y;
	((float4*)dL_drot)[idx] = drot;
}

}  // namespace

int launch_preprocess_backward(const gs2d_settings& s, const gs2d_inputs& in, const GeomState& g,
                               const int32_t* radii, const float* grad_rows, const gs2d_grads& out,
                               cudaStream_t st, float* mc_flat)
{
	const int P = s.P;
	const float focal_y = s.image_height / (2.0f * s.tanfovy);
	const float focal_x = s.image_width / (2.0f * s.tanfovx);
	{
		ProfScope ps(K_PREPROCESS_BWD, st);
		if (mc_flat)
			preprocess_backward_kernel<true><<<(P + 255) / 256, 256, 0, st>>>(
				P, s.sh_degree, s.sh_coeffs, (const float3*)in.means3D, radii, in.shs, g.clamped,
				(const float2*)in.scales, (const float4*)in.rotations, in.transMat_precomp, g.splat,
				in.viewmatrix, in.projmatrix, focal_x, focal_y, s.tanfovx, s.tanfovy, in.campos,
				g.tiles_touched, g.point_offsets, grad_rows, out.dL_dtransMat, out.dL_dcolors, out.dL_dopacity,
				out.dL_dsh, out.dL_dmeans2D, out.dL_dmeans3D, out.dL_dscales, out.dL_drotations, mc_flat);
		else
			preprocess_backward_kernel<false><<<(P + 255) / 256, 256, 0, st>>>(
				P, s.sh_degree, s.sh_coeffs, (const float3*)in.means3D, radii, in.shs, g.clamped,
				(const float2*)in.scales, (const float4*)in.rotations, in.transMat_precomp, g.splat,
				in.viewmatrix, in.projmatrix, focal_x, focal_y, s.tanfovx, s.tanfovy, in.campos,
				g.tiles_touched, g.point_offsets, grad_rows, out.dL_dtransMat, out.dL_dcolors, out.dL_dopacity,
				out.dL_dsh, out.dL_dmeans2D, out.dL_dmeans3D, out.dL_dscales, out.dL_drotations, nullptr);
		count_launch();
	}
	GS2D_CUDA_OK(cudaGetLastError());
	return GS2D_OK;
}

}  // namespace gs2d
