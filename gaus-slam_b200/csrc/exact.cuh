// exact.cuh -- the bit-critical part of the per-surfel transform, written as an explicit
// sequence of IEEE fp32 operations (__fmul_rn / __fmaf_rn / __fadd_rn never contract or reorder).
//
// Tile keys, radii, tile counts and therefore the sorted instance lists must be bit-identical to
// the reference (BASELINE.json), and those integers are decided by ceil()/trunc() of fp32 values
// computed here: view depth, the 3x3 ray-splat transform T, the 3-sigma bounding box.  The
// reference leaves the fusion of its a*b + c*d + ... expressions to nvcc's default -fmad=true, so
// "the reference's arithmetic" is the operation sequence its compiled kernel executes
// (CR/forward.cu:75-147, :176-238 and CR/auxiliary.h:78-97, :184-234 as built by the plain nvcc recipe of the test tree
// for sm_100a).  That sequence was read off the compiled kernel (SASS, after ptxas' own fusion) once and is restated below
// operation by operation; each helper names the source expression it stands for.  Products with
// the literal zeros of the reference's scale / ndc2pix matrices are dropped: for finite inputs
// x*0 only contributes a signed zero to a sum.
#pragma once
#include <cuda_runtime.h>

namespace gs2d {

// a0*b0 + a1*b1 + a2*b2 as the reference evaluates its 3-term dot products:
//   round(a1*b1) -> fma(a0, b0, .) -> fma(a2, b2, .)
__device__ __forceinline__ float dot3_ref(float a0, float b0, float a1, float b1, float a2, float b2)
{
	return __fmaf_rn(a2, b2, __fmaf_rn(a0, b0, __fmul_rn(a1, b1)));
}

// m * (p, 1), rows 0..2, m[4c + r] (transformPoint4x3, CR/auxiliary.h:78-86)
__device__ __forceinline__ float3 xform_point_exact(const float3 p, const float* __restrict__ m)
{
	return make_float3(__fadd_rn(m[12], dot3_ref(p.x, m[0], p.y, m[4], p.z, m[8])),
	                   __fadd_rn(m[13], dot3_ref(p.x, m[1], p.y, m[5], p.z, m[9])),
	                   __fadd_rn(m[14], dot3_ref(p.x, m[2], p.y, m[6], p.z, m[10])));
}

// m * (v, 0) (transformVec4x3, CR/auxiliary.h:100-108)
__device__ __forceinline__ float3 xform_vec_exact(const float3 v, const float* __restrict__ m)
{
	return make_float3(dot3_ref(v.x, m[0], v.y, m[4], v.z, m[8]),
	                   dot3_ref(v.x, m[1], v.y, m[5], v.z, m[9]),
	                   dot3_ref(v.x, m[2], v.y, m[6], v.z, m[10]));
}

struct SplatFrame {
	float3 Tu, Tv, Tw;   // rows of the ray-splat transform (transMat[0..8])
	float3 L0, L1, L2;   // columns of R * diag(sx, sy, 1)
	float3 R0, R1;       // first two columns of R (used by the backward pass)
};

// compute_transmat (CR/forward.cu:75-115) with quat_to_rotmat / scale_to_mat
// (CR/auxiliary.h:212-234, :285-291).  q = (w, x, y, z) as stored; mod = scale modifier.
__device__ __forceinline__ SplatFrame splat_frame_exact(const float3 p, const float2 scale, const float mod,
                                                        const float4 q, const float* __restrict__ pm,
                                                        const int W, const int H)
{
	SplatFrame f;
	const float n = __fmaf_rn(q.z, q.z, __fmaf_rn(q.y, q.y, __fmaf_rn(q.x, q.x, __fmul_rn(q.w, q.w))));
	const float s = rsqrtf(n);
	const float w = __fmul_rn(q.x, s), x = __fmul_rn(q.y, s), y = __fmul_rn(q.z, s), z = __fmul_rn(q.w, s);
	const float yy = __fmul_rn(y, y), zz = __fmul_rn(z, z);
	const float wz = __fmul_rn(w, z), wy = __fmul_rn(w, y), wx = __fmul_rn(w, x);
	// R columns (GLM is column-major): R0 = (R00, R01, R02) ...  The cross terms are
	// fma(a, b, +-round(c*d)); the diagonal is 1 - 2*fma(x, x, yy) etc., R00 is 1 - 2*(yy + zz).
	const float R00 = __fsub_rn(1.0f, __fmul_rn(2.0f, __fadd_rn(yy, zz)));
	const float R01 = __fmul_rn(2.0f, __fmaf_rn(x, y, wz));
	const float R02 = __fmul_rn(2.0f, __fmaf_rn(x, z, -wy));
	const float R10 = __fmul_rn(2.0f, __fmaf_rn(x, y, -wz));
	const float R11 = __fsub_rn(1.0f, __fmul_rn(2.0f, __fmaf_rn(x, x, zz)));
	const float R12 = __fmul_rn(2.0f, __fmaf_rn(y, z, wx));
	const float R20 = __fmul_rn(2.0f, __fmaf_rn(x, z, wy));
	const float R21 = __fmul_rn(2.0f, __fmaf_rn(y, z, -wx));
	const float R22 = __fsub_rn(1.0f, __fmul_rn(2.0f, __fmaf_rn(x, x, yy)));
	const float sx = __fmul_rn(mod, scale.x), sy = __fmul_rn(mod, scale.y);
	f.R0 = make_float3(R00, R01, R02);
	f.R1 = make_float3(R10, R11, R12);
	f.L0 = make_float3(__fmul_rn(sx, R00), __fmul_rn(sx, R01), __fmul_rn(sx, R02));
	f.L1 = make_float3(__fmul_rn(sy, R10), __fmul_rn(sy, R11), __fmul_rn(sy, R12));
	f.L2 = make_float3(R20, R21, R22);

	// X = splat2world^T * world2ndc: X[c][r] = <row r of (L0 | L1 | p), column c of the 4x4> (+ pm[12+c] for r = 2)
	float X[4][3];
#pragma unroll
	for (int c = 0; c < 4; c++) {
		X[c][0] = dot3_ref(f.L0.x, pm[c], f.L0.y, pm[4 + c], f.L0.z, pm[8 + c]);
		X[c][1] = dot3_ref(f.L1.x, pm[c], f.L1.y, pm[4 + c], f.L1.z, pm[8 + c]);
		X[c][2] = __fadd_rn(pm[12 + c], dot3_ref(p.x, pm[c], p.y, pm[4 + c], p.z, pm[8 + c]));
	}
	// T = X * ndc2pix with ndc2pix = [(W/2,0,0,(W-1)/2), (0,H/2,0,(H-1)/2), (0,0,0,1)]
	const float Wh = __fmul_rn((float)W, 0.5f), Wm = __fmul_rn((float)(W - 1), 0.5f);
	const float Hh = __fmul_rn((float)H, 0.5f), Hm = __fmul_rn((float)(H - 1), 0.5f);
	f.Tu = make_float3(__fmaf_rn(Wm, X[3][0], __fmul_rn(Wh, X[0][0])), __fmaf_rn(Wm, X[3][1], __fmul_rn(Wh, X[0][1])),
	                   __fmaf_rn(Wm, X[3][2], __fmul_rn(Wh, X[0][2])));
	f.Tv = make_float3(__fmaf_rn(Hm, X[3][0], __fmul_rn(Hh, X[1][0])), __fmaf_rn(Hm, X[3][1], __fmul_rn(Hh, X[1][1])),
	                   __fmaf_rn(Hm, X[3][2], __fmul_rn(Hh, X[1][2])));
	f.Tw = make_float3(X[3][0], X[3][1], X[3][2]);
	return f;
}

// compute_aabb at cutoff 3 (CR/forward.cu:119-147): bounding-box centre and half extent.
// Returns false when the reference returns false (distance == 0).
__device__ __forceinline__ bool aabb_exact(const float3 Tu, const float3 Tv, const float3 Tw, float2& centre,
                                           float2& extent)
{
	const float xx = __fmul_rn(Tw.x, Tw.x), yy = __fmul_rn(Tw.y, Tw.y);
	// sumf3(T3*T3*(9,9,-1)): round(yy*9) -> fma(xx, 9, .) -> fma(-Tw.z, Tw.z, .)
	const float distance = __fmaf_rn(-Tw.z, Tw.z, __fmaf_rn(xx, 9.0f, __fmul_rn(yy, 9.0f)));
	if (distance == 0.0f) return false;
	const float r = __frcp_rn(distance);
	const float nr = -r, f9 = __fmul_rn(r, 9.0f);
	const float a0 = __fmul_rn(f9, Tu.x), a1 = __fmul_rn(f9, Tu.y), a2 = __fmul_rn(Tu.z, nr);
	const float b0 = __fmul_rn(f9, Tv.x), b1 = __fmul_rn(f9, Tv.y), b2 = __fmul_rn(Tv.z, nr);
	// sumf3(f*T0*T3) and sumf3(f*T0*T0): round(x term) -> fma(y term) -> fma(z term)
	const float cx = __fmaf_rn(a2, Tw.z, __fmaf_rn(a1, Tw.y, __fmul_rn(a0, Tw.x)));
	const float cy = __fmaf_rn(b2, Tw.z, __fmaf_rn(b1, Tw.y, __fmul_rn(b0, Tw.x)));
	const float tx = __fmaf_rn(a2, Tu.z, __fmaf_rn(a1, Tu.y, __fmul_rn(a0, Tu.x)));
	const float ty = __fmaf_rn(b2, Tv.z, __fmaf_rn(b1, Tv.y, __fmul_rn(b0, Tv.x)));
	// point*point - temp is one fma
	const float hx = __fmaf_rn(cx, cx, -tx), hy = __fmaf_rn(cy, cy, -ty);
	centre = make_float2(cx, cy);
	extent = make_float2(__fsqrt_rn(fmaxf(hx, 1e-4f)), __fsqrt_rn(fmaxf(hy, 1e-4f)));
	return true;
}

// getRect (CR/auxiliary.h:66-76): int radius, /16 == * 0.0625 exactly, truncation toward zero
// before the clamp, "+ 16 - 1" as two fp32 additions.
__device__ __forceinline__ void tile_rect_exact(const float px, const float py, const int max_radius, const unsigned gx,
                                                const unsigned gy, uint2& rmin, uint2& rmax)
{
	const float r = (float)max_radius;
	rmin.x = min(gx, (unsigned)max(0, (int)__fmul_rn(__fsub_rn(px, r), 0.0625f)));
	rmin.y = min(gy, (unsigned)max(0, (int)__fmul_rn(__fsub_rn(py, r), 0.0625f)));
	rmax.x = min(gx, (unsigned)max(0, (int)__fmul_rn(__fadd_rn(__fadd_rn(__fadd_rn(px, r), 16.0f), -1.0f), 0.0625f)));
	rmax.y = min(gy, (unsigned)max(0, (int)__fmul_rn(__fadd_rn(__fadd_rn(__fadd_rn(py, r), 16.0f), -1.0f), 0.0625f)));
}

}  // namespace gs2d
