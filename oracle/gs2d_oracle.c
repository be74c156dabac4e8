/* gs2d_oracle.c -- CPU restatement of the reference 2DGS surfel rasterizer.
 *
 * TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline leg may load the library built from this file (oracle/build.py ->
 * oracle/_build/libgs2d_oracle.so).  The product (gaus-slam_b200/) never links or calls it and
 * has no CPU path at all.
 *
 * What it restates (file:line relative to /root/reference; RAST = submodules/gaus_2dgs_rasterization,
 * CR = RAST/cuda_rasterizer, KNN = submodules/simple-knn):
 *   oracle_preprocess          CR/forward.cu:150-253 (preprocessCUDA) with compute_transmat :75-115,
 *                              compute_aabb :119-147, in_frustum CR/auxiliary.h:184-209, getRect :66-76,
 *                              quat_to_rotmat :212-234, scale_to_mat :285-291, computeColorFromSH forward.cu:20-71
 *   oracle_scan / _duplicate / _sort / _ranges
 *                              CR/rasterizer_impl.cu:283 (InclusiveSum), :70-111 (duplicateWithKeys),
 *                              :309-314 (SortPairs == any stable sort on the low `bits` bits),
 *                              :116-138 + :316 (identifyTileRanges), :35-50 (getHigherMsb)
 *   oracle_render_forward      CR/forward.cu:258-467 (renderCUDA)
 *   oracle_render_backward     CR/backward.cu:143-463 (renderCUDA)
 *   oracle_preprocess_backward CR/backward.cu:466-664 (compute_transmat_aabb, preprocessCUDA),
 *                              quat_to_rotmat_vjp CR/auxiliary.h:237-281, SH backward backward.cu:20-139
 *   oracle_mark_visible        CR/rasterizer_impl.cu:54-66
 *   oracle_dist2               KNN/simple_knn.cu:46-222 (distCUDA2)
 *
 * Parity pin: the reference has no tests, golden vectors or fixtures for this path (SURVEY.md §4,
 * §8c).  This restatement is pinned against outputs of the reference itself: tests/golden/ (.npz files) were
 * produced by the UNMODIFIED reference CUDA code (oracle/_ref, built by oracle/build.py) on a B200 by
 * tests/golden/make_golden.py, and tests/test_oracle_golden.py checks every function here against them.
 *
 * Arithmetic: plain C fp32, compiled with -ffp-contract=off so that nothing is fused implicitly.
 * The per-surfel transform that decides radii / tile rectangles / depth keys spells out the fused
 * multiply-adds of the reference's compiled kernel with fmaf() (same sequence as
 * gaus-slam_b200/csrc/exact.cuh).  One operation cannot be reproduced on a CPU: the GPU's
 * approximate reciprocal square root (MUFU.RSQ) used to normalise the quaternion; 1/sqrtf is used
 * instead, which can move a transform entry by an ulp and, very rarely, a radius across an
 * integer boundary.  Per-pixel compositing is order-exact (front-to-back per pixel); gradient
 * sums differ from the GPU only by fp32 summation order.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define TILE 16
#define NEAR_N 0.2f
#define FAR_N 100.0f
#define FILTER_INV_SQUARE 100.0f

typedef struct { float x, y, z; } f3;

static inline float dot3_ref(float a0, float b0, float a1, float b1, float a2, float b2)
{
	return fmaf(a2, b2, fmaf(a0, b0, a1 * b1));
}

/* transformPoint4x3 (CR/auxiliary.h:78-86), m[4c + r] */
static inline f3 xform_point(f3 p, const float* m)
{
	f3 r = { m[12] + dot3_ref(p.x, m[0], p.y, m[4], p.z, m[8]),
	         m[13] + dot3_ref(p.x, m[1], p.y, m[5], p.z, m[9]),
	         m[14] + dot3_ref(p.x, m[2], p.y, m[6], p.z, m[10]) };
	return r;
}
/* transformVec4x3 (CR/auxiliary.h:100-108) */
static inline f3 xform_vec(f3 v, const float* m)
{
	f3 r = { dot3_ref(v.x, m[0], v.y, m[4], v.z, m[8]),
	         dot3_ref(v.x, m[1], v.y, m[5], v.z, m[9]),
	         dot3_ref(v.x, m[2], v.y, m[6], v.z, m[10]) };
	return r;
}
/* transformVec4x3Transpose (CR/auxiliary.h:110-118) */
static inline f3 xform_vec_T(f3 v, const float* m)
{
	f3 r = { m[0] * v.x + m[1] * v.y + m[2] * v.z,
	         m[4] * v.x + m[5] * v.y + m[6] * v.z,
	         m[8] * v.x + m[9] * v.y + m[10] * v.z };
	return r;
}

typedef struct {
	f3 Tu, Tv, Tw;
	f3 L0, L1, L2;
	f3 R0, R1, R2;
} frame_t;

/* quat_to_rotmat (CR/auxiliary.h:212-234), q = (w, x, y, z) as stored.  Column-major R. */
static inline void quat_to_rot(const float* q, f3* R0, f3* R1, f3* R2)
{
	const float n = fmaf(q[2], q[2], fmaf(q[1], q[1], fmaf(q[0], q[0], q[3] * q[3])));
	const float s = 1.0f / sqrtf(n);       /* GPU: rsqrtf (approximate) */
	const float w = q[0] * s, x = q[1] * s, y = q[2] * s, z = q[3] * s;
	const float yy = y * y, zz = z * z;
	const float wz = w * z, wy = w * y, wx = w * x;
	R0->x = 1.0f - 2.0f * (yy + zz);
	R0->y = 2.0f * fmaf(x, y, wz);
	R0->z = 2.0f * fmaf(x, z, -wy);
	R1->x = 2.0f * fmaf(x, y, -wz);
	R1->y = 1.0f - 2.0f * fmaf(x, x, zz);
	R1->z = 2.0f * fmaf(y, z, wx);
	R2->x = 2.0f * fmaf(x, z, wy);
	R2->y = 2.0f * fmaf(y, z, -wx);
	R2->z = 1.0f - 2.0f * fmaf(x, x, yy);
}

/* compute_transmat (CR/forward.cu:75-115): T = (splat2world)^T * world2ndc * ndc2pix. */
static inline void splat_frame(f3 p, const float* scale, float mod, const float* q, const float* pm,
                               int W, int H, frame_t* f)
{
	quat_to_rot(q, &f->R0, &f->R1, &f->R2);
	const float sx = mod * scale[0], sy = mod * scale[1];
	f->L0.x = sx * f->R0.x; f->L0.y = sx * f->R0.y; f->L0.z = sx * f->R0.z;
	f->L1.x = sy * f->R1.x; f->L1.y = sy * f->R1.y; f->L1.z = sy * f->R1.z;
	f->L2 = f->R2;
	float X[4][3];
	for (int c = 0; c < 4; c++) {
		X[c][0] = dot3_ref(f->L0.x, pm[c], f->L0.y, pm[4 + c], f->L0.z, pm[8 + c]);
		X[c][1] = dot3_ref(f->L1.x, pm[c], f->L1.y, pm[4 + c], f->L1.z, pm[8 + c]);
		X[c][2] = pm[12 + c] + dot3_ref(p.x, pm[c], p.y, pm[4 + c], p.z, pm[8 + c]);
	}
	const float Wh = (float)W * 0.5f, Wm = (float)(W - 1) * 0.5f;
	const float Hh = (float)H * 0.5f, Hm = (float)(H - 1) * 0.5f;
	f->Tu.x = fmaf(Wm, X[3][0], Wh * X[0][0]); f->Tu.y = fmaf(Wm, X[3][1], Wh * X[0][1]); f->Tu.z = fmaf(Wm, X[3][2], Wh * X[0][2]);
	f->Tv.x = fmaf(Hm, X[3][0], Hh * X[1][0]); f->Tv.y = fmaf(Hm, X[3][1], Hh * X[1][1]); f->Tv.z = fmaf(Hm, X[3][2], Hh * X[1][2]);
	f->Tw.x = X[3][0]; f->Tw.y = X[3][1]; f->Tw.z = X[3][2];
}

/* compute_aabb at cutoff 3 (CR/forward.cu:119-147) */
static inline int aabb(f3 Tu, f3 Tv, f3 Tw, float* cx_, float* cy_, float* ex, float* ey)
{
	const float xx = Tw.x * Tw.x, yy = Tw.y * Tw.y;
	const float distance = fmaf(-Tw.z, Tw.z, fmaf(xx, 9.0f, yy * 9.0f));
	if (distance == 0.0f) return 0;
	const float r = 1.0f / distance;
	const float nr = -r, f9 = r * 9.0f;
	const float a0 = f9 * Tu.x, a1 = f9 * Tu.y, a2 = Tu.z * nr;
	const float b0 = f9 * Tv.x, b1 = f9 * Tv.y, b2 = Tv.z * nr;
	const float cx = fmaf(a2, Tw.z, fmaf(a1, Tw.y, a0 * Tw.x));
	const float cy = fmaf(b2, Tw.z, fmaf(b1, Tw.y, b0 * Tw.x));
	const float tx = fmaf(a2, Tu.z, fmaf(a1, Tu.y, a0 * Tu.x));
	const float ty = fmaf(b2, Tv.z, fmaf(b1, Tv.y, b0 * Tv.x));
	const float hx = fmaf(cx, cx, -tx), hy = fmaf(cy, cy, -ty);
	*cx_ = cx; *cy_ = cy;
	*ex = sqrtf(fmaxf(hx, 1e-4f));
	*ey = sqrtf(fmaxf(hy, 1e-4f));
	return 1;
}

static inline int f2i_rz(float v)
{
	/* cvt.rzi.s32.f32: truncation, saturating, NaN -> 0 */
	if (!(v == v)) return 0;
	if (v >= 2147483648.0f) return 2147483647;
	if (v <= -2147483648.0f) return (-2147483647 - 1);
	return (int)v;
}
static inline unsigned umin_(unsigned a, unsigned b) { return a < b ? a : b; }
static inline int imax_(int a, int b) { return a > b ? a : b; }

/* getRect (CR/auxiliary.h:66-76) */
static inline void get_rect(float px, float py, int max_radius, unsigned gx, unsigned gy, unsigned* rmin, unsigned* rmax)
{
	const float r = (float)max_radius;
	rmin[0] = umin_(gx, (unsigned)imax_(0, f2i_rz((px - r) * 0.0625f)));
	rmin[1] = umin_(gy, (unsigned)imax_(0, f2i_rz((py - r) * 0.0625f)));
	rmax[0] = umin_(gx, (unsigned)imax_(0, f2i_rz((((px + r) + 16.0f) + -1.0f) * 0.0625f)));
	rmax[1] = umin_(gy, (unsigned)imax_(0, f2i_rz((((py + r) + 16.0f) + -1.0f) * 0.0625f)));
}

static const float SH_C0 = 0.28209479177387814f;
static const float SH_C1 = 0.4886025119029199f;
static const float SH_C2[5] = { 1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                                -1.0925484305920792f, 0.5462742152960396f };
static const float SH_C3[7] = { -0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f, 0.3731763325901154f,
                                -0.4570457994644658f, 1.445305721320277f, -0.5900435899266435f };

/* computeColorFromSH (CR/forward.cu:20-71) */
static void sh_to_rgb(int idx, int deg, int M, f3 pos, const float* campos, const float* shs, uint8_t* clamped, float* rgb)
{
	f3 dir = { pos.x - campos[0], pos.y - campos[1], pos.z - campos[2] };
	const float len = sqrtf(dir.x * dir.x + dir.y * dir.y + dir.z * dir.z);
	dir.x /= len; dir.y /= len; dir.z /= len;
	const float* sh = shs + (size_t)idx * M * 3;
	float res[3];
	for (int c = 0; c < 3; c++) res[c] = SH_C0 * sh[c];
	if (deg > 0) {
		const float x = dir.x, y = dir.y, z = dir.z;
		for (int c = 0; c < 3; c++)
			res[c] = res[c] - SH_C1 * y * sh[3 + c] + SH_C1 * z * sh[6 + c] - SH_C1 * x * sh[9 + c];
		if (deg > 1) {
			const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
			for (int c = 0; c < 3; c++)
				res[c] = res[c] + SH_C2[0] * xy * sh[12 + c] + SH_C2[1] * yz * sh[15 + c] +
				         SH_C2[2] * (2.0f * zz - xx - yy) * sh[18 + c] + SH_C2[3] * xz * sh[21 + c] +
				         SH_C2[4] * (xx - yy) * sh[24 + c];
			if (deg > 2) {
				for (int c = 0; c < 3; c++)
					res[c] = res[c] + SH_C3[0] * y * (3.0f * xx - yy) * sh[27 + c] + SH_C3[1] * xy * z * sh[30 + c] +
					         SH_C3[2] * y * (4.0f * zz - xx - yy) * sh[33 + c] +
					         SH_C3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * sh[36 + c] +
					         SH_C3[4] * x * (4.0f * zz - xx - yy) * sh[39 + c] + SH_C3[5] * z * (xx - yy) * sh[42 + c] +
					         SH_C3[6] * x * (xx - 3.0f * yy) * sh[45 + c];
			}
		}
	}
	for (int c = 0; c < 3; c++) {
		res[c] += 0.5f;
		clamped[3 * idx + c] = (res[c] < 0);
		rgb[c] = fmaxf(res[c], 0.0f);
	}
}

/* ------------------------------------------------------------------------------------------------
 * preprocessCUDA (CR/forward.cu:150-253).  Arrays are the reference's GeometryState members
 * (CR/rasterizer_impl.h:29-44).  Entries of culled surfels are left untouched except radii and
 * tiles_touched (zeroed), exactly like the kernel.  shs / colors_precomp: exactly one non-NULL;
 * (scales, rotations) / transMat_precomp: exactly one non-NULL.
 * ---------------------------------------------------------------------------------------------- */
void oracle_preprocess(int P, int D, int M, const float* means3D, const float* scales, float scale_modifier,
                       const float* rotations, const float* opacities, const float* shs, uint8_t* clamped,
                       const float* transMat_precomp, const float* colors_precomp, const float* viewmatrix,
                       const float* projmatrix, const float* campos, int W, int H,
                       int32_t* radii, float* means2D, float* depths, float* transMats, float* rgb,
                       float* normal_opacity, uint32_t* tiles_touched)
{
	const unsigned gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
#pragma omp parallel for schedule(static)
	for (int idx = 0; idx < P; idx++) {
		radii[idx] = 0;
		tiles_touched[idx] = 0;
		const f3 p = { means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2] };
		const f3 pv = xform_point(p, viewmatrix);
		if (pv.z <= 0.2f) continue;

		f3 Tu, Tv, Tw, normal;
		if (transMat_precomp == NULL) {
			frame_t f;
			splat_frame(p, scales + 2 * idx, scale_modifier, rotations + 4 * idx, projmatrix, W, H, &f);
			Tu = f.Tu; Tv = f.Tv; Tw = f.Tw;
			float* t = transMats + 9 * (size_t)idx;
			t[0] = Tu.x; t[1] = Tu.y; t[2] = Tu.z; t[3] = Tv.x; t[4] = Tv.y; t[5] = Tv.z; t[6] = Tw.x; t[7] = Tw.y; t[8] = Tw.z;
			normal = xform_vec(f.L2, viewmatrix);
		} else {
			const float* t = transMat_precomp + 9 * (size_t)idx;
			Tu.x = t[0]; Tu.y = t[1]; Tu.z = t[2]; Tv.x = t[3]; Tv.y = t[4]; Tv.z = t[5]; Tw.x = t[6]; Tw.y = t[7]; Tw.z = t[8];
			normal.x = 0.0f; normal.y = 0.0f; normal.z = 1.0f;
		}
		/* DUAL_VISIABLE (CR/forward.cu:211-216): cos = -sum(p_view * normal) */
		const float pdotn = dot3_ref(pv.x, normal.x, pv.y, normal.y, pv.z, normal.z);
		if (pdotn == 0.0f) continue;
		const float mult = pdotn < 0.0f ? 1.0f : -1.0f;
		normal.x *= mult; normal.y *= mult; normal.z *= mult;

		float cx, cy, ex, ey;
		if (!aabb(Tu, Tv, Tw, &cx, &cy, &ex, &ey)) continue;
		const float radius = ceilf(fmaxf(ex, ey));
		unsigned rmin[2], rmax[2];
		get_rect(cx, cy, f2i_rz(radius), gx, gy, rmin, rmax);
		if ((rmax[0] - rmin[0]) * (rmax[1] - rmin[1]) == 0) continue;

		if (colors_precomp == NULL) sh_to_rgb(idx, D, M, p, campos, shs, clamped, rgb + 3 * (size_t)idx);
		depths[idx] = pv.z;
		radii[idx] = f2i_rz(radius);
		means2D[2 * idx] = cx; means2D[2 * idx + 1] = cy;
		normal_opacity[4 * idx] = normal.x; normal_opacity[4 * idx + 1] = normal.y;
		normal_opacity[4 * idx + 2] = normal.z; normal_opacity[4 * idx + 3] = opacities[idx];
		tiles_touched[idx] = (rmax[1] - rmin[1]) * (rmax[0] - rmin[0]);
	}
}

/* checkFrustum (CR/rasterizer_impl.cu:54-66): present = !(z_view <= 0.2) */
void oracle_mark_visible(int P, const float* means3D, const float* viewmatrix, uint8_t* present)
{
	for (int idx = 0; idx < P; idx++) {
		const f3 p = { means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2] };
		present[idx] = xform_point(p, viewmatrix).z <= 0.2f ? 0 : 1;
	}
}

/* cub::DeviceScan::InclusiveSum (CR/rasterizer_impl.cu:283); returns num_rendered (:287). */
int64_t oracle_scan(int P, const uint32_t* tiles_touched, uint32_t* point_offsets)
{
	uint32_t run = 0;
	for (int i = 0; i < P; i++) { run += tiles_touched[i]; point_offsets[i] = run; }
	return P > 0 ? (int64_t)run : 0;
}

/* getHigherMsb (CR/rasterizer_impl.cu:35-50) */
uint32_t oracle_higher_msb(uint32_t n)
{
	uint32_t msb = sizeof(n) * 4, step = msb;
	while (step > 1) {
		step /= 2;
		if (n >> msb) msb += step; else msb -= step;
	}
	if (n >> msb) msb++;
	return msb;
}

/* duplicateWithKeys (CR/rasterizer_impl.cu:70-111) */
void oracle_duplicate(int P, int W, int H, const float* means2D, const float* depths, const uint32_t* point_offsets,
                      const int32_t* radii, uint64_t* keys_unsorted, uint32_t* values_unsorted)
{
	const unsigned gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
#pragma omp parallel for schedule(static)
	for (int idx = 0; idx < P; idx++) {
		if (radii[idx] <= 0) continue;
		uint32_t off = idx == 0 ? 0u : point_offsets[idx - 1];
		unsigned rmin[2], rmax[2];
		get_rect(means2D[2 * idx], means2D[2 * idx + 1], radii[idx], gx, gy, rmin, rmax);
		uint32_t dbits;
		memcpy(&dbits, depths + idx, 4);
		for (int y = (int)rmin[1]; y < (int)rmax[1]; y++)
			for (int x = (int)rmin[0]; x < (int)rmax[0]; x++) {
				uint64_t key = (uint64_t)((unsigned)y * gx + (unsigned)x);
				key <<= 32;
				key |= dbits;
				keys_unsorted[off] = key;
				values_unsorted[off] = (uint32_t)idx;
				off++;
			}
	}
}

/* cub::DeviceRadixSort::SortPairs on bits [0, end_bit) (CR/rasterizer_impl.cu:309-314): a stable
 * sort; implemented as an LSD radix sort with 16-bit digits (any stable sort gives the same result). */
void oracle_sort_pairs(int64_t n, const uint64_t* keys_in, const uint32_t* vals_in, uint64_t* keys_out,
                       uint32_t* vals_out, int end_bit)
{
	if (n <= 0) return;
	uint64_t* ka = (uint64_t*)malloc(sizeof(uint64_t) * (size_t)n);
	uint32_t* va = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)n);
	uint64_t* kb = (uint64_t*)malloc(sizeof(uint64_t) * (size_t)n);
	uint32_t* vb = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)n);
	size_t* cnt = (size_t*)malloc(sizeof(size_t) * 65537);
	memcpy(ka, keys_in, sizeof(uint64_t) * (size_t)n);
	memcpy(va, vals_in, sizeof(uint32_t) * (size_t)n);
	for (int shift = 0; shift < end_bit; shift += 16) {
		const int bits = end_bit - shift < 16 ? end_bit - shift : 16;
		const uint64_t mask = ((uint64_t)1 << bits) - 1;
		memset(cnt, 0, sizeof(size_t) * 65537);
		for (int64_t i = 0; i < n; i++) cnt[((ka[i] >> shift) & mask) + 1]++;
		for (int d = 0; d < 65536; d++) cnt[d + 1] += cnt[d];
		for (int64_t i = 0; i < n; i++) {
			const size_t pos = cnt[(ka[i] >> shift) & mask]++;
			kb[pos] = ka[i]; vb[pos] = va[i];
		}
		uint64_t* tk = ka; ka = kb; kb = tk;
		uint32_t* tv = va; va = vb; vb = tv;
	}
	memcpy(keys_out, ka, sizeof(uint64_t) * (size_t)n);
	memcpy(vals_out, va, sizeof(uint32_t) * (size_t)n);
	free(ka); free(va); free(kb); free(vb); free(cnt);
}

/* cudaMemset + identifyTileRanges (CR/rasterizer_impl.cu:116-138, :316) */
void oracle_tile_ranges(int64_t L, const uint64_t* keys_sorted, int W, int H, uint32_t* ranges)
{
	const size_t tiles = (size_t)((W + TILE - 1) / TILE) * ((H + TILE - 1) / TILE);
	memset(ranges, 0, sizeof(uint32_t) * 2 * tiles);
	for (int64_t idx = 0; idx < L; idx++) {
		const uint32_t curr = (uint32_t)(keys_sorted[idx] >> 32);
		if (idx == 0) ranges[2 * curr] = 0;
		else {
			const uint32_t prev = (uint32_t)(keys_sorted[idx - 1] >> 32);
			if (curr != prev) { ranges[2 * prev + 1] = (uint32_t)idx; ranges[2 * curr] = (uint32_t)idx; }
		}
		if (idx == L - 1) ranges[2 * curr + 1] = (uint32_t)L;
	}
}

/* ------------------------------------------------------------------------------------------------
 * renderCUDA forward (CR/forward.cu:258-467).  One (tile, pixel) at a time, instances front to back.
 * The CUDA block's collective fetch / exit vote does not influence any pixel's result, so a plain
 * loop is exact in operation order.  final_T = {T, M1, M2}[3,H,W]; n_contrib = {last, median}[2,H,W].
 * ---------------------------------------------------------------------------------------------- */
void oracle_render_forward(int W, int H, const uint32_t* ranges, const uint32_t* point_list, const float* means2D,
                           const float* features, const float* transMats, const float* normal_opacity,
                           const float* bg, int use_sa, float* final_T, uint32_t* n_contrib, float* out_color,
                           float* out_others, float* geo_median_depth, float* geo_depth_std)
{
	const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
	const size_t HW = (size_t)H * W;
#pragma omp parallel for schedule(dynamic, 4)
	for (int tile = 0; tile < gx * gy; tile++) {
		const int tx = tile % gx, ty = tile / gx;
		const uint32_t r0 = ranges[2 * tile], r1 = ranges[2 * tile + 1];
		for (int ly = 0; ly < TILE; ly++) for (int lx = 0; lx < TILE; lx++) {
			const unsigned pxi = tx * TILE + lx, pyi = ty * TILE + ly;
			if (pxi >= (unsigned)W || pyi >= (unsigned)H) continue;
			const size_t pix_id = (size_t)W * pyi + pxi;
			const float pixx = (float)pxi, pixy = (float)pyi;
			float T = 1.0f;
			uint32_t contributor = 0, last_contributor = 0;
			float C[3] = { 0, 0, 0 }, N[3] = { 0, 0, 0 };
			float D = 0, M1 = 0, M2 = 0, D2 = 0, distortion = 0, median_depth = 0;
			float median_contributor = -1;
			for (uint32_t q = r0; q < r1; q++) {
				contributor++;
				const uint32_t id = point_list[q];
				const float* t = transMats + 9 * (size_t)id;
				const f3 Tu = { t[0], t[1], t[2] }, Tv = { t[3], t[4], t[5] }, Tw = { t[6], t[7], t[8] };
				const float xyx = means2D[2 * id], xyy = means2D[2 * id + 1];
				const f3 k = { pixx * Tw.x - Tu.x, pixx * Tw.y - Tu.y, pixx * Tw.z - Tu.z };
				const f3 l = { pixy * Tw.x - Tv.x, pixy * Tw.y - Tv.y, pixy * Tw.z - Tv.z };
				const f3 p = { k.y * l.z - k.z * l.y, k.z * l.x - k.x * l.z, k.x * l.y - k.y * l.x };
				if (p.z == 0.0f) continue;
				const float sx = p.x / p.z, sy = p.y / p.z;
				/* fused multiply-adds exactly where the reference's compiled kernel has them (nvcc -fmad) */
				const float rho3d = fmaf(sx, sx, sy * sy);
				const float dx = xyx - pixx, dy = xyy - pixy;
				const float rho2d = FILTER_INV_SQUARE * fmaf(dy, dy, dx * dx);
				const float rho = fminf(rho3d, rho2d);
				float depth = (rho3d <= rho2d) ? fmaf(Tw.x, sx, Tw.y * sy) + Tw.z : Tw.z;
				if (depth < NEAR_N) continue;
				const float* no = normal_opacity + 4 * (size_t)id;
				const float opa = no[3];
				const float power = -0.5f * rho;
				if (power > 0.0f) continue;
				const float alpha = fminf(0.99f, opa * expf(power));
				if (alpha < 1.0f / 255.0f) continue;
				const float test_T = T * (1 - alpha);
				if (test_T < 0.0001f) break;   /* done = true: this pixel stops (CR/forward.cu:388-393) */
				const float w = alpha * T;
				if (T > 0.5) { median_depth = depth; median_contributor = (float)contributor; }
				if (use_sa) {
					if (D > 0) {
						const float exp_depth = median_depth;
						float exp_std = fmaf(exp_depth, exp_depth, (D2 - (D + D) * exp_depth) / (1 - T));
						exp_std = fmaxf(exp_std, 1e-7f);
						const float error = (exp_depth - depth) * (exp_depth - depth);
						const float conf = expf(error / (exp_std * -4.0f));
						depth = fmaf(depth, conf, exp_depth * (1 - conf));
					}
					D = fmaf(w, depth, D);
					D2 = fmaf(w, depth * depth, D2);
				} else {
					const float A = 1 - T;
					const float m = (FAR_N / (FAR_N - NEAR_N)) * (1 - NEAR_N / depth);
					const float mm = m * m;
					distortion = fmaf(fmaf(A, mm, M2) - M1 * (m + m), w, distortion);
					D = fmaf(depth, w, D);
					M1 = fmaf(m, w, M1);
					M2 = fmaf(mm, w, M2);
				}
				for (int ch = 0; ch < 3; ch++) N[ch] = fmaf(w, no[ch], N[ch]);
				for (int ch = 0; ch < 3; ch++) C[ch] = fmaf(w, features[3 * (size_t)id + ch], C[ch]);
				T = test_T;
				last_contributor = contributor;
			}
			final_T[pix_id] = T;
			n_contrib[pix_id] = last_contributor;
			for (int ch = 0; ch < 3; ch++) out_color[ch * HW + pix_id] = fmaf(T, bg[ch], C[ch]);
			/* float -> uint32 store (CR/forward.cu:449): cvt.rzi.u32.f32 saturates -1 to 0 */
			n_contrib[pix_id + HW] = median_contributor < 0 ? 0u : (uint32_t)median_contributor;
			final_T[pix_id + HW] = M1;
			final_T[pix_id + 2 * HW] = M2;
			out_others[pix_id + 0 * HW] = D;
			out_others[pix_id + 1 * HW] = 1 - T;
			for (int ch = 0; ch < 3; ch++) out_others[pix_id + (2 + ch) * HW] = N[ch];
			out_others[pix_id + 5 * HW] = median_depth;
			geo_median_depth[pix_id] = median_depth;
			const float var0 = D2 - (median_depth + median_depth) * D;
			geo_depth_std[pix_id] = fmaf(median_depth * median_depth, 1 - T, var0);
			if (use_sa) out_others[pix_id + 6 * HW] = fmaf(median_depth, median_depth * (1 - T), var0);
			else out_others[pix_id + 6 * HW] = distortion;
		}
	}
}

static inline void atomic_addf(float* p, float v)
{
#pragma omp atomic
	*p += v;
}

/* ------------------------------------------------------------------------------------------------
 * renderCUDA backward (CR/backward.cu:143-463).  Gradient arrays must be zero-initialised by the
 * caller (the reference's torch::zeros, RAST/rasterize_points.cu:192-200).
 * ---------------------------------------------------------------------------------------------- */
void oracle_render_backward(int W, int H, const uint32_t* ranges, const uint32_t* point_list, const float* bg,
                            const float* means2D, const float* normal_opacity, const float* transMats,
                            const float* colors, const float* final_Ts, const uint32_t* n_contrib,
                            const float* dL_dpixels, const float* dL_depths, const float* median_depth,
                            const float* depth_std, int use_sa, float* dL_dtransMat, float* dL_dmean2D /*[P,3]*/,
                            float* dL_dnormal3D, float* dL_dopacity, float* dL_dcolors)
{
	const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
	const size_t HW = (size_t)H * W;
#pragma omp parallel for schedule(dynamic, 4)
	for (int tile = 0; tile < gx * gy; tile++) {
		const int tx = tile % gx, ty = tile / gx;
		const uint32_t r0 = ranges[2 * tile], r1 = ranges[2 * tile + 1];
		for (int ly = 0; ly < TILE; ly++) for (int lx = 0; lx < TILE; lx++) {
			const unsigned pxi = tx * TILE + lx, pyi = ty * TILE + ly;
			if (pxi >= (unsigned)W || pyi >= (unsigned)H) continue;
			const size_t pix_id = (size_t)W * pyi + pxi;
			const float pixx = (float)pxi, pixy = (float)pyi;
			const float T_final = final_Ts[pix_id];
			float T = T_final;
			uint32_t contributor = r1 - r0;
			const int last_contributor = (int)n_contrib[pix_id];
			float accum_rec[3] = { 0, 0, 0 }, dL_dpixel[3], last_color[3] = { 0, 0, 0 };
			const float dL_ddepth = dL_depths[0 * HW + pix_id];
			const float dL_daccum = dL_depths[1 * HW + pix_id];
			const float dL_dreg = dL_depths[6 * HW + pix_id];
			float dL_dnormal2D[3];
			for (int i = 0; i < 3; i++) dL_dnormal2D[i] = dL_depths[(2 + i) * HW + pix_id];
			const int median_contributor = (int)n_contrib[pix_id + HW];
			const float dL_dmedian_depth = dL_depths[5 * HW + pix_id];
			const float mm = median_depth[pix_id];
			const float mstd = depth_std[pix_id];
			float last_depth = 0, last_normal[3] = { 0, 0, 0 };
			float accum_depth_rec = 0, accum_alpha_rec = 0, accum_normal_rec[3] = { 0, 0, 0 };
			const float final_D = final_Ts[pix_id + HW];
			const float final_D2 = final_Ts[pix_id + 2 * HW];
			const float final_A = 1 - T_final;
			float last_dL_dT = 0;
			for (int i = 0; i < 3; i++) dL_dpixel[i] = dL_dpixels[i * HW + pix_id];
			float last_alpha = 0;

			for (uint32_t q = r1; q-- > r0;) {
				contributor--;
				if (contributor >= (uint32_t)last_contributor) continue;
				const uint32_t id = point_list[q];
				const float* t = transMats + 9 * (size_t)id;
				const f3 Tu = { t[0], t[1], t[2] }, Tv = { t[3], t[4], t[5] }, Tw = { t[6], t[7], t[8] };
				const float xyx = means2D[2 * id], xyy = means2D[2 * id + 1];
				const f3 k = { pixx * Tw.x - Tu.x, pixx * Tw.y - Tu.y, pixx * Tw.z - Tu.z };
				const f3 l = { pixy * Tw.x - Tv.x, pixy * Tw.y - Tv.y, pixy * Tw.z - Tv.z };
				const f3 p = { k.y * l.z - k.z * l.y, k.z * l.x - k.x * l.z, k.x * l.y - k.y * l.x };
				if (p.z == 0.0f) continue;
				const float sx = p.x / p.z, sy = p.y / p.z;
				const float rho3d = (sx * sx + sy * sy);
				const float dx = xyx - pixx, dy = xyy - pixy;
				const float rho2d = FILTER_INV_SQUARE * (dx * dx + dy * dy);
				const float rho = fminf(rho3d, rho2d);
				float c_d = (rho3d <= rho2d) ? (sx * Tw.x + sy * Tw.y) + Tw.z : Tw.z;
				if (c_d < NEAR_N) continue;
				const float* no = normal_opacity + 4 * (size_t)id;
				const float opa = no[3];
				const float power = -0.5f * rho;
				if (power > 0.0f) continue;
				const float G = expf(power);
				const float alpha = fminf(0.99f, opa * G);
				if (alpha < 1.0f / 255.0f) continue;

				T = T / (1.f - alpha);
				const float dchannel_dcolor = alpha * T;
				const float w = alpha * T;
				float dL_dalpha = 0.0f;
				for (int ch = 0; ch < 3; ch++) {
					const float c = colors[3 * (size_t)id + ch];
					accum_rec[ch] = last_alpha * last_color[ch] + (1.f - last_alpha) * accum_rec[ch];
					last_color[ch] = c;
					const float dL_dchannel = dL_dpixel[ch];
					dL_dalpha += (c - accum_rec[ch]) * dL_dchannel;
					atomic_addf(&dL_dcolors[3 * (size_t)id + ch], dchannel_dcolor * dL_dchannel);
				}
				float conf = 1;
				if (use_sa) {
					/* double-precision expression in the reference (1e-7 and exp() promote) */
					conf = T < 0.5
					     ? (float)exp((double)(-(c_d - mm) * (c_d - mm)) / (4 * fmax((double)(mstd / (1 - T_final)), 1e-7)))
					     : 1.f;
					c_d = c_d * conf + mm * (1 - conf);
				}
				float dL_dz = 0.0f, dL_dweight = 0;
				const float m_d = FAR_N / (FAR_N - NEAR_N) * (1 - NEAR_N / c_d);
				const float dmd_dd = (FAR_N * NEAR_N) / ((FAR_N - NEAR_N) * c_d * c_d);
				if (contributor == (uint32_t)(median_contributor - 1)) dL_dz += dL_dmedian_depth;
				if (use_sa) dL_dweight += ((c_d - mm) * (c_d - mm)) * dL_dreg;
				else dL_dweight += (final_D2 + m_d * m_d * final_A - 2 * m_d * final_D) * dL_dreg;
				dL_dalpha += dL_dweight - last_dL_dT;
				last_dL_dT = dL_dweight * alpha + (1 - alpha) * last_dL_dT;
				if (use_sa) dL_dz += conf * 2.0f * w * (c_d - mm) * dL_dreg;
				else {
					const float dL_dmd = 2.0f * (T * alpha) * (m_d * final_A - final_D) * dL_dreg;
					dL_dz += dL_dmd * dmd_dd;
				}
				accum_depth_rec = last_alpha * last_depth + (1.f - last_alpha) * accum_depth_rec;
				last_depth = c_d;
				dL_dalpha += (c_d - accum_depth_rec) * dL_ddepth;
				accum_alpha_rec = (float)(last_alpha * 1.0 + (double)((1.f - last_alpha) * accum_alpha_rec));
				dL_dalpha += (1 - accum_alpha_rec) * dL_daccum;
				for (int ch = 0; ch < 3; ch++) {
					accum_normal_rec[ch] = last_alpha * last_normal[ch] + (1.f - last_alpha) * accum_normal_rec[ch];
					last_normal[ch] = no[ch];
					dL_dalpha += (no[ch] - accum_normal_rec[ch]) * dL_dnormal2D[ch];
					atomic_addf(&dL_dnormal3D[3 * (size_t)id + ch], alpha * T * dL_dnormal2D[ch]);
				}
				dL_dalpha *= T;
				last_alpha = alpha;
				float bg_dot_dpixel = 0;
				for (int i = 0; i < 3; i++) bg_dot_dpixel += bg[i] * dL_dpixel[i];
				dL_dalpha += (-T_final / (1.f - alpha)) * bg_dot_dpixel;
				const float dL_dG = opa * dL_dalpha;
				dL_dz += conf * alpha * T * dL_ddepth;

				float* dT = dL_dtransMat + 9 * (size_t)id;
				if (rho3d <= rho2d) {
					const float dL_dsx = dL_dG * -G * sx + dL_dz * Tw.x;
					const float dL_dsy = dL_dG * -G * sy + dL_dz * Tw.y;
					const float dsx_pz = dL_dsx / p.z, dsy_pz = dL_dsy / p.z;
					const f3 dL_dp = { dsx_pz, dsy_pz, -(dsx_pz * sx + dsy_pz * sy) };
					const f3 dL_dk = { l.y * dL_dp.z - l.z * dL_dp.y, l.z * dL_dp.x - l.x * dL_dp.z, l.x * dL_dp.y - l.y * dL_dp.x };
					const f3 dL_dl = { dL_dp.y * k.z - dL_dp.z * k.y, dL_dp.z * k.x - dL_dp.x * k.z, dL_dp.x * k.y - dL_dp.y * k.x };
					atomic_addf(dT + 0, -dL_dk.x); atomic_addf(dT + 1, -dL_dk.y); atomic_addf(dT + 2, -dL_dk.z);
					atomic_addf(dT + 3, -dL_dl.x); atomic_addf(dT + 4, -dL_dl.y); atomic_addf(dT + 5, -dL_dl.z);
					atomic_addf(dT + 6, pixx * dL_dk.x + pixy * dL_dl.x + dL_dz * sx);
					atomic_addf(dT + 7, pixx * dL_dk.y + pixy * dL_dl.y + dL_dz * sy);
					atomic_addf(dT + 8, pixx * dL_dk.z + pixy * dL_dl.z + dL_dz * 1.0f);
				} else {
					const float dG_ddelx = -G * FILTER_INV_SQUARE * dx;
					const float dG_ddely = -G * FILTER_INV_SQUARE * dy;
					atomic_addf(&dL_dmean2D[3 * (size_t)id + 0], dL_dG * dG_ddelx);
					atomic_addf(&dL_dmean2D[3 * (size_t)id + 1], dL_dG * dG_ddely);
					atomic_addf(dT + 8, dL_dz);
				}
				atomic_addf(&dL_dopacity[id], G * dL_dalpha);
			}
		}
	}
}

/* dnormvdv (CR/auxiliary.h:130-140) */
static inline f3 dnormvdv(f3 v, f3 dv)
{
	const float sum2 = v.x * v.x + v.y * v.y + v.z * v.z;
	const float inv = 1.0f / sqrtf(sum2 * sum2 * sum2);
	f3 o;
	o.x = ((+sum2 - v.x * v.x) * dv.x - v.y * v.x * dv.y - v.z * v.x * dv.z) * inv;
	o.y = (-v.x * v.y * dv.x + (sum2 - v.y * v.y) * dv.y - v.z * v.y * dv.z) * inv;
	o.z = (-v.x * v.z * dv.x - v.y * v.z * dv.y + (sum2 - v.z * v.z) * dv.z) * inv;
	return o;
}

/* computeColorFromSH backward (CR/backward.cu:20-139): writes dL_dsh[0 .. (deg+1)^2) and ADDS the
 * view-direction part to dL_dmeans[idx]. */
static void sh_backward(int idx, int deg, int M, const float* means, const float* campos, const float* shs,
                        const uint8_t* clamped, const float* dL_dcolor, float* dL_dmeans, float* dL_dshs)
{
	const f3 dir_orig = { means[3 * idx] - campos[0], means[3 * idx + 1] - campos[1], means[3 * idx + 2] - campos[2] };
	const float len = sqrtf(dir_orig.x * dir_orig.x + dir_orig.y * dir_orig.y + dir_orig.z * dir_orig.z);
	const float x = dir_orig.x / len, y = dir_orig.y / len, z = dir_orig.z / len;
	const float* sh = shs + (size_t)idx * M * 3;
	float* out = dL_dshs + (size_t)idx * M * 3;
	float g[3];
	for (int c = 0; c < 3; c++) g[c] = dL_dcolor[3 * (size_t)idx + c] * (clamped[3 * idx + c] ? 0.0f : 1.0f);
	float dx[3] = { 0, 0, 0 }, dy[3] = { 0, 0, 0 }, dz[3] = { 0, 0, 0 };
#define SH(k, c) sh[3 * (k) + (c)]
#define OUT(k, val) do { for (int c = 0; c < 3; c++) out[3 * (k) + c] = (val) * g[c]; } while (0)
	OUT(0, SH_C0);
	if (deg > 0) {
		OUT(1, -SH_C1 * y); OUT(2, SH_C1 * z); OUT(3, -SH_C1 * x);
		for (int c = 0; c < 3; c++) { dx[c] = -SH_C1 * SH(3, c); dy[c] = -SH_C1 * SH(1, c); dz[c] = SH_C1 * SH(2, c); }
		if (deg > 1) {
			const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
			OUT(4, SH_C2[0] * xy); OUT(5, SH_C2[1] * yz); OUT(6, SH_C2[2] * (2.f * zz - xx - yy));
			OUT(7, SH_C2[3] * xz); OUT(8, SH_C2[4] * (xx - yy));
			for (int c = 0; c < 3; c++) {
				dx[c] += SH_C2[0] * y * SH(4, c) + SH_C2[2] * 2.f * -x * SH(6, c) + SH_C2[3] * z * SH(7, c) + SH_C2[4] * 2.f * x * SH(8, c);
				dy[c] += SH_C2[0] * x * SH(4, c) + SH_C2[1] * z * SH(5, c) + SH_C2[2] * 2.f * -y * SH(6, c) + SH_C2[4] * 2.f * -y * SH(8, c);
				dz[c] += SH_C2[1] * y * SH(5, c) + SH_C2[2] * 2.f * 2.f * z * SH(6, c) + SH_C2[3] * x * SH(7, c);
			}
			if (deg > 2) {
				OUT(9, SH_C3[0] * y * (3.f * xx - yy)); OUT(10, SH_C3[1] * xy * z);
				OUT(11, SH_C3[2] * y * (4.f * zz - xx - yy)); OUT(12, SH_C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy));
				OUT(13, SH_C3[4] * x * (4.f * zz - xx - yy)); OUT(14, SH_C3[5] * z * (xx - yy));
				OUT(15, SH_C3[6] * x * (xx - 3.f * yy));
				for (int c = 0; c < 3; c++) {
					dx[c] += (SH_C3[0] * SH(9, c) * 3.f * 2.f * xy + SH_C3[1] * SH(10, c) * yz + SH_C3[2] * SH(11, c) * -2.f * xy +
					          SH_C3[3] * SH(12, c) * -3.f * 2.f * xz + SH_C3[4] * SH(13, c) * (-3.f * xx + 4.f * zz - yy) +
					          SH_C3[5] * SH(14, c) * 2.f * xz + SH_C3[6] * SH(15, c) * 3.f * (xx - yy));
					dy[c] += (SH_C3[0] * SH(9, c) * 3.f * (xx - yy) + SH_C3[1] * SH(10, c) * xz +
					          SH_C3[2] * SH(11, c) * (-3.f * yy + 4.f * zz - xx) + SH_C3[3] * SH(12, c) * -3.f * 2.f * yz +
					          SH_C3[4] * SH(13, c) * -2.f * xy + SH_C3[5] * SH(14, c) * -2.f * yz + SH_C3[6] * SH(15, c) * -3.f * 2.f * xy);
					dz[c] += (SH_C3[1] * SH(10, c) * xy + SH_C3[2] * SH(11, c) * 4.f * 2.f * yz +
					          SH_C3[3] * SH(12, c) * 3.f * (2.f * zz - xx - yy) + SH_C3[4] * SH(13, c) * 4.f * 2.f * xz +
					          SH_C3[5] * SH(14, c) * (xx - yy));
				}
			}
		}
	}
#undef SH
#undef OUT
	const f3 dL_ddir = { dx[0] * g[0] + dx[1] * g[1] + dx[2] * g[2], dy[0] * g[0] + dy[1] * g[1] + dy[2] * g[2],
	                     dz[0] * g[0] + dz[1] * g[1] + dz[2] * g[2] };
	const f3 dm = dnormvdv(dir_orig, dL_ddir);
	dL_dmeans[3 * idx] += dm.x; dL_dmeans[3 * idx + 1] += dm.y; dL_dmeans[3 * idx + 2] += dm.z;
}

/* ------------------------------------------------------------------------------------------------
 * BACKWARD::preprocess (CR/backward.cu:609-664) with compute_transmat_aabb (:466-607).
 * In/out like the reference: dL_dtransMats, dL_dmean2Ds hold the sums from the render backward on
 * entry; dL_dmean2Ds is OVERWRITTEN with the densification proxy (:660-663); dL_dmean3Ds,
 * dL_dscales, dL_drots must be zero-initialised by the caller.  transMats = transMat_precomp when
 * scales == NULL, else the forward's transMat array.
 * ---------------------------------------------------------------------------------------------- */
void oracle_preprocess_backward(int P, int D, int M, const float* means3D, const float* transMats,
                                const int32_t* radii, const float* shs, const uint8_t* clamped,
                                const float* scales, const float* rotations, const float* viewmatrix,
                                const float* projmatrix, float focal_x, float focal_y, float tan_fovx, float tan_fovy,
                                const float* campos, float* dL_dtransMats, const float* dL_dnormal3Ds,
                                float* dL_dcolors, float* dL_dshs, float* dL_dmean2Ds /*[P,3]*/, float* dL_dmean3Ds,
                                float* dL_dscales, float* dL_drots)
{
	const int W = (int)(focal_x * tan_fovx * 2);
	const int H = (int)(focal_y * tan_fovy * 2);
#pragma omp parallel for schedule(static)
	for (int idx = 0; idx < P; idx++) {
		if (!(radii[idx] > 0)) continue;
		const int precomp = scales == NULL;
		float T[3][3];      /* T[c] = column c = (Tu | Tv | Tw) */
		float Pm[3][4];     /* world2ndc * ndc2pix, 3 columns of 4 rows */
		f3 R0 = {0,0,0}, R1 = {0,0,0}, R2 = {0,0,0}, normal = {0,0,0}, p = {0,0,0};
		float sc[2] = {0, 0};
		const float* q = NULL;
		if (precomp) {
			for (int c = 0; c < 3; c++) for (int r = 0; r < 3; r++) T[c][r] = transMats[9 * (size_t)idx + 3 * c + r];
		} else {
			p.x = means3D[3 * idx]; p.y = means3D[3 * idx + 1]; p.z = means3D[3 * idx + 2];
			q = rotations + 4 * idx;
			sc[0] = scales[2 * idx]; sc[1] = scales[2 * idx + 1];
			quat_to_rot(q, &R0, &R1, &R2);
			const f3 L0 = { R0.x * sc[0], R0.y * sc[0], R0.z * sc[0] }, L1 = { R1.x * sc[1], R1.y * sc[1], R1.z * sc[1] };
			const float Mm[3][4] = { { L0.x, L0.y, L0.z, 0.f }, { L1.x, L1.y, L1.z, 0.f }, { p.x, p.y, p.z, 1.f } };
			const float n2p[3][4] = { { (float)W / 2.0f, 0.f, 0.f, (float)(W - 1) / 2.0f },
			                          { 0.f, (float)H / 2.0f, 0.f, (float)(H - 1) / 2.0f }, { 0.f, 0.f, 0.f, 1.f } };
			/* P[c][r] = sum_k w2n[k][r] * n2p[c][k], w2n[k][r] = projmatrix[k + 4r] (CR/backward.cu:513-527) */
			for (int c = 0; c < 3; c++) for (int r = 0; r < 4; r++) {
				float acc = 0.f;
				for (int k = 0; k < 4; k++) acc += projmatrix[k + 4 * r] * n2p[c][k];
				Pm[c][r] = acc;
			}
			for (int c = 0; c < 3; c++) for (int r = 0; r < 3; r++) {
				float acc = 0.f;
				for (int k = 0; k < 4; k++) acc += Mm[r][k] * Pm[c][k];
				T[c][r] = acc;
			}
			normal = xform_vec(R2, viewmatrix);
		}
		float dT[3][3];
		for (int c = 0; c < 3; c++) for (int r = 0; r < 3; r++) dT[c][r] = dL_dtransMats[9 * (size_t)idx + 3 * c + r];
		const float gmx = dL_dmean2Ds[3 * (size_t)idx], gmy = dL_dmean2Ds[3 * (size_t)idx + 1];
		if (gmx != 0 || gmy != 0) {
			const float* Tu = T[0]; const float* Tv = T[1]; const float* Tw = T[2];
			const float distance = Tw[0] * Tw[0] + Tw[1] * Tw[1] - Tw[2] * Tw[2];
			const float f = 1 / (distance);
			dT[0][0] += gmx * (f * Tw[0]); dT[0][1] += gmx * (f * Tw[1]); dT[0][2] += gmx * (-f * Tw[2]);
			dT[1][0] += gmy * (f * Tw[0]); dT[1][1] += gmy * (f * Tw[1]); dT[1][2] += gmy * (-f * Tw[2]);
			dT[2][0] += gmx * (Tu[0] * (f - 2 * f * f * Tw[0] * Tw[0])) + gmy * (Tv[0] * (f - 2 * f * f * Tw[0] * Tw[0]));
			dT[2][1] += gmx * (Tu[1] * (f - 2 * f * f * Tw[1] * Tw[1])) + gmy * (Tv[1] * (f - 2 * f * f * Tw[1] * Tw[1]));
			dT[2][2] += gmx * (-Tu[2] * (f + 2 * f * f * Tw[2] * Tw[2])) + gmy * (-Tv[2] * (f + 2 * f * f * Tw[2] * Tw[2]));
			if (precomp)
				for (int c = 0; c < 3; c++) for (int r = 0; r < 3; r++) dL_dtransMats[9 * (size_t)idx + 3 * c + r] = dT[c][r];
		}
		if (!precomp) {
			/* dL_dM[c][r] = sum_k P[k][r] * dT[k][c]  (P * transpose(dL_dT)) */
			float dM[3][4];
			for (int c = 0; c < 3; c++) for (int r = 0; r < 4; r++)
				dM[c][r] = Pm[0][r] * dT[0][c] + Pm[1][r] * dT[1][c] + Pm[2][r] * dT[2][c];
			const f3 dn = { dL_dnormal3Ds[3 * (size_t)idx], dL_dnormal3Ds[3 * (size_t)idx + 1], dL_dnormal3Ds[3 * (size_t)idx + 2] };
			f3 dtn = xform_vec_T(dn, viewmatrix);
			const f3 pv = xform_point(p, viewmatrix);
			const float cosv = -(pv.x * normal.x + pv.y * normal.y + pv.z * normal.z);
			const float mult = cosv > 0 ? 1.f : -1.f;
			dtn.x *= mult; dtn.y *= mult; dtn.z *= mult;
			const float dRS[3][3] = { { dM[0][0], dM[0][1], dM[0][2] }, { dM[1][0], dM[1][1], dM[1][2] }, { dtn.x, dtn.y, dtn.z } };
			float vR[3][3];
			for (int r = 0; r < 3; r++) { vR[0][r] = dRS[0][r] * sc[0]; vR[1][r] = dRS[1][r] * sc[1]; vR[2][r] = dRS[2][r]; }
			/* quat_to_rotmat_vjp (CR/auxiliary.h:237-281) */
			const float s = 1.0f / sqrtf(q[3] * q[3] + q[0] * q[0] + q[1] * q[1] + q[2] * q[2]);
			const float w = q[0] * s, x = q[1] * s, y = q[2] * s, z = q[3] * s;
			float* dq = dL_drots + 4 * (size_t)idx;
			dq[0] = 2.f * (x * (vR[1][2] - vR[2][1]) + y * (vR[2][0] - vR[0][2]) + z * (vR[0][1] - vR[1][0]));
			dq[1] = 2.f * (-2.f * x * (vR[1][1] + vR[2][2]) + y * (vR[0][1] + vR[1][0]) + z * (vR[0][2] + vR[2][0]) + w * (vR[1][2] - vR[2][1]));
			dq[2] = 2.f * (x * (vR[0][1] + vR[1][0]) - 2.f * y * (vR[0][0] + vR[2][2]) + z * (vR[1][2] + vR[2][1]) + w * (vR[2][0] - vR[0][2]));
			dq[3] = 2.f * (x * (vR[0][2] + vR[2][0]) + y * (vR[1][2] + vR[2][1]) - 2.f * z * (vR[0][0] + vR[1][1]) + w * (vR[0][1] - vR[1][0]));
			dL_dscales[2 * (size_t)idx] = dRS[0][0] * R0.x + dRS[0][1] * R0.y + dRS[0][2] * R0.z;
			dL_dscales[2 * (size_t)idx + 1] = dRS[1][0] * R1.x + dRS[1][1] * R1.y + dRS[1][2] * R1.z;
			dL_dmean3Ds[3 * (size_t)idx] = dM[2][0]; dL_dmean3Ds[3 * (size_t)idx + 1] = dM[2][1]; dL_dmean3Ds[3 * (size_t)idx + 2] = dM[2][2];
		}
		if (shs) sh_backward(idx, D, M, means3D, campos, shs, clamped, dL_dcolors, dL_dmean3Ds, dL_dshs);
		/* densification proxy (CR/backward.cu:660-663), double arithmetic like the reference */
		const float depth = transMats[9 * (size_t)idx + 8];
		dL_dmean2Ds[3 * (size_t)idx + 0] = (float)((double)(dL_dtransMats[9 * (size_t)idx + 2] * depth) * 0.5 * (double)(float)W);
		dL_dmean2Ds[3 * (size_t)idx + 1] = (float)((double)(dL_dtransMats[9 * (size_t)idx + 5] * depth) * 0.5 * (double)(float)H);
	}
}

/* ------------------------------------------------------------------------------------------------
 * distCUDA2 / SimpleKNN::knn (KNN/simple_knn.cu:46-222): mean squared distance to the 3 nearest
 * other points.  Same structure as the reference: bounding box (with the reference's quirk that the
 * reduction starts from (0,0,0), :191-200), 30-bit Morton codes, stable sort, boxes of 1024 points,
 * box-pruned exhaustive search.  The pruning never discards a true neighbour, so the result is the
 * exact 3-NN; the per-pair distance is fma(dz,dz, fma(dy,dy, dx*dx)) as compiled in the reference.
 * ---------------------------------------------------------------------------------------------- */
#define BOX_SIZE 1024
static inline uint32_t prep_morton(uint32_t x)
{
	x = (x | (x << 16)) & 0x030000FF;
	x = (x | (x << 8)) & 0x0300F00F;
	x = (x | (x << 4)) & 0x030C30C3;
	x = (x | (x << 2)) & 0x09249249;
	return x;
}
static inline uint32_t f2u_rz(float v)
{
	if (!(v == v) || v <= 0.0f) return 0u;
	if (v >= 4294967296.0f) return 0xffffffffu;
	return (uint32_t)v;
}
static inline void knn_update3(const float* ref, const float* pt, float* best)
{
	const float dx = pt[0] - ref[0], dy = pt[1] - ref[1], dz = pt[2] - ref[2];
	float dist = fmaf(dz, dz, fmaf(dx, dx, dy * dy));   /* the compiled order: round(dy*dy) first */
	for (int j = 0; j < 3; j++)
		if (best[j] > dist) { float t = best[j]; best[j] = dist; dist = t; }
}
static inline float box_dist(const float* mn, const float* mx, const float* p)
{
	float d[3] = { 0, 0, 0 };
	for (int a = 0; a < 3; a++)
		if (p[a] < mn[a] || p[a] > mx[a]) d[a] = fminf(fabsf(p[a] - mn[a]), fabsf(p[a] - mx[a]));
	return fmaf(d[2], d[2], fmaf(d[0], d[0], d[1] * d[1]));
}
#define FLT_MAX_ 3.402823466e+38F

void oracle_dist2(int P, const float* points, float* mean_dists)
{
	if (P <= 0) return;
	float mn[3] = { 0, 0, 0 }, mx[3] = { 0, 0, 0 };
	for (int i = 0; i < P; i++) for (int a = 0; a < 3; a++) {
		mn[a] = fminf(mn[a], points[3 * i + a]);
		mx[a] = fmaxf(mx[a], points[3 * i + a]);
	}
	uint64_t* keys = (uint64_t*)malloc(sizeof(uint64_t) * (size_t)P);
	uint32_t* idx = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)P);
	uint64_t* keys_s = (uint64_t*)malloc(sizeof(uint64_t) * (size_t)P);
	uint32_t* order = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)P);
	for (int i = 0; i < P; i++) {
		uint32_t c[3];
		for (int a = 0; a < 3; a++)
			c[a] = prep_morton(f2u_rz(((points[3 * i + a] - mn[a]) / (mx[a] - mn[a])) * (float)((1 << 10) - 1)));
		keys[i] = c[0] | (c[1] << 1) | (c[2] << 2);
		idx[i] = (uint32_t)i;
	}
	oracle_sort_pairs(P, keys, idx, keys_s, order, 32);
	const int nb = (P + BOX_SIZE - 1) / BOX_SIZE;
	float* bmn = (float*)malloc(sizeof(float) * 3 * (size_t)nb);
	float* bmx = (float*)malloc(sizeof(float) * 3 * (size_t)nb);
	for (int b = 0; b < nb; b++) {
		for (int a = 0; a < 3; a++) { bmn[3 * b + a] = FLT_MAX_; bmx[3 * b + a] = -FLT_MAX_; }
		const int end = (b + 1) * BOX_SIZE < P ? (b + 1) * BOX_SIZE : P;
		for (int i = b * BOX_SIZE; i < end; i++) for (int a = 0; a < 3; a++) {
			const float v = points[3 * (size_t)order[i] + a];
			bmn[3 * b + a] = fminf(bmn[3 * b + a], v);
			bmx[3 * b + a] = fmaxf(bmx[3 * b + a], v);
		}
	}
#pragma omp parallel for schedule(dynamic, 256)
	for (int i = 0; i < P; i++) {
		const float* pt = points + 3 * (size_t)order[i];
		float best[3] = { FLT_MAX_, FLT_MAX_, FLT_MAX_ };
		const int lo = i - 3 > 0 ? i - 3 : 0, hi = i + 3 < P - 1 ? i + 3 : P - 1;
		for (int j = lo; j <= hi; j++) if (j != i) knn_update3(pt, points + 3 * (size_t)order[j], best);
		const float reject = best[2];
		best[0] = best[1] = best[2] = FLT_MAX_;
		for (int b = 0; b < nb; b++) {
			const float d = box_dist(bmn + 3 * b, bmx + 3 * b, pt);
			if (d > reject || d > best[2]) continue;
			const int end = (b + 1) * BOX_SIZE < P ? (b + 1) * BOX_SIZE : P;
			for (int j = b * BOX_SIZE; j < end; j++) if (j != i) knn_update3(pt, points + 3 * (size_t)order[j], best);
		}
		mean_dists[order[i]] = ((best[0] + best[1]) + best[2]) / 3.0f;
	}
	free(keys); free(idx); free(keys_s); free(order); free(bmn); free(bmx);
}

int oracle_num_threads(void)
{
#ifdef _OPENMP
	return omp_get_max_threads();
#else
	return 1;
#endif
}
void oracle_set_num_threads(int n)
{
#ifdef _OPENMP
	if (n > 0) omp_set_num_threads(n);
#else
	(void)n;
#endif
}
