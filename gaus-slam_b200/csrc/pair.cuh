// pair.cuh -- the per-(pixel, instance) ray-splat evaluation shared by the footprint, forward and
// backward compositing kernels.
//
// pair_front() is the first half of the reference's inner loop body (CR/forward.cu:355-386,
// CR/backward.cu:283-316): intersect the two pixel planes with the splat plane, low-pass filter,
// depth, Gaussian weight, alpha, and the four "skip this instance" tests.  The expressions are kept
// in the reference's form so that nvcc fuses them the same way (the forward images are bit-identical
// to the reference build; tests/test_gpu_parity.py).
#pragma once
#include "common.cuh"

namespace gs2d {

// In-tile footprint box of an instance, packed in record word REC_MASK:
//   bits 0..3 x0, 4..7 y0, 8..11 x1, 12..15 y1 (tile-local pixel coordinates, inclusive), bit 16 valid.
struct TileBox {
	int x0, y0, w, h, area;
};
__device__ __forceinline__ TileBox unpack_box(uint32_t word)
{
	TileBox b;
	b.x0 = word & 15; b.y0 = (word >> 4) & 15;
	b.w = (int)((word >> 8) & 15) - b.x0 + 1;
	b.h = (int)((word >> 12) & 15) - b.y0 + 1;
	b.area = (word >> 16) & 1 ? b.w * b.h : 0;
	return b;
}
// ceil(65536 / w) for w = 1..16: floor(idx / w) == (idx * kInvW[w]) >> 16 for idx < 256.
__device__ __constant__ const uint32_t kInvW[17] = {0,     65536, 32768, 21846, 16384, 13108, 10923, 9363, 8192,
                                                    7282,  6554,  5958,  5462,  5042,  4682,  4370, 4096};
// idx -> tile-local pixel inside a box (row-major inside the box)
__device__ __forceinline__ void box_xy(const TileBox& b, int idx, int& lx, int& ly)
{
	const uint32_t q = ((uint32_t)idx * kInvW[b.w]) >> 16;
	lx = b.x0 + idx - (int)q * b.w;
	ly = b.y0 + (int)q;
}
// Compositing CTAs own one pixel per thread in tile row-major order: thread t <-> pixel (t & 15, t >> 4), so
// warp k owns rows 2k and 2k+1.  Footprint masks use the same order: word k, bit (ly & 1) * 16 + lx.
__device__ __forceinline__ int pixel_thread(int lx, int ly) { return (ly << 4) | lx; }

// Which 32 pixels of the tile a compositing warp owns (a per-kernel choice; the per-pixel words the forward pass
// leaves for the backward pass are indexed by pixel_thread(lx, ly), i.e. independent of it).  BLOCK = 0: tile rows 2k
// and 2k+1 (16 x 2, the mask word itself); 1: an 8 x 4 block -- warp k = block column k & 1, block row k >> 1, lane l =
// pixel (l & 7, l >> 3) of the block.  A compact block shares more of its instances between its pixels, so the
// per-pixel lists of a warp are more alike (a warp iterates as often as its longest list): measured -3 % on the
// forward kernel, +14 % on the backward kernel (its phase B wants an instance's pair slots in row order), hence
// forward 8 x 4, backward rows.  The footprint masks stay row-major; an 8 x 4 warp's word is byte bx of the four
// 16-bit row masks of its block rows.
#ifndef GS2D_FWD_BLOCK
#define GS2D_FWD_BLOCK 1
#endif
#ifndef GS2D_BWD_BLOCK
#define GS2D_BWD_BLOCK 0
#endif
template <int BLOCK>
__device__ __forceinline__ int thread_lx(int tid)
{
	return BLOCK ? (((tid >> 5) & 1) << 3) | (tid & 7) : (tid & 15);
}
template <int BLOCK>
__device__ __forceinline__ int thread_ly(int tid)
{
	return BLOCK ? ((tid >> 6) << 2) | ((tid >> 3) & 3) : (tid >> 4);
}
// word of warp `warp` from the 8 row-pair words of one instance's mask
template <int BLOCK>
__device__ __forceinline__ uint32_t warp_word(const uint32_t* __restrict__ mask8, int warp)
{
	if (BLOCK) {
		const uint2 rows = *reinterpret_cast<const uint2*>(mask8 + 2 * (warp >> 1));    // block rows 0,1 | 2,3
		return __byte_perm(rows.x, rows.y, (warp & 1) ? 0x7531 : 0x6420);
	}
	return mask8[warp];
}

// In-tile footprint box of an instance: the surfel's contribution bounds (preprocess.cu) clipped to the
// tile and to the image, packed as described in pair.cuh (x0 | y0<<4 | x1<<8 | y1<<12 | valid<<16).
__device__ __forceinline__ uint32_t tile_box_word(float4 cb, int tx, int ty, int W, int H)
{
	const int tx0 = tx * TILE, ty0 = ty * TILE;
	const int tx1 = min(tx0 + TILE - 1, W - 1), ty1 = min(ty0 + TILE - 1, H - 1);
	// clamp in float first: the bounds may be +-inf ("no culling")
	const int x0 = (int)ceilf(fminf(fmaxf(cb.x, (float)tx0), (float)(tx1 + 1)));
	const int y0 = (int)ceilf(fminf(fmaxf(cb.y, (float)ty0), (float)(ty1 + 1)));
	const int x1 = (int)floorf(fmaxf(fminf(cb.z, (float)tx1), (float)(tx0 - 1)));
	const int y1 = (int)floorf(fmaxf(fminf(cb.w, (float)ty1), (float)(ty0 - 1)));
	if (x0 > x1 || y0 > y1) return 0u;
	return (uint32_t)(x0 - tx0) | ((uint32_t)(y0 - ty0) << 4) | ((uint32_t)(x1 - tx0) << 8) |
	       ((uint32_t)(y1 - ty0) << 12) | (1u << 16);
}

struct PairFront {
	float3 k, l;      // pixel planes in splat space
	float pz;         // p.z of k x l
	float2 s;         // intersection in splat coordinates
	float2 d;         // centre - pixel
	float rho3d, rho2d;
	float depth;      // ray depth of the intersection (or of the centre for the 2D filter branch)
	float G;          // exp(-rho/2)
	float alpha;
};

// Returns false when the reference `continue`s (no contribution possible); fills f otherwise.
// ALPHA_MIN / DEPTH_MIN let the footprint kernel use slightly relaxed thresholds (superset).
__device__ __forceinline__ bool pair_front(const float4 r0, const float4 r1, const float4 r2, const float pxf,
                                           const float pyf, PairFront& f, const float alpha_min = 1.0f / 255.0f,
                                           const float depth_min = GS2D_NEAR)
{
	const float3 Tu = make_float3(r0.x, r0.y, r0.z);
	const float3 Tv = make_float3(r1.x, r1.y, r1.z);
	const float3 Tw = make_float3(r2.x, r2.y, r2.z);
	const float2 xy = make_float2(r0.w, r1.w);
	const float opa = r2.w;
	f.k = make_float3(pxf * Tw.x - Tu.x, pxf * Tw.y - Tu.y, pxf * Tw.z - Tu.z);
	f.l = make_float3(pyf * Tw.x - Tv.x, pyf * Tw.y - Tv.y, pyf * Tw.z - Tv.z);
	const float3 p = make_float3(f.k.y * f.l.z - f.k.z * f.l.y, f.k.z * f.l.x - f.k.x * f.l.z,
	                             f.k.x * f.l.y - f.k.y * f.l.x);
	f.pz = p.z;
	if (p.z == 0.0f) return false;   // (the reference writes 0.0: the promoted double compare decides the same)
	f.s = make_float2(p.x / p.z, p.y / p.z);
	f.rho3d = (f.s.x * f.s.x + f.s.y * f.s.y);
	f.d = make_float2(xy.x - pxf, xy.y - pyf);
	f.rho2d = GS2D_FILTER_INV_SQUARE * (f.d.x * f.d.x + f.d.y * f.d.y);
	const float rho = fminf(f.rho3d, f.rho2d);
	f.depth = (f.rho3d <= f.rho2d) ? (f.s.x * Tw.x + f.s.y * Tw.y) + Tw.z : Tw.z;
	if (f.depth < depth_min) return false;
	const float power = -0.5f * rho;
	if (power > 0.0f) return false;
	f.G = expf(power);
	f.alpha = fminf(0.99f, opa * f.G);
	if (f.alpha < alpha_min) return false;
	return true;
}

// The same quantities without the tests (phase B of the backward: the pair was accepted by pair_front already).
__device__ __forceinline__ void pair_geom(const float4 r0, const float4 r1, const float4 r2, const float pxf,
                                          const float pyf, PairFront& f)
{
	const float3 Tu = make_float3(r0.x, r0.y, r0.z);
	const float3 Tv = make_float3(r1.x, r1.y, r1.z);
	const float3 Tw = make_float3(r2.x, r2.y, r2.z);
	f.k = make_float3(pxf * Tw.x - Tu.x, pxf * Tw.y - Tu.y, pxf * Tw.z - Tu.z);
	f.l = make_float3(pyf * Tw.x - Tv.x, pyf * Tw.y - Tv.y, pyf * Tw.z - Tv.z);
	const float3 p = make_float3(f.k.y * f.l.z - f.k.z * f.l.y, f.k.z * f.l.x - f.k.x * f.l.z,
	                             f.k.x * f.l.y - f.k.y * f.l.x);
	f.pz = p.z;
	f.s = make_float2(p.x / p.z, p.y / p.z);
	f.rho3d = (f.s.x * f.s.x + f.s.y * f.s.y);
	f.d = make_float2(r0.w - pxf, r1.w - pyf);
	f.rho2d = GS2D_FILTER_INV_SQUARE * (f.d.x * f.d.x + f.d.y * f.d.y);
	f.G = expf(-0.5f * fminf(f.rho3d, f.rho2d));
}

// Backward-pass variant: MUFU reciprocal / exp2 instead of IEEE division / expf (the gradient tolerance is
// 1e-3 relative; these are ~1e-6), no tests: the forward pass recorded which pairs contributed.
__device__ __forceinline__ void pair_geom_fast(const float4 r0, const float4 r1, const float4 r2, const float pxf,
                                               const float pyf, PairFront& f, float& inv_pz)
{
	f.k = make_float3(pxf * r2.x - r0.x, pxf * r2.y - r0.y, pxf * r2.z - r0.z);
	f.l = make_float3(pyf * r2.x - r1.x, pyf * r2.y - r1.y, pyf * r2.z - r1.z);
	const float3 p = make_float3(f.k.y * f.l.z - f.k.z * f.l.y, f.k.z * f.l.x - f.k.x * f.l.z,
	                             f.k.x * f.l.y - f.k.y * f.l.x);
	f.pz = p.z;
	inv_pz = __fdividef(1.0f, p.z);
	f.s = make_float2(p.x * inv_pz, p.y * inv_pz);
	f.rho3d = (f.s.x * f.s.x + f.s.y * f.s.y);
	f.d = make_float2(r0.w - pxf, r1.w - pyf);
	f.rho2d = GS2D_FILTER_INV_SQUARE * (f.d.x * f.d.x + f.d.y * f.d.y);
	f.depth = (f.rho3d <= f.rho2d) ? (f.s.x * r2.x + f.s.y * r2.y) + r2.z : r2.z;
	f.G = __expf(-0.5f * fminf(f.rho3d, f.rho2d));
	f.alpha = fminf(0.99f, r2.w * f.G);
}

// Same tests with approximate reciprocal / exp2 (for the footprint kernel only, which compensates with
// relaxed thresholds): true when the pair may contribute.
__device__ __forceinline__ bool pair_may_contribute(const float4 r0, const float4 r1, const float4 r2, const float pxf,
                                                    const float pyf, const float alpha_min, const float depth_min)
{
	const float3 k = make_float3(pxf * r2.x - r0.x, pxf * r2.y - r0.y, pxf * r2.z - r0.z);
	const float3 l = make_float3(pyf * r2.x - r1.x, pyf * r2.y - r1.y, pyf * r2.z - r1.z);
	const float3 p = make_float3(k.y * l.z - k.z * l.y, k.z * l.x - k.x * l.z, k.x * l.y - k.y * l.x);
	// p.z == 0 is not an early-out here: inf / NaN fall through fminf to the 2D-filter branch (superset)
	const float inv = __frcp_rn(p.z);
	const float sx = p.x * inv, sy = p.y * inv;
	const float rho3d = sx * sx + sy * sy;
	const float dx = r0.w - pxf, dy = r1.w - pyf;
	const float rho2d = GS2D_FILTER_INV_SQUARE * (dx * dx + dy * dy);
	const float depth = (rho3d <= rho2d) ? (sx * r2.x + sy * r2.y) + r2.z : r2.z;
	if (depth < depth_min) return false;
	const float rho = fminf(rho3d, rho2d);
	if (rho < 0.0f) return false;
	const float alpha = fminf(0.99f, r2.w * exp2f(-0.72134752f * rho));
	return !(alpha < alpha_min);
}

}  // namespace gs2d
