// footprint.cuh -- closed-form pixel footprint of one instance inside its tile (see footprint.cu for the derivation).
#pragma once
#include "common.cuh"
#include "pair.cuh"

namespace gs2d {

constexpr float FOOT_MARGIN_PX = 0.03125f;     // widening of every interval end (pixels)
constexpr float FOOT_C_SCALE = 1.002f;         // relative widening of the rho threshold
constexpr float FOOT_C_BIAS = 0.002f;

__device__ __forceinline__ float3 cross3(const float3 a, const float3 b)
{
	return make_float3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}

// 16-bit mask of tile-local columns [xa, xb] clipped to [lo, hi] (all inclusive); empty when xa > xb.
__device__ __forceinline__ uint32_t span_bits(int xa, int xb, int lo, int hi)
{
	xa = max(xa, lo); xb = min(xb, hi);
	if (xa > xb) return 0u;
	return ((2u << xb) - 1u) & ~((1u << xa) - 1u);
}

// One thread per instance: calls emit(row, bits) for every tile row by0 .. by1 of the instance's footprint box with
// the row's 16-bit column mask (tile-local x), for the instance whose record is `rec` and box word `bw`, in its tile at pixel origin
// (ox, oy).  The row-independent part (the quadratic's leading coefficient, the expansion point) is evaluated once.
template <typename Emit>
__device__ __forceinline__ void footprint_rows(const float4* rec, const uint32_t bw, float ox, float oy, Emit emit)
{
	const float4 r0 = rec[0], r1 = rec[1], r2 = rec[2];
	const int bx0 = bw & 15, bx1 = (bw >> 8) & 15;
	const float a255 = 255.0f * r2.w;
	if (!(((bw >> 16) & 1u) && !(a255 * 1.001f < 1.0f))) return;
	const int by0 = (bw >> 4) & 15, by1 = (bw >> 12) & 15;
	const uint32_t full = span_bits(bx0, bx1, 0, 15);                 // fallback: the whole box row
	const bool nan_opacity = !(a255 == a255);   // NaN opacity: no culling
	const float c = 2.0f * __logf(fmaxf(a255, 1.0f)) * FOOT_C_SCALE + FOOT_C_BIAS;
	const float3 Tu = make_float3(r0.x, r0.y, r0.z), Tv = make_float3(r1.x, r1.y, r1.z),
	             Tw = make_float3(r2.x, r2.y, r2.z);
	// expansion point: the integer pixel nearest to the splat centre, clamped near the tile
	const float xc = fminf(fmaxf(rintf(r0.w), ox - 64.f), ox + 80.f);
	const float yc = fminf(fmaxf(rintf(r1.w), oy - 64.f), oy + 80.f);
	const float3 k0 = make_float3(xc * Tw.x - Tu.x, xc * Tw.y - Tu.y, xc * Tw.z - Tu.z);
	const float3 l0 = make_float3(yc * Tw.x - Tv.x, yc * Tw.y - Tv.y, yc * Tw.z - Tv.z);
	const float3 g = cross3(k0, l0), A = cross3(Tw, l0), B = cross3(k0, Tw);
	const float qa = A.x * A.x + A.y * A.y - c * A.z * A.z;
	// scale-free conditioning test: qa must be clearly positive (an ellipse section along the row)
	const float a2 = A.x * A.x + A.y * A.y;
	const bool conic = !nan_opacity && qa > 1e-3f * a2 && qa == qa;
	const float inv = 1.0f / qa;
	const float off = xc - ox;   // u -> tile-local x
	const float cx = r0.w - ox;
	for (int row = by0; row <= by1; row++) {
		uint32_t bits = full;
		const float yrow = oy + (float)row;
		const float v = yrow - yc;
		const float3 q = make_float3(g.x + v * B.x, g.y + v * B.y, g.z + v * B.z);
		const float qb = q.x * A.x + q.y * A.y - c * q.z * A.z;
		const float qc = q.x * q.x + q.y * q.y - c * q.z * q.z;
		if (conic && qb == qb && qc == qc) {
			const float disc = qb * qb - qa * qc;
			uint32_t m = 0;
			// widen the discriminant a little: roots of a nearly tangent row must not vanish
			const float dpos = disc + 1e-4f * (qb * qb + fabsf(qa * qc));
			if (dpos >= 0.0f) {
				const float sq = sqrtf(dpos);
				const float u1 = (-qb - sq) * inv, u2 = (-qb + sq) * inv;
				if (u1 == u1 && u2 == u2 && fabsf(u1) < 1e6f && fabsf(u2) < 1e6f) {
					const float slack = FOOT_MARGIN_PX + 1e-4f * (fabsf(u1) + fabsf(u2));
					m = span_bits((int)ceilf(u1 + off - slack), (int)floorf(u2 + off + slack), bx0, bx1);
				} else {
					m = full;
				}
			}
			// 2D low-pass filter disc around the stored centre: 100 * |d|^2 <= c
			const float dy = r1.w - yrow;
			const float r2d = 0.01f * c - dy * dy;
			if (r2d >= 0.0f) {
				const float h = sqrtf(r2d) + FOOT_MARGIN_PX;
				m |= span_bits((int)ceilf(cx - h), (int)floorf(cx + h), bx0, bx1);
			}
			bits = m;
		}
		emit(row, bits);
	}
}

}  // namespace gs2d
