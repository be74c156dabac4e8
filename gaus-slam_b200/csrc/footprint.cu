// footprint.cu -- per-instance pixel footprints (new stage; no counterpart in the reference).
//
// The reference evaluates every instance of a tile for all 256 pixels of the tile (CR/forward.cu:347-441,
// CR/backward.cu:272-461): on the BASELINE scene ~94 % of those evaluations end in one of the
// `continue`s.  This kernel finds, once per forward pass, WHICH pixels of its tile an instance can
// contribute to, in closed form: with k0 = xc*Tw - Tu, l0 = yc*Tw - Tv the pixel planes at the integer
// pixel (xc, yc) next to the splat centre, the intersection vector of pixel (xc+u, yc+v) is
//        p(u, v) = k0 x l0 + u * (Tw x l0) + v * (k0 x Tw)          (exactly bilinear-free: Tw x Tw = 0)
// so along a tile row the test rho3d = (p.x^2 + p.y^2) / p.z^2 <= c  (c = 2 ln(255 * opacity): alpha >= 1/255)
// is ONE quadratic in u.  A quarter-warp handles one instance, a lane per box row (two when the box is higher
// than 8 rows), after the row-independent part has been evaluated once: solve the quadratic,
// widen the root interval by a safety margin, union it with the (<= 0.34 px) disc of the 2D low-pass
// filter, clip to the instance's footprint box, emit the row's 16-bit mask.  Rows whose quadratic is not
// an ellipse section (grazing views) or not finite fall back to the whole box row.  The result is a
// SUPERSET of the contributing pixels (audited against the per-pixel tests by gs2d_footprint_audit), as a
// 256-bit mask per instance in TILE pixel order (word k = tile rows 2k and 2k+1, bit (y & 1) * 16 + x:
// exactly the 32 pixels warp k of a compositing CTA owns), 32 B per instance, written coalesced.
// Forward and backward compositing visit only pairs whose bit is set and re-apply the exact tests, so
// their results are unchanged.  Cost: ~100 instructions per instance, fully parallel over instances.
#include "common.cuh"
#include "pair.cuh"

namespace gs2d {

namespace {

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

__global__ void __launch_bounds__(256)
footprint_kernel(long long I, unsigned gx, float inv_gx, const float4* __restrict__ instances,
                 const uint64_t* __restrict__ keys_sorted, uint32_t* __restrict__ hit_words,
                 uint32_t* __restrict__ hit_count)
{
	// A quarter-warp per instance: the row-independent part (expansion point, the three cross products, the
	// quadratic's leading coefficient) is evaluated once, then lane l takes box rows by0 + l and by0 + l + 8.
	const int gl = threadIdx.x & 7, lane = threadIdx.x & 31;
	const long long i = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 3;
	const bool live = i < I;
	uint32_t packed = 0;          // bits of row by0 + gl (low half) and of row by0 + gl + 8 (high half)
	int by0 = 0, by1 = -1;
	if (live) {
		const float4* rec = instances + 5 * (size_t)i;
		const float4 r0 = rec[0], r1 = rec[1], r2 = rec[2];
		const uint32_t bw = __float_as_uint(rec[3].w);
		const int bx0 = bw & 15, bx1 = (bw >> 8) & 15;
		const float a255 = 255.0f * r2.w;
		if (((bw >> 16) & 1u) && !(a255 * 1.001f < 1.0f)) {
			by0 = (bw >> 4) & 15; by1 = (bw >> 12) & 15;
			const uint32_t tile = (uint32_t)(keys_sorted[i] >> 32);
			const uint32_t ty = (uint32_t)(((float)tile + 0.5f) * inv_gx);     // tile / gx, exact for tile < 2^20
			const float ox = (float)((tile - ty * gx) * TILE), oy = (float)(ty * TILE);
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
#pragma unroll
			for (int half = 0; half < 2; half++) {
				const int row = by0 + gl + 8 * half;
				if (row > by1) continue;
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
				packed |= bits << (16 * half);
			}
		}
	}
	// lane l of the group emits word l of the instance's mask: tile rows 2l (low half) and 2l + 1 (high half);
	// row r was computed by lane (r - by0) & 7 of the group, in half (r - by0) >> 3
	const int gbase = lane & ~7;
	const int da = 2 * gl - by0, db = da + 1;
	const uint32_t pa = __shfl_sync(0xffffffffu, packed, gbase + (da & 7));
	const uint32_t pb = __shfl_sync(0xffffffffu, packed, gbase + (db & 7));
	const int span = by1 - by0;
	const uint32_t ra = (da >= 0 && da <= span) ? ((pa >> (da & 8 ? 16 : 0)) & 0xffffu) : 0u;
	const uint32_t rb = (db >= 0 && db <= span) ? ((pb >> (db & 8 ? 16 : 0)) & 0xffffu) : 0u;
	uint32_t cnt = __popc(packed);
	cnt += __shfl_xor_sync(0xffffffffu, cnt, 4);
	cnt += __shfl_xor_sync(0xffffffffu, cnt, 2);
	cnt += __shfl_xor_sync(0xffffffffu, cnt, 1);
	if (live) {
		hit_words[8 * (size_t)i + gl] = ra | (rb << 16);
		if (gl == 0) hit_count[i] = cnt;
	}
}

// Audit (tests only): one thread per instance evaluates the exact tests (pair_front) at all 256 tile pixels
// and compares with the mask: counts[0] += exact hits, [1] += mask bits, [2] += exact hits MISSING in the mask.
__global__ void __launch_bounds__(128)
footprint_audit_kernel(long long I, unsigned gx, int W, int H, const float4* __restrict__ instances,
                       const uint64_t* __restrict__ keys_sorted, const uint16_t* __restrict__ hit_rows,
                       unsigned long long* counts)
{
	const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
	unsigned long long hits = 0, nbits = 0, missed = 0;
	if (i < I) {
		const float4* rec = instances + 5 * (size_t)i;
		const float4 r0 = rec[0], r1 = rec[1], r2 = rec[2];
		const uint32_t tile = (uint32_t)(keys_sorted[i] >> 32);
		const int ox = (int)(tile % gx) * TILE, oy = (int)(tile / gx) * TILE;
		for (int ly = 0; ly < TILE; ly++) {
			const uint32_t m = hit_rows[16 * (size_t)i + ly];
			nbits += __popc(m);
			for (int lx = 0; lx < TILE; lx++) {
				if (ox + lx >= W || oy + ly >= H) continue;
				PairFront f;
				if (pair_front(r0, r1, r2, (float)(ox + lx), (float)(oy + ly), f)) {
					hits++;
					if (!((m >> lx) & 1u)) missed++;
				}
			}
		}
	}
	for (int o = 16; o; o >>= 1) {
		hits += __shfl_xor_sync(0xffffffffu, hits, o);
		nbits += __shfl_xor_sync(0xffffffffu, nbits, o);
		missed += __shfl_xor_sync(0xffffffffu, missed, o);
	}
	if ((threadIdx.x & 31) == 0) {
		atomicAdd(&counts[0], hits);
		atomicAdd(&counts[1], nbits);
		atomicAdd(&counts[2], missed);
	}
}

}  // namespace

int launch_footprint_audit(const gs2d_settings& s, const float* instances, const uint64_t* keys_sorted,
                           const uint32_t* hit_inst, int64_t I, unsigned long long* counts, cudaStream_t st)
{
	if (I <= 0) return GS2D_OK;
	const unsigned gx = (s.image_width + TILE - 1) / TILE;
	footprint_audit_kernel<<<(unsigned)((I + 127) / 128), 128, 0, st>>>(
		I, gx, s.image_width, s.image_height, (const float4*)instances, keys_sorted, (const uint16_t*)hit_inst, counts);
	GS2D_CUDA_OK(cudaGetLastError());
	return GS2D_OK;
}

int launch_footprint(const gs2d_settings& s, const float* instances, const uint64_t* keys_sorted,
                     uint32_t* hit_inst, uint32_t* hit_count, int64_t I, cudaStream_t st)
{
	if (I <= 0) return GS2D_OK;
	const unsigned gx = (s.image_width + TILE - 1) / TILE;
	const long long threads = I * 8;
	ProfScope ps(K_FOOTPRINT, st);
	footprint_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, st>>>(I, gx, 1.0f / (float)gx, (const float4*)instances,
	                                                                    keys_sorted, hit_inst, hit_count);
	count_launch();
	GS2D_CUDA_OK(cudaGetLastError());
	return GS2D_OK;
}

}  // namespace gs2d
