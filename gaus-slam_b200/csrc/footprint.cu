// footprint.cu -- per-instance pixel footprints (new stage; no counterpart in the reference).
//
// The reference evaluates every instance of a tile for all 256 pixels of the tile (CR/forward.cu:347-441,
// CR/backward.cu:272-461): on the BASELINE scene ~94 % of those evaluations end in one of the
// `continue`s.  This kernel finds, once per forward pass, WHICH pixels of its tile an instance can
// contribute to, in closed form: with k0 = xc*Tw - Tu, l0 = yc*Tw - Tv the pixel planes at the integer
// pixel (xc, yc) next to the splat centre, the intersection vector of pixel (xc+u, yc+v) is
//        p(u, v) = k0 x l0 + u * (Tw x l0) + v * (k0 x Tw)          (exactly bilinear-free: Tw x Tw = 0)
// so along a tile row the test rho3d = (p.x^2 + p.y^2) / p.z^2 <= c  (c = 2 ln(255 * opacity): alpha >= 1/255)
// is ONE quadratic in u.  One thread handles one instance (footprint.cuh, called from tile_sort_kernel): the
// row-independent part is evaluated once, then for every row of the instance's box: solve the quadratic,
// widen the root interval by a safety margin, union it with the (<= 0.34 px) disc of the 2D low-pass
// filter, clip to the instance's footprint box, emit the row's 16-bit mask.  Rows whose quadratic is not
// an ellipse section (grazing views) or not finite fall back to the whole box row.  The result is a
// SUPERSET of the contributing pixels (audited against the per-pixel tests by gs2d_footprint_audit), as a
// 256-bit mask per instance in TILE pixel order (word k = tile rows 2k and 2k+1, bit (y & 1) * 16 + x:
// exactly the 32 pixels warp k of a compositing CTA owns), 32 B per instance, written coalesced.
// Forward and backward compositing visit only pairs whose bit is set and re-apply the exact tests, so
// their results are unchanged.  Cost: ~100 instructions per instance, fully parallel over instances.
#include "footprint.cuh"

namespace gs2d {

namespace {

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

}  // namespace gs2d
