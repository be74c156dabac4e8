// compact.cu -- stream compaction of the map state (second half).
//
// Densification removes surfels with `Gaussians.prune_optimizer(mask)` (scene/Gaussians.py:143-160, called from
// remove_gaussians_from_mask :280-297): every parameter group AND its two Adam moments are re-indexed with the same
// boolean mask -- 15 boolean-index launches (each a scan + a gather inside torch) and 15 fresh allocations.  Here the
// rank of every kept surfel is computed ONCE (plan: per-block counts, one-block scan of the counts) and all 15 arrays
// are moved by ONE kernel (move) that writes its output fully coalesced.  The order of the kept surfels is preserved
// (`tensor[mask]` semantics); the Adam step counter is untouched, as in the reference (the state dict keeps "step").
//
// HBM stream: P mask bytes twice + 39 (or 36 when isotropic) floats per kept surfel in and out.
#include "common.cuh"

namespace gs2d {

namespace {

constexpr int CMP_THREADS = 256;          // surfels per block
constexpr int CMP_SCAN_THREADS = 1024;
constexpr int CMP_ARRAYS = 15;            // 5 groups x (param, exp_avg, exp_avg_sq)

struct CompactArgs {
	const float* src[CMP_ARRAYS];
	float* dst[CMP_ARRAYS];
	int width[CMP_ARRAYS];
};

__global__ void __launch_bounds__(CMP_THREADS) compact_count_kernel(int P, const uint8_t* __restrict__ keep,
                                                                    int* __restrict__ block_count)
{
	const long long i = (long long)blockIdx.x * CMP_THREADS + threadIdx.x;
	const int k = (i < P) && keep[i] != 0;
	const int c = __syncthreads_count(k);
	if (threadIdx.x == 0) block_count[blockIdx.x] = c;
}

// One block: counts -> exclusive offsets (in place), total kept -> *total.
__global__ void __launch_bounds__(CMP_SCAN_THREADS) compact_scan_kernel(int nb, int* __restrict__ block_count,
                                                                        int* __restrict__ total)
{
	__shared__ int warp_sums[32];
	__shared__ int carry;
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	if (threadIdx.x == 0) carry = 0;
	__syncthreads();
	for (int base = 0; base < nb; base += CMP_SCAN_THREADS) {
		const int i = base + threadIdx.x;
		const int v = (i < nb) ? block_count[i] : 0;
		int x = v;                                       // inclusive scan inside the warp
#pragma unroll
		for (int d = 1; d < 32; d <<= 1) {
			const int y = __shfl_up_sync(0xffffffffu, x, d);
			if (lane >= d) x += y;
		}
		if (lane == 31) warp_sums[warp] = x;
		__syncthreads();
		if (warp == 0) {                                 // inclusive scan of the 32 warp totals
			int s = warp_sums[lane];
#pragma unroll
			for (int d = 1; d < 32; d <<= 1) {
				const int y = __shfl_up_sync(0xffffffffu, s, d);
				if (lane >= d) s += y;
			}
			warp_sums[lane] = s;
		}
		__syncthreads();
		const int exclusive = carry + (warp ? warp_sums[warp - 1] : 0) + x - v;
		if (i < nb) block_count[i] = exclusive;
		__syncthreads();                                 // carry and warp_sums have been read by everyone
		if (threadIdx.x == CMP_SCAN_THREADS - 1) carry = exclusive + v;
		__syncthreads();
	}
	if (threadIdx.x == 0) *total = carry;
}

__global__ void __launch_bounds__(CMP_THREADS) compact_move_kernel(int P, const uint8_t* __restrict__ keep,
                                                                   const int* __restrict__ block_offset, const CompactArgs a)
{
	__shared__ int src_of[CMP_THREADS];       // local index of the r-th kept surfel of this block
	__shared__ int warp_sums[CMP_THREADS / 32];
	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
	const long long base = (long long)blockIdx.x * CMP_THREADS;
	const int k = (base + tid < P) && keep[base + tid] != 0;
	const unsigned ballot = __ballot_sync(0xffffffffu, k);
	if (lane == 0) warp_sums[warp] = __popc(ballot);
	__syncthreads();
	int before = 0, cnt = 0;
#pragma unroll
	for (int w = 0; w < CMP_THREADS / 32; w++) {
		const int s = warp_sums[w];
		if (w < warp) before += s;
		cnt += s;
	}
	if (k) src_of[before + __popc(ballot & ((1u << lane) - 1u))] = tid;
	__syncthreads();
	if (cnt == 0) return;
	const long long off = block_offset[blockIdx.x];
#pragma unroll
	for (int t = 0; t < CMP_ARRAYS; t++) {
		const float* __restrict__ s = a.src[t];
		float* __restrict__ d = a.dst[t];
		const int w = a.width[t];
		if (s == nullptr) continue;
		for (int j = tid; j < cnt * w; j += CMP_THREADS) {
			const int r = j / w, c = j - r * w;
			d[off * w + j] = s[(base + src_of[r]) * w + c];
		}
	}
}

inline int compact_blocks(int P) { return (P + CMP_THREADS - 1) / CMP_THREADS; }

}  // namespace

}  // namespace gs2d

using namespace gs2d;

extern "C" {

size_t gs2d_map_compact_scratch_bytes(int32_t P)
{
	return P < 0 ? 0 : sizeof(int) * (size_t)(1 + compact_blocks(P));
}

int gs2d_map_compact_plan(int32_t P, const uint8_t* keep, void* scratch, void* stream)
{
	if (P < 0 || !scratch) return GS2D_ERR_INVALID;
	cudaStream_t st = (cudaStream_t)stream;
	int* total = static_cast<int*>(scratch);
	if (P == 0) {
		GS2D_CUDA_OK(cudaMemsetAsync(total, 0, sizeof(int), st));
		return GS2D_OK;
	}
	if (!keep) return GS2D_ERR_INVALID;
	const int nb = compact_blocks(P);
	{
		ProfScope ps(K_MAP_COMPACT, st);
		compact_count_kernel<<<nb, CMP_THREADS, 0, st>>>(P, keep, total + 1);
		count_launch();
		compact_scan_kernel<<<1, CMP_SCAN_THREADS, 0, st>>>(nb, total + 1, total);
		count_launch();
	}
	GS2D_CUDA_OK(cudaGetLastError());
	return GS2D_OK;
}

int gs2d_map_compact_move(int32_t P, int32_t scale_cols, const uint8_t* keep, const gs2d_map_state* src,
                          const gs2d_map_state* dst, const void* scratch, void* stream)
{
	if (P < 0 || (scale_cols != 1 && scale_cols != 2) || !src || !dst || !scratch) return GS2D_ERR_INVALID;
	if (P == 0) return GS2D_OK;
	if (!keep) return GS2D_ERR_INVALID;
	CompactArgs a{};
	const float* sp[5] = {src->xyz, src->opacity, src->scaling, src->rotation, src->rgb};
	float* dp[5] = {dst->xyz, dst->opacity, dst->scaling, dst->rotation, dst->rgb};
	const int width[5] = {3, 1, scale_cols, 4, 3};
	for (int g = 0; g < 5; g++) {
		const float* s3[3] = {sp[g], src->exp_avg[g], src->exp_avg_sq[g]};
		float* d3[3] = {dp[g], dst->exp_avg[g], dst->exp_avg_sq[g]};
		for (int k = 0; k < 3; k++) {
			// parameters are required; a moment may be absent on both sides (optimiser that has not stepped yet)
			if ((s3[k] == nullptr) != (d3[k] == nullptr) || (k == 0 && s3[k] == nullptr)) return GS2D_ERR_INVALID;
			if (s3[k] != nullptr && s3[k] == d3[k]) return GS2D_ERR_INVALID;      // out of place only
			a.src[3 * g + k] = s3[k];
			a.dst[3 * g + k] = d3[k];
			a.width[3 * g + k] = width[g];
		}
	}
	cudaStream_t st = (cudaStream_t)stream;
	{
		ProfScope ps(K_MAP_COMPACT, st);
		compact_move_kernel<<<compact_blocks(P), CMP_THREADS, 0, st>>>(P, keep, static_cast<const int*>(scratch) + 1, a);
		count_launch();
	}
	GS2D_CUDA_OK(cudaGetLastError());
	return GS2D_OK;
}

}  // extern "C"
