// common.cuh -- shared declarations for libgs2d (B200 / sm_100a).
//
// Internal to the library; the public surface is include/gs2d.h.
#pragma once

#include <cstddef>
#include <cstdint>
#include <cuda_runtime.h>

#include "../../include/gs2d.h"

namespace gs2d {

constexpr int TILE = GS2D_TILE;          // 16x16 pixel tiles (CR/config.h:16-17)
constexpr int TILE_PIXELS = TILE * TILE;
constexpr int REC_WORDS = 20;            // one splat / instance record = 5 x float4 = 80 B
constexpr int GRAD_WORDS = 20;           // one per-instance gradient row = 80 B

// Record word layout (per surfel in `splat`, per sorted instance in `instances`):
//   0..2  Tu     3  centre x        (transMat rows, CR/forward.cu:196-199; means2D)
//   4..6  Tv     7  centre y
//   8..10 Tw    11  opacity
//  12..14 normal 15 footprint box (instances only: in-tile pixel box, packed as in pair.cuh)
//  16..18 rgb    19 row (instances only: index of this instance's gradient row)
constexpr int REC_MASK = 15;
constexpr int REC_ROW = 19;

// Gradient row layout (backward compositing -> backward preprocess):
//   0..8 dL/dtransMat, 9..10 dL/dmean2D.xy, 11 dL/dopacity, 12..14 dL/dnormal, 15..17 dL/dcolor
constexpr int GR_T = 0, GR_M2D = 9, GR_OPA = 11, GR_NRM = 12, GR_COL = 15;
constexpr int GR_USED = 18;

// Constants of the reference (CR/auxiliary.h:37-39).
#define GS2D_NEAR 0.2f
#define GS2D_FAR 100.0f
#define GS2D_FILTER_INV_SQUARE 100.0f

constexpr size_t ALIGN = 256;

struct Carver {
	char* p;
	size_t off;
	explicit __host__ Carver(void* base) : p((char*)base), off(0) {}
	template <typename T>
	__host__ T* take(size_t count) {
		off = (off + ALIGN - 1) & ~(ALIGN - 1);
		T* r = p ? (T*)(p + off) : nullptr;
		off += count * sizeof(T);
		return r;
	}
	__host__ size_t total() const { return ((off + ALIGN - 1) & ~(ALIGN - 1)) + ALIGN; }
};

// ---- scan (decoupled look-back) geometry --------------------------------------------
// hit_pix holds, per tile, one word pair per 32 instances rounded up to whole compositing batches: HIT_PIX_PAD
// spare words per tile cover the rounding for batches of up to 32 * HIT_PIX_PAD instances.
constexpr int HIT_PIX_PAD = 8;
// The per-tile instance counters / scatter cursors are hit by one atomic per instance (~250 per tile and frame).  L2
// serialises atomics that fall into the same line, so every tile's word sits in a line of its own (TC_STRIDE words).
#ifndef GS2D_TC_STRIDE
#define GS2D_TC_STRIDE 32
#endif
constexpr int TC_STRIDE = GS2D_TC_STRIDE;
constexpr int SCAN_THREADS = 256;
#ifndef GS2D_SCAN_ITEMS
#define GS2D_SCAN_ITEMS 8
#endif
constexpr int SCAN_ITEMS = GS2D_SCAN_ITEMS;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;

// ---- onesweep radix sort geometry ----------------------------------------------------
constexpr int SORT_THREADS = 256;
#ifndef GS2D_SORT_ITEMS
#define GS2D_SORT_ITEMS 8
#endif
constexpr int SORT_ITEMS = GS2D_SORT_ITEMS;
constexpr int SORT_TILE = SORT_THREADS * SORT_ITEMS;   // pairs per block
constexpr int SORT_RADIX_BITS = 8;
constexpr int SORT_RADIX = 1 << SORT_RADIX_BITS;
constexpr int SORT_MAX_PASSES = 8;

struct GeomState {
	float* depths;            // [P]
	float* splat;             // [P,20]
	uint16_t* rect;           // [P,4]
	float* cull;              // [P,4]
	uint32_t* tiles_touched;  // [P]
	uint32_t* point_offsets;  // [P]
	uint8_t* clamped;         // [P,3]
	uint64_t* scan_status;    // [scan blocks]
	uint32_t* counters;       // [8]: 0 scan ticket, 1 num_rendered (u32), 2.. spare
	static __host__ GeomState carve(void* base, size_t P, size_t* bytes);
};

struct ImageState {
	float* final_T;           // [3,H,W]: T, M1, M2
	uint32_t* n_contrib;      // [2,H,W]: last contributor, median contributor
	float* median_depth;      // [H,W]
	float* depth_std;         // [H,W]
	uint32_t* ranges;         // [tiles,2]
	uint32_t* tile_count;     // [tiles * TC_STRIDE] instances per tile at word t * TC_STRIDE (counted by preprocess_forward; zeroed before it)
	uint32_t* tile_cursor;    // [tiles * TC_STRIDE] write cursor of the scatter at word t * TC_STRIDE (starts at the tile's range start)
	static __host__ ImageState carve(void* base, size_t W, size_t H, size_t* bytes);
};

struct SortScratch {
	uint32_t* hist;           // [SORT_MAX_PASSES, 256] -> exclusive digit bases
	uint32_t* tickets;        // [SORT_MAX_PASSES]
	uint32_t* status;         // [SORT_MAX_PASSES, blocks, 256]
	size_t blocks;
	size_t zero_bytes;        // bytes from `hist` that must be zeroed before a sort
	static __host__ SortScratch carve(void* base, size_t n, size_t* bytes);
};

struct BinningState {
	uint64_t* keys_a;         // [I] unsorted keys (pass ping)
	uint64_t* keys_b;         // [I] (pass pong)
	uint32_t* vals_a;         // [I]
	uint32_t* vals_b;         // [I]
	float* instances;         // [I,20]
	uint32_t* hit_inst;       // [I,8] per-instance footprint mask in tile pixel order (footprint.cu)
	uint32_t* hit_pix;        // [I/32 + HIT_PIX_PAD*tiles, 2, 256] per-pixel words written by the forward pass: tile t, batch b,
	                          // word g at ((((range.x >> 5) + 4*t + 4*b + g) * 2 + k) * 256 + pixel thread;
	                          // k = 0: pairs that contributed, k = 1: pairs of the footprint masks
	uint32_t* hit_count;      // [I] footprint-mask bits per instance
	uint32_t* pair_incl;      // [I] inclusive scan of hit_count = end of the instance's pair-slot range
	uint64_t* pair_status;    // scan look-back status
	uint32_t* pair_counters;  // [8]: [1] = total number of pair slots
	uint64_t* keys_dbg;       // [I] copy of the unsorted keys
	uint32_t* vals_dbg;       // [I]
	SortScratch sort;
	static __host__ BinningState carve(void* base, size_t I, size_t tiles, size_t* bytes);
};

inline int sort_passes(int end_bit) { return (end_bit + SORT_RADIX_BITS - 1) / SORT_RADIX_BITS; }

// Host-side launch bookkeeping -----------------------------------------------------------
extern thread_local int g_last_cuda_error;
void count_launch(int n = 1);
#define GS2D_CUDA_OK(expr)                                              \
	do {                                                                \
		cudaError_t e__ = (expr);                                       \
		if (e__ != cudaSuccess) { gs2d::g_last_cuda_error = (int)e__; return GS2D_ERR_CUDA; } \
	} while (0)

// Optional per-kernel CUDA-event timing (gs2d_timing_* in include/gs2d.h) -------------------------
enum KernelId {
	K_PREPROCESS_FWD = 0, K_SCAN, K_DUPLICATE, K_HISTOGRAM, K_DIGIT_BASE, K_ONESWEEP, K_GATHER_RANGES,
	K_COMPOSITE_FWD, K_COMPOSITE_BWD, K_PREPROCESS_BWD, K_MARK_VISIBLE, K_MEMSET, K_MEMCPY, K_KNN, K_FOOTPRINT, K_COMPOSITE_BWD_INST, K_TILE_OFFSETS,
	K_SCATTER, K_TILE_SORT, K_LOSS, K_ALLREDUCE, K_MAP_STEP, K_POSE, K_MAP_COMPACT, K_COUNT
};
extern bool g_timing_on;
void prof_mark(int id, cudaStream_t st, bool begin);
struct ProfScope {
	int id; cudaStream_t st;
	ProfScope(int id_, cudaStream_t st_) : id(id_), st(st_) { if (g_timing_on) prof_mark(id, st, true); }
	~ProfScope() { if (g_timing_on) prof_mark(id, st, false); }
};

// Kernel launchers (each returns a gs2d_status) ------------------------------------------
int launch_preprocess_forward(const gs2d_settings& s, const gs2d_inputs& in, const GeomState& g,
                              uint32_t* tile_count, int32_t* radii, cudaStream_t st);
int launch_scatter(const gs2d_settings& s, const GeomState& g, const ImageState& im, uint64_t* pairs, bool early,
                   int64_t capacity, cudaStream_t st);
int launch_tile_sort(const gs2d_settings& s, const GeomState& g, const ImageState& im, const BinningState& b,
                     uint64_t* pairs, int64_t I, cudaStream_t st, int tile0 = 0, int ntiles = -1);
// [scan_status, counters + 8) must be zero (preprocess_forward_kernel clears them); one extra block turns the per-tile counts into ranges + cursors
int launch_scan_tiles(const GeomState& g, int P, cudaStream_t st, int tiles, const uint32_t* tile_count,
                      uint32_t* ranges, uint32_t* tile_cursor);
int launch_scan_u32(const uint32_t* in, uint32_t* out, long long n, uint64_t* status, uint32_t* counters,
                    cudaStream_t st, bool clear = true, int side_n = 0, const uint32_t* side_in = nullptr,
                    uint32_t* side_ranges = nullptr, uint32_t* side_cursor = nullptr);
int launch_duplicate(const gs2d_settings& s, const GeomState& g, const int32_t* radii_unused,
                     const BinningState& b, int64_t I, cudaStream_t st);
int launch_sort_pairs(int64_t n, const uint64_t* keys_a, const uint32_t* vals_a, uint64_t* keys_b,
                      uint32_t* vals_b, uint64_t* keys_a_rw, uint32_t* vals_a_rw, int end_bit,
                      const SortScratch& ss, bool* result_in_b, cudaStream_t st);
int launch_footprint_audit(const gs2d_settings& s, const float* instances, const uint64_t* keys_sorted,
                           const uint32_t* hit_inst, int64_t I, unsigned long long* counts, cudaStream_t st);
int launch_composite_forward(const gs2d_settings& s, const gs2d_inputs& in, const ImageState& im,
                             const float* instances, const uint32_t* hit_inst, uint32_t* hit_pix, float* out_color,
                             float* out_others, cudaStream_t st, int tile0 = 0, int ntiles = -1);
size_t composite_backward_scratch_bytes(int64_t I, int64_t pairs, int64_t tiles);
int launch_composite_backward(const gs2d_settings& s, const gs2d_inputs& in, const ImageState& im,
                              const BinningState& b, int64_t I, int64_t pairs, const float* dL_dcolor,
                              const float* dL_dothers, void* scratch, float** grad_rows_out, cudaStream_t st);
int launch_preprocess_backward(const gs2d_settings& s, const gs2d_inputs& in, const GeomState& g,
                               const int32_t* radii, const float* grad_rows, const gs2d_grads& out,
                               cudaStream_t st, float* mc_flat = nullptr);
int launch_mark_visible(int P, const float* means3D, const float* view, const float* proj,
                        uint8_t* present, cudaStream_t st);

uint32_t higher_msb(uint32_t n);   // CR/rasterizer_impl.cu:35-50 semantics

}  // namespace gs2d
