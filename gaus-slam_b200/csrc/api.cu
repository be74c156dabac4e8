// api.cu -- the C ABI of libgs2d (include/gs2d.h): buffer layouts and stage orchestration.
//
// Mirrors the control flow of CudaRasterizer::Rasterizer::forward / backward
// (CR/rasterizer_impl.cu:201-350, :354-460) and the tensor bookkeeping of
// RAST/rasterize_points.cu:39-239, minus every memset the new kernels make unnecessary.
#include <atomic>
#include <cstring>
#include <mutex>
#include <vector>

#include "common.cuh"

namespace gs2d {

thread_local int g_last_cuda_error = 0;
static std::atomic<long long> g_launches{0};
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

// ---- per-kernel event timing ---------------------------------------------------------------
bool g_timing_on = false;
namespace {
struct ProfRec { int id; cudaEvent_t a, b; };
std::mutex g_prof_mu;
std::vector<ProfRec> g_prof_open;          // begun, not yet ended (LIFO)
std::vector<ProfRec> g_prof_done;          // both events recorded
std::vector<cudaEvent_t> g_prof_pool;
cudaEvent_t prof_event()
{
	if (!g_prof_pool.empty()) { cudaEvent_t e = g_prof_pool.back(); g_prof_pool.pop_back(); return e; }
	cudaEvent_t e = nullptr;
	cudaEventCreate(&e);
	return e;
}
const char* const kKernelNames[K_COUNT] = {
	"preprocess_forward", "scan_tiles", "duplicate_keys", "histogram", "digit_base", "onesweep_pass",
	"gather_ranges", "composite_forward", "composite_backward", "preprocess_backward", "mark_visible",
	"memset", "memcpy", "knn", "footprint", "composite_backward_instances", "tile_offsets", "scatter_pairs",
	"tile_sort_gather_footprint", "mapping_loss", "allreduce_multimem", "map_adam_step", "pose_transform", "map_compact"};
}  // namespace
void prof_mark(int id, cudaStream_t st, bool begin)
{
	std::lock_guard<std::mutex> lk(g_prof_mu);
	if (begin) {
		ProfRec r{id, prof_event(), prof_event()};
		cudaEventRecord(r.a, st);
		g_prof_open.push_back(r);
	} else if (!g_prof_open.empty()) {
		ProfRec r = g_prof_open.back();
		g_prof_open.pop_back();
		cudaEventRecord(r.b, st);
		g_prof_done.push_back(r);
	}
}

GeomState GeomState::carve(void* base, size_t P, size_t* bytes)
{
	Carver c(base);
	GeomState g;
	g.depths = c.take<float>(P);
	g.splat = c.take<float>(P * REC_WORDS);
	g.rect = c.take<uint16_t>(P * 4);
	g.cull = c.take<float>(P * 4);
	g.tiles_touched = c.take<uint32_t>(P);
	g.point_offsets = c.take<uint32_t>(P);
	g.clamped = c.take<uint8_t>(P * 3);
	g.scan_status = c.take<uint64_t>((P + SCAN_TILE - 1) / SCAN_TILE + 1);
	g.counters = c.take<uint32_t>(8);
	if (bytes) *bytes = c.total();
	return g;
}

ImageState ImageState::carve(void* base, size_t W, size_t H, size_t* bytes)
{
	Carver c(base);
	ImageState im;
	const size_t px = W * H;
	const size_t tiles = ((W + TILE - 1) / TILE) * ((H + TILE - 1) / TILE);
	im.final_T = c.take<float>(3 * px);
	im.n_contrib = c.take<uint32_t>(2 * px);
	im.median_depth = c.take<float>(px);
	im.depth_std = c.take<float>(px);
	im.ranges = c.take<uint32_t>(2 * tiles);
	im.tile_count = c.take<uint32_t>(tiles * TC_STRIDE);
	im.tile_cursor = c.take<uint32_t>(tiles * TC_STRIDE);
	if (bytes) *bytes = c.total();
	return im;
}

SortScratch SortScratch::carve(void* base, size_t n, size_t* bytes)
{
	Carver c(base);
	SortScratch s;
	s.blocks = (n + SORT_TILE - 1) / SORT_TILE;
	if (s.blocks == 0) s.blocks = 1;
	s.hist = c.take<uint32_t>(SORT_MAX_PASSES * SORT_RADIX);
	char* z0 = (char*)s.hist;
	s.tickets = c.take<uint32_t>(SORT_MAX_PASSES);
	s.status = c.take<uint32_t>((size_t)SORT_MAX_PASSES * s.blocks * SORT_RADIX);
	char* z1 = (char*)(s.status + (size_t)SORT_MAX_PASSES * s.blocks * SORT_RADIX);
	s.zero_bytes = (size_t)(z1 - z0);
	if (bytes) *bytes = c.total();
	return s;
}

BinningState BinningState::carve(void* base, size_t I, size_t tiles, size_t* bytes)
{
	Carver c(base);
	BinningState b;
	const size_t n = I ? I : 1;
	b.keys_a = c.take<uint64_t>(n);
	b.keys_b = c.take<uint64_t>(n);
	b.vals_a = c.take<uint32_t>(n);
	b.vals_b = c.take<uint32_t>(n);
	b.instances = c.take<float>(n * REC_WORDS);
	b.hit_inst = c.take<uint32_t>(n * 8 + 64);
	b.hit_pix = c.take<uint32_t>((n / 32 + HIT_PIX_PAD * tiles + 8) * 2 * TILE_PIXELS);
	b.hit_count = c.take<uint32_t>(n);
	b.pair_incl = c.take<uint32_t>(n);
	b.pair_status = c.take<uint64_t>(n / SCAN_TILE + 2);
	b.pair_counters = c.take<uint32_t>(8);
	b.keys_dbg = c.take<uint64_t>(n);
	b.vals_dbg = c.take<uint32_t>(n);
	size_t sort_bytes = 0;
	SortScratch::carve(nullptr, n, &sort_bytes);
	char* sbase = c.take<char>(sort_bytes);
	b.sort = SortScratch::carve(sbase, n, nullptr);
	if (bytes) *bytes = c.total();
	return b;
}

// Runs of tiles the forward pass is cut into (tile sort of run k+1 beside the compositing of run k, see
// gs2d_forward_render_scattered).  Measured on the BASELINE frame: 1 run 0.662 ms / step, 2 runs 0.681, 4 runs 0.739,
// 8 runs 0.828 -- the two kernels do not share an SM well (whichever gets there first fills it), so the default is 1.
#ifndef GS2D_FWD_PIPE
#define GS2D_FWD_PIPE 1
#endif

// A per-device side stream + a small pool of events for the one fork / join inside the forward pass (the pair-count
// scan runs beside the forward compositing).  The wait on an event is enqueued in the same critical section as its
// record, so a pooled event can be re-recorded by a later call without affecting an earlier wait.
struct SideStream {
	cudaStream_t stream = nullptr;
	cudaEvent_t ev[32] = {};
	unsigned next = 0;
};
static std::mutex g_side_mu;
static SideStream g_side[64];

// fork: `side` waits for everything enqueued on `st` so far.  Returns the side stream (nullptr = run serially on st).
static cudaStream_t side_fork(cudaStream_t st)
{
	int dev = 0;
	if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
	std::lock_guard<std::mutex> lk(g_side_mu);
	SideStream& ss = g_side[dev];
	if (!ss.stream) {
		if (cudaStreamCreateWithFlags(&ss.stream, cudaStreamNonBlocking) != cudaSuccess) { ss.stream = nullptr; return nullptr; }
		for (auto& e : ss.ev)
			if (cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess) return nullptr;
	}
	cudaEvent_t e = ss.ev[ss.next++ % 32];
	if (cudaEventRecord(e, st) != cudaSuccess || cudaStreamWaitEvent(ss.stream, e, 0) != cudaSuccess) return nullptr;
	return ss.stream;
}

// join: everything enqueued on `st` after this call waits for what was enqueued on the side stream so far
static int side_join(cudaStream_t st, cudaStream_t side)
{
	int dev = 0;
	GS2D_CUDA_OK(cudaGetDevice(&dev));
	std::lock_guard<std::mutex> lk(g_side_mu);
	SideStream& ss = g_side[dev];
	cudaEvent_t e = ss.ev[ss.next++ % 32];
	GS2D_CUDA_OK(cudaEventRecord(e, side));
	GS2D_CUDA_OK(cudaStreamWaitEvent(st, e, 0));
	return GS2D_OK;
}

static inline size_t tiles_of(int W, int H)
{
	return (size_t)((W + TILE - 1) / TILE) * (size_t)((H + TILE - 1) / TILE);
}

static int check_common(const gs2d_settings* s, const gs2d_inputs* in, const gs2d_buffers* b)
{
	if (!s || !in || !b) return GS2D_ERR_INVALID;
	if (s->P < 0 || s->image_width <= 0 || s->image_height <= 0) return GS2D_ERR_INVALID;
	if (!in->viewmatrix || !in->projmatrix || !in->background) return GS2D_ERR_INVALID;
	// an empty map (P == 0) has empty per-surfel tensors, which arrive as NULL: nothing to check, the image is the
	// background (the reference short-circuits too, RAST/rasterize_points.cu:101)
	if (s->P == 0) return GS2D_OK;
	if (!in->means3D || !in->opacities) return GS2D_ERR_INVALID;
	if ((in->shs == nullptr) == (in->colors_precomp == nullptr)) return GS2D_ERR_INVALID;
	const bool have_sr = in->scales != nullptr && in->rotations != nullptr;
	if (have_sr == (in->transMat_precomp != nullptr)) return GS2D_ERR_INVALID;
	if (in->shs && (s->sh_coeffs <= 0 || !in->campos)) return GS2D_ERR_INVALID;
	return GS2D_OK;
}

static int debug_sync(const gs2d_settings& s, cudaStream_t st)
{
	if (!s.debug) return GS2D_OK;
	GS2D_CUDA_OK(cudaStreamSynchronize(st));
	GS2D_CUDA_OK(cudaGetLastError());
	return GS2D_OK;
}

#define GS2D_TRY(expr) do { int rc__ = (expr); if (rc__ != GS2D_OK) return rc__; } while (0)

}  // namespace gs2d

using namespace gs2d;

extern "C" {

size_t gs2d_geom_bytes(int32_t P, int32_t) { size_t b = 0; GeomState::carve(nullptr, (size_t)(P > 0 ? P : 0), &b); return b; }
size_t gs2d_image_bytes(int32_t W, int32_t H) { size_t b = 0; ImageState::carve(nullptr, (size_t)W, (size_t)H, &b); return b; }
size_t gs2d_binning_bytes(int64_t I, int32_t W, int32_t H) { size_t b = 0; BinningState::carve(nullptr, (size_t)(I > 0 ? I : 0), tiles_of(W, H), &b); return b; }
size_t gs2d_backward_scratch_bytes(int32_t, int64_t I, int64_t pairs, int32_t W, int32_t H)
{
	return composite_backward_scratch_bytes(I, pairs, (int64_t)tiles_of(W, H));
}
size_t gs2d_sort_scratch_bytes(int64_t n)
{
	const size_t m = (size_t)(n > 0 ? n : 1);
	size_t sb = 0;
	SortScratch::carve(nullptr, m, &sb);
	return sb + ALIGN * 4 + m * (sizeof(uint64_t) + sizeof(uint32_t)) + 2 * ALIGN;
}

int gs2d_forward_preprocess(const gs2d_settings* s, const gs2d_inputs* in, const gs2d_buffers* buffers,
                            int32_t* radii, int64_t* num_rendered_host, void* stream)
{
	GS2D_TRY(check_common(s, in, buffers));
	cudaStream_t st = (cudaStream_t)stream;
	if (num_rendered_host) *num_rendered_host = 0;
	if (s->P == 0) return GS2D_OK;
	// "not there yet": the copy enqueued below overwrites the whole int64 (count, 0), so a host that polls the pinned
	// cell sees a negative value until the count has landed
	if (num_rendered_host) *num_rendered_host = -1;
	if (!buffers->geom || !radii || !buffers->image) return GS2D_ERR_INVALID;
	if (buffers->geom_bytes < gs2d_geom_bytes(s->P, s->sh_coeffs)) return GS2D_ERR_CAPACITY;
	if (buffers->image_bytes < gs2d_image_bytes(s->image_width, s->image_height)) return GS2D_ERR_CAPACITY;
	GeomState g = GeomState::carve(buffers->geom, (size_t)s->P, nullptr);
	ImageState im = ImageState::carve(buffers->image, (size_t)s->image_width, (size_t)s->image_height, nullptr);
	{
		ProfScope ps(K_MEMSET, st);
		GS2D_CUDA_OK(cudaMemsetAsync(im.tile_count, 0, sizeof(uint32_t) * TC_STRIDE * tiles_of(s->image_width, s->image_height), st));
	}
	GS2D_TRY(launch_preprocess_forward(*s, *in, g, im.tile_count, radii, st));
	GS2D_TRY(debug_sync(*s, st));
	GS2D_TRY(launch_scan_tiles(g, s->P, st, (int)tiles_of(s->image_width, s->image_height), im.tile_count, im.ranges,
	                           im.tile_cursor));
	GS2D_TRY(debug_sync(*s, st));
	if (num_rendered_host) {
		// 8-byte read-back: counters[1] = total, counters[2] = 0 -> a whole int64
		GS2D_CUDA_OK(cudaMemcpyAsync(num_rendered_host, g.counters + 1, 2 * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
	}
	return GS2D_OK;
}

int gs2d_forward_scatter_early(const gs2d_settings* s, const gs2d_buffers* buffers, uint64_t* pairs_arena,
                               int64_t capacity, void* stream)
{
	if (!s || !buffers || !pairs_arena || capacity <= 0) return GS2D_ERR_INVALID;
	if (s->P <= 0) return GS2D_OK;
	if (!buffers->geom || !buffers->image) return GS2D_ERR_INVALID;
	GeomState g = GeomState::carve(buffers->geom, (size_t)s->P, nullptr);
	ImageState im = ImageState::carve(buffers->image, (size_t)s->image_width, (size_t)s->image_height, nullptr);
	return launch_scatter(*s, g, im, pairs_arena, true, capacity, (cudaStream_t)stream);
}

int gs2d_forward_render(const gs2d_settings* s, const gs2d_inputs* in, const gs2d_buffers* buffers,
                        int64_t I, float* out_color, float* out_others, int64_t* num_pairs_host, void* stream)
{
	return gs2d_forward_render_scattered(s, in, buffers, I, out_color, out_others, num_pairs_host, nullptr, stream);
}

int gs2d_forward_render_scattered(const gs2d_settings* s, const gs2d_inputs* in, const gs2d_buffers* buffers,
                                  int64_t I, float* out_color, float* out_others, int64_t* num_pairs_host,
                                  uint64_t* scattered_pairs, void* stream)
{
	GS2D_TRY(check_common(s, in, buffers));
	cudaStream_t st = (cudaStream_t)stream;
	if (!out_color || !out_others || !buffers->image) return GS2D_ERR_INVALID;
	if (I < 0 || I >= (1ll << 30)) return GS2D_ERR_INVALID;
	if (buffers->image_bytes < gs2d_image_bytes(s->image_width, s->image_height)) return GS2D_ERR_CAPACITY;
	ImageState im = ImageState::carve(buffers->image, (size_t)s->image_width, (size_t)s->image_height, nullptr);
	const unsigned gx = (s->image_width + TILE - 1) / TILE, gy = (s->image_height + TILE - 1) / TILE;
	if (s->P == 0 || I == 0) {
		// nothing to blend: ranges all empty, the compositor still writes every pixel (background)
		GS2D_CUDA_OK(cudaMemsetAsync(im.ranges, 0, sizeof(uint32_t) * 2 * (size_t)gx * gy, st));
		return launch_composite_forward(*s, *in, im, nullptr, nullptr, nullptr, out_color, out_others, st);
	}
	if (!buffers->geom || !buffers->binning) return GS2D_ERR_INVALID;
	if (buffers->binning_bytes < gs2d_binning_bytes(I, s->image_width, s->image_height)) return GS2D_ERR_CAPACITY;
	GeomState g = GeomState::carve(buffers->geom, (size_t)s->P, nullptr);
	BinningState b = BinningState::carve(buffers->binning, (size_t)I, tiles_of(s->image_width, s->image_height), nullptr);

	if (s->debug) {
		// the reference's unsorted (key, surfel) list, for gs2d_state_view only
		GS2D_TRY(launch_duplicate(*s, g, nullptr, b, I, st));
		GS2D_TRY(debug_sync(*s, st));
	}
	// scatter into tile segments (unless gs2d_forward_scatter_early already did), per-tile sort + instance gather +
	// footprint masks
	uint64_t* pairs = scattered_pairs ? scattered_pairs : b.keys_a;
	if (!scattered_pairs) GS2D_TRY(launch_scatter(*s, g, im, pairs, false, I, st));
	const int tiles = (int)(gx * gy);
	// The pair-slot layout of the backward pass (inclusive scan of the footprint popcounts) and its total are not needed
	// by the forward compositing: they run BESIDE it on the library's side stream, which takes the scan's ~13 us off the
	// critical path.  (pair_status / pair_counters were zeroed by the tile-0 block of tile_sort_kernel.)
	// The compositing of a tile needs only the sort of THAT tile, so the frame can also be cut into GS2D_FWD_PIPE runs
	// of tiles -- sorts back to back on the caller's stream, each run's compositing on the side stream as soon as its
	// sort has finished -- but that measured slower (see GS2D_FWD_PIPE above) and is off by default.  One stream when
	// debugging or timing kernel by kernel.
	cudaStream_t side = (s->debug || g_timing_on) ? nullptr : side_fork(st);
	const int runs = (side && GS2D_FWD_PIPE > 1 && tiles >= GS2D_FWD_PIPE * 296) ? GS2D_FWD_PIPE : 1;
	if (!side || runs == 1) {
		GS2D_TRY(launch_tile_sort(*s, g, im, b, pairs, I, st));
		GS2D_TRY(debug_sync(*s, st));
		if (side) GS2D_TRY(side_join(side, st));     // the side stream waits for the sort (it forked before it)
		cudaStream_t scan_st = side ? side : st;
		GS2D_TRY(launch_scan_u32(b.hit_count, b.pair_incl, I, b.pair_status, b.pair_counters, scan_st, false));
		if (num_pairs_host)   // 8 bytes: counters[1] = total, counters[2] = 0 -> a whole int64
			GS2D_CUDA_OK(cudaMemcpyAsync(num_pairs_host, b.pair_counters + 1, 2 * sizeof(uint32_t), cudaMemcpyDeviceToHost, scan_st));
		GS2D_TRY(debug_sync(*s, st));
		GS2D_TRY(launch_composite_forward(*s, *in, im, b.instances, b.hit_inst, b.hit_pix, out_color, out_others, st));
		if (side) GS2D_TRY(side_join(st, side));
		GS2D_TRY(debug_sync(*s, st));
		return GS2D_OK;
	}
	for (int r = 0; r < runs; r++) {
		const int t0 = (int)((long long)tiles * r / runs), t1 = (int)((long long)tiles * (r + 1) / runs);
		GS2D_TRY(launch_tile_sort(*s, g, im, b, pairs, I, st, t0, t1 - t0));
		GS2D_TRY(side_join(side, st));          // the side stream waits for this run's sort
		GS2D_TRY(launch_composite_forward(*s, *in, im, b.instances, b.hit_inst, b.hit_pix, out_color, out_others, side,
		                                  t0, t1 - t0));
	}
	GS2D_TRY(launch_scan_u32(b.hit_count, b.pair_incl, I, b.pair_status, b.pair_counters, st, false));
	if (num_pairs_host)
		GS2D_CUDA_OK(cudaMemcpyAsync(num_pairs_host, b.pair_counters + 1, 2 * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
	GS2D_TRY(side_join(st, side));              // the caller's stream waits for the last compositing
	return GS2D_OK;
}

int gs2d_forward(const gs2d_settings* s, const gs2d_inputs* in, const gs2d_buffers* buffers, int32_t* radii,
                 int64_t* num_rendered_io, float* out_color, float* out_others, void* stream)
{
	if (!num_rendered_io) return GS2D_ERR_INVALID;
	cudaStream_t st = (cudaStream_t)stream;
	const int64_t capacity = *num_rendered_io;
	int64_t count = 0;
	GS2D_TRY(gs2d_forward_preprocess(s, in, buffers, radii, nullptr, stream));
	if (s->P > 0) {
		GeomState g = GeomState::carve(buffers->geom, (size_t)s->P, nullptr);
		uint32_t c32 = 0;
		GS2D_CUDA_OK(cudaMemcpyAsync(&c32, g.counters + 1, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
		GS2D_CUDA_OK(cudaStreamSynchronize(st));
		count = c32;
	}
	*num_rendered_io = count;
	if (count > capacity) return GS2D_ERR_CAPACITY;
	return gs2d_forward_render(s, in, buffers, count, out_color, out_others, nullptr, stream);
}

int gs2d_forward_pair_count(const gs2d_settings* s, const gs2d_buffers* buffers, int64_t I, int64_t* num_pairs_host,
                            void* stream)
{
	if (!s || !buffers || !num_pairs_host) return GS2D_ERR_INVALID;
	if (I <= 0) { *num_pairs_host = 0; return GS2D_OK; }
	if (!buffers->binning) return GS2D_ERR_INVALID;
	BinningState b = BinningState::carve(buffers->binning, (size_t)I, tiles_of(s->image_width, s->image_height), nullptr);
	GS2D_CUDA_OK(cudaMemcpyAsync(num_pairs_host, b.pair_counters + 1, 2 * sizeof(uint32_t), cudaMemcpyDeviceToHost,
	                             (cudaStream_t)stream));
	return GS2D_OK;
}

static int backward_impl(const gs2d_settings* s, const gs2d_inputs* in, const gs2d_buffers* buffers,
                         const int32_t* radii, int64_t I, int64_t num_pairs, const float* dL_dout_color,
                         const float* dL_dout_others, const gs2d_grads* grads, float* mc_flat, void* scratch, void* stream)
{
	GS2D_TRY(check_common(s, in, buffers));
	cudaStream_t st = (cudaStream_t)stream;
	if (!grads || !dL_dout_color || !dL_dout_others) return GS2D_ERR_INVALID;
	if (s->P == 0) return GS2D_OK;
	if (!radii || !buffers->geom || !buffers->image) return GS2D_ERR_INVALID;
	if (!grads->dL_dmeans2D || !grads->dL_dtransMat) return GS2D_ERR_INVALID;
	if (!mc_flat && (!grads->dL_dcolors || !grads->dL_dopacity || !grads->dL_dmeans3D || !grads->dL_dscales ||
	                 !grads->dL_drotations))
		return GS2D_ERR_INVALID;
	// the in-switch reduction adds 16-byte quaternion gradients at 6 P floats into the buffer: P % 4 == 0, 16-byte base
	if (mc_flat && ((s->P & 3) || ((uintptr_t)mc_flat & 15) || in->shs)) return GS2D_ERR_INVALID;
	if (in->shs && !grads->dL_dsh) return GS2D_ERR_INVALID;
	if (num_pairs < 0) return GS2D_ERR_INVALID;
	GeomState g = GeomState::carve(buffers->geom, (size_t)s->P, nullptr);
	ImageState im = ImageState::carve(buffers->image, (size_t)s->image_width, (size_t)s->image_height, nullptr);
	float* rows = nullptr;
	if (I > 0) {
		if (!scratch || !buffers->binning) return GS2D_ERR_INVALID;
		BinningState b = BinningState::carve(buffers->binning, (size_t)I, tiles_of(s->image_width, s->image_height), nullptr);
		// no zero-fill: every instance row is written exactly once
		GS2D_TRY(launch_composite_backward(*s, *in, im, b, I, num_pairs, dL_dout_color, dL_dout_others, scratch, &rows, st));
		GS2D_TRY(debug_sync(*s, st));
	}
	GS2D_TRY(launch_preprocess_backward(*s, *in, g, radii, rows, *grads, st, mc_flat));
	GS2D_TRY(debug_sync(*s, st));
	return GS2D_OK;
}

int gs2d_backward(const gs2d_settings* s, const gs2d_inputs* in, const gs2d_buffers* buffers,
                  const int32_t* radii, int64_t I, int64_t num_pairs, const float* dL_dout_color,
                  const float* dL_dout_others, const gs2d_grads* grads, void* scratch, void* stream)
{
	return backward_impl(s, in, buffers, radii, I, num_pairs, dL_dout_color, dL_dout_others, grads, nullptr, scratch, stream);
}

int gs2d_backward_reduce(const gs2d_settings* s, const gs2d_inputs* in, const gs2d_buffers* buffers,
                         const int32_t* radii, int64_t I, int64_t num_pairs, const float* dL_dout_color,
                         const float* dL_dout_others, const gs2d_grads* grads, float* multicast_flat, void* scratch,
                         void* stream)
{
	if (!multicast_flat) return GS2D_ERR_INVALID;
	return backward_impl(s, in, buffers, radii, I, num_pairs, dL_dout_color, dL_dout_others, grads, multicast_flat, scratch,
	                     stream);
}

int gs2d_backward_overflowed(const gs2d_settings* s, const gs2d_buffers* buffers, int64_t I, int32_t* flag_host,
                             void* stream)
{
	if (!s || !buffers || !flag_host) return GS2D_ERR_INVALID;
	*flag_host = 0;
	if (I <= 0) return GS2D_OK;
	if (!buffers->binning) return GS2D_ERR_INVALID;
	cudaStream_t st = (cudaStream_t)stream;
	BinningState b = BinningState::carve(buffers->binning, (size_t)I, tiles_of(s->image_width, s->image_height), nullptr);
	GS2D_CUDA_OK(cudaMemcpyAsync(flag_host, b.pair_counters + 3, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
	GS2D_CUDA_OK(cudaStreamSynchronize(st));
	return GS2D_OK;
}

int gs2d_mark_visible(int32_t P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                      uint8_t* present, void* stream)
{
	if (P < 0) return GS2D_ERR_INVALID;
	if (P == 0) return GS2D_OK;
	if (!means3D || !viewmatrix || !present) return GS2D_ERR_INVALID;
	return launch_mark_visible(P, means3D, viewmatrix, projmatrix, present, (cudaStream_t)stream);
}

int gs2d_state_view_get(const gs2d_settings* s, const gs2d_buffers* buffers, int64_t I, gs2d_state_view* out)
{
	if (!s || !buffers || !out) return GS2D_ERR_INVALID;
	memset(out, 0, sizeof(*out));
	if (buffers->geom && s->P > 0) {
		GeomState g = GeomState::carve(buffers->geom, (size_t)s->P, nullptr);
		out->depths = g.depths; out->splat = g.splat; out->rect = g.rect; out->cull = g.cull;
		out->tiles_touched = g.tiles_touched; out->point_offsets = g.point_offsets;
	}
	if (buffers->binning && I > 0) {
		BinningState b = BinningState::carve(buffers->binning, (size_t)I, tiles_of(s->image_width, s->image_height), nullptr);
		out->keys_unsorted = b.keys_dbg; out->values_unsorted = b.vals_dbg;
		out->keys_sorted = b.keys_b; out->point_list = b.vals_b; out->instances = b.instances;
	}
	if (buffers->image) {
		ImageState im = ImageState::carve(buffers->image, (size_t)s->image_width, (size_t)s->image_height, nullptr);
		out->ranges = im.ranges; out->final_T = im.final_T; out->n_contrib = im.n_contrib;
		out->median_depth = im.median_depth; out->depth_std = im.depth_std;
	}
	return GS2D_OK;
}

int gs2d_footprint_audit(const gs2d_settings* s, const gs2d_buffers* buffers, int64_t I, uint64_t* counts_host,
                         void* stream)
{
	if (!s || !buffers || !counts_host) return GS2D_ERR_INVALID;
	counts_host[0] = counts_host[1] = counts_host[2] = 0;
	if (I <= 0) return GS2D_OK;
	if (!buffers->binning) return GS2D_ERR_INVALID;
	cudaStream_t st = (cudaStream_t)stream;
	BinningState b = BinningState::carve(buffers->binning, (size_t)I, tiles_of(s->image_width, s->image_height), nullptr);
	unsigned long long* counts = (unsigned long long*)(b.hit_inst + 8 * (size_t)I + 8);   // padding of hit_inst
	GS2D_CUDA_OK(cudaMemsetAsync(counts, 0, 3 * sizeof(unsigned long long), st));
	GS2D_TRY(launch_footprint_audit(*s, b.instances, b.keys_b, b.hit_inst, I, counts, st));
	GS2D_CUDA_OK(cudaMemcpyAsync(counts_host, counts, 3 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
	GS2D_CUDA_OK(cudaStreamSynchronize(st));
	return GS2D_OK;
}

int gs2d_sort_pairs(int64_t n, const uint64_t* keys_in, const uint32_t* values_in, uint64_t* keys_out,
                    uint32_t* values_out, int32_t end_bit, void* scratch, void* stream)
{
	if (n < 0 || n >= (1ll << 30) || end_bit <= 0 || end_bit > 64) return GS2D_ERR_INVALID;
	if (n == 0) return GS2D_OK;
	if (!keys_in || !values_in || !keys_out || !values_out || !scratch) return GS2D_ERR_INVALID;
	cudaStream_t st = (cudaStream_t)stream;
	Carver c(scratch);
	size_t sb = 0;
	SortScratch::carve(nullptr, (size_t)n, &sb);
	char* sbase = c.take<char>(sb);
	SortScratch ss = SortScratch::carve(sbase, (size_t)n, nullptr);
	uint64_t* tk = c.take<uint64_t>((size_t)n);
	uint32_t* tv = c.take<uint32_t>((size_t)n);
	bool in_b = false;
	GS2D_TRY(launch_sort_pairs(n, keys_in, values_in, keys_out, values_out, tk, tv, end_bit, ss, &in_b, st));
	if (!in_b) {
		GS2D_CUDA_OK(cudaMemcpyAsync(keys_out, tk, sizeof(uint64_t) * (size_t)n, cudaMemcpyDeviceToDevice, st));
		GS2D_CUDA_OK(cudaMemcpyAsync(values_out, tv, sizeof(uint32_t) * (size_t)n, cudaMemcpyDeviceToDevice, st));
	}
	return GS2D_OK;
}

int gs2d_timing_enable(int on)
{
	std::lock_guard<std::mutex> lk(g_prof_mu);
	g_timing_on = on != 0;
	return GS2D_OK;
}

int gs2d_num_kernels(void) { return (int)K_COUNT; }
const char* gs2d_kernel_name(int id) { return (id >= 0 && id < (int)K_COUNT) ? kKernelNames[id] : "?"; }

int gs2d_timing_collect(double* total_ms, int64_t* launches)
{
	std::lock_guard<std::mutex> lk(g_prof_mu);
	if (!total_ms || !launches) return GS2D_ERR_INVALID;
	for (int i = 0; i < (int)K_COUNT; i++) { total_ms[i] = 0.0; launches[i] = 0; }
	for (ProfRec& r : g_prof_done) {
		GS2D_CUDA_OK(cudaEventSynchronize(r.b));
		float ms = 0.f;
		GS2D_CUDA_OK(cudaEventElapsedTime(&ms, r.a, r.b));
		total_ms[r.id] += (double)ms;
		launches[r.id] += 1;
		g_prof_pool.push_back(r.a);
		g_prof_pool.push_back(r.b);
	}
	g_prof_done.clear();
	return GS2D_OK;
}

const char* gs2d_strerror(int status)
{
	switch (status) {
	case GS2D_OK: return "ok";
	case GS2D_ERR_INVALID: return "invalid argument";
	case GS2D_ERR_CUDA: return "CUDA runtime error";
	case GS2D_ERR_CAPACITY: return "workspace too small";
	case GS2D_ERR_UNSUPPORTED: return "unsupported";
	default: return "unknown gs2d status";
	}
}

int gs2d_last_cuda_error(void) { return g_last_cuda_error; }
int gs2d_abi_version(void) { return GS2D_ABI_VERSION; }
int64_t gs2d_launch_count(void) { return (int64_t)g_launches.load(); }

}  // extern "C"
