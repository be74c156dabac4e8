/* gs2d.h -- C ABI of the B200-native 2D-Gaussian-surfel rasterizer (libgs2d.so).
 *
 * Drop-in boundary for the hot path of vasabi-root/gaus-slam: the entry points below
 * are what a binding for the reference's `gaus_2dgs_rasterization._C` and
 * `simple_knn._C` extension modules would call.  File:line citations are relative to
 * the reference checkout (RAST = submodules/gaus_2dgs_rasterization, CR = RAST/cuda_rasterizer,
 * KNN = submodules/simple-knn).
 *
 *   gs2d_forward_*  replace  RasterizeGaussiansCUDA            RAST/rasterize_points.cu:39-138
 *                            CudaRasterizer::Rasterizer::forward  CR/rasterizer_impl.cu:201-350
 *   gs2d_backward   replaces RasterizeGaussiansBackwardCUDA    RAST/rasterize_points.cu:140-239
 *                            CudaRasterizer::Rasterizer::backward CR/rasterizer_impl.cu:354-460
 *   gs2d_mark_visible replaces markVisible                     RAST/rasterize_points.cu:241-260
 *   gs2d_dist2      replaces distCUDA2 / SimpleKNN::knn        KNN/spatial.cu:15-26, KNN/simple_knn.cu:186-222
 *
 * Rules of the ABI
 *   - plain C: pointers, sizes, scalars; no torch types, no exceptions, no state that a result depends on: calls on
 *     different streams / devices / threads are independent.  Process-wide are only diagnostics (the launch counter,
 *     the opt-in kernel timing of gs2d_timing_*), the one-time per-device kernel attribute opt-ins, one internal
 *     side stream per device (the forward pass forks one scan onto it and joins it again), and the static scratch of
 *     gs2d_sort_pairs_static.
 *   - every pointer is a DEVICE pointer owned by the caller unless its name ends in
 *     `_host`; nothing is allocated or freed behind the caller's back.
 *   - every function that launches work takes the CUDA stream to launch on
 *     (a `cudaStream_t` passed as void*; NULL = legacy default stream).
 *   - return value: 0 (GS2D_OK) or a negative gs2d_status; gs2d_strerror() names it.
 *   - there is no CPU fallback: without a CUDA device every launch fails with
 *     GS2D_ERR_CUDA.
 */
#ifndef GS2D_H_INCLUDED
#define GS2D_H_INCLUDED

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* 3: gs2d_adam_params gained group_step[5] (and negative lr = "no gradient"); gs2d_map_compact_* added */
#define GS2D_ABI_VERSION 4

typedef enum gs2d_status {
	GS2D_OK = 0,
	GS2D_ERR_INVALID = -1,   /* bad argument (null pointer, negative size, ...)          */
	GS2D_ERR_CUDA = -2,      /* a CUDA runtime call failed; see gs2d_last_cuda_error()    */
	GS2D_ERR_CAPACITY = -3,  /* binning workspace smaller than num_rendered needs         */
	GS2D_ERR_UNSUPPORTED = -4
} gs2d_status;

/* Tile geometry is part of the contract (CR/config.h:15-17): keys, sort order and tile
 * ranges are bit-exact with the reference only for 16x16 tiles. */
#define GS2D_TILE 16
#define GS2D_NUM_CHANNELS 3

/* Scalar settings: the non-tensor fields of GaussianRasterizationSettings
 * (RAST/gaus_2dgs_rasterization/__init__.py:163-176) plus the sizes. */
typedef struct gs2d_settings {
	int32_t P;              /* number of surfels                                         */
	int32_t image_width;
	int32_t image_height;
	int32_t sh_degree;      /* D  (active SH degree)                                      */
	int32_t sh_coeffs;      /* M  (coefficients per surfel in `shs`; 0 when shs == NULL)  */
	int32_t use_sa;         /* surface-aware depth (GauS-SLAM) vs 2DGS distortion         */
	int32_t prefiltered;
	int32_t debug;          /* 1: synchronise + check after every stage (CHECK_CUDA)      */
	float tanfovx;
	float tanfovy;
	float scale_modifier;
	float reserved_;
} gs2d_settings;

/* Inputs shared by forward and backward.  A NULL pointer selects the same branch an
 * empty tensor selects in the reference (CR/forward.cu:194, :241):
 * exactly one of {shs, colors_precomp}; exactly one of {(scales, rotations), transMat_precomp}. */
typedef struct gs2d_inputs {
	const float* background;        /* [3]                                        */
	const float* means3D;           /* [P,3]                                      */
	const float* shs;               /* [P,M,3] or NULL                            */
	const float* colors_precomp;    /* [P,3]   or NULL                            */
	const float* opacities;         /* [P,1]                                      */
	const float* scales;            /* [P,2]   or NULL                            */
	const float* rotations;         /* [P,4] (w,x,y,z) or NULL                    */
	const float* transMat_precomp;  /* [P,9]   or NULL                            */
	const float* viewmatrix;        /* [16], memory = column-major V (w2c^T rows) */
	const float* projmatrix;        /* [16], memory = column-major P*V            */
	const float* campos;            /* [3]                                        */
} gs2d_inputs;

/* The three opaque state buffers kept between forward and backward: the counterparts of
 * geomBuffer / binningBuffer / imgBuffer (RAST/rasterize_points.cu:93-98).  Their layout is
 * private to this library; gs2d_state_view() exposes the parity-relevant arrays. */
typedef struct gs2d_buffers {
	void* geom;    size_t geom_bytes;
	void* binning; size_t binning_bytes;
	void* image;   size_t image_bytes;
} gs2d_buffers;

/* Sizes of the state buffers.  `num_rendered` may be an upper bound (capacity). */
size_t gs2d_geom_bytes(int32_t P, int32_t sh_coeffs);
size_t gs2d_image_bytes(int32_t image_width, int32_t image_height);
size_t gs2d_binning_bytes(int64_t num_rendered, int32_t image_width, int32_t image_height);

/* Stage 1 of forward: per-surfel preprocess (world->ray-splat transform, screen bounds,
 * tile counts; CR/forward.cu:150-253) and the inclusive scan of tile counts
 * (CR/rasterizer_impl.cu:283).  Writes `radii` [P] int32 (zero for culled surfels).
 * Also counts the instances of every tile and turns the counts into the tile ranges
 * (the output of identifyTileRanges, CR/rasterizer_impl.cu:116-138) for the tile-local sort.
 * Enqueues an async copy of num_rendered to `num_rendered_host` (pinned host int64) when
 * that pointer is not NULL: the cell is set to -1 by the call and becomes the count (>= 0) when the copy has
 * landed, so the host may poll it instead of synchronising (0 at once for an empty map). Needs buffers->geom (gs2d_geom_bytes) and buffers->image
 * (gs2d_image_bytes). */
int gs2d_forward_preprocess(const gs2d_settings* s, const gs2d_inputs* in,
                            const gs2d_buffers* buffers, int32_t* radii,
                            int64_t* num_rendered_host, void* stream);

/* Stage 2 of forward: binning -- the sorted per-tile instance lists that duplicateWithKeys +
 * cub::DeviceRadixSort::SortPairs + identifyTileRanges produce (CR/rasterizer_impl.cu:295-323), here by a
 * scatter into tile segments and a sort inside each tile, same order bit for bit -- and per-tile
 * compositing (CR/rasterizer_impl.cu:325-347, CR/forward.cu:258-467).  `num_rendered` is the exact
 * instance count read back after stage 1; buffers->binning must hold
 * gs2d_binning_bytes(num_rendered, W, H) and buffers->image gs2d_image_bytes(W, H).
 * out_color [3,H,W], out_others [7,H,W] are fully written (no pre-zeroing needed).
 * When `num_pairs_host` (pinned host int64) is not NULL the pair-slot count (see gs2d_forward_pair_count) is
 * copied into it asynchronously as soon as it is known, i.e. before the compositing kernel runs. */
int gs2d_forward_render(const gs2d_settings* s, const gs2d_inputs* in,
                        const gs2d_buffers* buffers, int64_t num_rendered,
                        float* out_color, float* out_others, int64_t* num_pairs_host, void* stream);

/* Optional overlap of the host read-back with device work: between stage 1 and the host's wait for
 * num_rendered, gs2d_forward_scatter_early launches the scatter of stage 2 into a caller-owned arena of
 * `capacity` 8-byte words (sized from earlier frames).  The kernel reads num_rendered on the device and does
 * nothing when it exceeds the capacity.  Once the host knows num_rendered it calls
 * gs2d_forward_render_scattered with the arena when num_rendered <= capacity (stage 2 then skips its own
 * scatter; the arena must stay untouched until that call's kernels have run -- stream order is enough), else
 * gs2d_forward_render as usual.  (No counterpart in the reference: it blocks at CR/rasterizer_impl.cu:287 with
 * the device idle.) */
int gs2d_forward_scatter_early(const gs2d_settings* s, const gs2d_buffers* buffers, uint64_t* pairs_arena,
                               int64_t capacity, void* stream);
int gs2d_forward_render_scattered(const gs2d_settings* s, const gs2d_inputs* in, const gs2d_buffers* buffers,
                                  int64_t num_rendered, float* out_color, float* out_others,
                                  int64_t* num_pairs_host, uint64_t* scattered_pairs, void* stream);

/* Convenience: stage 1, stream synchronise, stage 2 -- the reference's control flow
 * (CR/rasterizer_impl.cu:287 blocks on the same 4-byte read-back).  On entry
 * *num_rendered_io is the binning capacity (in instances) the caller provisioned; on
 * return it is the actual count.  Returns GS2D_ERR_CAPACITY (after stage 1 only) when
 * the capacity is too small: re-provision and call gs2d_forward_render. */
int gs2d_forward(const gs2d_settings* s, const gs2d_inputs* in, const gs2d_buffers* buffers,
                 int32_t* radii, int64_t* num_rendered_io,
                 float* out_color, float* out_others, void* stream);

/* Gradients, all caller-allocated, fully written (no pre-zeroing needed), in the layout of
 * the reference's return tuple (RAST/rasterize_points.cu:192-200, :238). */
typedef struct gs2d_grads {
	float* dL_dmeans2D;    /* [P,3]  densification proxy (CR/backward.cu:660-663)  */
	float* dL_dcolors;     /* [P,3]                                                */
	float* dL_dopacity;    /* [P,1]                                                */
	float* dL_dmeans3D;    /* [P,3]                                                */
	float* dL_dtransMat;   /* [P,9]                                                */
	float* dL_dsh;         /* [P,M,3] or NULL when M == 0                          */
	float* dL_dscales;     /* [P,2]                                                */
	float* dL_drotations;  /* [P,4]                                                */
} gs2d_grads;

/* Number of (pixel, instance) pair slots of the last forward pass (sum of the footprint-mask popcounts): the
 * backward pass keeps 16 bytes per pair in its scratch.  Enqueues an async 4-byte copy into `num_pairs_host`
 * (pinned host int64; all 8 bytes are written); valid once the stream has passed gs2d_forward_render. */
int gs2d_forward_pair_count(const gs2d_settings* s, const gs2d_buffers* buffers, int64_t num_rendered,
                            int64_t* num_pairs_host, void* stream);

/* Backward (CR/backward.cu:143-463 render, :466-664 preprocess).  `radii` and the three
 * buffers are those of the matching forward; dL_dout_color [3,H,W], dL_dout_others [7,H,W].
 * `scratch` must hold gs2d_backward_scratch_bytes(P, num_rendered, num_pairs, W, H) bytes.  `num_pairs` must be the
 * count of THAT forward (gs2d_forward_render's num_pairs_host, or gs2d_forward_pair_count).  A smaller value cannot
 * corrupt memory: a tile whose pair slots would end beyond `num_pairs` writes nothing (its gradient rows stay
 * undefined) and sets the overflow word of the binning state, which gs2d_backward_overflowed() reads back. */
size_t gs2d_backward_scratch_bytes(int32_t P, int64_t num_rendered, int64_t num_pairs, int32_t image_width,
                                   int32_t image_height);
int gs2d_backward(const gs2d_settings* s, const gs2d_inputs* in, const gs2d_buffers* buffers,
                  const int32_t* radii, int64_t num_rendered, int64_t num_pairs,
                  const float* dL_dout_color, const float* dL_dout_others,
                  const gs2d_grads* grads, void* scratch, void* stream);
/* gs2d_backward fused with the all-reduce of the view-sharded mapping step (new; the reference is single-GPU,
 * SURVEY.md 8e).  The five map-parameter gradients -- dL/d(means3D, opacity, scales, rotations, colors) -- are not
 * stored: the backward-preprocess epilogue ADDS them with multimem.red, i.e. inside the NVSwitch, into every rank's copy
 * of a flat [13 * P] fp32 buffer laid out group by group (xyz 3P | opacity P | scale 2P | rotation 4P | rgb 3P).
 * `multicast_flat` is the NVLS multicast address of that buffer (symmetric memory bound to a multicast object; same
 * offset on every rank; 16-byte aligned), P % 4 == 0, colours precomputed (no SH).  Protocol: every rank zeroes its
 * copy, a cross-GPU barrier, every rank calls this for its own view, a second cross-GPU barrier -- then every rank's
 * copy holds the SUM over the ranks; no separate all-reduce pass exists.  The five map-group pointers of `grads` are
 * ignored (may be NULL); dL_dmeans2D, dL_dtransMat (and nothing else) are written locally as in gs2d_backward. */
int gs2d_backward_reduce(const gs2d_settings* s, const gs2d_inputs* in, const gs2d_buffers* buffers,
                         const int32_t* radii, int64_t num_rendered, int64_t num_pairs, const float* dL_dout_color,
                         const float* dL_dout_others, const gs2d_grads* grads, float* multicast_flat, void* scratch,
                         void* stream);

/* 1 in *flag_host (pinned or pageable host int32) if the last gs2d_backward on these buffers skipped tiles because
 * num_pairs was too small, else 0.  Synchronises the stream. */
int gs2d_backward_overflowed(const gs2d_settings* s, const gs2d_buffers* buffers, int64_t num_rendered,
                             int32_t* flag_host, void* stream);

/* present[i] = view-space z of means3D[i] > 0.2 (CR/rasterizer_impl.cu:54-66, CR/auxiliary.h:184-209).
 * `present` is one byte per surfel (torch.bool). */
int gs2d_mark_visible(int32_t P, const float* means3D, const float* viewmatrix,
                      const float* projmatrix, uint8_t* present, void* stream);

/* distCUDA2: mean squared distance to the 3 nearest other points (KNN/simple_knn.cu:148-222).
 * points [P,3]; mean_dists [P]; scratch of gs2d_dist2_scratch_bytes(P). */
size_t gs2d_dist2_scratch_bytes(int32_t P);
int gs2d_dist2(int32_t P, const float* points, float* mean_dists, void* scratch, void* stream);

/* Parity accessors: device pointers into the opaque buffers (valid after the stage that
 * writes them).  Tests use these to compare binning bit-for-bit with the reference. */
typedef struct gs2d_state_view {
	/* per surfel (stage 1) */
	const float*    depths;           /* [P]   view-space z                          */
	const float*    splat;            /* [P,20] Tu,x | Tv,y | Tw,opacity | n,- | rgb,- */
	const uint16_t* rect;             /* [P,4] tile rect x0,y0,x1,y1                  */
	const float*    cull;             /* [P,4] contribution bounds x0,y0,x1,y1 (pixels) */
	const uint32_t* tiles_touched;    /* [P]                                          */
	const uint32_t* point_offsets;    /* [P]   inclusive scan                         */
	/* per instance (stage 2) */
	const uint64_t* keys_unsorted;    /* [I]   tile<<32 | float bits of depth         */
	const uint32_t* values_unsorted;  /* [I]   surfel index                           */
	const uint64_t* keys_sorted;      /* [I]                                          */
	const uint32_t* point_list;       /* [I]   sorted surfel indices                  */
	const float*    instances;        /* [I,20] per-instance records in sorted order   */
	/* per tile / per pixel */
	const uint32_t* ranges;           /* [tiles,2]                                    */
	const float*    final_T;          /* [3,H,W] T, M1, M2                            */
	const uint32_t* n_contrib;        /* [2,H,W] last contributor, median contributor */
	const float*    median_depth;     /* [H,W]                                        */
	const float*    depth_std;        /* [H,W]                                        */
} gs2d_state_view;

int gs2d_state_view_get(const gs2d_settings* s, const gs2d_buffers* buffers,
                        int64_t num_rendered, gs2d_state_view* out);

/* Parity accessor for the footprint stage (csrc/footprint.cu): re-evaluates the reference's per-pixel
 * contribution tests for every (instance, tile pixel) pair of the last forward and compares them with the
 * footprint masks.  counts_host[0] = pairs that pass the tests, [1] = mask bits set, [2] = passing pairs
 * whose bit is NOT set (must be 0: the masks are a superset).  Synchronises the stream. */
int gs2d_footprint_audit(const gs2d_settings* s, const gs2d_buffers* buffers, int64_t num_rendered,
                         uint64_t* counts_host, void* stream);

/* Stand-alone stable LSD radix sort of (u64 key, u32 value) pairs on bits [0, end_bit) -- sorted in whole 8-bit
 * digits, i.e. on bits [0, 8 * ceil(end_bit / 8)): keys must have no bit set at or above end_bit (the rasterizer's
 * and the Morton keys do not) for the result to equal a sort on [0, end_bit) --:
 * the hand-written onesweep that replaces cub::DeviceRadixSort::SortPairs
 * (CR/rasterizer_impl.cu:309-314).  scratch: gs2d_sort_scratch_bytes(n). */
size_t gs2d_sort_scratch_bytes(int64_t n);
int gs2d_sort_pairs(int64_t n, const uint64_t* keys_in, const uint32_t* values_in,
                    uint64_t* keys_out, uint32_t* values_out, int32_t end_bit,
                    void* scratch, void* stream);

/* gs2d_sort_pairs without a scratch argument and without any allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  The scratch is a static device array inside the library
 * (GS2D_SORT_STATIC_SCRATCH_BYTES, held by every GPU that loads the library until exit); the call issues only kernel
 * launches, cudaMemsetAsync and cudaMemcpyAsync on `stream`.  n > GS2D_SORT_STATIC_MAX_PAIRS returns
 * GS2D_ERR_CAPACITY ("workspace too small") and launches nothing.  All calls on a device share the array: what must
 * start at zero is cleared on the stream at the start of every call, and calls that overlap on the GPU (different
 * streams without ordering between them) race -- serialise them.  Same result as gs2d_sort_pairs, bit for bit. */
#define GS2D_SORT_STATIC_MAX_PAIRS (1 << 18)
#define GS2D_SORT_STATIC_SCRATCH_BYTES (4718592u) /* 4.5 MiB >= gs2d_sort_scratch_bytes(GS2D_SORT_STATIC_MAX_PAIRS) */
int gs2d_sort_pairs_static(int64_t n, const uint64_t* keys_in, const uint32_t* values_in,
                           uint64_t* keys_out, uint32_t* values_out, int32_t end_bit, void* stream);

/* ---- "next" row f2: fused mapping loss + image gradients --------------------------------
 * The mapping branch of get_loss (slam/Loss.py:22-25, :36, :51-63) on the rasterizer's outputs, including the
 * weight normalisation of the depth that Renderer_* applies first (render/__init__.py:46-49):
 *   depth = allmap[0] / (allmap[1] + eps), zeroed outside [depth_near, depth_far]; NaN/inf -> nan_to_num(x,0,0);
 *   mask = gt_depth > 1e-5 and depth > 1e-5;
 *   loss = w_color * mean_mask,3ch |color - gt_color| + w_depth * mean_mask |depth - gt_depth|
 *        + w_dist * mean_mask allmap[6].
 * Defaults of every shipped config: use_edge_growth = use_normal_loss = enable_exposure = False (other
 * branches: not built).  color [3,H,W], allmap [7,H,W], gt_depth [H,W] (or [1,H,W]); gt_color [3,H,W], or
 * [H,W,3] (the reference's layout) when gt_color_hwc != 0.  `scratch`: gs2d_mapping_loss_scratch_bytes bytes,
 * ZERO-INITIALISED once by the caller and then reused call after call (forward leaves the mask count there for
 * backward).  Forward writes the scalar loss to loss_out (device float).  Backward writes dL/dcolor [3,H,W] and
 * dL/dallmap [7,H,W] entirely (channels 2..5 are zeros), scaled by *grad_loss (device float; NULL = 1). */
typedef struct gs2d_loss_params {
	float w_color, w_depth, w_dist;   /* configs: 0.5, 1.0, 0.1                                  */
	float eps;                        /* render.eps = 1e-6                                        */
	float depth_near, depth_far;      /* render.depth_near = 1e-2, render.depth_far = 1e2         */
	int32_t tracking;                 /* 1: the tracking branch (slam/Loss.py:35-47, ignore_outliners = False):
	                                   * mask also needs nan_to_num(alpha) > silmask_th; loss = w_color * SUM |dc|
	                                   * + w_depth * SUM |dd| over the mask (sums, not means; no distortion term) */
	float silmask_th;                 /* loss.silmask_th = 0.9 (tracking only)                    */
	float edge_thres;                 /* ABI v4.  < 0: densify.use_edge_growth = False (every shipped config) -- colour and
	                                   * distortion are averaged over the depth mask.  >= 0 (densify.edge_thres = 0.4):
	                                   * over nan_to_num(alpha, 0, 0) > edge_thres instead (slam/Loss.py:53, mapping only) */
} gs2d_loss_params;

size_t gs2d_mapping_loss_scratch_bytes(int32_t image_width, int32_t image_height);
int gs2d_mapping_loss_forward(int32_t image_width, int32_t image_height, const float* color, const float* allmap,
                              const float* gt_color, const float* gt_depth, int32_t gt_color_hwc,
                              const gs2d_loss_params* params, void* scratch, float* loss_out, void* stream);
int gs2d_mapping_loss_backward(int32_t image_width, int32_t image_height, const float* color, const float* allmap,
                               const float* gt_color, const float* gt_depth, int32_t gt_color_hwc,
                               const gs2d_loss_params* params, const void* scratch, const float* grad_loss,
                               float* dL_dcolor, float* dL_dallmap, void* stream);

/* ---- "next" row f3: fused map optimiser step -------------------------------------------------
 * The reference's map optimiser: five raw parameter tensors (scene/Gaussians.py:106-114), rendered through
 * sigmoid / exp / F.normalize (get_render_params, :336-358) and stepped by torch.optim.Adam(eps=1e-15) with one
 * learning rate per group (training_setup, :124-137).  gs2d_map_adam_step is one launch for: the chain rule of
 * the three activations, the Adam update of all five groups (torch/optim/adam.py, no amsgrad / weight decay)
 * and the activations of the updated parameters for the next forward pass.
 * Group order everywhere: xyz, opacity, scaling, rotation, rgb.  `flat_grad` is the gradient w.r.t. the RENDER
 * parameters in the flat [13 * P] layout of the view-sharded all-reduce (gs2d_b200/mapping.py):
 * means3D 3P | opacities P | scales 2P | rotations 4P | colors 3P.  scale_cols = 2 (anisotropic, every shipped
 * config) or 1 (isotropic: _scaling is [P,1] and the render scales are exp(s) in both columns). */
typedef struct gs2d_map_state {
	float* xyz;             /* [P,3]  _xyz                                   */
	float* opacity;         /* [P,1]  _opacity (logit)                       */
	float* scaling;         /* [P,scale_cols] _scaling (log)                 */
	float* rotation;        /* [P,4]  _rotation (unnormalised, w first)      */
	float* rgb;             /* [P,3]  _rgb                                   */
	float* exp_avg[5];      /* Adam first moments, shapes as above           */
	float* exp_avg_sq[5];   /* Adam second moments                           */
} gs2d_map_state;

typedef struct gs2d_adam_params {
	double lr[5];           /* xyz_lr, opacity_lr, scaling_lr, rotation_lr, rgb_lr; a NEGATIVE value = "this group
	                         * has no gradient in this step": parameter, moments and activation are left untouched,
	                         * as torch.optim.Adam does for a parameter whose .grad is None                       */
	double beta1, beta2;    /* 0.9, 0.999                                                  */
	double eps;             /* 1e-15 (scene/Gaussians.py:137)                              */
	int64_t step;           /* 1 for the first update (torch increments before using it)   */
	int64_t group_step[5];  /* per-group step counts (torch keeps one per parameter: a group skipped in earlier
	                         * steps is behind the others); 0 = use `step`                 */
} gs2d_adam_params;

/* Render parameters of the current raw parameters: opacities [P,1] = sigmoid, scales [P,2] = exp,
 * rotations [P,4] = q / max(|q|, 1e-12).  (means3D and colors are the raw tensors themselves.)  The moments of
 * `ms` are not touched and may be NULL. */
int gs2d_map_activate(int32_t P, int32_t scale_cols, const gs2d_map_state* ms, float* opacities, float* scales,
                      float* rotations, void* stream);
/* One optimiser step in place; opacities / scales / rotations (all three or none) receive the activations of the
 * updated parameters. */
int gs2d_map_adam_step(int32_t P, int32_t scale_cols, const gs2d_map_state* ms, const float* flat_grad,
                       const gs2d_adam_params* params, float* opacities, float* scales, float* rotations,
                       void* stream);

/* Stream compaction of the map state: what Gaussians.prune_optimizer(mask) does to the five parameter tensors and
 * their Adam moments (scene/Gaussians.py:143-160; remove_gaussians_from_mask :280-297), in two steps so that the
 * caller can size the destination exactly:
 *   plan:  keep [P] bytes (torch.bool storage: 0 / non-zero) -> scratch (gs2d_map_compact_scratch_bytes(P) bytes,
 *          need not be initialised): int32 scratch[0] = number of kept surfels, the rest are per-block offsets;
 *   move:  every array of `src` -> the same array of `dst` (row r of dst = the r-th kept row of src, order
 *          preserved as in tensor[mask]).  Out of place.  A moment pointer may be NULL on BOTH sides (optimiser that
 *          has not stepped); parameters are required.  `dst` arrays hold at least scratch[0] rows.
 * The Adam step counter is not touched (the reference keeps the state dict's "step"). */
size_t gs2d_map_compact_scratch_bytes(int32_t P);
int gs2d_map_compact_plan(int32_t P, const uint8_t* keep, void* scratch, void* stream);
int gs2d_map_compact_move(int32_t P, int32_t scale_cols, const uint8_t* keep, const gs2d_map_state* src,
                          const gs2d_map_state* dst, const void* scratch, void* stream);

/* ---- "next" row f1: fused pose transform + pose-gradient reduction ---------------------------
 * Renderer_tracking / Renderer_BA move the map into the camera frame before rendering with an identity camera
 * (render/__init__.py:35-36, :97-98):
 *   means3D'   = (w2c[:3,:3] @ means3D.T + w2c[:3,3:]).T
 *   rotations' = quaternion_multiply(matrix_to_quaternion(w2c[None,:3,:3]), rotations)        (no gradient)
 * The pose is given EITHER as the optimiser's own scalars (scene/Frame.py:66-67, :87-96: cam_rot [4] unnormalised
 * quaternion, real first; cam_trans [3]; w2c = [quaternion_to_matrix(normalize(cam_rot)) | cam_trans]) OR as a
 * row-major 4x4 `w2c` (fix_w2c / LocalMap.get_frame_w2c) -- the other form's pointers NULL.  All device pointers.
 * rotations / rotations_out may both be NULL (means only); when given they must be 16-byte aligned.
 * Backward: g = dL/dmeans3D' [P,3] -> dL/dcam_rot [4] + dL/dcam_trans [3] (quaternion form) or dL/dw2c [16]
 * (matrix form; last row zero), and, when dL_dxyz is not NULL, dL/dmeans3D = R^T g [P,3] (bundle adjustment).
 * `scratch`: gs2d_pose_transform_scratch_bytes() bytes, ZERO-INITIALISED once by the caller, then reusable. */
int gs2d_pose_transform_forward(int32_t P, const float* xyz, const float* rotations, const float* cam_rot,
                                const float* cam_trans, const float* w2c, float* means3D_out, float* rotations_out,
                                void* stream);
size_t gs2d_pose_transform_scratch_bytes(void);
int gs2d_pose_transform_backward(int32_t P, const float* xyz, const float* cam_rot, const float* cam_trans,
                                 const float* w2c, const float* dL_dmeans3D, void* scratch, float* dL_dcam_rot,
                                 float* dL_dcam_trans, float* dL_dw2c, float* dL_dxyz, void* stream);

/* Adam step of the tracker's seven pose scalars -- scene/Frame.py:66-73: torch.optim.Adam over `cam_rot` [4] and
 * `cam_trans` [3], one learning rate each (cameras.*_lr), betas (0.7, 0.99), stepped once per tracking iteration
 * (slam/Frontend.py:92) -- in torch's operation order, as ONE launch.  `state` = 14 device floats, exp_avg[7] |
 * exp_avg_sq[7] (cam_rot first), zero before the first step.  A NULL gradient = that parameter has no gradient in
 * this step: it, its moments and its step counter stay untouched, like torch.  step[k] = 1 for a parameter's
 * first update. */
typedef struct gs2d_pose_adam_params {
	double lr[2];           /* cam_rot, cam_trans                      */
	double beta1, beta2;    /* 0.7, 0.99 (scene/Frame.py:70-73)        */
	double eps;             /* 1e-8                                    */
	int64_t step[2];        /* per-parameter step counts, 1-based      */
} gs2d_pose_adam_params;
int gs2d_pose_adam_step(float* cam_rot, float* cam_trans, const float* dL_dcam_rot, const float* dL_dcam_trans,
                        float* state, const gs2d_pose_adam_params* params, void* stream);

/* ---- multi-GPU (new; the reference is single-GPU) -----------------------------------------
 * In-switch all-reduce (sum, fp32) of a buffer that lives in symmetric memory bound to an NVLS multicast
 * object: `multicast_ptr` is the multicast address of the buffer (same offset on every rank), n_floats a
 * multiple of 4.  Rank r reduces and re-broadcasts the r-th of `world` slices with multimem.ld_reduce /
 * multimem.st.  The caller must separate the launch from the producers and consumers of the buffer on ALL
 * ranks with cross-GPU barriers (before: every rank's data is written; after: every slice is stored).
 * max_blocks <= 0 picks a default grid. */
int gs2d_allreduce_multimem(void* multicast_ptr, int64_t n_floats, int32_t rank, int32_t world,
                            int32_t max_blocks, void* stream);

/* Optional per-kernel timing for bench.py's roofline line: while enabled, every kernel launch
 * (and memset) of this library is bracketed by CUDA events on the launching stream.
 * gs2d_timing_collect() waits for the recorded events, writes the summed device time (ms) and
 * the launch count per kernel id [0, gs2d_num_kernels()) and clears the record. */
int gs2d_timing_enable(int on);
int gs2d_num_kernels(void);
const char* gs2d_kernel_name(int id);
int gs2d_timing_collect(double* total_ms, int64_t* launches);

const char* gs2d_strerror(int status);
int gs2d_last_cuda_error(void);          /* cudaError_t of the last failing call in this thread */
int gs2d_abi_version(void);
/* Number of kernels this library has launched in this process (for bench.py gpu_launches). */
int64_t gs2d_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* GS2D_H_INCLUDED */
