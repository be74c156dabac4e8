// oracle/ref_shim.cu -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// A C-ABI wrapper (written for this repo) around the UNMODIFIED reference CUDA
// rasterizer and simple-knn.  It is compiled by oracle/build.py together with
// the reference's own source files *where they lie* under /root/reference
// (nothing is copied into this repo); the output goes to oracle/_ref/ only.
//
// It calls exactly the entry points the reference's torch glue calls:
//   CudaRasterizer::Rasterizer::forward   (RAST/rasterize_points.cu:109,  CR/rasterizer_impl.cu:201)
//   CudaRasterizer::Rasterizer::backward  (RAST/rasterize_points.cu:204,  CR/rasterizer_impl.cu:354)
//   CudaRasterizer::Rasterizer::markVisible (RAST/rasterize_points.cu:252, CR/rasterizer_impl.cu:141)
//   SimpleKNN::knn                         (KNN/spatial.cu:23, KNN/simple_knn.cu:186)
// and reproduces what that glue does around them: zero-filled outputs
// (rasterize_points.cu:86-89), the three growable byte buffers (:31-37, :93-98)
// and the nine zero-filled gradient tensors (:192-200).  All work is issued on
// the legacy default stream, as the reference does (CR/forward.cu:489 etc.).
//
// Only tests/, __graft_entry__.smoke() and bench.py may load the resulting
// library (as the parity checker and as the timed "reference" arm).

#include <cstdint>
#include <cstdio>
#include <cstring>
#include <functional>
#include <cuda_runtime.h>

#include "rasterizer.h"        // CR/rasterizer.h (reference, via -I)
#include "rasterizer_impl.h"   // CR/rasterizer_impl.h: GeometryState / BinningState / ImageState
#include "simple_knn.h"        // KNN/simple_knn.h (reference, via -I)

namespace {

struct GrowBuf {
	char* ptr = nullptr;
	size_t cap = 0;
	size_t size = 0;
	char* resize(size_t n) {
		if (n > cap) {
			if (ptr) cudaFree(ptr);
			size_t want = n + n / 4 + 256;
			cudaMalloc(&ptr, want);
			cap = want;
		}
		size = n;
		return ptr;
	}
	void release() { if (ptr) cudaFree(ptr); ptr = nullptr; cap = size = 0; }
};

struct RefCtx {
	GrowBuf geom, binning, img, out_mask;
	// gradient scratch that the torch glue allocates but does not return
	GrowBuf dL_dnormal;
	int P = 0, W = 0, H = 0, R = 0;
};

}  // namespace

extern "C" {

// Must mirror oracle/oracle.py::RefStatePtrs.
struct RefStatePtrs {
	// GeometryState (CR/rasterizer_impl.h:29-44)
	const float* depths;
	const int* internal_radii;
	const float* means2D;          // float2[P]
	const float* transMat;         // float[9P]
	const float* normal_opacity;   // float4[P]
	const uint32_t* tiles_touched;
	const uint32_t* point_offsets;
	// BinningState (:57-67)
	const uint64_t* point_list_keys_unsorted;
	const uint64_t* point_list_keys;
	const uint32_t* point_list_unsorted;
	const uint32_t* point_list;
	// ImageState (:46-55)
	const uint32_t* ranges;        // uint2[tiles]
	const uint32_t* n_contrib;     // [2*W*H]
	const float* accum_alpha;      // [3*W*H] = {final T, M1, M2}
	const float* median_depth;
	const float* depth_std;
};

void* ref_create() { return new RefCtx(); }

void ref_destroy(void* h) {
	RefCtx* c = (RefCtx*)h;
	if (!c) return;
	c->geom.release(); c->binning.release(); c->img.release();
	c->out_mask.release(); c->dL_dnormal.release();
	delete c;
}

// Mirrors RasterizeGaussiansCUDA (RAST/rasterize_points.cu:39-138).  All pointers are
// device pointers; shs / colors / scales / rotations / transMat_precomp may be null
// exactly where the reference would receive an empty tensor's null data pointer.
int ref_forward(void* h,
	int P, int D, int M,
	const float* background, int W, int H,
	const float* means3D, const float* shs, const float* colors_precomp,
	const float* opacities, const float* scales, float scale_modifier,
	const float* rotations, const float* transMat_precomp,
	const float* viewmatrix, const float* projmatrix, const float* campos,
	float tan_fovx, float tan_fovy, int prefiltered,
	float* out_color, float* out_others, int* radii,
	int use_sa, int debug)
{
	RefCtx* c = (RefCtx*)h;
	c->P = P; c->W = W; c->H = H; c->R = 0;
	// torch::full(...) of the outputs (rasterize_points.cu:86-89)
	cudaMemsetAsync(out_color, 0, sizeof(float) * 3 * (size_t)W * H, 0);
	cudaMemsetAsync(out_others, 0, sizeof(float) * 7 * (size_t)W * H, 0);
	if (P > 0) cudaMemsetAsync(radii, 0, sizeof(int) * (size_t)P, 0);
	cudaMemsetAsync(c->out_mask.resize(sizeof(float) * (size_t)W * H), 0, sizeof(float) * (size_t)W * H, 0);
	if (P == 0) return 0;

	std::function<char*(size_t)> geomF = [c](size_t n) { return c->geom.resize(n); };
	std::function<char*(size_t)> binF  = [c](size_t n) { return c->binning.resize(n); };
	std::function<char*(size_t)> imgF  = [c](size_t n) { return c->img.resize(n); };

	int rendered = CudaRasterizer::Rasterizer::forward(
		geomF, binF, imgF, P, D, M, background, W, H,
		means3D, shs, colors_precomp, opacities, scales, scale_modifier,
		rotations, transMat_precomp, viewmatrix, projmatrix, campos,
		tan_fovx, tan_fovy, prefiltered != 0,
		out_color, (float*)c->out_mask.ptr, out_others, radii,
		use_sa != 0, debug != 0);
	c->R = rendered;
	return rendered;
}

// Decode the opaque byte buffers with the reference's own fromChunk functions.
int ref_state(void* h, RefStatePtrs* out)
{
	RefCtx* c = (RefCtx*)h;
	if (!c || !c->geom.ptr || !c->img.ptr) return -1;
	char* g = c->geom.ptr;
	CudaRasterizer::GeometryState gs = CudaRasterizer::GeometryState::fromChunk(g, c->P);
	char* im = c->img.ptr;
	CudaRasterizer::ImageState is = CudaRasterizer::ImageState::fromChunk(im, (size_t)c->W * c->H);
	memset(out, 0, sizeof(*out));
	out->depths = gs.depths;
	out->internal_radii = gs.internal_radii;
	out->means2D = (const float*)gs.means2D;
	out->transMat = gs.transMat;
	out->normal_opacity = (const float*)gs.normal_opacity;
	out->tiles_touched = gs.tiles_touched;
	out->point_offsets = gs.point_offsets;
	if (c->binning.ptr) {
		char* b = c->binning.ptr;
		CudaRasterizer::BinningState bs = CudaRasterizer::BinningState::fromChunk(b, c->R);
		out->point_list_keys_unsorted = bs.point_list_keys_unsorted;
		out->point_list_keys = bs.point_list_keys;
		out->point_list_unsorted = bs.point_list_unsorted;
		out->point_list = bs.point_list;
	}
	out->ranges = (const uint32_t*)is.ranges;
	out->n_contrib = is.n_contrib;
	out->accum_alpha = is.accum_alpha;
	out->median_depth = is.m_media_depth;
	out->depth_std = is.depth_std;
	return 0;
}

// Mirrors RasterizeGaussiansBackwardCUDA (RAST/rasterize_points.cu:140-239); uses the
// buffers left in the context by the preceding ref_forward.
void ref_backward(void* h,
	int P, int D, int M,
	const float* background, int W, int H,
	const float* means3D, const float* shs, const float* colors_precomp,
	const float* scales, float scale_modifier, const float* rotations,
	const float* transMat_precomp, const float* viewmatrix, const float* projmatrix,
	const float* campos, float tan_fovx, float tan_fovy, const int* radii,
	const float* dL_dpix, const float* dL_dothers,
	float* dL_dmeans2D,   // [P,3]
	float* dL_dcolors,    // [P,3]
	float* dL_dopacity,   // [P,1]
	float* dL_dmeans3D,   // [P,3]
	float* dL_dtransMat,  // [P,9]
	float* dL_dsh,        // [P,M,3]
	float* dL_dscales,    // [P,2]
	float* dL_drotations, // [P,4]
	int use_sa, int debug)
{
	RefCtx* c = (RefCtx*)h;
	size_t p = (size_t)P;
	// torch::zeros(...) x9 (rasterize_points.cu:192-200)
	cudaMemsetAsync(dL_dmeans3D, 0, 12 * p, 0);
	cudaMemsetAsync(dL_dmeans2D, 0, 12 * p, 0);
	cudaMemsetAsync(dL_dcolors, 0, 12 * p, 0);
	cudaMemsetAsync(c->dL_dnormal.resize(12 * p + 16), 0, 12 * p, 0);
	cudaMemsetAsync(dL_dopacity, 0, 4 * p, 0);
	cudaMemsetAsync(dL_dtransMat, 0, 36 * p, 0);
	if (M > 0 && dL_dsh) cudaMemsetAsync(dL_dsh, 0, 12 * p * (size_t)M, 0);
	cudaMemsetAsync(dL_dscales, 0, 8 * p, 0);
	cudaMemsetAsync(dL_drotations, 0, 16 * p, 0);
	if (P == 0) return;

	CudaRasterizer::Rasterizer::backward(P, D, M, c->R, background, W, H,
		means3D, shs, colors_precomp, scales, scale_modifier, rotations,
		transMat_precomp, viewmatrix, projmatrix, campos, tan_fovx, tan_fovy,
		radii, c->geom.ptr, c->binning.ptr, c->img.ptr,
		dL_dpix, dL_dothers,
		dL_dmeans2D, (float*)c->dL_dnormal.ptr, dL_dopacity, dL_dcolors,
		dL_dmeans3D, dL_dtransMat, dL_dsh, dL_dscales, dL_drotations,
		use_sa != 0, debug != 0);
}

// Mirrors markVisible (RAST/rasterize_points.cu:241-260). present: bool[P] (1 byte each).
void ref_mark_visible(int P, float* means3D, float* viewmatrix, float* projmatrix, unsigned char* present)
{
	cudaMemsetAsync(present, 0, (size_t)P, 0);
	if (P == 0) return;
	CudaRasterizer::Rasterizer::markVisible(P, means3D, viewmatrix, projmatrix, (bool*)present);
}

// Mirrors distCUDA2 (KNN/spatial.cu:15-26).
void ref_dist_cuda2(int P, const float* points, float* mean_dists)
{
	cudaMemsetAsync(mean_dists, 0, sizeof(float) * (size_t)P, 0);
	if (P == 0) return;
	SimpleKNN::knn(P, (float3*)points, mean_dists);
}

// Blocking device-to-device copy (lets the Python side snapshot the state arrays).
int ref_copy_d2d(void* dst, const void* src, size_t bytes)
{
	cudaDeviceSynchronize();
	return (int)cudaMemcpy(dst, src, bytes, cudaMemcpyDeviceToDevice);
}

int ref_last_error() { return (int)cudaGetLastError(); }

}  // extern "C"
