// collective.cu -- all-reduce of the flat [13 * P] map-gradient buffer over NVSwitch multicast (NVLS).
//
// The view-sharded mapping step ends with one all-reduce of the gradient buffer the backward pass has just
// written (gs2d_b200/mapping.py).  With the buffer in symmetric memory that is bound to a multicast object
// (torch.distributed._symmetric_memory allocates and binds it), the reduction runs IN the switch:
// rank r owns the r-th slice; for every 16 bytes of its slice it issues one multimem.ld_reduce -- the switch
// fetches the 16 bytes from all ranks' copies and returns their fp32 sum -- and one multimem.st, which the switch
// broadcasts to all ranks' copies.  Every byte crosses each GPU's links once in each direction (the two-shot
// pattern, without an intermediate buffer and without SM-side adds).  The caller brackets the launch with
// cross-GPU barriers (all gradients written / all slices stored).
#include "common.cuh"

namespace gs2d {

namespace {

__global__ void __launch_bounds__(512)
allreduce_multimem_kernel(float4* mc, long long first, long long last)
{
	for (long long i = first + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < last;
	     i += (long long)gridDim.x * blockDim.x) {
		float4 v;
		float4* p = mc + i;
		asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
		             : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
		             : "l"(p)
		             : "memory");
		asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y),
		             "f"(v.z), "f"(v.w)
		             : "memory");
	}
}

}  // namespace

}  // namespace gs2d

using namespace gs2d;

extern "C" int gs2d_allreduce_multimem(void* multicast_ptr, int64_t n_floats, int32_t rank, int32_t world,
                                       int32_t max_blocks, void* stream)
{
	if (!multicast_ptr || n_floats < 0 || world <= 0 || rank < 0 || rank >= world) return GS2D_ERR_INVALID;
	if ((n_floats & 3) || ((uintptr_t)multicast_ptr & 15)) return GS2D_ERR_INVALID;
	const long long n4 = n_floats / 4;
	const long long chunk = (n4 + world - 1) / world;
	const long long first = chunk * rank, last = first + chunk < n4 ? first + chunk : n4;
	if (first >= last) return GS2D_OK;
	cudaStream_t st = (cudaStream_t)stream;
	long long blocks = (last - first + 511) / 512;
	const long long cap = max_blocks > 0 ? max_blocks : 148 * 2;
	if (blocks > cap) blocks = cap;
	ProfScope ps(K_ALLREDUCE, st);
	allreduce_multimem_kernel<<<(unsigned)blocks, 512, 0, st>>>((float4*)multicast_ptr, first, last);
	count_launch();
	GS2D_CUDA_OK(cudaGetLastError());
	return GS2D_OK;
}
