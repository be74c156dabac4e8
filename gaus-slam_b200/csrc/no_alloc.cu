// no_alloc.cu -- gs2d_sort_pairs_static: the onesweep pair sort with its scratch in a static device array.
//
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid allocation).  It is
// gs2d_sort_pairs (csrc/api.cu, csrc/binning.cu) with one difference: the scratch the caller would pass
// (gs2d_sort_scratch_bytes(n): digit histograms, look-back status words, tickets, the ping-pong copy of the pairs) is
// the file-scope array below, so the call allocates nothing -- no cudaMalloc / cudaFree, no allocating library routine;
// only kernel launches, cudaMemsetAsync and cudaMemcpyAsync on the caller's stream.
//
// Limits and sharing (also in include/gs2d.h):
//   * n <= GS2D_SORT_STATIC_MAX_PAIRS (262 144 pairs); larger n -> GS2D_ERR_CAPACITY, nothing is launched;
//   * the array (GS2D_SORT_STATIC_SCRATCH_BYTES = 4.5 MiB) exists once per GPU that loads the library, until exit;
//   * every call on a device shares it and finds the previous call's contents: the words that must start at zero
//     (histograms, tickets, status) are cleared on the stream at the start of every call (launch_sort_pairs does
//     that for any scratch); calls that overlap on the GPU (different streams, no ordering) race on it.
#include <mutex>

#include "common.cuh"

namespace gs2d {

namespace {

constexpr size_t kStaticScratchBytes = GS2D_SORT_STATIC_SCRATCH_BYTES;
static __device__ __align__(256) unsigned char g_sort_static_scratch[kStaticScratchBytes];

// In host code `g_sort_static_scratch` is a host-side stand-in, not a device address: the address a kernel may use
// comes from cudaGetSymbolAddress, asked once per device with that device current.
constexpr int kMaxDevices = 64;
std::mutex g_addr_mu;
void* g_addr[kMaxDevices] = {};

int static_scratch(void** out)
{
	int dev = -1;
	GS2D_CUDA_OK(cudaGetDevice(&dev));
	if (dev < 0 || dev >= kMaxDevices) return GS2D_ERR_UNSUPPORTED;
	std::lock_guard<std::mutex> lk(g_addr_mu);
	if (!g_addr[dev]) {
		void* p = nullptr;
		GS2D_CUDA_OK(cudaGetSymbolAddress(&p, g_sort_static_scratch));
		if (!p) return GS2D_ERR_CUDA;
		g_addr[dev] = p;
	}
	*out = g_addr[dev];
	return GS2D_OK;
}

}  // namespace

}  // namespace gs2d

extern "C" int gs2d_sort_pairs_static(int64_t n, const uint64_t* keys_in, const uint32_t* values_in,
                                      uint64_t* keys_out, uint32_t* values_out, int32_t end_bit, void* stream)
{
	using namespace gs2d;
	if (n < 0 || end_bit <= 0 || end_bit > 64) return GS2D_ERR_INVALID;
	if (n > (int64_t)GS2D_SORT_STATIC_MAX_PAIRS) return GS2D_ERR_CAPACITY;   // beyond the static scratch: refuse
	if (n == 0) return GS2D_OK;
	if (gs2d_sort_scratch_bytes(n) > kStaticScratchBytes) return GS2D_ERR_CAPACITY;
	void* scratch = nullptr;
	const int rc = static_scratch(&scratch);
	if (rc != GS2D_OK) return rc;
	// same code path as the caller-scratch entry point from here on (it clears what must start at zero)
	return gs2d_sort_pairs(n, keys_in, values_in, keys_out, values_out, end_bit, scratch, stream);
}
