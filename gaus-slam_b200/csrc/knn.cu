// knn.cu -- distCUDA2 (placeholder until the Morton/box-pruned 3-NN kernels land).
#include "common.cuh"
extern "C" {
size_t gs2d_dist2_scratch_bytes(int32_t P) { return (size_t)(P > 0 ? P : 0) * 64 + 4096; }
int gs2d_dist2(int32_t, const float*, float*, void*, void*) { return GS2D_ERR_UNSUPPORTED; }
}
