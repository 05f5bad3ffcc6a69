// tcgen05 Zero-DCE path (placeholder until the implicit-GEMM kernel lands): reports UNSUPPORTED.
#include "common.cuh"
#include "dce_tc.cuh"

namespace b200 {
size_t dce_tc_packed_bytes(int, int) { return 0; }
int dce_tc_pack_weights(const float* const*, const float* const*, void*, int, int, cudaStream_t) {
    return B200_ERR_UNSUPPORTED;
}
size_t dce_tc_workspace_bytes(int, int, int, int) { return 0; }
int dce_tc_forward(const float*, const void*, float*, float*, int, int, int, int, int, void*, size_t, cudaStream_t) {
    return B200_ERR_UNSUPPORTED;
}
}  // namespace b200
