// Internal interface of the tcgen05 (bf16 / split-bf16) Zero-DCE path; implemented in dce_tc.cu.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>

namespace b200 {
size_t dce_tc_packed_bytes(int precision, int iters);
int dce_tc_pack_weights(const float* const* w, const float* const* b, void* packed, int precision, int iters,
                        cudaStream_t s);
size_t dce_tc_workspace_bytes(int B, int H, int W, int precision);
int dce_tc_forward(const float* x, const void* packed, float* adjust, float* enhanced, int B, int H, int W, int iters,
                   int precision, void* workspace, size_t workspace_bytes, cudaStream_t s);
}  // namespace b200
