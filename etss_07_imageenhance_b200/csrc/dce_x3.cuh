// Internal interface of the split-bf16 (B200_PREC_BF16X3) Zero-DCE path; implemented in dce_x3.cu.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>

namespace b200 {
size_t dce_x3_packed_bytes(int iters);
int dce_x3_pack_weights(const float* const* w, const float* const* b, void* packed, int iters, cudaStream_t s);
size_t dce_x3_workspace_bytes(int B, int H, int W);
int dce_x3_forward(const float* x, const void* packed, float* adjust, float* enhanced, int B, int H, int W, int iters,
                   void* workspace, size_t workspace_bytes, cudaStream_t s);
}  // namespace b200
