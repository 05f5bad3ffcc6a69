// guided_filter.cu's first stage for other translation units (dcev_tail.cu)
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>

namespace b200 {
size_t gfl_tot_doubles(int planes, int H, int W);
// A, b = the linear model of every (2r + 1)^2 window of (x, y) [planes][H][W]; H, W >= 2r + 1, W <= 4096
int gfl_coefficients(const float* x, const float* y, float* A, float* Bc, int planes, int H, int W, int r, float eps,
                     float* d1, size_t map_stride, double* tot, cudaStream_t s);
}  // namespace b200
