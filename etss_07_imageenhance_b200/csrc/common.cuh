// Shared helpers for the b200dce kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/b200dce.h"

#define B200_NUM_SMS 148

namespace b200 {

__host__ inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

int curve_per_iter_unaligned_ok(const float* x, const float* a, float* y, int B, int C, int H, int W, int iters,
                                cudaStream_t s);  // curve.cu

__host__ inline int launch_status() {
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? B200_OK : B200_ERR_LAUNCH;
}

// streaming 128-bit accesses: data that is touched exactly once bypasses L1
__device__ __forceinline__ float4 ld_stream4(const float* p) {
    float4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
                 : "l"(p));
    return r;
}
__device__ __forceinline__ void st_stream4(float* p, const float4& v) {
    asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z),
                 "f"(v.w)
                 : "memory");
}

// One light-enhancement-curve step with ATen's rounding sequence
// (pow(y,2) -> sub -> mul -> add; zero_dcepp.py:203), no FMA contraction.
__device__ __forceinline__ float curve_step(float y, float a) {
    return __fadd_rn(y, __fmul_rn(a, __fsub_rn(__fmul_rn(y, y), y)));
}

// Two curve steps at once on the packed fp32 pipe (FMUL2 / FFMA2 / FADD2), bit-identical to curve_step:
//   d = rn(y - rn(y*y)) = -rn(rn(y*y) - y)   (the FFMA2 by -1 rounds once; negation is exact)
//   y' = rn(y + rn((-a) * d))
// `na` is the NEGATED curve parameter.  Halves the instruction count of the 8-iteration loop.
__device__ __forceinline__ float2 curve_step2(float2 y, float2 na) {
    const float2 p = __fmul2_rn(y, y);
    const float2 d = __ffma2_rn(p, make_float2(-1.f, -1.f), y);
    return __fadd2_rn(y, __fmul2_rn(na, d));
}

// ATen area_pixel_compute_source_index, align_corners=True:  src = scale*dst
struct Tap {
    int i0, i1;
    float w0, w1;
};
__device__ __forceinline__ Tap tap_align_corners(int dst, float scale, int in_size) {
    float src = __fmul_rn(scale, (float)dst);
    int i0 = (int)src;
    if (i0 > in_size - 1) i0 = in_size - 1;
    Tap t;
    t.i0 = i0;
    t.i1 = i0 + (i0 < in_size - 1 ? 1 : 0);
    float l1 = __fsub_rn(src, (float)i0);
    l1 = fminf(fmaxf(l1, 0.f), 1.f);
    t.w1 = l1;
    t.w0 = __fsub_rn(1.f, l1);
    return t;
}
// align_corners=False with an explicit scale (F.interpolate(scale_factor=1/sf)):
//   src = sf*(dst+0.5)-0.5, clamped at 0
__device__ __forceinline__ Tap tap_half_pixel(int dst, float scale, int in_size) {
    float src = __fsub_rn(__fmul_rn(scale, __fadd_rn((float)dst, 0.5f)), 0.5f);
    src = fmaxf(src, 0.f);
    int i0 = (int)src;
    if (i0 > in_size - 1) i0 = in_size - 1;
    Tap t;
    t.i0 = i0;
    t.i1 = i0 + (i0 < in_size - 1 ? 1 : 0);
    float l1 = __fsub_rn(src, (float)i0);
    l1 = fminf(fmaxf(l1, 0.f), 1.f);
    t.w1 = l1;
    t.w0 = __fsub_rn(1.f, l1);
    return t;
}

}  // namespace b200
