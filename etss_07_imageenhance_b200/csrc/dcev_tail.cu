// Tail of ZeroDCEV.forward (src/mon/vision/enhance/llie/zero_dce/zero_dce_v.py:193-197, :208-230):
//   image_v_fixed = clip(FastGuidedFilter(radius=1)(image_v_lr, image_v_fixed_lr, image_v), 0, 1)
//                   (src/mon/vision/filtering/guided_filter.py:171-203, box filter: box_filter.py:26-117)
//   enhanced      = hsv_to_rgb(H, S of the input, V = image_v_fixed) / max over the frame   (kornia's conversions)
//
// The guided filter fits y = A x + b in 3x3 windows of the low-resolution pair and applies the bicubically upsampled
// (align_corners=True) A, b to the full-resolution V.  A = cov / (var + 1e-8) is ill-conditioned where the window is
// flat (cov and var are differences of nearly equal numbers), so the low-resolution part repeats the reference's
// rounding sequence exactly:
//   box_filter(t) = diff_y(diff_x(t.cumsum(2)).cumsum(3)): ATen's CPU cumsum accumulates fp32 input in DOUBLE and
//   rounds every prefix to fp32; the window sums are differences of those rounded prefixes (not exact window sums).
//   Sums of <= 2^11 fp32 values >= 2^-19 are exact in double in any order, so the prefixes here equal ATen's.
// Everything is HBM-bound streaming except the S x S coefficient stage (256 x 256 per frame: three small launches of
// guided_filter.cu, ~30 us for 16 frames).
#include <math.h>

#include "common.cuh"
#include "guided_filter.cuh"

namespace b200 {

constexpr float GF_EPS = 1e-8f;
constexpr int GF_R = 1;  // ZeroDCEV.filter_up(radius=1)

__host__ __device__ inline size_t gf_align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// The S x S coefficient stage (box filters of x, y, x*y, x*x as differences of rounded double-accumulated running sums,
// then A = cov / (var + eps), b = mean_y - A mean_x) is guided_filter.cu's first stage: gfl_coefficients().

// ---- bicubic upsample (ATen upsample_bicubic2d, align_corners=True, A = -0.75) -------------------------------------
__device__ __forceinline__ float cubic1(float x) {  // |x| <= 1
    return __fadd_rn(__fmul_rn(__fmul_rn(__fsub_rn(__fmul_rn(1.25f, x), 2.25f), x), x), 1.f);
}
__device__ __forceinline__ float cubic2(float x) {  // 1 < |x| < 2
    return __fsub_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fsub_rn(__fmul_rn(-0.75f, x), -3.75f), x), -6.f), x), -3.f);
}
struct CubicTap {
    int i[4];
    float w[4];
};
__device__ __forceinline__ CubicTap cubic_tap(int dst, float scale, int in_size) {
    const float real = __fmul_rn(scale, (float)dst);
    const float fl = floorf(real);
    const float t = __fsub_rn(real, fl);
    const int i0 = (int)fl;
    CubicTap c;
#pragma unroll
    for (int k = 0; k < 4; ++k) c.i[k] = min(max(i0 - 1 + k, 0), in_size - 1);
    c.w[0] = cubic2(__fadd_rn(t, 1.f));
    c.w[1] = cubic1(t);
    c.w[2] = cubic1(__fsub_rn(1.f, t));
    c.w[3] = cubic2(__fsub_rn(2.f, t));
    return c;
}

// kornia.color.rgb_to_hsv / hsv_to_rgb as published (eps = 1e-8), with V replaced in between; fp32 op for op as the
// oracle's restatement (oracle/oracle.py rgb_to_hsv_kornia / hsv_to_rgb_kornia)
// Python / torch `%` on the two ranges that occur (no fmodf: the library routine is ~40 instructions):
__device__ __forceinline__ float pymod1(float a) {  // a % 1.0 for |a| < 2^23: the fractional part is exact
    float r = __fsub_rn(a, truncf(a));
    if (r < 0.f) r = __fadd_rn(r, 1.f);
    return r;
}
__device__ __forceinline__ float pymod6(float a) {  // a % 6.0 for 0 <= a < 12: the subtraction is exact
    return a >= 6.f ? __fsub_rn(a, 6.f) : a;
}
__device__ __forceinline__ void recolor(float r, float g, float b, float v_new, float& ro, float& go, float& bo) {
    const float mx = fmaxf(fmaxf(r, g), b), mn = fminf(fminf(r, g), b);
    const int arg = (r >= g && r >= b) ? 0 : (g >= b ? 1 : 2);  // torch.max returns the first maximal index
    float delta = __fsub_rn(mx, mn);
    const float s = __fdiv_rn(delta, __fadd_rn(mx, 1e-8f));
    if (delta == 0.f) delta = 1.f;
    const float rc = __fsub_rn(mx, r), gc = __fsub_rn(mx, g), bc = __fsub_rn(mx, b);
    float h = arg == 0 ? __fsub_rn(bc, gc)
            : arg == 1 ? __fadd_rn(__fsub_rn(rc, bc), __fmul_rn(2.f, delta))
                       : __fadd_rn(__fsub_rn(gc, rc), __fmul_rn(4.f, delta));
    h = __fdiv_rn(h, delta);
    h = pymod1(__fdiv_rn(h, 6.f));
    h = __fmul_rn(6.2831854820251465f, h);          // 2 * math.pi as the fp32 scalar torch multiplies by
    // hsv_to_rgb
    const float hh = __fdiv_rn(h, 6.2831854820251465f);
    const float h6 = __fmul_rn(hh, 6.f);
    const float hi = pymod6(floorf(h6));
    const float f = __fsub_rn(pymod6(h6), hi);
    const float p = __fmul_rn(v_new, __fsub_rn(1.f, s));
    const float q = __fmul_rn(v_new, __fsub_rn(1.f, __fmul_rn(f, s)));
    const float t = __fmul_rn(v_new, __fsub_rn(1.f, __fmul_rn(__fsub_rn(1.f, f), s)));
    switch ((int)hi) {
        case 0: ro = v_new, go = t, bo = p; break;
        case 1: ro = q, go = v_new, bo = p; break;
        case 2: ro = p, go = v_new, bo = t; break;
        case 3: ro = p, go = q, bo = v_new; break;
        case 4: ro = t, go = p, bo = v_new; break;
        default: ro = v_new, go = p, bo = q; break;
    }
}

// Full resolution, first pass: image_v_fixed = clip(A_up * v + b_up) and its maximum over the frame (non-negative floats
// order like their bit patterns).  That maximum IS the maximum of the recoloured frame the reference divides by
// (zero_dce_v.py:228-229): hsv_to_rgb hands one channel the new V unchanged and the other two V times a factor in
// [0, 1] (p, q, t above; s and f lie in [0, 1]), so max(r, g, b) == V bit for bit -- the recoloured frame need not be
// written, read back and rescaled.  Thread = column, CTA = 256 columns x
// GF_ROWS rows: the column's four x taps are computed once, and the four x-interpolated source rows of A and b are
// kept while consecutive output rows fall between the same source rows (every ~4 rows change at 256 -> 1080).
constexpr int GF_ROWS = 32;
__global__ void __launch_bounds__(256) dcev_gf_apply_kernel(const float* __restrict__ v, const float* __restrict__ A,
                                                            const float* __restrict__ Bc, float* __restrict__ v_fixed,
                                                            unsigned int* __restrict__ fmax, int H, int W, int S,
                                                            float sh, float sw) {
    const int f = blockIdx.z, x = blockIdx.x * 256 + threadIdx.x;
    const size_t hw = (size_t)H * W, ss = (size_t)S * S;
    const float* a = A + f * ss;
    const float* b = Bc + f * ss;
    float local_max = 0.f;
    if (x < W) {
        const CubicTap tx = cubic_tap(x, sw, S);
        float ra[4], rb[4];
        int have = -0x7fffffff;  // first source row of the cached window
        const int y1 = min(H, (blockIdx.y + 1) * GF_ROWS);
        for (int y = blockIdx.y * GF_ROWS; y < y1; ++y) {
            const CubicTap ty = cubic_tap(y, sh, S);
            const int first = (int)floorf(__fmul_rn(sh, (float)y)) - 1;  // unclamped: identifies the window
            if (first != have) {
                have = first;
#pragma unroll
                for (int j = 0; j < 4; ++j) {  // ATen: a row is interpolated along x first
                    const float* ar = a + (size_t)ty.i[j] * S;
                    const float* br = b + (size_t)ty.i[j] * S;
                    ra[j] = __fmul_rn(__ldg(ar + tx.i[0]), tx.w[0]);
                    rb[j] = __fmul_rn(__ldg(br + tx.i[0]), tx.w[0]);
#pragma unroll
                    for (int i = 1; i < 4; ++i) {
                        ra[j] = __fadd_rn(ra[j], __fmul_rn(__ldg(ar + tx.i[i]), tx.w[i]));
                        rb[j] = __fadd_rn(rb[j], __fmul_rn(__ldg(br + tx.i[i]), tx.w[i]));
                    }
                }
            }
            float ua = __fmul_rn(ra[0], ty.w[0]), ub = __fmul_rn(rb[0], ty.w[0]);
#pragma unroll
            for (int j = 1; j < 4; ++j) {
                ua = __fadd_rn(ua, __fmul_rn(ra[j], ty.w[j]));
                ub = __fadd_rn(ub, __fmul_rn(rb[j], ty.w[j]));
            }
            const size_t p = (size_t)y * W + x;
            const float vf = fminf(fmaxf(__fadd_rn(__fmul_rn(ua, v[f * hw + p]), ub), 0.f), 1.f);
            v_fixed[f * hw + p] = vf;
            local_max = fmaxf(local_max, vf);
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) local_max = fmaxf(local_max, __shfl_xor_sync(0xffffffffu, local_max, o));
    if ((threadIdx.x & 31) == 0 && local_max > 0.f) atomicMax(fmax + f, __float_as_uint(local_max));
}

// Second pass: enhanced = hsv_to_rgb(H, S of the input, V = image_v_fixed) / the frame's maximum (IEEE division, as
// torch's).  grid (blocks, B); four pixels per thread with 128-bit accesses when the frame allows.
__global__ void __launch_bounds__(256) dcev_recolor_kernel(const float* __restrict__ rgb, const float* __restrict__ v_fixed,
                                                           float* __restrict__ out, const unsigned int* __restrict__ fmax,
                                                           long long hw, int vec) {
    const long long f = blockIdx.y;
    const float m = __uint_as_float(fmax[f]);
    const float* r = rgb + f * 3 * hw;
    const float* vf = v_fixed + f * hw;
    float* o = out + f * 3 * hw;
    const long long step = (long long)gridDim.x * blockDim.x, i0 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (vec) {
        for (long long i = i0; i < (hw >> 2); i += step) {
            const float4 r4 = reinterpret_cast<const float4*>(r)[i], g4 = reinterpret_cast<const float4*>(r + hw)[i],
                         b4 = reinterpret_cast<const float4*>(r + 2 * hw)[i], v4 = reinterpret_cast<const float4*>(vf)[i];
            float4 ro, go, bo;
            recolor(r4.x, g4.x, b4.x, v4.x, ro.x, go.x, bo.x);
            recolor(r4.y, g4.y, b4.y, v4.y, ro.y, go.y, bo.y);
            recolor(r4.z, g4.z, b4.z, v4.z, ro.z, go.z, bo.z);
            recolor(r4.w, g4.w, b4.w, v4.w, ro.w, go.w, bo.w);
            ro.x = __fdiv_rn(ro.x, m), ro.y = __fdiv_rn(ro.y, m), ro.z = __fdiv_rn(ro.z, m), ro.w = __fdiv_rn(ro.w, m);
            go.x = __fdiv_rn(go.x, m), go.y = __fdiv_rn(go.y, m), go.z = __fdiv_rn(go.z, m), go.w = __fdiv_rn(go.w, m);
            bo.x = __fdiv_rn(bo.x, m), bo.y = __fdiv_rn(bo.y, m), bo.z = __fdiv_rn(bo.z, m), bo.w = __fdiv_rn(bo.w, m);
            reinterpret_cast<float4*>(o)[i] = ro;
            reinterpret_cast<float4*>(o + hw)[i] = go;
            reinterpret_cast<float4*>(o + 2 * hw)[i] = bo;
        }
    } else {
        for (long long i = i0; i < hw; i += step) {
            float ro, go, bo;
            recolor(r[i], r[hw + i], r[2 * hw + i], vf[i], ro, go, bo);
            o[i] = __fdiv_rn(ro, m), o[hw + i] = __fdiv_rn(go, m), o[2 * hw + i] = __fdiv_rn(bo, m);
        }
    }
}

}  // namespace b200

// scratch: the four difference maps of the coefficient stage + A + b (B x S x S fp32 each), its chunk totals, one max word
// per frame
extern "C" size_t b200_dcev_tail_workspace_bytes(int B, int S) {
    if (B <= 0 || S <= 0) return 0;
    return 6 * b200::gf_align_up((size_t)B * S * S * sizeof(float), 256) +
           b200::gf_align_up(b200::gfl_tot_doubles(B, S, S) * sizeof(double), 256) + b200::gf_align_up((size_t)B * 4, 256);
}

extern "C" int b200_dcev_tail(const float* image_rgb, const float* image_v, const float* image_v_lr,
                              const float* image_v_fixed_lr, float* image_v_fixed, float* enhanced, int B, int H, int W,
                              int S, void* workspace, size_t workspace_bytes, void* stream) {
    using namespace b200;
    if (!image_rgb || !image_v || !image_v_lr || !image_v_fixed_lr || !image_v_fixed || !enhanced || !workspace)
        return B200_ERR_NULL;
    if (B <= 0 || H <= 0 || W <= 0 || S <= 0) return B200_ERR_BAD_SHAPE;
    // FastGuidedFilter asserts h, w > 2 * radius + 1 (guided_filter.py:182)
    if (S <= 2 * GF_R + 1 || S > 4096 || B > 65535) return B200_ERR_UNSUPPORTED;
    if (workspace_bytes < b200_dcev_tail_workspace_bytes(B, S)) return B200_ERR_WORKSPACE;
    if ((reinterpret_cast<uintptr_t>(workspace) & 15u) != 0) return B200_ERR_MISALIGNED;
    cudaStream_t s = (cudaStream_t)stream;
    const size_t plane_set = gf_align_up((size_t)B * S * S * sizeof(float), 256) / sizeof(float);
    float* d1 = (float*)workspace;  // 4 maps
    float* A = d1 + 4 * plane_set;
    float* Bc = A + plane_set;
    double* tot = (double*)(Bc + plane_set);
    unsigned int* fmax = (unsigned int*)((char*)tot + gf_align_up(gfl_tot_doubles(B, S, S) * sizeof(double), 256));
    if (cudaMemsetAsync(fmax, 0, (size_t)B * 4, s) != cudaSuccess) return B200_ERR_LAUNCH;
    if (int rc = gfl_coefficients(image_v_lr, image_v_fixed_lr, A, Bc, B, S, S, GF_R, GF_EPS, d1, plane_set, tot, s)) return rc;
    // ATen area_pixel_compute_scale<float>(in, out, align_corners=true) = (in - 1) / (out - 1), 0 when out == 1
    const float sh = H > 1 ? (float)(S - 1) / (float)(H - 1) : 0.f;
    const float sw = W > 1 ? (float)(S - 1) / (float)(W - 1) : 0.f;
    dim3 grid((W + 255) / 256, (H + GF_ROWS - 1) / GF_ROWS, B);
    dcev_gf_apply_kernel<<<grid, 256, 0, s>>>(image_v, A, Bc, image_v_fixed, fmax, H, W, S, sh, sw);
    const long long hw = (long long)H * W;
    const int vec = (hw & 3) == 0 && ((reinterpret_cast<uintptr_t>(image_rgb) | reinterpret_cast<uintptr_t>(image_v_fixed) |
                                       reinterpret_cast<uintptr_t>(enhanced)) & 15u) == 0;
    long long blocks = ((vec ? hw / 4 : hw) + 255) / 256;
    if (blocks > B200_NUM_SMS * 4) blocks = B200_NUM_SMS * 4;
    dcev_recolor_kernel<<<dim3((unsigned)blocks, B), 256, 0, s>>>(image_rgb, image_v_fixed, enhanced, fmax, hw, vec);
    return launch_status();
}
