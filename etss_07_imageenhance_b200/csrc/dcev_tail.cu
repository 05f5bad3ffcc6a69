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
// Everything is HBM-bound streaming except the S x S coefficient stage (256 x 256 per frame: microseconds).
#include <math.h>

#include "common.cuh"

namespace b200 {

constexpr float GF_EPS = 1e-8f;
constexpr int GF_R = 1;  // ZeroDCEV.filter_up(radius=1)

__host__ __device__ inline size_t gf_align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// diff_x / diff_y of box_filter.py on a prefix array `p` with stride `st`: window [i - r, i + r] clipped to the ends
__device__ __forceinline__ float gf_diff(const float* p, int i, int n, int st) {
    if (i <= GF_R) return p[(size_t)(i + GF_R) * st];
    const float lo = p[(size_t)(i - GF_R - 1) * st];
    const float hi = i < n - GF_R ? p[(size_t)(i + GF_R) * st] : p[(size_t)(n - 1) * st];
    return __fsub_rn(hi, lo);
}

// Prefix planes of one of the four maps m (x, y, x*y, x*x) of one frame; grid (B, 4), 256 threads:
//   phase 1 (thread = column, coalesced over the columns): P1 = cumsum down the rows
//   phase 2 (thread = row): P2 = cumsum along the row of diff_x(P1); the warp brings 32 x 32 blocks of diff_x(P1)
//           through shared memory (coalesced 128-byte rows in, conflict-free scan, coalesced rows out)
__global__ void __launch_bounds__(256) dcev_gf_prefix_kernel(const float* __restrict__ x_lr, const float* __restrict__ y_lr,
                                                             float* __restrict__ scratch, int S) {
    __shared__ float tile[8][32][33];
    const int f = blockIdx.x, m = blockIdx.y;
    const size_t ss = (size_t)S * S;
    const float* x = x_lr + f * ss;
    const float* y = y_lr + f * ss;
    float* p1 = scratch + ((size_t)f * 8 + m) * ss;
    float* p2 = scratch + ((size_t)f * 8 + 4 + m) * ss;
    auto value = [&](size_t o) {
        const float xv = x[o];
        if (m == 0) return xv;
        if (m == 3) return __fmul_rn(xv, xv);
        const float yv = y[o];
        return m == 1 ? yv : __fmul_rn(xv, yv);
    };
    for (int w = threadIdx.x; w < S; w += blockDim.x) {
        double acc = 0.0;
        int h = 0;
        for (; h + 8 <= S; h += 8) {  // eight independent loads in flight, the adds stay in order
            float v[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) v[k] = value((size_t)(h + k) * S + w);
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                acc += (double)v[k];
                p1[(size_t)(h + k) * S + w] = (float)acc;
            }
        }
        for (; h < S; ++h) {
            acc += (double)value((size_t)h * S + w);
            p1[(size_t)h * S + w] = (float)acc;
        }
    }
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int r0 = warp * 32; r0 < S; r0 += 256) {  // this warp's 32 rows; lane = row in the scan, column in the copies
        double acc = 0.0;
        for (int c0 = 0; c0 < S; c0 += 32) {
            const int w = c0 + lane;
#pragma unroll 8
            for (int r = 0; r < 32; ++r)
                tile[warp][r][lane] = (r0 + r < S && w < S) ? gf_diff(p1 + w, r0 + r, S, S) : 0.f;
            __syncwarp();
            float o[32];
#pragma unroll
            for (int k = 0; k < 32; ++k) {
                acc += (double)tile[warp][lane][k];
                o[k] = (float)acc;
            }
            __syncwarp();
#pragma unroll
            for (int k = 0; k < 32; ++k) tile[warp][lane][k] = o[k];
            __syncwarp();
#pragma unroll 8
            for (int r = 0; r < 32; ++r)
                if (r0 + r < S && w < S) p2[(size_t)(r0 + r) * S + w] = tile[warp][r][lane];
            __syncwarp();
        }
    }
}

// window sums = diff_y(P2), then the linear model of every window: A = cov / (var + eps), b = mean_y - A mean_x
__global__ void __launch_bounds__(256) dcev_gf_ab_kernel(const float* __restrict__ scratch, float* __restrict__ A,
                                                         float* __restrict__ Bc, int S, long long total) {
    const size_t ss = (size_t)S * S;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int f = (int)(i / (long long)ss);
        const size_t j = (size_t)(i - (long long)f * (long long)ss);
        const int h = (int)(j / S), w = (int)(j - (size_t)h * S);
        const int nh = min(h + GF_R, S - 1) - max(h - GF_R, 0) + 1, nw = min(w + GF_R, S - 1) - max(w - GF_R, 0) + 1;
        const float n = (float)(nh * nw);  // box_filter of ones: exact
        const float* p2 = scratch + ((size_t)f * 8 + 4) * ss;
        float b[4];
#pragma unroll
        for (int m = 0; m < 4; ++m) b[m] = gf_diff(p2 + m * ss + (size_t)h * S, w, S, 1);
        const float mean_x = __fdiv_rn(b[0], n), mean_y = __fdiv_rn(b[1], n);
        const float cov = __fsub_rn(__fdiv_rn(b[2], n), __fmul_rn(mean_x, mean_y));
        const float var = __fsub_rn(__fdiv_rn(b[3], n), __fmul_rn(mean_x, mean_x));
        const float a = __fdiv_rn(cov, __fadd_rn(var, GF_EPS));
        A[i] = a;
        Bc[i] = __fsub_rn(mean_y, __fmul_rn(a, mean_x));
    }
}

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

// Full resolution: image_v_fixed = clip(A_up * v + b_up), the recoloured pixel (not yet divided by the frame maximum)
// and the frame maximum (non-negative floats order like their bit patterns).  Thread = column, CTA = 256 columns x
// GF_ROWS rows: the column's four x taps are computed once, and the four x-interpolated source rows of A and b are
// kept while consecutive output rows fall between the same source rows (every ~4 rows change at 256 -> 1080).
constexpr int GF_ROWS = 32;
__global__ void __launch_bounds__(256) dcev_gf_apply_kernel(const float* __restrict__ rgb, const float* __restrict__ v,
                                                            const float* __restrict__ A, const float* __restrict__ Bc,
                                                            float* __restrict__ v_fixed, float* __restrict__ out,
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
            float ro, go, bo;
            recolor(rgb[(f * 3 + 0) * hw + p], rgb[(f * 3 + 1) * hw + p], rgb[(f * 3 + 2) * hw + p], vf, ro, go, bo);
            out[(f * 3 + 0) * hw + p] = ro;
            out[(f * 3 + 1) * hw + p] = go;
            out[(f * 3 + 2) * hw + p] = bo;
            local_max = fmaxf(local_max, fmaxf(fmaxf(ro, go), bo));
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) local_max = fmaxf(local_max, __shfl_xor_sync(0xffffffffu, local_max, o));
    if ((threadIdx.x & 31) == 0 && local_max > 0.f) atomicMax(fmax + f, __float_as_uint(local_max));
}

__global__ void __launch_bounds__(256) dcev_normalise_kernel(float* __restrict__ out, const unsigned int* __restrict__ fmax,
                                                             long long per_frame, long long total) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x)
        out[i] = __fdiv_rn(out[i], __uint_as_float(fmax[i / per_frame]));
}

}  // namespace b200

// scratch: 8 S x S fp32 prefix planes + A + b per frame, one max word per frame
extern "C" size_t b200_dcev_tail_workspace_bytes(int B, int S) {
    if (B <= 0 || S <= 0) return 0;
    return b200::gf_align_up((size_t)B * 10 * S * S * sizeof(float), 256) + b200::gf_align_up((size_t)B * 4, 256);
}

extern "C" int b200_dcev_tail(const float* image_rgb, const float* image_v, const float* image_v_lr,
                              const float* image_v_fixed_lr, float* image_v_fixed, float* enhanced, int B, int H, int W,
                              int S, void* workspace, size_t workspace_bytes, void* stream) {
    using namespace b200;
    if (!image_rgb || !image_v || !image_v_lr || !image_v_fixed_lr || !image_v_fixed || !enhanced || !workspace)
        return B200_ERR_NULL;
    if (B <= 0 || H <= 0 || W <= 0 || S <= 0) return B200_ERR_BAD_SHAPE;
    // FastGuidedFilter asserts h, w > 2 * radius + 1 (guided_filter.py:182)
    if (S <= 2 * GF_R + 1 || B > 65535) return B200_ERR_UNSUPPORTED;
    if (workspace_bytes < b200_dcev_tail_workspace_bytes(B, S)) return B200_ERR_WORKSPACE;
    if ((reinterpret_cast<uintptr_t>(workspace) & 15u) != 0) return B200_ERR_MISALIGNED;
    cudaStream_t s = (cudaStream_t)stream;
    const size_t ss = (size_t)S * S;
    float* scratch = (float*)workspace;
    float* A = scratch + (size_t)B * 8 * ss;
    float* Bc = A + (size_t)B * ss;
    unsigned int* fmax = (unsigned int*)((char*)workspace + gf_align_up((size_t)B * 10 * ss * sizeof(float), 256));
    if (cudaMemsetAsync(fmax, 0, (size_t)B * 4, s) != cudaSuccess) return B200_ERR_LAUNCH;
    dcev_gf_prefix_kernel<<<dim3(B, 4), 256, 0, s>>>(image_v_lr, image_v_fixed_lr, scratch, S);
    {
        const long long tot = (long long)B * (long long)ss;
        long long blk = (tot + 255) / 256;
        if (blk > B200_NUM_SMS * 16) blk = B200_NUM_SMS * 16;
        dcev_gf_ab_kernel<<<(unsigned)blk, 256, 0, s>>>(scratch, A, Bc, S, tot);
    }
    // ATen area_pixel_compute_scale<float>(in, out, align_corners=true) = (in - 1) / (out - 1), 0 when out == 1
    const float sh = H > 1 ? (float)(S - 1) / (float)(H - 1) : 0.f;
    const float sw = W > 1 ? (float)(S - 1) / (float)(W - 1) : 0.f;
    dim3 grid((W + 255) / 256, (H + GF_ROWS - 1) / GF_ROWS, B);
    dcev_gf_apply_kernel<<<grid, 256, 0, s>>>(image_rgb, image_v, A, Bc, image_v_fixed, enhanced, fmax, H, W, S, sh, sw);
    const long long per = 3LL * H * W, total = per * B;
    long long blocks = (total + 255) / 256;
    if (blocks > B200_NUM_SMS * 16) blocks = B200_NUM_SMS * 16;
    dcev_normalise_kernel<<<(unsigned)blocks, 256, 0, s>>>(enhanced, fmax, per, total);
    return launch_status();
}
