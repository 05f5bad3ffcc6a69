// Light-enhancement curve kernels and u8<->fp32 frame conversion.
// HBM-bound streaming kernels: 128-bit accesses, grid sized in multiples of the SM count.
#include "common.cuh"

namespace b200 {

// PER_ITER=false: a is [B,C,H,W]     (zero_dcepp.py:201-203)
// PER_ITER=true : a is [B,C*iters,H,W], iteration i uses plane i*C+c (zero_dce.py:183-188)
template <bool PER_ITER>
__global__ void __launch_bounds__(256) curve_vec4_kernel(const float* __restrict__ x, const float* __restrict__ a,
                                                         float* __restrict__ y, long long planes, int C,
                                                         long long hw4, int iters) {
    // one item = 4 consecutive pixels of one (b,c) plane
    const long long total = planes * hw4;
    for (long long it = (long long)blockIdx.x * blockDim.x + threadIdx.x; it < total;
         it += (long long)gridDim.x * blockDim.x) {
        const long long plane = it / hw4;
        const long long p4 = it - plane * hw4;
        const float* xp = x + (plane * hw4 + p4) * 4;
        float4 v = ld_stream4(xp);
        if (!PER_ITER) {
            float4 m = ld_stream4(a + (plane * hw4 + p4) * 4);
            for (int i = 0; i < iters; ++i) {
                v.x = curve_step(v.x, m.x);
                v.y = curve_step(v.y, m.y);
                v.z = curve_step(v.z, m.z);
                v.w = curve_step(v.w, m.w);
            }
        } else {
            const long long b = plane / C;
            const int c = (int)(plane - b * C);
            const float* ap = a + ((b * C * iters + c) * hw4 + p4) * 4;
            const long long stride = (long long)C * hw4 * 4;
            float4 m[8];
#pragma unroll
            for (int i = 0; i < 8; ++i)
                if (i < iters) m[i] = ld_stream4(ap + i * stride);
#pragma unroll
            for (int i = 0; i < 8; ++i)
                if (i < iters) {
                    v.x = curve_step(v.x, m[i].x);
                    v.y = curve_step(v.y, m[i].y);
                    v.z = curve_step(v.z, m[i].z);
                    v.w = curve_step(v.w, m[i].w);
                }
            for (int i = 8; i < iters; ++i) {
                float4 mm = ld_stream4(ap + i * stride);
                v.x = curve_step(v.x, mm.x);
                v.y = curve_step(v.y, mm.y);
                v.z = curve_step(v.z, mm.z);
                v.w = curve_step(v.w, mm.w);
            }
        }
        st_stream4(y + (plane * hw4 + p4) * 4, v);
    }
}

// scalar tail version for H*W not divisible by 4
template <bool PER_ITER>
__global__ void __launch_bounds__(256) curve_scalar_kernel(const float* __restrict__ x, const float* __restrict__ a,
                                                           float* __restrict__ y, long long planes, int C,
                                                           long long hw, int iters) {
    const long long total = planes * hw;
    for (long long it = (long long)blockIdx.x * blockDim.x + threadIdx.x; it < total;
         it += (long long)gridDim.x * blockDim.x) {
        const long long plane = it / hw;
        const long long p = it - plane * hw;
        float v = x[it];
        if (!PER_ITER) {
            const float m = a[it];
            for (int i = 0; i < iters; ++i) v = curve_step(v, m);
        } else {
            const long long b = plane / C;
            const int c = (int)(plane - b * C);
            for (int i = 0; i < iters; ++i) v = curve_step(v, a[((b * iters + i) * C + c) * hw + p]);
        }
        y[it] = v;
    }
}

template <bool PER_ITER>
static int launch_curve(const float* x, const float* a, float* y, int B, int C, int H, int W, int iters, void* stream,
                        bool strict = true) {
    if (!x || !a || !y) return B200_ERR_NULL;
    if (B <= 0 || C <= 0 || H <= 0 || W <= 0 || iters < 0) return B200_ERR_BAD_SHAPE;
    // the public ABI requires 16-byte aligned bases; internal callers (frame sub-batches of a ragged shape start
    // at 4-byte-aligned offsets) pass strict=false and fall back to the scalar kernel
    if (strict && (!aligned16(x) || !aligned16(a) || !aligned16(y))) return B200_ERR_MISALIGNED;
    if ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(y)) & 3u)
        return B200_ERR_MISALIGNED;
    cudaStream_t s = (cudaStream_t)stream;
    const long long planes = (long long)B * C;
    const long long hw = (long long)H * W;
    if (hw % 4 == 0 && aligned16(x) && aligned16(a) && aligned16(y)) {
        const long long total = planes * (hw / 4);
        long long blocks = (total + 255) / 256;
        const long long cap = (long long)B200_NUM_SMS * 16;
        if (blocks > cap) blocks = cap;
        curve_vec4_kernel<PER_ITER><<<(unsigned)blocks, 256, 0, s>>>(x, a, y, planes, C, hw / 4, iters);
    } else {
        const long long total = planes * hw;
        long long blocks = (total + 255) / 256;
        const long long cap = (long long)B200_NUM_SMS * 16;
        if (blocks > cap) blocks = cap;
        curve_scalar_kernel<PER_ITER><<<(unsigned)blocks, 256, 0, s>>>(x, a, y, planes, C, hw, iters);
    }
    return launch_status();
}

// ---- GCENet's curve loop with the brightness attention map (gcenet.py:469-484) ------------------------------------
//   bam = (1 - V)^gamma, V = max(R, G, B)                       (nn/modules/prior/image.py:269-272; kornia's HSV value)
//   iters times:  bright = e * (1 - bam);  dark = e * bam;  e = bright + dark + a * (dark^2 - dark)
// with ONE 3-channel map `a` for all iterations.  Same rounding sequence as the ATen ops (no contraction); the only
// operation that is not correctly rounded on either side is the pow.
struct GceOut {
    float* enhanced;
    float* bam;     // [B,1,H,W] or nullptr
    float* bright;  // [B,3,H,W] or nullptr: the last iteration's
    float* dark;    // [B,3,H,W] or nullptr
};

__device__ __forceinline__ void gce_step(float& e, float a, float bam, float one_minus_bam, float& bright, float& dark) {
    bright = __fmul_rn(e, one_minus_bam);
    dark = __fmul_rn(e, bam);
    e = __fadd_rn(__fadd_rn(bright, dark), __fmul_rn(a, __fsub_rn(__fmul_rn(dark, dark), dark)));
}

// VEC: one item = 4 consecutive pixels of one frame in all three channels (H*W % 4 == 0, 16-byte aligned bases)
template <bool VEC>
__global__ void __launch_bounds__(256) gce_curve_kernel(const float* __restrict__ image, const float* __restrict__ adjust,
                                                        GceOut out, long long frames, long long hw, int iters,
                                                        float gamma) {
    constexpr int V = VEC ? 4 : 1;
    const long long per = hw / V, total = frames * per;
    for (long long it = (long long)blockIdx.x * blockDim.x + threadIdx.x; it < total;
         it += (long long)gridDim.x * blockDim.x) {
        const long long f = it / per, p = (it - f * per) * V;
        const long long o = f * 3 * hw + p;
        float e[3][4], a[3][4], br[3][4], dk[3][4], bam[4];
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            if (VEC) {
                const float4 ev = ld_stream4(image + o + c * hw), av = ld_stream4(adjust + o + c * hw);
                e[c][0] = ev.x, e[c][1] = ev.y, e[c][2] = ev.z, e[c][3] = ev.w;
                a[c][0] = av.x, a[c][1] = av.y, a[c][2] = av.z, a[c][3] = av.w;
            } else {
                e[c][0] = image[o + c * hw];
                a[c][0] = adjust[o + c * hw];
            }
        }
#pragma unroll
        for (int k = 0; k < V; ++k) bam[k] = powf(__fsub_rn(1.f, fmaxf(fmaxf(e[0][k], e[1][k]), e[2][k])), gamma);
        if (VEC) {  // two pixels per instruction on the packed fp32 pipe: the same roundings as gce_step; the six
                    // (pixel pair, channel) chains advance together, one iteration at a time (a chain alone is five
                    // dependent operations per iteration)
            float2 ev[2][3], b2[2][3], d2[2][3], av[2][3], bm[2], omb[2];
#pragma unroll
            for (int k = 0; k < 2; ++k) {
                bm[k] = make_float2(bam[2 * k], bam[2 * k + 1]);
                omb[k] = make_float2(__fsub_rn(1.f, bm[k].x), __fsub_rn(1.f, bm[k].y));
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    ev[k][c] = make_float2(e[c][2 * k], e[c][2 * k + 1]);
                    av[k][c] = make_float2(a[c][2 * k], a[c][2 * k + 1]);
                    b2[k][c] = d2[k][c] = make_float2(0.f, 0.f);
                }
            }
            for (int i = 0; i < iters; ++i) {
#pragma unroll
                for (int k = 0; k < 2; ++k)
#pragma unroll
                    for (int c = 0; c < 3; ++c) {
                        b2[k][c] = __fmul2_rn(ev[k][c], omb[k]);
                        d2[k][c] = __fmul2_rn(ev[k][c], bm[k]);
                        // rn(d*d - d) as rn(rn(d*d) + (-1)*d): the FFMA2 by -1 rounds once
                        const float2 q = __ffma2_rn(d2[k][c], make_float2(-1.f, -1.f), __fmul2_rn(d2[k][c], d2[k][c]));
                        ev[k][c] = __fadd2_rn(__fadd2_rn(b2[k][c], d2[k][c]), __fmul2_rn(av[k][c], q));
                    }
            }
#pragma unroll
            for (int k = 0; k < 2; ++k)
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    e[c][2 * k] = ev[k][c].x, e[c][2 * k + 1] = ev[k][c].y;
                    br[c][2 * k] = b2[k][c].x, br[c][2 * k + 1] = b2[k][c].y;
                    dk[c][2 * k] = d2[k][c].x, dk[c][2 * k + 1] = d2[k][c].y;
                }
        } else {
            const float omb = __fsub_rn(1.f, bam[0]);
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                br[c][0] = dk[c][0] = 0.f;
                for (int i = 0; i < iters; ++i) gce_step(e[c][0], a[c][0], bam[0], omb, br[c][0], dk[c][0]);
            }
        }
        auto put = [&](float* base, const float* v, long long off) {
            if (VEC) {
                st_stream4(base + off, make_float4(v[0], v[1], v[2], v[3]));
            } else {
                base[off] = v[0];
            }
        };
        if (out.bam) put(out.bam, bam, f * hw + p);
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            put(out.enhanced, e[c], o + c * hw);
            if (out.bright) put(out.bright, br[c], o + c * hw);
            if (out.dark) put(out.dark, dk[c], o + c * hw);
        }
    }
}

// ---- u8 HWC <-> fp32 NCHW (C = 3) ------------------------------------------------
// to_image_tensor(normalize=True) -> normalize_image_by_range (core/image/photometry.py:262-263):
//   image = (image - 0.0) * ratio + 0.0   with ratio = 1/255 (a Python double) applied to an fp32 tensor,
// i.e. ONE fp32 multiply by fl32(1/255) -- not a division (the two differ in the last bit for 126 byte values).
__device__ __forceinline__ float from_u8(uint32_t byte) { return __fmul_rn((float)byte, 1.f / 255.f); }

// One thread handles 4 consecutive pixels = 12 interleaved bytes (3 x 32-bit words).
__global__ void __launch_bounds__(256) u8hwc_to_f32nchw_kernel(const uint8_t* __restrict__ src, float* __restrict__ dst,
                                                               long long frames, long long hw) {
    const long long hw4 = hw >> 2;
    const long long total = frames * hw4;
    for (long long it = (long long)blockIdx.x * blockDim.x + threadIdx.x; it < total;
         it += (long long)gridDim.x * blockDim.x) {
        const long long f = it / hw4;
        const long long p4 = it - f * hw4;
        const uint32_t* s = reinterpret_cast<const uint32_t*>(src + (f * hw + p4 * 4) * 3);
        const uint32_t w0 = __ldg(s), w1 = __ldg(s + 1), w2 = __ldg(s + 2);
        // bytes: r0 g0 b0 r1 | g1 b1 r2 g2 | b2 r3 g3 b3
        float4 r, g, b;
        r.x = from_u8(w0 & 0xff);
        g.x = from_u8((w0 >> 8) & 0xff);
        b.x = from_u8((w0 >> 16) & 0xff);
        r.y = from_u8(w0 >> 24);
        g.y = from_u8(w1 & 0xff);
        b.y = from_u8((w1 >> 8) & 0xff);
        r.z = from_u8((w1 >> 16) & 0xff);
        g.z = from_u8(w1 >> 24);
        b.z = from_u8(w2 & 0xff);
        r.w = from_u8((w2 >> 8) & 0xff);
        g.w = from_u8((w2 >> 16) & 0xff);
        b.w = from_u8(w2 >> 24);
        float* d = dst + f * 3 * hw + p4 * 4;
        st_stream4(d, r);
        st_stream4(d + hw, g);
        st_stream4(d + 2 * hw, b);
    }
}

// H*W not a multiple of 4 (frames are then not word aligned): one pixel per thread
__global__ void __launch_bounds__(256) u8hwc_to_f32nchw_scalar_kernel(const uint8_t* __restrict__ src,
                                                                      float* __restrict__ dst, long long frames,
                                                                      long long hw) {
    const long long total = frames * hw;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const long long f = i / hw, p = i - f * hw;
        const uint8_t* s = src + i * 3;
        float* d = dst + f * 3 * hw + p;
        d[0] = from_u8(s[0]);
        d[hw] = from_u8(s[1]);
        d[2 * hw] = from_u8(s[2]);
    }
}

__device__ __forceinline__ uint32_t to_u8(float v) {
    // to_image_nparray(denormalize=True): v*255 then astype(uint8) (truncation toward zero).  Values outside [0, 1]
    // are implementation-defined in the reference (C float -> uint8 cast); they are clamped here.
    float t = __fmul_rn(fminf(fmaxf(v, 0.f), 1.f), 255.f);
    return (uint32_t)(int)t;
}

__global__ void __launch_bounds__(256) f32nchw_to_u8hwc_kernel(const float* __restrict__ src, uint8_t* __restrict__ dst,
                                                               long long frames, long long hw) {
    const long long hw4 = hw >> 2;
    const long long total = frames * hw4;
    for (long long it = (long long)blockIdx.x * blockDim.x + threadIdx.x; it < total;
         it += (long long)gridDim.x * blockDim.x) {
        const long long f = it / hw4;
        const long long p4 = it - f * hw4;
        const float* s = src + f * 3 * hw + p4 * 4;
        const float4 r = ld_stream4(s), g = ld_stream4(s + hw), b = ld_stream4(s + 2 * hw);
        const uint32_t w0 = to_u8(r.x) | (to_u8(g.x) << 8) | (to_u8(b.x) << 16) | (to_u8(r.y) << 24);
        const uint32_t w1 = to_u8(g.y) | (to_u8(b.y) << 8) | (to_u8(r.z) << 16) | (to_u8(g.z) << 24);
        const uint32_t w2 = to_u8(b.z) | (to_u8(r.w) << 8) | (to_u8(g.w) << 16) | (to_u8(b.w) << 24);
        uint32_t* d = reinterpret_cast<uint32_t*>(dst + (f * hw + p4 * 4) * 3);
        d[0] = w0;
        d[1] = w1;
        d[2] = w2;
    }
}

__global__ void __launch_bounds__(256) f32nchw_to_u8hwc_scalar_kernel(const float* __restrict__ src,
                                                                      uint8_t* __restrict__ dst, long long frames,
                                                                      long long hw) {
    const long long total = frames * hw;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const long long f = i / hw, p = i - f * hw;
        const float* s = src + f * 3 * hw + p;
        uint8_t* d = dst + i * 3;
        d[0] = (uint8_t)to_u8(s[0]);
        d[1] = (uint8_t)to_u8(s[hw]);
        d[2] = (uint8_t)to_u8(s[2 * hw]);
    }
}

// internal entry for sub-batch pointers that are only float-aligned (dce.cu)
int curve_per_iter_unaligned_ok(const float* x, const float* a, float* y, int B, int C, int H, int W, int iters,
                                cudaStream_t s) {
    return launch_curve<true>(x, a, y, B, C, H, W, iters, (void*)s, false);
}


// ------------------------------------------------------------------------------------------------
// Bilinear resize, align_corners=False, no antialias: core.resize -> kornia resize -> F.interpolate(size=...,
// mode="bilinear") (core/image/geometry/transform.py:218-228), the pre/post step of VisionModel.infer
// (vision/model.py:100-121).  ATen's index math (area_pixel_compute_source_index with scale = in / out) and its
// evaluation order  h0*(w0*v00 + w1*v01) + h1*(w0*v10 + w1*v11).  One thread per output pixel, all planes of it.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
resize_bilinear_kernel(const float* __restrict__ src, float* __restrict__ dst, int planes, int Hi, int Wi, int Ho,
                       int Wo, float sh, float sw) {
    const long long n = (long long)Ho * Wo;
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n; i += (long long)gridDim.x * 256) {
        const int X = (int)(i % Wo), Y = (int)(i / Wo);
        const Tap ty = tap_half_pixel(Y, sh, Hi), tx = tap_half_pixel(X, sw, Wi);
        const size_t o00 = (size_t)ty.i0 * Wi + tx.i0, o01 = (size_t)ty.i0 * Wi + tx.i1;
        const size_t o10 = (size_t)ty.i1 * Wi + tx.i0, o11 = (size_t)ty.i1 * Wi + tx.i1;
        for (int p = 0; p < planes; ++p) {
            const float* s = src + (size_t)p * Hi * Wi;
            const float top = __fadd_rn(__fmul_rn(tx.w0, __ldg(s + o00)), __fmul_rn(tx.w1, __ldg(s + o01)));
            const float bot = __fadd_rn(__fmul_rn(tx.w0, __ldg(s + o10)), __fmul_rn(tx.w1, __ldg(s + o11)));
            dst[(size_t)p * n + i] = __fadd_rn(__fmul_rn(ty.w0, top), __fmul_rn(ty.w1, bot));
        }
    }
}

}  // namespace b200

extern "C" int b200_curve_shared(const float* x, const float* a, float* y, int B, int C, int H, int W, int iters,
                                 void* stream) {
    return b200::launch_curve<false>(x, a, y, B, C, H, W, iters, stream);
}

extern "C" int b200_curve_per_iter(const float* x, const float* a, float* y, int B, int C, int H, int W, int iters,
                                   void* stream) {
    return b200::launch_curve<true>(x, a, y, B, C, H, W, iters, stream);
}

extern "C" int b200_gce_curve(const float* image, const float* adjust, float* enhanced, float* bam, float* bright,
                              float* dark, int B, int H, int W, int iters, float gamma, void* stream) {
    if (!image || !adjust || !enhanced) return B200_ERR_NULL;
    if (B <= 0 || H <= 0 || W <= 0 || iters < 0) return B200_ERR_BAD_SHAPE;
    // gcenet.py:470-475: without a gamma the loop is the plain shared-map curve and there is no map to return
    if (gamma == 0.f) return b200::launch_curve<false>(image, adjust, enhanced, B, 3, H, W, iters, stream);
    const float* in[2] = {image, adjust};
    float* outs[4] = {enhanced, bam, bright, dark};
    bool al = true;
    for (const float* q : in) al = al && b200::aligned16(q);
    for (float* q : outs) al = al && (!q || b200::aligned16(q));
    if (!al) return B200_ERR_MISALIGNED;
    const long long hw = (long long)H * W;
    const bool vec = hw % 4 == 0;
    const long long total = (long long)B * (vec ? hw / 4 : hw);
    long long blocks = (total + 255) / 256;
    const long long cap = (long long)B200_NUM_SMS * 16;
    if (blocks > cap) blocks = cap;
    const b200::GceOut o{enhanced, bam, bright, dark};
    if (vec)
        b200::gce_curve_kernel<true><<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(image, adjust, o, B, hw, iters, gamma);
    else
        b200::gce_curve_kernel<false><<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(image, adjust, o, B, hw, iters, gamma);
    return b200::launch_status();
}

extern "C" int b200_u8hwc_to_f32nchw(const uint8_t* src, float* dst, int B, int H, int W, void* stream) {
    if (!src || !dst) return B200_ERR_NULL;
    if (B <= 0 || H <= 0 || W <= 0) return B200_ERR_BAD_SHAPE;
    const long long hw = (long long)H * W;
    if (hw % 4 != 0 || !b200::aligned16(dst) || (reinterpret_cast<uintptr_t>(src) & 3u)) {
        long long blocks = ((long long)B * hw + 255) / 256;
        if (blocks > B200_NUM_SMS * 16) blocks = B200_NUM_SMS * 16;
        b200::u8hwc_to_f32nchw_scalar_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(src, dst, B, hw);
        return b200::launch_status();
    }
    long long blocks = ((long long)B * (hw / 4) + 255) / 256;
    if (blocks > B200_NUM_SMS * 16) blocks = B200_NUM_SMS * 16;
    b200::u8hwc_to_f32nchw_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(src, dst, B, hw);
    return b200::launch_status();
}

extern "C" int b200_f32nchw_to_u8hwc(const float* src, uint8_t* dst, int B, int H, int W, void* stream) {
    if (!src || !dst) return B200_ERR_NULL;
    if (B <= 0 || H <= 0 || W <= 0) return B200_ERR_BAD_SHAPE;
    const long long hw = (long long)H * W;
    if (hw % 4 != 0 || !b200::aligned16(src) || (reinterpret_cast<uintptr_t>(dst) & 3u)) {
        long long blocks = ((long long)B * hw + 255) / 256;
        if (blocks > B200_NUM_SMS * 16) blocks = B200_NUM_SMS * 16;
        b200::f32nchw_to_u8hwc_scalar_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(src, dst, B, hw);
        return b200::launch_status();
    }
    long long blocks = ((long long)B * (hw / 4) + 255) / 256;
    if (blocks > B200_NUM_SMS * 16) blocks = B200_NUM_SMS * 16;
    b200::f32nchw_to_u8hwc_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(src, dst, B, hw);
    return b200::launch_status();
}

extern "C" int b200_resize_bilinear(const float* src, float* dst, int planes, int Hin, int Win, int Hout, int Wout,
                                    void* stream) {
    if (!src || !dst) return B200_ERR_NULL;
    if (planes <= 0 || Hin <= 0 || Win <= 0 || Hout <= 0 || Wout <= 0) return B200_ERR_BAD_SHAPE;
    // ATen area_pixel_compute_scale<float>(in, out, align_corners=false, scale=nullopt) = (float)in / out
    const float sh = (float)Hin / (float)Hout, sw = (float)Win / (float)Wout;
    long long blocks = ((long long)Hout * Wout + 255) / 256;
    if (blocks > B200_NUM_SMS * 16) blocks = B200_NUM_SMS * 16;
    b200::resize_bilinear_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(src, dst, planes, Hin, Win, Hout,
                                                                                    Wout, sh, sw);
    return b200::launch_status();
}
