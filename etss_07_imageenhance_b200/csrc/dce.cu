// Zero-DCE (ZeroDCE_RE.forward, zero_dce.py:173-191) for sm_100a — host dispatch and the fp32 path.
//
// precision = B200_PREC_FP32: direct 3x3 convolution on CUDA cores (packed FFMA2, fp32 accumulate), one launch
//   per layer, torch.cat never materialised (two source pointers), tanh fused into layer 7, then the
//   per-iteration curve kernel.  This is the <=1e-3 parity path of BASELINE.json's "fp32 mode".
// precision = B200_PREC_BF16 / BF16X3: the tcgen05 implicit-GEMM path in dce_tc.cu.
#include "common.cuh"
#include "dce_tc.cuh"
#include "dce_x3.cuh"
#include "tcgen05.cuh"

namespace b200 {

// ------------------------------------------------------------------------------------------------
// fp32 packed weights: per layer  w[cin][9][32] | b[32]   (cout zero-padded to 32)
// ------------------------------------------------------------------------------------------------
__host__ __device__ constexpr int dce_cin(int l) { return l == 0 ? 3 : (l <= 3 ? 32 : 64); }
__host__ __device__ constexpr int dce_f32_layer_floats(int l) { return dce_cin(l) * 9 * 32 + 32; }
__host__ __device__ constexpr int dce_f32_layer_offset(int l) {
    int o = 0;
    for (int i = 0; i < l; ++i) o += dce_f32_layer_floats(i);
    return o;
}
constexpr int DCE_F32_TOTAL = dce_f32_layer_offset(7);

struct DcePackArgs {
    const float* w[7];
    const float* b[7];
};

__global__ void dce_pack_f32_kernel(DcePackArgs a, float* __restrict__ packed, int cout_last) {
    const int l = blockIdx.x;
    const int cin = dce_cin(l);
    const int cout = l == 6 ? cout_last : 32;
    float* out = packed + dce_f32_layer_offset(l);
    const int nw = cin * 9 * 32;
    for (int i = threadIdx.x; i < nw; i += blockDim.x) {
        const int o = i & 31;
        const int tap = (i >> 5) % 9;
        const int c = (i >> 5) / 9;
        out[i] = o < cout ? a.w[l][(o * cin + c) * 9 + tap] : 0.f;  // OIHW -> [cin][tap][cout]
    }
    for (int i = threadIdx.x; i < 32; i += blockDim.x) out[nw + i] = i < cout ? a.b[l][i] : 0.f;
}

// ------------------------------------------------------------------------------------------------
// fp32 direct conv3x3.  CTA tile = 8 rows x 32 cols, 256 threads; thread = 4 px (along x) x 8 couts.
// Input channels are streamed in chunks of 8 through shared memory (halo tile + that chunk's weights).
// ------------------------------------------------------------------------------------------------
constexpr int C_TH = 8, C_TW = 32, C_HS = 36, C_CS = 8;

template <bool TANH>
__global__ void __launch_bounds__(256, 2)
dce_conv_f32_kernel(const float* __restrict__ srcA, int ca, const float* __restrict__ srcB, int cb,
                    float* __restrict__ dst, const float* __restrict__ wl, int cout, int H, int W) {
    // two stages: the next chunk's halo tile and weights arrive by cp.async while this one is computed
    __shared__ __align__(16) float s_in_all[2][C_CS * 10 * C_HS];
    __shared__ __align__(16) float s_w_all[2][C_CS * 9 * 32];
    const int t = threadIdx.x;
    const int b = blockIdx.z;
    const int y0 = blockIdx.y * C_TH, x0 = blockIdx.x * C_TW;
    const int cin = ca + cb;
    const size_t plane = (size_t)H * W;

    const int lane = t & 31, warp = t >> 5;
    const int pg = (warp & 1) * 32 + lane;
    const int og = warp >> 1;
    const int r = pg >> 3, c4 = (pg & 7) * 4;
    const bool active = og * 8 < cout;  // warp-uniform

    float2 acc[4][4];
#pragma unroll
    for (int p = 0; p < 4; ++p)
#pragma unroll
        for (int o = 0; o < 4; ++o) acc[p][o] = make_float2(0.f, 0.f);

    // staging map, the same for every chunk: warp k copies channel c0 + k, a lane the elements lane, lane + 32, ...
    // of its 10 x 34 halo tile (offset -1 = outside the image: zero fill)
    int goff[11];
#pragma unroll
    for (int i = 0; i < 11; ++i) {
        const int e = lane + 32 * i;
        const int rr = e / 34, cc = e - rr * 34;
        const int yy = y0 - 1 + rr, xx = x0 - 1 + cc;
        goff[i] = (e < 340 && yy >= 0 && yy < H && xx >= 0 && xx < W) ? yy * W + xx : -1;
    }
    auto stage = [&](int c0, int buf) {
        const int nvalid = min(C_CS, cin - c0);
        const int c = c0 + warp;
        const bool have = warp < nvalid;
        const float* src = srcA;
        if (have) src = c < ca ? srcA + ((size_t)b * ca + c) * plane : srcB + ((size_t)b * cb + (c - ca)) * plane;
        const uint32_t sdst = tc::smem_u32(&s_in_all[buf][warp * 10 * C_HS]);
#pragma unroll
        for (int i = 0; i < 11; ++i) {
            const int e = lane + 32 * i;
            if (e < 340) {
                const bool ok = have && goff[i] >= 0;
                tc::cp_async4(sdst + 4 * (e + 2 * (e / 34)), ok ? src + goff[i] : src, ok);  // (e / 34) * 36 + e % 34
            }
        }
        const float* wsrc = wl + (size_t)c0 * 9 * 32;
        const uint32_t wdst = tc::smem_u32(&s_w_all[buf][0]);
        const int n4 = nvalid * 9 * 32 / 4;
        for (int i = t; i < n4; i += 256) tc::cp_async16(wdst + 16 * i, wsrc + 4 * i, true);
        tc::cp_async_commit();
    };

    stage(0, 0);
    int it = 0;
    for (int c0 = 0; c0 < cin; c0 += C_CS, ++it) {
        const int nvalid = min(C_CS, cin - c0);
        if (c0 + C_CS < cin) {
            stage(c0 + C_CS, (it + 1) & 1);  // its buffer was released by the barrier that ended chunk it - 1
            tc::cp_async_wait<1>();
        } else {
            tc::cp_async_wait<0>();
        }
        __syncthreads();
        const float* s_in = s_in_all[it & 1];
        const float* s_w = s_w_all[it & 1];
        if (active) {
#pragma unroll 2
            for (int k = 0; k < nvalid; ++k) {
#pragma unroll
                for (int ky = 0; ky < 3; ++ky) {
                    const float* ip = s_in + (k * 10 + r + ky) * C_HS + c4;
                    const float4 i0 = *reinterpret_cast<const float4*>(ip);
                    const float2 i1 = *reinterpret_cast<const float2*>(ip + 4);
                    const float v[6] = {i0.x, i0.y, i0.z, i0.w, i1.x, i1.y};
#pragma unroll
                    for (int kx = 0; kx < 3; ++kx) {
                        const float* wp = s_w + (k * 9 + ky * 3 + kx) * 32 + og * 8;
                        const float4 w0 = *reinterpret_cast<const float4*>(wp);
                        const float4 w1 = *reinterpret_cast<const float4*>(wp + 4);
                        const float2 wv[4] = {make_float2(w0.x, w0.y), make_float2(w0.z, w0.w),
                                              make_float2(w1.x, w1.y), make_float2(w1.z, w1.w)};
#pragma unroll
                        for (int p = 0; p < 4; ++p) {
                            const float2 aa = make_float2(v[p + kx], v[p + kx]);
#pragma unroll
                            for (int o = 0; o < 4; ++o) acc[p][o] = __ffma2_rn(aa, wv[o], acc[p][o]);
                        }
                    }
                }
            }
        }
        __syncthreads();  // this chunk's buffer is free for the stage issued two iterations on
    }

    const int yy = y0 + r, xx = x0 + c4;
    if (!active || yy >= H || xx >= W) return;
    const float* bias = wl + (size_t)cin * 9 * 32;
#pragma unroll
    for (int o = 0; o < 8; ++o) {
        const int oc = og * 8 + o;
        if (oc >= cout) break;
        const float bv = __ldg(bias + oc);
        float v[4];
#pragma unroll
        for (int p = 0; p < 4; ++p) {
            const float u = ((o & 1) ? acc[p][o >> 1].y : acc[p][o >> 1].x) + bv;
            v[p] = TANH ? tanhf(u) : fmaxf(u, 0.f);
        }
        float* d = dst + ((size_t)b * cout + oc) * plane + (size_t)yy * W + xx;
        if ((W & 3) == 0 && xx + 3 < W) {
            *reinterpret_cast<float4*>(d) = make_float4(v[0], v[1], v[2], v[3]);
        } else {
#pragma unroll
            for (int p = 0; p < 4; ++p)
                if (xx + p < W) d[p] = v[p];
        }
    }
}

static int launch_conv_f32(const float* a, int ca, const float* b2, int cb, float* dst, const float* wl, int cout,
                           bool tanh_out, int B, int H, int W, cudaStream_t s) {
    dim3 grid((W + C_TW - 1) / C_TW, (H + C_TH - 1) / C_TH, B);
    if (tanh_out)
        dce_conv_f32_kernel<true><<<grid, 256, 0, s>>>(a, ca, b2, cb, dst, wl, cout, H, W);
    else
        dce_conv_f32_kernel<false><<<grid, 256, 0, s>>>(a, ca, b2, cb, dst, wl, cout, H, W);
    return launch_status();
}

static size_t dce_align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }
constexpr int DCE_F32_FRAMES_PER_PASS = 2;

static int dce_forward_f32(const float* x, const float* packed, float* adjust, float* enhanced, int B, int H, int W,
                           int iters, void* workspace, size_t workspace_bytes, cudaStream_t s) {
    const int fb = B < DCE_F32_FRAMES_PER_PASS ? B : DCE_F32_FRAMES_PER_PASS;
    const size_t plane = (size_t)H * W;
    const size_t fmap = dce_align_up((size_t)fb * 32 * plane * sizeof(float), 256);
    const size_t amap = dce_align_up((size_t)fb * 3 * iters * plane * sizeof(float), 256);
    if (workspace_bytes < 5 * fmap + amap) return B200_ERR_WORKSPACE;
    char* ws = (char*)workspace;
    float* x1 = (float*)(ws + 0 * fmap);
    float* x2 = (float*)(ws + 1 * fmap);
    float* x3 = (float*)(ws + 2 * fmap);
    float* x4 = (float*)(ws + 3 * fmap);
    float* x5 = (float*)(ws + 4 * fmap);
    float* x6 = x4;
    float* a_ws = (float*)(ws + 5 * fmap);
    const int cout7 = 3 * iters;
    for (int b0 = 0; b0 < B; b0 += fb) {
        const int nb = B - b0 < fb ? B - b0 : fb;
        const float* xb = x + (size_t)b0 * 3 * plane;
        float* ab = adjust ? adjust + (size_t)b0 * cout7 * plane : a_ws;
        int rc;
        if ((rc = launch_conv_f32(xb, 3, nullptr, 0, x1, packed + dce_f32_layer_offset(0), 32, false, nb, H, W, s))) return rc;
        if ((rc = launch_conv_f32(x1, 32, nullptr, 0, x2, packed + dce_f32_layer_offset(1), 32, false, nb, H, W, s))) return rc;
        if ((rc = launch_conv_f32(x2, 32, nullptr, 0, x3, packed + dce_f32_layer_offset(2), 32, false, nb, H, W, s))) return rc;
        if ((rc = launch_conv_f32(x3, 32, nullptr, 0, x4, packed + dce_f32_layer_offset(3), 32, false, nb, H, W, s))) return rc;
        if ((rc = launch_conv_f32(x3, 32, x4, 32, x5, packed + dce_f32_layer_offset(4), 32, false, nb, H, W, s))) return rc;
        if ((rc = launch_conv_f32(x2, 32, x5, 32, x6, packed + dce_f32_layer_offset(5), 32, false, nb, H, W, s))) return rc;
        if ((rc = launch_conv_f32(x1, 32, x6, 32, ab, packed + dce_f32_layer_offset(6), cout7, true, nb, H, W, s))) return rc;
        if ((rc = curve_per_iter_unaligned_ok(xb, ab, enhanced + (size_t)b0 * 3 * plane, nb, 3, H, W, iters, s))) return rc;
    }
    return B200_OK;
}

// ------------------------------------------------------------------------------------------------
// ZeroDCE-V core (ZeroDCEV.forward, zero_dce_v.py:172-197): V channel of the RGB image, nearest resize
// to down_size x down_size, DCE-Net with in_channels = 1 -> num_iters (15) curve maps, num_iters curve steps on V.
// The convolutions are the fp32 CUDA-core kernel above (its source / output channel counts are run-time values);
// packed layout as for Zero-DCE with cin(layer 1) = 1.
// ------------------------------------------------------------------------------------------------
__host__ __device__ constexpr int dcev_cin(int l) { return l == 0 ? 1 : (l <= 3 ? 32 : 64); }
__host__ __device__ constexpr int dcev_layer_offset(int l) {
    int o = 0;
    for (int i = 0; i < l; ++i) o += dcev_cin(i) * 9 * 32 + 32;
    return o;
}
constexpr int DCEV_TOTAL = dcev_layer_offset(7);

__global__ void dcev_pack_kernel(DcePackArgs a, float* __restrict__ packed, int cout_last) {
    const int l = blockIdx.x;
    const int cin = dcev_cin(l);
    const int cout = l == 6 ? cout_last : 32;
    float* out = packed + dcev_layer_offset(l);
    const int nw = cin * 9 * 32;
    for (int i = threadIdx.x; i < nw; i += blockDim.x) {
        const int o = i & 31;
        const int tap = (i >> 5) % 9;
        const int c = (i >> 5) / 9;
        out[i] = o < cout ? a.w[l][(o * cin + c) * 9 + tap] : 0.f;  // OIHW -> [cin][tap][cout]
    }
    for (int i = threadIdx.x; i < 32; i += blockDim.x) out[nw + i] = i < cout ? a.b[l][i] : 0.f;
}

// image_v = max(R, G, B) (kornia rgb_to_hsv: v = image.max(-3)) at full resolution, and its nearest-neighbour
// resize to S x S (F.interpolate(size=(S, S)), mode "nearest": src = min(floor(dst * scale), in - 1),
// scale = (float)in / out, ATen nearest_neighbor_compute_source_index)
__global__ void __launch_bounds__(256)
dcev_v_kernel(const float* __restrict__ rgb, float* __restrict__ v, long long n, long long plane) {
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n; i += (long long)gridDim.x * 256) {
        const long long b = i / plane, p = i - b * plane;
        const float* s = rgb + b * 3 * plane + p;
        v[i] = fmaxf(fmaxf(s[0], s[plane]), s[2 * plane]);
    }
}
__global__ void __launch_bounds__(256)
dcev_nearest_kernel(const float* __restrict__ v, float* __restrict__ v_lr, int B, int H, int W, int S, float sh, float sw) {
    const long long n = (long long)B * S * S;
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n; i += (long long)gridDim.x * 256) {
        const int X = (int)(i % S), Y = (int)((i / S) % S);
        const long long b = i / ((long long)S * S);
        const int ys = min((int)floorf(__fmul_rn((float)Y, sh)), H - 1);
        const int xs = min((int)floorf(__fmul_rn((float)X, sw)), W - 1);
        v_lr[i] = v[(b * H + ys) * (long long)W + xs];
    }
}

}  // namespace b200

using namespace b200;

extern "C" size_t b200_dcev_packed_floats(void) { return (size_t)DCEV_TOTAL; }

extern "C" int b200_dcev_pack_weights(const float* const* w, const float* const* b, float* packed, int iters,
                                      void* stream) {
    if (!w || !b || !packed) return B200_ERR_NULL;
    if (iters < 1 || iters > 32) return B200_ERR_UNSUPPORTED;
    DcePackArgs a;
    for (int l = 0; l < 7; ++l) {
        if (!w[l] || !b[l]) return B200_ERR_NULL;
        a.w[l] = w[l];
        a.b[l] = b[l];
    }
    dcev_pack_kernel<<<7, 256, 0, (cudaStream_t)stream>>>(a, packed, iters);
    return launch_status();
}

extern "C" size_t b200_dcev_workspace_bytes(int B, int S) {
    if (B <= 0 || S <= 0) return 0;
    return 5 * dce_align_up((size_t)B * 32 * S * S * sizeof(float), 256);
}

extern "C" int b200_dcev_core(const float* image_rgb, const float* packed, float* image_v, float* image_v_lr,
                              float* adjust, float* image_v_fixed_lr, int B, int H, int W, int S, int iters,
                              void* workspace, size_t workspace_bytes, void* stream) {
    if (!image_rgb || !packed || !image_v || !image_v_lr || !adjust || !image_v_fixed_lr || !workspace)
        return B200_ERR_NULL;
    if (B <= 0 || H <= 0 || W <= 0 || S <= 0) return B200_ERR_BAD_SHAPE;
    if (iters < 1 || iters > 32 || B > 65535) return B200_ERR_UNSUPPORTED;
    if (!aligned16(image_rgb) || !aligned16(packed) || !aligned16(image_v) || !aligned16(image_v_lr) ||
        !aligned16(adjust) || !aligned16(image_v_fixed_lr) || !aligned16(workspace))
        return B200_ERR_MISALIGNED;
    if (workspace_bytes < b200_dcev_workspace_bytes(B, S)) return B200_ERR_WORKSPACE;
    cudaStream_t s = (cudaStream_t)stream;
    const long long plane = (long long)H * W, n = (long long)B * plane;
    long long blocks = (n + 255) / 256;
    if (blocks > B200_NUM_SMS * 16) blocks = B200_NUM_SMS * 16;
    dcev_v_kernel<<<(unsigned)blocks, 256, 0, s>>>(image_rgb, image_v, n, plane);
    blocks = ((long long)B * S * S + 255) / 256;
    if (blocks > B200_NUM_SMS * 16) blocks = B200_NUM_SMS * 16;
    // ATen: scale = (float)input_size / output_size (no scale_factor given)
    dcev_nearest_kernel<<<(unsigned)blocks, 256, 0, s>>>(image_v, image_v_lr, B, H, W, S, (float)H / (float)S,
                                                        (float)W / (float)S);
    int rc = launch_status();
    if (rc) return rc;
    const size_t fmap = dce_align_up((size_t)B * 32 * S * S * sizeof(float), 256);
    char* ws = (char*)workspace;
    float* x1 = (float*)(ws + 0 * fmap);
    float* x2 = (float*)(ws + 1 * fmap);
    float* x3 = (float*)(ws + 2 * fmap);
    float* x4 = (float*)(ws + 3 * fmap);
    float* x5 = (float*)(ws + 4 * fmap);
    float* x6 = x4;
    auto wl = [&](int l) { return packed + dcev_layer_offset(l); };
    if ((rc = launch_conv_f32(image_v_lr, 1, nullptr, 0, x1, wl(0), 32, false, B, S, S, s))) return rc;
    if ((rc = launch_conv_f32(x1, 32, nullptr, 0, x2, wl(1), 32, false, B, S, S, s))) return rc;
    if ((rc = launch_conv_f32(x2, 32, nullptr, 0, x3, wl(2), 32, false, B, S, S, s))) return rc;
    if ((rc = launch_conv_f32(x3, 32, nullptr, 0, x4, wl(3), 32, false, B, S, S, s))) return rc;
    if ((rc = launch_conv_f32(x3, 32, x4, 32, x5, wl(4), 32, false, B, S, S, s))) return rc;
    if ((rc = launch_conv_f32(x2, 32, x5, 32, x6, wl(5), 32, false, B, S, S, s))) return rc;
    if ((rc = launch_conv_f32(x1, 32, x6, 32, adjust, wl(6), iters, true, B, S, S, s))) return rc;
    return curve_per_iter_unaligned_ok(image_v_lr, adjust, image_v_fixed_lr, B, 1, S, S, iters, s);
}

extern "C" size_t b200_dce_packed_bytes(int precision, int iters) {
    if (iters < 1 || iters > 8) return 0;
    if (precision == B200_PREC_FP32) return (size_t)DCE_F32_TOTAL * sizeof(float);
    if (precision == B200_PREC_BF16X3_MMA) return dce_x3_packed_bytes(iters);
    return dce_tc_packed_bytes(precision, iters);
}

extern "C" int b200_dce_pack_weights(const float* const* w, const float* const* b, void* packed, int precision,
                                     int iters, void* stream) {
    if (!w || !b || !packed) return B200_ERR_NULL;
    if (iters < 1 || iters > 8) return B200_ERR_UNSUPPORTED;
    for (int l = 0; l < 7; ++l)
        if (!w[l] || !b[l]) return B200_ERR_NULL;
    if (!aligned16(packed)) return B200_ERR_MISALIGNED;
    if (precision == B200_PREC_FP32) {
        DcePackArgs a;
        for (int l = 0; l < 7; ++l) {
            a.w[l] = w[l];
            a.b[l] = b[l];
        }
        dce_pack_f32_kernel<<<7, 256, 0, (cudaStream_t)stream>>>(a, (float*)packed, 3 * iters);
        return launch_status();
    }
    if (precision == B200_PREC_BF16X3_MMA) return dce_x3_pack_weights(w, b, packed, iters, (cudaStream_t)stream);
    return dce_tc_pack_weights(w, b, packed, precision, iters, (cudaStream_t)stream);
}

extern "C" size_t b200_dce_workspace_bytes(int B, int H, int W, int precision) {
    if (B <= 0 || H <= 0 || W <= 0) return 0;
    if (precision == B200_PREC_FP32) {
        const size_t fb = B < DCE_F32_FRAMES_PER_PASS ? B : DCE_F32_FRAMES_PER_PASS;
        const size_t plane = (size_t)H * W;
        const size_t fmap = dce_align_up(fb * 32 * plane * sizeof(float), 256);
        const size_t amap = dce_align_up(fb * 24 * plane * sizeof(float), 256);
        return 5 * fmap + amap;
    }
    if (precision == B200_PREC_BF16X3_MMA) return dce_x3_workspace_bytes(B, H, W);
    return dce_tc_workspace_bytes(B, H, W, precision);
}

extern "C" int b200_dce_forward(const float* x, const void* packed, float* adjust, float* enhanced, int B, int H,
                                int W, int iters, int precision, void* workspace, size_t workspace_bytes,
                                void* stream) {
    if (!x || !packed || !enhanced || !workspace) return B200_ERR_NULL;
    if (B <= 0 || H <= 0 || W <= 0) return B200_ERR_BAD_SHAPE;
    if (iters < 1 || iters > 8) return B200_ERR_UNSUPPORTED;
    if (B > 65535) return B200_ERR_UNSUPPORTED;
    if (!aligned16(x) || !aligned16(packed) || !aligned16(enhanced) || (adjust && !aligned16(adjust)) ||
        !aligned16(workspace))
        return B200_ERR_MISALIGNED;
    if (precision == B200_PREC_FP32)
        return dce_forward_f32(x, (const float*)packed, adjust, enhanced, B, H, W, iters, workspace, workspace_bytes,
                               (cudaStream_t)stream);
    if (precision == B200_PREC_BF16X3_MMA)
        return dce_x3_forward(x, packed, adjust, enhanced, B, H, W, iters, workspace, workspace_bytes,
                              (cudaStream_t)stream);
    return dce_tc_forward(x, packed, adjust, enhanced, B, H, W, iters, precision, workspace, workspace_bytes,
                          (cudaStream_t)stream);
}
