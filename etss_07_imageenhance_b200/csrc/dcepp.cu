// Zero-DCE++ (ZeroDCEpp_RE.forward, zero_dcepp.py:182-208) for sm_100a.
//
// Two stages:
//   K1  low-resolution curve-estimation net: 7 x [depthwise 3x3 -> pointwise 1x1] (+ReLU / tanh), one
//       launch per layer; the bilinear downsample (align_corners=False) is fused into layer 1's tile load.
//       fp32 CUDA-core math: depthwise as a register sliding window, pointwise as an smem-staged
//       register-tiled GEMM on packed FFMA2 (fma.rn.f32x2).  Feature maps (B x 32 x h x w) stay L2-resident
//       at scale_factor=12.
//   K2  full-resolution pass (the HBM-bound one): reads x once (128-bit), bilinear-upsamples the low-res map
//       (align_corners=True) from shared memory, applies the 8-iteration curve, writes `enhanced` and
//       `adjust` once each.  Persistent CTAs, grid = multiple of the SM count.
#include <stdlib.h>

#include <mutex>

#include "common.cuh"
#include "dcepp_mma.cuh"

namespace b200 {

// ------------------------------------------------------------------------------------------------
// packed weight layout (floats):  for each layer l:
//   dw_w[Cin][9] | dw_b[Cin] | (pad to a multiple of 4) | pw_wT[Cin][CoutPad] | pw_b[CoutPad]
// ------------------------------------------------------------------------------------------------
struct PpLayer {
    int cin, cout, cout_pad;
};
__host__ __device__ constexpr PpLayer pp_layer(int l) {
    return l == 0 ? PpLayer{3, 32, 32} : (l <= 3 ? PpLayer{32, 32, 32} : (l <= 5 ? PpLayer{64, 32, 32} : PpLayer{64, 3, 4}));
}
__host__ __device__ constexpr int pp_layer_floats(int l) {
    return (pp_layer(l).cin * 10 + 3) / 4 * 4 + pp_layer(l).cin * pp_layer(l).cout_pad + pp_layer(l).cout_pad;
}
__host__ __device__ constexpr int pp_layer_offset(int l) {
    int o = 0;
    for (int i = 0; i < l; ++i) o += (pp_layer_floats(i) + 3) / 4 * 4;
    return o;
}
constexpr int PP_TOTAL_FLOATS = pp_layer_offset(7);

struct PackArgs {
    const float* dw_w[7];
    const float* dw_b[7];
    const float* pw_w[7];
    const float* pw_b[7];
};

__global__ void dcepp_pack_kernel(PackArgs args, float* __restrict__ packed) {
    const int l = blockIdx.x;
    const PpLayer L = pp_layer(l);
    float* out = packed + pp_layer_offset(l);
    const int n_dw = L.cin * 9, n_db = L.cin, n_pw = L.cin * L.cout_pad, n_pb = L.cout_pad;
    for (int i = threadIdx.x; i < n_dw; i += blockDim.x) out[i] = args.dw_w[l][i];
    out += n_dw;
    for (int i = threadIdx.x; i < n_db; i += blockDim.x) out[i] = args.dw_b[l][i];
    out = packed + pp_layer_offset(l) + (L.cin * 10 + 3) / 4 * 4;
    for (int i = threadIdx.x; i < n_pw; i += blockDim.x) {
        const int k = i / L.cout_pad, o = i - k * L.cout_pad;
        out[i] = o < L.cout ? args.pw_w[l][o * L.cin + k] : 0.f;  // [Cout,Cin,1,1] -> [Cin][CoutPad]
    }
    out += n_pw;
    for (int i = threadIdx.x; i < n_pb; i += blockDim.x) out[i] = i < L.cout ? args.pw_b[l][i] : 0.f;
    // layers 2..7 also get the tensor-core layout (depthwise taps + split-bf16 pointwise fragments)
    if (l >= 1)
        mm_pack_layer(l, args.dw_w[l], args.dw_b[l], args.pw_w[l], args.pw_b[l],
                      reinterpret_cast<uint32_t*>(packed + PP_TOTAL_FLOATS) + mm_offset(l));
}

// ------------------------------------------------------------------------------------------------
// K1: one depthwise-separable layer.  CTA tile = 8 rows x 32 cols of output pixels, 256 threads.
//   NSRC sources of CS channels each (torch.cat is never materialised), NOUT outputs per thread
//   (COUT_PAD = 4*NOUT: 32 for the hidden layers, 4 for the 3-channel last layer).
// ------------------------------------------------------------------------------------------------
constexpr int K1_TH = 8, K1_TW = 32, K1_HS = 36;  // halo row stride (34 used)

template <int CS, int NSRC, int NOUT, bool FIRST, bool TANH>
__global__ void __launch_bounds__(256, 2)
dcepp_layer_kernel(const float* __restrict__ srcA, const float* __restrict__ srcB, float* __restrict__ dst,
                   const float* __restrict__ wl, int h, int w, int H, int W, float sf_scale, int* __restrict__ sync) {
    constexpr int CIN = CS * NSRC;
    constexpr int COUT_PAD = 4 * NOUT;
    constexpr int COUT = TANH ? 3 : 32;
    extern __shared__ __align__(16) float smem[];
    float* s_dww = smem;                       // [CIN][9]
    float* s_dwb = s_dww + CIN * 9;            // [CIN]
    constexpr int PW_OFF = (CIN * 10 + 3) / 4 * 4;
    float* s_pww = smem + PW_OFF;              // [CIN][COUT_PAD]   (16-byte aligned)
    float* s_pwb = s_pww + CIN * COUT_PAD;     // [COUT_PAD]
    float* s_halo = s_pwb + COUT_PAD;          // [CS][10][K1_HS]
    float* s_dw = s_halo + CS * 10 * K1_HS;    // [CS][256]

    const int t = threadIdx.x;
    const int b = blockIdx.z;
    const int y0 = blockIdx.y * K1_TH, x0 = blockIdx.x * K1_TW;

    // layer 1 shares its tile grid with dcepp_net_kernel: clear that launch's work counter and this tile's five
    // "published" flags (the net launch follows this one in stream order)
    if (FIRST && sync != nullptr) {
        static_assert(K1_TH == MM_TH && K1_TW == MM_TW, "layer 1 and the net kernel tile the map identically");
        const int tpf = gridDim.x * gridDim.y, tile = (b * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
        const int flag0 = mm_flag0(gridDim.z);
        if (t < 5) sync[flag0 + t * (gridDim.z * tpf) + tile] = 0;
        if (tile == 0)  // counters, abort word, per-SM CTA counts, per-frame "layer 7 done" counts
            for (int i = t; i < flag0; i += 256) sync[i] = 0;
    }

    // weights -> smem  (layout in global == layout in smem, except the 16B alignment gap for CIN=3)
    {
        constexpr int NW = CIN * 9 + CIN;
        for (int i = t; i < NW; i += 256) s_dww[i] = wl[i];
        for (int i = t; i < CIN * COUT_PAD + COUT_PAD; i += 256) s_pww[i] = wl[PW_OFF + i];
    }

    // pointwise accumulators: 4 px x NOUT outputs, packed in pairs along the output dim
    float2 acc[4][(NOUT + 1) / 2];
#pragma unroll
    for (int p = 0; p < 4; ++p)
#pragma unroll
        for (int o = 0; o < (NOUT + 1) / 2; ++o) acc[p][o] = make_float2(0.f, 0.f);

    const int lane = t & 31, warp = t >> 5;
    const int pg = (warp & 1) * 32 + lane;  // pixel group (4 px along x)
    const int og = warp >> 1;               // output group

#pragma unroll 1
    for (int s = 0; s < NSRC; ++s) {
        const float* src = s == 0 ? srcA : srcB;
        __syncthreads();  // previous chunk's s_halo / s_dw consumers are done; weights visible on first pass
        // ---- phase 1: halo tile load (zero padding outside the low-res image) ----
        if (FIRST && (H != h || W != w)) {
            // F.interpolate(scale_factor=1/sf, mode="bilinear"), align_corners=False (zero_dcepp.py:188): four
            // full-resolution taps per low-res pixel, gathered straight from HBM.  All of a thread's taps (4 elements
            // x 4 loads) are issued before the first one is used, so one DRAM round trip covers the whole tile.
            constexpr int N_EL = CS * 10 * 34, PER = (N_EL + 255) / 256;
            float v00[PER], v01[PER], v10[PER], v11[PER], wy0[PER], wy1[PER], wx0[PER], wx1[PER];
#pragma unroll
            for (int e = 0; e < PER; ++e) {
                const int idx = t + e * 256;
                const int k = idx / 340, rem = idx - k * 340;
                const int rr = rem / 34, cc = rem - rr * 34;
                const int yy = y0 - 1 + rr, xx = x0 - 1 + cc;
                v00[e] = v01[e] = v10[e] = v11[e] = 0.f;
                wy0[e] = wy1[e] = wx0[e] = wx1[e] = 0.f;
                if (idx < N_EL && yy >= 0 && yy < h && xx >= 0 && xx < w) {
                    const Tap ty = tap_half_pixel(yy, sf_scale, H), tx = tap_half_pixel(xx, sf_scale, W);
                    const float* p = src + ((size_t)(b * CS + k) * H) * W;
                    v00[e] = __ldg(p + (size_t)ty.i0 * W + tx.i0);
                    v01[e] = __ldg(p + (size_t)ty.i0 * W + tx.i1);
                    v10[e] = __ldg(p + (size_t)ty.i1 * W + tx.i0);
                    v11[e] = __ldg(p + (size_t)ty.i1 * W + tx.i1);
                    wy0[e] = ty.w0, wy1[e] = ty.w1, wx0[e] = tx.w0, wx1[e] = tx.w1;
                }
            }
#pragma unroll
            for (int e = 0; e < PER; ++e) {
                const int idx = t + e * 256;
                if (idx < N_EL) {
                    const int k = idx / 340, rem = idx - k * 340;
                    const int rr = rem / 34, cc = rem - rr * 34;
                    s_halo[(k * 10 + rr) * K1_HS + cc] =
                        wy0[e] * (wx0[e] * v00[e] + wx1[e] * v01[e]) + wy1[e] * (wx0[e] * v10[e] + wx1[e] * v11[e]);
                }
            }
        } else {
            for (int idx = t; idx < CS * 10 * 34; idx += 256) {
                const int k = idx / 340;
                const int rem = idx - k * 340;
                const int rr = rem / 34, cc = rem - rr * 34;
                const int yy = y0 - 1 + rr, xx = x0 - 1 + cc;
                float v = 0.f;
                if (yy >= 0 && yy < h && xx >= 0 && xx < w) v = __ldg(src + ((size_t)(b * CS + k) * h + yy) * w + xx);
                s_halo[(k * 10 + rr) * K1_HS + cc] = v;
            }
        }
        __syncthreads();
        // ---- phase 2: depthwise 3x3 (+bias), sliding window down the 8 tile rows ----
        {
            constexpr int KPG = (CS + 7) / 8;  // channels per thread group
            const int c = t & 31, g = t >> 5;
#pragma unroll
            for (int kk = 0; kk < KPG; ++kk) {
                const int k = g * KPG + kk;
                if (k < CS) {
                    const float* wk = s_dww + (s * CS + k) * 9;
                    float wr[9];
#pragma unroll
                    for (int i = 0; i < 9; ++i) wr[i] = wk[i];
                    const float bias = s_dwb[s * CS + k];
                    float o[8];
#pragma unroll
                    for (int r = 0; r < 8; ++r) o[r] = bias;
                    const float* hp = s_halo + k * 10 * K1_HS + c;
#pragma unroll
                    for (int rr = 0; rr < 10; ++rr) {
                        const float v0 = hp[rr * K1_HS], v1 = hp[rr * K1_HS + 1], v2 = hp[rr * K1_HS + 2];
#pragma unroll
                        for (int dy = 0; dy < 3; ++dy) {
                            const int r = rr - dy;
                            if (r >= 0 && r < 8) {
                                o[r] = fmaf(wr[dy * 3 + 0], v0, o[r]);
                                o[r] = fmaf(wr[dy * 3 + 1], v1, o[r]);
                                o[r] = fmaf(wr[dy * 3 + 2], v2, o[r]);
                            }
                        }
                    }
#pragma unroll
                    for (int r = 0; r < 8; ++r) s_dw[k * 256 + r * 32 + c] = o[r];
                }
            }
        }
        __syncthreads();
        // ---- phase 3: pointwise 1x1 as a register-tiled GEMM: [256 px x CS] . [CS x COUT_PAD] ----
        {
            const float* ap = s_dw + pg * 4;
            const float* wp = s_pww + (s * CS) * COUT_PAD + og * NOUT;
#pragma unroll 4
            for (int k = 0; k < CS; ++k) {
                const float4 a = *reinterpret_cast<const float4*>(ap + k * 256);
                if (NOUT == 8) {
                    const float4 w0 = *reinterpret_cast<const float4*>(wp + k * COUT_PAD);
                    const float4 w1 = *reinterpret_cast<const float4*>(wp + k * COUT_PAD + 4);
                    const float2 wv[4] = {make_float2(w0.x, w0.y), make_float2(w0.z, w0.w), make_float2(w1.x, w1.y),
                                          make_float2(w1.z, w1.w)};
                    const float av[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
                    for (int p = 0; p < 4; ++p) {
                        const float2 aa = make_float2(av[p], av[p]);
#pragma unroll
                        for (int o = 0; o < 4; ++o) acc[p][o] = __ffma2_rn(aa, wv[o], acc[p][o]);
                    }
                } else {
                    const float wv = wp[k * COUT_PAD];
                    acc[0][0].x = fmaf(a.x, wv, acc[0][0].x);
                    acc[1][0].x = fmaf(a.y, wv, acc[1][0].x);
                    acc[2][0].x = fmaf(a.z, wv, acc[2][0].x);
                    acc[3][0].x = fmaf(a.w, wv, acc[3][0].x);
                }
            }
        }
    }

    // ---- epilogue: bias + activation; 4 consecutive px per thread ----
    const int r = pg >> 3, c4 = (pg & 7) * 4;
    const int yy = y0 + r, xx = x0 + c4;
    if (yy >= h) return;
    if (!TANH) {
        // channel-pair interleaved feature map [B][16][h][wp] float2 (dcepp_mma.cuh): this thread's NOUT outputs are
        // NOUT/2 whole pairs; 4 px of a pair = 32 contiguous bytes
        const int wp = mm_pitch(w);
#pragma unroll
        for (int o2 = 0; o2 < NOUT / 2; ++o2) {
            const float2 bias = *reinterpret_cast<const float2*>(s_pwb + og * NOUT + 2 * o2);
            float2 v[4];
#pragma unroll
            for (int p = 0; p < 4; ++p)
                v[p] = make_float2(fmaxf(acc[p][o2].x + bias.x, 0.f), fmaxf(acc[p][o2].y + bias.y, 0.f));
            float2* d = reinterpret_cast<float2*>(dst) + ((size_t)(b * 16 + og * (NOUT / 2) + o2) * h + yy) * wp + xx;
            if (xx + 3 < w) {
                reinterpret_cast<float4*>(d)[0] = make_float4(v[0].x, v[0].y, v[1].x, v[1].y);
                reinterpret_cast<float4*>(d)[1] = make_float4(v[2].x, v[2].y, v[3].x, v[3].y);
            } else {
#pragma unroll
                for (int p = 0; p < 4; ++p)
                    if (xx + p < w) d[p] = v[p];
            }
        }
    } else {
#pragma unroll
        for (int o = 0; o < NOUT; ++o) {
            const int oc = og * NOUT + o;
            if (oc >= COUT) continue;
            const float bias = s_pwb[oc];
            float* d = dst + ((size_t)(b * COUT + oc) * h + yy) * w + xx;
#pragma unroll
            for (int p = 0; p < 4; ++p) {
                const float u = ((o & 1) ? acc[p][o >> 1].y : acc[p][o >> 1].x) + bias;
                if (xx + p < w) d[p] = tanhf(u);
            }
        }
    }
}

template <int CS, int NSRC, int NOUT>
constexpr size_t k1_smem_bytes() {
    return sizeof(float) *
           (size_t)((CS * NSRC * 10 + 3) / 4 * 4 + CS * NSRC * 4 * NOUT + 4 * NOUT + CS * 10 * K1_HS + CS * 256);
}

template <int CS, int NSRC, int NOUT, bool FIRST, bool TANH>
static int launch_k1(const float* a, const float* b2, float* dst, const float* wl, int B, int h, int w, int H, int W,
                     float sf_scale, int* sync, cudaStream_t s) {
    auto kern = dcepp_layer_kernel<CS, NSRC, NOUT, FIRST, TANH>;
    constexpr size_t smem = k1_smem_bytes<CS, NSRC, NOUT>();
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return B200_ERR_LAUNCH;
    }
    dim3 grid((w + K1_TW - 1) / K1_TW, (h + K1_TH - 1) / K1_TH, B);
    kern<<<grid, 256, smem, s>>>(a, b2, dst, wl, h, w, H, W, sf_scale, sync);
    return launch_status();
}

// ------------------------------------------------------------------------------------------------
// K2: full-resolution upsample (align_corners=True) + curve.  Persistent CTAs over row strips.
// ------------------------------------------------------------------------------------------------
template <bool WRITE_ADJ>
__global__ void __launch_bounds__(256)
dcepp_apply_vec4_kernel(const float* __restrict__ x, const float* __restrict__ a_lr, float* __restrict__ enh,
                        float* __restrict__ adj, int B, int H, int W, int h, int w, int iters, float sh, float sw,
                        int R, int strips_per_frame) {
    extern __shared__ __align__(16) float vrow[];  // [R][3][w]: vertically interpolated low-res rows
    const int t = threadIdx.x;
    const int W4 = W >> 2;
    const int total = B * strips_per_frame;
    for (int strip = blockIdx.x; strip < total; strip += gridDim.x) {
        const int b = strip / strips_per_frame;
        const int Y0 = (strip - b * strips_per_frame) * R;
        const int rows = min(R, H - Y0);
        const size_t plane = (size_t)H * W;
        const int n_it = rows * W4;
        // the x values of this thread's NEXT group are always in flight: the first group's loads are issued before
        // the low-resolution rows of the strip are prepared, the following ones while the current group is computed
        float4 vn[3];
        if (t < n_it) {
            const int rr = t / W4;
            const size_t base = ((size_t)(b * 3) * H + (Y0 + rr)) * W + (t - rr * W4) * 4;
            vn[0] = ld_stream4(x + base);
            vn[1] = ld_stream4(x + base + plane);
            vn[2] = ld_stream4(x + base + 2 * plane);
        }
        __syncthreads();
        for (int idx = t; idx < rows * 3 * w; idx += 256) {
            const int rr = idx / (3 * w);
            const int rem = idx - rr * 3 * w;
            const int c = rem / w, xx = rem - c * w;
            const Tap ty = tap_align_corners(Y0 + rr, sh, h);
            const float* ap = a_lr + ((size_t)(b * 3 + c) * h) * w + xx;
            vrow[idx] = mm_lerp(ty.w0, __ldg(ap + (size_t)ty.i0 * w), ty.w1, __ldg(ap + (size_t)ty.i1 * w));
        }
        __syncthreads();
        for (int it = t; it < n_it; it += 256) {
            const int rr = it / W4;
            const int X = (it - rr * W4) * 4;
            const size_t base = ((size_t)(b * 3) * H + (Y0 + rr)) * W + X;
            float4 v[3] = {vn[0], vn[1], vn[2]};
            if (it + 256 < n_it) {
                const int rn = (it + 256) / W4;
                const size_t bn = ((size_t)(b * 3) * H + (Y0 + rn)) * W + (it + 256 - rn * W4) * 4;
                vn[0] = ld_stream4(x + bn);
                vn[1] = ld_stream4(x + bn + plane);
                vn[2] = ld_stream4(x + bn + 2 * plane);
            }
            Tap tx[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) tx[j] = tap_align_corners(X + j, sw, w);
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                const float* vr = vrow + (rr * 3 + c) * w;
                float4 a;
                a.x = mm_lerp(tx[0].w0, vr[tx[0].i0], tx[0].w1, vr[tx[0].i1]);
                a.y = mm_lerp(tx[1].w0, vr[tx[1].i0], tx[1].w1, vr[tx[1].i1]);
                a.z = mm_lerp(tx[2].w0, vr[tx[2].i0], tx[2].w1, vr[tx[2].i1]);
                a.w = mm_lerp(tx[3].w0, vr[tx[3].i0], tx[3].w1, vr[tx[3].i1]);
                float2 y0 = make_float2(v[c].x, v[c].y), y1 = make_float2(v[c].z, v[c].w);
                const float2 na0 = make_float2(-a.x, -a.y), na1 = make_float2(-a.z, -a.w);
                for (int i = 0; i < iters; ++i) {
                    y0 = curve_step2(y0, na0);
                    y1 = curve_step2(y1, na1);
                }
                st_stream4(enh + base + c * plane, make_float4(y0.x, y0.y, y1.x, y1.y));
                if (WRITE_ADJ) st_stream4(adj + base + c * plane, a);
            }
        }
    }
}

// generic fallback (W not a multiple of 4): one pixel per thread, taps straight from global/L2
__global__ void __launch_bounds__(256)
dcepp_apply_scalar_kernel(const float* __restrict__ x, const float* __restrict__ a_lr, float* __restrict__ enh,
                          float* __restrict__ adj, int B, int H, int W, int h, int w, int iters, float sh, float sw) {
    const long long total = (long long)B * 3 * H * W;
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < total; i += (long long)gridDim.x * 256) {
        const int X = (int)(i % W);
        const long long r1 = i / W;
        const int Y = (int)(r1 % H);
        const long long bc = r1 / H;
        const Tap ty = tap_align_corners(Y, sh, h), tx = tap_align_corners(X, sw, w);
        const float* ap = a_lr + (size_t)bc * h * w;
        const float a = mm_lerp(tx.w0, mm_lerp(ty.w0, ap[(size_t)ty.i0 * w + tx.i0], ty.w1, ap[(size_t)ty.i1 * w + tx.i0]),
                                tx.w1, mm_lerp(ty.w0, ap[(size_t)ty.i0 * w + tx.i1], ty.w1, ap[(size_t)ty.i1 * w + tx.i1]));
        float y = x[i];
        for (int k = 0; k < iters; ++k) y = curve_step(y, a);
        enh[i] = y;
        if (adj) adj[i] = a;
    }
}

static size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }
// one feature map: [B][16 channel pairs][h][mm_pitch(w)] float2
static size_t fmap_bytes(int B, int h, int w) { return align_up((size_t)B * 16 * h * mm_pitch(w) * sizeof(float2), 256); }
// the low-res curve map [B][3][h][pitch], pitch = w rounded up to a multiple of 4 (a tensor map reads it in the fused launch)
static size_t amap_bytes(int B, int h, int w) { return align_up((size_t)B * 3 * h * ((w + 3) & ~3) * sizeof(float), 256); }

}  // namespace b200

using namespace b200;

extern "C" size_t b200_dcepp_packed_floats(void) { return (size_t)PP_TOTAL_FLOATS + MM_TOTAL_WORDS; }

extern "C" int b200_dcepp_pack_weights(const float* const* dw_w, const float* const* dw_b, const float* const* pw_w,
                                       const float* const* pw_b, float* packed, void* stream) {
    if (!dw_w || !dw_b || !pw_w || !pw_b || !packed) return B200_ERR_NULL;
    PackArgs args;
    for (int l = 0; l < 7; ++l) {
        if (!dw_w[l] || !dw_b[l] || !pw_w[l] || !pw_b[l]) return B200_ERR_NULL;
        args.dw_w[l] = dw_w[l];
        args.dw_b[l] = dw_b[l];
        args.pw_w[l] = pw_w[l];
        args.pw_b[l] = pw_b[l];
    }
    dcepp_pack_kernel<<<7, 256, 0, (cudaStream_t)stream>>>(args, packed);
    return launch_status();
}

// workspace: 5 feature-map buffers [B,32,h,w] | the low-res / full-res 3-channel map | the net launch's sync words
extern "C" size_t b200_dcepp_workspace_bytes(int B, int H, int W, int sf) {
    if (B <= 0 || H <= 0 || W <= 0 || sf <= 0) return 0;
    const size_t h = H / sf, w = W / sf;
    const size_t fmap = fmap_bytes(B, (int)h, (int)w);
    const size_t amap = amap_bytes(B, (int)h, (int)w);
    return 5 * fmap + amap + align_up(mm_sync_words(B, (int)h, (int)w) * sizeof(int), 256);
}

// Layer 1 launch + the persistent launch for layers 2..7.  With `fuse` != nullptr the full-resolution pass rides in the
// second launch (blocks of `x` in the tile queue) and writes fuse->enh / fuse->adj.
static int dcepp_net(const float* x, const float* packed, float* a_lr, int B, int H, int W, int sf, void* workspace,
                     size_t workspace_bytes, const MmApplyArgs* fuse, cudaStream_t s, cudaEvent_t mid = nullptr) {
    const int h = H / sf, w = W / sf;
    const size_t fmap = fmap_bytes(B, h, w);
    const size_t amap = amap_bytes(B, h, w);
    if (workspace_bytes < b200_dcepp_workspace_bytes(B, H, W, sf)) return B200_ERR_WORKSPACE;
    char* ws = (char*)workspace;
    float* x1 = (float*)(ws + 0 * fmap);
    float* x2 = (float*)(ws + 1 * fmap);
    float* x3 = (float*)(ws + 2 * fmap);
    float* x4 = (float*)(ws + 3 * fmap);
    float* x5 = (float*)(ws + 4 * fmap);
    float* x6 = x4;  // x4 is dead once layer 5 has run
    const float sfs = (float)sf;
    int* sync = (int*)(ws + 5 * fmap + amap);
    MmApplyArgs ap{};
    if (fuse) ap = *fuse;
    // debug aid (not part of the ABI): with B200DCE_TRACE=1 and >= 1 MiB of workspace beyond what the call needs, the
    // fused launch records a per-item timeline there (tools/trace_dcepp.py reads it)
    if (fuse && getenv("B200DCE_TRACE") && workspace_bytes >= b200_dcepp_workspace_bytes(B, H, W, sf) + (1u << 20)) {
        ap.trace = (uint32_t*)(ws + b200_dcepp_workspace_bytes(B, H, W, sf));
        ap.trace_cap = (int)((workspace_bytes - b200_dcepp_workspace_bytes(B, H, W, sf)) / 16);
        cudaMemsetAsync(ap.trace, 0, 16, s);
    }
    int rc;
    if ((rc = launch_k1<3, 1, 8, true, false>(x, nullptr, x1, packed + pp_layer_offset(0), B, h, w, H, W, sfs, sync, s)))
        return rc;
    if (mid != nullptr && cudaEventRecord(mid, s) != cudaSuccess) return B200_ERR_LAUNCH;
    const uint32_t* mm = reinterpret_cast<const uint32_t*>(packed + PP_TOTAL_FLOATS);
    MmNetArgs a;  // e_conv2..7 (zero_dcepp.py:191-196): x2 x3 x4 | cat(x3,x4) -> x5 | cat(x2,x5) -> x6 | cat(x1,x6) -> x_r
    float* ds[6] = {x2, x3, x4, x5, x6, a_lr};
    for (int l = 0; l < 6; ++l) {
        a.dst[l] = ds[l];
        a.wl[l] = mm + mm_offset(l + 1);
    }
    float* maps[5] = {x1, x2, x3, x4, x5};
    return launch_mm_net(a, ap, x, maps, B, h, w, sync, s);
}

extern "C" int b200_dcepp_lowres(const float* x, const float* packed, float* a_lr, int B, int H, int W, int sf,
                                 void* workspace, size_t workspace_bytes, void* stream) {
    if (!x || !packed || !a_lr || !workspace) return B200_ERR_NULL;
    if (B <= 0 || H <= 0 || W <= 0 || sf <= 0) return B200_ERR_BAD_SHAPE;
    if (H % sf != 0 || W % sf != 0) return B200_ERR_BAD_SHAPE;
    if (B > 65535) return B200_ERR_UNSUPPORTED;
    if (!aligned16(x) || !aligned16(packed) || !aligned16(a_lr) || !aligned16(workspace)) return B200_ERR_MISALIGNED;
    return dcepp_net(x, packed, a_lr, B, H, W, sf, workspace, workspace_bytes, nullptr, (cudaStream_t)stream);
}

extern "C" int b200_dcepp_apply(const float* x, const float* a_lr, float* enhanced, float* adjust, int B, int H, int W,
                                int h, int w, int iters, void* stream) {
    if (!x || !a_lr || !enhanced) return B200_ERR_NULL;
    if (B <= 0 || H <= 0 || W <= 0 || h <= 0 || w <= 0 || iters < 0) return B200_ERR_BAD_SHAPE;
    if (!aligned16(x) || !aligned16(a_lr) || !aligned16(enhanced) || (adjust && !aligned16(adjust)))
        return B200_ERR_MISALIGNED;
    cudaStream_t s = (cudaStream_t)stream;
    // ATen area_pixel_compute_scale<float>(in, out, align_corners=true) = (in-1)/(out-1), 0 when out == 1
    const float sh = H > 1 ? (float)(h - 1) / (float)(H - 1) : 0.f;
    const float sw = W > 1 ? (float)(w - 1) / (float)(W - 1) : 0.f;
    if ((W & 3) == 0 && (size_t)3 * w * sizeof(float) <= 96 * 1024) {
        // rows per strip: as many as fit ~32 KB of smem, at most 8, and enough strips for >= 4 waves
        int R = (int)((32 * 1024) / ((size_t)3 * w * sizeof(float)));
        if (R < 1) R = 1;
        if (R > 8) R = 8;
        while (R > 1 && (long long)B * ((H + R - 1) / R) < (long long)B200_NUM_SMS * 8 * 4) R >>= 1;
        const int spf = (H + R - 1) / R;
        const size_t smem = (size_t)R * 3 * w * sizeof(float);
        // persistent CTAs: exactly one resident wave (SM count x achievable CTAs per SM)
        int occ = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(
                &occ, adjust ? (const void*)dcepp_apply_vec4_kernel<true> : (const void*)dcepp_apply_vec4_kernel<false>,
                256, smem) != cudaSuccess || occ < 1)
            occ = 4;
        long long grid = (long long)B * spf;
        if (grid > (long long)B200_NUM_SMS * occ) grid = (long long)B200_NUM_SMS * occ;
        if (adjust) {
            auto k = dcepp_apply_vec4_kernel<true>;
            if (smem > 48 * 1024 &&
                cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
                return B200_ERR_LAUNCH;
            k<<<(unsigned)grid, 256, smem, s>>>(x, a_lr, enhanced, adjust, B, H, W, h, w, iters, sh, sw, R, spf);
        } else {
            auto k = dcepp_apply_vec4_kernel<false>;
            if (smem > 48 * 1024 &&
                cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
                return B200_ERR_LAUNCH;
            k<<<(unsigned)grid, 256, smem, s>>>(x, a_lr, enhanced, nullptr, B, H, W, h, w, iters, sh, sw, R, spf);
        }
    } else {
        long long blocks = ((long long)B * 3 * H * W + 255) / 256;
        if (blocks > B200_NUM_SMS * 16) blocks = B200_NUM_SMS * 16;
        dcepp_apply_scalar_kernel<<<(unsigned)blocks, 256, 0, s>>>(x, a_lr, enhanced, adjust, B, H, W, h, w, iters, sh,
                                                                  sw);
    }
    return launch_status();
}

// B200DCE_DCEPP_FUSE (read per call; tuning / A-B aid, not part of the ABI): 0 = three launches (layer 1, net, full-
// resolution pass), 1 / 2 = the full-resolution pass rides in the net launch and 1 / 2 of an SM's two CTAs prefer its
// blocks over net tiles.  Default 1.
static int dcepp_fuse_mode() {
    const char* e = getenv("B200DCE_DCEPP_FUSE");
    if (!e || !*e) return 1;
    const int v = atoi(e);
    return v < 0 ? 0 : (v > 2 ? 2 : v);
}

static int dcepp_forward(const float* x, const float* packed, float* enhanced, float* adjust, int B, int H, int W, int sf,
                         int iters, void* workspace, size_t workspace_bytes, void* stream, cudaEvent_t mid) {
    if (!x || !packed || !enhanced || !workspace) return B200_ERR_NULL;
    if (B <= 0 || H <= 0 || W <= 0 || sf <= 0 || iters < 0) return B200_ERR_BAD_SHAPE;
    if (H % sf != 0 || W % sf != 0) return B200_ERR_BAD_SHAPE;
    if (workspace_bytes < b200_dcepp_workspace_bytes(B, H, W, sf)) return B200_ERR_WORKSPACE;
    const int h = H / sf, w = W / sf;
    const size_t fmap = fmap_bytes(B, h, w);
    float* a_ws = (float*)((char*)workspace + 5 * fmap);
    int rc;
    if (sf == 1) {
        // no resize: the net's output IS the full-resolution map (zero_dcepp.py:186-187,198)
        float* a = adjust ? adjust : a_ws;
        if ((rc = b200_dcepp_lowres(x, packed, a, B, H, W, 1, workspace, workspace_bytes, stream))) return rc;
        return b200_curve_shared(x, a, enhanced, B, 3, H, W, iters, stream);
    }
    const int mode = dcepp_fuse_mode();
    if (mode > 0 && B <= 65535 && aligned16(x) && aligned16(packed) && aligned16(enhanced) && aligned16(workspace) &&
        (!adjust || aligned16(adjust))) {
        MmApplyArgs ap{};
        ap.enh = enhanced;
        ap.adj = adjust;
        ap.H = H, ap.W = W, ap.iters = iters;
        ap.sh = H > 1 ? (float)(h - 1) / (float)(H - 1) : 0.f;  // as b200_dcepp_apply
        ap.sw = W > 1 ? (float)(w - 1) / (float)(W - 1) : 0.f;
        ap.prefer = mode;
        if (mm_plan_blocks(ap, B, w))
            return dcepp_net(x, packed, a_ws, B, H, W, sf, workspace, workspace_bytes, &ap, (cudaStream_t)stream, mid);
    }
    if ((rc = b200_dcepp_lowres(x, packed, a_ws, B, H, W, sf, workspace, workspace_bytes, stream))) return rc;
    if (mid != nullptr && cudaEventRecord(mid, (cudaStream_t)stream) != cudaSuccess) return B200_ERR_LAUNCH;
    return b200_dcepp_apply(x, a_ws, enhanced, adjust, B, H, W, h, w, iters, stream);
}

extern "C" int b200_dcepp_forward(const float* x, const float* packed, float* enhanced, float* adjust, int B, int H,
                                  int W, int sf, int iters, void* workspace, size_t workspace_bytes, void* stream) {
    return dcepp_forward(x, packed, enhanced, adjust, B, H, W, sf, iters, workspace, workspace_bytes, stream, nullptr);
}

extern "C" int b200_dcepp_forward_marked(const float* x, const float* packed, float* enhanced, float* adjust, int B,
                                         int H, int W, int sf, int iters, void* workspace, size_t workspace_bytes,
                                         void* stream, void* mid_event) {
    return dcepp_forward(x, packed, enhanced, adjust, B, H, W, sf, iters, workspace, workspace_bytes, stream,
                         (cudaEvent_t)mid_event);
}

// ------------------------------------------------------------------------------------------------
// b200_dcepp_forward_static: the same forward with its scratch (feature maps, low-res curve map, tile flags) in a
// file-scope device array instead of a caller-provided workspace -- for callers that cannot call cudaMalloc (CUDA graph
// capture, tight serving loops).  Every GPU that loads the library holds the array until the process exits.
// `g_dcepp_static_scratch` names a device symbol: in host code it is only a handle for cudaGetSymbolAddress, never
// an address to hand to a kernel.
// ------------------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned char g_dcepp_static_scratch[B200_DCEPP_STATIC_SCRATCH_BYTES];

// device address of the array on the CURRENT device, looked up once per device (with that device current)
static int dcepp_static_scratch_ptr(void** out) {
    constexpr int MAX_DEV = 64;
    static std::mutex mu;
    static void* ptrs[MAX_DEV] = {};
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= MAX_DEV) {
        (void)cudaGetLastError();
        return B200_ERR_NO_DEVICE;
    }
    std::lock_guard<std::mutex> lock(mu);
    if (!ptrs[dev]) {
        void* p = nullptr;
        if (cudaGetSymbolAddress(&p, g_dcepp_static_scratch) != cudaSuccess || !p) {
            (void)cudaGetLastError();
            return B200_ERR_LAUNCH;
        }
        ptrs[dev] = p;
    }
    *out = ptrs[dev];
    return B200_OK;
}

extern "C" size_t b200_dcepp_static_scratch_bytes(void) { return (size_t)B200_DCEPP_STATIC_SCRATCH_BYTES; }

extern "C" int b200_dcepp_forward_static(const float* x, const float* packed, float* enhanced, float* adjust, int B,
                                         int H, int W, int sf, int iters, void* stream) {
    if (!x || !packed || !enhanced) return B200_ERR_NULL;
    if (B <= 0 || H <= 0 || W <= 0 || sf <= 0 || iters < 0) return B200_ERR_BAD_SHAPE;
    if (H % sf != 0 || W % sf != 0) return B200_ERR_BAD_SHAPE;
    const size_t need = b200_dcepp_workspace_bytes(B, H, W, sf);
    if (need > (size_t)B200_DCEPP_STATIC_SCRATCH_BYTES) return B200_ERR_TOO_LARGE;
    void* scratch = nullptr;
    const int rc = dcepp_static_scratch_ptr(&scratch);
    if (rc) return rc;
    // Nothing in the array has to start at zero from the caller's side: the layer-1 launch of every forward clears the
    // tile counters and dependency flags it is going to use before the persistent launch reads them; feature maps and
    // the curve map are fully written before they are read.  (`need`, not the array size: no debug trace area.)
    return dcepp_forward(x, packed, enhanced, adjust, B, H, W, sf, iters, scratch, need, stream, nullptr);
}
