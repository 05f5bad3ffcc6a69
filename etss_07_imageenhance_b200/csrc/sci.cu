// SCI illumination-estimation enhancer (Finetunemodel.forward, sci/model.py:155-159 over
// EnhanceNetwork.forward :36-45, layers=1, channels=3, BatchNorm2d eval) as ONE fused pass:
//   fea1 = relu(conv3x3(x));  fea2 = fea1 + relu(bn(conv3x3(fea1)));  fea3 = sigmoid(conv3x3(fea2));
//   illu = clamp(fea3 + x, 1e-4, 1);  r = clamp(x / illu, 0, 1)
// CTA tile 28x60 output pixels, 8 warps; the three stacked 3x3 layers are evaluated on shrinking halo regions held in
// shared memory (x: 34x66, fea1: 32x64, fea2: 32x62), so x is read once from HBM and illu / r written once.
//
// The work is 3 x 81 MAC per position on the fp32 pipe (~550 FLOP per output pixel with the halo recompute): the
// kernel is bound by the FMA pipe and by instruction issue (both ~65 % busy), not by HBM, so the inner loop is built
// to be (almost) nothing but FMAs, with as few other instructions per FMA as the data flow allows:
//   * all three planes (x, fea1, fea2) are stored ROW-PAIR INTERLEAVED: element [pair row p][column] is the float2
//     (row 2p, row 2p+1), so a vertical pair of inputs is one aligned 64-bit operand of an FFMA2 and two adjacent
//     columns are one 128-bit shared access (a warp's accesses are consecutive: conflict free);
//   * a thread computes an item of FOUR rows x TWO columns x 3 output channels: per input channel six 128-bit loads
//     bring the 6 x 4 input window; the taps ky = 0 / 2 read aligned row pairs (FFMA2), the tap ky = 1 reads the pair
//     (r+1, r+2) that straddles two row pairs: two scalar FFMAs on the halves -- the same FMA-lane cycles as one
//     FFMA2 and no register moves (ptxas rebuilt the odd pairs with ~3.5 moves per FFMA2 saved; a second, odd-aligned
//     copy of the planes would double shared memory and its traffic);
//   * every FMA carries one weight as a scalar broadcast from a uniform register; the 252 weights travel as a by-value
//     kernel parameter in the constant bank, reordered on the host to the order the loop consumes them;
//   * tile width 60: with two halo columns per side layer 1 is exactly 32 column pairs = one warp per row group, no
//     left-over columns; 8 / 8 / 7 row groups of 4 rows = 8 warps (two per scheduler), one round per layer; 1080p is
//     39 x 32 tiles per frame = 45.0 waves of 3 CTAs per SM for 16 frames;
//   * sigmoid and the division on the SFU in their flush-to-zero forms (see sfu_ex2); tiles whose halo lies inside
//     the image take a path without masks; the x values for the epilogue come by cp.async under the last conv.
// History of the measurements (16 x 1080p): profiles/r02_summary.md section 6.
#include <math.h>

#include "common.cuh"
#include "tcgen05.cuh"

namespace b200 {

// b200_sci_pack_weights_host's layout (the ABI: OIHW as in the state dict, BatchNorm folded into w2 / b2)
struct SciWeights {
    float w1[81], b1[3];  // in_conv.0
    float w2[81], b2[3];  // conv.0 with BatchNorm folded in
    float w3[81], b3[3];  // out_conv.0
};
static_assert(sizeof(SciWeights) == B200_SCI_PACKED_FLOATS * sizeof(float), "packed size");
// the kernel parameter: each layer's weights as [ci][kx][co][ky] (the order the loop consumes them), every input
// channel's 27 padded to 28 floats (16-byte blocks; lets the loop over ci be rolled, see SCI_UNROLL_*)
constexpr int SCI_WCI = 28, SCI_WL = 3 * SCI_WCI + 4;  // per layer: 3 x 28 weights, 3 biases (+1)
struct __align__(16) SciKernelWeights {
    float w[3][SCI_WL];
};

#ifndef SCI_UNROLL_L1
#define SCI_UNROLL_L1 true
#endif
#ifndef SCI_UNROLL_L3
#define SCI_UNROLL_L3 true
#endif
#ifndef SCI_TH
#define SCI_TH 28
#endif
#ifndef SCI_CTAS
#define SCI_CTAS 3
#endif
constexpr int S_TH = SCI_TH, S_TW = 60;           // with the two halo columns per side layer 1 is exactly 32 column pairs wide
constexpr int SCI_NT = 32 * (S_TH / 4 + 1);   // one warp per row group of the tallest layer (S_TH / 4 + 1 groups of 4 rows)
constexpr int S1_G = S_TH / 4 + 1, S2_G = S_TH / 4 + 1, S3_G = S_TH / 4;  // row groups (4 rows) per layer
constexpr int SX_RP = 2 * S1_G + 1, SX_W = S_TW + 6;  // x:    rows y0-3 .. y0+S_TH+2 as 2*S1_G+1 row pairs, columns x0-3 .. x0+62
constexpr int S1_RP = 2 * S2_G + 1, S1_W = S_TW + 4;  // fea1: rows y0-2 .. y0+S_TH+1 (+ one pair never written, read by the
                                                      //       two fea2 rows that feed nothing), columns x0-2 .. x0+61
constexpr int S2_RP = 2 * S2_G, S2_W = S_TW + 2;      // fea2: rows y0-1 .. y0+S_TH+2, columns x0-1 .. x0+60
static_assert(S1_W / 2 <= 32 && S_TW % 4 == 0, "a warp covers a row group; tiles start on 16-byte boundaries");
constexpr int SCI_SX_BYTES = 8 * 3 * SX_RP * SX_W, SCI_S1_BYTES = 8 * 3 * S1_RP * S1_W, SCI_S2_BYTES = 8 * 3 * S2_RP * S2_W;
// fea2 reuses the x planes' bytes (x is only read by layer 1); 58 KB: three CTAs per SM.  During layer 3 the dead fea1
// region holds each thread's own x values for the epilogue (fetched by cp.async under the conv)
constexpr int SCI_SMEM = (SCI_SX_BYTES > SCI_S2_BYTES ? SCI_SX_BYTES : SCI_S2_BYTES) + SCI_S1_BYTES;
static_assert(12 * 32 * S3_G * 8 <= SCI_S1_BYTES, "layer 3's x slots fit the fea1 region");
static_assert(SCI_SX_BYTES % 16 == 0 && SCI_S1_BYTES % 16 == 0 && (SX_W * 8) % 16 == 0 && (S1_W * 8) % 16 == 0 && (S2_W * 8) % 16 == 0, "128-bit rows");

// One item: output rows b .. b+3 (two row pairs) x columns j, j+1 x 3 channels from a row-pair plane whose origin lies
// one row above and one column left of the output's origin.  `in` points at (pair row b/2, column j) of channel 0;
// W_IN / PLANE in float2 units.  out[col][co][pair].  UNROLL: the loop over the input channel unrolled (default for
// all layers; rolled it is 1/3 of the code -- that mattered while the epilogue was 4x its present size and a fifth of
// all issue slots went to instruction fetch; with the lean epilogue unrolled is 7 % faster).
template <int W_IN, int PLANE, bool RESIDUAL, bool UNROLL>
__device__ __forceinline__ void conv3_item(const float2* __restrict__ in, const float* __restrict__ w, float2 out[2][3][2],
                                           float2 res[2][3][2]) {
#pragma unroll
    for (int col = 0; col < 2; ++col)
#pragma unroll
        for (int co = 0; co < 3; ++co) out[col][co][0] = out[col][co][1] = make_float2(w[3 * SCI_WCI + co], w[3 * SCI_WCI + co]);
    // RESIDUAL (layer 2 only): unrolled, so that res[col][ci][pair] = the input at the output's own position (rows b+1 ..
    // b+4, columns j+1, j+2) is taken from the registers of iteration ci
#pragma unroll(UNROLL ? 3 : 1)
    for (int ci = 0; ci < 3; ++ci) {
        float2 a[3][4];  // a[t][k] = rows (b+2t, b+2t+1) at column j+k
#pragma unroll
        for (int t = 0; t < 3; ++t) {
            const float4 lo = *reinterpret_cast<const float4*>(in + ci * PLANE + t * W_IN);
            const float4 hi = *reinterpret_cast<const float4*>(in + ci * PLANE + t * W_IN + 2);
            a[t][0] = make_float2(lo.x, lo.y), a[t][1] = make_float2(lo.z, lo.w);
            a[t][2] = make_float2(hi.x, hi.y), a[t][3] = make_float2(hi.z, hi.w);
        }
        if (RESIDUAL) {
#pragma unroll
            for (int col = 0; col < 2; ++col) {
                res[col][ci][0] = make_float2(a[0][1 + col].y, a[1][1 + col].x);
                res[col][ci][1] = make_float2(a[1][1 + col].y, a[2][1 + col].x);
            }
        }
        const float* wc = w + ci * SCI_WCI;
#pragma unroll
        for (int kx = 0; kx < 3; ++kx) {
#pragma unroll
            for (int co = 0; co < 3; ++co) {
                const float* wp = wc + (kx * 3 + co) * 3;
                const float2 w0 = make_float2(wp[0], wp[0]), w2 = make_float2(wp[2], wp[2]);
                const float w1 = wp[1];
#pragma unroll
                for (int col = 0; col < 2; ++col) {
                    const int k = col + kx;
                    // ky = 0 / 2: the input pairs are aligned row pairs -> FFMA2; ky = 1: the pair (b+1, b+2) straddles
                    // two row pairs -> two scalar FFMAs on the halves (the same FMA-lane cycles, no register moves)
                    out[col][co][0] = __ffma2_rn(w0, a[0][k], out[col][co][0]);
                    out[col][co][1] = __ffma2_rn(w0, a[1][k], out[col][co][1]);
                    out[col][co][0].x = fmaf(w1, a[0][k].y, out[col][co][0].x);
                    out[col][co][0].y = fmaf(w1, a[1][k].x, out[col][co][0].y);
                    out[col][co][1].x = fmaf(w1, a[1][k].y, out[col][co][1].x);
                    out[col][co][1].y = fmaf(w1, a[2][k].x, out[col][co][1].y);
                    out[col][co][0] = __ffma2_rn(w2, a[1][k], out[col][co][0]);
                    out[col][co][1] = __ffma2_rn(w2, a[2][k], out[col][co][1]);
                }
            }
        }
    }
}

// sigmoid and x / illu on the SFU: ex2.approx / rcp.approx (~2 ulp), flush-to-zero forms -- the default forms wrap
// every MUFU in a denormal-range fix-up (two multiplies and a compare each), the IEEE-rounded ones cost as many
// FMA-pipe instructions per pixel as a whole conv layer; the contract is 1e-3.  Neither operand range comes near a
// denormal: 1 + 2^t >= 1, illu is clamped to [1e-4, 1].
__device__ __forceinline__ float sfu_ex2(float v) {
    float r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(v));
    return r;
}
__device__ __forceinline__ float sfu_rcp(float v) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(v));
    return r;
}

// item (row group gi, column pair) of thread t for a layer with G row groups of NP <= 32 column pairs: warp gi, lane pair
template <int G, int NP>
__device__ __forceinline__ bool sci_item(int t, int& gi, int& ci2) {
    static_assert(G * 32 <= SCI_NT && NP <= 32, "one round");
    gi = t >> 5, ci2 = t & 31;
    return gi < G && ci2 < NP;
}

// relu / optional residual / (MASKED: zero outside the image = the next conv's zero padding), then the two row pairs of
// the two columns of channel c as two 128-bit stores; (yy, xx) = image position of the item's first row / column
template <int W_OUT, int RP, bool RESIDUAL, bool MASKED>
__device__ __forceinline__ void sci_store_item(float2* __restrict__ plane, int gi, int j, const float2 out[2][3][2],
                                               const float2 res[2][3][2], int yy, int xx, int H, int W) {
    bool rin[4], cin[2];
#pragma unroll
    for (int q = 0; q < 4; ++q) rin[q] = !MASKED || (yy + q >= 0 && yy + q < H);
#pragma unroll
    for (int col = 0; col < 2; ++col) cin[col] = !MASKED || (xx + col >= 0 && xx + col < W);
#pragma unroll
    for (int c = 0; c < 3; ++c)
#pragma unroll
        for (int pr = 0; pr < 2; ++pr) {
            float v[2][2];
#pragma unroll
            for (int col = 0; col < 2; ++col) {
                float e0 = fmaxf(out[col][c][pr].x, 0.f), e1 = fmaxf(out[col][c][pr].y, 0.f);
                if (RESIDUAL) e0 += res[col][c][pr].x, e1 += res[col][c][pr].y;
                v[col][0] = cin[col] && rin[2 * pr] ? e0 : 0.f;
                v[col][1] = cin[col] && rin[2 * pr + 1] ? e1 : 0.f;
            }
            *reinterpret_cast<float4*>(plane + (c * RP + 2 * gi + pr) * W_OUT + j) =
                make_float4(v[0][0], v[0][1], v[1][0], v[1][1]);
        }
}

__global__ void __launch_bounds__(SCI_NT, SCI_CTAS)
sci_fused_kernel(const float* __restrict__ x, float* __restrict__ illu, float* __restrict__ r,
                 const __grid_constant__ SciKernelWeights wt, int H, int W) {
    extern __shared__ __align__(16) float2 sci_smem[];
    float2* sx = sci_smem;                                   // [3][SX_RP][SX_W]
    float2* s2 = sci_smem;                                   // [3][S2_RP][S2_W], over the dead x planes
    float2* s1 = sci_smem + (SCI_SMEM - SCI_S1_BYTES) / 8;   // [3][S1_RP][S1_W]
    const int t = threadIdx.x;
    const int b = blockIdx.z;
    const int y0 = blockIdx.y * S_TH, x0 = blockIdx.x * S_TW;
    const size_t plane = (size_t)H * W;
    const float* xb = x + (size_t)b * 3 * plane;
    // tile + halo completely inside the image (block-uniform): no zero padding, no masks, 128-bit global accesses
    const bool interior = (W & 3) == 0 && y0 >= 3 && y0 + S_TH + 3 <= H && x0 >= 4 && x0 + S_TW + 4 <= W;
    {   // x halo tile into row pairs: a task = (row pair p, 4 columns from x0 - 4 + 4q) in all three channels; the loads
        // of a task are in flight together; zeros outside the image are the conv's zero padding
        constexpr int NQ = (SX_W + 2 + 3) / 4;  // columns x0-4 .. (the outer two unused)
        constexpr int NPQ = SX_RP * NQ;
        static_assert(NPQ <= 2 * SCI_NT, "two rounds of tasks");
        auto xtask = [&](int pq) {
            const int p = pq / NQ, q = pq - p * NQ;
            const int gx = x0 - 4 + 4 * q, r0 = y0 - 3 + 2 * p;
            float4 v0[3], v1[3];
            if (interior) {
                const float* g0 = xb + (size_t)r0 * W + gx;
#pragma unroll
                for (int u = 0; u < 3; ++u) {
                    v0[u] = __ldg(reinterpret_cast<const float4*>(g0 + u * plane));
                    v1[u] = __ldg(reinterpret_cast<const float4*>(g0 + u * plane + W));
                }
            } else {
                const bool vec = (W & 3) == 0;
                const bool in0 = r0 >= 0 && r0 < H, in1 = r0 + 1 >= 0 && r0 + 1 < H;
#pragma unroll
                for (int u = 0; u < 3; ++u) {
                    v0[u] = v1[u] = make_float4(0.f, 0.f, 0.f, 0.f);
                    const float* g0 = xb + ((long long)u * (long long)plane + (long long)r0 * W + gx);
                    if (vec) {
                        if (gx >= 0 && gx < W) {
                            if (in0) v0[u] = __ldg(reinterpret_cast<const float4*>(g0));
                            if (in1) v1[u] = __ldg(reinterpret_cast<const float4*>(g0 + W));
                        }
                    } else {
                        float e0[4] = {0.f, 0.f, 0.f, 0.f}, e1[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
                        for (int e = 0; e < 4; ++e)
                            if (gx + e >= 0 && gx + e < W) {
                                if (in0) e0[e] = __ldg(g0 + e);
                                if (in1) e1[e] = __ldg(g0 + W + e);
                            }
                        v0[u] = make_float4(e0[0], e0[1], e0[2], e0[3]);
                        v1[u] = make_float4(e1[0], e1[1], e1[2], e1[3]);
                    }
                }
            }
            float2* dst = sx + p * SX_W + 4 * q - 1;  // plane column k <-> image column x0 - 3 + k
#pragma unroll
            for (int u = 0; u < 3; ++u) {
                float2* d = dst + u * SX_RP * SX_W;
                if (q > 0) d[0] = make_float2(v0[u].x, v1[u].x);
                *reinterpret_cast<float4*>(d + 1) = make_float4(v0[u].y, v1[u].y, v0[u].z, v1[u].z);
                if (4 * q + 2 < SX_W) d[3] = make_float2(v0[u].w, v1[u].w);
            }
        };
        xtask(t);
        if (t < NPQ - SCI_NT) xtask(SCI_NT + t);
    }
    __syncthreads();
    int gi, ip;
    // fea1 = relu(conv(x)) on 28 x 68
    if (sci_item<S1_G, S1_W / 2>(t, gi, ip)) {
        const int j = 2 * ip, yy = y0 - 2 + 4 * gi, xx = x0 - 2 + j;
        float2 o[2][3][2];
        conv3_item<SX_W, SX_RP * SX_W, false, SCI_UNROLL_L1>(sx + (2 * gi) * SX_W + j, wt.w[0], o, o);
        if (interior) {
            sci_store_item<S1_W, S1_RP, false, false>(s1, gi, j, o, o, yy, xx, H, W);
        } else {
            sci_store_item<S1_W, S1_RP, false, true>(s1, gi, j, o, o, yy, xx, H, W);
        }
    }
    __syncthreads();
    // fea2 = fea1 + relu(bn(conv(fea1))) on 28 x 66 (rows y0-1 ..; the last two rows feed nothing)
    if (sci_item<S2_G, S2_W / 2>(t, gi, ip)) {
        const int j = 2 * ip, yy = y0 - 1 + 4 * gi, xx = x0 - 1 + j;
        float2 o[2][3][2], res[2][3][2];
        conv3_item<S1_W, S1_RP * S1_W, true, true>(s1 + (2 * gi) * S1_W + j, wt.w[1], o, res);
        if (interior) {
            sci_store_item<S2_W, S2_RP, true, false>(s2, gi, j, o, res, yy, xx, H, W);
        } else {
            sci_store_item<S2_W, S2_RP, true, true>(s2, gi, j, o, res, yy, xx, H, W);
        }
    }
    __syncthreads();
    if (sci_item<S3_G, S_TW / 2>(t, gi, ip)) {
        const int j = 2 * ip, xx = x0 + j;
        if (xx < W && y0 + 4 * gi < H) {
            const bool pairs = (W & 1) == 0;  // then xx + 1 < W and every (row, xx) is 8-byte aligned
            // the x values under this thread's outputs: cp.async into the thread's own slots of the (dead) fea1 region,
            // in flight under the conv; slot k of thread t at [k][t]: conflict-free
            float2* xs = s1 + t;
            // 32-bit element offsets inside the frame (the host checks 3 * H * W < 2^31): one IMAD.WIDE per address
            const int hw = H * W, off0 = (y0 + 4 * gi) * W + xx;
            if (pairs) {
                const uint32_t sh = tc::smem_u32(xs);
#pragma unroll
                for (int c = 0; c < 3; ++c)
#pragma unroll
                    for (int q = 0; q < 4; ++q)
                        if (y0 + 4 * gi + q < H) tc::cp_async8(sh + (c * 4 + q) * 32 * S3_G * 8, xb + (off0 + c * hw + q * W));
                tc::cp_async_commit();
            }
            float2 o[2][3][2];
            conv3_item<S2_W, S2_RP * S2_W, false, SCI_UNROLL_L3>(s2 + (2 * gi) * S2_W + j, wt.w[2], o, o);
            if (pairs) {
                tc::cp_async_wait<0>();
                float* ip = illu ? illu + (size_t)b * 3 * plane : nullptr;
                float* rp = r + (size_t)b * 3 * plane;
#pragma unroll
                for (int c = 0; c < 3; ++c) {
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        if (y0 + 4 * gi + q >= H) continue;
                        const float2 xv = xs[(c * 4 + q) * 32 * S3_G];
                        const float xin[2] = {xv.x, xv.y};
                        float il[2], rv[2];
#pragma unroll
                        for (int col = 0; col < 2; ++col) {
                            const float ov = (q & 1) ? o[col][c][q >> 1].y : o[col][c][q >> 1].x;
                            const float sg = sfu_rcp(1.f + sfu_ex2(ov * -1.4426950408889634f));  // 1 / (1 + e^-ov)
                            il[col] = fminf(fmaxf(sg + xin[col], 0.0001f), 1.f);
                            rv[col] = fminf(fmaxf(xin[col] * sfu_rcp(il[col]), 0.f), 1.f);
                        }
                        if (ip) *reinterpret_cast<float2*>(ip + (off0 + c * hw + q * W)) = make_float2(il[0], il[1]);
                        *reinterpret_cast<float2*>(rp + (off0 + c * hw + q * W)) = make_float2(rv[0], rv[1]);
                    }
                }
            } else {  // odd widths: scalar accesses
#pragma unroll 1
                for (int c = 0; c < 3; ++c) {
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const int yy = y0 + 4 * gi + q;
                        if (yy >= H) continue;
                        const size_t off = ((size_t)(b * 3 + c)) * plane + (size_t)yy * W + xx;
#pragma unroll
                        for (int col = 0; col < 2; ++col) {
                            if (xx + col >= W) continue;
                            const float xin = __ldg(x + off + col);
                            const float ov = c == 0 ? ((q & 1) ? o[col][0][q >> 1].y : o[col][0][q >> 1].x)
                                           : c == 1 ? ((q & 1) ? o[col][1][q >> 1].y : o[col][1][q >> 1].x)
                                                    : ((q & 1) ? o[col][2][q >> 1].y : o[col][2][q >> 1].x);
                            const float sg = sfu_rcp(1.f + sfu_ex2(ov * -1.4426950408889634f));
                            const float il = fminf(fmaxf(sg + xin, 0.0001f), 1.f);
                            if (illu) illu[off + col] = il;
                            r[off + col] = fminf(fmaxf(xin * sfu_rcp(il), 0.f), 1.f);
                        }
                    }
                }
            }
        }
    }
}

}  // namespace b200

extern "C" int b200_sci_pack_weights_host(const float* in_w, const float* in_b, const float* mid_w, const float* mid_b,
                                          const float* bn_gamma, const float* bn_beta, const float* bn_mean,
                                          const float* bn_var, float bn_eps, const float* out_w, const float* out_b,
                                          float* packed_host) {
    if (!in_w || !in_b || !mid_w || !mid_b || !bn_gamma || !bn_beta || !bn_mean || !bn_var || !out_w || !out_b ||
        !packed_host)
        return B200_ERR_NULL;
    b200::SciWeights* p = reinterpret_cast<b200::SciWeights*>(packed_host);
    for (int i = 0; i < 81; ++i) p->w1[i] = in_w[i];
    for (int i = 0; i < 3; ++i) p->b1[i] = in_b[i];
    for (int co = 0; co < 3; ++co) {
        // eval-mode BatchNorm2d: y = (t - mean) / sqrt(var + eps) * gamma + beta, folded into conv.0
        const double s = (double)bn_gamma[co] / sqrt((double)bn_var[co] + (double)bn_eps);
        for (int i = 0; i < 27; ++i) p->w2[co * 27 + i] = (float)((double)mid_w[co * 27 + i] * s);
        p->b2[co] = (float)(((double)mid_b[co] - (double)bn_mean[co]) * s + (double)bn_beta[co]);
    }
    for (int i = 0; i < 81; ++i) p->w3[i] = out_w[i];
    for (int i = 0; i < 3; ++i) p->b3[i] = out_b[i];
    return B200_OK;
}

extern "C" int b200_sci_forward(const float* x, const float* packed_host, float* illu, float* r, int B, int H, int W,
                                void* stream) {
    if (!x || !packed_host || !r) return B200_ERR_NULL;
    if (B <= 0 || H <= 0 || W <= 0) return B200_ERR_BAD_SHAPE;
    if (B > 65535 || 3LL * H * W + W >= (1LL << 31)) return B200_ERR_UNSUPPORTED;  // grid.z; 32-bit offsets inside a frame
    if (!b200::aligned16(x) || !b200::aligned16(r) || (illu && !b200::aligned16(illu))) return B200_ERR_MISALIGNED;
    b200::SciWeights pk;
    memcpy(&pk, packed_host, sizeof(pk));
    b200::SciKernelWeights wt{};  // OIHW -> [ci][kx][co][ky], the order the kernel's loop consumes
    const float* src_w[3] = {pk.w1, pk.w2, pk.w3};
    const float* src_b[3] = {pk.b1, pk.b2, pk.b3};
    for (int l = 0; l < 3; ++l) {
        for (int co = 0; co < 3; ++co)
            for (int ci = 0; ci < 3; ++ci)
                for (int ky = 0; ky < 3; ++ky)
                    for (int kx = 0; kx < 3; ++kx)
                        wt.w[l][ci * b200::SCI_WCI + (kx * 3 + co) * 3 + ky] = src_w[l][((co * 3 + ci) * 3 + ky) * 3 + kx];
        for (int co = 0; co < 3; ++co) wt.w[l][3 * b200::SCI_WCI + co] = src_b[l][co];
    }
    dim3 grid((W + b200::S_TW - 1) / b200::S_TW, (H + b200::S_TH - 1) / b200::S_TH, B);
    if (cudaFuncSetAttribute(b200::sci_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, b200::SCI_SMEM) !=
        cudaSuccess) {
        (void)cudaGetLastError();
        return B200_ERR_LAUNCH;
    }
    b200::sci_fused_kernel<<<grid, b200::SCI_NT, b200::SCI_SMEM, (cudaStream_t)stream>>>(x, illu, r, wt, H, W);
    return b200::launch_status();
}
