// SCI illumination-estimation enhancer (Finetunemodel.forward, sci/model.py:155-159 over
// EnhanceNetwork.forward :36-45, layers=1, channels=3, BatchNorm2d eval) as ONE fused pass:
//   fea1 = relu(conv3x3(x));  fea2 = fea1 + relu(bn(conv3x3(fea1)));  fea3 = sigmoid(conv3x3(fea2));
//   illu = clamp(fea3 + x, 1e-4, 1);  r = clamp(x / illu, 0, 1)
// CTA tile 24x32 output pixels; the three stacked 3x3 layers are evaluated on shrinking halo regions held in
// shared memory (x: 30x38, fea1: 28x36, fea2: 26x34), so x is read once from HBM and illu / r written once.
//
// The work is 3 x 81 MAC per position on the fp32 pipe (~580 FLOP per output pixel with the halo recompute): the
// kernel is bound by the packed-FMA pipe, not by HBM, so the inner loop is built to be (almost) nothing but FFMA2:
//   * a thread computes FOUR vertically adjacent positions x 3 output channels; every FFMA2 carries one weight (scalar
//     broadcast from a uniform register; all 252 weights travel as a by-value kernel parameter in the constant bank)
//     times a vertical value pair (rows r, r+1);
//   * fea1 / fea2 are stored ROW-PAIR INTERLEAVED, twice: copy A holds (row 2p, 2p+1), copy B holds (row 2p+1, 2p+2),
//     so all five pairs a 4-row item needs per (channel, column) are aligned 64-bit shared loads -- no register moves
//     to assemble operands (the row-major version spent as many MOVs as FFMA2s);
//   * item counts per layer (7 x 36, 7 x 34, 6 x 32) are just under the 256 threads: one round per layer, a warp
//     per row group so that its 64-bit shared accesses are conflict free.
#include <math.h>

#include "common.cuh"
#include "tcgen05.cuh"

namespace b200 {

struct SciWeights {
    float w1[81], b1[3];  // in_conv.0
    float w2[81], b2[3];  // conv.0 with BatchNorm folded in
    float w3[81], b3[3];  // out_conv.0
};
static_assert(sizeof(SciWeights) == B200_SCI_PACKED_FLOATS * sizeof(float), "packed size");

constexpr int S_TH = 24, S_TW = 32;
constexpr int SX_H = S_TH + 6, SX_W = 40;  // 30 rows x (4 + 32 + 4) columns: image column x0 - 4 at column 0, so every
                                           // row is ten 16-byte cp.async chunks
constexpr int S1_G = 7, S1_W = S_TW + 4;   // fea1: rows y0-2 .. y0+25 (7 groups of 4), columns x0-2 .. x0+33
constexpr int S2_G = 7, S2_W = S_TW + 2;   // fea2: rows y0-1 .. y0+26 (the last two are never read), columns x0-1 .. x0+32
constexpr int S3_G = S_TH / 4;
constexpr int S1_RP = 2 * S1_G + 1, S2_RP = 2 * S2_G + 1;  // row pairs per plane (+1: the B copy's last half pair, and
                                                            // fea2's unused rows read one pair past fea1's end)
// fea2 reuses the x tile's bytes (x is only read by layer 1; the epilogue re-reads its 12 centre values from L2)
constexpr int SCI_SX_BYTES = 4 * 3 * SX_H * SX_W, SCI_S2_BYTES = 8 * 2 * 3 * S2_RP * S2_W;
constexpr int SCI_SMEM = (SCI_SX_BYTES > SCI_S2_BYTES ? SCI_SX_BYTES : SCI_S2_BYTES) + 8 * 2 * 3 * S1_RP * S1_W;

// four vertically adjacent outputs (rows b .. b+3 as two pairs) x 3 channels from row-pair planes:
//   a = copy A at pair row b/2 (rows b, b+1), b2 = copy B at pair row b/2 (rows b+1, b+2); planes are PLANE float2 apart
template <int W_IN, int PLANE>
__device__ __forceinline__ void conv3_quad(const float2* __restrict__ a, const float2* __restrict__ b2, const float* w,
                                           const float* bias, float2 out[3][2]) {
#pragma unroll
    for (int co = 0; co < 3; ++co) out[co][0] = out[co][1] = make_float2(bias[co], bias[co]);
#pragma unroll
    for (int ci = 0; ci < 3; ++ci) {
#pragma unroll
        for (int kx = 0; kx < 3; ++kx) {
            const float2 a0 = a[ci * PLANE + kx], a1 = a[ci * PLANE + W_IN + kx], a2 = a[ci * PLANE + 2 * W_IN + kx];
            const float2 b0 = b2[ci * PLANE + kx], b1 = b2[ci * PLANE + W_IN + kx];
#pragma unroll
            for (int co = 0; co < 3; ++co) {
                const float w0 = w[((co * 3 + ci) * 3 + 0) * 3 + kx], w1 = w[((co * 3 + ci) * 3 + 1) * 3 + kx],
                            w2 = w[((co * 3 + ci) * 3 + 2) * 3 + kx];
                out[co][0] = __ffma2_rn(make_float2(w0, w0), a0, out[co][0]);
                out[co][1] = __ffma2_rn(make_float2(w0, w0), a1, out[co][1]);
                out[co][0] = __ffma2_rn(make_float2(w1, w1), b0, out[co][0]);
                out[co][1] = __ffma2_rn(make_float2(w1, w1), b1, out[co][1]);
                out[co][0] = __ffma2_rn(make_float2(w2, w2), a1, out[co][0]);
                out[co][1] = __ffma2_rn(make_float2(w2, w2), a2, out[co][1]);
            }
        }
    }
}

// the same from a row-major plane (the x tile): `in` points at row b, left tap column, channel 0
template <int IN_W, int IN_PLANE>
__device__ __forceinline__ void conv3_quad_rowmajor(const float* __restrict__ in, const float* w, const float* bias,
                                                    float2 out[3][2]) {
#pragma unroll
    for (int co = 0; co < 3; ++co) out[co][0] = out[co][1] = make_float2(bias[co], bias[co]);
#pragma unroll
    for (int ci = 0; ci < 3; ++ci) {
#pragma unroll
        for (int kx = 0; kx < 3; ++kx) {
            float v[6];
#pragma unroll
            for (int rr = 0; rr < 6; ++rr) v[rr] = in[ci * IN_PLANE + rr * IN_W + kx];
#pragma unroll
            for (int ky = 0; ky < 3; ++ky) {
                const float2 p0 = make_float2(v[ky], v[ky + 1]), p1 = make_float2(v[ky + 2], v[ky + 3]);
#pragma unroll
                for (int co = 0; co < 3; ++co) {
                    const float wv = w[((co * 3 + ci) * 3 + ky) * 3 + kx];
                    out[co][0] = __ffma2_rn(make_float2(wv, wv), p0, out[co][0]);
                    out[co][1] = __ffma2_rn(make_float2(wv, wv), p1, out[co][1]);
                }
            }
        }
    }
}

// store rows b .. b+3 (b = 4 * gi) of one channel at column c into the two row-pair copies
template <int W_OUT>
__device__ __forceinline__ void store_quad(float2* __restrict__ A, float2* __restrict__ B, int gi, int c, float o0,
                                           float o1, float o2, float o3) {
    A[(2 * gi) * W_OUT + c] = make_float2(o0, o1);
    A[(2 * gi + 1) * W_OUT + c] = make_float2(o2, o3);
    B[(2 * gi) * W_OUT + c] = make_float2(o1, o2);
    if (gi > 0) B[(2 * gi - 1) * W_OUT + c].y = o0;
    B[(2 * gi + 1) * W_OUT + c].x = o3;
}

// item (row group gi, column cc) of thread t for a layer with G groups of COLS >= 32 columns: warp gi takes columns
// 0..31 of its group (consecutive lanes = consecutive columns: conflict-free 64-bit shared accesses), the last warp
// takes the COLS - 32 leftover columns of every group
template <int G, int COLS>
__device__ __forceinline__ bool sci_item(int t, int& gi, int& cc) {
    static_assert(G <= 7 && COLS >= 32 && (COLS - 32) * G <= 32, "one round of 8 warps");
    const int warp = t >> 5, lane = t & 31;
    if (warp < G) {
        gi = warp, cc = lane;
        return true;
    }
    if (COLS == 32 || warp != 7 || lane >= (COLS - 32) * G) return false;
    gi = lane / (COLS - 32 > 0 ? COLS - 32 : 1), cc = 32 + lane % (COLS - 32 > 0 ? COLS - 32 : 1);
    return true;
}

__global__ void __launch_bounds__(256, 4)
sci_fused_kernel(const float* __restrict__ x, float* __restrict__ illu, float* __restrict__ r,
                 const __grid_constant__ SciWeights wt, int H, int W) {
    extern __shared__ __align__(16) float sci_smem[];
    float* sx = sci_smem;                                                     // [3][SX_H][SX_W]
    float2* s2a = reinterpret_cast<float2*>(sci_smem);                        // [3][S2_RP][S2_W], over the dead x tile
    float2* s2b = s2a + 3 * S2_RP * S2_W;
    float2* s1a = reinterpret_cast<float2*>(sci_smem + (SCI_SMEM - 8 * 2 * 3 * S1_RP * S1_W) / 4);  // [3][S1_RP][S1_W]
    float2* s1b = s1a + 3 * S1_RP * S1_W;
    const int t = threadIdx.x;
    const int b = blockIdx.z;
    const int y0 = blockIdx.y * S_TH, x0 = blockIdx.x * S_TW;
    const size_t plane = (size_t)H * W;
    const float* xb = x + (size_t)b * 3 * plane;

    {   // x halo tile via cp.async (all loads in flight at once); zero fill outside the image = the conv's zero padding
        const uint32_t sh = tc::smem_u32(sx);
        if ((W & 3) == 0) {
            for (int i = t; i < 3 * SX_H * (SX_W / 4); i += 256) {
                const int p = i / (SX_W / 4), j = i - p * (SX_W / 4);
                const int c = p / SX_H, rr = p - c * SX_H;
                const int yy = y0 - 3 + rr, xx = x0 - 4 + 4 * j;
                const bool ok = yy >= 0 && yy < H && xx >= 0 && xx < W;
                tc::cp_async16(sh + (p * SX_W + 4 * j) * 4, xb + (ok ? c * plane + (size_t)yy * W + xx : (size_t)0), ok);
            }
        } else {
            for (int i = t; i < 3 * SX_H * SX_W; i += 256) {
                const int p = i / SX_W, cc = i - p * SX_W;
                const int c = p / SX_H, rr = p - c * SX_H;
                const int yy = y0 - 3 + rr, xx = x0 - 4 + cc;
                const bool ok = yy >= 0 && yy < H && xx >= 0 && xx < W;
                tc::cp_async4(sh + i * 4, xb + (ok ? c * plane + (size_t)yy * W + xx : (size_t)0), ok);
            }
        }
        tc::cp_async_commit();
        tc::cp_async_wait<0>();
    }
    __syncthreads();
    // fea1 = relu(conv(x)) on 28 x 36; zero outside the image (that is the next conv's zero padding)
    int gi, cc;
    if (sci_item<S1_G, S1_W>(t, gi, cc)) {
        const int yy = y0 - 2 + 4 * gi, xx = x0 - 2 + cc;
        float2 o[3][2];
        conv3_quad_rowmajor<SX_W, SX_H * SX_W>(sx + (4 * gi) * SX_W + cc + 1, wt.w1, wt.b1, o);
        const bool xin = xx >= 0 && xx < W;
        const bool in0 = xin && yy >= 0 && yy < H, in1 = xin && yy + 1 >= 0 && yy + 1 < H;
        const bool in2 = xin && yy + 2 >= 0 && yy + 2 < H, in3 = xin && yy + 3 >= 0 && yy + 3 < H;
#pragma unroll
        for (int c = 0; c < 3; ++c)
            store_quad<S1_W>(s1a + c * S1_RP * S1_W, s1b + c * S1_RP * S1_W, gi, cc, in0 ? fmaxf(o[c][0].x, 0.f) : 0.f,
                             in1 ? fmaxf(o[c][0].y, 0.f) : 0.f, in2 ? fmaxf(o[c][1].x, 0.f) : 0.f,
                             in3 ? fmaxf(o[c][1].y, 0.f) : 0.f);
    }
    __syncthreads();
    // fea2 = fea1 + relu(bn(conv(fea1))) on 28 x 34 (rows y0-1 ..; the last two rows feed nothing)
    if (sci_item<S2_G, S2_W>(t, gi, cc)) {
        const int yy = y0 - 1 + 4 * gi, xx = x0 - 1 + cc;
        float2 o[3][2];
        conv3_quad<S1_W, S1_RP * S1_W>(s1a + (2 * gi) * S1_W + cc, s1b + (2 * gi) * S1_W + cc, wt.w2, wt.b2, o);
        const bool xin = xx >= 0 && xx < W;
        const bool in0 = xin && yy >= 0 && yy < H, in1 = xin && yy + 1 >= 0 && yy + 1 < H;
        const bool in2 = xin && yy + 2 >= 0 && yy + 2 < H, in3 = xin && yy + 3 >= 0 && yy + 3 < H;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            // residual: fea1 at the same position = fea1 rows 4gi+1 .. 4gi+4, column cc+1: copy B pair rows 2gi, 2gi+1
            const float2 ra = s1b[(c * S1_RP + 2 * gi) * S1_W + cc + 1], rb = s1b[(c * S1_RP + 2 * gi + 1) * S1_W + cc + 1];
            store_quad<S2_W>(s2a + c * S2_RP * S2_W, s2b + c * S2_RP * S2_W, gi, cc,
                             in0 ? ra.x + fmaxf(o[c][0].x, 0.f) : 0.f, in1 ? ra.y + fmaxf(o[c][0].y, 0.f) : 0.f,
                             in2 ? rb.x + fmaxf(o[c][1].x, 0.f) : 0.f, in3 ? rb.y + fmaxf(o[c][1].y, 0.f) : 0.f);
        }
    }
    __syncthreads();
    if (sci_item<S3_G, S_TW>(t, gi, cc)) {
        const int xx = x0 + cc;
        if (xx < W) {
            float2 o[3][2];
            conv3_quad<S2_W, S2_RP * S2_W>(s2a + (2 * gi) * S2_W + cc, s2b + (2 * gi) * S2_W + cc, wt.w3, wt.b3, o);
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int yy = y0 + 4 * gi + q;
                if (yy >= H) continue;
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    const float xin = __ldg(xb + c * plane + (size_t)yy * W + xx);
                    const float ov = (q & 1) ? o[c][q >> 1].y : o[c][q >> 1].x;
                    // sigmoid and x / illu on the SFU (ex2.approx / rcp.approx, ~2 ulp): the IEEE-rounded forms cost as
                    // many FMA-pipe instructions per pixel as a whole conv layer; the contract is 1e-3
                    const float sg = __fdividef(1.f, 1.f + __expf(-ov));
                    const float il = fminf(fmaxf(sg + xin, 0.0001f), 1.f);
                    const float rv = fminf(fmaxf(__fdividef(xin, il), 0.f), 1.f);
                    const size_t off = ((size_t)(b * 3 + c)) * plane + (size_t)yy * W + xx;
                    if (illu) illu[off] = il;
                    r[off] = rv;
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
    if (B > 65535) return B200_ERR_UNSUPPORTED;
    if (!b200::aligned16(x) || !b200::aligned16(r) || (illu && !b200::aligned16(illu))) return B200_ERR_MISALIGNED;
    b200::SciWeights wt;
    memcpy(&wt, packed_host, sizeof(wt));
    dim3 grid((W + b200::S_TW - 1) / b200::S_TW, (H + b200::S_TH - 1) / b200::S_TH, B);
    if (cudaFuncSetAttribute(b200::sci_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, b200::SCI_SMEM) !=
        cudaSuccess) {
        (void)cudaGetLastError();
        return B200_ERR_LAUNCH;
    }
    b200::sci_fused_kernel<<<grid, 256, b200::SCI_SMEM, (cudaStream_t)stream>>>(x, illu, r, wt, H, W);
    return b200::launch_status();
}
