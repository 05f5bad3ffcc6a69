// SCI illumination-estimation enhancer (Finetunemodel.forward, sci/model.py:155-159 over
// EnhanceNetwork.forward :36-45, layers=1, channels=3, BatchNorm2d eval) as ONE fused pass:
//   fea1 = relu(conv3x3(x));  fea2 = fea1 + relu(bn(conv3x3(fea1)));  fea3 = sigmoid(conv3x3(fea2));
//   illu = clamp(fea3 + x, 1e-4, 1);  r = clamp(x / illu, 0, 1)
// CTA tile 16x32 output pixels; the three stacked 3x3 layers are evaluated on shrinking halo regions held in
// shared memory (x: 22x38, fea1: 20x36, fea2: 18x34), so x is read once from HBM and illu / r written once.
// All 252 weights travel as a by-value kernel parameter: every FFMA takes its weight straight from the
// constant bank (no weight loads at all).
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

constexpr int S_TH = 16, S_TW = 32;
constexpr int SX_H = S_TH + 6, SX_W = 44;         // 22 rows x (4 + 32 + 8) columns: image column x0 - 4 at column 0,
                                                  // so every row is eleven 16-byte cp.async chunks
constexpr int S1_H = S_TH + 4, S1_W = S_TW + 4;  // 20 x 36
constexpr int S2_H = S_TH + 2, S2_W = S_TW + 2;  // 18 x 34

// 3x3x3 -> 3 convolution at TWO vertically adjacent positions (rows r, r+1 of the same column) on the packed fp32
// pipe: each FFMA2 carries one weight (scalar broadcast) times the value pair (in[r+ky][..], in[r+1+ky][..]).
// `in` points at the top-left tap of channel 0 for the upper position.  out[co] = (upper, lower).
template <int IN_W, int IN_PLANE>
__device__ __forceinline__ void conv3_pair(const float* __restrict__ in, const float* w, const float* bias,
                                           float2 out[3]) {
#pragma unroll
    for (int co = 0; co < 3; ++co) out[co] = make_float2(bias[co], bias[co]);
#pragma unroll
    for (int ci = 0; ci < 3; ++ci) {
#pragma unroll
        for (int kx = 0; kx < 3; ++kx) {
            float v[4];
#pragma unroll
            for (int rr = 0; rr < 4; ++rr) v[rr] = in[ci * IN_PLANE + rr * IN_W + kx];
#pragma unroll
            for (int ky = 0; ky < 3; ++ky) {
                const float2 vp = make_float2(v[ky], v[ky + 1]);
#pragma unroll
                for (int co = 0; co < 3; ++co) {
                    const float wv = w[((co * 3 + ci) * 3 + ky) * 3 + kx];
                    out[co] = __ffma2_rn(make_float2(wv, wv), vp, out[co]);
                }
            }
        }
    }
}

__global__ void __launch_bounds__(256)
sci_fused_kernel(const float* __restrict__ x, float* __restrict__ illu, float* __restrict__ r,
                 const __grid_constant__ SciWeights wt, int H, int W) {
    __shared__ __align__(16) float sx[3 * SX_H * SX_W];
    __shared__ float s1[3 * S1_H * S1_W];
    __shared__ float s2[3 * S2_H * S2_W];
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
    // fea1 = relu(conv(x)) on the 20x36 region; zero outside the image (that is the next conv's zero padding)
    for (int idx = t; idx < (S1_H / 2) * S1_W; idx += 256) {
        const int rp = idx / S1_W, cc = idx - rp * S1_W;
        const int rr = 2 * rp;
        const int yy = y0 - 2 + rr, xx = x0 - 2 + cc;
        float2 o[3];
        conv3_pair<SX_W, SX_H * SX_W>(sx + rr * SX_W + cc + 1, wt.w1, wt.b1, o);
        const bool xin = xx >= 0 && xx < W;
        const bool in0 = xin && yy >= 0 && yy < H, in1 = xin && yy + 1 >= 0 && yy + 1 < H;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            s1[c * S1_H * S1_W + rr * S1_W + cc] = in0 ? fmaxf(o[c].x, 0.f) : 0.f;
            s1[c * S1_H * S1_W + (rr + 1) * S1_W + cc] = in1 ? fmaxf(o[c].y, 0.f) : 0.f;
        }
    }
    __syncthreads();
    // fea2 = fea1 + relu(bn(conv(fea1))) on the 18x34 region
    for (int idx = t; idx < (S2_H / 2) * S2_W; idx += 256) {
        const int rp = idx / S2_W, cc = idx - rp * S2_W;
        const int rr = 2 * rp;
        const int yy = y0 - 1 + rr, xx = x0 - 1 + cc;
        float2 o[3];
        conv3_pair<S1_W, S1_H * S1_W>(s1 + rr * S1_W + cc, wt.w2, wt.b2, o);
        const bool xin = xx >= 0 && xx < W;
        const bool in0 = xin && yy >= 0 && yy < H, in1 = xin && yy + 1 >= 0 && yy + 1 < H;
        const int ctr = (rr + 1) * S1_W + cc + 1;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            s2[c * S2_H * S2_W + rr * S2_W + cc] = in0 ? s1[c * S1_H * S1_W + ctr] + fmaxf(o[c].x, 0.f) : 0.f;
            s2[c * S2_H * S2_W + (rr + 1) * S2_W + cc] =
                in1 ? s1[c * S1_H * S1_W + ctr + S1_W] + fmaxf(o[c].y, 0.f) : 0.f;
        }
    }
    __syncthreads();
    for (int idx = t; idx < (S_TH / 2) * S_TW; idx += 256) {
        const int rp = idx / S_TW, cc = idx - rp * S_TW;
        const int rr = 2 * rp;
        const int xx = x0 + cc;
        if (xx >= W) continue;
        float2 o[3];
        conv3_pair<S2_W, S2_H * S2_W>(s2 + rr * S2_W + cc, wt.w3, wt.b3, o);
#pragma unroll
        for (int half = 0; half < 2; ++half) {
            const int yy = y0 + rr + half;
            if (yy >= H) continue;
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                const float xin = sx[c * SX_H * SX_W + (rr + half + 3) * SX_W + cc + 4];
                const float ov = half ? o[c].y : o[c].x;
                const float sg = 1.f / (1.f + expf(-ov));
                const float il = fminf(fmaxf(sg + xin, 0.0001f), 1.f);
                const float rv = fminf(fmaxf(__fdiv_rn(xin, il), 0.f), 1.f);
                const size_t off = ((size_t)(b * 3 + c)) * plane + (size_t)yy * W + xx;
                if (illu) illu[off] = il;
                r[off] = rv;
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
    b200::sci_fused_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(x, illu, r, wt, H, W);
    return b200::launch_status();
}
