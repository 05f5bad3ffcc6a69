// RUAS inference (upstream Network.forward, ruas/model.py:202-206): three IEM iterations (EnhanceNetwork.forward
// :117-125, IEM.forward :84-95) followed by the optional denoise net (DenoiseNetwork.forward :142-149).  Both nets are
// stacks of SearchBlock cells (:46-66) whose ops are fixed by genotypes.py:15-16:
//   IEM cell (3 ch): c1_d skip, c1_r res 1x1, c2_d res dilated 3x3, c2_r conv 3x3, c3_d conv 3x3, c3_r skip, c4 conv 3x3
//   NRM cell (6 ch): c1_d res 1x1, c1_r res 1x1, c2_d res dilated 3x3, c2_r skip, c3_d res 1x1, c3_r res 1x1, c4 skip
// ONE launch per cell: a CTA owns a 12 x 32 pixel tile, keeps the cell's intermediates (r_c1 with a 2-pixel halo,
// r_c2 / r_c3 with a 1-pixel halo) in shared memory and writes only the cell's output, so every cell reads its input
// once and writes its result once.  Like SCI the arithmetic (~370 / ~610 MAC per pixel and cell on 3 / 6 channels) is
// bound by the fp32 FMA pipe: a thread computes a 1 x 2 pixel strip for all output channels, activations come from
// 64-bit shared loads, and the cell's weights travel as a by-value kernel parameter so every FFMA takes its weight
// straight from the constant bank.
#include <math.h>
#include <string.h>

#include "common.cuh"

namespace b200 {

struct RuasIemCell {  // 387 floats
    float c1r_w[9], c1r_b[3], c2d_w[81], c2d_b[3], c2r_w[81], c2r_b[3], c3d_w[81], c3d_b[3], c4_w[81], c4_b[3],
        c5_w[36], c5_b[3];
};
struct RuasNrmCell {  // 648 floats
    float c1d_w[36], c1d_b[6], c1r_w[36], c1r_b[6], c2d_w[324], c2d_b[6], c3d_w[36], c3d_b[6], c3r_w[36], c3r_b[6],
        c5_w[144], c5_b[6];
};
struct RuasStem {  // 168 floats: conv 3x3, 3 -> 6
    float w[162], b[6];
};
struct RuasTail {  // 165 floats: conv 3x3, 6 -> 3
    float w[162], b[3];
};
struct RuasPacked {
    RuasIemCell iem[3];
    RuasStem stem;
    RuasNrmCell nrm[3];
    RuasTail tail;
};
static_assert(sizeof(RuasPacked) == B200_RUAS_PACKED_FLOATS * sizeof(float), "packed size");

constexpr int R_TH = 12, R_TW = 32, R_THREADS = 256;
constexpr int R1_H = R_TH + 4, R1_W = R_TW + 4;  // r_c1 region: rows y0-2 .. y0+13, columns x0-2 .. x0+33
constexpr int R2_H = R_TH + 2, R2_W = R_TW + 2;  // r_c2 region: rows y0-1 .. y0+12, columns x0-1 .. x0+32
constexpr int R_STRIPS = R_TH * (R_TW / 2);      // 192 centre strips of 1 x 2 pixels

__device__ __forceinline__ float lrelu(float v) { return v > 0.f ? v : v * 0.05f; }

// acc[co][j] += sum_{ci,ky,kx} w[co][ci][ky][kx] * src[ci][row + (ky-1)*DIL][col + j + (kx-1)*DIL],  j = 0, 1.
// `s` points at src[0][row][col - DIL] and must be 8-byte aligned (even RS, PS and column).
template <int CI, int CO, int DIL, int PS, int RS>
__device__ __forceinline__ void conv3_strip2(const float* __restrict__ s, const float* __restrict__ w,
                                             float (&acc)[CO][2]) {
#pragma unroll
    for (int ci = 0; ci < CI; ++ci) {
#pragma unroll
        for (int ky = 0; ky < 3; ++ky) {
            const float* p = s + ci * PS + (ky - 1) * DIL * RS;
            float v[2 + 2 * DIL];
#pragma unroll
            for (int q = 0; q < 1 + DIL; ++q) {
                const float2 t = *reinterpret_cast<const float2*>(p + 2 * q);
                v[2 * q] = t.x;
                v[2 * q + 1] = t.y;
            }
#pragma unroll
            for (int kx = 0; kx < 3; ++kx) {
#pragma unroll
                for (int co = 0; co < CO; ++co) {
                    const float wv = w[((co * CI + ci) * 3 + ky) * 3 + kx];
                    acc[co][0] = fmaf(wv, v[kx * DIL], acc[co][0]);
                    acc[co][1] = fmaf(wv, v[kx * DIL + 1], acc[co][1]);
                }
            }
        }
    }
}

// out[co] = b[co] + sum_ci w[co][ci] * in[ci]
template <int CI, int CO>
__device__ __forceinline__ void conv1(const float* __restrict__ w, const float* __restrict__ b, const float (&in)[CI],
                                      float (&out)[CO]) {
#pragma unroll
    for (int co = 0; co < CO; ++co) {
        float a = b[co];
#pragma unroll
        for (int ci = 0; ci < CI; ++ci) a = fmaf(w[co * CI + ci], in[ci], a);
        out[co] = a;
    }
}

// out[co] += sum_ci w[co][OFF + ci] * in[ci]  (one quarter of the 4C -> C fusion conv c5)
template <int C, int OFF>
__device__ __forceinline__ void c5_part(const float* __restrict__ w, const float (&in)[C], float (&out)[C]) {
#pragma unroll
    for (int co = 0; co < C; ++co)
#pragma unroll
        for (int ci = 0; ci < C; ++ci) out[co] = fmaf(w[co * 4 * C + OFF + ci], in[ci], out[co]);
}

// ------------------------------------------------------------------------------------------------ IEM iteration
// K0: first iteration (t_hat = max_operation(y));  else t_hat = max_operation(u) - 0.5 (u - y)   (model.py:84-89)
template <bool K0>
__global__ void __launch_bounds__(R_THREADS) ruas_iem_kernel(const float* __restrict__ y, const float* __restrict__ u,
                                                             float* __restrict__ u_out, float* __restrict__ t_out,
                                                             const RuasIemCell c, int H, int W) {
    __shared__ __align__(16) float s_t[3 * R1_H * R1_W];   // t_hat
    __shared__ __align__(16) float s_r1[3 * R1_H * R1_W];  // r_c1, zero outside the image (= the convs' padding)
    __shared__ __align__(16) float s_r2[3 * R2_H * R2_W];  // r_c2
    __shared__ __align__(16) float s_r3[3 * R2_H * R2_W];  // r_c3 = act(r_c2 + r_c2)  (c3_r is a skip)
    const int tid = threadIdx.x, x0 = blockIdx.x * R_TW, y0 = blockIdx.y * R_TH;
    const size_t plane = (size_t)H * W;
    const float* yb = y + (size_t)blockIdx.z * 3 * plane;
    const float* ub = K0 ? yb : u + (size_t)blockIdx.z * 3 * plane;

    // phase 0: t_hat and r_c1 = act((conv1x1(t) + t) + t) over the halo-2 region
    for (int i = tid; i < R1_H * R1_W; i += R_THREADS) {
        const int ry = i / R1_W, rx = i - ry * R1_W, gy = y0 - 2 + ry, gx = x0 - 2 + rx;
        float t[3] = {0.f, 0.f, 0.f}, r1[3] = {0.f, 0.f, 0.f};
        if (gy >= 0 && gy < H && gx >= 0 && gx < W) {
            const bool hr = gx + 1 < W, hd = gy + 1 < H;
#pragma unroll
            for (int ch = 0; ch < 3; ++ch) {
                const float* p = ub + ch * plane + (size_t)gy * W + gx;
                const float v = __ldg(p);
                float m = fmaxf(v, hr ? __ldg(p + 1) : 0.f);
                m = fmaxf(m, fmaxf(hd ? __ldg(p + W) : 0.f, hr && hd ? __ldg(p + W + 1) : 0.f));
                t[ch] = K0 ? m : m - 0.5f * (v - __ldg(yb + ch * plane + (size_t)gy * W + gx));
            }
            float o[3];
            conv1<3, 3>(c.c1r_w, c.c1r_b, t, o);
#pragma unroll
            for (int ch = 0; ch < 3; ++ch) r1[ch] = lrelu((o[ch] + t[ch]) + t[ch]);
        }
#pragma unroll
        for (int ch = 0; ch < 3; ++ch) {
            s_t[ch * R1_H * R1_W + i] = t[ch];
            s_r1[ch * R1_H * R1_W + i] = r1[ch];
        }
    }
    __syncthreads();

    // phase 1: r_c2 = act(conv3x3(r_c1) + r_c1), r_c3 = act(2 r_c2) over the halo-1 region (14 x 17 strips)
    if (tid < R2_H * (R2_W / 2)) {
        const int ry = tid / (R2_W / 2), j = tid - ry * (R2_W / 2);
        const int gy = y0 - 1 + ry, gx = x0 - 1 + 2 * j;
        float acc[3][2];
#pragma unroll
        for (int co = 0; co < 3; ++co) acc[co][0] = acc[co][1] = c.c2r_b[co];
        const float* s = s_r1 + (ry + 1) * R1_W + 2 * j;  // region column 2j = r_c1 column 2j+1; minus the tap offset
        conv3_strip2<3, 3, 1, R1_H * R1_W, R1_W>(s, c.c2r_w, acc);
#pragma unroll
        for (int co = 0; co < 3; ++co)
#pragma unroll
            for (int q = 0; q < 2; ++q) {
                const bool in = gy >= 0 && gy < H && gx + q >= 0 && gx + q < W;
                const float r2 = in ? lrelu(acc[co][q] + s[co * R1_H * R1_W + 1 + q]) : 0.f;
                s_r2[co * R2_H * R2_W + ry * R2_W + 2 * j + q] = r2;
                s_r3[co * R2_H * R2_W + ry * R2_W + 2 * j + q] = lrelu(r2 + r2);
            }
    }
    __syncthreads();

    // phase 2: the four distilled maps at the centre, c5, sigmoid, clamp, u = clamp(y / t)
    if (tid < R_STRIPS) {
        const int ry = tid / (R_TW / 2), j = tid - ry * (R_TW / 2);
        const int gy = y0 + ry, gx = x0 + 2 * j;
        if (gy >= H || gx >= W) return;
        float o[2][3];
        float d2[3][2], d3[3][2], r4[3][2];
#pragma unroll
        for (int co = 0; co < 3; ++co) {
            d2[co][0] = d2[co][1] = c.c2d_b[co];
            d3[co][0] = d3[co][1] = c.c3d_b[co];
            r4[co][0] = r4[co][1] = c.c4_b[co];
        }
        const float* s1 = s_r1 + (ry + 2) * R1_W + 2 * j;  // centre column 2j = r_c1 column 2j+2; minus dilation 2
        conv3_strip2<3, 3, 2, R1_H * R1_W, R1_W>(s1, c.c2d_w, d2);
        const int o2 = (ry + 1) * R2_W + 2 * j;  // centre column 2j = r_c2 column 2j+1; minus the tap offset
        conv3_strip2<3, 3, 1, R2_H * R2_W, R2_W>(s_r2 + o2, c.c3d_w, d3);
        conv3_strip2<3, 3, 1, R2_H * R2_W, R2_W>(s_r3 + o2, c.c4_w, r4);
#pragma unroll
        for (int q = 0; q < 2; ++q) {
            float d1[3], e2[3], e3[3], e4[3];
#pragma unroll
            for (int ch = 0; ch < 3; ++ch) {
                d1[ch] = lrelu(s_t[ch * R1_H * R1_W + (ry + 2) * R1_W + 2 * j + 2 + q]);
                e2[ch] = lrelu(d2[ch][q] + s1[ch * R1_H * R1_W + 2 + q]);
                e3[ch] = lrelu(d3[ch][q]);
                e4[ch] = lrelu(r4[ch][q]);
                o[q][ch] = c.c5_b[ch];
            }
            c5_part<3, 0>(c.c5_w, d1, o[q]);
            c5_part<3, 3>(c.c5_w, e2, o[q]);
            c5_part<3, 6>(c.c5_w, e3, o[q]);
            c5_part<3, 9>(c.c5_w, e4, o[q]);
        }
        const size_t base = (size_t)blockIdx.z * 3 * plane + (size_t)gy * W + gx;
#pragma unroll
        for (int ch = 0; ch < 3; ++ch)
#pragma unroll
            for (int q = 0; q < 2; ++q) {
                if (gx + q >= W) continue;
                float t = 1.f / (1.f + __expf(-o[q][ch]));
                t = fminf(fmaxf(t, 0.001f), 1.f);
                const float yv = __ldg(y + base + ch * plane + q);
                const float uv = fminf(fmaxf(yv / t, 0.f), 1.f);
                if (t_out) t_out[base + ch * plane + q] = t;
                u_out[base + ch * plane + q] = uv;
            }
    }
}

// ------------------------------------------------------------------------------------------------ NRM cell
__global__ void __launch_bounds__(R_THREADS) ruas_nrm_kernel(const float* __restrict__ f, float* __restrict__ out,
                                                             const RuasNrmCell c, int H, int W) {
    __shared__ __align__(16) float s_r1[6 * R1_H * R1_W];
    const int tid = threadIdx.x, x0 = blockIdx.x * R_TW, y0 = blockIdx.y * R_TH;
    const size_t plane = (size_t)H * W;
    const float* fb = f + (size_t)blockIdx.z * 6 * plane;

    // phase 0: r_c1 = act((conv1x1(f) + f) + f) over the halo-2 region, zero outside the image
    for (int i = tid; i < R1_H * R1_W; i += R_THREADS) {
        const int ry = i / R1_W, rx = i - ry * R1_W, gy = y0 - 2 + ry, gx = x0 - 2 + rx;
        float r1[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        if (gy >= 0 && gy < H && gx >= 0 && gx < W) {
            float v[6], o[6];
#pragma unroll
            for (int ch = 0; ch < 6; ++ch) v[ch] = __ldg(fb + ch * plane + (size_t)gy * W + gx);
            conv1<6, 6>(c.c1r_w, c.c1r_b, v, o);
#pragma unroll
            for (int ch = 0; ch < 6; ++ch) r1[ch] = lrelu((o[ch] + v[ch]) + v[ch]);
        }
#pragma unroll
        for (int ch = 0; ch < 6; ++ch) s_r1[ch * R1_H * R1_W + i] = r1[ch];
    }
    __syncthreads();

    if (tid < R_STRIPS) {
        const int ry = tid / (R_TW / 2), j = tid - ry * (R_TW / 2);
        const int gy = y0 + ry, gx = x0 + 2 * j;
        if (gy >= H || gx >= W) return;
        float d2[6][2];
#pragma unroll
        for (int co = 0; co < 6; ++co) d2[co][0] = d2[co][1] = c.c2d_b[co];
        const float* s1 = s_r1 + (ry + 2) * R1_W + 2 * j;
        conv3_strip2<6, 6, 2, R1_H * R1_W, R1_W>(s1, c.c2d_w, d2);
        const size_t base = (size_t)gy * W + gx;
#pragma unroll
        for (int q = 0; q < 2; ++q) {
            if (gx + q >= W) continue;
            float v[6], r1[6], r2[6], t[6], o[6];
#pragma unroll
            for (int ch = 0; ch < 6; ++ch) {
                v[ch] = __ldg(fb + ch * plane + base + q);
                r1[ch] = s1[ch * R1_H * R1_W + 2 + q];
                o[ch] = c.c5_b[ch];
            }
            conv1<6, 6>(c.c1d_w, c.c1d_b, v, t);  // distilled_c1 = act(conv1x1(f) + f)
#pragma unroll
            for (int ch = 0; ch < 6; ++ch) t[ch] = lrelu(t[ch] + v[ch]);
            c5_part<6, 0>(c.c5_w, t, o);
#pragma unroll
            for (int ch = 0; ch < 6; ++ch) {  // distilled_c2 = act(dilconv(r_c1) + r_c1);  r_c2 = act(r_c1 + r_c1)
                t[ch] = lrelu(d2[ch][q] + r1[ch]);
                r2[ch] = lrelu(r1[ch] + r1[ch]);
            }
            c5_part<6, 6>(c.c5_w, t, o);
            conv1<6, 6>(c.c3d_w, c.c3d_b, r2, t);  // distilled_c3 = act(conv1x1(r_c2) + r_c2)
#pragma unroll
            for (int ch = 0; ch < 6; ++ch) t[ch] = lrelu(t[ch] + r2[ch]);
            c5_part<6, 12>(c.c5_w, t, o);
            conv1<6, 6>(c.c3r_w, c.c3r_b, r2, t);  // r_c3 = act((conv1x1(r_c2) + r_c2) + r_c2);  r_c4 = act(r_c3)
#pragma unroll
            for (int ch = 0; ch < 6; ++ch) t[ch] = lrelu(lrelu((t[ch] + r2[ch]) + r2[ch]));
            c5_part<6, 18>(c.c5_w, t, o);
#pragma unroll
            for (int ch = 0; ch < 6; ++ch) out[(size_t)blockIdx.z * 6 * plane + ch * plane + base + q] = o[ch];
        }
    }
}

// ------------------------------------------------------------------------------------------------ stem / tail
// 3x3 convolution CI -> CO with zero padding.  TAIL: n = conv(src); out = sub_from - n (model.py:147-148), n optional.
template <int CI, int CO, bool TAIL, typename WT>
__global__ void __launch_bounds__(R_THREADS) ruas_conv3_kernel(const float* __restrict__ src, float* __restrict__ out,
                                                               const float* __restrict__ sub_from,
                                                               float* __restrict__ n_out, const WT c, int H, int W) {
    __shared__ __align__(16) float s_in[CI * R2_H * R2_W];
    const int tid = threadIdx.x, x0 = blockIdx.x * R_TW, y0 = blockIdx.y * R_TH;
    const size_t plane = (size_t)H * W;
    const float* sb = src + (size_t)blockIdx.z * CI * plane;
    for (int i = tid; i < CI * R2_H * R2_W; i += R_THREADS) {
        const int ch = i / (R2_H * R2_W), r = i - ch * (R2_H * R2_W);
        const int ry = r / R2_W, rx = r - ry * R2_W, gy = y0 - 1 + ry, gx = x0 - 1 + rx;
        s_in[i] = gy >= 0 && gy < H && gx >= 0 && gx < W ? __ldg(sb + ch * plane + (size_t)gy * W + gx) : 0.f;
    }
    __syncthreads();
    if (tid < R_STRIPS) {
        const int ry = tid / (R_TW / 2), j = tid - ry * (R_TW / 2);
        const int gy = y0 + ry, gx = x0 + 2 * j;
        if (gy >= H || gx >= W) return;
        float acc[CO][2];
#pragma unroll
        for (int co = 0; co < CO; ++co) acc[co][0] = acc[co][1] = c.b[co];
        conv3_strip2<CI, CO, 1, R2_H * R2_W, R2_W>(s_in + (ry + 1) * R2_W + 2 * j, c.w, acc);
        const size_t base = (size_t)blockIdx.z * CO * plane + (size_t)gy * W + gx;
#pragma unroll
        for (int co = 0; co < CO; ++co)
#pragma unroll
            for (int q = 0; q < 2; ++q) {
                if (gx + q >= W) continue;
                const size_t o = base + co * plane + q;
                if (TAIL) {
                    if (n_out) n_out[o] = acc[co][q];
                    out[o] = __ldg(sub_from + o) - acc[co][q];
                } else {
                    out[o] = acc[co][q];
                }
            }
    }
}

}  // namespace b200

extern "C" size_t b200_ruas_workspace_bytes(int B, int H, int W, int with_denoise) {
    if (B <= 0 || H <= 0 || W <= 0 || !with_denoise) return 0;
    return 2 * (size_t)B * 6 * H * W * sizeof(float);
}

extern "C" int b200_ruas_forward(const float* x, const float* packed_host, float* u, float* t_or_null,
                                 float* denoised_or_null, float* noise_or_null, int B, int H, int W, void* workspace,
                                 size_t workspace_bytes, void* stream) {
    using namespace b200;
    if (!x || !packed_host || !u) return B200_ERR_NULL;
    if (B <= 0 || H <= 0 || W <= 0) return B200_ERR_BAD_SHAPE;
    if (B > 65535) return B200_ERR_UNSUPPORTED;
    if (!aligned16(x) || !aligned16(u) || (t_or_null && !aligned16(t_or_null)) ||
        (denoised_or_null && !aligned16(denoised_or_null)) || (noise_or_null && !aligned16(noise_or_null)))
        return B200_ERR_MISALIGNED;
    if (noise_or_null && !denoised_or_null) return B200_ERR_NULL;
    if (denoised_or_null && (!workspace || workspace_bytes < b200_ruas_workspace_bytes(B, H, W, 1)))
        return B200_ERR_WORKSPACE;
    RuasPacked pk;  // by-value kernel parameters are taken from this host copy
    memcpy(&pk, packed_host, sizeof(pk));
    const RuasPacked* p = &pk;
    cudaStream_t s = (cudaStream_t)stream;
    const dim3 grid((W + R_TW - 1) / R_TW, (H + R_TH - 1) / R_TH, B);
    const size_t img = (size_t)B * 3 * H * W;
    for (int k = 0; k < 3; ++k) {
        float* uo = u + k * img;
        float* to = t_or_null ? t_or_null + k * img : nullptr;
        if (k == 0)
            ruas_iem_kernel<true><<<grid, R_THREADS, 0, s>>>(x, nullptr, uo, to, p->iem[0], H, W);
        else
            ruas_iem_kernel<false><<<grid, R_THREADS, 0, s>>>(x, u + (k - 1) * img, uo, to, p->iem[k], H, W);
    }
    if (denoised_or_null) {
        float* fa = static_cast<float*>(workspace);
        float* fb = fa + (size_t)B * 6 * H * W;
        const float* u3 = u + 2 * img;
        ruas_conv3_kernel<3, 6, false, RuasStem><<<grid, R_THREADS, 0, s>>>(u3, fa, nullptr, nullptr, p->stem, H, W);
        ruas_nrm_kernel<<<grid, R_THREADS, 0, s>>>(fa, fb, p->nrm[0], H, W);
        ruas_nrm_kernel<<<grid, R_THREADS, 0, s>>>(fb, fa, p->nrm[1], H, W);
        ruas_nrm_kernel<<<grid, R_THREADS, 0, s>>>(fa, fb, p->nrm[2], H, W);
        ruas_conv3_kernel<6, 3, true, RuasTail>
            <<<grid, R_THREADS, 0, s>>>(fb, denoised_or_null, u3, noise_or_null, p->tail, H, W);
    }
    return launch_status();
}
