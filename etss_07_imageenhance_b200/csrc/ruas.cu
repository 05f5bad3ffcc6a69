// RUAS inference (upstream Network.forward, ruas/model.py:202-206): three IEM iterations (EnhanceNetwork.forward
// :117-125, IEM.forward :84-95) followed by the optional denoise net (DenoiseNetwork.forward :142-149).  Both nets are
// stacks of SearchBlock cells (:46-66) whose ops are fixed by genotypes.py:15-16:
//   IEM cell (3 ch): c1_d skip, c1_r res 1x1, c2_d res dilated 3x3, c2_r conv 3x3, c3_d conv 3x3, c3_r skip, c4 conv 3x3
//   NRM cell (6 ch): c1_d res 1x1, c1_r res 1x1, c2_d res dilated 3x3, c2_r skip, c3_d res 1x1, c3_r res 1x1, c4 skip
// ONE launch per cell: a CTA owns a 12 x 32 (IEM) or 16 x 32 (NRM, stem, tail) pixel tile, keeps the cell's intermediates (r_c1 with a 2-pixel halo,
// r_c2 / r_c3 with a 1-pixel halo) in shared memory and writes only the cell's output, so every cell reads its input
// once and writes its result once.  Like SCI the arithmetic (~370 / ~610 MAC per pixel and cell on 3 / 6 channels) is
// bound by the fp32 FMA pipe: a thread computes a 1 x 2 pixel strip for all output channels on the PACKED pipe (every
// FFMA2 = one weight x a horizontally adjacent activation pair, which is the 64-bit word the shared load returned), and
// the cell's weights travel as a by-value kernel parameter, i.e. they are read from the constant bank.
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

constexpr int R_TW = 32, R_THREADS = 256;
constexpr int IEM_TH = 12;  // r_c2 region = 14 x 17 strips (238 of 256 threads), centre = 192 strips
constexpr int NRM_TH = 16;  // centre = 256 strips
// tile geometry for a tile height TH: r_c1 region rows y0-2 .. y0+TH+1, columns x0-2 .. x0+33;
// r_c2 region rows y0-1 .. y0+TH, columns x0-1 .. x0+32
constexpr int R1_W = R_TW + 4, R2_W = R_TW + 2;

// LeakyReLU(0.05) (model.py:43) = max(v, 0.05 v): one FMUL + one FMNMX
__device__ __forceinline__ float lrelu(float v) { return fmaxf(v, v * 0.05f); }

__device__ __forceinline__ float2 lrelu2(float2 v) { return make_float2(lrelu(v.x), lrelu(v.y)); }
__device__ __forceinline__ float2 add2(float2 a, float2 b) { return __fadd2_rn(a, b); }
__device__ __forceinline__ float2 bcast(float w) { return make_float2(w, w); }

// Packed strip convolution: acc[co] = (pixel col, pixel col+1) accumulators,
//   acc[co] += sum_{ci,ky,kx} w[co][ci][ky][kx] * src[ci][row + (ky-1)*DIL][col + {0,1} + (kx-1)*DIL].
// `s` points at src[0][row][col - DIL] and must be 8-byte aligned (even RS, PS and column).  Every FFMA2 multiplies a
// horizontally adjacent value pair by one weight broadcast from the constant bank; with DIL = 2 all three tap pairs are
// the 64-bit words as loaded, with DIL = 1 the centre tap pair is assembled from two of them.
template <int CI, int CO, int DIL, int PS, int RS>
__device__ __forceinline__ void conv3_strip2(const float* __restrict__ s, const float* __restrict__ w,
                                             float2 (&acc)[CO]) {
#pragma unroll
    for (int ci = 0; ci < CI; ++ci) {
#pragma unroll
        for (int ky = 0; ky < 3; ++ky) {
            const float2* p = reinterpret_cast<const float2*>(s + ci * PS + (ky - 1) * DIL * RS);
            float2 v[3];
            if (DIL == 2) {
                v[0] = p[0], v[1] = p[1], v[2] = p[2];
            } else {
                v[0] = p[0], v[2] = p[1];
                v[1] = make_float2(v[0].y, v[2].x);
            }
#pragma unroll
            for (int kx = 0; kx < 3; ++kx) {
#pragma unroll
                for (int co = 0; co < CO; ++co)
                    acc[co] = __ffma2_rn(bcast(w[((co * CI + ci) * 3 + ky) * 3 + kx]), v[kx], acc[co]);
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
// the same on pixel pairs
template <int CI, int CO>
__device__ __forceinline__ void conv1(const float* __restrict__ w, const float* __restrict__ b, const float2 (&in)[CI],
                                      float2 (&out)[CO]) {
#pragma unroll
    for (int co = 0; co < CO; ++co) {
        float2 a = bcast(b[co]);
#pragma unroll
        for (int ci = 0; ci < CI; ++ci) a = __ffma2_rn(bcast(w[co * CI + ci]), in[ci], a);
        out[co] = a;
    }
}

// out[co] += sum_ci w[co][OFF + ci] * in[ci]  (one quarter of the 4C -> C fusion conv c5), on pixel pairs
template <int C, int OFF>
__device__ __forceinline__ void c5_part(const float* __restrict__ w, const float2 (&in)[C], float2 (&out)[C]) {
#pragma unroll
    for (int co = 0; co < C; ++co)
#pragma unroll
        for (int ci = 0; ci < C; ++ci) out[co] = __ffma2_rn(bcast(w[co * 4 * C + OFF + ci]), in[ci], out[co]);
}

// ------------------------------------------------------------------------------------------------ IEM iteration
// K0: first iteration (t_hat = max_operation(y));  else t_hat = max_operation(u) - 0.5 (u - y)   (model.py:84-89)
template <bool K0>
__global__ void __launch_bounds__(R_THREADS, 6) ruas_iem_kernel(const float* __restrict__ y, const float* __restrict__ u,
                                                             float* __restrict__ u_out, float* __restrict__ t_out,
                                                             const RuasIemCell c, int H, int W) {
    constexpr int TH = IEM_TH, R1_H = TH + 4, R2_H = TH + 2, P1 = R1_H * R1_W, P2 = R2_H * R2_W;
    __shared__ __align__(16) float s_t[3 * P1];   // t_hat
    __shared__ __align__(16) float s_r1[3 * P1];  // r_c1, zero outside the image (= the convs' padding)
    __shared__ __align__(16) float s_r2[3 * P2];  // r_c2
    __shared__ __align__(16) float s_r3[3 * P2];  // r_c3 = act(r_c2 + r_c2)  (c3_r is a skip)
    const int tid = threadIdx.x, x0 = blockIdx.x * R_TW, y0 = blockIdx.y * TH;
    const size_t plane = (size_t)H * W;
    const float* yb = y + (size_t)blockIdx.z * 3 * plane;
    const float* ub = K0 ? yb : u + (size_t)blockIdx.z * 3 * plane;

    // phase 0: t_hat and r_c1 = act((conv1x1(t) + t) + t) over the halo-2 region, two horizontally adjacent positions
    // per thread (max_operation of the pair needs columns gx .. gx+2 of rows gy, gy+1)
    const bool w_even = (W & 1) == 0;
    for (int i = tid; i < R1_H * (R1_W / 2); i += R_THREADS) {
        const int ry = i / (R1_W / 2), jx = i - ry * (R1_W / 2), gy = y0 - 2 + ry, gx = x0 - 2 + 2 * jx;
        float2 t[3], r1[3];
        const bool rowin = gy >= 0 && gy < H;
        const bool in0 = rowin && gx >= 0 && gx < W, in1 = rowin && gx + 1 >= 0 && gx + 1 < W;
        if (w_even && rowin && gx >= 0 && gx + 3 < W && gy + 1 < H) {  // interior: 64-bit loads
#pragma unroll
            for (int ch = 0; ch < 3; ++ch) {
                const float* p = ub + ch * plane + (size_t)gy * W + gx;
                const float2 a = __ldg(reinterpret_cast<const float2*>(p));
                const float2 bb = __ldg(reinterpret_cast<const float2*>(p + W));
                const float a2 = __ldg(p + 2), b2 = __ldg(p + W + 2);
                const float mid = fmaxf(a.y, bb.y);
                float2 m = make_float2(fmaxf(fmaxf(a.x, bb.x), mid), fmaxf(mid, fmaxf(a2, b2)));
                if (!K0) {
                    const float2 yv = __ldg(reinterpret_cast<const float2*>(yb + ch * plane + (size_t)gy * W + gx));
                    m = make_float2(m.x - 0.5f * (a.x - yv.x), m.y - 0.5f * (a.y - yv.y));
                }
                t[ch] = m;
            }
        } else {
            auto at = [&](const float* base, int r, int cc) {
                return r >= 0 && r < H && cc >= 0 && cc < W ? __ldg(base + (size_t)r * W + cc) : 0.f;
            };
#pragma unroll
            for (int ch = 0; ch < 3; ++ch) {
                const float* pu = ub + ch * plane;
                const float a0 = at(pu, gy, gx), a1 = at(pu, gy, gx + 1), a2 = at(pu, gy, gx + 2);
                const float b0 = at(pu, gy + 1, gx), b1 = at(pu, gy + 1, gx + 1), b2 = at(pu, gy + 1, gx + 2);
                float2 m = make_float2(fmaxf(fmaxf(a0, a1), fmaxf(b0, b1)), fmaxf(fmaxf(a1, a2), fmaxf(b1, b2)));
                if (!K0)
                    m = make_float2(m.x - 0.5f * (a0 - at(yb + ch * plane, gy, gx)),
                                    m.y - 0.5f * (a1 - at(yb + ch * plane, gy, gx + 1)));
                t[ch] = m;
            }
        }
        conv1<3, 3>(c.c1r_w, c.c1r_b, t, r1);
#pragma unroll
        for (int ch = 0; ch < 3; ++ch) {
            const float2 v = lrelu2(add2(add2(r1[ch], t[ch]), t[ch]));
            *reinterpret_cast<float2*>(s_t + ch * P1 + ry * R1_W + 2 * jx) = t[ch];
            *reinterpret_cast<float2*>(s_r1 + ch * P1 + ry * R1_W + 2 * jx) = make_float2(in0 ? v.x : 0.f, in1 ? v.y : 0.f);
        }
    }
    __syncthreads();

    // phase 1: r_c2 = act(conv3x3(r_c1) + r_c1), r_c3 = act(2 r_c2) over the halo-1 region (R2_H x 17 strips)
    if (tid < R2_H * (R2_W / 2)) {
        const int ry = tid / (R2_W / 2), j = tid - ry * (R2_W / 2);
        const int gy = y0 - 1 + ry, gx = x0 - 1 + 2 * j;
        float2 acc[3];
#pragma unroll
        for (int co = 0; co < 3; ++co) acc[co] = bcast(c.c2r_b[co]);
        const float* s = s_r1 + (ry + 1) * R1_W + 2 * j;  // region column 2j = r_c1 column 2j+1; minus the tap offset
        conv3_strip2<3, 3, 1, P1, R1_W>(s, c.c2r_w, acc);
        const bool iny = gy >= 0 && gy < H, in0 = iny && gx >= 0 && gx < W, in1 = iny && gx + 1 >= 0 && gx + 1 < W;
#pragma unroll
        for (int co = 0; co < 3; ++co) {
            const float a0 = in0 ? lrelu(acc[co].x + s[co * P1 + 1]) : 0.f;
            const float a1 = in1 ? lrelu(acc[co].y + s[co * P1 + 2]) : 0.f;
            *reinterpret_cast<float2*>(s_r2 + co * P2 + ry * R2_W + 2 * j) = make_float2(a0, a1);
            *reinterpret_cast<float2*>(s_r3 + co * P2 + ry * R2_W + 2 * j) = make_float2(lrelu(a0 + a0), lrelu(a1 + a1));
        }
    }
    __syncthreads();

    // phase 2: the four distilled maps at the centre, c5, sigmoid, clamp, u = clamp(y / t)
    if (tid < TH * (R_TW / 2)) {
        const int ry = tid / (R_TW / 2), j = tid - ry * (R_TW / 2);
        const int gy = y0 + ry, gx = x0 + 2 * j;
        if (gy >= H || gx >= W) return;
        float2 d1[3], d2[3], d3[3], r4[3], o[3];
#pragma unroll
        for (int co = 0; co < 3; ++co) {
            d2[co] = bcast(c.c2d_b[co]);
            d3[co] = bcast(c.c3d_b[co]);
            r4[co] = bcast(c.c4_b[co]);
            o[co] = bcast(c.c5_b[co]);
        }
        const float* s1 = s_r1 + (ry + 2) * R1_W + 2 * j;  // centre column 2j = r_c1 column 2j+2; minus dilation 2
        conv3_strip2<3, 3, 2, P1, R1_W>(s1, c.c2d_w, d2);
        const int o2 = (ry + 1) * R2_W + 2 * j;  // centre column 2j = r_c2 column 2j+1; minus the tap offset
        conv3_strip2<3, 3, 1, P2, R2_W>(s_r2 + o2, c.c3d_w, d3);
        conv3_strip2<3, 3, 1, P2, R2_W>(s_r3 + o2, c.c4_w, r4);
#pragma unroll
        for (int ch = 0; ch < 3; ++ch) {
            d1[ch] = lrelu2(*reinterpret_cast<const float2*>(s_t + ch * P1 + (ry + 2) * R1_W + 2 * j + 2));
            d2[ch] = lrelu2(add2(d2[ch], *reinterpret_cast<const float2*>(s1 + ch * P1 + 2)));
            d3[ch] = lrelu2(d3[ch]);
            r4[ch] = lrelu2(r4[ch]);
        }
        c5_part<3, 0>(c.c5_w, d1, o);
        c5_part<3, 3>(c.c5_w, d2, o);
        c5_part<3, 6>(c.c5_w, d3, o);
        c5_part<3, 9>(c.c5_w, r4, o);
        const size_t base = (size_t)blockIdx.z * 3 * plane + (size_t)gy * W + gx;
#pragma unroll
        for (int ch = 0; ch < 3; ++ch) {
            const float ov[2] = {o[ch].x, o[ch].y};
#pragma unroll
            for (int q = 0; q < 2; ++q) {
                if (gx + q >= W) continue;
                float t = __fdividef(1.f, 1.f + __expf(-ov[q]));
                t = fminf(fmaxf(t, 0.001f), 1.f);
                const float yv = __ldg(y + base + ch * plane + q);
                const float uv = fminf(fmaxf(__fdividef(yv, t), 0.f), 1.f);
                if (t_out) t_out[base + ch * plane + q] = t;
                u_out[base + ch * plane + q] = uv;
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------ NRM cell
__global__ void __launch_bounds__(R_THREADS) ruas_nrm_kernel(const float* __restrict__ f, float* __restrict__ out,
                                                             const RuasNrmCell c, int H, int W) {
    constexpr int TH = NRM_TH, R1_H = TH + 4, P1 = R1_H * R1_W;
    __shared__ __align__(16) float s_r1[6 * P1];
    const int tid = threadIdx.x, x0 = blockIdx.x * R_TW, y0 = blockIdx.y * TH;
    const size_t plane = (size_t)H * W;
    const float* fb = f + (size_t)blockIdx.z * 6 * plane;

    // phase 0: r_c1 = act((conv1x1(f) + f) + f) over the halo-2 region (position pairs), zero outside the image
    const bool w_even = (W & 1) == 0;
    for (int i = tid; i < R1_H * (R1_W / 2); i += R_THREADS) {
        const int ry = i / (R1_W / 2), jx = i - ry * (R1_W / 2), gy = y0 - 2 + ry, gx = x0 - 2 + 2 * jx;
        const bool rowin = gy >= 0 && gy < H;
        const bool in0 = rowin && gx >= 0 && gx < W, in1 = rowin && gx + 1 >= 0 && gx + 1 < W;
        float2 v[6], o[6];
        const float* p = fb + (size_t)gy * W + gx;
        if (w_even && in0) {  // even W and even gx: both positions are inside, 64-bit loads
#pragma unroll
            for (int ch = 0; ch < 6; ++ch) v[ch] = __ldg(reinterpret_cast<const float2*>(p + ch * plane));
        } else {
#pragma unroll
            for (int ch = 0; ch < 6; ++ch)
                v[ch] = make_float2(in0 ? __ldg(p + ch * plane) : 0.f, in1 ? __ldg(p + ch * plane + 1) : 0.f);
        }
        conv1<6, 6>(c.c1r_w, c.c1r_b, v, o);
#pragma unroll
        for (int ch = 0; ch < 6; ++ch) {
            const float2 r = lrelu2(add2(add2(o[ch], v[ch]), v[ch]));
            *reinterpret_cast<float2*>(s_r1 + ch * P1 + ry * R1_W + 2 * jx) = make_float2(in0 ? r.x : 0.f, in1 ? r.y : 0.f);
        }
    }
    __syncthreads();

    {
        const int ry = tid / (R_TW / 2), j = tid - ry * (R_TW / 2);
        const int gy = y0 + ry, gx = x0 + 2 * j;
        if (gy >= H || gx >= W) return;
        const bool has1 = gx + 1 < W;
        float2 d2[6], v[6], r1[6], r2[6], t[6], o[6];
#pragma unroll
        for (int co = 0; co < 6; ++co) d2[co] = bcast(c.c2d_b[co]);
        const float* s1 = s_r1 + (ry + 2) * R1_W + 2 * j;
        conv3_strip2<6, 6, 2, P1, R1_W>(s1, c.c2d_w, d2);
        const size_t base = (size_t)gy * W + gx;
#pragma unroll
        for (int ch = 0; ch < 6; ++ch) {
            const float* p = fb + ch * plane + base;
            v[ch] = make_float2(__ldg(p), has1 ? __ldg(p + 1) : 0.f);
            r1[ch] = *reinterpret_cast<const float2*>(s1 + ch * P1 + 2);
            o[ch] = bcast(c.c5_b[ch]);
        }
        conv1<6, 6>(c.c1d_w, c.c1d_b, v, t);  // distilled_c1 = act(conv1x1(f) + f)
#pragma unroll
        for (int ch = 0; ch < 6; ++ch) t[ch] = lrelu2(add2(t[ch], v[ch]));
        c5_part<6, 0>(c.c5_w, t, o);
#pragma unroll
        for (int ch = 0; ch < 6; ++ch) {  // distilled_c2 = act(dilconv(r_c1) + r_c1);  r_c2 = act(r_c1 + r_c1)
            t[ch] = lrelu2(add2(d2[ch], r1[ch]));
            r2[ch] = lrelu2(add2(r1[ch], r1[ch]));
        }
        c5_part<6, 6>(c.c5_w, t, o);
        conv1<6, 6>(c.c3d_w, c.c3d_b, r2, t);  // distilled_c3 = act(conv1x1(r_c2) + r_c2)
#pragma unroll
        for (int ch = 0; ch < 6; ++ch) t[ch] = lrelu2(add2(t[ch], r2[ch]));
        c5_part<6, 12>(c.c5_w, t, o);
        conv1<6, 6>(c.c3r_w, c.c3r_b, r2, t);  // r_c3 = act((conv1x1(r_c2) + r_c2) + r_c2);  r_c4 = act(r_c3)
#pragma unroll
        for (int ch = 0; ch < 6; ++ch) t[ch] = lrelu2(lrelu2(add2(add2(t[ch], r2[ch]), r2[ch])));
        c5_part<6, 18>(c.c5_w, t, o);
        float* ob = out + (size_t)blockIdx.z * 6 * plane + base;
#pragma unroll
        for (int ch = 0; ch < 6; ++ch) {
            ob[ch * plane] = o[ch].x;
            if (has1) ob[ch * plane + 1] = o[ch].y;
        }
    }
}

// ------------------------------------------------------------------------------------------------ stem / tail
// 3x3 convolution CI -> CO with zero padding.  TAIL: n = conv(src); out = sub_from - n (model.py:147-148), n optional.
template <int CI, int CO, bool TAIL, typename WT>
__global__ void __launch_bounds__(R_THREADS) ruas_conv3_kernel(const float* __restrict__ src, float* __restrict__ out,
                                                               const float* __restrict__ sub_from,
                                                               float* __restrict__ n_out, const WT c, int H, int W) {
    constexpr int TH = NRM_TH, R2_H = TH + 2, P2 = R2_H * R2_W;
    __shared__ __align__(16) float s_in[CI * P2];
    const int tid = threadIdx.x, x0 = blockIdx.x * R_TW, y0 = blockIdx.y * TH;
    const size_t plane = (size_t)H * W;
    const float* sb = src + (size_t)blockIdx.z * CI * plane;
    for (int r = tid; r < P2; r += R_THREADS) {
        const int ry = r / R2_W, rx = r - ry * R2_W, gy = y0 - 1 + ry, gx = x0 - 1 + rx;
        const bool in = gy >= 0 && gy < H && gx >= 0 && gx < W;
        const float* p = sb + (size_t)gy * W + gx;
#pragma unroll
        for (int ch = 0; ch < CI; ++ch) s_in[ch * P2 + r] = in ? __ldg(p + ch * plane) : 0.f;
    }
    __syncthreads();
    {
        const int ry = tid / (R_TW / 2), j = tid - ry * (R_TW / 2);
        const int gy = y0 + ry, gx = x0 + 2 * j;
        if (gy >= H || gx >= W) return;
        float2 acc[CO];
#pragma unroll
        for (int co = 0; co < CO; ++co) acc[co] = bcast(c.b[co]);
        conv3_strip2<CI, CO, 1, P2, R2_W>(s_in + (ry + 1) * R2_W + 2 * j, c.w, acc);
        const size_t base = (size_t)blockIdx.z * CO * plane + (size_t)gy * W + gx;
#pragma unroll
        for (int co = 0; co < CO; ++co) {
            const float av[2] = {acc[co].x, acc[co].y};
#pragma unroll
            for (int q = 0; q < 2; ++q) {
                if (gx + q >= W) continue;
                const size_t o = base + co * plane + q;
                if (TAIL) {
                    if (n_out) n_out[o] = av[q];
                    out[o] = __ldg(sub_from + o) - av[q];
                } else {
                    out[o] = av[q];
                }
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
    const dim3 grid((W + R_TW - 1) / R_TW, (H + IEM_TH - 1) / IEM_TH, B);
    const dim3 grid_n((W + R_TW - 1) / R_TW, (H + NRM_TH - 1) / NRM_TH, B);
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
        ruas_conv3_kernel<3, 6, false, RuasStem><<<grid_n, R_THREADS, 0, s>>>(u3, fa, nullptr, nullptr, p->stem, H, W);
        ruas_nrm_kernel<<<grid_n, R_THREADS, 0, s>>>(fa, fb, p->nrm[0], H, W);
        ruas_nrm_kernel<<<grid_n, R_THREADS, 0, s>>>(fb, fa, p->nrm[1], H, W);
        ruas_nrm_kernel<<<grid_n, R_THREADS, 0, s>>>(fa, fb, p->nrm[2], H, W);
        ruas_conv3_kernel<6, 3, true, RuasTail>
            <<<grid_n, R_THREADS, 0, s>>>(fb, denoised_or_null, u3, noise_or_null, p->tail, H, W);
    }
    return launch_status();
}
