// Zero-DCE++ layers 2..7 (depthwise 3x3 -> pointwise 1x1, zero_dcepp.py:191-196 / dsconv.py:145-149) for sm_100a.
//
// The depthwise stage stays on the fp32 CUDA cores; the pointwise stage is a [pixels x Cin] . [Cin x Cout] GEMM
// that runs on warp-level tensor-core MMAs (mma.sync m16n8k16, bf16 operands, fp32 accumulate) with a 3-term
// split so the result keeps fp32-class accuracy (the path's contract is max-abs <= 1e-3 against the fp32 reference):
//     a = a_hi + a_lo,  w = w_hi + w_lo  (bf16 each)      a.w ~= a_hi.w_hi + a_lo.w_hi + a_hi.w_lo      (|err| ~ 2^-16)
// There is no operand exchange between the stages: every lane computes exactly the depthwise outputs that ARE its
// A fragment — for an m16 tile laid out as 2 image rows x 8 columns, lane (g = lane/4, c = lane%4) owns pixels
// (row 0/1, column g) x channels {2c, 2c+1, 2c+8, 2c+9}; two stacked m-tiles give it 4 vertically adjacent pixels,
// i.e. a 6-row sliding window per channel.  The work is small (B*h*w low-res pixels) and L2 resident; this kernel
// exists so that the HBM-bound full-resolution pass dominates the step.
//
// CTA = 256 threads = 8 warps, tile = 8 rows x 32 columns; warp w covers rows 4*(w/4).., columns 8*(w%4)..
// Shared memory: halo tile [32 ch][10 rows][40] fp32 (+4 floats per channel: with a channel stride of 404 the (g, c)
// access pattern is bank-conflict free, and the 32 interior columns of every row start 16-byte aligned so the tile is
// filled with 16-byte cp.async -- all of a source's loads are in flight at once),
// depthwise weights [Cin][12] (9 taps, bias, pad), pointwise bias [32], pointwise B fragments (bf16 hi/lo, one
// 16-byte vector per lane per (k16, n8) tile).
#pragma once
#include <cuda_bf16.h>

#include "common.cuh"
#include "tcgen05.cuh"

namespace b200 {

constexpr int MM_TH = 8, MM_TW = 32, MM_P = 40, MM_CH = 32;  // tile rows / cols, halo row pitch, channels per source
constexpr int MM_S = 10 * MM_P + 4;                          // channel stride (floats)
constexpr int MM_C0 = 3;                                     // halo column of image column x0 - 1
constexpr int MM_HALO_FLOATS = MM_CH * MM_S;

// per-layer packed block (32-bit words): dw [Cin][12] | pw bias [32] | frags [KT][NT][32 lanes][4]
__host__ __device__ constexpr int mm_cin(int l) { return l <= 3 ? 32 : 64; }  // l = 1..6 (e_conv2..7)
__host__ __device__ constexpr int mm_nt(int l) { return l == 6 ? 1 : 4; }
__host__ __device__ constexpr int mm_words(int l) { return mm_cin(l) * 12 + 32 + (mm_cin(l) / 16) * mm_nt(l) * 128; }
__host__ __device__ constexpr int mm_offset(int l) {
    int o = 0;
    for (int i = 1; i < l; ++i) o += mm_words(i);
    return o;
}
constexpr int MM_TOTAL_WORDS = mm_offset(7);

__device__ __forceinline__ uint32_t mm_pack_bf16x2(float lo_elem, float hi_elem) {
    __nv_bfloat162 h = __floats2bfloat162_rn(lo_elem, hi_elem);  // .x (low 16 bits) = lo_elem
    return *reinterpret_cast<uint32_t*>(&h);
}
// split two fp32 values into packed bf16 hi and lo parts
__device__ __forceinline__ void mm_split2(float v0, float v1, uint32_t& hi, uint32_t& lo) {
    hi = mm_pack_bf16x2(v0, v1);
    const float h0 = __uint_as_float(hi << 16), h1 = __uint_as_float(hi & 0xffff0000u);
    lo = mm_pack_bf16x2(v0 - h0, v1 - h1);
}

// one layer's block; called by dcepp_pack_kernel (block l = 1..6)
__device__ void mm_pack_layer(int l, const float* dw_w, const float* dw_b, const float* pw_w, const float* pw_b,
                              uint32_t* out) {
    const int cin = mm_cin(l), nt_n = mm_nt(l), cout = l == 6 ? 3 : 32;
    float* f = reinterpret_cast<float*>(out);
    for (int i = threadIdx.x; i < cin * 12; i += blockDim.x) {
        const int ch = i / 12, j = i - ch * 12;
        f[i] = j < 9 ? dw_w[ch * 9 + j] : (j == 9 ? dw_b[ch] : 0.f);
    }
    for (int i = threadIdx.x; i < 32; i += blockDim.x) f[cin * 12 + i] = i < cout ? pw_b[i] : 0.f;
    uint32_t* fr = out + cin * 12 + 32;
    const int n_fr = (cin / 16) * nt_n * 128;
    for (int i = threadIdx.x; i < n_fr; i += blockDim.x) {
        const int j = i & 3, lane = (i >> 2) & 31, tile = i >> 7;
        const int kt = tile / nt_n, nt = tile - kt * nt_n;
        const int g = lane >> 2, c = lane & 3;
        const int n = nt * 8 + g;
        const int k0 = kt * 16 + 2 * c + ((j & 1) ? 8 : 0);
        float w0 = 0.f, w1 = 0.f;
        if (n < cout) {
            w0 = pw_w[n * cin + k0];
            w1 = pw_w[n * cin + k0 + 1];
        }
        uint32_t hi, lo;
        mm_split2(w0, w1, hi, lo);
        fr[i] = j < 2 ? hi : lo;
    }
}

__device__ __forceinline__ void mm_mma_bf16(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

template <int NSRC, int NT>
constexpr int mm_smem_bytes() {
    return 4 * (MM_HALO_FLOATS + NSRC * 32 * 12 + 32 + NSRC * 2 * NT * 128);
}

// One 8 x 32 output tile of one layer.  srcA / srcB: [B,32,h,w] fp32 planes (torch.cat never materialised);
// dst: [B,COUT,h,w]; wl: this layer's packed block; mm_smem: mm_smem_bytes<NSRC, NT>() of shared memory.
// Every thread of the 256-thread CTA must call it (it synchronises the CTA).
template <int NSRC, int NT, bool TANH>
__device__ __forceinline__ void mm_tile(const float* __restrict__ srcA, const float* __restrict__ srcB,
                                        float* __restrict__ dst, const uint32_t* __restrict__ wl, int h, int w, int b,
                                        int y0, int x0, uint32_t* mm_smem) {
    constexpr int CIN = 32 * NSRC, KT = CIN / 16, COUT = TANH ? 3 : 32;
    float* s_halo = reinterpret_cast<float*>(mm_smem);
    float* s_dw = s_halo + MM_HALO_FLOATS;            // [CIN][12]
    float* s_pb = s_dw + CIN * 12;                    // [32]
    const uint4* s_fr = reinterpret_cast<const uint4*>(s_pb + 32);  // [KT][NT][32]

    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int g = lane >> 2, c = lane & 3;
    const int wy = (warp >> 2) * 4, wx = (warp & 3) * 8;

    {  // weights -> smem (same layout as the packed block)
        constexpr int NW = CIN * 12 + 32 + KT * NT * 128;
        uint32_t* s = reinterpret_cast<uint32_t*>(s_dw);
        for (int i = t; i < NW / 4; i += 256)
            reinterpret_cast<uint4*>(s)[i] = __ldg(reinterpret_cast<const uint4*>(wl) + i);
    }

    float acc[2][NT][4];
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
        for (int nt = 0; nt < NT; ++nt)
#pragma unroll
            for (int i = 0; i < 4; ++i) acc[mt][nt][i] = 0.f;

    const size_t plane = (size_t)h * w;
#pragma unroll 1
    for (int s = 0; s < NSRC; ++s) {
        const float* src = (s == 0 ? srcA : srcB) + (size_t)b * 32 * plane;
        __syncthreads();  // previous source's halo consumers are done (and weights are visible on the first pass)
        // ---- halo tile: 32 channels x 10 rows x (1 + 32 + 1) columns, zero padded, via cp.async ----
        {
            const uint32_t sh = tc::smem_u32(s_halo);
            if ((w & 3) == 0) {
                // interior: thread = (channel t/8, 16-byte chunk t%8), one chunk per halo row: no index division
                const int ch = t >> 3, j = t & 7;
                const int xa = x0 + 4 * j;
                const bool xok = xa < w;
                const float* gp = src + (size_t)ch * plane + xa;
                const uint32_t sp = sh + (ch * MM_S + MM_C0 + 1 + 4 * j) * 4;
#pragma unroll
                for (int rr = 0; rr < 10; ++rr) {
                    const int yy = y0 - 1 + rr;
                    const bool ok = xok && yy >= 0 && yy < h;
                    tc::cp_async16(sp + rr * MM_P * 4, ok ? gp + (ptrdiff_t)yy * w : src, ok);
                }
                // the two edge columns x0 - 1 and x0 + 32: thread = (channel, 8 of the 20 (row, side) slots)
                for (int e = j; e < 20; e += 8) {
                    const int rr = e >> 1, side = e & 1;
                    const int yy = y0 - 1 + rr, xe = side ? x0 + 32 : x0 - 1;
                    const bool ok = yy >= 0 && yy < h && xe >= 0 && xe < w;
                    tc::cp_async4(sh + (ch * MM_S + rr * MM_P + MM_C0 + (side ? 33 : 0)) * 4,
                                  ok ? src + (size_t)ch * plane + (size_t)yy * w + xe : src, ok);
                }
            } else {
                for (int i = t; i < 320 * 34; i += 256) {
                    const int p = i / 34, col = i - p * 34;
                    const int ch = p / 10, rr = p - ch * 10;
                    const int yy = y0 - 1 + rr, xa = x0 - 1 + col;
                    const bool ok = yy >= 0 && yy < h && xa >= 0 && xa < w;
                    const float* gp = src + (ok ? ((size_t)ch * h + yy) * w + xa : (size_t)0);
                    tc::cp_async4(sh + (ch * MM_S + rr * MM_P + MM_C0 + col) * 4, gp, ok);
                }
            }
            tc::cp_async_commit();
            tc::cp_async_wait<0>();
        }
        __syncthreads();
#pragma unroll
        for (int kt2 = 0; kt2 < 2; ++kt2) {  // two k16 blocks per 32-channel source
            const int kt = s * 2 + kt2;
            // ---- depthwise 3x3 for this lane's fragment entries: 4 channels x 4 vertically adjacent pixels ----
            float val[4][4];  // [channel index j][output row o]
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int chl = kt2 * 16 + 2 * c + (j & 1) + (j >> 1) * 8;  // channel within the source
                const float4 w0 = *reinterpret_cast<const float4*>(s_dw + (s * 32 + chl) * 12);
                const float4 w1 = *reinterpret_cast<const float4*>(s_dw + (s * 32 + chl) * 12 + 4);
                const float4 w2 = *reinterpret_cast<const float4*>(s_dw + (s * 32 + chl) * 12 + 8);
                const float wk[9] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w, w2.x};
                float o[4] = {w2.y, w2.y, w2.y, w2.y};  // bias
                const float* hp = s_halo + chl * MM_S + wy * MM_P + MM_C0 + wx + g;
#pragma unroll
                for (int r = 0; r < 6; ++r) {
                    const float v0 = hp[r * MM_P], v1 = hp[r * MM_P + 1], v2 = hp[r * MM_P + 2];
#pragma unroll
                    for (int dy = 0; dy < 3; ++dy) {
                        const int oo = r - dy;
                        if (oo >= 0 && oo < 4) {
                            o[oo] = fmaf(wk[dy * 3 + 0], v0, o[oo]);
                            o[oo] = fmaf(wk[dy * 3 + 1], v1, o[oo]);
                            o[oo] = fmaf(wk[dy * 3 + 2], v2, o[oo]);
                        }
                    }
                }
#pragma unroll
                for (int oo = 0; oo < 4; ++oo) val[j][oo] = o[oo];
            }
            // ---- A fragments (hi / lo) of the two m-tiles, then 3 MMAs per n-tile ----
#pragma unroll
            for (int mt = 0; mt < 2; ++mt) {
                uint32_t ah[4], al[4];
                mm_split2(val[0][2 * mt], val[1][2 * mt], ah[0], al[0]);          // row g    , k = 2c, 2c+1
                mm_split2(val[0][2 * mt + 1], val[1][2 * mt + 1], ah[1], al[1]);  // row g + 8, k = 2c, 2c+1
                mm_split2(val[2][2 * mt], val[3][2 * mt], ah[2], al[2]);          // row g    , k = 2c+8, 2c+9
                mm_split2(val[2][2 * mt + 1], val[3][2 * mt + 1], ah[3], al[3]);  // row g + 8, k = 2c+8, 2c+9
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) {
                    const uint4 bf = s_fr[(kt * NT + nt) * 32 + lane];  // {hi b0, hi b1, lo b0, lo b1}
                    mm_mma_bf16(acc[mt][nt], al, bf.x, bf.y);
                    mm_mma_bf16(acc[mt][nt], ah, bf.z, bf.w);
                    mm_mma_bf16(acc[mt][nt], ah, bf.x, bf.y);
                }
            }
        }
    }

    // ---- epilogue: bias + activation, NCHW stores.  acc[mt][nt] = {(row g, n 2c), (g, 2c+1), (g+8, 2c), (g+8, 2c+1)}
    // one base pointer per thread, then constant multiples of the plane / row pitch
    const int xx = x0 + wx + g;
    if (xx < w) {
        float* dp = dst + ((size_t)b * COUT + 2 * c) * plane + (size_t)(y0 + wy) * w + xx;
        const int rows = h - (y0 + wy);  // valid rows of this warp's 4
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int oc = nt * 8 + 2 * c + e;
                if (COUT < 32 && oc >= COUT) continue;
                const float pb = s_pb[oc];
                float* q = dp + (size_t)(nt * 8 + e) * plane;
#pragma unroll
                for (int r = 0; r < 4; ++r) {  // image row y0 + wy + r = m-tile r / 2, half r % 2
                    const float u = acc[r >> 1][nt][(r & 1) * 2 + e] + pb;
                    if (r < rows) q[(size_t)r * w] = TANH ? tanhf(u) : fmaxf(u, 0.f);
                }
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// Layers 2..7 as ONE persistent launch.  The 6 x (B x tiles) output tiles form a single layer-major work
// list; CTAs draw tile numbers from a global counter and a tile of layer l starts as soon as the (up to)
// nine layer l-1 tiles under its halo have been published -- no launch boundary, no wave tail between
// layers, and the last tiles of layer l overlap the first tiles of layer l+1.  Deadlock-free: a CTA
// only ever waits on tiles with smaller numbers, and those were drawn earlier by CTAs that are resident
// (grid <= one resident wave) and wait only on still smaller numbers.
//
// sync words (caller's workspace, zeroed by layer 1's launch): [0] = work counter, [MM_FLAG0 + q] = tile q
// of layers 2..6 published (layer 7 has no consumer in this launch).
// ------------------------------------------------------------------------------------------------
constexpr int MM_FLAG0 = 32;  // flags start one 128-byte line after the counter

struct MmNetArgs {
    const float* srcA[6];
    const float* srcB[6];
    float* dst[6];
    const uint32_t* wl[6];
};

__host__ __device__ inline size_t mm_sync_words(int B, int h, int w) {
    const size_t tiles = (size_t)B * ((h + MM_TH - 1) / MM_TH) * ((w + MM_TW - 1) / MM_TW);
    return MM_FLAG0 + 5 * tiles;
}

__device__ __forceinline__ int mm_ld_acquire(const int* p) {
    int v;
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void mm_st_release(int* p, int v) {
    asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

constexpr int MM_NET_SMEM = mm_smem_bytes<2, 4>();

__global__ void __launch_bounds__(256, 3)
dcepp_net_kernel(MmNetArgs a, int B, int h, int w, int* __restrict__ sync) {
    extern __shared__ __align__(16) uint32_t mm_smem[];
    __shared__ int s_q;
    const int t = threadIdx.x;
    const int ntx = (w + MM_TW - 1) / MM_TW, nty = (h + MM_TH - 1) / MM_TH;
    const int tpf = ntx * nty, tpl = B * tpf, total = 6 * tpl;
    int* flags = sync + MM_FLAG0;
    for (;;) {
        __syncthreads();  // every thread has read s_q / finished with the previous tile's smem
        if (t == 0) s_q = atomicAdd(sync, 1);
        __syncthreads();
        const int q = s_q;
        if (q >= total) break;
        const int l = q / tpl, r = q - l * tpl;
        const int b = r / tpf, r2 = r - b * tpf;
        const int tyi = r2 / ntx, txi = r2 - tyi * ntx;
        if (l > 0) {
            if (t < 9) {  // the nine layer l-1 tiles under this tile's halo
                const int ny = tyi + t / 3 - 1, nx = txi + t % 3 - 1;
                if (ny >= 0 && ny < nty && nx >= 0 && nx < ntx) {
                    const int* f = flags + (size_t)(l - 1) * tpl + b * tpf + ny * ntx + nx;
                    while (mm_ld_acquire(f) == 0) __nanosleep(40);
                }
            }
            __syncthreads();
        }
        const int y0 = tyi * MM_TH, x0 = txi * MM_TW;
        const float *sa = nullptr, *sb = nullptr;  // selected with constant indices: the kernel parameters stay in
        float* ds = nullptr;                       // the constant bank (a dynamic index would spill them to the stack)
        const uint32_t* wl = nullptr;
#pragma unroll
        for (int i = 0; i < 6; ++i)
            if (i == l) sa = a.srcA[i], sb = a.srcB[i], ds = a.dst[i], wl = a.wl[i];
        if (l < 3)
            mm_tile<1, 4, false>(sa, nullptr, ds, wl, h, w, b, y0, x0, mm_smem);
        else if (l < 5)
            mm_tile<2, 4, false>(sa, sb, ds, wl, h, w, b, y0, x0, mm_smem);
        else
            mm_tile<2, 1, true>(sa, sb, ds, wl, h, w, b, y0, x0, mm_smem);
        if (l < 5) {
            __syncthreads();  // all of this tile's stores are issued
            if (t == 0) mm_st_release(flags + q, 1);
        }
    }
}

static int launch_mm_net(const MmNetArgs& a, int B, int h, int w, int* sync, cudaStream_t s) {
    auto kern = dcepp_net_kernel;
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, MM_NET_SMEM) != cudaSuccess) {
        (void)cudaGetLastError();
        return B200_ERR_LAUNCH;
    }
    int occ = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, 256, MM_NET_SMEM) != cudaSuccess || occ < 1) {
        (void)cudaGetLastError();
        return B200_ERR_LAUNCH;  // the wait-on-earlier-tiles scheme needs every CTA resident
    }
    int dev = 0, sms = B200_NUM_SMS;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const long long total = 6LL * B * ((h + MM_TH - 1) / MM_TH) * ((w + MM_TW - 1) / MM_TW);
    long long grid = (long long)sms * occ;
    if (grid > total) grid = total;
    kern<<<(unsigned)grid, 256, MM_NET_SMEM, s>>>(a, B, h, w, sync);
    return launch_status();
}

}  // namespace b200
