// Zero-DCE++ layers 2..7 (depthwise 3x3 -> pointwise 1x1, zero_dcepp.py:191-196 / dsconv.py:145-149) for sm_100a,
// as ONE persistent, TMA-fed launch.
//
// Arithmetic.  The depthwise stage stays on the fp32 CUDA cores, two channels per instruction (fma.rn.f32x2); the
// pointwise stage is a [pixels x Cin] . [Cin x Cout] GEMM on warp-level tensor-core MMAs (mma.sync m16n8k16, bf16
// operands, fp32 accumulate) with a 3-term split so the result keeps fp32-class accuracy (the path's contract is
// max-abs <= 1e-3 against the fp32 reference):
//     a = a_hi + a_lo,  w = w_hi + w_lo  (bf16 each)      a.w ~= a_hi.w_hi + a_lo.w_hi + a_hi.w_lo      (|err| ~ 2^-16)
// There is no operand exchange between the stages: every lane computes exactly the depthwise outputs that ARE its
// A fragment -- for an m16 tile laid out as 2 image rows x 8 columns, lane (g = lane/4, c = lane%4) owns pixels
// (row 0/1, column g) x channel pairs {c, c+4} of a 16-channel block; two stacked m-tiles give it 4 vertically
// adjacent pixels, i.e. a 6-row sliding window per channel pair.
//
// Layout.  The feature maps x1..x6 are private to the path (caller's workspace), so they are stored the way both
// stages want them: CHANNEL-PAIR INTERLEAVED, [B][16 pairs][h][wp] float2 (wp = w rounded up to even, so rows are
// 16-byte multiples).  A float2 is (channel 2p, channel 2p+1) of one pixel = one FFMA2 operand = one A-fragment
// register pair before the split = one C-fragment register pair after the MMA (stored with one 64-bit store).
//
// Data movement.  A 3-D tensor map per feature map (8-byte elements: x, y, pair + 16*b) lets one elected lane fetch a
// stage = 8 pairs x 10 rows x 38 columns (a tile's halo for one k16 block, 24.3 KB) with a single
// cp.async.bulk.tensor; coordinates outside the map (x < 0, x >= w, ...) are zero-filled = the conv's zero padding,
// so there is no per-thread halo addressing at all.  With 380 float2 per pair plane the (g, c) access pattern of the
// 64-bit shared loads is bank-conflict free.
//
// Schedule.  The 6 x (B x tiles) output tiles form a single layer-major work list.  Each CTA has a producer warp
// that draws tile numbers from a global counter, waits until the (up to) nine layer l-1 tiles under the tile's halo
// are published (acquire), and streams the tile's 2 or 4 stages through a 4-slot ring (full / empty mbarriers) --
// running up to a whole tile ahead of the 8 consumer warps, so tile-fetch, dependency and load latencies are off the
// critical path.  There is no launch boundary or wave tail between layers.  Deadlock-free: a producer only waits on
// tiles with smaller numbers, drawn earlier by CTAs that are resident (grid <= one resident wave) and whose consumers
// never wait on anything but their own producer.
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <string.h>

#include "common.cuh"
#include "tcgen05.cuh"
#include "tma_host.cuh"

namespace b200 {

constexpr int MM_TH = 8, MM_TW = 32;                 // output tile
// halo rows / box columns.  TMA wants the box's first byte 16-byte aligned in the innermost dimension (measured:
// tools/exp/tma_cfg_probe.cu -- an odd x start with 8-byte elements raises an illegal-instruction fault), so the box
// starts at x0 - 2 instead of x0 - 1; 38 columns (2 + 32 + 1, +3 unused) keep the pair-plane stride = 4 mod 16.
constexpr int MM_HR = MM_TH + 2, MM_HC = 38;
constexpr int MM_XOFF = 2;                           // image column of box column 0 = x0 - MM_XOFF
constexpr int MM_PS = MM_HR * MM_HC;                 // float2 per pair plane of a stage
constexpr int MM_STAGE_PAIRS = 8;                    // one k16 block
constexpr int MM_STAGE_BYTES = MM_STAGE_PAIRS * MM_PS * 8;
constexpr int MM_NS = 4;          // ring slots
constexpr int MM_THREADS = 288;   // 8 consumer warps + the producer warp
static_assert(MM_STAGE_BYTES % 128 == 0, "TMA destinations stay 128-byte aligned");
static_assert(MM_PS % 16 == 4 || MM_PS % 16 == 12, "pair-plane stride keeps the (g, c) 64-bit loads conflict free");

// per-layer packed block (32-bit words): dw [Cin/2][20] | pw bias [32] | frags [KT][NT][32 lanes][4]
//   dw row of pair p: 9 x (tap of channel 2p, tap of channel 2p+1), then (bias 2p, bias 2p+1)
__host__ __device__ constexpr int mm_cin(int l) { return l <= 3 ? 32 : 64; }  // l = 1..6 (e_conv2..7)
__host__ __device__ constexpr int mm_nt(int l) { return l == 6 ? 1 : 4; }
__host__ __device__ constexpr int mm_words(int l) { return mm_cin(l) * 10 + 32 + (mm_cin(l) / 16) * mm_nt(l) * 128; }
__host__ __device__ constexpr int mm_offset(int l) {
    int o = 0;
    for (int i = 1; i < l; ++i) o += mm_words(i);
    return o;
}
constexpr int MM_TOTAL_WORDS = mm_offset(7);
constexpr int MM_W_WORDS_MAX = mm_words(4);
constexpr int MM_NET_SMEM = MM_NS * MM_STAGE_BYTES + MM_W_WORDS_MAX * 4 + 2 * MM_NS * 8 + 4 * 16 + 16;

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
    for (int i = threadIdx.x; i < cin * 10; i += blockDim.x) {
        const int p = i / 20, j = i - p * 20;
        const int tap = j >> 1, ch = 2 * p + (j & 1);
        f[i] = tap < 9 ? dw_w[ch * 9 + tap] : dw_b[ch];
    }
    for (int i = threadIdx.x; i < 32; i += blockDim.x) f[cin * 10 + i] = i < cout ? pw_b[i] : 0.f;
    uint32_t* fr = out + cin * 10 + 32;
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

__device__ __forceinline__ int mm_ld_acquire(const int* p) {
    int v;
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void mm_st_release(int* p, int v) {
    asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// sync words (caller's workspace, zeroed by layer 1's launch): [0] = work counter, [MM_FLAG0 + q] = tile q of
// layers 2..6 published (layer 7 has no consumer in this launch)
constexpr int MM_FLAG0 = 32;  // flags start one 128-byte line after the counter
__host__ __device__ inline size_t mm_sync_words(int B, int h, int w) {
    const size_t tiles = (size_t)B * ((h + MM_TH - 1) / MM_TH) * ((w + MM_TW - 1) / MM_TW);
    return MM_FLAG0 + 5 * tiles;
}
__host__ __device__ inline int mm_pitch(int w) { return (w + 1) & ~1; }  // float2 per feature-map row

struct MmNetArgs {
    float* dst[6];          // x2 x3 x4 x5 x6 (pair interleaved) | a_lr (planar [B,3,h,w])
    const uint32_t* wl[6];  // packed blocks of e_conv2..7
};

// One 8 x 32 output tile, executed by the 8 consumer warps.  `ck` = this CTA's running stage counter.
template <int NSRC, int NT, bool TANH>
__device__ __forceinline__ void mm_tile(const uint8_t* ring, uint64_t* full, uint64_t* empty, int& ck,
                                        const float* s_w, float* __restrict__ dst, int h, int w, int wp, int b, int y0,
                                        int x0) {
    constexpr int CIN = 32 * NSRC, KT = CIN / 16;
    const float* s_dw = s_w;                                        // [CIN/2][20]
    const float* s_pb = s_w + CIN * 10;                             // [32]
    const uint4* s_fr = reinterpret_cast<const uint4*>(s_pb + 32);  // [KT][NT][32]

    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int g = lane >> 2, c = lane & 3;
    const int wy = (warp >> 2) * 4, wx = (warp & 3) * 8;

    // accumulators start at the pointwise bias: acc[mt][nt] = {(row g, n 2c), (g, 2c+1), (g+8, 2c), (g+8, 2c+1)}
    float acc[2][NT][4];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
        const float2 pb = *reinterpret_cast<const float2*>(s_pb + nt * 8 + 2 * c);
#pragma unroll
        for (int mt = 0; mt < 2; ++mt) acc[mt][nt][0] = acc[mt][nt][2] = pb.x, acc[mt][nt][1] = acc[mt][nt][3] = pb.y;
    }

#pragma unroll 1
    for (int kt = 0; kt < KT; ++kt) {
        const int slot = ck % MM_NS;
        tc::mbar_wait(&full[slot], (ck / MM_NS) & 1);
        const float2* st = reinterpret_cast<const float2*>(ring + slot * MM_STAGE_BYTES);
        // ---- depthwise 3x3 for this lane's fragment entries: 2 channel pairs x 4 vertically adjacent pixels ----
        float2 val[2][4];  // [pair c / c+4][output row]
#pragma unroll
        for (int j2 = 0; j2 < 2; ++j2) {
            const int p = j2 * 4 + c;
            const float4* wq = reinterpret_cast<const float4*>(s_dw + (kt * 8 + p) * 20);
            const float4 q0 = wq[0], q1 = wq[1], q2 = wq[2], q3 = wq[3], q4 = wq[4];
            const float2 wk[9] = {make_float2(q0.x, q0.y), make_float2(q0.z, q0.w), make_float2(q1.x, q1.y),
                                  make_float2(q1.z, q1.w), make_float2(q2.x, q2.y), make_float2(q2.z, q2.w),
                                  make_float2(q3.x, q3.y), make_float2(q3.z, q3.w), make_float2(q4.x, q4.y)};
            // three independent partial sums per output (one per filter column): 12 short dependency chains instead
            // of 4 long ones keep the packed-FMA pipe fed at 4-5 warps per scheduler
            float2 o[4][3];
            const float2* hp = st + p * MM_PS + wy * MM_HC + wx + g + (MM_XOFF - 1);
#pragma unroll
            for (int r = 0; r < 6; ++r) {
                const float2 v[3] = {hp[r * MM_HC], hp[r * MM_HC + 1], hp[r * MM_HC + 2]};
#pragma unroll
                for (int dy = 0; dy < 3; ++dy) {
                    const int oo = r - dy;
                    if (oo >= 0 && oo < 4) {
#pragma unroll
                        for (int dx = 0; dx < 3; ++dx) {
                            if (dy == 0)
                                o[oo][dx] = dx == 0 ? __ffma2_rn(wk[dx], v[dx], make_float2(q4.z, q4.w))  // + bias
                                                    : __fmul2_rn(wk[dx], v[dx]);
                            else
                                o[oo][dx] = __ffma2_rn(wk[dy * 3 + dx], v[dx], o[oo][dx]);
                        }
                    }
                }
            }
#pragma unroll
            for (int oo = 0; oo < 4; ++oo) val[j2][oo] = __fadd2_rn(__fadd2_rn(o[oo][0], o[oo][1]), o[oo][2]);
        }
        __syncwarp();  // every lane's shared loads of this stage have returned
        if (lane == 0) tc::mbar_arrive(&empty[slot]);
        ++ck;
        // ---- A fragments (hi / lo) of the two m-tiles, then 3 MMAs per n-tile ----
#pragma unroll
        for (int mt = 0; mt < 2; ++mt) {
            uint32_t ah[4], al[4];
            mm_split2(val[0][2 * mt].x, val[0][2 * mt].y, ah[0], al[0]);          // row g    , k = 2c, 2c+1
            mm_split2(val[0][2 * mt + 1].x, val[0][2 * mt + 1].y, ah[1], al[1]);  // row g + 8, k = 2c, 2c+1
            mm_split2(val[1][2 * mt].x, val[1][2 * mt].y, ah[2], al[2]);          // row g    , k = 2c+8, 2c+9
            mm_split2(val[1][2 * mt + 1].x, val[1][2 * mt + 1].y, ah[3], al[3]);  // row g + 8, k = 2c+8, 2c+9
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) {
                const uint4 bf = s_fr[(kt * NT + nt) * 32 + lane];  // {hi b0, hi b1, lo b0, lo b1}
                mm_mma_bf16(acc[mt][nt], al, bf.x, bf.y);
                mm_mma_bf16(acc[mt][nt], ah, bf.z, bf.w);
                mm_mma_bf16(acc[mt][nt], ah, bf.x, bf.y);
            }
        }
    }

    // ---- epilogue: activation + stores; image row y0 + wy + r = m-tile r / 2, half r % 2 ----
    const int xx = x0 + wx + g;
    const int rows = h - (y0 + wy);  // valid rows of this warp's 4
    if (xx < w) {
        if (!TANH) {  // pair plane nt*4 + c, one 64-bit store per (pixel, pair)
            const size_t plane = (size_t)h * wp;
            float2* dp = reinterpret_cast<float2*>(dst) + ((size_t)b * 16 + c) * plane + (size_t)(y0 + wy) * wp + xx;
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) {
                float2* q = dp + (size_t)(nt * 4) * plane;
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    const float2 u = make_float2(fmaxf(acc[r >> 1][nt][(r & 1) * 2], 0.f),
                                                 fmaxf(acc[r >> 1][nt][(r & 1) * 2 + 1], 0.f));
                    if (r < rows) q[(size_t)r * wp] = u;
                }
            }
        } else {  // x_r: planar [B,3,h,w], channels 2c and 2c+1
            const size_t plane = (size_t)h * w;
            float* dp = dst + ((size_t)b * 3 + 2 * c) * plane + (size_t)(y0 + wy) * w + xx;
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                if (2 * c + e >= 3) continue;
#pragma unroll
                for (int r = 0; r < 4; ++r)
                    if (r < rows) dp[(size_t)e * plane + (size_t)r * w] = tanhf(acc[r >> 1][0][(r & 1) * 2 + e]);
            }
        }
    }
}

__global__ void __launch_bounds__(MM_THREADS, 2)
dcepp_net_kernel(const MmNetArgs a, const __grid_constant__ CUtensorMap tm1, const __grid_constant__ CUtensorMap tm2,
                 const __grid_constant__ CUtensorMap tm3, const __grid_constant__ CUtensorMap tm4,
                 const __grid_constant__ CUtensorMap tm5, int B, int h, int w, int* __restrict__ sync) {
    extern __shared__ __align__(128) uint8_t mm_smem[];
    uint8_t* ring = mm_smem;
    float* s_w = reinterpret_cast<float*>(mm_smem + MM_NS * MM_STAGE_BYTES);
    uint64_t* full = reinterpret_cast<uint64_t*>(s_w + MM_W_WORDS_MAX);
    uint64_t* empty = full + MM_NS;
    // tile descriptors, producer -> consumers: {tile number q (-1 = end), layer + 8 * frame, y0, x0}
    volatile int4* s_tq = reinterpret_cast<volatile int4*>(empty + MM_NS);

    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int ntx = (w + MM_TW - 1) / MM_TW, nty = (h + MM_TH - 1) / MM_TH;
    const int tpf = ntx * nty, tpl = B * tpf, total = 6 * tpl;
    const int wp = mm_pitch(w);
    int* flags = sync + MM_FLAG0;

    if (t == 0) {
        for (int i = 0; i < MM_NS; ++i) {
            tc::mbar_init(&full[i], 1);   // the producer's expect_tx arrive (+ the stage's bytes)
            tc::mbar_init(&empty[i], 8);  // one arrive per consumer warp
        }
        tc::mbar_fence_init();
    }
    __syncthreads();

    if (warp == 8) {
        // ================= producer warp =================
        if (lane == 0) {
            tc::tma_prefetch_desc(&tm1);
            tc::tma_prefetch_desc(&tm2);
            tc::tma_prefetch_desc(&tm3);
            tc::tma_prefetch_desc(&tm4);
            tc::tma_prefetch_desc(&tm5);
        }
        int pk = 0, tk = 0;
        for (;;) {
            int q = 0;
            if (lane == 0) q = atomicAdd(sync, 1);
            q = __shfl_sync(0xffffffffu, q, 0);
            if (q >= total) {  // end marker: an empty stage whose tile number is -1
                const int slot = pk % MM_NS;
                if (pk >= MM_NS) tc::mbar_wait(&empty[slot], ((pk / MM_NS) - 1) & 1);
                if (lane == 0) {
                    s_tq[tk & 3].x = -1;
                    tc::mbar_arrive(&full[slot]);
                }
                break;
            }
            const int l = q / tpl, r = q - l * tpl;
            const int b = r / tpf, r2 = r - b * tpf;
            const int tyi = r2 / ntx, txi = r2 - tyi * ntx;
            if (l > 0) {
                if (lane < 9) {  // the nine layer l-1 tiles under this tile's halo
                    const int ny = tyi + lane / 3 - 1, nx = txi + lane % 3 - 1;
                    if (ny >= 0 && ny < nty && nx >= 0 && nx < ntx) {
                        const int* f = flags + (size_t)(l - 1) * tpl + b * tpf + ny * ntx + nx;
                        while (mm_ld_acquire(f) == 0) __nanosleep(32);
                    }
                }
                __syncwarp();
            }
            // sources (zero_dcepp.py:191-196): x1 | x2 | x3 | cat(x3,x4) | cat(x2,x5) | cat(x1,x6 = x4's buffer)
            const CUtensorMap* mA = l == 0 || l == 5 ? &tm1 : (l == 1 || l == 4 ? &tm2 : &tm3);
            const CUtensorMap* mB = l == 4 ? &tm5 : &tm4;
            const int nst = l < 3 ? 2 : 4;
            for (int i = 0; i < nst; ++i, ++pk) {
                const int slot = pk % MM_NS;
                if (pk >= MM_NS) tc::mbar_wait(&empty[slot], ((pk / MM_NS) - 1) & 1);
                if (lane == 0) {
                    if (i == 0) {
                        volatile int4* d = &s_tq[tk & 3];
                        d->x = q, d->y = l + 8 * b, d->z = tyi * MM_TH, d->w = txi * MM_TW;
                        tc::fence_proxy_async_all();  // the acquired tiles were written through the generic proxy
                    }
                    tc::mbar_arrive_expect_tx(&full[slot], MM_STAGE_BYTES);
                    tc::tma_load_3d(tc::smem_u32(ring + slot * MM_STAGE_BYTES), i < 2 ? mA : mB, txi * MM_TW - MM_XOFF,
                                    tyi * MM_TH - 1, b * 16 + (i & 1) * 8, &full[slot]);
                }
                __syncwarp();
            }
            ++tk;
        }
        return;
    }

    // ================= consumer warps =================
    int ck = 0, tk = 0, cur_l = -1;
    for (;;) {
        tc::mbar_wait(&full[ck % MM_NS], (ck / MM_NS) & 1);  // the tile's first stage (or the end marker)
        const int q = s_tq[tk & 3].x;
        if (q < 0) break;
        const int lb = s_tq[tk & 3].y, y0 = s_tq[tk & 3].z, x0 = s_tq[tk & 3].w;
        const int l = lb & 7, b = lb >> 3;
        const uint32_t* wl = nullptr;  // selected with constant indices: the kernel parameters stay in the constant
        float* ds = nullptr;           // bank (a dynamic index would copy them to the stack)
#pragma unroll
        for (int i = 0; i < 6; ++i)
            if (i == l) wl = a.wl[i], ds = a.dst[i];
        if (l != cur_l) {  // every consumer warp is past the previous tile's closing barrier: the old block is dead
            const int nw4 = (l < 3 ? mm_words(1) : (l < 5 ? mm_words(4) : mm_words(6))) / 4;
            for (int i = t; i < nw4; i += 256)
                reinterpret_cast<uint4*>(s_w)[i] = __ldg(reinterpret_cast<const uint4*>(wl) + i);
            tc::named_bar_sync(1, 256);
            cur_l = l;
        }
        if (l < 3)
            mm_tile<1, 4, false>(ring, full, empty, ck, s_w, ds, h, w, wp, b, y0, x0);
        else if (l < 5)
            mm_tile<2, 4, false>(ring, full, empty, ck, s_w, ds, h, w, wp, b, y0, x0);
        else
            mm_tile<2, 1, true>(ring, full, empty, ck, s_w, ds, h, w, wp, b, y0, x0);
        tc::named_bar_sync(1, 256);  // all of this tile's stores are issued
        if (l < 5 && t == 0) mm_st_release(flags + q, 1);
        ++tk;
    }
}

// [B][16 pairs][h][wp] float2 as a 3-D tensor of 8-byte elements (x, y, pair + 16*b); box = one stage
static int mm_make_map(CUtensorMap* m, const void* base, int B, int h, int w) {
    tma_encode_fn enc = tma_encoder();
    if (!enc) return B200_ERR_LAUNCH;
    const int wp = mm_pitch(w);
    const cuuint64_t dims[3] = {(cuuint64_t)w, (cuuint64_t)h, (cuuint64_t)16 * B};
    const cuuint64_t strides[2] = {(cuuint64_t)wp * 8, (cuuint64_t)h * wp * 8};
    const cuuint32_t box[3] = {MM_HC, MM_HR, MM_STAGE_PAIRS};
    const cuuint32_t estr[3] = {1, 1, 1};
    const CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_UINT64, 3, const_cast<void*>(base), dims, strides, box, estr,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? B200_OK : B200_ERR_LAUNCH;
}

// x[0..4] = x1 x2 x3 x4 x5 (x6 reuses x4's buffer)
static int launch_mm_net(const MmNetArgs& a, float* const* x, int B, int h, int w, int* sync, cudaStream_t s) {
    auto kern = dcepp_net_kernel;
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, MM_NET_SMEM) != cudaSuccess) {
        (void)cudaGetLastError();
        return B200_ERR_LAUNCH;
    }
    int occ = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, MM_THREADS, MM_NET_SMEM) != cudaSuccess || occ < 1) {
        (void)cudaGetLastError();
        return B200_ERR_LAUNCH;  // the wait-on-earlier-tiles scheme needs every CTA resident
    }
    int dev = 0, sms = B200_NUM_SMS;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    CUtensorMap m[5];
    memset(m, 0, sizeof(m));
    for (int i = 0; i < 5; ++i) {
        const int rc = mm_make_map(&m[i], x[i], B, h, w);
        if (rc) return rc;
    }
    const long long total = 6LL * B * ((h + MM_TH - 1) / MM_TH) * ((w + MM_TW - 1) / MM_TW);
    long long grid = (long long)sms * occ;
    if (grid > total) grid = total;
    kern<<<(unsigned)grid, MM_THREADS, MM_NET_SMEM, s>>>(a, m[0], m[1], m[2], m[3], m[4], B, h, w, sync);
    return launch_status();
}

}  // namespace b200
